// decomp.h -- 2-D subdomain decomposition and halo-message geometry (host-only, no CUDA).
//
// Replaces mpi_setup.F:15-262 (rank grid, neighbour ranks with periodic wrap) and the message
// geometry of mpi_exchanges.F:528-670 (8 neighbours, halo width 2).  One GPU = one subdomain =
// one "tile" of the reference (compute_tile_bounds.h with NSUB_X=NSUB_E=1).
//
// Directions d=0..7 enumerate (dx,dy) in {-1,0,1}^2 \ (0,0):
//   0:W(-1,0) 1:E(+1,0) 2:S(0,-1) 3:N(0,+1) 4:SW 5:SE 6:NW 7:NE ; opposite(d) flips both signs.
#pragma once

struct DirBox { int i0, j0, ni, nj; };   // inclusive start, extents (model indices, halo = -1,0 / L+1,L+2)

static inline void decomp_dir(int d, int* dx, int* dy) {
  static const int DX[8] = {-1, +1, 0, 0, -1, +1, -1, +1};
  static const int DY[8] = {0, 0, -1, +1, -1, -1, +1, +1};
  *dx = DX[d]; *dy = DY[d];
}
static inline int decomp_opposite(int d) {
  static const int OPP[8] = {1, 0, 3, 2, 7, 6, 5, 4};
  return OPP[d];
}
// rank = jnode*np_xi + inode (mpi_setup.F:61-70).  Returns the neighbour rank in direction d or -1 when
// the subdomain touches a non-periodic physical edge there (mpi_setup.F:231-250 wraps periodic ones).
static inline int decomp_neighbor(int np_xi, int np_eta, int inode, int jnode, int ew_periodic, int ns_periodic, int d) {
  int dx, dy;
  decomp_dir(d, &dx, &dy);
  int in = inode + dx, jn = jnode + dy;
  if (in < 0 || in >= np_xi) {
    if (!ew_periodic) return -1;
    in = (in + np_xi) % np_xi;
  }
  if (jn < 0 || jn >= np_eta) {
    if (!ns_periodic) return -1;
    jn = (jn + np_eta) % np_eta;
  }
  return jn * np_xi + in;
}
// Block extent along one direction (mpi_setup.F:117-160): every block gets L0 = ceil(LL/np) interior points, the
// surplus off = np*L0 - LL is trimmed from the two physical-edge blocks (off/2 from the first, (off+1)/2 from the
// last); corn = global index offset of the block (iSW_corn / jSW_corn: local i=1 is global corn+1).
static inline void decomp_extent(int LL, int np, int node, int* L, int* corn) {
  const int L0 = (LL + np - 1) / np, off = np * L0 - LL;
  int l = L0;
  if (node == 0) l -= off / 2;
  if (node == np - 1) l -= (off + 1) / 2;
  *L = l;
  *corn = node == 0 ? 0 : node * L0 - off / 2;
}
// interior cells this subdomain SENDS to its neighbour in direction d
// swall/nwall: this subdomain touches a closed southern/northern physical edge; E-W messages then also carry
// the boundary rows j=0 / j=Mm+1 (mpi_exchanges.F:538-543: "if (jnode==0) jl0=0", "if (jnode==NP_ETA-1) jl1=nyl+1");
// wwall/ewall: likewise N-S messages carry the boundary columns i=0 / i=Lm+1.
static inline DirBox decomp_send_box(int Lm, int Mm, int d, int swall = 0, int nwall = 0, int wwall = 0, int ewall = 0) {
  int dx, dy;
  decomp_dir(d, &dx, &dy);
  DirBox b;
  b.ni = dx == 0 ? Lm : 2; b.nj = dy == 0 ? Mm : 2;
  b.i0 = dx > 0 ? Lm - 1 : 1;
  b.j0 = dy > 0 ? Mm - 1 : 1;
  if (dx != 0 && dy == 0) { b.j0 -= swall ? 1 : 0; b.nj += (swall ? 1 : 0) + (nwall ? 1 : 0); }
  if (dy != 0 && dx == 0) { b.i0 -= wwall ? 1 : 0; b.ni += (wwall ? 1 : 0) + (ewall ? 1 : 0); }
  return b;
}
// halo cells this subdomain RECEIVES from its neighbour in direction d
static inline DirBox decomp_recv_box(int Lm, int Mm, int d, int swall = 0, int nwall = 0, int wwall = 0, int ewall = 0) {
  int dx, dy;
  decomp_dir(d, &dx, &dy);
  DirBox b;
  b.ni = dx == 0 ? Lm : 2; b.nj = dy == 0 ? Mm : 2;
  b.i0 = dx < 0 ? -1 : (dx > 0 ? Lm + 1 : 1);
  b.j0 = dy < 0 ? -1 : (dy > 0 ? Mm + 1 : 1);
  if (dx != 0 && dy == 0) { b.j0 -= swall ? 1 : 0; b.nj += (swall ? 1 : 0) + (nwall ? 1 : 0); }
  if (dy != 0 && dx == 0) { b.i0 -= wwall ? 1 : 0; b.ni += (wwall ? 1 : 0) + (ewall ? 1 : 0); }
  return b;
}
