// tracer_tma.cuh -- horizontal tracer advection (pre_step3d4S.F:152-168 predictor C4, step3d_t_ISO.F:196-213 corrector UP3;
// compute_horiz_tracer_fluxes.h:1-246) with every operand staged by the TMA unit (sm_100a).
//
// k_pre_t_horiz / k_t_horiz (kernels.cuh) fetch the nine tracer values, four volume fluxes and the own-point operands of a
// point with ~25 scalar loads per level; ncu shows them waiting on those loads (long-scoreboard stall 13-24 cycles per issued
// instruction, IPC 1.1-1.3, 33-45 % of the HBM peak).  Here a block owns a 32x8 tile and marches over KB levels.  A producer
// warp issues, per level, cp.async.bulk.tensor copies of
//     FlxU, FlxV                 (34x9 box: the faces c, c+1 / c, c+SX of every cell)
//     [Hz, FlxDiv]               (32x8, predictor)
//     per tracer  t(nrhs)        (36x12: the tile and its 2-cell halo)
//                 t(nnew)  or, predictor, t(nstp), t(indx)   (32x8)
// into a ring of NS shared-memory stages, each completing on an mbarrier, NS-1 levels ahead of the arithmetic; the 256 compute
// threads only read shared memory and store results.  ONE tensor map per box shape describes the whole arena as a 2-D tensor
// (SX columns x all rows of all planes: every field is a stack of (Mm+4)-row planes with the same pitch), so an operand is
// addressed by (column, plane_row0 + row): three CUtensorMap objects serve every field, tracer and time level.
// Points within two cells of a closed wall and river cells take the global-memory path of the original kernels (same
// device functions).  Arithmetic: expression for expression that of k_pre_t_horiz / k_t_horiz.
#pragma once
#include <cuda.h>
#include <cstdint>

struct TmaMaps {
  CUtensorMap t36x12, f34x9, o32x8;      // 2-D boxes over the arena seen as SX x (all rows)
};
template <int TB>
struct TrcRows {      // first arena row (plane index x (Mm+4)) of level k=1 of every operand
  int flxu, flxv, hz, fd;
  int rhs[TB], tnew[TB], stp[TB], indx[TB];
};
#define TT_W 36
#define TT_H 12
#define TF_W 34
#define TF_H 9
#define TT_N (TT_W * TT_H)       // 432 doubles = 3456 B (multiple of 128)
#define TF_N 320                 // 34*9 = 306 doubles, padded to 2560 B
#define TO_N 256                 // 32*8
template <bool PRE, int TB>
struct TtmaLayout {
  static constexpr int flxu = 0, flxv = TF_N, hz = 2 * TF_N, fd = 2 * TF_N + TO_N;
  static constexpr int trc0 = PRE ? 2 * TF_N + 2 * TO_N : 2 * TF_N;
  static constexpr int per_trc = PRE ? TT_N + 2 * TO_N : TT_N + TO_N;      // predictor: t(nrhs) tile, t(nstp), t(indx); corrector: t(nrhs) tile, t(nnew)
  static constexpr int stage = trc0 + TB * per_trc;                          // doubles per stage
  static constexpr unsigned bytes = (2 * TF_W * TF_H + (PRE ? 2 * TO_N : 0) + TB * (TT_N + (PRE ? 2 : 1) * TO_N)) * 8;   // bytes that arrive per stage
};
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int x, int y, uint64_t* bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(s2m_u32(dst)), "l"(map), "r"(x),
               "r"(y), "r"(s2m_u32(bar))
               : "memory");
}

// The wall path of k_rhs_uv_tma is OUT OF LINE: inlined it costs the hot path its registers (144-176 B of spills at the 72-register
// bound; UP3 1.61 -> 1.39 ms, C4 1.38 -> 1.14 ms on the same box); only the threads next to a wall pay for the call.  (The same move
// made k_t_horiz_tma slower, 1.20 -> 1.29 ms, and left k_prsgrd_tma unchanged: those keep their wall paths inline.)
template <bool UP>
__device__ __noinline__ void rhs_uv_point_near_ool(const KP& p, const double* __restrict__ u, const double* __restrict__ v, ixr_t c, int i, int j, int k, bool dov,
                                                   double& ru, double& rv) {
  rhs_uv_point<UP, true>(p, u, v, c, i, j, k, dov, ru, rv);
}

template <bool PRE, bool WALL, bool RIV, int TB, int KB, int NS>
__global__ void __launch_bounds__(288, (PRE ? 2 : 3))
k_t_horiz_tma(const __grid_constant__ KP p, const __grid_constant__ TrcBatch<TB> tb, const __grid_constant__ TmaMaps tm, const __grid_constant__ TrcRows<TB> tr,
              double dtau, double cf_stp, double cf_bak) {
  typedef TtmaLayout<PRE, TB> L;
  extern __shared__ __align__(128) unsigned char ttma_raw[];
  double* st = (double*)ttma_raw;                              // [NS][L::stage]
  uint64_t* full = (uint64_t*)(st + NS * L::stage);            // [NS]
  const int tid = threadIdx.x;
  const int bx = blockIdx.x, by = blockIdx.y;
  const int j0 = p.jstr + by * 8;
  const int k0 = blockIdx.z * KB + 1;
  const int nlev = min(KB, p.N - k0 + 1);
  const int R = p.Mm + 4;
  if (tid == 0) {
    for (int q = 0; q < NS; ++q) s2m_mbar_init(&full[q], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (tid >= 256) {
    // ---- producer warp: lane 0 arms the barrier of the stage, then the copies of one level
    auto issue = [&](int q) {
      if (q >= nlev || tid != 256) return;
      double* sg = st + (q % NS) * L::stage;
      uint64_t* bar = &full[q % NS];
      const int krow = (k0 + q - 1) * R;
      s2m_expect_tx(bar, L::bytes);
      const int xo = bx * 32, yo = krow + j0 + 1;              // own column / row of the tile's first point
      tma_load_2d(sg + L::flxu, &tm.f34x9, xo, tr.flxu + yo, bar);
      tma_load_2d(sg + L::flxv, &tm.f34x9, xo, tr.flxv + yo, bar);
      if (PRE) {
        tma_load_2d(sg + L::hz, &tm.o32x8, xo, tr.hz + yo, bar);
        tma_load_2d(sg + L::fd, &tm.o32x8, xo, tr.fd + yo, bar);
      }
#pragma unroll
      for (int t = 0; t < TB; ++t) {
        double* s0 = sg + L::trc0 + t * L::per_trc;
        tma_load_2d(s0, &tm.t36x12, xo - 2, tr.rhs[t] + yo - 2, bar);
        if (PRE) {
          tma_load_2d(s0 + TT_N, &tm.o32x8, xo, tr.stp[t] + yo, bar);
          tma_load_2d(s0 + TT_N + TO_N, &tm.o32x8, xo, tr.indx[t] + yo, bar);
        } else {
          tma_load_2d(s0 + TT_N, &tm.o32x8, xo, tr.tnew[t] + yo, bar);
        }
      }
    };
    for (int q = 0; q < NS; ++q) issue(q);
    for (int q = 0; q < nlev; ++q) {
      __syncthreads();                                         // every compute thread is done with level q: its stage is free
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      issue(q + NS);
    }
    return;
  }
  // ---- compute threads
  const int tx = tid & 31, ty = tid >> 5;
  const int i = bx * 32 + tx - IOFF, j = j0 + ty;
  const bool active = i >= p.istr && i <= p.iend && j <= p.jend;
  const ix_t c = CIDX(active ? i : p.istr, active ? j : p.jstr);
  const ix_t SX = p.SX;
  (void)SX;
  TMask m;
  m.u0 = m.u1 = m.u2 = m.u3 = m.v0 = m.v1 = m.v2 = m.v3 = 0.0;
  double cdt = 0.0;
  bool near = false;
  RivFaces rf;
  rf.any = false;
  if (active) {
    m = tmask_load(p, c);
    cdt = (PRE ? dtau : p.dt) * p.pm[c] * p.pn[c];
    near = WALL && near_wall(p, i, j);
    if (RIV) rf = riv_faces(p, c);
  }
  for (int q = 0; q < nlev; ++q) {
    const int k = k0 + q;
    const double* sg = st + (q % NS) * L::stage;
    s2m_wait(&full[q % NS], (uint32_t)((q / NS) & 1));
    if (active) {
      Flx4 f;
      f.u0 = sg[L::flxu + ty * TF_W + tx]; f.u1 = sg[L::flxu + ty * TF_W + tx + 1];
      f.v0 = sg[L::flxv + ty * TF_W + tx]; f.v1 = sg[L::flxv + (ty + 1) * TF_W + tx];
      double Hz = 0.0, Hz_bak = 0.0;
      if (PRE) { Hz = sg[L::hz + ty * 32 + tx]; Hz_bak = Hz + sg[L::fd + ty * 32 + tx]; }
#pragma unroll
      for (int t = 0; t < TB; ++t) {
        const double* s0 = sg + L::trc0 + t * L::per_trc;
        double dv;
        if ((RIV && rf.any) || near) {
          const double* __restrict__ t_rhs = tb.rhs[t];
          if (RIV && rf.any) dv = tdiv_riv<!PRE>(p, t_rhs, c, i, j, k, m, f, near, rf, tb.itrc[t]);
          else dv = tdiv<!PRE, true>(p, t_rhs, c, i, j, k, m, f);
        } else {
          // tflux4<UP, false> with the five values along each axis read from the staged tile
          const double* sc = s0 + (ty + 2) * TT_W + tx + 2;
          const double a0 = sc[-2], a1 = sc[-1], t2 = sc[0], a3 = sc[1], a4 = sc[2];
          const double dA = (a1 - a0) * m.u0, dB = (t2 - a1) * m.u1, dC = (a3 - t2) * m.u2, dD = (a4 - a3) * m.u3;
          const double FXi = tflux_fin<!PRE>(a1, t2, dA, dB, dC, f.u0);
          const double FXip = tflux_fin<!PRE>(t2, a3, dB, dC, dD, f.u1);
          const double b0 = sc[-2 * TT_W], b1 = sc[-TT_W], b3 = sc[TT_W], b4 = sc[2 * TT_W];
          const double eA = (b1 - b0) * m.v0, eB = (t2 - b1) * m.v1, eC = (b3 - t2) * m.v2, eD = (b4 - b3) * m.v3;
          const double FEj = tflux_fin<!PRE>(b1, t2, eA, eB, eC, f.v0);
          const double FEjp = tflux_fin<!PRE>(t2, b3, eB, eC, eD, f.v1);
          dv = FXip - FXi + FEjp - FEj;
        }
        if (PRE) {
          const double ts = s0[TT_N + ty * 32 + tx];
          R3(tb.tnew[t], c, k) = Hz_bak * (cf_stp * ts + cf_bak * s0[TT_N + TO_N + ty * 32 + tx]) - cdt * (dv);
          R3(tb.indx[t], c, k) = Hz * ts;
        } else {
          R3(tb.tnew[t], c, k) = s0[TT_N + ty * 32 + tx] - cdt * (dv);
        }
      }
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------------
// Horizontal momentum right-hand side (compute_horiz_rhs_uv_terms.h:1-291: Coriolis + UP3 / C4 advection of u and v) with the
// same TMA staging: per level the 36x12 boxes of u, v, FlxU, FlxV, Hz and the
// 32x8 boxes of ru, rv (the pressure-gradient term left there by prsgrd).  Replaces k_rhs_uv_horiz (~110 scalar loads per
// point through L1).  Points within two cells of a closed wall take rhs_uv_point<UP, true> on global memory.
struct RhsRows { int u, v, flxu, flxv, hz, ru, rv; };
struct RhsLayout {
  static constexpr int u = 0, v = TT_N, fu = 2 * TT_N, fv = 3 * TT_N, hz = 4 * TT_N, ru = 5 * TT_N, rv = 5 * TT_N + TO_N;
  static constexpr int stage = 5 * TT_N + 2 * TO_N;                        // 2672 doubles = 21376 B (multiple of 128)
  static constexpr unsigned bytes = (5 * TT_N + 2 * TO_N) * 8;
};
template <bool UP, bool WALL, int KB, int NS>
__global__ void __launch_bounds__(288, 3)
k_rhs_uv_tma(const __grid_constant__ KP p, const __grid_constant__ TmaMaps tm, const __grid_constant__ RhsRows tr, const double* __restrict__ u, const double* __restrict__ v) {
  typedef RhsLayout L;
  extern __shared__ __align__(128) unsigned char ttma_raw[];
  double* st = (double*)ttma_raw;
  uint64_t* full = (uint64_t*)(st + NS * L::stage);
  const int tid = threadIdx.x;
  const int bx = blockIdx.x, by = blockIdx.y;
  const int j0 = p.jstr + by * 8;
  const int k0 = blockIdx.z * KB + 1;
  const int nlev = min(KB, p.N - k0 + 1);
  const int R = p.Mm + 4;
  if (tid == 0) {
    for (int q = 0; q < NS; ++q) s2m_mbar_init(&full[q], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (tid >= 256) {
    auto issue = [&](int q) {
      if (q >= nlev || tid != 256) return;
      double* sg = st + (q % NS) * L::stage;
      uint64_t* bar = &full[q % NS];
      const int xo = bx * 32, yo = (k0 + q - 1) * R + j0 + 1;
      s2m_expect_tx(bar, L::bytes);
      tma_load_2d(sg + L::u, &tm.t36x12, xo - 2, tr.u + yo - 2, bar);
      tma_load_2d(sg + L::v, &tm.t36x12, xo - 2, tr.v + yo - 2, bar);
      tma_load_2d(sg + L::fu, &tm.t36x12, xo - 2, tr.flxu + yo - 2, bar);
      tma_load_2d(sg + L::fv, &tm.t36x12, xo - 2, tr.flxv + yo - 2, bar);
      tma_load_2d(sg + L::hz, &tm.t36x12, xo - 2, tr.hz + yo - 2, bar);     // (a box must start on a 16-byte boundary: even column)
      tma_load_2d(sg + L::ru, &tm.o32x8, xo, tr.ru + yo, bar);
      tma_load_2d(sg + L::rv, &tm.o32x8, xo, tr.rv + yo, bar);
    };
    for (int q = 0; q < NS; ++q) issue(q);
    for (int q = 0; q < nlev; ++q) {
      __syncthreads();
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      issue(q + NS);
    }
    return;
  }
  const int tx = tid & 31, ty = tid >> 5;
  const int i = bx * 32 + tx - IOFF, j = j0 + ty;
  const bool active = i >= p.istr && i <= p.iend && j <= p.jend;
  const ixr_t c = CIDX(active ? i : p.istr, active ? j : p.jstr);
  const bool near = active && WALL && near_wall(p, i, j);
  double fomn_c = 0.0, fomn_w = 0.0, fomn_s = 0.0;
  if (active) { fomn_c = p.fomn[c]; fomn_w = p.fomn[c - 1]; fomn_s = p.fomn[c - p.SX]; }
  const int cl = (ty + 2) * TT_W + tx + 2;          // this thread's point inside a 36x12 box
  for (int q = 0; q < nlev; ++q) {
    const int k = k0 + q;
    const double* sg = st + (q % NS) * L::stage;
    s2m_wait(&full[q % NS], (uint32_t)((q / NS) & 1));
    if (active) {
      double ru, rv;
      bool dou = true, dov = true;
      if (near) {
        dov = j >= p.jstrV; dou = i >= p.istrU;
        rhs_uv_point_near_ool<UP>(p, u, v, c, i, j, k, dov, ru, rv);
      } else {
        // rhs_uv_point<UP, false> on the staged boxes (row pitch TT_W instead of SX, level offset 0)
        const double* U = sg + L::u;
        const double* V = sg + L::v;
        const double* FU = sg + L::fu;
        const double* FV = sg + L::fv;
        const double* sh = sg + L::hz + cl;
        const double cf0 = 0.5 * sh[0] * (fomn_c);
        const double cfw = 0.5 * sh[-1] * (fomn_w);
        const double v00 = V[cl], u00 = U[cl];
        const double UFx_c = cf0 * (v00 + V[cl + TT_W]);
        const double UFx_w = cfw * (V[cl - 1] + V[cl - 1 + TT_W]);
        ru = sg[L::ru + ty * 32 + tx] + 0.5 * (UFx_c + UFx_w);
        const double UFx_i = flux_along<UP, false>(p, U, FU, cl, 1, 1);
        const double UFx_im = flux_along<UP, false>(p, U, FU, cl - 1, 1, 1);
        const double UFe_jp = flux_cross<UP, false>(p, U, FV, cl + TT_W, TT_W, 1, 1);
        const double UFe_j = flux_cross<UP, false>(p, U, FV, cl, TT_W, 1, 1);
        ru = ru - UFx_i + UFx_im - UFe_jp + UFe_j;
        const double cfs = 0.5 * sh[-TT_W] * (fomn_s);
        const double VFe_c = cf0 * (u00 + U[cl + 1]);
        const double VFe_s = cfs * (U[cl - TT_W] + U[cl - TT_W + 1]);
        rv = sg[L::rv + ty * 32 + tx] - 0.5 * (VFe_c + VFe_s);
        const double VFx_ip = flux_cross<UP, false>(p, V, FU, cl + 1, 1, TT_W, 1);
        const double VFx_i = flux_cross<UP, false>(p, V, FU, cl, 1, TT_W, 1);
        const double VFe_j = flux_along<UP, false>(p, V, FV, cl, TT_W, 1);
        const double VFe_jm = flux_along<UP, false>(p, V, FV, cl - TT_W, TT_W, 1);
        rv = rv - VFx_ip + VFx_i - VFe_j + VFe_jm;
      }
      if (dou) R3(p.ru, c, k) = ru;
      if (dov) R3(p.rv, c, k) = rv;
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------------
// prsgrd.F:231-421, horizontal part (nonlinear split EOS), with TMA staging: per level the 36x12 boxes of z_r, rho1, qp1
// (the four points c-2d..c+d of both directions) and the 34x9 boxes of P, rho, Hz (origin (i0-2, j0-1): the points c-d, c).
// Replaces k_prsgrd_horiz<true, *, *>.  Points within two cells of a closed wall take prs_dir_cl on global memory.
struct PrsRows { int zr, r1, q1, P, rho, hz; };
struct PrsLayout {
  static constexpr int zr = 0, r1 = TT_N, q1 = 2 * TT_N, P = 3 * TT_N, rho = 3 * TT_N + TF_N, hz = 3 * TT_N + 2 * TF_N;
  static constexpr int stage = 3 * TT_N + 3 * TF_N;                        // 2256 doubles = 18048 B (multiple of 128)
  static constexpr unsigned bytes = (3 * TT_N + 3 * TF_W * TF_H) * 8;
};
// prs_dir<true> on staged boxes: zr, r1, q1 share the 36-wide layout (centre cl, stride d), P, rho, Hz the 34-wide one (centre ch, stride dh)
__device__ __forceinline__ double prs_dir_s(const KP& p, const double* __restrict__ zr, const double* __restrict__ r1, const double* __restrict__ q1v, int cl, int d,
                                            const double* __restrict__ sP, const double* __restrict__ srho, const double* __restrict__ sHz, int ch, int dh, double m1,
                                            double m2, double m3, double dn) {
  const double qp2 = 0.0000172;
  const double z0 = zr[cl - 2 * d], z1 = zr[cl - d], z2 = zr[cl], z3 = zr[cl + d];
  const double FCa = (z1 - z0) * m1, FCb = (z2 - z1) * m2, FCc = (z3 - z2) * m3;   // FC(i-1), FC(i), FC(i+1)
  const double a0 = r1[cl - 2 * d], a1 = r1[cl - d], a2 = r1[cl], a3 = r1[cl + d];
  const double b0 = q1v[cl - 2 * d], b1 = q1v[cl - d], b2 = q1v[cl], b3 = q1v[cl + d];
  double dpth = -0.5 * (z1 + z0);
  const double rxa = (a1 - a0 + (b1 - b0) * dpth * (1.0 - qp2 * dpth)) * m1;
  dpth = -0.5 * (z2 + z1);
  const double rxb = (a2 - a1 + (b2 - b1) * dpth * (1.0 - qp2 * dpth)) * m2;
  dpth = -0.5 * (z3 + z2);
  const double rxc = (a3 - a2 + (b3 - b2) * dpth * (1.0 - qp2 * dpth)) * m3;
  return prs_tail_core<true>(p, dn, FCa, FCb, FCc, rxa, rxb, rxc, b1, b2, z1, z2, srho[ch], srho[ch - dh], sHz[ch], sHz[ch - dh], sP[ch - dh], sP[ch]);
}
template <bool WALL, int KB, int NS, int MINB>
__global__ void __launch_bounds__(288, MINB) k_prsgrd_tma(const __grid_constant__ KP p, const __grid_constant__ TmaMaps tm, const __grid_constant__ PrsRows tr) {
  typedef PrsLayout L;
  extern __shared__ __align__(128) unsigned char ttma_raw[];
  double* st = (double*)ttma_raw;
  uint64_t* full = (uint64_t*)(st + NS * L::stage);
  const int tid = threadIdx.x;
  const int bx = blockIdx.x, by = blockIdx.y;
  const int j0 = p.jstr + by * 8;
  const int k0 = blockIdx.z * KB + 1;
  const int nlev = min(KB, p.N - k0 + 1);
  const int R = p.Mm + 4;
  if (tid == 0) {
    for (int q = 0; q < NS; ++q) s2m_mbar_init(&full[q], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (tid >= 256) {
    auto issue = [&](int q) {
      if (q >= nlev || tid != 256) return;
      double* sg = st + (q % NS) * L::stage;
      uint64_t* bar = &full[q % NS];
      const int xo = bx * 32, yo = (k0 + q - 1) * R + j0 + 1;
      s2m_expect_tx(bar, L::bytes);
      tma_load_2d(sg + L::zr, &tm.t36x12, xo - 2, tr.zr + yo - 2, bar);
      tma_load_2d(sg + L::r1, &tm.t36x12, xo - 2, tr.r1 + yo - 2, bar);
      tma_load_2d(sg + L::q1, &tm.t36x12, xo - 2, tr.q1 + yo - 2, bar);
      tma_load_2d(sg + L::P, &tm.f34x9, xo - 2, tr.P + yo - 1, bar);
      tma_load_2d(sg + L::rho, &tm.f34x9, xo - 2, tr.rho + yo - 1, bar);
      tma_load_2d(sg + L::hz, &tm.f34x9, xo - 2, tr.hz + yo - 1, bar);
    };
    for (int q = 0; q < NS; ++q) issue(q);
    for (int q = 0; q < nlev; ++q) {
      __syncthreads();
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      issue(q + NS);
    }
    return;
  }
  const int tx = tid & 31, ty = tid >> 5;
  const int i = bx * 32 + tx - IOFF, j = j0 + ty;
  const bool active = i >= p.istr && i <= p.iend && j <= p.jend;
  const ix_t c = CIDX(active ? i : p.istr, active ? j : p.jstr);
  const ix_t SX = p.SX;
  const bool near = active && WALL && near_wall(p, i, j);
  double dnu = 0, dmv = 0, mu1 = 0, mu2 = 0, mu3 = 0, mv1 = 0, mv2 = 0, mv3 = 0;
  ix_t oA = 0, oB = 0, oC = 0, iA = 0, iB = 0, iC = 0;
  if (active) {
    dnu = p.dn_u[c]; dmv = p.dm_v[c];
    mu1 = p.umask[c - 1]; mu2 = p.umask[c]; mu3 = p.umask[c + 1];
    mv1 = p.vmask[c - SX]; mv2 = p.vmask[c]; mv3 = p.vmask[c + SX];
    if (near) {
      oA = jclamp(p, j - 1, p.jstrV, p.jend); oB = jclamp(p, j, p.jstrV, p.jend); oC = jclamp(p, j + 1, p.jstrV, p.jend);
      iA = iclamp(p, i - 1, p.istrU, p.iend); iB = iclamp(p, i, p.istrU, p.iend); iC = iclamp(p, i + 1, p.istrU, p.iend);
    }
  }
  const int cl = (ty + 2) * TT_W + tx + 2, ch = (ty + 1) * TF_W + tx + 2;
  for (int q = 0; q < nlev; ++q) {
    const int k = k0 + q;
    const double* sg = st + (q % NS) * L::stage;
    s2m_wait(&full[q % NS], (uint32_t)((q / NS) & 1));
    if (active) {
      if (near) {     // prsgrd.F:288-310: FC,rx exist on rows jstrV..jend / columns istrU..iend between walls, copied outward
        if (i >= p.istrU) R3(p.ru, c, k) = prs_dir_cl<true>(p, c, 1, k, p.umask, dnu, p.rho, iA, iB, iC);
        if (j >= p.jstrV) R3(p.rv, c, k) = prs_dir_cl<true>(p, c, SX, k, p.vmask, dmv, p.rho, oA, oB, oC);
      } else {
        const double ru = prs_dir_s(p, sg + L::zr, sg + L::r1, sg + L::q1, cl, 1, sg + L::P, sg + L::rho, sg + L::hz, ch, 1, mu1, mu2, mu3, dnu);
        const double rv = prs_dir_s(p, sg + L::zr, sg + L::r1, sg + L::q1, cl, TT_W, sg + L::P, sg + L::rho, sg + L::hz, ch, TF_W, mv1, mv2, mv3, dmv);
        R3(p.ru, c, k) = ru;
        R3(p.rv, c, k) = rv;
      }
    }
    __syncthreads();
  }
}
