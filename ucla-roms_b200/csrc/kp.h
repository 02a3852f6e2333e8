// kp.h -- kernel parameter block and device indexing for libromsb200.
//
// HBM layout: every field keeps the reference's (i,j,k) order, i fastest, 2-cell halo
// (set_global_definitions.h:146-150), but rows are padded so that interior point i=1 of every
// row of every plane sits on a 128-byte boundary: element (i,j[,k]) lives at
//   base[ (i + IOFF) + (j+1)*SX + kplane*SXY ],   IOFF = 15, SX = roundup(Lm+18, 16), SXY = SX*(Mm+4)
// so a warp reading 32 consecutive i of one row touches exactly two 128-B lines.
#pragma once
#include <cstddef>

// element index inside one field: 32-bit is always enough on one GPU (a field of 2^31 doubles would need 16 GiB per 3-D
// array and the step keeps 22+3*NT of them resident); roms_create checks it.  Halves the integer address arithmetic.
typedef int ix_t;

#define IOFF 15
#define MAXNT 40
#define MAXRIV 8
#define RIVROW (1 + MAXNT)
#define MAXPIP 8
#define PIPROW(N) ((N) + MAXNT)

struct KP {
  int Lm, Mm, N, NT;
  int SX;            // padded row pitch (doubles)
  ix_t SXY;          // padded plane size (doubles)
  int istr, iend, jstr, jend;   // compute_tile_bounds.h with NSUB_X=NSUB_E=1
  int swall, nwall;             // this subdomain touches a closed (non-periodic) southern / northern physical edge
  int jstrR, jendR, jstrV;      // compute_auxiliary_bounds.h: jstr-1 / jend+1 / jstr+1 on such an edge, else jstr / jend / jstr
  int wwall, ewall;             // likewise a closed western / eastern physical edge
  int istrR, iendR, istrU;      // istr-1 / iend+1 / istr+1 on such an edge, else istr / iend / istr
  int nonlin, salinity;
  int wrap_x, wrap_y;           // single subdomain along a periodic direction: the fused 2-D kernel stores the halo images itself
  // scalars
  double dt, dtfast, g, rho0, Tcoef, T0, Scoef, S0, rdrg, rdrg2, Zob, gamma2, hc;
  // static 2-D
  double *h, *pm, *pn, *f, *rmask, *umask, *vmask, *dm_u, *dn_u, *dm_v, *dn_v, *dm_r, *dn_r, *fomn, *hinv;
  double *pmask, *dm_p, *dn_p, *pmon_u, *pnom_v, *visc2_r, *visc2_p;   // lateral mixing (visc3d_S.F, t3dmix_S.F)
  // 2-D state
  double *rhoA, *rhoS, *rufrc, *rvfrc, *Zt_avg1, *DU_avg1, *DV_avg1, *DU_avg2, *DV_avg2, *DU_avg_bak, *DV_avg_bak, *r_D;
  double *sustr, *svstr, *swflx;
  // 3-D state
  double *Hz, *z_r, *z_w, *FlxU, *FlxV, *We, *Wi, *rho1, *qp1, *Akv, *Akt[2], *ru, *rv;
  // device-owned scratch (the reference's A3d(1,3:4) and A2d planes)
  double *FlxDiv;    // Hz_bak = Hz + FlxDiv, Hz_fwd = Hz - FlxDiv (pre_step3d4S.F:138-147)
  double *P, *rho;   // prsgrd.F:12 P and in-situ rho
  double *scr;       // column scratch, 3 arrays of (N+1) planes (global-memory fallback)
  double *zeta_new;  // step2d_FB.F private zeta_new
  double *Cs_w, *Cs_r;   // scoord.F:38 (k=0:N, k=1:N at index k)
  int use_smem;      // column scratch in shared memory ([k][thread]) instead of global
  // river_frc.F: river faces (10*iriver + fraction of the river's volume flux), per-river table [iriver-1][0]=volume flux,
  // [iriver-1][1+itrc]=tracer value; rivers = river_source
  double *riv_uflx, *riv_vflx, *riv_tab;
  int rivers;
  // pipe_frc.F: pipe cells (index > 0) and their volume flux; per-pipe row of pipe_tab: release profile prf(1:N) then NT tracer values
  double *pipe_idx, *pipe_flx, *pipe_tab;
  int pipes;
  double* ptide;     // tides.F:26 equilibrium tidal potential
  int pot_tides;
  // LMD/KPP vertical mixing (lmd_vmix.F, lmd_kpp.F): allocated by roms_set_kpp
  int lmd_kpp;
  double *bvf, *swr_frac, *ghat;     // (0:N)
  double *hbls, *hbbl, *srflx;       // 2-D
  double *k_ustar, *k_Bo, *k_Bosol, *k_hbl, *k_bbl;   // 2-D work planes of lmd_kpp
  // open boundaries (obc.cuh): bit 0 west, 1 east, 2 south, 3 north edge is open (only looked at where the matching *wall flag says
  // that this subdomain touches that non-periodic physical edge).  bry2d / bry3d: the boundary points of ALL non-periodic edges are
  // set by the general boundary kernels of obc.cuh instead of the closed-wall code inside the interior kernels (always on when an
  // edge is open; ROMS_BRY=1 forces it for closed basins, which validates those kernels against the in-kernel wall code).
  int obc, bry2d, bry3d;
  int obc_m2, obc_m3, obc_t;   // 0 Flather / Orlanski / upstream (default), 1 specified, 2 zero-gradient (roms_cfg)
  double ubind;
  // boundary.F:21-60 boundary data, side 0 west, 1 east (lines over j = 0:Mm+1), 2 south, 3 north (lines over i = 0:Lm+1), the reference's
  // extents (boundary.F:111-131); the 3-D ones level-major [k-1][q]; bry_t: [itrc][k-1][q]
  double *bry_zeta[4], *bry_ubar[4], *bry_vbar[4], *bry_u[4], *bry_v[4], *bry_t[4];
};
#define BRYL(p, side) ((side) < 2 ? (p).Mm + 2 : (p).Lm + 2)

// column base index of (i,j) and relative addressing
#define CIDX(i, j) ((ix_t)((i) + IOFF) + (ix_t)((j) + 1) * p.SX)
#define R3(a, c, k) (a)[(c) + (ix_t)((k)-1) * p.SXY]    // rho/u/v-type, k=1:N
#define RW(a, c, k) (a)[(c) + (ix_t)(k) * p.SXY]        // w-type, k=0:N
