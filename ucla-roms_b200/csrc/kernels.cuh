// kernels.cuh -- fp64 CUDA kernels (sm_100a) for the UCLA-ROMS explicit stencil time step.
//
// Design: two kernel shapes only.
//  * "plane" kernels: one thread per (i,j,k) point, x along i (coalesced, 128-B aligned rows),
//    every horizontal flux recomputed in registers from the L1/L2-cached neighbourhood, so none
//    of the reference's 2-D scratch planes (UFx,UFe,VFx,VFe,wrk1,wrk2,FX,FE,...) exist.
//  * "column" kernels: one thread per (i,j) column, x along i; the k-recurrences (parabolic
//    splines, Thomas solves, omega integration) run in registers with a sliding window, and the
//    per-level coefficients that a back-substitution needs live in a full-size global scratch slab laid out
//    [k][j][i] (coalesced); a shared-memory variant exists (ROMS_COL_SMEM=1) but loses on occupancy.
// Operation order inside each expression follows the reference (file:line cited per kernel) so
// that results agree with the CPU oracle to rounding of fp64 division/log only.
#pragma once
#include "kp.h"

// minimum resident 256-thread blocks per SM asked of ptxas for the latency-bound plane kernels (tuned on B200)
#ifndef LB_PRS
#define LB_PRS 4
#endif
#ifndef LB_RHS
#define LB_RHS 4
#endif
// levels per thread (2-D operands stay in registers across them) and resident blocks of the k-blocked plane kernels
#ifndef KB_T
#define KB_T 8
#endif
#ifndef KB_PRS
#define KB_PRS 1
#endif
#ifndef LB_TP
#define LB_TP 5
#endif
#ifndef LB_TC
#define LB_TC 4
#endif
#ifndef LB_S2D
#define LB_S2D 4
#endif
#define DMAX(a, b) fmax((a), (b))
#define DMIN(a, b) fmin((a), (b))

struct ColScr {
  double* b;
  ix_t ks, as;
};
__device__ __forceinline__ ColScr col_scratch(const KP& p, ix_t c, double* smem) {
  ColScr s;
  if (p.use_smem) {
    const int nth = blockDim.x * blockDim.y;
    s.b = smem + threadIdx.y * blockDim.x + threadIdx.x;
    s.ks = nth;
    s.as = (ix_t)nth * (p.N + 1);
  } else {
    s.b = p.scr + c;
    s.ks = p.SXY;
    s.as = p.SXY * (p.N + 1);
  }
  return s;
}
#define S(n, k) sc.b[(ix_t)(n)*sc.as + (ix_t)(k)*sc.ks]

// Element-index width per kernel group, chosen by measurement on B200 (profiles/r1_tuning_log.txt): 32-bit indices
// cut the integer address arithmetic of the column and pressure-gradient kernels, while the momentum-RHS, barotropic
// and uv2/set_HUV1 kernels schedule their loads better with 64-bit indices.
#ifndef IXR_T
#define IXR_T long
#endif
#ifndef IX2_T
#define IX2_T long
#endif
#ifndef IXC_T
#define IXC_T long
#endif
typedef IXR_T ixr_t;
typedef IX2_T ix2_t;
typedef IXC_T ixc_t;
#define THREAD_IJT(T, i0, i1, j0, j1)                               \
  const int i = blockIdx.x * blockDim.x + threadIdx.x - IOFF;       \
  const int j = (j0) + blockIdx.y * blockDim.y + threadIdx.y;       \
  if (i < (i0) || i > (i1) || j > (j1)) return;                     \
  const T c = CIDX(i, j);                                           \
  const T SX = p.SX; (void)SX;
#define THREAD_IJ(i0, i1, j0, j1) THREAD_IJT(ix_t, i0, i1, j0, j1)

// Closed N/S walls: the reference extends its one-dimensional difference arrays across a wall by copying
// the nearest computed row (e.g. "wrk1(i,jstr-1)=wrk1(i,jstr)", compute_horiz_rhs_uv_terms.h:110-140,
// "FE(i,jstr-1)=FE(i,jstr)", compute_horiz_tracer_fluxes.h:150-160, "FC(i,jmin-1)=FC(i,jmin)", prsgrd.F:300).
// Kernels that recompute those differences in registers evaluate them at the clamped row instead:
// element offset that moves row j to clamp(j, lo, hi); lo/hi only bind on a wall.
__device__ __forceinline__ ix_t jclamp(const KP& p, int j, int lo_wall, int hi_wall) {
  int jc = j;
  if (p.swall && jc < lo_wall) jc = lo_wall;
  if (p.nwall && jc > hi_wall) jc = hi_wall;
  return (ix_t)(jc - j) * p.SX;
}
// the same along i on closed western/eastern walls ("wrk1(istrU-1,j)=wrk1(istrU,j)", "FX(istr-1,j)=FX(istr,j)", "FC(imin-1,j)=FC(imin,j)")
__device__ __forceinline__ ix_t iclamp(const KP& p, int i, int lo_wall, int hi_wall) {
  int ic = i;
  if (p.wwall && ic < lo_wall) ic = lo_wall;
  if (p.ewall && ic > hi_wall) ic = hi_wall;
  return (ix_t)(ic - i);
}
// a point within two cells of a closed wall: its difference stencils need the clamped forms
__device__ __forceinline__ bool near_wall(const KP& p, int i, int j) {
  return (p.swall && j <= p.jstr + 2) || (p.nwall && j >= p.jend - 2) || (p.wwall && i <= p.istr + 2) || (p.ewall && i >= p.iend - 2);
}

// ------------------------------------------------------------------------------------------------
// exchange_xxx on one subdomain (mpi_exchanges.F:34-40 with the periodic wrap of mpi_setup.F:231-250):
// every halo cell copies its periodic image from the interior.  Up to 4 fields per launch.
struct HaloList {
  double* a[4];
  int nk[4];
  int n;
};
__global__ void k_halo_periodic(const __grid_constant__ KP p, const __grid_constant__ HaloList hl, int ew, int ns) {
  const int Lm = p.Lm, Mm = p.Mm;
  const int SXh = Lm + 4;
  // halo cells enumerated as: 4 full rows (j=-1,0,Mm+1,Mm+2) then 4 columns x Mm rows
  const int nrow = 4 * SXh, ncol = 4 * Mm;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nrow + ncol) return;
  int i, j;
  if (t < nrow) {
    const int r = t / SXh;
    i = t - r * SXh - 1;
    j = (r < 2) ? (r - 1) : (Mm + r - 1);
  } else {
    const int q = t - nrow;
    const int r = q / Mm;
    j = q - r * Mm + 1;
    i = (r < 2) ? (r - 1) : (Lm + r - 1);
  }
  int is = i, js = j;
  if (ew) { if (i < 1) is = i + Lm; else if (i > Lm) is = i - Lm; }
  if (ns) { if (j < 1) js = j + Mm; else if (j > Mm) js = j - Mm; }
  const int jlo = p.swall ? 0 : 1, jhi = p.nwall ? Mm + 1 : Mm, ilo = p.wwall ? 0 : 1, ihi = p.ewall ? Lm + 1 : Lm;
  if ((is < ilo || is > ihi) || (js < jlo || js > jhi)) return;   // non-periodic edge: untouched
  if (is == i && js == j) return;
  const ix_t cd = CIDX(i, j), cs = CIDX(is, js);
  const int f = blockIdx.y;
  double* a = hl.a[f];
  for (int k = blockIdx.z; k < hl.nk[f]; k += gridDim.z) a[cd + (ix_t)k * p.SXY] = a[cs + (ix_t)k * p.SXY];
}

// ------------------------------------------------------------------------------------------------
// set_depth.F:16-186 set_depth_tile: z_w, z_r, Hz from zeta(knew).  Plane kernel over k=0:N.
__global__ void k_set_depth(const __grid_constant__ KP p, const double* __restrict__ zeta) {
  THREAD_IJ(p.istrR, p.iendR, p.jstrR, p.jendR)
  const int k = blockIdx.z;
  const int N = p.N;
  const double h = p.h[c];
  if (k == 0) { RW(p.z_w, c, 0) = -h; return; }
  const double ds = 1.0 / double(N);
  const double z = zeta[c], hinv = p.hinv[c];
  const double cff_w = p.hc * ds * double(k - N);
  const double cff_r = p.hc * ds * (double(k - N) - 0.5);
  const double zw = z + (z + h) * (cff_w + p.Cs_w[k] * h) * hinv;
  const double zr = z + (z + h) * (cff_r + p.Cs_r[k] * h) * hinv;
  double zwm;
  if (k == 1) zwm = -h;
  else {
    const double cff_wm = p.hc * ds * double(k - 1 - N);
    zwm = z + (z + h) * (cff_wm + p.Cs_w[k - 1] * h) * hinv;
  }
  RW(p.z_w, c, k) = zw;
  R3(p.z_r, c, k) = zr;
  R3(p.Hz, c, k) = zw - zwm;
}
// set_depth.F iic==0 branch: hinv, DU_avg1, DV_avg1
__global__ void k_set_depth_init(const __grid_constant__ KP p, const double* __restrict__ zeta, const double* __restrict__ ubar,
                                 const double* __restrict__ vbar) {
  THREAD_IJ(p.istrR, p.iendR, p.jstrR, p.jendR)
  p.hinv[c] = 1.0 / (p.h[c] + p.hc);
  if (i >= p.istr) p.DU_avg1[c] = 0.5 * (p.h[c - 1] + p.h[c] + zeta[c] + zeta[c - 1]) * p.dn_u[c] * (ubar[c]);
  if (j >= p.jstr) p.DV_avg1[c] = 0.5 * (p.h[c] + p.h[c - SX] + zeta[c] + zeta[c - SX]) * p.dm_v[c] * (vbar[c]);
}

// set_depth.F:201-234 set_HUV_tile
__global__ void k_set_huv(const __grid_constant__ KP p, const double* __restrict__ u, const double* __restrict__ v) {
  THREAD_IJ(p.istrR, p.iendR, p.jstrR, p.jendR)
  const int k = blockIdx.z + 1;
  const double Hz = R3(p.Hz, c, k);
  if (i >= p.istr) R3(p.FlxU, c, k) = 0.5 * (Hz + R3(p.Hz, c - 1, k)) * p.dn_u[c] * (R3(u, c, k));
  if (j >= p.jstr) R3(p.FlxV, c, k) = 0.5 * (Hz + R3(p.Hz, c - SX, k)) * p.dm_v[c] * (R3(v, c, k));
}

// set_depth.F:252-422 set_HUV1_tile.  DIR 0: u/FlxU, 1: v/FlxV.  Column kernel, two sweeps, no scratch.
template <int DIR>
__global__ void k_set_huv1(const __grid_constant__ KP p, double* __restrict__ u, int first) {
  THREAD_IJT(ixc_t, (DIR == 0 ? p.istr : p.istrR), p.iendR, (DIR == 0 ? p.jstrR : p.jstr), p.jendR)
  const ixc_t d = DIR == 0 ? 1 : SX;
  const double dn = DIR == 0 ? p.dn_u[c] : p.dm_v[c];
  const double msk = DIR == 0 ? p.umask[c] : p.vmask[c];
  double* Flx = DIR == 0 ? p.FlxU : p.FlxV;
  const int N = p.N;
  double DCN = 0.5 * (R3(p.Hz, c, N) + R3(p.Hz, c - d, N)) * dn;
  double DC0 = DCN;
  double FC0 = DCN * R3(u, c, N);
  for (int k = N - 1; k >= 1; --k) {
    const double DCk = 0.5 * (R3(p.Hz, c, k) + R3(p.Hz, c - d, k)) * dn;
    DC0 = DC0 + DCk;
    FC0 = FC0 + DCk * R3(u, c, k);
  }
  const double* a1 = DIR == 0 ? p.DU_avg1 : p.DV_avg1;
  const double* a2 = DIR == 0 ? p.DU_avg2 : p.DV_avg2;
  const double* ab = DIR == 0 ? p.DU_avg_bak : p.DV_avg_bak;
  if (first) FC0 = (FC0 - a1[c]) / DC0;
  else FC0 = (FC0 - 3.63 * a1[c] + 4.47 * a2[c] - 2.05 * ab[c]) / DC0;
  for (int k = 1; k <= N; ++k) {
    const double DCk = 0.5 * (R3(p.Hz, c, k) + R3(p.Hz, c - d, k)) * dn;
    const double un = (R3(u, c, k) - FC0) * msk;
    R3(u, c, k) = un;
    R3(Flx, c, k) = DCk * (un);
  }
}

// omega.F:17-236 omega_tile.  Column kernel: up-sweep integrates continuity, down-sweep removes the
// grid motion and splits W into explicit We and implicit Wi by the local Courant number.
#ifndef KCO
#define KCO 4
#endif
__global__ void k_omega(const __grid_constant__ KP p, double dtau) {
  extern __shared__ double smem[];
  THREAD_IJ(p.istr, p.iend, p.jstr, p.jend)
  ColScr sc = col_scratch(p, c, smem);
  const int N = p.N;
  const double cu_min = 0.6, cu_max = 1.0, cmnx_ratio = cu_min / cu_max, cutoff = 2.0 - cmnx_ratio, r4cmx = 0.25 / (1.0 - cmnx_ratio);
  double W = 0.0;
  RW(p.Wi, c, 0) = 0.0;
  const int pidx = p.pipes ? (int)p.pipe_idx[c] : 0;
  const double pflx = pidx > 0 ? p.pipe_flx[c] : 0.0;
  const double* __restrict__ gFu = p.FlxU;
  const double* __restrict__ gFv = p.FlxV;
  const double* __restrict__ gzw = p.z_w;
  const double* __restrict__ gHz = p.Hz;
  for (int k0 = 1; k0 <= N; k0 += KCO) {        // chunked: all loads of KCO levels are issued before the recurrence
    double a1[KCO], a0[KCO], b1[KCO], b0[KCO];
#pragma unroll
    for (int q = 0; q < KCO; ++q)
      if (k0 + q <= N) { a1[q] = R3(gFu, c + 1, k0 + q); a0[q] = R3(gFu, c, k0 + q); b1[q] = R3(gFv, c + SX, k0 + q); b0[q] = R3(gFv, c, k0 + q); }
#pragma unroll
    for (int q = 0; q < KCO; ++q) {
      const int k = k0 + q;
      if (k <= N) {
        const double fu1 = a1[q], fu0 = a0[q], fv1 = b1[q], fv0 = b0[q];
        W = W - fu1 + fu0 - fv1 + fv0;
        if (pidx > 0) W = W + pflx * p.pipe_tab[(pidx - 1) * PIPROW(N) + k - 1];   // omega.F:102-108
        S(0, k) = DMAX(fu1, 0.0) - DMIN(fu0, 0.0) + DMAX(fv1, 0.0) - DMIN(fv0, 0.0);
        S(1, k) = W;
      }
    }
  }
  W = W + p.swflx[c] * p.dm_r[c] * p.dn_r[c];
  const double zw0 = RW(gzw, c, 0);
  const double wrk = W / (RW(gzw, c, N) - zw0);
  RW(p.Wi, c, N) = 0.0;
  RW(p.We, c, N) = 0.0;
  RW(p.We, c, 0) = 0.0;
  const double CX0 = dtau * p.pm[c] * p.pn[c];
  double CXkp = S(0, N), Hzkp = R3(gHz, c, N);
  for (int k0 = N - 1; k0 >= 1; k0 -= KCO) {
    double sw[KCO], sx[KCO], zw[KCO], hz[KCO];
#pragma unroll
    for (int q = 0; q < KCO; ++q)
      if (k0 - q >= 1) { sw[q] = S(1, k0 - q); sx[q] = S(0, k0 - q); zw[q] = RW(gzw, c, k0 - q); hz[q] = R3(gHz, c, k0 - q); }
#pragma unroll
    for (int q = 0; q < KCO; ++q) {
      const int k = k0 - q;
      if (k >= 1) {
        double Wik = sw[q] - wrk * (zw[q] - zw0);
        const double CXk = sx[q], Hzk = hz[q];
        const double c2d = DMAX(CXk, CXkp);
        const double dh = DMIN(Hzk, Hzkp);
        const double cw_max = cu_max * dh - c2d * CX0;
        double Wek;
        if (cw_max > 0.0) {
          const double cw_max2 = cw_max * cw_max;
          const double cw_min = cw_max * cmnx_ratio;
          const double cw = fabs(Wik) * CX0;
          double cff;
          if (cw < cw_min) cff = cw_max2;
          else if (cw < cutoff * cw_max) cff = cw_max2 + r4cmx * ((cw - cw_min) * (cw - cw_min));
          else cff = cw_max * cw;
          Wek = cw_max2 * Wik / cff;
          Wik = Wik - Wek;
        } else Wek = 0.0;
        RW(p.We, c, k) = Wek;
        RW(p.Wi, c, k) = Wik;
        CXkp = CXk; Hzkp = Hzk;
      }
    }
  }
}

// rho_eos.F:26-409 rho_eos_tile over the extended range (whole padded subdomain), one down-sweep:
// rho1,qp1 per level and the vertical integrals rhoA, rhoS (rho_eos.F:364-395).
template <bool NONLIN>
__global__ void k_rho_eos(const __grid_constant__ KP p, const double* __restrict__ T, const double* __restrict__ Sal) {
  THREAD_IJ((p.wwall ? p.istr - 1 : p.istr - 2), (p.ewall ? p.iend + 1 : p.iend + 2), (p.swall ? p.jstr - 1 : p.jstr - 2), (p.nwall ? p.jend + 1 : p.jend + 2))
  const int N = p.N;
  const double rho0 = p.rho0, qp2 = 0.0000172;
  const double r00 = 999.842594, r01 = 6.793952e-2, r02 = -9.095290e-3, r03 = 1.001685e-4, r04 = -1.120083e-6, r05 = 6.536332e-9,
               r10 = 0.824493, r11 = -4.08990e-3, r12 = 7.64380e-5, r13 = -8.24670e-7, r14 = 5.38750e-9, rS0 = -5.72466e-3,
               rS1 = 1.02270e-4, rS2 = -1.65460e-6, r20 = 4.8314e-4, K00 = 19092.56, K01 = 209.8925, K02 = -3.041638,
               K03 = -1.852732e-3, K04 = -1.361629e-5, K10 = 104.4077, K11 = -6.500517, K12 = 0.1553190, K13 = 2.326469e-4,
               KS0 = -5.587545, KS1 = +0.7390729, KS2 = -1.909078e-2;
  double K0_Duk = 0, dr00 = 0, cffl = 0;
  if (NONLIN) {
    const double Tt = 3.8, Ts = 34.5, sqrtTs = sqrt(Ts);
    K0_Duk = Tt * (K01 + Tt * (K02 + Tt * (K03 + Tt * K04))) + Ts * (K10 + Tt * (K11 + Tt * (K12 + Tt * K13)) + sqrtTs * (KS0 + Tt * (KS1 + Tt * KS2)));
    dr00 = r00 - rho0;
  } else {
    cffl = p.salinity ? (p.Tcoef * p.T0 - p.Scoef * p.S0) : (p.Tcoef * p.T0);
  }
  const double rmask = p.rmask[c];
  double rhoA = 0, rhoS = 0;
  double zr_up = 0, r1_up = 0, q1_up = 0;   // level k+1 (bvf)
  for (int k = N; k >= 1; --k) {
    double r1, cff;
    const double Hz = R3(p.Hz, c, k);
    if (NONLIN) {
      const double Tt = R3(T, c, k), Ts = R3(Sal, c, k), sqrtTs = sqrt(DMAX(0.0, Ts));
      r1 = (dr00 + Tt * (r01 + Tt * (r02 + Tt * (r03 + Tt * (r04 + Tt * r05)))) +
            Ts * (r10 + Tt * (r11 + Tt * (r12 + Tt * (r13 + Tt * r14))) + sqrtTs * (rS0 + Tt * (rS1 + Tt * rS2)) + Ts * r20)) * rmask;
      const double K0 = Tt * (K01 + Tt * (K02 + Tt * (K03 + Tt * K04))) + Ts * (K10 + Tt * (K11 + Tt * (K12 + Tt * K13)) + sqrtTs * (KS0 + Tt * (KS1 + Tt * KS2)));
      const double q1 = 0.1 * (rho0 + r1) * (K0_Duk - K0) / ((K00 + K0) * (K00 + K0_Duk)) * rmask;
      R3(p.qp1, c, k) = q1;
      const double zr = R3(p.z_r, c, k);
      const double dpth = -zr;
      cff = Hz * (r1 + q1 * dpth * (1.0 - qp2 * dpth));
      if (p.lmd_kpp) {      // rho_eos.F (LMD_KPP): Brunt-Vaisala frequency between levels k and k+1
        if (k < N) {
          const double dp2 = -0.5 * (zr_up + zr);
          RW(p.bvf, c, k) = -(p.g / rho0) * (r1_up - r1 + (q1_up - q1) * dp2 * (1. - qp2 * dp2)) / (zr_up - zr) * rmask;
        }
        zr_up = zr; r1_up = r1; q1_up = q1;
      }
    } else {
      r1 = cffl - p.Tcoef * R3(T, c, k);
      if (p.salinity) r1 = r1 + p.Scoef * R3(Sal, c, k);
      r1 = r1 * rmask;
      cff = Hz * r1;
    }
    R3(p.rho1, c, k) = r1;
    if (k == N) { rhoS = 0.5 * cff * Hz; rhoA = cff; }
    else { rhoS = rhoS + Hz * (rhoA + 0.5 * cff); rhoA = rhoA + cff; }
  }
  if (NONLIN && p.lmd_kpp) { RW(p.bvf, c, N) = RW(p.bvf, c, N - 1); RW(p.bvf, c, 0) = RW(p.bvf, c, 1); }
  const double cff1 = 1.0 / rho0;
  const double cff = 1.0 / (RW(p.z_w, c, N) - RW(p.z_w, c, 0));
  p.rhoA[c] = cff * cff1 * rhoA;
  p.rhoS[c] = 2.0 * cff * cff * cff1 * rhoS;
}

// prsgrd.F:160-226: vertical part.  The harmonic means dZ,dR of the reference depend only on levels
// k-1,k,k+1, so the whole column is one downward sweep with a 3-level register window: writes the
// hydrostatic pressure P and (SPLIT_EOS) the in-situ density rho.  Columns i=0:Lm, j=0:Mm.
template <bool NONLIN>
__global__ void k_prsgrd_vert(const __grid_constant__ KP p) {
  THREAD_IJ(0, p.Lm, 0, p.Mm)
  const int N = p.N;
  const double qp2 = 0.0000172, OneFifth = 0.2, OneTwelfth = 1.0 / 12.0, epsil = 0.0;
  const double grho = p.g / p.rho0, HalfGRho = 0.5 * grho;
  // window: level k+1 ("p"), k ("c"), k-1 ("m")
  double zr_c = R3(p.z_r, c, N), r1_c = R3(p.rho1, c, N), q1_c = NONLIN ? R3(p.qp1, c, N) : 0.0;
  double zr_m = R3(p.z_r, c, N - 1), r1_m = R3(p.rho1, c, N - 1), q1_m = NONLIN ? R3(p.qp1, c, N - 1) : 0.0;
  double zr_p = 0, rho_p = 0, dZh_p = 0, dRh_p = 0, P_p = 0;
  // raw differences: rawZ_c = dZ(k) (between k and k+1), rawZ_m = dZ(k-1)
  double rawZ_c, rawR_c, rawZ_m, rawR_m;
  {
    rawZ_m = zr_c - zr_m;
    if (NONLIN) { const double dpth = -0.5 * (zr_c + zr_m); rawR_m = r1_c - r1_m + (q1_c - q1_m) * dpth * (1.0 - qp2 * dpth); }
    else rawR_m = r1_c - r1_m;
    rawZ_c = rawZ_m; rawR_c = rawR_m;   // dZ(N)=dZ(N-1), dR(N)=dR(N-1)
  }
  for (int k = N; k >= 1; --k) {
    // harmonic averages at level k from raw(k), raw(k-1)
    double cff = 2.0 * rawZ_c * rawZ_m;
    const double dZh = cff / (rawZ_c + rawZ_m);
    const double cfr = 2.0 * rawR_c * rawR_m;
    double dRh = (cfr > epsil) ? cfr / (rawR_c + rawR_m) : 0.0;
    double rho_c;
    if (NONLIN) {
      const double dpth = -zr_c;
      dRh = dRh - q1_c * dZh * (1.0 - 2.0 * qp2 * dpth);
      rho_c = r1_c + q1_c * dpth * (1.0 - qp2 * dpth);
      R3(p.rho, c, k) = rho_c;
    } else rho_c = r1_c;
    double Pk;
    if (k == N) {
      double rho_m;
      if (NONLIN) { const double dpm = -zr_m; rho_m = r1_m + q1_m * dpm * (1.0 - qp2 * dpm); }
      else rho_m = r1_m;
      const double zwN = RW(p.z_w, c, N);
      Pk = p.g * zwN + grho * (rho_c + 0.5 * (rho_c - rho_m) * (zwN - zr_c) / (zr_c - zr_m)) * (zwN - zr_c);
      if (p.pot_tides) Pk = Pk - p.g * p.ptide[c];   // prsgrd.F:209-211
    } else {
      Pk = P_p + HalfGRho * ((rho_p + rho_c) * (zr_p - zr_c) -
                             OneFifth * ((dRh_p - dRh) * (zr_p - zr_c - OneTwelfth * (dZh_p + dZh)) - (dZh_p - dZh) * (rho_p - rho_c - OneTwelfth * (dRh_p + dRh))));
    }
    R3(p.P, c, k) = Pk;
    // slide the window down
    zr_p = zr_c; rho_p = rho_c; dZh_p = dZh; dRh_p = dRh; P_p = Pk;
    rawZ_c = rawZ_m; rawR_c = rawR_m;
    zr_c = zr_m; r1_c = r1_m; q1_c = q1_m;
    if (k - 2 >= 1) {
      zr_m = R3(p.z_r, c, k - 2); r1_m = R3(p.rho1, c, k - 2); if (NONLIN) q1_m = R3(p.qp1, c, k - 2);
      rawZ_m = zr_c - zr_m;
      if (NONLIN) { const double dpth = -0.5 * (zr_c + zr_m); rawR_m = r1_c - r1_m + (q1_c - q1_m) * dpth * (1.0 - qp2 * dpth); }
      else rawR_m = r1_c - r1_m;
    }  // else dZ(0)=dZ(1): rawZ_m keeps the value just shifted into rawZ_c
  }
}

// rho_eos_tile FUSED with the vertical part of prsgrd_tile (used when prsgrd follows a rho_eos of the same time level, as in
// main.F:397-413 and :430-445): both are downward sweeps of the same column, so the equation of state feeds the 3-level
// window of the pressure integration directly and rho1, qp1, z_r are not read back (8 instead of 11 words per point for
// the pair).  Every value is produced by the expressions of k_rho_eos / k_prsgrd_vert in the same order (the EOS levels
// are visited N, N-1, ..., 1, so the running integrals rhoA, rhoS round identically).  Runs over the extended range of
// rho_eos; columns outside prsgrd's range only do the EOS part.
template <bool NONLIN>
__global__ void k_rho_prs_vert(const __grid_constant__ KP p, const double* __restrict__ T, const double* __restrict__ Sal) {
  THREAD_IJ((p.wwall ? p.istr - 1 : p.istr - 2), (p.ewall ? p.iend + 1 : p.iend + 2), (p.swall ? p.jstr - 1 : p.jstr - 2), (p.nwall ? p.jend + 1 : p.jend + 2))
  const bool doP = i >= 0 && i <= p.Lm && j >= 0 && j <= p.Mm;
  const int N = p.N;
  const double rho0 = p.rho0, qp2 = 0.0000172;
  const double r00 = 999.842594, r01 = 6.793952e-2, r02 = -9.095290e-3, r03 = 1.001685e-4, r04 = -1.120083e-6, r05 = 6.536332e-9,
               r10 = 0.824493, r11 = -4.08990e-3, r12 = 7.64380e-5, r13 = -8.24670e-7, r14 = 5.38750e-9, rS0 = -5.72466e-3,
               rS1 = 1.02270e-4, rS2 = -1.65460e-6, r20 = 4.8314e-4, K00 = 19092.56, K01 = 209.8925, K02 = -3.041638,
               K03 = -1.852732e-3, K04 = -1.361629e-5, K10 = 104.4077, K11 = -6.500517, K12 = 0.1553190, K13 = 2.326469e-4,
               KS0 = -5.587545, KS1 = +0.7390729, KS2 = -1.909078e-2;
  double K0_Duk = 0, dr00 = 0, cffl = 0;
  if (NONLIN) {
    const double Tt = 3.8, Ts = 34.5, sqrtTs = sqrt(Ts);
    K0_Duk = Tt * (K01 + Tt * (K02 + Tt * (K03 + Tt * K04))) + Ts * (K10 + Tt * (K11 + Tt * (K12 + Tt * K13)) + sqrtTs * (KS0 + Tt * (KS1 + Tt * KS2)));
    dr00 = r00 - rho0;
  } else {
    cffl = p.salinity ? (p.Tcoef * p.T0 - p.Scoef * p.S0) : (p.Tcoef * p.T0);
  }
  const double rmask = p.rmask[c];
  double rhoA = 0, rhoS = 0;
  double zr_up = 0, r1_up = 0, q1_up = 0;   // level k+1 (bvf)
  // equation of state at level k (k_rho_eos body): stores rho1, qp1, advances the integrals, returns r1, q1, z_r
  auto eos = [&](int k, double& r1, double& q1, double& zr) {
    double cff;
    const double Hz = R3(p.Hz, c, k);
    zr = R3(p.z_r, c, k);
    if (NONLIN) {
      const double Tt = R3(T, c, k), Ts = R3(Sal, c, k), sqrtTs = sqrt(DMAX(0.0, Ts));
      r1 = (dr00 + Tt * (r01 + Tt * (r02 + Tt * (r03 + Tt * (r04 + Tt * r05)))) +
            Ts * (r10 + Tt * (r11 + Tt * (r12 + Tt * (r13 + Tt * r14))) + sqrtTs * (rS0 + Tt * (rS1 + Tt * rS2)) + Ts * r20)) * rmask;
      const double K0 = Tt * (K01 + Tt * (K02 + Tt * (K03 + Tt * K04))) + Ts * (K10 + Tt * (K11 + Tt * (K12 + Tt * K13)) + sqrtTs * (KS0 + Tt * (KS1 + Tt * KS2)));
      q1 = 0.1 * (rho0 + r1) * (K0_Duk - K0) / ((K00 + K0) * (K00 + K0_Duk)) * rmask;
      R3(p.qp1, c, k) = q1;
      const double dpth = -zr;
      cff = Hz * (r1 + q1 * dpth * (1.0 - qp2 * dpth));
      if (p.lmd_kpp) {      // rho_eos.F (LMD_KPP): Brunt-Vaisala frequency between levels k and k+1
        if (k < N) {
          const double dp2 = -0.5 * (zr_up + zr);
          RW(p.bvf, c, k) = -(p.g / rho0) * (r1_up - r1 + (q1_up - q1) * dp2 * (1. - qp2 * dp2)) / (zr_up - zr) * rmask;
        }
        zr_up = zr; r1_up = r1; q1_up = q1;
      }
    } else {
      r1 = cffl - p.Tcoef * R3(T, c, k);
      if (p.salinity) r1 = r1 + p.Scoef * R3(Sal, c, k);
      r1 = r1 * rmask;
      q1 = 0.0;
      cff = Hz * r1;
    }
    R3(p.rho1, c, k) = r1;
    if (k == N) { rhoS = 0.5 * cff * Hz; rhoA = cff; }
    else { rhoS = rhoS + Hz * (rhoA + 0.5 * cff); rhoA = rhoA + cff; }
  };
  const double OneFifth = 0.2, OneTwelfth = 1.0 / 12.0, epsil = 0.0;
  const double grho = p.g / p.rho0, HalfGRho = 0.5 * grho;
  double zr_c, r1_c, q1_c, zr_m, r1_m, q1_m;
  eos(N, r1_c, q1_c, zr_c);
  eos(N - 1, r1_m, q1_m, zr_m);
  double zr_p = 0, rho_p = 0, dZh_p = 0, dRh_p = 0, P_p = 0;
  double rawZ_c, rawR_c, rawZ_m, rawR_m;
  {
    rawZ_m = zr_c - zr_m;
    if (NONLIN) { const double dpth = -0.5 * (zr_c + zr_m); rawR_m = r1_c - r1_m + (q1_c - q1_m) * dpth * (1.0 - qp2 * dpth); }
    else rawR_m = r1_c - r1_m;
    rawZ_c = rawZ_m; rawR_c = rawR_m;   // dZ(N)=dZ(N-1), dR(N)=dR(N-1)
  }
  for (int k = N; k >= 1; --k) {
    if (doP) {
      double cff = 2.0 * rawZ_c * rawZ_m;
      const double dZh = cff / (rawZ_c + rawZ_m);
      const double cfr = 2.0 * rawR_c * rawR_m;
      double dRh = (cfr > epsil) ? cfr / (rawR_c + rawR_m) : 0.0;
      double rho_c;
      if (NONLIN) {
        const double dpth = -zr_c;
        dRh = dRh - q1_c * dZh * (1.0 - 2.0 * qp2 * dpth);
        rho_c = r1_c + q1_c * dpth * (1.0 - qp2 * dpth);
        R3(p.rho, c, k) = rho_c;
      } else rho_c = r1_c;
      double Pk;
      if (k == N) {
        double rho_m;
        if (NONLIN) { const double dpm = -zr_m; rho_m = r1_m + q1_m * dpm * (1.0 - qp2 * dpm); }
        else rho_m = r1_m;
        const double zwN = RW(p.z_w, c, N);
        Pk = p.g * zwN + grho * (rho_c + 0.5 * (rho_c - rho_m) * (zwN - zr_c) / (zr_c - zr_m)) * (zwN - zr_c);
        if (p.pot_tides) Pk = Pk - p.g * p.ptide[c];   // prsgrd.F:209-211
      } else {
        Pk = P_p + HalfGRho * ((rho_p + rho_c) * (zr_p - zr_c) -
                               OneFifth * ((dRh_p - dRh) * (zr_p - zr_c - OneTwelfth * (dZh_p + dZh)) - (dZh_p - dZh) * (rho_p - rho_c - OneTwelfth * (dRh_p + dRh))));
      }
      R3(p.P, c, k) = Pk;
      zr_p = zr_c; rho_p = rho_c; dZh_p = dZh; dRh_p = dRh; P_p = Pk;
    }
    rawZ_c = rawZ_m; rawR_c = rawR_m;
    zr_c = zr_m; r1_c = r1_m; q1_c = q1_m;
    if (k - 2 >= 1) {
      eos(k - 2, r1_m, q1_m, zr_m);
      rawZ_m = zr_c - zr_m;
      if (NONLIN) { const double dpth = -0.5 * (zr_c + zr_m); rawR_m = r1_c - r1_m + (q1_c - q1_m) * dpth * (1.0 - qp2 * dpth); }
      else rawR_m = r1_c - r1_m;
    }
  }
  if (NONLIN && p.lmd_kpp) { RW(p.bvf, c, N) = RW(p.bvf, c, N - 1); RW(p.bvf, c, 0) = RW(p.bvf, c, 1); }
  const double cff1 = 1.0 / rho0;
  const double cff = 1.0 / (RW(p.z_w, c, N) - RW(p.z_w, c, 0));
  p.rhoA[c] = cff * cff1 * rhoA;
  p.rhoS[c] = 2.0 * cff * cff * cff1 * rhoS;
}

// prsgrd.F:231-421: horizontal part, plane kernel.  ru, rv at (i,j,k) from the density-Jacobian with
// harmonic-mean reconstruction of the horizontal differences.
template <bool NONLIN>
__device__ __forceinline__ void prs_fc_rx(const KP& p, ix_t X, ix_t d, int k, const double* __restrict__ msk, const double* __restrict__ rho,
                                          double& FC, double& rx) {
  const double qp2 = 0.0000172;
  const double zu = R3(p.z_r, X, k), zl = R3(p.z_r, X - d, k), m = msk[X];
  FC = (zu - zl) * m;
  if (NONLIN) {
    const double dpth = -0.5 * (zu + zl);
    rx = (R3(p.rho1, X, k) - R3(p.rho1, X - d, k) + (R3(p.qp1, X, k) - R3(p.qp1, X - d, k)) * dpth * (1.0 - qp2 * dpth)) * m;
  } else rx = (R3(rho, X, k) - R3(rho, X - d, k)) * m;
}
// second half of prsgrd.F:330-421 for one direction: harmonic means of the three first differences -> ru / rv
// prs_tail_core: the arithmetic, operands in registers (shared by the global-memory and the TMA-staged kernels)
template <bool NONLIN>
__device__ __forceinline__ double prs_tail_core(const KP& p, double dn, double FCa, double FCb, double FCc, double rxa, double rxb, double rxc, double q1, double q2v,
                                                double z1, double z2, double rho_c, double rho_m, double Hz_c, double Hz_m, double P_m, double P_c) {
  const double qp2 = 0.0000172, OneFifth = 0.2, OneTwelfth = 1.0 / 12.0, epsil = 0.0;
  const double HalfGRho = 0.5 * (p.g / p.rho0);
  // dZx, dRx at c-d ("m") and c ("c"): harmonic means (prsgrd.F:330-345).  The quotient is formed unconditionally and
  // selected afterwards (same value where it is used): no divergent branch around the fp64 division.
  double cff = 2.0 * FCa * FCb, quo = cff / (FCa + FCb);
  const double dZx_m = (cff > epsil) ? quo : 0.0;
  cff = 2.0 * FCb * FCc; quo = cff / (FCb + FCc);
  const double dZx_c = (cff > epsil) ? quo : 0.0;
  cff = 2.0 * rxa * rxb; quo = cff / (rxa + rxb);
  double dRx_m = (cff > epsil) ? quo : 0.0;
  cff = 2.0 * rxb * rxc; quo = cff / (rxb + rxc);
  double dRx_c = (cff > epsil) ? quo : 0.0;
  if (NONLIN) {
    dRx_m = dRx_m - q1 * dZx_m * (1.0 + 2.0 * qp2 * z1);
    dRx_c = dRx_c - q2v * dZx_c * (1.0 + 2.0 * qp2 * z2);
  }
  return 0.5 * (Hz_c + Hz_m) * dn *
         (P_m - P_c -
          HalfGRho * ((rho_c + rho_m) * (z2 - z1) -
                      OneFifth * ((dRx_c - dRx_m) * (z2 - z1 - OneTwelfth * (dZx_c + dZx_m)) - (dZx_c - dZx_m) * (rho_c - rho_m - OneTwelfth * (dRx_c + dRx_m)))));
}
template <bool NONLIN>
__device__ __forceinline__ double prs_tail(const KP& p, ix_t c, ix_t d, int k, double dn, const double* __restrict__ rho, double FCa, double FCb, double FCc,
                                           double rxa, double rxb, double rxc, double q1, double q2v, double z1, double z2) {
  return prs_tail_core<NONLIN>(p, dn, FCa, FCb, FCc, rxa, rxb, rxc, q1, q2v, z1, z2, R3(rho, c, k), R3(rho, c - d, k), R3(p.Hz, c, k), R3(p.Hz, c - d, k),
                               R3(p.P, c - d, k), R3(p.P, c, k));
}
// rows next to a closed wall: the three difference positions (c-d, c, c+d) are displaced by oA, oB, oC (see jclamp)
template <bool NONLIN>
__device__ __forceinline__ double prs_dir_cl(const KP& p, ix_t c, ix_t d, int k, const double* __restrict__ msk, double dn, const double* __restrict__ rho,
                                             ix_t oA, ix_t oB, ix_t oC) {
  const double z1 = R3(p.z_r, c - d, k), z2 = R3(p.z_r, c, k);
  double FCa, FCb, FCc, rxa, rxb, rxc, q1 = 0, q2v = 0;
  prs_fc_rx<NONLIN>(p, c - d + oA, d, k, msk, rho, FCa, rxa);
  prs_fc_rx<NONLIN>(p, c + oB, d, k, msk, rho, FCb, rxb);
  prs_fc_rx<NONLIN>(p, c + d + oC, d, k, msk, rho, FCc, rxc);
  if (NONLIN) { q1 = R3(p.qp1, c - d, k); q2v = R3(p.qp1, c, k); }
  return prs_tail<NONLIN>(p, c, d, k, dn, rho, FCa, FCb, FCc, rxa, rxb, rxc, q1, q2v, z1, z2);
}
// interior: points c-2d, c-d, c, c+d; the masks m1..m3 = msk(c-d), msk(c), msk(c+d) are level-independent and come in registers
template <bool NONLIN>
__device__ __forceinline__ double prs_dir(const KP& p, ix_t c, ix_t d, int k, double m1, double m2, double m3, double dn, const double* __restrict__ rho) {
  const double qp2 = 0.0000172;
  const double z0 = R3(p.z_r, c - 2 * d, k), z1 = R3(p.z_r, c - d, k), z2 = R3(p.z_r, c, k), z3 = R3(p.z_r, c + d, k);
  const double FCa = (z1 - z0) * m1, FCb = (z2 - z1) * m2, FCc = (z3 - z2) * m3;   // FC(i-1), FC(i), FC(i+1)
  double rxa, rxb, rxc, q1 = 0, q2v = 0;
  if (NONLIN) {
    const double a0 = R3(p.rho1, c - 2 * d, k), a1 = R3(p.rho1, c - d, k), a2 = R3(p.rho1, c, k), a3 = R3(p.rho1, c + d, k);
    const double b0 = R3(p.qp1, c - 2 * d, k), b1 = R3(p.qp1, c - d, k), b2 = R3(p.qp1, c, k), b3 = R3(p.qp1, c + d, k);
    double dpth = -0.5 * (z1 + z0); rxa = (a1 - a0 + (b1 - b0) * dpth * (1.0 - qp2 * dpth)) * m1;
    dpth = -0.5 * (z2 + z1); rxb = (a2 - a1 + (b2 - b1) * dpth * (1.0 - qp2 * dpth)) * m2;
    dpth = -0.5 * (z3 + z2); rxc = (a3 - a2 + (b3 - b2) * dpth * (1.0 - qp2 * dpth)) * m3;
    q1 = b1; q2v = b2;
  } else {
    const double a0 = R3(rho, c - 2 * d, k), a1 = R3(rho, c - d, k), a2 = R3(rho, c, k), a3 = R3(rho, c + d, k);
    rxa = (a1 - a0) * m1; rxb = (a2 - a1) * m2; rxc = (a3 - a2) * m3;
  }
  return prs_tail<NONLIN>(p, c, d, k, dn, rho, FCa, FCb, FCc, rxa, rxb, rxc, q1, q2v, z1, z2);
}
// Each thread computes KB consecutive levels of one (i,j): the six masks and two metric factors of the point stay in
// registers across the levels (as 2-D operands re-fetched per level they cost up to 3.6 words/point of HBM traffic).
template <bool NONLIN, bool WALL, int KB>
__global__ void __launch_bounds__(256, LB_PRS) k_prsgrd_horiz(const __grid_constant__ KP p) {
  THREAD_IJ(p.istr, p.iend, p.jstr, p.jend)
  const int k0 = blockIdx.z * KB + 1;
  const double* __restrict__ rho = NONLIN ? p.rho : p.rho1;
  const double dnu = p.dn_u[c], dmv = p.dm_v[c];
  if (WALL && near_wall(p, i, j)) {   // prsgrd.F:288-310: FC,rx exist on rows jstrV..jend / columns istrU..iend between walls, copied outward
    const ix_t oA = jclamp(p, j - 1, p.jstrV, p.jend), oB = jclamp(p, j, p.jstrV, p.jend), oC = jclamp(p, j + 1, p.jstrV, p.jend);
    const ix_t iA = iclamp(p, i - 1, p.istrU, p.iend), iB = iclamp(p, i, p.istrU, p.iend), iC = iclamp(p, i + 1, p.istrU, p.iend);
#pragma unroll 1
    for (int k = k0; k < k0 + KB && k <= p.N; ++k) {
      if (i >= p.istrU) R3(p.ru, c, k) = prs_dir_cl<NONLIN>(p, c, 1, k, p.umask, dnu, rho, iA, iB, iC);
      if (j >= p.jstrV) R3(p.rv, c, k) = prs_dir_cl<NONLIN>(p, c, SX, k, p.vmask, dmv, rho, oA, oB, oC);
    }
    return;
  }
  const double mu1 = p.umask[c - 1], mu2 = p.umask[c], mu3 = p.umask[c + 1];
  const double mv1 = p.vmask[c - SX], mv2 = p.vmask[c], mv3 = p.vmask[c + SX];
#pragma unroll
  for (int q = 0; q < KB; ++q) {
    const int k = k0 + q;
    if (k <= p.N) {
      const double ru = prs_dir<NONLIN>(p, c, 1, k, mu1, mu2, mu3, dnu, rho);
      const double rv = prs_dir<NONLIN>(p, c, SX, k, mv1, mv2, mv3, dmv, rho);
      R3(p.ru, c, k) = ru;
      R3(p.rv, c, k) = rv;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// compute_horiz_rhs_uv_terms.h:1-291.  Fluxes evaluated in registers at the flux point `c`.
// UP=false: 4th-order centred (pre_step3d4S.F:178); UP=true: 3rd-order upstream (step3d_uv1.F:124).
#define HDELTA 0.1666666666666667
#define HGAMMA 0.3333333333333333
// along-axis flux of the along-axis velocity: UFx (d=1,u,FlxU) / VFe (d=SX,v,FlxV) at rho point c
template <bool UP, bool CL>
__device__ __forceinline__ double flux_along(const KP& p, const double* __restrict__ u, const double* __restrict__ Flx, ixr_t c, ixr_t d, int k,
                                             ixr_t oa = 0, ixr_t ob = 0) {
  const double u0 = R3(u, c, k), u1 = R3(u, c + d, k);
  const double f0 = R3(Flx, c, k), f1 = R3(Flx, c + d, k);
  double w1a, w1b, w2a, w2b;     // wrk1(c), wrk1(c+d), wrk2(c), wrk2(c+d)
  if (CL) {                      // second differences taken at the wall-clamped rows c+oa, c+d+ob
    const ixr_t ca = c + oa, cb = c + d + ob;
    w1a = R3(u, ca - d, k) - 2.0 * R3(u, ca, k) + R3(u, ca + d, k);
    w1b = R3(u, cb - d, k) - 2.0 * R3(u, cb, k) + R3(u, cb + d, k);
    w2a = R3(Flx, ca - d, k) - 2.0 * R3(Flx, ca, k) + R3(Flx, ca + d, k);
    w2b = R3(Flx, cb - d, k) - 2.0 * R3(Flx, cb, k) + R3(Flx, cb + d, k);
  } else {
    const double um = R3(u, c - d, k), u2 = R3(u, c + 2 * d, k);
    const double fm = R3(Flx, c - d, k), f2 = R3(Flx, c + 2 * d, k);
    w1a = um - 2.0 * u0 + u1; w1b = u0 - 2.0 * u1 + u2;
    w2a = fm - 2.0 * f0 + f1; w2b = f0 - 2.0 * f1 + f2;
  }
  if (UP) {
    const double cff = f0 + f1 - HDELTA * (w2a + w2b);
    return 0.25 * (cff * (u0 + u1) - HGAMMA * (DMAX(cff, 0.0) * w1a + DMIN(cff, 0.0) * w1b));
  } else
    return 0.25 * (u0 + u1 - HDELTA * (w1a + w1b)) * (f0 + f1 - HDELTA * (w2a + w2b));
}
// cross-axis flux at psi point c: UFe (u advected by FlxV: du=SX, df=1) / VFx (v advected by FlxU: du=1, df=SX)
template <bool UP, bool CL>
__device__ __forceinline__ double flux_cross(const KP& p, const double* __restrict__ u, const double* __restrict__ Flx, ixr_t c, ixr_t du, ixr_t df, int k,
                                             ixr_t om = 0, ixr_t oc = 0) {
  // u second differences along du at c and c-du; Flx second differences along df at c and c-df
  const double ub = R3(u, c - du, k), uc = R3(u, c, k);
  double w1m, w1c;     // wrk1(c-du), wrk1(c)
  if (CL) {
    const ixr_t cm = c - du + om, cc = c + oc;
    w1m = R3(u, cm - du, k) - 2.0 * R3(u, cm, k) + R3(u, cm + du, k);
    w1c = R3(u, cc - du, k) - 2.0 * R3(u, cc, k) + R3(u, cc + du, k);
  } else {
    const double ua = R3(u, c - 2 * du, k), ud = R3(u, c + du, k);
    w1m = ua - 2.0 * ub + uc; w1c = ub - 2.0 * uc + ud;
  }
  const double fa = R3(Flx, c - 2 * df, k), fb = R3(Flx, c - df, k), fc = R3(Flx, c, k), fd = R3(Flx, c + df, k);
  const double w2m = fa - 2.0 * fb + fc, w2c = fb - 2.0 * fc + fd;     // wrk2(c-df), wrk2(c)
  if (UP) {
    const double cff = fc + fb - HDELTA * (w2c + w2m);
    return 0.25 * (cff * (uc + ub) - HGAMMA * (DMAX(cff, 0.0) * w1m + DMIN(cff, 0.0) * w1c));
  } else
    return 0.25 * (uc + ub - HDELTA * (w1c + w1m)) * (fc + fb - HDELTA * (w2c + w2m));
}
// one (i,j,k) point of the horizontal momentum RHS; NEAR: a row within two rows of a closed wall (clamped differences)
template <bool UP, bool NEAR>
__device__ __forceinline__ void rhs_uv_point(const KP& p, const double* __restrict__ u, const double* __restrict__ v, ixr_t c, int i, int j, int k, bool dov,
                                             double& ru, double& rv) {
  const ixr_t SX = p.SX;
  // Coriolis (no CURVGRID): compute_horiz_rhs_uv_terms.h:4-45
  const double cf0 = 0.5 * R3(p.Hz, c, k) * (p.fomn[c]);
  const double cfw = 0.5 * R3(p.Hz, c - 1, k) * (p.fomn[c - 1]);
  const double v00 = R3(v, c, k), u00 = R3(u, c, k);
  const double UFx_c = cf0 * (v00 + R3(v, c + SX, k));
  const double UFx_w = cfw * (R3(v, c - 1, k) + R3(v, c - 1 + SX, k));
  ru = R3(p.ru, c, k) + 0.5 * (UFx_c + UFx_w);
  // advection of u
  double UFx_i, UFx_im;
  if (NEAR) {   // u, FlxU second differences in i exist on columns istrU..iend between walls (compute_horiz_rhs_uv_terms.h:60-83)
    UFx_i = flux_along<UP, true>(p, u, p.FlxU, c, 1, k, iclamp(p, i, p.istrU, p.iend), iclamp(p, i + 1, p.istrU, p.iend));
    UFx_im = flux_along<UP, true>(p, u, p.FlxU, c - 1, 1, k, iclamp(p, i - 1, p.istrU, p.iend), iclamp(p, i, p.istrU, p.iend));
  } else {
    UFx_i = flux_along<UP, false>(p, u, p.FlxU, c, 1, k);
    UFx_im = flux_along<UP, false>(p, u, p.FlxU, c - 1, 1, k);
  }
  double UFe_jp, UFe_j;
  if (NEAR) {   // u second differences in j exist on rows jstr..jend between walls (compute_horiz_rhs_uv_terms.h:186-199)
    UFe_jp = flux_cross<UP, true>(p, u, p.FlxV, c + SX, SX, 1, k, jclamp(p, j, p.jstr, p.jend), jclamp(p, j + 1, p.jstr, p.jend));
    UFe_j = flux_cross<UP, true>(p, u, p.FlxV, c, SX, 1, k, jclamp(p, j - 1, p.jstr, p.jend), jclamp(p, j, p.jstr, p.jend));
  } else {
    UFe_jp = flux_cross<UP, false>(p, u, p.FlxV, c + SX, SX, 1, k);
    UFe_j = flux_cross<UP, false>(p, u, p.FlxV, c, SX, 1, k);
  }
  ru = ru - UFx_i + UFx_im - UFe_jp + UFe_j;
  if (dov) {
    const double cfs = 0.5 * R3(p.Hz, c - SX, k) * (p.fomn[c - SX]);
    const double VFe_c = cf0 * (u00 + R3(u, c + 1, k));
    const double VFe_s = cfs * (R3(u, c - SX, k) + R3(u, c - SX + 1, k));
    rv = R3(p.rv, c, k) - 0.5 * (VFe_c + VFe_s);
    double VFx_ip, VFx_i;
    if (NEAR) {   // v second differences in i exist on columns istr..iend between walls (compute_horiz_rhs_uv_terms.h:225-250)
      VFx_ip = flux_cross<UP, true>(p, v, p.FlxU, c + 1, 1, SX, k, iclamp(p, i, p.istr, p.iend), iclamp(p, i + 1, p.istr, p.iend));
      VFx_i = flux_cross<UP, true>(p, v, p.FlxU, c, 1, SX, k, iclamp(p, i - 1, p.istr, p.iend), iclamp(p, i, p.istr, p.iend));
    } else {
      VFx_ip = flux_cross<UP, false>(p, v, p.FlxU, c + 1, 1, SX, k);
      VFx_i = flux_cross<UP, false>(p, v, p.FlxU, c, 1, SX, k);
    }
    double VFe_j, VFe_jm;
    if (NEAR) {   // v, FlxV second differences exist on rows jstrV..jend between walls (compute_horiz_rhs_uv_terms.h:128-141)
      VFe_j = flux_along<UP, true>(p, v, p.FlxV, c, SX, k, jclamp(p, j, p.jstrV, p.jend), jclamp(p, j + 1, p.jstrV, p.jend));
      VFe_jm = flux_along<UP, true>(p, v, p.FlxV, c - SX, SX, k, jclamp(p, j - 1, p.jstrV, p.jend), jclamp(p, j, p.jstrV, p.jend));
    } else {
      VFe_j = flux_along<UP, false>(p, v, p.FlxV, c, SX, k);
      VFe_jm = flux_along<UP, false>(p, v, p.FlxV, c - SX, SX, k);
    }
    rv = rv - VFx_ip + VFx_i - VFe_j + VFe_jm;
  }
}
// Each thread computes JB consecutive rows: the neighbour loads and second differences shared by the rows are
// evaluated once (results are stored only after the last row so that nothing forces a reload).
// Measured on B200 (1024x1024x64): C4 (UP=false) 1.86 ms at JB=1 -> 1.59 ms at JB=4 with 128 registers; UP3 is best at JB=1.
template <bool UP, bool WALL, int JB>
__global__ void __launch_bounds__(256, (JB >= 4 ? 2 : LB_RHS)) k_rhs_uv_horiz(const __grid_constant__ KP p, const double* __restrict__ u, const double* __restrict__ v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x - IOFF;
  const int jb = p.jstr + (blockIdx.y * blockDim.y + threadIdx.y) * JB;
  if (i < p.istr || i > p.iend || jb > p.jend) return;
  const int k = blockIdx.z + 1;
  const bool near = WALL && ((p.swall && jb <= p.jstr + 2) || (p.nwall && jb + JB - 1 >= p.jend - 2) || (p.wwall && i <= p.istr + 2) || (p.ewall && i >= p.iend - 2));
  const ixr_t c0 = CIDX(i, jb), SX = p.SX;
  if (near || jb + JB - 1 > p.jend) {   // points next to a wall / ragged last block: one row at a time
#pragma unroll 1
    for (int q = 0; q < JB && jb + q <= p.jend; ++q) {
      const int j = jb + q;
      const ixr_t c = c0 + q * SX;
      const bool dov = !WALL || j >= p.jstrV, dou = !WALL || i >= p.istrU;
      double ru, rv;
      if (WALL) rhs_uv_point<UP, true>(p, u, v, c, i, j, k, dov, ru, rv);
      else rhs_uv_point<UP, false>(p, u, v, c, i, j, k, dov, ru, rv);
      if (dou) R3(p.ru, c, k) = ru;
      if (dov) R3(p.rv, c, k) = rv;
    }
    return;
  }
  double ru[JB], rv[JB];
#pragma unroll
  for (int q = 0; q < JB; ++q) rhs_uv_point<UP, false>(p, u, v, c0 + q * SX, i, jb + q, k, true, ru[q], rv[q]);
#pragma unroll
  for (int q = 0; q < JB; ++q) {
    R3(p.ru, c0 + q * SX, k) = ru[q];
    R3(p.rv, c0 + q * SX, k) = rv[q];
  }
}

// compute_horiz_tracer_fluxes.h:1-246: FX at u-point c (d=1, FlxU, umask) or FE at v-point c (d=SX)
template <bool UP>
__device__ __forceinline__ double tflux_fin(double t1, double t2, double Fa, double Fb, double Fc, double fl) {
  if (UP) {
    const double wm = Fb - Fa, wc = Fc - Fb;   // wrk1(c-d), wrk1(c)
    return 0.5 * (t2 + t1) * fl - 0.1666666666666666 * (wm * DMAX(fl, 0.0) + wc * DMIN(fl, 0.0));
  } else {
    const double wm = 0.5 * (Fb + Fa), wc = 0.5 * (Fc + Fb);
    return 0.5 * (t2 + t1 - 0.3333333333333333 * (wc - wm)) * fl;
  }
}
// rows next to a closed wall: first differences taken at the wall-clamped rows (offsets oA, oB, oC, see jclamp); fl = Flx(c,k)
template <bool UP>
__device__ __forceinline__ double tflux_cl(const KP& p, const double* __restrict__ t, double fl, const double* __restrict__ msk, ix_t c, ix_t d,
                                           int k, ix_t oA, ix_t oB, ix_t oC) {
  const double t1 = R3(t, c - d, k), t2 = R3(t, c, k);
  const ix_t XA = c - d + oA, XB = c + oB, XC = c + d + oC;
  const double Fa = (R3(t, XA, k) - R3(t, XA - d, k)) * msk[XA];
  const double Fb = (R3(t, XB, k) - R3(t, XB - d, k)) * msk[XB];
  const double Fc = (R3(t, XC, k) - R3(t, XC - d, k)) * msk[XC];
  return tflux_fin<UP>(t1, t2, Fa, Fb, Fc, fl);
}

// pre_step3d4S.F:138-147: FlxDiv such that Hz_bak = Hz + FlxDiv, Hz_fwd = Hz - FlxDiv
__global__ void k_pre_flxdiv(const __grid_constant__ KP p, double dtau) {
  THREAD_IJ(p.istrU - 1, p.iend, p.jstrV - 1, p.jend)
  const int k = blockIdx.z + 1;
  const double cff = 0.5 * dtau;
  R3(p.FlxDiv, c, k) = cff * p.pm[c] * p.pn[c] *
                       (R3(p.FlxU, c + 1, k) - R3(p.FlxU, c, k) + R3(p.FlxV, c + SX, k) - R3(p.FlxV, c, k) + RW(p.We, c, k) + RW(p.Wi, c, k) -
                        RW(p.We, c, k - 1) - RW(p.Wi, c, k - 1));
}
// pre_step3d4S.F:152-168: tracer predictor, horizontal part (C4), and t(indx)=Hz*t(nstp)
// step3d_t_ISO.F:196-213: tracer corrector, horizontal part (UP3)
// Horizontal flux divergence  FX(i+1)-FX(i)+FE(j+1)-FE(j)  at (i,j,k).  Between walls the first differences in j exist on
// rows jstr..jend+1 only (compute_horiz_tracer_fluxes.h:118-160); NEAR rows take the clamped form.  The eight mask values
// of the point are level-independent and are passed in registers; the four volume fluxes of the cell's faces are loaded
// ONCE per level and shared by every tracer of the batch (they, the masks and the metrics do not depend on the tracer).
struct TMask { double u0, u1, u2, u3, v0, v1, v2, v3; };   // umask(c-1..c+2), vmask(c-SX..c+2SX)
__device__ __forceinline__ TMask tmask_load(const KP& p, ix_t c) {
  const ix_t SX = p.SX;
  TMask m;
  m.u0 = p.umask[c - 1]; m.u1 = p.umask[c]; m.u2 = p.umask[c + 1]; m.u3 = p.umask[c + 2];
  m.v0 = p.vmask[c - SX]; m.v1 = p.vmask[c]; m.v2 = p.vmask[c + SX]; m.v3 = p.vmask[c + 2 * SX];
  return m;
}
struct Flx4 { double u0, u1, v0, v1; };                     // FlxU(c), FlxU(c+1), FlxV(c), FlxV(c+SX) at one level
__device__ __forceinline__ Flx4 flx4_load(const KP& p, ix_t c, int k) {
  const double* __restrict__ gU = p.FlxU;
  const double* __restrict__ gV = p.FlxV;
  Flx4 f;
  f.u0 = R3(gU, c, k); f.u1 = R3(gU, c + 1, k); f.v0 = R3(gV, c, k); f.v1 = R3(gV, c + p.SX, k);
  return f;
}
template <bool UP, bool NEAR>
__device__ __forceinline__ void tflux4(const KP& p, const double* __restrict__ t, ix_t c, int i, int j, int k, const TMask& m, const Flx4& f, double& FXi,
                                       double& FXip, double& FEj, double& FEjp) {
  const ix_t SX = p.SX;
  if (NEAR) {   // first differences in i exist on columns istr..iend+1 between walls (compute_horiz_tracer_fluxes.h:50-84)
    FXi = tflux_cl<UP>(p, t, f.u0, p.umask, c, 1, k, iclamp(p, i - 1, p.istr, p.iend + 1), iclamp(p, i, p.istr, p.iend + 1), iclamp(p, i + 1, p.istr, p.iend + 1));
    FXip = tflux_cl<UP>(p, t, f.u1, p.umask, c + 1, 1, k, iclamp(p, i, p.istr, p.iend + 1), iclamp(p, i + 1, p.istr, p.iend + 1),
                        iclamp(p, i + 2, p.istr, p.iend + 1));
    FEj = tflux_cl<UP>(p, t, f.v0, p.vmask, c, SX, k, jclamp(p, j - 1, p.jstr, p.jend + 1), jclamp(p, j, p.jstr, p.jend + 1), jclamp(p, j + 1, p.jstr, p.jend + 1));
    FEjp = tflux_cl<UP>(p, t, f.v1, p.vmask, c + SX, SX, k, jclamp(p, j, p.jstr, p.jend + 1), jclamp(p, j + 1, p.jstr, p.jend + 1),
                        jclamp(p, j + 2, p.jstr, p.jend + 1));
  } else {      // interior: the five values along each axis are loaded once and serve both faces
    const double a0 = R3(t, c - 2, k), a1 = R3(t, c - 1, k), t2 = R3(t, c, k), a3 = R3(t, c + 1, k), a4 = R3(t, c + 2, k);
    const double dA = (a1 - a0) * m.u0, dB = (t2 - a1) * m.u1, dC = (a3 - t2) * m.u2, dD = (a4 - a3) * m.u3;
    FXi = tflux_fin<UP>(a1, t2, dA, dB, dC, f.u0);
    FXip = tflux_fin<UP>(t2, a3, dB, dC, dD, f.u1);
    const double b0 = R3(t, c - 2 * SX, k), b1 = R3(t, c - SX, k), b3 = R3(t, c + SX, k), b4 = R3(t, c + 2 * SX, k);
    const double eA = (b1 - b0) * m.v0, eB = (t2 - b1) * m.v1, eC = (b3 - t2) * m.v2, eD = (b4 - b3) * m.v3;
    FEj = tflux_fin<UP>(b1, t2, eA, eB, eC, f.v0);
    FEjp = tflux_fin<UP>(t2, b3, eB, eC, eD, f.v1);
  }
}
template <bool UP, bool NEAR>
__device__ __forceinline__ double tdiv(const KP& p, const double* __restrict__ t, ix_t c, int i, int j, int k, const TMask& m, const Flx4& f) {
  double FXi, FXip, FEj, FEjp;
  tflux4<UP, NEAR>(p, t, c, i, j, k, m, f, FXi, FXip, FEj, FEjp);
  return FXip - FXi + FEjp - FEj;
}
// compute_horiz_tracer_fluxes.h:217-246: the flux through a river face is the river's tracer value times its volume flux,
// spread over the column in proportion to the layer thickness.  d = 1: u-face (i,j), d = SX: v-face.
__device__ __forceinline__ bool riv_on(double code) { return fabs(code) > 1e-3; }
__device__ __forceinline__ double riv_speed(const KP& p, double code, ix_t c, ix_t d, int& iriver) {   // vol*fraction/riv_depth
  const double riv_depth = 0.5 * (RW(p.z_w, c - d, p.N) - RW(p.z_w, c - d, 0) + RW(p.z_w, c, p.N) - RW(p.z_w, c, 0));
  iriver = (int)lround(code / 10);
  return p.riv_tab[(iriver - 1) * RIVROW] * (code - 10 * iriver) / riv_depth;
}
__device__ __forceinline__ double riv_tflux(const KP& p, double code, ix_t c, ix_t d, int k, int itrc) {
  int iriver;
  const double vel = riv_speed(p, code, c, d, iriver);
  return p.riv_tab[(iriver - 1) * RIVROW + itrc] * 0.5 * (R3(p.Hz, c - d, k) + R3(p.Hz, c, k)) * vel;
}
struct RivFaces { double u0, u1, v0, v1; bool any; };   // codes of the four faces of a cell
__device__ __forceinline__ RivFaces riv_faces(const KP& p, ix_t c) {
  RivFaces r;
  r.u0 = p.riv_uflx[c]; r.u1 = p.riv_uflx[c + 1]; r.v0 = p.riv_vflx[c]; r.v1 = p.riv_vflx[c + p.SX];
  r.any = riv_on(r.u0) | riv_on(r.u1) | riv_on(r.v0) | riv_on(r.v1);
  return r;
}
template <bool UP>
__device__ __forceinline__ double tdiv_riv(const KP& p, const double* __restrict__ t, ix_t c, int i, int j, int k, const TMask& m, const Flx4& f, bool near,
                                           const RivFaces& r, int itrc) {
  double FXi, FXip, FEj, FEjp;
  if (near) tflux4<UP, true>(p, t, c, i, j, k, m, f, FXi, FXip, FEj, FEjp);
  else tflux4<UP, false>(p, t, c, i, j, k, m, f, FXi, FXip, FEj, FEjp);
  if (riv_on(r.u0)) FXi = riv_tflux(p, r.u0, c, 1, k, itrc);
  if (riv_on(r.u1)) FXip = riv_tflux(p, r.u1, c + 1, 1, k, itrc);
  if (riv_on(r.v0)) FEj = riv_tflux(p, r.v0, c, p.SX, k, itrc);
  if (riv_on(r.v1)) FEjp = riv_tflux(p, r.v1, c + p.SX, p.SX, k, itrc);
  return FXip - FXi + FEjp - FEj;
}
// A batch of TB tracers handled by one launch (the reference loops "do itrc=1,nt" around the same code, pre_step3d4S.F:152,
// step3d_t_ISO.F:179): everything that does not depend on the tracer -- FlxU, FlxV, Hz, FlxDiv, We, Wi, the masks, the
// Hz-only spline elimination, the Thomas coefficients of tracers that share a diffusivity -- is read / computed once per batch.
template <int TB>
struct TrcBatch {
  const double* rhs[TB];    // t(:,:,:,nrhs,itrc)
  double* tnew[TB];         // t(:,:,:,nnew,itrc)
  const double* stp[TB];    // t(:,:,:,nstp,itrc)   (predictor)
  double* indx[TB];         // t(:,:,:,indx,itrc)   (predictor)
  const double* akt[TB];    // Akt(:,:,:,itrc<=2 ? itrc : isalt)
  const double* stflx[TB];  // stflx(:,:,itrc)      (corrector)
  int itrc[TB];             // 1-based tracer index: itemp = 1, isalt = 2
};
template <bool WALL, int KB, bool RIV, int TB>
__global__ void __launch_bounds__(256, (TB == 1 ? LB_TP : (TB == 2 ? 4 : 3))) k_pre_t_horiz(const __grid_constant__ KP p, const __grid_constant__ TrcBatch<TB> tb, double dtau, double cf_stp,
                                                            double cf_bak) {
  THREAD_IJ(p.istr, p.iend, p.jstr, p.jend)
  const int k0 = blockIdx.z * KB + 1;
  const TMask m = tmask_load(p, c);
  const double cdt = dtau * p.pm[c] * p.pn[c];
  const bool near = WALL && near_wall(p, i, j);
  RivFaces rf;
  rf.any = false;
  if (RIV) rf = riv_faces(p, c);
  const double* __restrict__ gHz = p.Hz;
  const double* __restrict__ gFd = p.FlxDiv;
#pragma unroll
  for (int q = 0; q < KB; ++q) {
    const int k = k0 + q;
    if (k <= p.N) {
      const Flx4 f = flx4_load(p, c, k);
      const double Hz = R3(gHz, c, k), Hz_bak = Hz + R3(gFd, c, k);
#pragma unroll
      for (int t = 0; t < TB; ++t) {
        const double* __restrict__ t_rhs = tb.rhs[t];
        const double dv = (RIV && rf.any) ? tdiv_riv<false>(p, t_rhs, c, i, j, k, m, f, near, rf, tb.itrc[t])
                                          : (near ? tdiv<false, true>(p, t_rhs, c, i, j, k, m, f) : tdiv<false, false>(p, t_rhs, c, i, j, k, m, f));
        const double ts = R3(tb.stp[t], c, k);
        R3(tb.tnew[t], c, k) = Hz_bak * (cf_stp * ts + cf_bak * R3(tb.indx[t], c, k)) - cdt * (dv);
        R3(tb.indx[t], c, k) = Hz * ts;
      }
    }
  }
}
template <bool WALL, int KB, bool RIV, int TB>
__global__ void __launch_bounds__(256, (TB == 1 ? LB_TC : (TB == 2 ? 3 : 2))) k_t_horiz(const __grid_constant__ KP p, const __grid_constant__ TrcBatch<TB> tb) {
  THREAD_IJ(p.istr, p.iend, p.jstr, p.jend)
  const int k0 = blockIdx.z * KB + 1;
  const TMask m = tmask_load(p, c);
  const double cdt = p.dt * p.pm[c] * p.pn[c];
  const bool near = WALL && near_wall(p, i, j);
  RivFaces rf;
  rf.any = false;
  if (RIV) rf = riv_faces(p, c);
#pragma unroll
  for (int q = 0; q < KB; ++q) {
    const int k = k0 + q;
    if (k <= p.N) {
      const Flx4 f = flx4_load(p, c, k);
#pragma unroll
      for (int t = 0; t < TB; ++t) {
        const double* __restrict__ t_rhs = tb.rhs[t];
        const double dv = (RIV && rf.any) ? tdiv_riv<true>(p, t_rhs, c, i, j, k, m, f, near, rf, tb.itrc[t])
                                          : (near ? tdiv<true, true>(p, t_rhs, c, i, j, k, m, f) : tdiv<true, false>(p, t_rhs, c, i, j, k, m, f));
        R3(tb.tnew[t], c, k) = R3(tb.tnew[t], c, k) - cdt * (dv);
      }
    }
  }
}

// pre_step3d4S.F:493-522 / step3d_uv2.F:689-717: the velocity through a river face is the river's, uniform over the column.
// One thread per (i,j); iu0/jv0 = first u-column / v-row of the reference's loops (istrU,jstrV in pre_step3d; istr,jstr in uv2).
__global__ void k_river_uv(const __grid_constant__ KP p, double* __restrict__ u, double* __restrict__ v, int iu0, int jv0) {
  THREAD_IJ(p.istr, p.iend, p.jstr, p.jend)
  const double cu = p.riv_uflx[c], cv = p.riv_vflx[c];
  int iriver;
  if (i >= iu0 && riv_on(cu)) {
    const double riv_depth = 0.5 * (RW(p.z_w, c - 1, p.N) - RW(p.z_w, c - 1, 0) + RW(p.z_w, c, p.N) - RW(p.z_w, c, 0));
    iriver = (int)lround(cu / 10);
    const double vel = p.riv_tab[(iriver - 1) * RIVROW] * (cu - 10 * iriver) / (p.dn_u[c] * riv_depth);
    for (int k = 1; k <= p.N; ++k) R3(u, c, k) = vel;
  }
  if (j >= jv0 && riv_on(cv)) {
    const double riv_depth = 0.5 * (RW(p.z_w, c - SX, p.N) - RW(p.z_w, c - SX, 0) + RW(p.z_w, c, p.N) - RW(p.z_w, c, 0));
    iriver = (int)lround(cv / 10);
    const double vel = p.riv_tab[(iriver - 1) * RIVROW] * (cv - 10 * iriver) / (p.dm_v[c] * riv_depth);
    for (int k = 1; k <= p.N; ++k) R3(v, c, k) = vel;
  }
}

// compute_rd_bott_drag.h:1-74 (log-layer drag when Zob>0, else linear rdrg capped by 0.8*Hz/dt)
__global__ void k_rd_bott_drag(const __grid_constant__ KP p, const double* __restrict__ u, const double* __restrict__ v) {
  THREAD_IJ(p.istrU - 1, p.iend, p.jstrV - 1, p.jend)
  const double vonKar = 0.41;
  if (p.Zob > 0.0) {
    const double u0 = R3(u, c, 1), u1 = R3(u, c + 1, 1), v0 = R3(v, c, 1), v1 = R3(v, c + SX, 1);
    const double cff = sqrt(0.333333333333 * (u0 * u0 + u1 * u1 + u0 * u1 + v0 * v0 + v1 * v1 + v0 * v1));
    const double lg = vonKar / log(1.0 + 0.5 * R3(p.Hz, c, 1) / p.Zob);
    p.r_D[c] = cff * (lg * lg);
  } else {
    p.r_D[c] = DMIN(p.rdrg, 0.8 * R3(p.Hz, c, 1) / p.dt);
  }
}

// ------------------------------------------------------------------------------------------------
// Tracer column kernel: compute_vert_tracer_fluxes.h:1-140 (SPLINE_TS) + flux divergence + implicit
// vertical diffusion/advection Thomas solve, for a batch of TB tracers per thread.
// MODE 0: predictor (pre_step3d4S.F:201-263, dtau, Hz_fwd);
// MODE 1: corrector (step3d_t_ISO.F:907-1100: dt, Hz, swflx/stflx at the surface, the LMD_KPP solar-penetration and
//         non-local transport terms :961-998, pipes :927-935, rmask).
// Four sweeps (up, down, up, down).  Scratch (global slab, [k][j][i]): array 0 = spline CF(2:N), shared by the batch (it
// depends on Hz only); array 1+t = spline FC(0:N-1) of tracer t.  The right-hand side and the Thomas DC of every tracer are
// parked IN t(nnew) itself (each level is read before it is overwritten), and the Thomas CF re-uses the dead spline slots:
// array 0 when the batch shares one diffusivity (SHARE), else array 1+t.
#define KC 4   // levels per register chunk: all global/scratch loads of a chunk are issued before its arithmetic
#ifndef LB_TV1
#define LB_TV1 12
#endif
#ifndef LB_TV2
#define LB_TV2 8
#endif
#ifndef LB_TV4
#define LB_TV4 5
#endif
template <int MODE, int TB, bool SHARE>
__global__ void __launch_bounds__(64, (TB == 1 ? LB_TV1 : (TB == 2 ? LB_TV2 : LB_TV4)))
k_t_vert(const __grid_constant__ KP p, const __grid_constant__ TrcBatch<TB> tb, double dtau) {
  extern __shared__ double smem[];
  THREAD_IJ(p.istr, p.iend, p.jstr, p.jend)
  ColScr sc = col_scratch(p, c, smem);
  constexpr int TA = SHARE ? 1 : TB;     // distinct diffusivities in the batch
  const int N = p.N;
  const double* __restrict__ gHz = p.Hz;
  const double* __restrict__ gWe = p.We;
  const double* __restrict__ gWi = p.Wi;
  const double* __restrict__ gFd = p.FlxDiv;
  // ---- sweep 1 (up): spline forward elimination
  double Hzk = R3(gHz, c, 1), CFk = 1.0;
  double tk[TB], FCkm[TB];
#pragma unroll
  for (int t = 0; t < TB; ++t) {
    tk[t] = R3(tb.rhs[t], c, 1);
    FCkm[t] = 2.0 * tk[t];
    S(1 + t, 0) = FCkm[t];
  }
  for (int k0 = 1; k0 <= N - 1; k0 += KC) {
    double hz[KC], tt[TB][KC];
#pragma unroll
    for (int q = 0; q < KC; ++q)
      if (k0 + q <= N - 1) {
        hz[q] = R3(gHz, c, k0 + q + 1);
#pragma unroll
        for (int t = 0; t < TB; ++t) { const double* __restrict__ r = tb.rhs[t]; tt[t][q] = R3(r, c, k0 + q + 1); }
      }
#pragma unroll
    for (int q = 0; q < KC; ++q) {
      const int k = k0 + q;
      if (k <= N - 1) {
        const double Hzp = hz[q];
        const double cff = 1.0 / (2.0 * Hzk + Hzp * (2.0 - CFk));
        CFk = cff * Hzk;                                            // CF(k+1)
        S(0, k + 1) = CFk;
#pragma unroll
        for (int t = 0; t < TB; ++t) {
          const double tp = tt[t][q];
          FCkm[t] = cff * (3.0 * (Hzk * tp + Hzp * tk[t]) - Hzp * FCkm[t]);   // FC(k)
          S(1 + t, k) = FCkm[t];
          tk[t] = tp;
        }
        Hzk = Hzp;
      }
    }
  }
  // ---- sweep 2 (down): back-substitution, vertical flux and its divergence -> rhs(k) parked in t(nnew)
  const double cdt = dtau * p.pm[c] * p.pn[c];
  const int pidx = (MODE == 1 && p.pipes) ? (int)p.pipe_idx[c] : 0;
  const double pflx = pidx > 0 ? p.pipe_flx[c] : 0.0;
  const bool kpp = MODE == 1 && p.lmd_kpp;
  const double srflx = kpp ? p.srflx[c] : 0.0;
  double FCs_p[TB], flux_p[TB], ptrc[TB], sfl[TB], kcf_up[TB];
#pragma unroll
  for (int t = 0; t < TB; ++t) {
    FCs_p[t] = (2.0 * tk[t] - FCkm[t]) / (1.0 - CFk);   // FC(N) spline value
    flux_p[t] = 0.0;                                     // FC(N)=0 after the sweep
    ptrc[t] = pidx > 0 ? p.pipe_tab[(pidx - 1) * PIPROW(N) + N + tb.itrc[t] - 1] : 0.0;
    sfl[t] = MODE == 1 ? tb.stflx[t][c] : 0.0;
    kcf_up[t] = 0.0;
  }
  for (int k0 = N - 1; k0 >= 0; k0 -= KC) {
    double a0[KC], we[KC], b1[TB][KC], tn[TB][KC];
#pragma unroll
    for (int q = 0; q < KC; ++q) {
      const int k = k0 - q;
      if (k >= 0) {
        if (k >= 1) { a0[q] = S(0, k + 1); we[q] = RW(gWe, c, k); }
#pragma unroll
        for (int t = 0; t < TB; ++t) {
          tn[t][q] = R3(tb.tnew[t], c, k + 1);
          if (k >= 1) b1[t][q] = S(1 + t, k);
        }
      }
    }
#pragma unroll
    for (int q = 0; q < KC; ++q) {
      const int k = k0 - q;
      if (k >= 0) {
        double gh = 0.0, sw = 0.0;
        if (kpp && k >= 1) { gh = RW(p.ghat, c, k); sw = RW(p.swr_frac, c, k); }
#pragma unroll
        for (int t = 0; t < TB; ++t) {
          double FCs = 0.0, flux = 0.0;                        // FC(0) of the spline is never used (flux(0)=0)
          if (k >= 1) { FCs = b1[t][q] - a0[q] * FCs_p[t]; flux = FCs * we[q]; }
          double r = tn[t][q] - cdt * (flux_p[t] - flux);      // rhs(k+1) = t_new(k+1) - cdt*(flux(k+1)-flux(k))
          if (MODE == 1 && pidx > 0) r = r + cdt * pflx * p.pipe_tab[(pidx - 1) * PIPROW(N) + k] * ptrc[t];   // step3d_t_ISO.F:927-935 (level k+1)
          if (MODE == 1 && k + 1 == N) {
            if (tb.itrc[t] == 1) r = r + p.dt * p.swflx[c] * r / R3(gHz, c, N);
            r = r + p.dt * sfl[t];
          }
          if (kpp && tb.itrc[t] <= 2) {   // step3d_t_ISO.F:961-998, reference order k=N-1..1: level k+1 first loses dt*cff(k+1)... see below
            // the reference's loop visits interface k+1 (adds +dt*cff(k+1) to level k+1) BEFORE interface k (subtracts dt*cff(k))
            if (tb.itrc[t] == 1) {
              if (k + 1 <= N - 1) r = r + p.dt * kcf_up[t];
              double cf = 0.0;
              if (k >= 1) { cf = srflx * sw - gh * (sfl[t] - srflx); r = r - p.dt * cf; }
              kcf_up[t] = cf;
            } else {
              if (k + 1 <= N - 1) r = r + kcf_up[t];
              double cf = 0.0;
              if (k >= 1) { cf = -p.dt * gh * sfl[t]; r = r - cf; }
              kcf_up[t] = cf;
            }
          }
          R3(tb.tnew[t], c, k + 1) = r;
          FCs_p[t] = FCs; flux_p[t] = flux;
        }
      }
    }
  }
  // ---- sweep 3 (up): Thomas forward elimination
  const double DC0 = cdt;
  double Hzf1 = R3(gHz, c, 1), Hzf2 = R3(gHz, c, 2);
  if (MODE == 0) { Hzf1 -= R3(gFd, c, 1); Hzf2 -= R3(gFd, c, 2); }
  double WC = DC0 * RW(gWi, c, 1);
  double FCt[TA], CFt[TA], DCt[TB];
  {
    double cffa[TA];
#pragma unroll
    for (int a = 0; a < TA; ++a) {
      FCt[a] = 2.0 * dtau * RW(tb.akt[a], c, 1) / (MODE == 0 ? (Hzf2 + Hzf1) : (Hzf1 + Hzf2));
      cffa[a] = 1.0 / (Hzf1 + FCt[a] + DMAX(WC, 0.0));
      CFt[a] = cffa[a] * (FCt[a] - DMIN(WC, 0.0));
      S(SHARE ? 0 : 1 + a, 1) = CFt[a];
    }
#pragma unroll
    for (int t = 0; t < TB; ++t) {
      DCt[t] = cffa[SHARE ? 0 : t] * R3(tb.tnew[t], c, 1);
      R3(tb.tnew[t], c, 1) = DCt[t];
    }
  }
  double Hzfk = Hzf2;
  for (int k0 = 2; k0 <= N - 1; k0 += KC) {
    double hz[KC], wi[KC], ak[TA][KC], rh[TB][KC];
#pragma unroll
    for (int q = 0; q < KC; ++q) {
      const int k = k0 + q;
      if (k <= N - 1) {
        hz[q] = R3(gHz, c, k + 1);
        if (MODE == 0) hz[q] -= R3(gFd, c, k + 1);
        wi[q] = RW(gWi, c, k);
#pragma unroll
        for (int a = 0; a < TA; ++a) { const double* __restrict__ A = tb.akt[a]; ak[a][q] = RW(A, c, k); }
#pragma unroll
        for (int t = 0; t < TB; ++t) rh[t][q] = R3(tb.tnew[t], c, k);
      }
    }
#pragma unroll
    for (int q = 0; q < KC; ++q) {
      const int k = k0 + q;
      if (k <= N - 1) {
        const double Hzfp = hz[q];
        const double WCk = DC0 * wi[q];
        double cffa[TA], FCtk[TA];
#pragma unroll
        for (int a = 0; a < TA; ++a) {
          FCtk[a] = 2.0 * dtau * ak[a][q] / (MODE == 0 ? (Hzfp + Hzfk) : (Hzfk + Hzfp));
          cffa[a] = 1.0 / (Hzfk + FCtk[a] + DMAX(WCk, 0.0) + FCt[a] - DMIN(WC, 0.0) - CFt[a] * (FCt[a] + DMAX(WC, 0.0)));
        }
#pragma unroll
        for (int t = 0; t < TB; ++t) {
          const int a = SHARE ? 0 : t;
          DCt[t] = cffa[a] * (rh[t][q] + DCt[t] * (FCt[a] + DMAX(WC, 0.0)));
          R3(tb.tnew[t], c, k) = DCt[t];
        }
#pragma unroll
        for (int a = 0; a < TA; ++a) {
          CFt[a] = cffa[a] * (FCtk[a] - DMIN(WCk, 0.0));
          S(SHARE ? 0 : 1 + a, k) = CFt[a];
          FCt[a] = FCtk[a];
        }
        WC = WCk; Hzfk = Hzfp;
      }
    }
  }
  // ---- sweep 4 (down): back-substitution
  const double rm = MODE == 1 ? p.rmask[c] : 1.0;
  double tkp[TB];
#pragma unroll
  for (int t = 0; t < TB; ++t) {
    const int a = SHARE ? 0 : t;
    tkp[t] = (R3(tb.tnew[t], c, N) + DCt[t] * (FCt[a] + DMAX(WC, 0.0))) / (Hzfk + FCt[a] - DMIN(WC, 0.0) - CFt[a] * (FCt[a] + DMAX(WC, 0.0)));
    if (MODE == 1) tkp[t] = tkp[t] * rm;
    R3(tb.tnew[t], c, N) = tkp[t];
  }
  for (int k0 = N - 1; k0 >= 1; k0 -= KC) {
    double dc[TB][KC], cf[TA][KC];
#pragma unroll
    for (int q = 0; q < KC; ++q)
      if (k0 - q >= 1) {
#pragma unroll
        for (int a = 0; a < TA; ++a) cf[a][q] = S(SHARE ? 0 : 1 + a, k0 - q);
#pragma unroll
        for (int t = 0; t < TB; ++t) dc[t][q] = R3(tb.tnew[t], c, k0 - q);
      }
#pragma unroll
    for (int q = 0; q < KC; ++q) {
      const int k = k0 - q;
      if (k >= 1) {
#pragma unroll
        for (int t = 0; t < TB; ++t) {
          tkp[t] = dc[t][q] + cf[SHARE ? 0 : t][q] * tkp[t];
          if (MODE == 1) tkp[t] = tkp[t] * rm;
          R3(tb.tnew[t], c, k) = tkp[t];
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Momentum column kernel: compute_vert_rhs_uv_terms.h:1-188 (SPLINE_UV vertical advection) fused with
// the implicit vertical viscosity/advection Thomas solve.  DIR 0: u, 1: v.
// MODE 0: predictor (pre_step3d4S.F:276-488): rhs from Hz_bak*(cf_stp*u(nstp)+cf_bak*u(indx)), Hz_fwd,
//         dtau; also u(indx) = Hz_u*u(nstp).
// MODE 1: corrector (step3d_uv1.F:136-280): rhs = u(nnew) [=Hz_u*u(nstp)] + dt*..*ru, Hz, dt, bottom
//         drag r_D, output Hz_u*u and the barotropic forcing rufrc/rvfrc.
// Sweeps: up (spline elimination), down (spline back-substitution + final ru + Thomas elimination),
// up (Thomas back-substitution).  A: spline CF(2:N) -> Thomas CF(k-1) in slot k; B: spline FC(0:N-1) -> DC(2:N).
#define KC2 2
template <int MODE, int DIR>
__global__ void __launch_bounds__(64, 8)
k_uv_vert(const __grid_constant__ KP p, const double* __restrict__ u_rhs, double* __restrict__ u_new, const double* __restrict__ u_stp,
          double* __restrict__ u_indx, double dtau, double cf_stp, double cf_bak) {
  extern __shared__ double smem[];
  THREAD_IJ((DIR == 0 ? p.istrU : p.istr), p.iend, (DIR == 0 ? p.jstr : p.jstrV), p.jend)
  ColScr sc = col_scratch(p, c, smem);
  const int N = p.N;
  const ix_t d = DIR == 0 ? 1 : SX;
  double* __restrict__ r = DIR == 0 ? p.ru : p.rv;
  const double* msk = DIR == 0 ? p.umask : p.vmask;
  const double strs = DIR == 0 ? p.sustr[c] : p.svstr[c];
  const double* __restrict__ gHz = p.Hz;
  const double* __restrict__ gWe = p.We;
  const double* __restrict__ gWi = p.Wi;
  const double* __restrict__ gFd = p.FlxDiv;
  const double* __restrict__ gAkv = p.Akv;
  // ---- sweep 1 (up): spline elimination
#define HZ4(k) (0.5625 * (R3(gHz, c, k) + R3(gHz, c - d, k)) - 0.0625 * (R3(gHz, c + d, k) + R3(gHz, c - 2 * d, k)))
  double DCk = HZ4(1), uk = R3(u_rhs, c, 1);
  double CFk = 1.0, FCkm = 2.0 * uk;
  S(1, 0) = FCkm;
  for (int k0 = 1; k0 <= N - 1; k0 += KC) {
    double h0[KC], h1[KC], h2[KC], h3[KC], uu[KC];
#pragma unroll
    for (int q = 0; q < KC; ++q)
      if (k0 + q <= N - 1) {
        const int k = k0 + q + 1;
        h0[q] = R3(gHz, c, k); h1[q] = R3(gHz, c - d, k); h2[q] = R3(gHz, c + d, k); h3[q] = R3(gHz, c - 2 * d, k);
        uu[q] = R3(u_rhs, c, k);
      }
#pragma unroll
    for (int q = 0; q < KC; ++q) {
      const int k = k0 + q;
      if (k <= N - 1) {
        const double DCp = 0.5625 * (h0[q] + h1[q]) - 0.0625 * (h2[q] + h3[q]), up = uu[q];
        const double cff = 1.0 / (2.0 * DCk + DCp * (2.0 - CFk));
        CFk = cff * DCk;
        FCkm = cff * (3.0 * (DCk * up + DCp * uk) - DCp * FCkm);
        S(0, k + 1) = CFk;
        S(1, k) = FCkm;
        DCk = DCp; uk = up;
      }
    }
  }
#undef HZ4
  // ---- sweep 2 (down): spline back-substitution -> vertical flux DCf(k) -> ru(k+1) final -> Thomas row k+1
  const double mp = msk[c + d], mm = msk[c - d];
  const double DC0 = dtau * 0.25 * (p.pm[c] + p.pm[c - d]) * (p.pn[c] + p.pn[c - d]);
  const double rD = 0.5 * dtau * (p.r_D[c] + p.r_D[c - d]);
  double FCs_p = (2.0 * uk - FCkm) / (1.0 - CFk);   // FC(N)
  double DCf_p = 0.0;                                // DC(N)=0
  double FCt_k = 0, WC_k = 0, CFt_k = 0, DCt_kp = 0;   // Thomas FC(k), WC(k), CF(k), DC(k+1) of the row above
  // raw thicknesses of the row being eliminated (level kk), carried from the previous iteration's loads
  double Hz_c = R3(gHz, c, N), Hz_d = R3(gHz, c - d, N), Fd_c = 0.0, Fd_d = 0.0;
  if (MODE == 0) { Fd_c = R3(gFd, c, N); Fd_d = R3(gFd, c - d, N); }
  for (int k0 = N - 1; k0 >= 0; k0 -= KC2) {
    double sB[KC2], sA[KC2], w0[KC2], w1[KC2], w2[KC2], w3[KC2], rr[KC2], us[KC2], ui[KC2];
    double hcm[KC2], hdm[KC2], fcm[KC2], fdm[KC2], akc[KC2], akd[KC2], wic[KC2], wid[KC2];
#pragma unroll
    for (int q = 0; q < KC2; ++q) {
      const int k = k0 - q, kk = k + 1;
      if (k >= 0) {
        if (k >= 1) {
          sB[q] = S(1, k); sA[q] = S(0, k + 1);
          w0[q] = RW(gWe, c, k); w1[q] = RW(gWe, c - d, k); w2[q] = RW(gWe, c + d, k); w3[q] = RW(gWe, c - 2 * d, k);
          hcm[q] = R3(gHz, c, kk - 1); hdm[q] = R3(gHz, c - d, kk - 1);
          if (MODE == 0) { fcm[q] = R3(gFd, c, kk - 1); fdm[q] = R3(gFd, c - d, kk - 1); }
          akc[q] = RW(gAkv, c, kk - 1); akd[q] = RW(gAkv, c - d, kk - 1);
          wic[q] = RW(gWi, c, kk - 1); wid[q] = RW(gWi, c - d, kk - 1);
        }
        rr[q] = R3(r, c, kk);
        if (MODE == 0) { us[q] = R3(u_stp, c, kk); ui[q] = R3(u_indx, c, kk); }
        else us[q] = R3(u_new, c, kk);
      }
    }
#pragma unroll
    for (int q = 0; q < KC2; ++q) {
      const int k = k0 - q, kk = k + 1;   // kk: row whose final ru becomes known in this iteration
      if (k >= 0) {
        double DCf = 0.0;
        if (k >= 1) {
          const double FCs = sB[q] - sA[q] * FCs_p;
          DCf = FCs * 0.5 * (w0[q] + w1[q] - 0.125 * ((w2[q] - w0[q]) * mp - (w1[q] - w3[q]) * mm));
          FCs_p = FCs;
        }
        const double ruf = (k >= 1) ? (rr[q] - DCf_p + DCf) : (rr[q] - DCf_p);
        if (MODE == 1) R3(r, c, kk) = ruf;   // kept for the ascending rufrc sum (step3d_uv1.F:194-206)
        DCf_p = DCf;
        double rhs;
        if (MODE == 0) {
          const double Hzb = 0.5 * ((Hz_c + Fd_c) + (Hz_d + Fd_d));
          rhs = Hzb * (cf_stp * us[q] + cf_bak * ui[q]) + DC0 * ruf;
          R3(u_indx, c, kk) = 0.5 * (Hz_c + Hz_d) * us[q];
        } else {
          rhs = us[q] + DC0 * ruf;
        }
        const double Hzf_c = MODE == 0 ? (Hz_c - Fd_c) : Hz_c, Hzf_d = MODE == 0 ? (Hz_d - Fd_d) : Hz_d;
        if (kk == 1) {
          const double u1 = (rhs + DCt_kp * (FCt_k - DMIN(WC_k, 0.0))) /
                            (0.5 * (Hzf_c + Hzf_d) + rD + FCt_k + DMAX(WC_k, 0.0) - CFt_k * (FCt_k - DMIN(WC_k, 0.0)));
          S(1, 1) = u1;
        } else {
          const double Hzf_cm = MODE == 0 ? (hcm[q] - fcm[q]) : hcm[q], Hzf_dm = MODE == 0 ? (hdm[q] - fdm[q]) : hdm[q];
          const double FCt_m = 2.0 * dtau * (akc[q] + akd[q]) / (Hzf_c + Hzf_d + Hzf_cm + Hzf_dm);
          const double WC_m = DC0 * 0.5 * (wic[q] + wid[q]);
          double cff, DCt;
          if (kk == N) {
            cff = 1.0 / (0.5 * (Hzf_c + Hzf_d) + FCt_m - DMIN(WC_m, 0.0));
            DCt = cff * (rhs + dtau * strs);
          } else {
            cff = 1.0 / (0.5 * (Hzf_c + Hzf_d) + FCt_m - DMIN(WC_m, 0.0) + FCt_k + DMAX(WC_k, 0.0) - CFt_k * (FCt_k - DMIN(WC_k, 0.0)));
            DCt = cff * (rhs + DCt_kp * (FCt_k - DMIN(WC_k, 0.0)));
          }
          const double CFt_m = cff * (FCt_m + DMAX(WC_m, 0.0));
          S(0, kk) = CFt_m;         // Thomas CF(kk-1) goes to slot A(kk): spline CF(kk) was consumed above, A(kk-1) is still live
          S(1, kk) = DCt;           // Thomas DC(kk): spline FC(kk) is dead
          FCt_k = FCt_m; WC_k = WC_m; CFt_k = CFt_m; DCt_kp = DCt;
          Hz_c = hcm[q]; Hz_d = hdm[q];
          if (MODE == 0) { Fd_c = fcm[q]; Fd_d = fdm[q]; }
        }
      }
    }
  }
  // ---- sweep 3 (up): back-substitution
  double uprev = S(1, 1);
  if (MODE == 0) {
    R3(u_new, c, 1) = uprev;
    for (int k0 = 2; k0 <= N; k0 += KC) {
      double dc[KC], cf[KC];
#pragma unroll
      for (int q = 0; q < KC; ++q)
        if (k0 + q <= N) { dc[q] = S(1, k0 + q); cf[q] = S(0, k0 + q); }
#pragma unroll
      for (int q = 0; q < KC; ++q)
        if (k0 + q <= N) { uprev = dc[q] + cf[q] * uprev; R3(u_new, c, k0 + q) = uprev; }
    }
  } else {
    const double dmn = DIR == 0 ? p.dm_u[c] * p.dn_u[c] : p.dm_v[c] * p.dn_v[c];
    double* rfrc = DIR == 0 ? p.rufrc : p.rvfrc;
    R3(u_new, c, 1) = uprev * 0.5 * (R3(gHz, c, 1) + R3(gHz, c - d, 1));
    double rf = R3(r, c, 1) + dmn * (strs - 0.5 * (p.r_D[c - d] + p.r_D[c]) * uprev);
    for (int k0 = 2; k0 <= N; k0 += KC) {
      double dc[KC], cf[KC], ha[KC], hb[KC], rk[KC];
#pragma unroll
      for (int q = 0; q < KC; ++q)
        if (k0 + q <= N) {
          dc[q] = S(1, k0 + q); cf[q] = S(0, k0 + q);
          ha[q] = R3(gHz, c, k0 + q); hb[q] = R3(gHz, c - d, k0 + q); rk[q] = R3(r, c, k0 + q);
        }
#pragma unroll
      for (int q = 0; q < KC; ++q)
        if (k0 + q <= N) {
          uprev = dc[q] + cf[q] * uprev;
          R3(u_new, c, k0 + q) = uprev * 0.5 * (ha[q] + hb[q]);
          rf = rf + rk[q];
        }
    }
    rfrc[c] = rf;
  }
}

// ------------------------------------------------------------------------------------------------
// step2d_FB.F:24-576, stage 1: zeta_new over [istr-1:iend]x[jstr-1:jend] (kept in a private plane),
// zeta(knew) and the fast-time averages Zt_avg1, DU_avg2, DV_avg2 (+ DU_avg_bak on the first substep)
// on the interior.  DUon/DVom of the reference are recomputed in registers.
struct S2D {
  const double *zs, *zb, *zo, *us, *ub, *uo, *vs, *vb, *vo;   // kstp, kbak, kold levels
  double *zn, *un, *vn;                                       // knew
  double fwd, fwd1, fwd2, bkw_new, bkw, bkw1, bkw2, w1, w2;
  int iif, nfast;
  int region;     // 0: every interior point; 1: all but the ring of width 2 along the subdomain's edges; 2: that ring only (halo-swap overlap)
};
// the ring of width 2 whose results travel in the halo messages (and whose inputs come from them)
__device__ __forceinline__ bool s2d_in_ring(const KP& p, int i, int j) { return i < p.istr + 2 || i > p.iend - 2 || j < p.jstr + 2 || j > p.jend - 2; }
__device__ __forceinline__ double s2d_Drhs(const KP& p, const S2D& s, ix2_t c) { return p.h[c] + s.fwd * s.zs[c] + s.fwd1 * s.zb[c] + s.fwd2 * s.zo[c]; }
__device__ __forceinline__ double s2d_DUon(const KP& p, const S2D& s, ix2_t c) {
  const double urhs = s.fwd * s.us[c] + s.fwd1 * s.ub[c] + s.fwd2 * s.uo[c];
  return 0.5 * (s2d_Drhs(p, s, c) + s2d_Drhs(p, s, c - 1)) * p.dn_u[c] * (urhs);
}
__device__ __forceinline__ double s2d_DVom(const KP& p, const S2D& s, ix2_t c) {
  const double vrhs = s.fwd * s.vs[c] + s.fwd1 * s.vb[c] + s.fwd2 * s.vo[c];
  return 0.5 * (s2d_Drhs(p, s, c) + s2d_Drhs(p, s, c - p.SX)) * p.dm_v[c] * (vrhs);
}
__device__ __forceinline__ double s2d_zeta_new(const KP& p, const S2D& s, ix2_t c) {
  const double DUon = s2d_DUon(p, s, c), DUon1 = s2d_DUon(p, s, c + 1), DVom = s2d_DVom(p, s, c), DVom1 = s2d_DVom(p, s, c + p.SX);
  double zn = s.zs[c] + p.dtfast * p.pm[c] * p.pn[c] * (DUon - DUon1 + DVom - DVom1) + p.dtfast * p.swflx[c];
  if (p.pipes && p.pipe_idx[c] > 0.) zn = zn + p.dtfast * p.pm[c] * p.pn[c] * p.pipe_flx[c];   // step2d_FB.F:155-159
  return zn * p.rmask[c];
}
// pressure-gradient terms rubar,rvbar (step2d_FB.F:245-268, first-substep correction :313-384)
struct S2Pt { double UFx, UFe, VFx, VFe; };
__device__ __forceinline__ double s2d_pgf(const KP& p, double cff, double dn, ix2_t c, ix2_t d, const S2Pt& a, const S2Pt& b) {
  // a at c-d, b at c
  return cff * dn *
         ((p.h[c - d] + p.h[c]) * (a.UFe - b.UFe) + a.VFe - b.VFe +
          (p.h[c - d] - p.h[c]) * (a.VFx + b.VFx + 0.333333333333 * (p.rhoA[c - d] - p.rhoA[c]) * (a.UFx - b.UFx)));
}
// ---- fused barotropic substep ---------------------------------------------------------------------------------
// One launch per substep: a 32x8 tile first evaluates zeta_new on its points plus the row below and the column to
// the left (33x9 values in shared memory -- the reference's private zeta_new plane never exists in HBM), then every
// thread finishes its own point: zeta(knew), the fast-time averages, rubar/rvbar, ubar/vbar(knew).  The thread next
// to a closed wall also owns the wall point beside it (zetabc / u2dbc / v2dbc, step2d_FB.F:449-529), and on a single
// subdomain the results are stored to their periodic images as well, so no exchange launch follows.
#ifndef S2_TX
#define S2_TX 32
#endif
#ifndef S2_TY
#define S2_TY 8
#endif
__device__ __forceinline__ double s2d_zn_at(const KP& p, const S2D& s, ix2_t c, int j) {
  if (p.swall && j < p.jstr) return s2d_zeta_new(p, s, c + p.SX) * p.rmask[c];
  if (p.nwall && j > p.jend) return s2d_zeta_new(p, s, c - p.SX) * p.rmask[c];
  return s2d_zeta_new(p, s, c);
}
__device__ __forceinline__ S2Pt s2d_ptz(const KP& p, const S2D& s, ix2_t c, double zn, bool corr) {
  S2Pt r;
  const double zs = s.zs[c];
  if (!corr) {
    r.UFx = s.bkw_new * zn + s.bkw * zs + s.bkw1 * s.zb[c] + s.bkw2 * s.zo[c];
    r.UFe = (1.0 + p.rhoS[c]) * r.UFx;
    r.VFx = r.UFx * (p.rhoS[c] - p.rhoA[c]);
    r.VFe = r.UFe * r.UFx;
  } else {
    r.UFx = zn - zs;
    r.UFe = (1.0 + p.rhoS[c]) * r.UFx;
    r.VFe = r.UFe * (zn + zs);
    r.VFx = r.UFx * (p.rhoS[c] - p.rhoA[c]);
  }
  return r;
}
// store to (i,j) and, on a single periodic subdomain, to its images in the halo (mpi_setup.F:231-250 wrap)
__device__ __forceinline__ void s2d_store(const KP& p, double* __restrict__ a, int i, int j, ix2_t c, double v) {
  a[c] = v;
  if (p.wrap_x | p.wrap_y) {
    const ix2_t ox = !p.wrap_x ? 0 : (i <= 2 ? (ix2_t)p.Lm : (i >= p.Lm - 1 ? -(ix2_t)p.Lm : 0));
    const ix2_t oy = !p.wrap_y ? 0 : (j <= 2 ? (ix2_t)p.Mm * p.SX : (j >= p.Mm - 1 ? -(ix2_t)p.Mm * p.SX : 0));
    if (ox) a[c + ox] = v;
    if (oy) a[c + oy] = v;
    if (ox && oy) a[c + ox + oy] = v;
  }
}
// zeta(knew), Zt_avg1, DU_avg2, DV_avg2 (and the first-substep resets) at one point; returns the running DU_avg1/DV_avg1
__device__ __forceinline__ void s2d_avg_out(const KP& p, const S2D& s, int i, int j, ix2_t c, double zn, double DUon, double DVom, double& DU1, double& DV1,
                                            double& Zt) {
  s2d_store(p, s.zn, i, j, c, zn);
  if (s.iif == 1) {
    p.DU_avg_bak[c] = p.DU_avg1[c] - 0.1024390243902439 * p.DU_avg2[c];
    p.DV_avg_bak[c] = p.DV_avg1[c] - 0.1024390243902439 * p.DV_avg2[c];
    Zt = s.w1 * zn;
    DU1 = 0.0;
    DV1 = 0.0;
    p.DU_avg2[c] = s.w2 * DUon;
    p.DV_avg2[c] = s.w2 * DVom;
  } else {
    Zt = p.Zt_avg1[c] + s.w1 * zn;
    DU1 = p.DU_avg1[c];
    DV1 = p.DV_avg1[c];
    p.DU_avg2[c] = p.DU_avg2[c] + s.w2 * DUon;
    p.DV_avg2[c] = p.DV_avg2[c] + s.w2 * DVom;
  }
  p.Zt_avg1[c] = Zt;
}
__global__ void __launch_bounds__(S2_TX * S2_TY, LB_S2D) k_step2d(const __grid_constant__ KP p, const __grid_constant__ S2D s) {
  __shared__ double zt[S2_TY + 1][S2_TX + 1];
  const int tx = threadIdx.x, ty = threadIdx.y;
  const int i0 = blockIdx.x * S2_TX - IOFF, j0 = p.jstr + blockIdx.y * S2_TY;
  const int i = i0 + tx, j = j0 + ty;
  const ix2_t SX = p.SX;
  // ring-only launch: tiles that hold no ring point leave at once
  if (s.region == 2 && !(i0 <= p.istr + 1 || i0 + S2_TX - 1 >= p.iend - 1 || j0 <= p.jstr + 1 || j0 + S2_TY - 1 >= p.jend - 1)) return;
  // ---- phase 1: zeta_new (step2d_FB.F:151-216) on the tile, the row below it and the column to its left
  if (i >= p.istrU - 1 && i <= p.iend && j <= p.jend) zt[ty + 1][tx + 1] = s2d_zeta_new(p, s, CIDX(i, j));
  {
    const int t = ty * S2_TX + tx;
    if (t < S2_TX + 1) {
      const int ii = i0 - 1 + t, jj = j0 - 1;
      if (ii >= p.istrU - 1 && ii <= p.iend) zt[0][t] = s2d_zn_at(p, s, CIDX(ii, jj), jj);
    } else if (t < S2_TX + 1 + S2_TY) {
      const int r = t - (S2_TX + 1), ii = i0 - 1, jj = j0 + r;
      if (ii >= p.istrU - 1 && jj <= p.jend) zt[r + 1][0] = s2d_zeta_new(p, s, CIDX(ii, jj));
    }
  }
  __syncthreads();
  if (i < p.istr || i > p.iend || j > p.jend) return;
  if (s.region && (s2d_in_ring(p, i, j) != (s.region == 2))) return;
  // ---- phase 2: own point
  const ix2_t c = CIDX(i, j);
  const double zn = zt[ty + 1][tx + 1], znw = zt[ty + 1][tx], zns = zt[ty][tx + 1];
  double DU1, DV1, Zt;
  s2d_avg_out(p, s, i, j, c, zn, s2d_DUon(p, s, c), s2d_DVom(p, s, c), DU1, DV1, Zt);
  const double cffg = 0.5 * p.g;
  const bool sv = (j - 1 >= p.jstrV - 1);     // row j-1 carries a zeta_new
  const bool dov = (j >= p.jstrV);
  const bool dou = (i >= p.istrU);            // u(istr) sits on a closed western wall: u2dbc, not the momentum update
  const S2Pt zero = {0.0, 0.0, 0.0, 0.0};     // row below a southern wall: zero-initialised scratch in the reference
  const S2Pt a0 = s2d_ptz(p, s, c, zn, false), aw = dou ? s2d_ptz(p, s, c - 1, znw, false) : zero, as = sv ? s2d_ptz(p, s, c - SX, zns, false) : zero;
  double rubar = s2d_pgf(p, cffg, p.dn_u[c], c, 1, aw, a0);
  double rvbar = s2d_pgf(p, cffg, p.dm_v[c], c, SX, as, a0);
  double rufrc = p.rufrc[c], rvfrc = p.rvfrc[c];
  if (s.iif == 1) {
    rufrc = rufrc - rubar;
    rvfrc = rvfrc - rvbar;
    p.rufrc[c] = rufrc;
    p.rvfrc[c] = rvfrc;
    const S2Pt b0 = s2d_ptz(p, s, c, zn, true), bw = dou ? s2d_ptz(p, s, c - 1, znw, true) : zero, bs = sv ? s2d_ptz(p, s, c - SX, zns, true) : zero;
    rubar = rubar + s2d_pgf(p, cffg, p.dn_u[c], c, 1, bw, b0);
    rvbar = rvbar + s2d_pgf(p, cffg, p.dm_v[c], c, SX, bs, b0);
  }
  const double D0 = s.zs[c] + p.h[c], Dw = s.zs[c - 1] + p.h[c - 1], Ds = s.zs[c - SX] + p.h[c - SX];
  const double Dn0 = zn + p.h[c], Dnw = znw + p.h[c - 1], Dns = zns + p.h[c - SX];
  const double cff = 0.5 * p.dtfast, cff1 = 0.5 * s.w1;
  double un = 0.0;      // u2dbc_im.F on a closed western wall: ubar(istr)=0 (its DU_avg1 term is identically zero)
  if (dou) {
    const double DUnew = ((D0 + Dw) * s.us[c] + cff * (p.pm[c] + p.pm[c - 1]) * (p.pn[c] + p.pn[c - 1]) * (rubar + rufrc)) * p.umask[c];
    un = DUnew / (Dn0 + Dnw);
    double DUa = DU1 + cff1 * p.dn_u[c] * (DUnew);
    if (p.rivers) {        // step2d_FB.F:531-554: a river face carries the river's volume flux
      const double code = p.riv_uflx[c];
      if (riv_on(code)) {
        const int iriver = (int)lround(code / 10);
        const double river_flux = p.riv_tab[(iriver - 1) * RIVROW] * (code - 10 * iriver);
        un = river_flux * 2 / (p.dn_u[c] * (Dnw + Dn0));
        DUa = river_flux;
      }
    }
    p.DU_avg1[c] = DUa;
  } else if (s.iif == 1 && !p.bry2d) p.DU_avg1[c] = DU1;
  if (dou || !p.bry2d) s2d_store(p, s.un, i, j, c, un);
  if (dov) {
    const double DVnew = ((D0 + Ds) * s.vs[c] + cff * (p.pm[c] + p.pm[c - SX]) * (p.pn[c] + p.pn[c - SX]) * (rvbar + rvfrc)) * p.vmask[c];
    double vn = DVnew / (Dn0 + Dns);
    double DVa = DV1 + cff1 * p.dm_v[c] * (DVnew);
    if (p.rivers) {
      const double code = p.riv_vflx[c];
      if (riv_on(code)) {
        const int iriver = (int)lround(code / 10);
        const double river_flux = p.riv_tab[(iriver - 1) * RIVROW] * (code - 10 * iriver);
        vn = river_flux * 2 / (p.dm_v[c] * (Dns + Dn0));
        DVa = river_flux;
      }
    }
    s2d_store(p, s.vn, i, j, c, vn);
    p.DV_avg1[c] = DVa;
  } else if (!p.bry2d) {   // row jstr on a southern wall: v2dbc (no normal flow) and its DV_avg1 term (step2d_FB.F:449-529)
    const double vn = 0.0;
    s2d_store(p, s.vn, i, j, c, vn);
    p.DV_avg1[c] = DV1 + cff1 * (Dn0 + Dns) * (vn) * p.dm_v[c];
  }
  if (s.iif == s.nfast) s2d_store(p, s.zn, i, j, c, Zt);
  // ---- wall point beside this one (bry2d: every boundary point belongs to k_step2d_bry, obc.cuh)
  const bool sw = p.swall && j == p.jstr && !p.bry2d, nw = p.nwall && j == p.jend && !p.bry2d;
  if (sw | nw) {
#pragma unroll 1
    for (int side = 0; side < 2; ++side) {
      if (side == 0 ? !sw : !nw) continue;
      const int jw = side == 0 ? j - 1 : j + 1;
      const ix2_t w = side == 0 ? c - SX : c + SX;
      // zetabc.F:3: gradient condition; the neighbour in -i likewise (needed for D at the u-point)
      const double znb = side == 0 ? zns : zn * p.rmask[w];
      const double znbw = side == 0 ? zt[ty][tx] : znw * p.rmask[w - 1];
      double DU1w, DV1w, Ztw;
      // DVom below a southern wall is never filled by the reference (zero scratch)
      s2d_avg_out(p, s, i, jw, w, znb, s2d_DUon(p, s, w), side == 0 ? 0.0 : s2d_DVom(p, s, w), DU1w, DV1w, Ztw);
      const double unw = p.gamma2 * un * p.umask[w];          // u2dbc_im.F:150-152
      s2d_store(p, s.un, i, jw, w, unw);
      if (dou) p.DU_avg1[w] = DU1w + cff1 * ((p.h[w] + znb) + (p.h[w - 1] + znbw)) * (unw)*p.dn_u[w];
      else if (s.iif == 1) p.DU_avg1[w] = DU1w;     // on the western wall the term is identically zero (ubar=0)
      if (side == 1) {                                        // v2dbc_im.F: no normal flow through the northern wall
        const double vnw = 0.0;
        s2d_store(p, s.vn, i, jw, w, vnw);
        p.DV_avg1[w] = DV1w + cff1 * ((znb + p.h[w]) + (zn + p.h[c])) * (vnw)*p.dm_v[w];
      } else if (s.iif == 1) p.DV_avg1[w] = DV1w;
      if (s.iif == s.nfast) s2d_store(p, s.zn, i, jw, w, Ztw);
    }
  }
}
#include "step2d_tma.cuh"
#include "step2d_march.cuh"
#include "step2d_hyb.cuh"
#include "tracer_tma.cuh"

// zeta_new at any point of the rho-range incl. wall rows, wall columns and basin corners (zetabc.F:39-81)
__device__ __forceinline__ double s2d_zn_full(const KP& p, const S2D& s, int i, int j) {
  const int ii = (p.wwall && i < p.istr) ? p.istr : ((p.ewall && i > p.iend) ? p.iend : i);
  const int jj = (p.swall && j < p.jstr) ? p.jstr : ((p.nwall && j > p.jend) ? p.jend : j);
  const double zin = s2d_zeta_new(p, s, CIDX(ii, jj));
  if (ii == i && jj == j) return zin;
  if (ii != i && jj != j) return 0.5 * (zin * p.rmask[CIDX(ii, j)] + zin * p.rmask[CIDX(i, jj)]);   // corner: mean of its two edge neighbours
  return zin * p.rmask[CIDX(i, j)];
}
// Closed western/eastern walls of the barotropic substep, launched after k_step2d: one thread per wall-column point
// (x = row index, y = 0 west / 1 east).  zetabc on the column and the corners, the fast-time averages there,
// v2dbc slip columns and their DV_avg1 term (step2d_FB.F:259-303), u2dbc zero flow through the eastern wall.
__global__ void k_step2d_ew(const __grid_constant__ KP p, const __grid_constant__ S2D s) {
  const int west = blockIdx.y == 0;
  if (west ? !p.wwall : !p.ewall) return;
  const int j = p.jstrR + blockIdx.x * blockDim.x + threadIdx.x;
  if (j > p.jendR) return;
  const int iw = west ? p.istr - 1 : p.iend + 1, ia = west ? p.istr : p.iend;
  const ix2_t c = CIDX(iw, j), SX = p.SX;
  const double zn = s2d_zn_full(p, s, iw, j);
  // DUon at i=istr-1 is never filled by the reference (zero scratch), likewise DVom below a southern wall
  const double DUon = west ? 0.0 : s2d_DUon(p, s, c);
  const double DVom = (j >= p.jstrV - 1) ? s2d_DVom(p, s, c) : 0.0;
  double DU1, DV1, Zt;
  s2d_avg_out(p, s, iw, j, c, zn, DUon, DVom, DU1, DV1, Zt);
  if (s.iif == 1) p.DU_avg1[c] = DU1;
  const double cff1 = 0.5 * s.w1;
  if (j >= p.jstr) {     // v2dbc_im.F: gamma2 slip; v on the S/N wall rows has just been zeroed
    const bool zeroed = (p.swall && j == p.jstr) || (p.nwall && j == p.jend + 1);
    const double vsrc = zeroed ? 0.0 : s.vn[CIDX(ia, j)];
    const double vn = p.gamma2 * vsrc * p.vmask[c];
    s.vn[c] = vn;
    if (j >= p.jstrV && j <= p.jend) {
      const double Dn = p.h[c] + zn, Dns = p.h[c - SX] + s2d_zn_full(p, s, iw, j - 1);
      p.DV_avg1[c] = DV1 + cff1 * (Dn + Dns) * (vn)*p.dm_v[c];
    } else if (s.iif == 1) p.DV_avg1[c] = DV1;
  } else if (s.iif == 1) p.DV_avg1[c] = DV1;
  if (!west) {           // u2dbc_im.F: no flow through the eastern wall (u-point iend+1); slip rule on the wall rows gives gamma2*0*umask
    const bool row = (j < p.jstr) || (j > p.jend);
    s.un[c] = row ? p.gamma2 * 0.0 * p.umask[c] : 0.0;
  }
  if (s.iif == s.nfast) s.zn[c] = Zt;
}

// Lateral boundary conditions of the 3-D fields on closed walls, one launch per field: u3dbc_im.F:293-295 (u: zero on the
// W/E wall, gamma2 slip rows), v3dbc_im.F (v: zero on the S/N wall, slip columns), t3dbc_im.F:86-144 (gradient + the
// mask-weighted corner mean), omega.F:171-231 (We/Wi copies + corners).  Every boundary point is written by one thread as a
// pure function of interior values, so the result does not depend on the order in which the reference applies the edges
// (where one edge rule reads a value another edge rule has just set, that value is substituted: it is always 0).
// grid: x = position along the edge, y = side (0 S row, 1 N row, 2 W column, 3 E column), z = level.
enum { BC_U = 0, BC_V = 1, BC_T = 2, BC_W = 3 };
template <int KIND>
__global__ void k_bc3d(const __grid_constant__ KP p, double* __restrict__ a, double gamma2) {
  const int side = blockIdx.y, q = blockIdx.x * blockDim.x + threadIdx.x;
  const ix_t ko = (ix_t)blockIdx.z * p.SXY;
  if (side < 2) {
    if (side == 0 ? !p.swall : !p.nwall) return;
    const int i = p.istr - 1 + q;
    if (i > p.iend + 1) return;
    const int jw = side == 0 ? p.jstr - 1 : p.jend + 1, ja = side == 0 ? p.jstr : p.jend;
    const bool wc = i < p.istr, ec = i > p.iend;       // corner columns
    if (KIND == BC_U) {
      if (i < p.istr || i > p.iendR) return;
      const bool onwall = (p.wwall && i == p.istr) || (p.ewall && i == p.iend + 1);      // u there has just been zeroed
      const double src = onwall ? 0.0 : a[CIDX(i, ja) + ko];
      a[CIDX(i, jw) + ko] = gamma2 * src * p.umask[CIDX(i, jw)];
    } else if (KIND == BC_V) {
      if (wc || ec) return;
      a[CIDX(i, side == 0 ? p.jstr : p.jend + 1) + ko] = 0.0;
    } else if (KIND == BC_T) {
      if (!wc && !ec) a[CIDX(i, jw) + ko] = a[CIDX(i, ja) + ko] * p.rmask[CIDX(i, jw)];
      else if ((wc && p.wwall) || (ec && p.ewall)) {
        const int ii = wc ? p.istr : p.iend;
        const double tin = a[CIDX(ii, ja) + ko];
        const double ra = p.rmask[CIDX(ii, jw)], rb = p.rmask[CIDX(i, ja)];
        const double ta = tin * ra, tb = tin * rb;    // the two edge values next to the corner
        double cff = ra + rb;
        if (cff > 0.0) {
          cff = 1.0 / cff;
          a[CIDX(i, jw) + ko] = cff * (ra * ta + rb * tb);
        } else a[CIDX(i, jw) + ko] = 0.0;
      }
    } else {
      if (!wc && !ec) a[CIDX(i, jw) + ko] = a[CIDX(i, ja) + ko];
      else if ((wc && p.wwall) || (ec && p.ewall)) a[CIDX(i, jw) + ko] = a[CIDX(wc ? p.istr : p.iend, ja) + ko];
    }
  } else {
    if (side == 2 ? !p.wwall : !p.ewall) return;
    const int j = p.jstr + q;
    const int iw = side == 2 ? p.istr - 1 : p.iend + 1, ia = side == 2 ? p.istr : p.iend;
    if (KIND == BC_U) {
      if (j > p.jend) return;
      a[CIDX(side == 2 ? p.istr : p.iend + 1, j) + ko] = 0.0;
    } else if (KIND == BC_V) {
      if (j > p.jendR) return;
      const bool zeroed = (p.swall && j == p.jstr) || (p.nwall && j == p.jend + 1);       // v there has just been zeroed
      const double src = zeroed ? 0.0 : a[CIDX(ia, j) + ko];
      a[CIDX(iw, j) + ko] = gamma2 * src * p.vmask[CIDX(iw, j)];
    } else if (KIND == BC_T) {
      if (j > p.jend) return;
      a[CIDX(iw, j) + ko] = a[CIDX(ia, j) + ko] * p.rmask[CIDX(iw, j)];
    } else {
      if (j > p.jend) return;
      a[CIDX(iw, j) + ko] = a[CIDX(ia, j) + ko];
    }
  }
}

// ------------------------------------------------------------------------------------------------
#include "obc.cuh"

// step3d_uv2.F:18-786.  PASS 0 (:244-268): u <- (u/Hz_u - vertical-mean mismatch vs DU_avg1)*umask.
// PASS 1 (:521-577): ubar(knew)=DU_avg1/D, re-correct u, FlxU <- DELTA*FlxU+EPSIL*DC*(u(nstp)+u(nnew))
// minus its mismatch with DU_avg2.  Column kernels without scratch (re-read through L1/L2).
template <int DIR>
__global__ void k_uv2_pass0(const __grid_constant__ KP p, double* __restrict__ u) {
  THREAD_IJT(ixc_t, (DIR == 0 ? p.istrU : p.istr), p.iend, (DIR == 0 ? p.jstr : p.jstrV), p.jend)
  const ixc_t d = DIR == 0 ? 1 : SX;
  const int N = p.N;
  const double dn = DIR == 0 ? p.dn_u[c] : p.dm_v[c];
  const double msk = DIR == 0 ? p.umask[c] : p.vmask[c];
  const double Davg1 = DIR == 0 ? p.DU_avg1[c] : p.DV_avg1[c];
  double CF0 = 0.5 * (R3(p.Hz, c, N) + R3(p.Hz, c - d, N));
  double DC0 = R3(u, c, N);
  R3(u, c, N) = DC0 / CF0;
  for (int k = N - 1; k >= 1; --k) {
    const double cff = 0.5 * (R3(p.Hz, c, k) + R3(p.Hz, c - d, k));
    CF0 = CF0 + cff;
    const double uk = R3(u, c, k);
    DC0 = DC0 + uk;
    R3(u, c, k) = uk / cff;
  }
  DC0 = (DC0 * dn - Davg1) / (CF0 * dn);
  for (int k = 1; k <= N; ++k) R3(u, c, k) = (R3(u, c, k) - DC0) * msk;
}
template <int DIR>
__global__ void k_uv2_pass1(const __grid_constant__ KP p, double* __restrict__ u, const double* __restrict__ u_stp, double* __restrict__ bar) {
  THREAD_IJT(ixc_t, (DIR == 0 ? p.istr : p.istrR), p.iendR, (DIR == 0 ? p.jstrR : p.jstr), p.jendR)
  const ixc_t d = DIR == 0 ? 1 : SX;
  const int N = p.N;
  const double dn = DIR == 0 ? p.dn_u[c] : p.dm_v[c];
  const double msk = DIR == 0 ? p.umask[c] : p.vmask[c];
  const double Davg1 = DIR == 0 ? p.DU_avg1[c] : p.DV_avg1[c];
  const double Davg2 = DIR == 0 ? p.DU_avg2[c] : p.DV_avg2[c];
  double* Flx = DIR == 0 ? p.FlxU : p.FlxV;
  double DCN = 0.5 * (R3(p.Hz, c, N) + R3(p.Hz, c - d, N)) * dn;
  double DC0 = DCN;
  double FC0 = DCN * R3(u, c, N);
  for (int k = N - 1; k >= 1; --k) {
    const double DCk = 0.5 * (R3(p.Hz, c, k) + R3(p.Hz, c - d, k)) * dn;
    DC0 = DC0 + DCk;
    FC0 = FC0 + DCk * R3(u, c, k);
  }
  DC0 = 1.0 / DC0;
  bar[c] = DC0 * Davg1;
  FC0 = DC0 * (FC0 - Davg1);
  double CF0 = 0.0;
  for (int k = N; k >= 1; --k) {
    const double DCk = 0.5 * (R3(p.Hz, c, k) + R3(p.Hz, c - d, k)) * dn;
    const double un = (R3(u, c, k) - FC0) * msk;
    R3(u, c, k) = un;
    const double CFk = 0.28 * R3(Flx, c, k) + 0.36 * DCk * (R3(u_stp, c, k) + un);
    R3(Flx, c, k) = CFk;      // parked; corrected in the next loop
    CF0 = CF0 + CFk;
  }
  CF0 = DC0 * (CF0 - Davg2);
  for (int k = 1; k <= N; ++k) {
    const double DCk = 0.5 * (R3(p.Hz, c, k) + R3(p.Hz, c - d, k)) * dn;
    R3(Flx, c, k) = R3(Flx, c, k) - DCk * CF0;
  }
}

// ------------------------------------------------------------------------------------------------
// k_uv2_pass1 with the column parked in SHARED memory between its three sweeps (one warp-row per block, [k][lane]) and
// the operands of each sweep fetched one register chunk ahead of the arithmetic: every 3-D operand crosses HBM once
// (6 words per point instead of ~10.7: Hz and u are not re-read twice, Flx is not written twice), 1.0 -> 0.78 ms on
// B200.  The sums have short dependency chains, so the 6 resident warps per SM (shared memory holds 2(N+1) doubles per
// column) are enough here -- unlike the spline/Thomas column kernels, where the same design is 20-80 % SLOWER than the
// global scratch slab (profiles/r1_scratch_policy.txt, third experiment).  Arithmetic is expression-for-expression
// that of k_uv2_pass1.
#ifndef CP_K
#define CP_K 16
#endif
#define COLSM_PROLOGUE                                                                   \
  const int nth = blockDim.x * blockDim.y;                                               \
  double* __restrict__ sA = smem + threadIdx.y * blockDim.x + threadIdx.x;               \
  double* __restrict__ sB = sA + (p.N + 1) * nth;
#define SA(k) sA[(k)*nth]
#define SB(k) sB[(k)*nth]
// summing sweep: A(k) = DC(k) = Hz_u*dn, B(k) = u(k); returns sum DC, sum DC*u
template <int DIR>
__device__ __forceinline__ void colsm_sum_sweep(const KP& p, ixc_t c, ixc_t d, double dn, const double* __restrict__ u, double* __restrict__ sA,
                                                double* __restrict__ sB, int nth, double& DC0, double& FC0) {
  const int N = p.N;
  const double* __restrict__ gHz = p.Hz;
  double na[CP_K], nb[CP_K], nu[CP_K];
#pragma unroll
  for (int q = 0; q < CP_K; ++q)
    if (N - q >= 1) { na[q] = R3(gHz, c, N - q); nb[q] = R3(gHz, c - d, N - q); nu[q] = R3(u, c, N - q); }
  for (int k0 = N; k0 >= 1; k0 -= CP_K) {
    double ha[CP_K], hb[CP_K], uu[CP_K];
#pragma unroll
    for (int q = 0; q < CP_K; ++q) { ha[q] = na[q]; hb[q] = nb[q]; uu[q] = nu[q]; }
    const int kn = k0 - CP_K;
#pragma unroll
    for (int q = 0; q < CP_K; ++q)
      if (kn - q >= 1) { na[q] = R3(gHz, c, kn - q); nb[q] = R3(gHz, c - d, kn - q); nu[q] = R3(u, c, kn - q); }
#pragma unroll
    for (int q = 0; q < CP_K; ++q) {
      const int k = k0 - q;
      if (k >= 1) {
        const double DCk = 0.5 * (ha[q] + hb[q]) * dn;
        if (k == N) { DC0 = DCk; FC0 = DCk * uu[q]; }
        else { DC0 = DC0 + DCk; FC0 = FC0 + DCk * uu[q]; }
        SA(k) = DCk;
        SB(k) = uu[q];
      }
    }
  }
}
template <int DIR>
__global__ void __launch_bounds__(64, 3) k_uv2_pass1_sm(const __grid_constant__ KP p, double* __restrict__ u, const double* __restrict__ u_stp, double* __restrict__ bar) {
  extern __shared__ double smem[];
  THREAD_IJT(ixc_t, (DIR == 0 ? p.istr : p.istrR), p.iendR, (DIR == 0 ? p.jstrR : p.jstr), p.jendR)
  COLSM_PROLOGUE
  const ixc_t d = DIR == 0 ? 1 : SX;
  const int N = p.N;
  const double dn = DIR == 0 ? p.dn_u[c] : p.dm_v[c];
  const double msk = DIR == 0 ? p.umask[c] : p.vmask[c];
  const double Davg1 = DIR == 0 ? p.DU_avg1[c] : p.DV_avg1[c];
  const double Davg2 = DIR == 0 ? p.DU_avg2[c] : p.DV_avg2[c];
  double* __restrict__ Flx = DIR == 0 ? p.FlxU : p.FlxV;
  double DC0 = 0.0, FC0 = 0.0;
  colsm_sum_sweep<DIR>(p, c, d, dn, u, sA, sB, nth, DC0, FC0);
  DC0 = 1.0 / DC0;
  bar[c] = DC0 * Davg1;
  FC0 = DC0 * (FC0 - Davg1);
  double CF0 = 0.0;
  {
    double nf[CP_K], ns[CP_K];
#pragma unroll
    for (int q = 0; q < CP_K; ++q)
      if (N - q >= 1) { nf[q] = R3(Flx, c, N - q); ns[q] = R3(u_stp, c, N - q); }
    for (int k0 = N; k0 >= 1; k0 -= CP_K) {
      double fl[CP_K], us[CP_K];
#pragma unroll
      for (int q = 0; q < CP_K; ++q) { fl[q] = nf[q]; us[q] = ns[q]; }
      const int kn = k0 - CP_K;
#pragma unroll
      for (int q = 0; q < CP_K; ++q)
        if (kn - q >= 1) { nf[q] = R3(Flx, c, kn - q); ns[q] = R3(u_stp, c, kn - q); }
#pragma unroll
      for (int q = 0; q < CP_K; ++q) {
        const int k = k0 - q;
        if (k >= 1) {
          const double DCk = SA(k);
          const double un = (SB(k) - FC0) * msk;
          R3(u, c, k) = un;
          const double CFk = 0.28 * fl[q] + 0.36 * DCk * (us[q] + un);
          SB(k) = CFk;      // parked; corrected in the next loop
          CF0 = CF0 + CFk;
        }
      }
    }
  }
  CF0 = DC0 * (CF0 - Davg2);
  for (int k = 1; k <= N; ++k) R3(Flx, c, k) = SB(k) - SA(k) * CF0;
}

// step3d_uv2.F passes 0 and 1 (:244-268 and :521-577) in ONE kernel: the column never leaves shared memory between the
// reference's two column loops, so u(nnew) is read once and written once (6 words per point for the pair instead of ~11).
//   sweep A (down): q(k) = (Hz_u*u)(k)/Hz_u(k), sums -> vertical-mean mismatch of pass 0
//   sweep B (down): u0(k) = (q(k)-mismatch)*mask [= pass 0's result], DC(k) = Hz_u*dn, pass 1's two sums
//   sweep C (down): u(k) final, FlxU candidate parked, its sum;   sweep D (up): FlxU
// EDGE=false: the columns both passes cover (pass 0's range).  EDGE=true: the columns only pass 1 covers -- wall rows and wall
// columns, where the reference applies u3dbc/v3dbc between the passes (step3d_uv2.F:471-472; u3dbc_im.F:293-295 closed walls):
// the value there is gamma2*(pass-0 result of the neighbouring interior column)*mask, or zero on the wall itself, so the
// thread repeats pass 0 for that neighbour from the ORIGINAL u(nnew).  The EDGE launches therefore run BEFORE the interior
// launch (they read interior columns and write edge columns only; the interior launch reads and writes its own column only).
template <int DIR, bool EDGE>
__global__ void __launch_bounds__(64, 3) k_uv2_sm(const __grid_constant__ KP p, double* __restrict__ u, const double* __restrict__ u_stp, double* __restrict__ bar,
                                                   double gamma2, int i0, int i1, int j0, int j1) {
  extern __shared__ double smem[];
  THREAD_IJT(ixc_t, i0, i1, j0, j1)
  COLSM_PROLOGUE
  const ixc_t d = DIR == 0 ? 1 : SX;
  const int N = p.N;
  const double dn = DIR == 0 ? p.dn_u[c] : p.dm_v[c];
  const double msk = DIR == 0 ? p.umask[c] : p.vmask[c];
  const double Davg1 = DIR == 0 ? p.DU_avg1[c] : p.DV_avg1[c];
  const double Davg2 = DIR == 0 ? p.DU_avg2[c] : p.DV_avg2[c];
  double* __restrict__ Flx = DIR == 0 ? p.FlxU : p.FlxV;
  const double* __restrict__ gHz = p.Hz;
  double DC0 = 0.0, FC0 = 0.0;
  if (!EDGE) {
    // ---- sweep A: pass 0, first loop (step3d_uv2.F:244-256)
    double CFs = 0.0, DCs = 0.0;
    {
      double na[CP_K], nb[CP_K], nu[CP_K];
#pragma unroll
      for (int q = 0; q < CP_K; ++q)
        if (N - q >= 1) { na[q] = R3(gHz, c, N - q); nb[q] = R3(gHz, c - d, N - q); nu[q] = R3(u, c, N - q); }
      for (int k0 = N; k0 >= 1; k0 -= CP_K) {
        double ha[CP_K], hb[CP_K], uu[CP_K];
#pragma unroll
        for (int q = 0; q < CP_K; ++q) { ha[q] = na[q]; hb[q] = nb[q]; uu[q] = nu[q]; }
        const int kn = k0 - CP_K;
#pragma unroll
        for (int q = 0; q < CP_K; ++q)
          if (kn - q >= 1) { na[q] = R3(gHz, c, kn - q); nb[q] = R3(gHz, c - d, kn - q); nu[q] = R3(u, c, kn - q); }
#pragma unroll
        for (int q = 0; q < CP_K; ++q) {
          const int k = k0 - q;
          if (k >= 1) {
            const double cff = 0.5 * (ha[q] + hb[q]);
            if (k == N) { CFs = cff; DCs = uu[q]; }
            else { CFs = CFs + cff; DCs = DCs + uu[q]; }
            SA(k) = cff;
            SB(k) = uu[q] / cff;
          }
        }
      }
    }
    const double mis = (DCs * dn - Davg1) / (CFs * dn);
    // ---- sweep B: pass 0's second loop (:258-268) feeding pass 1's first loop (:521-540)
    for (int k = N; k >= 1; --k) {
      const double u0 = (SB(k) - mis) * msk;
      const double DCk = SA(k) * dn;
      if (k == N) { DC0 = DCk; FC0 = DCk * u0; }
      else { DC0 = DC0 + DCk; FC0 = FC0 + DCk * u0; }
      SA(k) = DCk;
      SB(k) = u0;
    }
  } else {
    // the value u3dbc/v3dbc leaves in this column: gamma2*src*mask with src the pass-0 result of the interior neighbour
    // (0 where that neighbour sits on a wall and has just been zeroed), or 0 on the wall line itself
    bool slip;      // true: slip row/column beside the wall; false: the wall line (no normal flow)
    bool srcz;      // slip point whose source has just been zeroed
    ixc_t cs = c;
    if (DIR == 0) {
      slip = (j < p.jstr || j > p.jend);
      srcz = (p.wwall && i == p.istr) || (p.ewall && i == p.iend + 1);
      if (slip) cs = CIDX(i, j < p.jstr ? p.jstr : p.jend);
    } else {
      slip = (i < p.istr || i > p.iend);
      srcz = (p.swall && j == p.jstr) || (p.nwall && j == p.jend + 1);
      if (slip) cs = CIDX(i < p.istr ? p.istr : p.iend, j);
    }
    if (slip && !srcz) {
      const double dns = DIR == 0 ? p.dn_u[cs] : p.dm_v[cs];
      const double msks = DIR == 0 ? p.umask[cs] : p.vmask[cs];
      const double Davg1s = DIR == 0 ? p.DU_avg1[cs] : p.DV_avg1[cs];
      double CFs = 0.0, DCs = 0.0;
      for (int k = N; k >= 1; --k) {
        const double cff = 0.5 * (R3(gHz, cs, k) + R3(gHz, cs - d, k));
        const double uk = R3(u, cs, k);
        if (k == N) { CFs = cff; DCs = uk; }
        else { CFs = CFs + cff; DCs = DCs + uk; }
        SB(k) = uk / cff;
      }
      const double mis = (DCs * dns - Davg1s) / (CFs * dns);
      for (int k = 1; k <= N; ++k) SB(k) = gamma2 * ((SB(k) - mis) * msks) * msk;
    } else {
      const double v0 = slip ? gamma2 * 0.0 * msk : 0.0;
      for (int k = 1; k <= N; ++k) SB(k) = v0;
    }
    for (int k = N; k >= 1; --k) {
      const double DCk = 0.5 * (R3(gHz, c, k) + R3(gHz, c - d, k)) * dn;
      if (k == N) { DC0 = DCk; FC0 = DCk * SB(k); }
      else { DC0 = DC0 + DCk; FC0 = FC0 + DCk * SB(k); }
      SA(k) = DCk;
    }
  }
  DC0 = 1.0 / DC0;
  bar[c] = DC0 * Davg1;
  FC0 = DC0 * (FC0 - Davg1);
  double CF0 = 0.0;
  {
    double nf[CP_K], ns[CP_K];
#pragma unroll
    for (int q = 0; q < CP_K; ++q)
      if (N - q >= 1) { nf[q] = R3(Flx, c, N - q); ns[q] = R3(u_stp, c, N - q); }
    for (int k0 = N; k0 >= 1; k0 -= CP_K) {
      double fl[CP_K], us[CP_K];
#pragma unroll
      for (int q = 0; q < CP_K; ++q) { fl[q] = nf[q]; us[q] = ns[q]; }
      const int kn = k0 - CP_K;
#pragma unroll
      for (int q = 0; q < CP_K; ++q)
        if (kn - q >= 1) { nf[q] = R3(Flx, c, kn - q); ns[q] = R3(u_stp, c, kn - q); }
#pragma unroll
      for (int q = 0; q < CP_K; ++q) {
        const int k = k0 - q;
        if (k >= 1) {
          const double DCk = SA(k);
          const double un = (SB(k) - FC0) * msk;
          R3(u, c, k) = un;
          const double CFk = 0.28 * fl[q] + 0.36 * DCk * (us[q] + un);
          SB(k) = CFk;      // parked; corrected in the next loop
          CF0 = CF0 + CFk;
        }
      }
    }
  }
  CF0 = DC0 * (CF0 - Davg2);
  for (int k = 1; k <= N; ++k) R3(Flx, c, k) = SB(k) - SA(k) * CF0;
}

#undef SA
#undef SB

// ------------------------------------------------------------------------------------------------
// LMD/KPP vertical mixing closure: lmd_swr_frac.F, lmd_vmix.F:31-433 (LMD_RIMIX), lmd_kpp.F (INT_AT_RHO_POINTS,
// SMOOTH_HBL, LMD_BKPP, LMD_NONLOCAL).  Kernels follow the reference's phases; Kv,Kt,Ks live in Akv,Akt(:,:,:,1:2) themselves.
__global__ void k_swr_frac(const __grid_constant__ KP p) {
  THREAD_IJ(p.istr, p.iend, p.jstr, p.jend)
  const double attn1 = -1. / 0.35, attn2 = -1. / 23.0;
  double swdk1 = 0.58, swdk2 = 1. - swdk1;
  RW(p.swr_frac, c, p.N) = 1.;
  for (int k = p.N; k >= 1; --k) {
    const double xi1 = attn1 * R3(p.Hz, c, k);
    if (xi1 > -20.) swdk1 = swdk1 * exp(xi1); else swdk1 = 0.;
    const double xi2 = attn2 * R3(p.Hz, c, k);
    if (xi2 > -20.) swdk2 = swdk2 * exp(xi2); else swdk2 = 0.;
    RW(p.swr_frac, c, k - 1) = swdk1 + swdk2;
  }
}
// Richardson number (divided by its critical value) on levels 1..N-1 over [imin:imax]x[jmin:jmax] (blockIdx.z = k-1)
__global__ void k_lmd_rig(const __grid_constant__ KP p, const double* __restrict__ u, const double* __restrict__ v, double* __restrict__ Rig, int imin, int imax,
                          int jmin, int jmax) {
  THREAD_IJ(imin, imax, jmin, jmax)
  const int k = blockIdx.z + 1;
  const double cff = 0.5 / (R3(p.z_r, c, k + 1) - R3(p.z_r, c, k));
  const double dudz = cff * (R3(u, c, k + 1) - R3(u, c, k) + R3(u, c + 1, k + 1) - R3(u, c + 1, k));
  const double dvdz = cff * (R3(v, c, k + 1) - R3(v, c, k) + R3(v, c + SX, k + 1) - R3(v, c + SX, k));
  RW(Rig, c, k) = RW(p.bvf, c, k) / (0.7 * DMAX(dudz * dudz + dvdz * dvdz, 1.E-10));
}
// copies of the first/last computed line across a closed edge, incl. the corners (lmd_vmix.F:159-200, lmd_kpp_smooth_hbl.h)
__global__ void k_lmd_edges(const __grid_constant__ KP p, double* __restrict__ a, int imin, int imax, int jmin, int jmax) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  double* pl = a + (ix_t)blockIdx.z * p.SXY;
  const int side = blockIdx.y;     // 0 W, 1 E, 2 S, 3 N, 4 corners
  if (side == 0 && p.wwall) { const int j = jmin + t; if (j <= jmax) pl[CIDX(p.istr - 1, j)] = pl[CIDX(p.istr, j)]; }
  if (side == 1 && p.ewall) { const int j = jmin + t; if (j <= jmax) pl[CIDX(p.iend + 1, j)] = pl[CIDX(p.iend, j)]; }
  if (side == 2 && p.swall) { const int i = imin + t; if (i <= imax) pl[CIDX(i, p.jstr - 1)] = pl[CIDX(i, p.jstr)]; }
  if (side == 3 && p.nwall) { const int i = imin + t; if (i <= imax) pl[CIDX(i, p.jend + 1)] = pl[CIDX(i, p.jend)]; }
  if (side == 4 && t == 0) {
    if (p.wwall && p.swall) pl[CIDX(p.istr - 1, p.jstr - 1)] = pl[CIDX(p.istr, p.jstr)];
    if (p.wwall && p.nwall) pl[CIDX(p.istr - 1, p.jend + 1)] = pl[CIDX(p.istr, p.jend)];
    if (p.ewall && p.swall) pl[CIDX(p.iend + 1, p.jstr - 1)] = pl[CIDX(p.iend, p.jstr)];
    if (p.ewall && p.nwall) pl[CIDX(p.iend + 1, p.jend + 1)] = pl[CIDX(p.iend, p.jend)];
  }
}
// the masked isotropic smoothing (lmd_vmix.F:201-246): out-of-place, every intermediate of the reference (FX, FE1, FE) is
// re-evaluated in registers from the 3x3 neighbourhood with the same expressions
__global__ void k_lmd_smooth(const __grid_constant__ KP p, const double* __restrict__ a, double* __restrict__ out, int mask_result) {
  THREAD_IJ(p.istr, p.iend, p.jstr, p.jend)
  const double* __restrict__ q = a + (ix_t)blockIdx.z * p.SXY;
  const double cff = 1. / 12., cff1 = 3. / 16.;
  auto fx0 = [&](ix_t x) { return (q[x] - q[x - 1]) * p.umask[x]; };         // FX before the transversal terms
  auto fe1 = [&](ix_t x) { return (q[x] - q[x - SX]) * p.vmask[x]; };
  auto fe = [&](ix_t x) { return fe1(x) + cff * (fx0(x + 1) + fx0(x - SX) - fx0(x) - fx0(x + 1 - SX)); };
  auto fx = [&](ix_t x) { return fx0(x) + cff * (fe1(x + SX) + fe1(x - 1) - fe1(x) - fe1(x - 1 + SX)); };
  double r = q[c] + cff1 * (fx(c + 1) - fx(c) + fe(c + SX) - fe(c));
  if (mask_result) r = r * p.rmask[c];
  out[c + (ix_t)blockIdx.z * p.SXY] = r;
}
// interior mixing from the smoothed Rig, bottom taper, end values and the ascending in-place 1-2-1 filter (lmd_vmix.F:253-407)
__global__ void k_lmd_vmix_col(const __grid_constant__ KP p, const double* __restrict__ Rig, double Akv_bak, double Akt_bak0, double Akt_bak1, int convec) {
  THREAD_IJ(p.istr, p.iend, p.jstr, p.jend)
  const int N = p.N;
  const double nu0m = 1.e-2, nu0s = 1.e-2, nuwm = 1.0e-4, nuws = 0.1e-4, Lturb = 10., pi = 3.14159265358979323;
  double* Kv = p.Akv; double* Kt = p.Akt[0]; double* Ks = p.Akt[1];
  const double zw0 = RW(p.z_w, c, 0);
  for (int k = 1; k <= N - 1; ++k) {
    const double cff = DMIN(1., DMAX(0., RW(Rig, c, k)));
    double nu_sx = 1. - cff * cff;
    nu_sx = nu_sx * nu_sx * nu_sx;
    double kv = nuwm + nu0m * nu_sx, kt = nuws + nu0s * nu_sx;
    if (convec && RW(Rig, c, k) < 0.) { kv = kv + 0.1; kt = kt + 0.1; }      // LMD_CONVEC (lmd_vmix.F:269-274): nu0c
    double ks = kt;
    const double dist = RW(p.z_w, c, k) - zw0;
    if (dist < Lturb) {
      const double mult = sin(0.5 * pi * (RW(p.z_w, c, k) - zw0) / Lturb);
      kv = kv * mult; kt = kt * mult; ks = ks * mult;
    }
    RW(Kv, c, k) = kv; RW(Kt, c, k) = kt; RW(Ks, c, k) = ks;
  }
  RW(Kv, c, N) = RW(Kv, c, N - 1) + Akv_bak; RW(Ks, c, N) = RW(Ks, c, N - 1) + Akt_bak1; RW(Kt, c, N) = RW(Kt, c, N - 1) + Akt_bak0;
  RW(Kv, c, 0) = RW(Kv, c, 1) + Akv_bak; RW(Ks, c, 0) = RW(Ks, c, 1) + Akt_bak1; RW(Kt, c, 0) = RW(Kt, c, 1) + Akt_bak0;
  for (int k = 1; k <= N - 1; ++k) {
    RW(Kv, c, k) = 0.5 * RW(Kv, c, k) + 0.25 * RW(Kv, c, k - 1) + 0.25 * RW(Kv, c, k + 1) + Akv_bak;
    RW(Kt, c, k) = 0.5 * RW(Kt, c, k) + 0.25 * RW(Kt, c, k - 1) + 0.25 * RW(Kt, c, k + 1) + Akt_bak0;
    RW(Ks, c, k) = 0.5 * RW(Ks, c, k) + 0.25 * RW(Ks, c, k - 1) + 0.25 * RW(Ks, c, k + 1) + Akt_bak1;
  }
}
// lmd_kpp.F first j-loop: surface forcing, the integral bulk-Richardson functional FC(0:N) (kept in FCs), hbl and bbl
__global__ void k_lmd_kpp1(const __grid_constant__ KP p, const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ alpha,
                           const double* __restrict__ beta, const double* __restrict__ stflxT, const double* __restrict__ stflxS, double* __restrict__ FCs,
                           int imin, int imax, int jmin, int jmax) {
  THREAD_IJ(imin, imax, jmin, jmax)
  const int N = p.N;
  const double Ricr = 0.15, Ri_inv = 1. / Ricr, epssfc = 0.1, betaT = -0.2, Cv = 1.8, C_Ek = 258., zeta_s = -1.0, a_s = -28.86, c_s = 98.96, r2 = 0.5, r3 = 1. / 3.,
               vonKar = 0.41;
  const double Vtc = Cv * sqrt(-betaT / (c_s * epssfc)) / (Ricr * (vonKar * vonKar));
  const double g = p.g;
  const double Bo = g * (alpha[c] * (stflxT[c] - p.srflx[c]) - beta[c] * (stflxS ? stflxS[c] : 0.0));
  const double Bosol = g * alpha[c] * p.srflx[c];
  const double su0 = p.sustr[c], su1 = p.sustr[c + 1], sv0 = p.svstr[c], sv1 = p.svstr[c + SX];
  const double us = sqrt(sqrt(0.333333333333 * (su0 * su0 + su1 * su1 + su0 * su1 + sv0 * sv0 + sv1 * sv1 + sv0 * sv1)));   // lmd_kpp.F:176-180 (no BULK_FRC)
  p.k_Bo[c] = Bo; p.k_Bosol[c] = Bosol; p.k_ustar[c] = us;
  const double hbl0 = p.hbls[c], bbl0 = p.hbbl[c];
  const double zwN = RW(p.z_w, c, N), zw0 = RW(p.z_w, c, 0), ff = p.f[c];
  const double eh = epssfc * hbl0, eb = epssfc * bbl0;
  const double rm = p.rmask[c], ustar3 = us * us * us;
  // Cr(k) = FC(k) + Vtsq(k) is formed in the same down-sweep that integrates FC (the reference's separate k=N..1 loop reads
  // nothing that the sweep has not produced yet); kbls = first level from the top with Cr < 0
  int kbls = 0;
  double Cr_up = 0., Cr_hit = 0., Cr_hit_up = 0.;
  auto crit = [&](int k, double FCk) {
    const double w1 = sqrt(RW(p.swr_frac, c, k) * RW(p.swr_frac, c, k - 1));
    double zscale = zwN - R3(p.z_r, c, k);
    const double Bfsfc = Bo + Bosol * (1. - w1);
    zscale = DMIN(zscale, hbl0 * epssfc);
    zscale = zscale * rm;
    const double zetahat = vonKar * zscale * Bfsfc;
    double ws;
    if (zetahat >= 0.) ws = vonKar * us * ustar3 / DMAX(ustar3 + 5. * zetahat, 1.E-20);
    else if (zetahat > zeta_s * ustar3) ws = vonKar * pow((ustar3 - 16. * zetahat) / us, r2);
    else ws = vonKar * pow(a_s * ustar3 - c_s * zetahat, r3);
    const double Vtsq = 1.8 * Vtc * ws * sqrt(DMAX(1.e-5, RW(p.bvf, c, k - 1)));
    const double Cr = FCk + Vtsq;
    if (kbls == 0 && Cr < 0.) { kbls = k; Cr_hit = Cr; Cr_hit_up = Cr_up; }
    Cr_up = Cr;
  };
  double FC = 0.;
  RW(FCs, c, N) = 0.;
  crit(N, 0.);
  for (int k = N - 1; k >= 1; --k) {
    const double du = zwN - RW(p.z_w, c, k), dd = RW(p.z_w, c, k) - zw0;
    const double cff_up = du * du, cff_dn = dd * dd;
    const double Kern = cff_up * cff_dn / ((cff_up + eh * eh) * (cff_dn + eb * eb));
    const double su = R3(u, c, k + 1) + R3(u, c + 1, k + 1) - R3(u, c, k) - R3(u, c + 1, k);
    const double sv = R3(v, c, k + 1) + R3(v, c + SX, k + 1) - R3(v, c, k) - R3(v, c + SX, k);
    const double hh = R3(p.Hz, c, k) + R3(p.Hz, c, k + 1);
    FC = FC + Kern * (0.5 * (su * su + sv * sv) / hh - 0.5 * hh * (Ri_inv * RW(p.bvf, c, k) + C_Ek * ff * ff));
    RW(FCs, c, k) = FC;
    crit(k, FC);
  }
  {
    const double z_bl = zw0 + 0.25 * R3(p.Hz, c, 1);
    const double du = zwN - z_bl, dd = z_bl - zw0;
    const double cff_up = du * du, cff_dn = dd * dd;
    const double Kern = cff_up * cff_dn / ((cff_up + eh * eh) * (cff_dn + eb * eb));
    const double su = R3(u, c, 1) + R3(u, c + 1, 1), sv = R3(v, c, 1) + R3(v, c + SX, 1);
    FC = FC + Kern * (0.5 * (su * su + sv * sv) / R3(p.Hz, c, 1) - 0.5 * R3(p.Hz, c, 1) * (Ri_inv * RW(p.bvf, c, 1) + C_Ek * ff * ff));
  }
  const double FC0 = FC;
  double hbl;
  if (kbls > 0) {
    if (kbls == N) hbl = zwN - R3(p.z_r, c, N);
    else hbl = zwN - (R3(p.z_r, c, kbls) * Cr_hit_up - R3(p.z_r, c, kbls + 1) * Cr_hit) / (Cr_hit_up - Cr_hit);
  } else hbl = zwN - zw0;
  p.k_hbl[c] = hbl * rm;
  // bottom boundary layer: first level from the bottom with FC(k) - FC(0) > 0 (the search stops there)
  int kbbl = 0;
  double Cr_dn = 0., Cb = 0., Cb_dn = 0.;
  for (int k = 1; k <= N && kbbl == 0; ++k) {
    const double Cr = RW(FCs, c, k) - FC0;
    if (Cr > 0.) { kbbl = k; Cb = Cr; Cb_dn = Cr_dn; }
    Cr_dn = Cr;
  }
  double bbl;
  if (kbbl > 0) {
    if (kbbl == 1) bbl = R3(p.z_r, c, 1) - zw0;
    else bbl = (R3(p.z_r, c, kbbl - 1) * Cb - R3(p.z_r, c, kbbl) * Cb_dn) / (Cb - Cb_dn) - zw0;
  } else bbl = zwN - zw0;
  p.k_bbl[c] = bbl * rm;
}
// lmd_kpp.F second j-loop: time filter of hbl/bbl, KPP shape functions blended into Kv,Kt,Ks (= Akv,Akt), ghat, hbls/hbbl
__global__ void k_lmd_kpp2(const __grid_constant__ KP p, const double* __restrict__ hbl_s, const double* __restrict__ bbl_s, const double* __restrict__ un,
                           const double* __restrict__ vn, int first) {
  THREAD_IJ(p.istr, p.iend, p.jstr, p.jend)
  const int N = p.N;
  const double epssfc = 0.1, Cstar = 10., zeta_m = -0.2, a_m = 1.257, c_m = 8.360, zeta_s = -1.0, a_s = -28.86, c_s = 98.96, r2 = 0.5, r3 = 1. / 3., r4 = 0.25,
               EPS = 1.E-20, vonKar = 0.41;
  const double Cg = Cstar * vonKar * pow(c_s * vonKar * epssfc, 1. / 3.);
  double* Kv = p.Akv; double* Kt = p.Akt[0]; double* Ks = p.Akt[1];
  double hbl = hbl_s[c], bbl = bbl_s[c];
  if (!first) { hbl = 0.5 * (hbl + p.hbls[c]); bbl = 0.5 * (bbl + p.hbbl[c]); }
  const double zwN = RW(p.z_w, c, N), zw0 = RW(p.z_w, c, 0), rm = p.rmask[c];
  const double Bo = p.k_Bo[c], Bosol = p.k_Bosol[c], us = p.k_ustar[c], ustar3 = us * us * us;
  int kbls = N;
  for (int k = N - 1; k >= 1; --k)
    if (RW(p.z_w, c, k) > zwN - hbl) kbls = k;
  double Bfsfc;
  {
    const int k = kbls;
    const double z_bl = zwN - hbl;
    if (RW(p.swr_frac, c, k - 1) > 0.)
      Bfsfc = Bo + Bosol * (1. - RW(p.swr_frac, c, k - 1) * RW(p.swr_frac, c, k) * (RW(p.z_w, c, k) - RW(p.z_w, c, k - 1)) /
                                     (RW(p.swr_frac, c, k) * (RW(p.z_w, c, k) - z_bl) + RW(p.swr_frac, c, k - 1) * (z_bl - RW(p.z_w, c, k - 1))));
    else Bfsfc = Bo + Bosol;
  }
  for (int k = N; k >= 0; --k) {
    double zscale = zwN - RW(p.z_w, c, k);
    zscale = DMIN(zscale, hbl * epssfc);
    zscale = zscale * rm;
    const double zetahat = vonKar * zscale * Bfsfc;
    double wm, ws;
    if (zetahat >= 0.) {
      wm = vonKar * us * ustar3 / DMAX(ustar3 + 5. * zetahat, 1.E-20);
      ws = wm;
    } else {
      if (zetahat > zeta_m * ustar3) wm = vonKar * pow(us * (ustar3 - 16. * zetahat), r4);
      else wm = vonKar * pow(a_m * ustar3 - c_m * zetahat, r3);
      if (zetahat > zeta_s * ustar3) ws = vonKar * pow((ustar3 - 16. * zetahat) / us, r2);
      else ws = vonKar * pow(a_s * ustar3 - c_s * zetahat, r3);
    }
    const double ssgm = (zwN - RW(p.z_w, c, k)) / DMAX(hbl, EPS);
    double gh = 0.;
    if (ssgm < 1.) {
      double cff;
      if (ssgm < 0.07) cff = 0.5 * ((ssgm - 0.07) * (ssgm - 0.07)) / 0.07;
      else cff = 0.;
      cff = cff + ssgm * ((1. - ssgm) * (1. - ssgm));
      const double amp = ssgm * ssgm;
      const double akv = amp * RW(Kv, c, k), akt = amp * RW(Kt, c, k), aks = amp * RW(Ks, c, k);
      const double bm = wm * hbl * cff, bs = ws * hbl * cff;
      RW(Kv, c, k) = sqrt(akv * akv + bm * bm);
      RW(Kt, c, k) = sqrt(akt * akt + bs * bs);
      RW(Ks, c, k) = sqrt(aks * aks + bs * bs);
      if (Bfsfc < 0.) gh = -Cg * ssgm * ((1. - ssgm) * (1. - ssgm));
    }
    RW(p.ghat, c, k) = gh;
  }
  {
    const double u0 = R3(un, c, 1), u1 = R3(un, c + 1, 1), v0 = R3(vn, c, 1), v1 = R3(vn, c + SX, 1);
    const double wmb = vonKar * vonKar * sqrt(0.333333333333 * (u0 * u0 + u1 * u1 + u0 * u1 + v0 * v0 + v1 * v1 + v0 * v1)) / log(1. + 0.5 * R3(p.Hz, c, 1) / p.Zob);
    const double wsb = wmb;
    for (int k = 0; k <= N; ++k) {
      const double sgmb = (RW(p.z_w, c, k) - zw0 + p.Zob) / (bbl + p.Zob);
      if (sgmb < 1.) {
        const double cff1 = sgmb * ((1. - sgmb) * (1. - sgmb));
        const double bm = wmb * bbl * cff1, bs = wsb * bbl * cff1;
        RW(Kv, c, k) = sqrt(RW(Kv, c, k) * RW(Kv, c, k) + bm * bm);
        RW(Kt, c, k) = sqrt(RW(Kt, c, k) * RW(Kt, c, k) + bs * bs);
        RW(Ks, c, k) = sqrt(RW(Ks, c, k) * RW(Ks, c, k) + bs * bs);
      }
    }
  }
  if (!(rm > 0.5))
    for (int k = 0; k <= N; ++k) { RW(Kv, c, k) = 0.; RW(Kt, c, k) = 0.; RW(Ks, c, k) = 0.; }
  p.hbls[c] = hbl;
  p.hbbl[c] = bbl;
}

// ------------------------------------------------------------------------------------------------
// diag.F:50-658: per-column kinetic energy, barotropic KE, free-surface volume and max advective
// Courant number.  Stage 1: vertical means ub,vb over [istr:iend+1]; stage 2: per-block partials.
__global__ void k_diag_bar(const __grid_constant__ KP p, const double* __restrict__ u, const double* __restrict__ v, double* __restrict__ ub, double* __restrict__ vb) {
  THREAD_IJ(p.istr, p.iend + 1, p.jstr, p.jend + 1)
  const int N = p.N;
  double a = (R3(p.Hz, c, N) + R3(p.Hz, c - 1, N)) * R3(u, c, N);
  double b = (R3(p.Hz, c, N) + R3(p.Hz, c - SX, N)) * R3(v, c, N);
  for (int k = N - 1; k >= 2; --k) {
    a = a + (R3(p.Hz, c, k) + R3(p.Hz, c - 1, k)) * R3(u, c, k);
    b = b + (R3(p.Hz, c, k) + R3(p.Hz, c - SX, k)) * R3(v, c, k);
  }
  ub[c] = (a + (R3(p.Hz, c, 1) + R3(p.Hz, c - 1, 1)) * R3(u, c, 1)) / (RW(p.z_w, c, N) + RW(p.z_w, c - 1, N) - RW(p.z_w, c, 0) - RW(p.z_w, c - 1, 0));
  vb[c] = (b + (R3(p.Hz, c, 1) + R3(p.Hz, c - SX, 1)) * R3(v, c, 1)) / (RW(p.z_w, c, N) + RW(p.z_w, c - SX, N) - RW(p.z_w, c, 0) - RW(p.z_w, c - SX, 0));
}
// one thread per column writes its 5 values; k_diag_rows / k_diag_final reduce them on the device in a fixed order
__global__ void k_diag_col(const __grid_constant__ KP p, const double* __restrict__ u, const double* __restrict__ v, const double* __restrict__ ub,
                           const double* __restrict__ vb, double* __restrict__ o_ke, double* __restrict__ o_ke2b, double* __restrict__ o_vol,
                           double* __restrict__ o_cx, double* __restrict__ o_cw) {
  THREAD_IJ(p.istr, p.iend, p.jstr, p.jend)
  const int N = p.N;
  double v2 = 0.5 * (ub[c] * ub[c] + ub[c + 1] * ub[c + 1] + vb[c] * vb[c] + vb[c + SX] * vb[c + SX]);
  const double ke2b = 0.5 * (RW(p.z_w, c, N) - RW(p.z_w, c, 0)) * v2;
  double ke = 0.0, cxm = 0.0, cwm = 0.0;
  for (int k = N; k >= 1; --k) {
    const double u0 = R3(u, c, k), u1 = R3(u, c + 1, k), v0 = R3(v, c, k), v1 = R3(v, c + SX, k);
    v2 = 0.5 * (u0 * u0 + u1 * u1 + v0 * v0 + v1 * v1);
    const double Hz = R3(p.Hz, c, k);
    const double ciV = p.dt * p.rmask[c] * p.pm[c] * p.pn[c] / Hz;
    const double cw = ciV * (DMAX(RW(p.We, c, k) + RW(p.Wi, c, k), 0.0) - DMIN(RW(p.We, c, k - 1) + RW(p.Wi, c, k - 1), 0.0));
    const double cx = cw + ciV * (DMAX(R3(p.FlxU, c + 1, k), 0.0) - DMIN(R3(p.FlxU, c, k), 0.0) + DMAX(R3(p.FlxV, c + SX, k), 0.0) - DMIN(R3(p.FlxV, c, k), 0.0));
    if (cx > cxm) { cxm = cx; cwm = cw; }
    ke = ke + 0.5 * v2 * Hz;
  }
  const double dA = p.rmask[c] / (p.pm[c] * p.pn[c]);
  o_vol[c] = dA * RW(p.z_w, c, N);
  o_ke[c] = dA * ke;
  o_ke2b[c] = dA * ke2b;
  o_cx[c] = cxm;
  o_cw[c] = cwm;
}

// diag.F:256-325: the global sums, on the device.  The summation tree is fixed -- halving down to runs of at most 8 terms, the
// same tree whatever the launch geometry -- so the diagnostics do not depend on the number of threads: one thread per row j
// sums its row of the three per-column planes with that tree, accumulates the row's area and volume left to right and finds the
// row's first maximum of the advective Courant number; k_diag_final then combines the rows with the same tree.  Only the 7
// results (56 bytes) travel to the host.
__device__ double d_pairsum(const double* __restrict__ a, int n) {
  if (n <= 8) { double s = 0; for (int i = 0; i < n; ++i) s += a[i]; return s; }
  return d_pairsum(a, n / 2) + d_pairsum(a + n / 2, n - n / 2);
}
// rows: [7][Mm] = ke, ke2b, vol(zeta part), area, volume(h part), max cx, its cw
__global__ void k_diag_rows(const __grid_constant__ KP p, const double* __restrict__ o_ke, const double* __restrict__ o_ke2b, const double* __restrict__ o_vol,
                            const double* __restrict__ o_cx, const double* __restrict__ o_cw, double* __restrict__ rows) {
  const int j = p.jstr + blockIdx.x * blockDim.x + threadIdx.x;
  if (j > p.jend) return;
  const int Lm = p.Lm, Mm = p.Mm;
  const ix_t c0 = CIDX(1, j);
  rows[0 * Mm + j - 1] = d_pairsum(o_ke + c0, Lm);
  rows[1 * Mm + j - 1] = d_pairsum(o_ke2b + c0, Lm);
  rows[2 * Mm + j - 1] = d_pairsum(o_vol + c0, Lm);
  double ar = 0, vr = 0, cx = 0.0, cw = 0.0;
  for (int i = 0; i < Lm; ++i) {
    const ix_t c = c0 + i;
    const double dA = p.rmask[c] / (p.pm[c] * p.pn[c]);
    ar += dA; vr += dA * p.h[c];
    if (o_cx[c] > cx) { cx = o_cx[c]; cw = o_cw[c]; }
  }
  rows[3 * Mm + j - 1] = ar; rows[4 * Mm + j - 1] = vr;
  rows[5 * Mm + j - 1] = cx; rows[6 * Mm + j - 1] = cw;
}
__global__ void k_diag_final(const double* __restrict__ rows, int Mm, double* __restrict__ red) {
  const int q = threadIdx.x;
  if (q < 5) red[q] = d_pairsum(rows + q * Mm, Mm);
  if (q == 5) {
    double cx = 0.0, cw = 0.0;
    for (int j = 0; j < Mm; ++j)
      if (rows[5 * Mm + j] > cx) { cx = rows[5 * Mm + j]; cw = rows[6 * Mm + j]; }
    red[5] = cx; red[6] = cw;
  }
}

// alfabeta.F:4-79
__global__ void k_alfabeta(const __grid_constant__ KP p, const double* __restrict__ T, const double* __restrict__ Sal, double* __restrict__ alpha, double* __restrict__ beta) {
  THREAD_IJ(p.istr - 2, p.iend + 2, p.jstr - 2, p.jend + 2)
  const int N = p.N;
  if (p.nonlin) {
    const double r01 = 6.793952e-2, r02 = -9.095290e-3, r03 = +1.001685e-4, r04 = -1.120083e-6, r05 = +6.536332e-9, r10 = +0.824493,
                 r11 = -4.08990e-3, r12 = +7.64380e-5, r13 = -8.24670e-7, r14 = +5.38750e-9, rS0 = -5.72466e-3, rS1 = +1.02270e-4,
                 rS2 = -1.65460e-6, r20 = +4.8314e-4;
    const double cff = 1.0 / p.rho0;
    const double Tt = R3(T, c, N), Ts = R3(Sal, c, N), sqrtTs = sqrt(DMAX(0.0, Ts));
    alpha[c] = -cff * (r01 + Tt * (2.0 * r02 + Tt * (3.0 * r03 + Tt * (4.0 * r04 + Tt * 5.0 * r05))) +
                       Ts * (r11 + Tt * (2.0 * r12 + Tt * (3.0 * r13 + Tt * 4.0 * r14)) + sqrtTs * (rS1 + Tt * 2.0 * rS2)));
    beta[c] = cff * (r10 + Tt * (r11 + Tt * (r12 + Tt * (r13 + Tt * r14))) + 1.5 * (rS0 + Tt * (rS1 + Tt * rS2)) * sqrtTs + 2.0 * r20 * Ts);
  } else {
    alpha[c] = fabs(p.Tcoef);
    beta[c] = p.salinity ? fabs(p.Scoef) : 0.0;
  }
}

// ------------------------------------------------------------------------------------------------
// Lateral mixing along S-surfaces.
//
// visc3d_S.F:18-134 visc3d_tile.  The reference sweeps four 2-D scratch planes (UFx,VFe at rho points, UFe,VFx at psi
// points) per level; here a thread owns the (i,j) column of u- and v-points, keeps every level-independent metric
// product of its three rho-point and three psi-point stresses in registers and walks k, so that rufrc/rvfrc receive
// the levels in the reference's order (k=1..N, one rounding per level) without atomics.
struct VisR { double A1, pnE, pnW, B1, pmN, pmS, nu, dn, dm; };      // rho point: (dn_r*pm), pn(i)+pn(i+1), pn(i-1)+pn(i), (dm_r*pn), ...
struct VisP { double C1, pnA, pnB, C2, pmA, pmB, nu, msk, dn, dm; }; // psi point
__device__ __forceinline__ VisR visr_load(const KP& p, ix_t c) {
  const ix_t SX = p.SX;
  VisR r;
  const double pm = p.pm[c], pn = p.pn[c];
  r.dn = p.dn_r[c]; r.dm = p.dm_r[c];
  r.A1 = r.dn * pm; r.pnE = pn + p.pn[c + 1]; r.pnW = p.pn[c - 1] + pn;
  r.B1 = r.dm * pn; r.pmN = pm + p.pm[c + SX]; r.pmS = p.pm[c - SX] + pm;
  r.nu = p.visc2_r[c];
  return r;
}
__device__ __forceinline__ VisP visp_load(const KP& p, ix_t c) {
  const ix_t SX = p.SX;
  VisP q;
  const double pm00 = p.pm[c], pmw = p.pm[c - 1], pms = p.pm[c - SX], pmsw = p.pm[c - 1 - SX];
  const double pn00 = p.pn[c], pnw = p.pn[c - 1], pns = p.pn[c - SX], pnsw = p.pn[c - 1 - SX];
  q.dn = p.dn_p[c]; q.dm = p.dm_p[c];
  q.C1 = 0.25 * (pmw + pm00 + pmsw + pms) * q.dn; q.pnA = pns + pn00; q.pnB = pnsw + pnw;
  q.C2 = 0.25 * (pnw + pn00 + pnsw + pns) * q.dm; q.pmA = pmw + pm00; q.pmB = pmsw + pms;
  q.nu = p.visc2_p[c]; q.msk = p.pmask[c];
  return q;
}
// cff of visc3d_S.F:64-74 at rho point c, level k (UFx = cff*dn*dn, VFe = -cff*dm*dm)
__device__ __forceinline__ double visr_cff(const KP& p, const double* __restrict__ u, const double* __restrict__ v, ix_t c, int k, const VisR& r) {
  return 0.5 * R3(p.Hz, c, k) * r.nu * (r.A1 * (r.pnE * R3(u, c + 1, k) - r.pnW * R3(u, c, k)) - r.B1 * (r.pmN * R3(v, c + p.SX, k) - r.pmS * R3(v, c, k)));
}
// cff of visc3d_S.F:80-97 at psi point c (UFe = cff*dm_p*dm_p, VFx = cff*dn_p*dn_p)
__device__ __forceinline__ double visp_cff(const KP& p, const double* __restrict__ u, const double* __restrict__ v, ix_t c, int k, const VisP& q) {
  const ix_t SX = p.SX;
  return 0.125 * (R3(p.Hz, c - 1, k) + R3(p.Hz, c, k) + R3(p.Hz, c - 1 - SX, k) + R3(p.Hz, c - SX, k)) * q.nu *
         (q.C1 * (q.pnA * R3(v, c, k) - q.pnB * R3(v, c - 1, k)) + q.C2 * (q.pmA * R3(u, c, k) - q.pmB * R3(u, c - SX, k))) * q.msk;
}
__global__ void __launch_bounds__(256, 2) k_visc3d(const __grid_constant__ KP p, const double* __restrict__ u, const double* __restrict__ v,
                                                    double* __restrict__ ui, double* __restrict__ vi) {
  THREAD_IJ(p.istr, p.iend, p.jstr, p.jend)
  const int N = p.N;
  const bool dou = i >= p.istrU, dov = j >= p.jstrV;
  const VisR r0 = visr_load(p, c), rw = visr_load(p, c - 1), rs = visr_load(p, c - SX);
  const VisP q0 = visp_load(p, c), qn = visp_load(p, c + SX), qe = visp_load(p, c + 1);
  const double pm0 = p.pm[c], pn0 = p.pn[c];
  const double pmu = p.pm[c - 1] + pm0, pnu = p.pn[c - 1] + pn0;        // u point (i,j)
  const double pmv = pm0 + p.pm[c - SX], pnv = pn0 + p.pn[c - SX];      // v point (i,j)
  const double cu = 0.125 * pmu * pnu, cv = 0.125 * pmv * pnv;
  const double dt = p.dt;
  double ruf = p.rufrc[c], rvf = p.rvfrc[c];
#pragma unroll 2
  for (int k = 1; k <= N; ++k) {
    const double s0 = visr_cff(p, u, v, c, k, r0), sw = visr_cff(p, u, v, c - 1, k, rw), ss = visr_cff(p, u, v, c - SX, k, rs);
    const double t0 = visp_cff(p, u, v, c, k, q0), tn = visp_cff(p, u, v, c + SX, k, qn), te = visp_cff(p, u, v, c + 1, k, qe);
    if (dou) {
      const double UFx_c = s0 * r0.dn * r0.dn, UFx_w = sw * rw.dn * rw.dn;
      const double UFe_n = tn * qn.dm * qn.dm, UFe_c = t0 * q0.dm * q0.dm;
      const double cff = cu * (pnu * (UFx_c - UFx_w) + pmu * (UFe_n - UFe_c));
      ruf = ruf + cff;
      R3(ui, c, k) = R3(ui, c, k) + dt * cff;
    }
    if (dov) {
      const double VFx_e = te * qe.dn * qe.dn, VFx_c = t0 * q0.dn * q0.dn;
      const double VFe_c = -s0 * r0.dm * r0.dm, VFe_s = -ss * rs.dm * rs.dm;
      const double cff = cv * (pnv * (VFx_e - VFx_c) + pmv * (VFe_c - VFe_s));
      rvf = rvf + cff;
      R3(vi, c, k) = R3(vi, c, k) + dt * cff;
    }
  }
  if (dou) p.rufrc[c] = ruf;
  if (dov) p.rvfrc[c] = rvf;
}

// t3dmix_S.F:54-105 (TS_DIF2): t(nnew) += dt*pm*pn*(FX(i+1)-FX(i)+FE(j+1)-FE(j))/Hz with the Laplacian fluxes recomputed in
// registers at the four faces of the point; KB levels per thread share the level-independent face coefficients and masks.
template <int KB>
__global__ void __launch_bounds__(256, 4) k_t3dmix(const __grid_constant__ KP p, const double* __restrict__ t_rhs, double* __restrict__ t_new,
                                                    const double* __restrict__ d2) {
  THREAD_IJ(p.istr, p.iend, p.jstr, p.jend)
  const int k0 = blockIdx.z * KB + 1;
  const double dc = d2[c];
  const double xw = 0.25 * (dc + d2[c - 1]) * p.pmon_u[c], xe = 0.25 * (d2[c + 1] + dc) * p.pmon_u[c + 1];
  const double ys = 0.25 * (dc + d2[c - SX]) * p.pnom_v[c], yn = 0.25 * (d2[c + SX] + dc) * p.pnom_v[c + SX];
  const double mw = p.umask[c], me = p.umask[c + 1], ms = p.vmask[c], mn = p.vmask[c + SX];
  const double cdt = p.dt * p.pm[c] * p.pn[c];
#pragma unroll
  for (int q = 0; q < KB; ++q) {
    const int k = k0 + q;
    if (k <= p.N) {
      const double H0 = R3(p.Hz, c, k), t0 = R3(t_rhs, c, k);
      const double FXw = xw * (H0 + R3(p.Hz, c - 1, k)) * (t0 - R3(t_rhs, c - 1, k)) * mw;
      const double FXe = xe * (R3(p.Hz, c + 1, k) + H0) * (R3(t_rhs, c + 1, k) - t0) * me;
      const double FEs = ys * (H0 + R3(p.Hz, c - SX, k)) * (t0 - R3(t_rhs, c - SX, k)) * ms;
      const double FEn = yn * (R3(p.Hz, c + SX, k) + H0) * (R3(t_rhs, c + SX, k) - t0) * mn;
      R3(t_new, c, k) = R3(t_new, c, k) + cdt * (FXe - FXw + FEn - FEs) / H0;
    }
  }
}
