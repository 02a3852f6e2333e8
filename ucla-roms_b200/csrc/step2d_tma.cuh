// step2d_tma.cuh -- the barotropic substep (step2d_FB.F:24-576) as a TMA-fed marching kernel (sm_100a).
//
// k_step2d (kernels.cuh) gives every thread one point and lets it fetch its ~110 operands through L1; ncu shows it bound by
// load issue and latency, not by DRAM (profiles/r1_ncu_full_v2_step2d.txt).  Here a block of S2M_TX threads owns a strip of
// S2M_TX-1 columns and marches north over `rows` rows.  One elected warp streams whole row segments of the 27 operand planes
// into a ring of S2M_S shared-memory stages with cp.async.bulk (the TMA unit: one instruction per plane and row, no address
// arithmetic in the compute threads, completion on an mbarrier), two rows ahead of the arithmetic.  Every operand crosses
// L2->SM exactly once per block; east/west neighbours are shared-memory reads, the southern neighbour's values are carried in
// registers from the previous row, zeta_new of the western neighbour comes through a 2-row exchange buffer (one
// __syncthreads per row).  The copies are issued by a producer warp of their own.  Thread 0 of a block is a ghost: it only evaluates zeta_new of the column left of the strip.
// The arithmetic is expression for expression that of k_step2d (bit-identical results; tests/test_gpu_parity.py).
#pragma once
#include <cstdint>

#define S2M_NA 27
// plane order: [0,4) suffice for row ja-2, [0,8) for the row north of the strip, [0,18) for the warm-up row ja-1
enum { A_H, A_ZS, A_ZB, A_ZO, A_VS, A_VB, A_VO, A_DMV, A_US, A_UB, A_UO, A_DNU, A_PM, A_PN, A_RM, A_SWF, A_RHOA, A_RHOS,
       A_UM, A_VM, A_RUF, A_RVF, A_ZT, A_DU1, A_DV1, A_DU2, A_DV2 };
struct S2Planes { const double* a[S2M_NA]; };

__device__ __forceinline__ uint32_t s2m_u32(const void* q) { return (uint32_t)__cvta_generic_to_shared(q); }
__device__ __forceinline__ void s2m_mbar_init(uint64_t* b, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s2m_u32(b)), "r"(count));
}
__device__ __forceinline__ void s2m_expect_tx(uint64_t* b, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s2m_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void s2m_wait(uint64_t* b, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(done) : "r"(s2m_u32(b)), "r"(parity) : "memory");
  } while (!done);
}
__device__ __forceinline__ void s2m_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* b) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s2m_u32(dst)), "l"(src), "r"(bytes),
               "r"(s2m_u32(b))
               : "memory");
}

// S2M_TX: compute threads (columns incl. the ghost) per block; S2M_S: ring depth in rows (rows j, j+1 in use, the rest in flight)
template <int S2M_TX, int S2M_S>
__global__ void __launch_bounds__(S2M_TX + 32) k_step2d_tma(const __grid_constant__ KP p, const __grid_constant__ S2D s, const __grid_constant__ S2Planes ap, int rows) {
  constexpr int S2M_W = S2M_TX + 4;      // doubles per staged row segment (16-byte aligned start, even length)
  extern __shared__ __align__(128) unsigned char s2m_raw[];
  double* st = (double*)s2m_raw;                                  // [S2M_S][S2M_NA][S2M_W]
  double* znx = st + S2M_S * S2M_NA * S2M_W;                      // [2][S2M_TX] zeta_new exchange
  uint64_t* full = (uint64_t*)(znx + 2 * S2M_TX);                 // [S2M_S]
  const int tx = threadIdx.x;
  const int i0 = p.istr + blockIdx.x * (S2M_TX - 1);              // first column the block finishes; thread tx <-> column i0-1+tx
  const int ja = p.jstr + blockIdx.y * rows;
  const int jb = min(ja + rows - 1, p.jend);
  const int i = i0 - 1 + tx;
  const int xoff = (i0 - 2 + IOFF) & 1;                           // staged segment starts at column i0-2-xoff (16-byte aligned)
  const int e0 = i0 - 2 - xoff + IOFF;                            // its element offset inside a row
  const int nW = min(S2M_W, p.SX - e0);                           // even: SX and e0 are
  const int x = tx + 1 + xoff;                                    // this thread's column inside a staged row
  const int r0 = ja - 2, rlast = jb + 1;                          // rows streamed: r0 .. rlast
  const ix2_t SX = p.SX;

  if (tx == 0) {
    for (int q = 0; q < S2M_S; ++q) s2m_mbar_init(&full[q], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  // producer warp (threads S2M_TX..S2M_TX+31; one lane works): streams row r of the planes it needs into slot (r-r0)%S, then joins the
  // compute warps' per-row barrier, after which the slot of row j-1 is free for row j-1+S
  if (tx >= S2M_TX) {
    // lane a owns plane a: its source pointer walks north one row per copy; the compiler serialises the lanes' UBLKCP (uniform
    // datapath instruction) with an ELECT loop of ~8 instructions per copy
    const int lane = tx - S2M_TX;
    const double* src0 = ap.a[lane < S2M_NA ? lane : 0] + e0;
    auto issue = [&](int r) {
      if (r > rlast) return;
      const int sl = (r - r0) % S2M_S;
      const int na = (r == r0) ? 4 : ((r == rlast) ? 8 : ((r == ja - 1) ? 18 : S2M_NA));
      if (lane == 0) s2m_expect_tx(&full[sl], (uint32_t)(na * nW * 8));
      __syncwarp();
      if (lane < na) s2m_bulk_g2s(st + (sl * S2M_NA + lane) * S2M_W, src0 + (ix2_t)(r + 1) * SX, (uint32_t)(nW * 8), &full[sl]);
    };
    for (int q = 0; q < S2M_S; ++q) issue(r0 + q);
    for (int j = ja - 1; j <= jb; ++j) {
      __syncthreads();
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      issue(j - 1 + S2M_S);
    }
    return;
  }
#define SM_(sl, A, dx) st[((sl)*S2M_NA + (A)) * S2M_W + x + (dx)]
#define DRHS_(sl, dx) (SM_(sl, A_H, dx) + s.fwd * SM_(sl, A_ZS, dx) + s.fwd1 * SM_(sl, A_ZB, dx) + s.fwd2 * SM_(sl, A_ZO, dx))
  // ---- rows ja-2 and ja-1: Drhs, DVom of the warm-up row
  s2m_wait(&full[0], 0);
  s2m_wait(&full[1 % S2M_S], 0);
  double Drhs_c, DVom_c;
  {
    const double Drhs_ss = DRHS_(0, 0);
    Drhs_c = DRHS_(1, 0);
    const double vrhs = s.fwd * SM_(1, A_VS, 0) + s.fwd1 * SM_(1, A_VB, 0) + s.fwd2 * SM_(1, A_VO, 0);
    DVom_c = 0.5 * (Drhs_c + Drhs_ss) * SM_(1, A_DMV, 0) * (vrhs);
  }
  // values of the southern neighbour (row j-1) carried in registers
  double zn_s = 0.0, h_s = 0.0, zs_s = 0.0, rhoA_s = 0.0, pm_s = 0.0, pn_s = 0.0;
  S2Pt as = {0.0, 0.0, 0.0, 0.0}, bs = {0.0, 0.0, 0.0, 0.0};
  const double cffg = 0.5 * p.g;
  const double cff = 0.5 * p.dtfast, cff1 = 0.5 * s.w1;
  const bool dou = (i >= p.istrU);            // u(istr) sits on a closed western wall: u2dbc, not the momentum update
  // single periodic subdomain: results are stored to their halo images too (s2d_store); the offsets are fixed per thread / per row
  const ix2_t ox = !p.wrap_x ? 0 : (i <= 2 ? (ix2_t)p.Lm : (i >= p.Lm - 1 ? -(ix2_t)p.Lm : 0));
#define S2ST_(a, cc, v, oyy)                                  \
  do {                                                        \
    (a)[cc] = (v);                                            \
    if (ox | (oyy)) {                                         \
      if (ox) (a)[(cc) + ox] = (v);                           \
      if (oyy) (a)[(cc) + (oyy)] = (v);                       \
      if (ox && (oyy)) (a)[(cc) + ox + (oyy)] = (v);          \
    }                                                         \
  } while (0)
  for (int j = ja - 1; j <= jb; ++j) {
    const ix2_t oy = !p.wrap_y ? 0 : (j <= 2 ? (ix2_t)p.Mm * SX : (j >= p.Mm - 1 ? -(ix2_t)p.Mm * SX : 0));
    const int C = (j - r0) % S2M_S, Nn = (j + 1 - r0) % S2M_S;
    s2m_wait(&full[Nn], (uint32_t)(((j + 1 - r0) / S2M_S) & 1));
    // ---- zeta_new(i,j) (step2d_FB.F:151-216)
    const double Drhs_n = DRHS_(Nn, 0);
    const double vrhs_n = s.fwd * SM_(Nn, A_VS, 0) + s.fwd1 * SM_(Nn, A_VB, 0) + s.fwd2 * SM_(Nn, A_VO, 0);
    const double DVom_n = 0.5 * (Drhs_n + Drhs_c) * SM_(Nn, A_DMV, 0) * (vrhs_n);
    const double Drhs_w = DRHS_(C, -1), Drhs_e = DRHS_(C, 1);
    const double urhs_c = s.fwd * SM_(C, A_US, 0) + s.fwd1 * SM_(C, A_UB, 0) + s.fwd2 * SM_(C, A_UO, 0);
    const double urhs_e = s.fwd * SM_(C, A_US, 1) + s.fwd1 * SM_(C, A_UB, 1) + s.fwd2 * SM_(C, A_UO, 1);
    const double DUon_c = 0.5 * (Drhs_c + Drhs_w) * SM_(C, A_DNU, 0) * (urhs_c);
    const double DUon_e = 0.5 * (Drhs_e + Drhs_c) * SM_(C, A_DNU, 1) * (urhs_e);
    const double zs = SM_(C, A_ZS, 0), pm0 = SM_(C, A_PM, 0), pn0 = SM_(C, A_PN, 0), h0 = SM_(C, A_H, 0);
    const ix2_t c = CIDX(i, j);
    double zn = zs + p.dtfast * pm0 * pn0 * (DUon_c - DUon_e + DVom_c - DVom_n) + p.dtfast * SM_(C, A_SWF, 0);
    if (p.pipes && i <= p.iend && p.pipe_idx[c] > 0.) zn = zn + p.dtfast * pm0 * pn0 * p.pipe_flx[c];   // step2d_FB.F:155-159
    zn = zn * SM_(C, A_RM, 0);
    znx[(j & 1) * S2M_TX + tx] = zn;
    __syncthreads();     // also tells the producer warp that every thread is past row j-1
    const double rhoA0 = SM_(C, A_RHOA, 0), rhoS0 = SM_(C, A_RHOS, 0);
    const double zb0 = SM_(C, A_ZB, 0), zo0 = SM_(C, A_ZO, 0);
    // s2d_ptz(c, zn, false) and, on the first substep, s2d_ptz(c, zn, true)
    S2Pt a0, b0 = {0.0, 0.0, 0.0, 0.0};
    a0.UFx = s.bkw_new * zn + s.bkw * zs + s.bkw1 * zb0 + s.bkw2 * zo0;
    a0.UFe = (1.0 + rhoS0) * a0.UFx;
    a0.VFx = a0.UFx * (rhoS0 - rhoA0);
    a0.VFe = a0.UFe * a0.UFx;
    if (s.iif == 1) {
      b0.UFx = zn - zs;
      b0.UFe = (1.0 + rhoS0) * b0.UFx;
      b0.VFe = b0.UFe * (zn + zs);
      b0.VFx = b0.UFx * (rhoS0 - rhoA0);
    }
    if (j >= ja && tx >= 1 && i <= p.iend && !(s.region == 1 && s2d_in_ring(p, i, j))) {
      // ---- own point (k_step2d phase 2)
      const double znw = znx[(j & 1) * S2M_TX + tx - 1];
      const bool swr = p.swall && j == p.jstr, nwr = p.nwall && j == p.jend;
      double zns = zn_s;
      if (swr) zns = zn * p.rmask[c - SX];       // zetabc.F: gradient condition on the row below the southern wall (s2d_zn_at)
      double DU1, DV1, Zt;
      S2ST_(s.zn, c, zn, oy);
      if (s.iif == 1) {
        p.DU_avg_bak[c] = SM_(C, A_DU1, 0) - 0.1024390243902439 * SM_(C, A_DU2, 0);
        p.DV_avg_bak[c] = SM_(C, A_DV1, 0) - 0.1024390243902439 * SM_(C, A_DV2, 0);
        Zt = s.w1 * zn;
        DU1 = 0.0;
        DV1 = 0.0;
        p.DU_avg2[c] = s.w2 * DUon_c;
        p.DV_avg2[c] = s.w2 * DVom_c;
      } else {
        Zt = SM_(C, A_ZT, 0) + s.w1 * zn;
        DU1 = SM_(C, A_DU1, 0);
        DV1 = SM_(C, A_DV1, 0);
        p.DU_avg2[c] = SM_(C, A_DU2, 0) + s.w2 * DUon_c;
        p.DV_avg2[c] = SM_(C, A_DV2, 0) + s.w2 * DVom_c;
      }
      p.Zt_avg1[c] = Zt;
      const bool sv = (j - 1 >= p.jstrV - 1);     // row j-1 carries a zeta_new
      const bool dov = (j >= p.jstrV);
      const S2Pt zero = {0.0, 0.0, 0.0, 0.0};     // row below a southern wall: zero-initialised scratch in the reference
      const double h_w = SM_(C, A_H, -1), rhoA_w = SM_(C, A_RHOA, -1), rhoS_w = SM_(C, A_RHOS, -1), zs_w = SM_(C, A_ZS, -1);
      S2Pt aw = zero, bw = zero;
      if (dou) {
        aw.UFx = s.bkw_new * znw + s.bkw * zs_w + s.bkw1 * SM_(C, A_ZB, -1) + s.bkw2 * SM_(C, A_ZO, -1);
        aw.UFe = (1.0 + rhoS_w) * aw.UFx;
        aw.VFx = aw.UFx * (rhoS_w - rhoA_w);
        aw.VFe = aw.UFe * aw.UFx;
      }
      S2Pt asv;      // component-wise selects: a struct select may be lowered through local memory (64 B of stack, -45 %)
      asv.UFx = sv ? as.UFx : 0.0; asv.UFe = sv ? as.UFe : 0.0; asv.VFx = sv ? as.VFx : 0.0; asv.VFe = sv ? as.VFe : 0.0;
      const double dnu = SM_(C, A_DNU, 0), dmv = SM_(C, A_DMV, 0);
      // s2d_pgf(cffg, dn, c, d, a, b): a at c-d, b at c
      double rubar = cffg * dnu *
                     ((h_w + h0) * (aw.UFe - a0.UFe) + aw.VFe - a0.VFe + (h_w - h0) * (aw.VFx + a0.VFx + 0.333333333333 * (rhoA_w - rhoA0) * (aw.UFx - a0.UFx)));
      double rvbar = cffg * dmv *
                     ((h_s + h0) * (asv.UFe - a0.UFe) + asv.VFe - a0.VFe + (h_s - h0) * (asv.VFx + a0.VFx + 0.333333333333 * (rhoA_s - rhoA0) * (asv.UFx - a0.UFx)));
      double rufrc = SM_(C, A_RUF, 0), rvfrc = SM_(C, A_RVF, 0);
      if (s.iif == 1) {
        rufrc = rufrc - rubar;
        rvfrc = rvfrc - rvbar;
        p.rufrc[c] = rufrc;
        p.rvfrc[c] = rvfrc;
        if (dou) {
          bw.UFx = znw - zs_w;
          bw.UFe = (1.0 + rhoS_w) * bw.UFx;
          bw.VFe = bw.UFe * (znw + zs_w);
          bw.VFx = bw.UFx * (rhoS_w - rhoA_w);
        }
        S2Pt bsv;
        bsv.UFx = sv ? bs.UFx : 0.0; bsv.UFe = sv ? bs.UFe : 0.0; bsv.VFx = sv ? bs.VFx : 0.0; bsv.VFe = sv ? bs.VFe : 0.0;
        rubar = rubar + cffg * dnu *
                            ((h_w + h0) * (bw.UFe - b0.UFe) + bw.VFe - b0.VFe + (h_w - h0) * (bw.VFx + b0.VFx + 0.333333333333 * (rhoA_w - rhoA0) * (bw.UFx - b0.UFx)));
        rvbar = rvbar + cffg * dmv *
                            ((h_s + h0) * (bsv.UFe - b0.UFe) + bsv.VFe - b0.VFe + (h_s - h0) * (bsv.VFx + b0.VFx + 0.333333333333 * (rhoA_s - rhoA0) * (bsv.UFx - b0.UFx)));
      }
      const double D0 = zs + h0, Dw = zs_w + h_w, Ds = zs_s + h_s;
      const double Dn0 = zn + h0, Dnw = znw + h_w, Dns = zns + h_s;
      double un = 0.0;      // u2dbc_im.F on a closed western wall: ubar(istr)=0 (its DU_avg1 term is identically zero)
      if (dou) {
        const double DUnew = ((D0 + Dw) * SM_(C, A_US, 0) + cff * (pm0 + SM_(C, A_PM, -1)) * (pn0 + SM_(C, A_PN, -1)) * (rubar + rufrc)) * SM_(C, A_UM, 0);
        un = DUnew / (Dn0 + Dnw);
        double DUa = DU1 + cff1 * dnu * (DUnew);
        if (p.rivers) {        // step2d_FB.F:531-554: a river face carries the river's volume flux
          const double code = p.riv_uflx[c];
          if (riv_on(code)) {
            const int iriver = (int)lround(code / 10);
            const double river_flux = p.riv_tab[(iriver - 1) * RIVROW] * (code - 10 * iriver);
            un = river_flux * 2 / (dnu * (Dnw + Dn0));
            DUa = river_flux;
          }
        }
        p.DU_avg1[c] = DUa;
      } else if (s.iif == 1) p.DU_avg1[c] = DU1;
      S2ST_(s.un, c, un, oy);
      if (dov) {
        const double DVnew = ((D0 + Ds) * SM_(C, A_VS, 0) + cff * (pm0 + pm_s) * (pn0 + pn_s) * (rvbar + rvfrc)) * SM_(C, A_VM, 0);
        double vn = DVnew / (Dn0 + Dns);
        double DVa = DV1 + cff1 * dmv * (DVnew);
        if (p.rivers) {
          const double code = p.riv_vflx[c];
          if (riv_on(code)) {
            const int iriver = (int)lround(code / 10);
            const double river_flux = p.riv_tab[(iriver - 1) * RIVROW] * (code - 10 * iriver);
            vn = river_flux * 2 / (dmv * (Dns + Dn0));
            DVa = river_flux;
          }
        }
        S2ST_(s.vn, c, vn, oy);
        p.DV_avg1[c] = DVa;
      } else {   // row jstr on a southern wall: v2dbc (no normal flow) and its DV_avg1 term (step2d_FB.F:449-529)
        const double vn = 0.0;
        S2ST_(s.vn, c, vn, oy);
        p.DV_avg1[c] = DV1 + cff1 * (Dn0 + Dns) * (vn)*dmv;
      }
      if (s.iif == s.nfast) S2ST_(s.zn, c, Zt, oy);
      // ---- wall point beside this one (k_step2d, same expressions; operands straight from global memory: two rows per domain)
      if (swr | nwr) {
#pragma unroll 1
        for (int side = 0; side < 2; ++side) {
          if (side == 0 ? !swr : !nwr) continue;
          const int jw = side == 0 ? j - 1 : j + 1;
          const ix2_t w = side == 0 ? c - SX : c + SX;
          const double znb = side == 0 ? zns : zn * p.rmask[w];
          const double znbw = znw * p.rmask[w - 1];      // side 0: s2d_zn_at(i-1, jstr-1) = zeta_new(i-1,jstr)*rmask
          double DU1w, DV1w, Ztw;
          s2d_avg_out(p, s, i, jw, w, znb, s2d_DUon(p, s, w), side == 0 ? 0.0 : s2d_DVom(p, s, w), DU1w, DV1w, Ztw);
          const double unw = p.gamma2 * un * p.umask[w];          // u2dbc_im.F:150-152
          S2ST_(s.un, w, unw, (ix2_t)0);
          if (dou) p.DU_avg1[w] = DU1w + cff1 * ((p.h[w] + znb) + (p.h[w - 1] + znbw)) * (unw)*p.dn_u[w];
          else if (s.iif == 1) p.DU_avg1[w] = DU1w;
          if (side == 1) {
            const double vnw = 0.0;
            S2ST_(s.vn, w, vnw, (ix2_t)0);
            p.DV_avg1[w] = DV1w + cff1 * ((znb + p.h[w]) + (zn + p.h[c])) * (vnw)*p.dm_v[w];
          } else if (s.iif == 1) p.DV_avg1[w] = DV1w;
          if (s.iif == s.nfast) S2ST_(s.zn, w, Ztw, (ix2_t)0);
        }
      }
    }
    // ---- this row becomes the southern neighbour
    zn_s = zn; h_s = h0; zs_s = zs; rhoA_s = rhoA0; pm_s = pm0; pn_s = pn0;
    as = a0; bs = b0;
    Drhs_c = Drhs_n; DVom_c = DVom_n;
  }
#undef SM_
#undef S2ST_
#undef DRHS_
}
