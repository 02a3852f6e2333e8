// step2d_hyb.cuh -- k_step2d_hyb: the TMA-fed marching substep (step2d_tma.cuh) on a shared-memory diet.
//
// Only the 12 operand planes whose east/west neighbours are read (h, zeta x3, ubar x3, dn_u, pm, pn, rhoA, rhoS) go through the
// cp.async.bulk ring; the 15 planes of which a thread reads its own column only (vbar x3, dm_v, masks, forcing, the five
// fast-time averages) are requested with plain loads one row iteration ahead and live in registers, as in k_step2d_march.
// A stage shrinks from 14.7 to 6.5 KB, the producer warp is gone (lanes 0..11 of warp 0 issue the copies of a row right after the
// row barrier), so 6 blocks of 64 threads fit an SM instead of 3 blocks of 96: twice the resident compute warps.
// (The row body is generated from step2d_tma.cuh; arithmetic expression for expression that of k_step2d.)
#pragma once
enum { HA_H, HA_ZS, HA_ZB, HA_ZO, HA_US, HA_UB, HA_UO, HA_DNU, HA_PM, HA_PN, HA_RHOA, HA_RHOS, HA_N };
struct S2PlanesH { const double* a[HA_N]; };

template <int TX, int NS, int MINB>
__global__ void __launch_bounds__(TX, MINB) k_step2d_hyb(const __grid_constant__ KP p, const __grid_constant__ S2D s, const __grid_constant__ S2PlanesH ap, int rows) {
  constexpr int W = TX + 4;
  extern __shared__ __align__(128) unsigned char s2m_raw[];
  double* st = (double*)s2m_raw;                                  // [NS][HA_N][W]
  double* znx = st + NS * HA_N * W;                               // [2][TX] zeta_new exchange
  uint64_t* full = (uint64_t*)(znx + 2 * TX);                     // [NS]
  const int tx = threadIdx.x;
  const int i0 = p.istr + blockIdx.x * (TX - 1);                  // first column the block finishes; thread tx <-> column i0-1+tx
  const int ja = p.jstr + blockIdx.y * rows;
  const int jb = min(ja + rows - 1, p.jend);
  const int i = i0 - 1 + tx;
  const int il = min(i, p.iend + 1);                              // threads right of the domain recompute column iend+1 (never stored)
  const int xoff = (i0 - 2 + IOFF) & 1;                           // staged segment starts at column i0-2-xoff (16-byte aligned)
  const int e0 = i0 - 2 - xoff + IOFF;
  const int nW = min(W, p.SX - e0);
  const int x = min(tx, p.iend + 2 - i0) + 1 + xoff;              // this thread's column inside a staged row (clamped like il)
  const int r0 = ja - 2, rlast = jb + 1;
  const ix2_t SX = p.SX;
  if (tx == 0) {
    for (int q = 0; q < NS; ++q) s2m_mbar_init(&full[q], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  // lane a of warp 0 owns plane a; rows ja-2 and jb+1 only need h and the three zeta levels
  const double* src0 = ap.a[tx < HA_N ? tx : 0] + e0;
  // (a macro, not a lambda: when the by-reference lambda was not inlined its captures lived on the stack -- 64 B, -45 %)
#define S2H_ISSUE(r_)                                                                                                          \
  do {                                                                                                                         \
    const int rr_ = (r_);                                                                                                      \
    if (rr_ <= rlast) {                                                                                                        \
      const int sl_ = (rr_ - r0) % NS;                                                                                         \
      const int na_ = (rr_ == r0 || rr_ == rlast) ? 4 : HA_N;                                                                  \
      if (tx == 0) s2m_expect_tx(&full[sl_], (uint32_t)(na_ * nW * 8));                                                        \
      __syncwarp();                                                                                                            \
      if (tx < na_) s2m_bulk_g2s(st + (sl_ * HA_N + tx) * W, src0 + (ix2_t)(rr_ + 1) * SX, (uint32_t)(nW * 8), &full[sl_]);    \
    }                                                                                                                          \
  } while (0)
  if (tx < 32)
    for (int q = 0; q < NS; ++q) S2H_ISSUE(r0 + q);
#define SMH_(sl, A, dx) st[((sl)*HA_N + (A)) * W + x + (dx)]
#define DRHS_(sl, dx) (SMH_(sl, HA_H, dx) + s.fwd * SMH_(sl, HA_ZS, dx) + s.fwd1 * SMH_(sl, HA_ZB, dx) + s.fwd2 * SMH_(sl, HA_ZO, dx))
  // own-column operands kept in registers: row ja-1 (c*), the v-type ones of row ja (n*)
  double cRM, cSWF, cUM, cVM, cRUF, cRVF, cZT, cDU1, cDV1, cDU2, cDV2, cVS, cDMV;
  double nVS, nVB, nVO, nDMV;
  double Drhs_c, DVom_c;
  {
    const ix2_t c1 = CIDX(il, ja - 1);
    cVS = s.vs[c1]; cDMV = p.dm_v[c1];
    const double vb1 = s.vb[c1], vo1 = s.vo[c1];
    cRM = p.rmask[c1]; cSWF = p.swflx[c1]; cUM = p.umask[c1]; cVM = p.vmask[c1]; cRUF = p.rufrc[c1]; cRVF = p.rvfrc[c1]; cZT = p.Zt_avg1[c1]; cDU1 = p.DU_avg1[c1]; cDV1 = p.DV_avg1[c1]; cDU2 = p.DU_avg2[c1]; cDV2 = p.DV_avg2[c1];
    nVS = s.vs[c1 + SX]; nVB = s.vb[c1 + SX]; nVO = s.vo[c1 + SX]; nDMV = p.dm_v[c1 + SX];
    s2m_wait(&full[0], 0);
    s2m_wait(&full[1 % NS], 0);
    const double Drhs_ss = DRHS_(0, 0);
    Drhs_c = DRHS_(1, 0);
    const double vrhs = s.fwd * cVS + s.fwd1 * vb1 + s.fwd2 * vo1;
    DVom_c = 0.5 * (Drhs_c + Drhs_ss) * cDMV * (vrhs);
  }
  double zn_s = 0.0, h_s = 0.0, zs_s = 0.0, rhoA_s = 0.0, pm_s = 0.0, pn_s = 0.0;
  S2Pt as = {0.0, 0.0, 0.0, 0.0}, bs = {0.0, 0.0, 0.0, 0.0};
  const double cffg = 0.5 * p.g;
  const double cff = 0.5 * p.dtfast, cff1 = 0.5 * s.w1;
  const bool dou = (i >= p.istrU);
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
    const int C = (j - r0) % NS, Nn = (j + 1 - r0) % NS;
    const ix2_t c = CIDX(il, j);
    // requests for the next iteration: own-column operands of row j+1, v-type operands of row j+2
    const double pRM = p.rmask[c + SX]; const double pSWF = p.swflx[c + SX];
    // the nine planes that only the second half of the row needs (masks, forcing, fast-time averages): asked into L2 one row ahead
    // (no register cost), loaded here, used after the row barrier -- their registers do not live across the loop back-edge
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p.umask + c + SX)); asm volatile("prefetch.global.L2 [%0];" ::"l"(p.vmask + c + SX)); asm volatile("prefetch.global.L2 [%0];" ::"l"(p.rufrc + c + SX)); asm volatile("prefetch.global.L2 [%0];" ::"l"(p.rvfrc + c + SX)); asm volatile("prefetch.global.L2 [%0];" ::"l"(p.Zt_avg1 + c + SX)); asm volatile("prefetch.global.L2 [%0];" ::"l"(p.DU_avg1 + c + SX)); asm volatile("prefetch.global.L2 [%0];" ::"l"(p.DV_avg1 + c + SX)); asm volatile("prefetch.global.L2 [%0];" ::"l"(p.DU_avg2 + c + SX)); asm volatile("prefetch.global.L2 [%0];" ::"l"(p.DV_avg2 + c + SX));
    cUM = p.umask[c]; cVM = p.vmask[c]; cRUF = p.rufrc[c]; cRVF = p.rvfrc[c]; cZT = p.Zt_avg1[c]; cDU1 = p.DU_avg1[c]; cDV1 = p.DV_avg1[c]; cDU2 = p.DU_avg2[c]; cDV2 = p.DV_avg2[c];
    const double qVS = s.vs[c + 2 * SX]; const double qVB = s.vb[c + 2 * SX]; const double qVO = s.vo[c + 2 * SX]; const double qDMV = p.dm_v[c + 2 * SX];
    s2m_wait(&full[Nn], (uint32_t)(((j + 1 - r0) / NS) & 1));
    // ---- zeta_new(i,j) (step2d_FB.F:151-216)
    const double Drhs_n = DRHS_(Nn, 0);
    const double vrhs_n = s.fwd * nVS + s.fwd1 * nVB + s.fwd2 * nVO;
    const double DVom_n = 0.5 * (Drhs_n + Drhs_c) * nDMV * (vrhs_n);
    const double Drhs_w = DRHS_(C, -1), Drhs_e = DRHS_(C, 1);
    const double urhs_c = s.fwd * SMH_(C, HA_US, 0) + s.fwd1 * SMH_(C, HA_UB, 0) + s.fwd2 * SMH_(C, HA_UO, 0);
    const double urhs_e = s.fwd * SMH_(C, HA_US, 1) + s.fwd1 * SMH_(C, HA_UB, 1) + s.fwd2 * SMH_(C, HA_UO, 1);
    const double DUon_c = 0.5 * (Drhs_c + Drhs_w) * SMH_(C, HA_DNU, 0) * (urhs_c);
    const double DUon_e = 0.5 * (Drhs_e + Drhs_c) * SMH_(C, HA_DNU, 1) * (urhs_e);
    const double zs = SMH_(C, HA_ZS, 0), pm0 = SMH_(C, HA_PM, 0), pn0 = SMH_(C, HA_PN, 0), h0 = SMH_(C, HA_H, 0);
    double zn = zs + p.dtfast * pm0 * pn0 * (DUon_c - DUon_e + DVom_c - DVom_n) + p.dtfast * cSWF;
    if (p.pipes && i <= p.iend && p.pipe_idx[c] > 0.) zn = zn + p.dtfast * pm0 * pn0 * p.pipe_flx[c];   // step2d_FB.F:155-159
    zn = zn * cRM;
    znx[(j & 1) * TX + tx] = zn;
    __syncthreads();     // every thread is past row j-1: its slot is free
    if (tx < 32) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      S2H_ISSUE(j - 1 + NS);
    }
    const double rhoA0 = SMH_(C, HA_RHOA, 0), rhoS0 = SMH_(C, HA_RHOS, 0);
    const double zb0 = SMH_(C, HA_ZB, 0), zo0 = SMH_(C, HA_ZO, 0);
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
      const double znw = znx[(j & 1) * TX + tx - 1];
      const bool swr = p.swall && j == p.jstr, nwr = p.nwall && j == p.jend;
      double zns = zn_s;
      if (swr) zns = zn * p.rmask[c - SX];       // zetabc.F: gradient condition on the row below the southern wall (s2d_zn_at)
      double DU1, DV1, Zt;
      S2ST_(s.zn, c, zn, oy);
      if (s.iif == 1) {
        p.DU_avg_bak[c] = cDU1 - 0.1024390243902439 * cDU2;
        p.DV_avg_bak[c] = cDV1 - 0.1024390243902439 * cDV2;
        Zt = s.w1 * zn;
        DU1 = 0.0;
        DV1 = 0.0;
        p.DU_avg2[c] = s.w2 * DUon_c;
        p.DV_avg2[c] = s.w2 * DVom_c;
      } else {
        Zt = cZT + s.w1 * zn;
        DU1 = cDU1;
        DV1 = cDV1;
        p.DU_avg2[c] = cDU2 + s.w2 * DUon_c;
        p.DV_avg2[c] = cDV2 + s.w2 * DVom_c;
      }
      p.Zt_avg1[c] = Zt;
      const bool sv = (j - 1 >= p.jstrV - 1);     // row j-1 carries a zeta_new
      const bool dov = (j >= p.jstrV);
      const S2Pt zero = {0.0, 0.0, 0.0, 0.0};     // row below a southern wall: zero-initialised scratch in the reference
      const double h_w = SMH_(C, HA_H, -1), rhoA_w = SMH_(C, HA_RHOA, -1), rhoS_w = SMH_(C, HA_RHOS, -1), zs_w = SMH_(C, HA_ZS, -1);
      S2Pt aw = zero, bw = zero;
      if (dou) {
        aw.UFx = s.bkw_new * znw + s.bkw * zs_w + s.bkw1 * SMH_(C, HA_ZB, -1) + s.bkw2 * SMH_(C, HA_ZO, -1);
        aw.UFe = (1.0 + rhoS_w) * aw.UFx;
        aw.VFx = aw.UFx * (rhoS_w - rhoA_w);
        aw.VFe = aw.UFe * aw.UFx;
      }
      S2Pt asv;      // component-wise selects: a struct select may be lowered through local memory (64 B of stack, -45 %)
      asv.UFx = sv ? as.UFx : 0.0; asv.UFe = sv ? as.UFe : 0.0; asv.VFx = sv ? as.VFx : 0.0; asv.VFe = sv ? as.VFe : 0.0;
      const double dnu = SMH_(C, HA_DNU, 0), dmv = cDMV;
      // s2d_pgf(cffg, dn, c, d, a, b): a at c-d, b at c
      double rubar = cffg * dnu *
                     ((h_w + h0) * (aw.UFe - a0.UFe) + aw.VFe - a0.VFe + (h_w - h0) * (aw.VFx + a0.VFx + 0.333333333333 * (rhoA_w - rhoA0) * (aw.UFx - a0.UFx)));
      double rvbar = cffg * dmv *
                     ((h_s + h0) * (asv.UFe - a0.UFe) + asv.VFe - a0.VFe + (h_s - h0) * (asv.VFx + a0.VFx + 0.333333333333 * (rhoA_s - rhoA0) * (asv.UFx - a0.UFx)));
      double rufrc = cRUF, rvfrc = cRVF;
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
        const double DUnew = ((D0 + Dw) * SMH_(C, HA_US, 0) + cff * (pm0 + SMH_(C, HA_PM, -1)) * (pn0 + SMH_(C, HA_PN, -1)) * (rubar + rufrc)) * cUM;
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
        const double DVnew = ((D0 + Ds) * cVS + cff * (pm0 + pm_s) * (pn0 + pn_s) * (rvbar + rvfrc)) * cVM;
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
    cVS = nVS; cDMV = nDMV;
    cRM = pRM; cSWF = pSWF;
    nVS = qVS; nVB = qVB; nVO = qVO; nDMV = qDMV;
  }
#undef SMH_
#undef S2H_ISSUE
#undef DRHS_
#undef S2ST_
}
