// step2d_march.cuh -- the barotropic substep (step2d_FB.F:24-576) as a register-marching kernel.
//
// Same decomposition as k_step2d_tma (strips of TX-1 columns, a block marches north over `rows` rows, thread 0 is the ghost that
// only evaluates zeta_new of the column left of the strip), but without shared-memory staging: every thread keeps the operands
// of its own column in registers -- the 27 values of the next row are requested from HBM one full row iteration before they
// are used, the southern neighbour's values are carried from the previous iteration -- and reads its east/west neighbours'
// operands (16 per row) through L1, where the lines already are.  ~43 loads per point instead of the ~110 of the tile kernel
// k_step2d, and no smem capacity limit on the number of resident warps.  zeta_new of the western neighbour comes through a
// 2-row exchange buffer (one __syncthreads per row).  Arithmetic: expression for expression that of k_step2d.
// (This file's row body is generated from step2d_tma.cuh by replacing staged operands with registers / L1 loads.)
#pragma once

template <int TX, int MINB>
__global__ void __launch_bounds__(TX, MINB) k_step2d_march(const __grid_constant__ KP p, const __grid_constant__ S2D s, int rows) {
  __shared__ double znx[2 * TX];
  const int tx = threadIdx.x;
  const int i0 = p.istr + blockIdx.x * (TX - 1);              // first column the block finishes; thread tx <-> column i0-1+tx
  const int ja = p.jstr + blockIdx.y * rows;
  const int jb = min(ja + rows - 1, p.jend);
  const int i = i0 - 1 + tx;
  const int il = min(i, p.iend + 1);                          // threads right of the domain recompute column iend+1 (never stored)
  const ix2_t SX = p.SX;
  // ---- rows ja-2 and ja-1: Drhs, DVom of the warm-up row; operands of rows ja-1 (C) and ja (N)
  double Drhs_c, DVom_c;
  double cH, cZS, cZB, cZO, cVS, cDMV;
  double cUS, cUB, cUO, cDNU, cPM, cPN, cRM, cSWF, cRHOA, cRHOS, cUM, cVM, cRUF, cRVF, cZT, cDU1, cDV1, cDU2, cDV2;
  double nH, nZS, nZB, nZO, nVS, nVB, nVO, nDMV;
  {
    const ix2_t c1 = CIDX(il, ja - 1);
    const double Drhs_ss = p.h[c1 - SX] + s.fwd * s.zs[c1 - SX] + s.fwd1 * s.zb[c1 - SX] + s.fwd2 * s.zo[c1 - SX];
    cH = p.h[c1]; cZS = s.zs[c1]; cZB = s.zb[c1]; cZO = s.zo[c1]; cVS = s.vs[c1]; cDMV = p.dm_v[c1];
    Drhs_c = cH + s.fwd * cZS + s.fwd1 * cZB + s.fwd2 * cZO;
    const double vrhs = s.fwd * cVS + s.fwd1 * s.vb[c1] + s.fwd2 * s.vo[c1];
    DVom_c = 0.5 * (Drhs_c + Drhs_ss) * cDMV * (vrhs);
    cUS = s.us[c1]; cUB = s.ub[c1]; cUO = s.uo[c1]; cDNU = p.dn_u[c1]; cPM = p.pm[c1]; cPN = p.pn[c1]; cRM = p.rmask[c1]; cSWF = p.swflx[c1]; cRHOA = p.rhoA[c1]; cRHOS = p.rhoS[c1]; cUM = p.umask[c1]; cVM = p.vmask[c1]; cRUF = p.rufrc[c1]; cRVF = p.rvfrc[c1]; cZT = p.Zt_avg1[c1]; cDU1 = p.DU_avg1[c1]; cDV1 = p.DV_avg1[c1]; cDU2 = p.DU_avg2[c1]; cDV2 = p.DV_avg2[c1];
    nH = p.h[c1 + SX]; nZS = s.zs[c1 + SX]; nZB = s.zb[c1 + SX]; nZO = s.zo[c1 + SX]; nVS = s.vs[c1 + SX]; nVB = s.vb[c1 + SX]; nVO = s.vo[c1 + SX]; nDMV = p.dm_v[c1 + SX];
  }
  // values of the southern neighbour (row j-1) carried in registers
  double zn_s = 0.0, h_s = 0.0, zs_s = 0.0, rhoA_s = 0.0, pm_s = 0.0, pn_s = 0.0;
  S2Pt as = {0.0, 0.0, 0.0, 0.0}, bs = {0.0, 0.0, 0.0, 0.0};
  const double cffg = 0.5 * p.g;
  const double cff = 0.5 * p.dtfast, cff1 = 0.5 * s.w1;
  const bool dou = (i >= p.istrU);            // u(istr) sits on a closed western wall: u2dbc, not the momentum update
  for (int j = ja - 1; j <= jb; ++j) {
    const ix2_t c = CIDX(il, j);
    // ---- requests for the next iteration: own-column operands of row j+1, northern operands of row j+2
    const double pUS = s.us[c + SX]; const double pUB = s.ub[c + SX]; const double pUO = s.uo[c + SX]; const double pDNU = p.dn_u[c + SX]; const double pPM = p.pm[c + SX]; const double pPN = p.pn[c + SX]; const double pRM = p.rmask[c + SX]; const double pSWF = p.swflx[c + SX]; const double pRHOA = p.rhoA[c + SX]; const double pRHOS = p.rhoS[c + SX]; const double pUM = p.umask[c + SX]; const double pVM = p.vmask[c + SX]; const double pRUF = p.rufrc[c + SX]; const double pRVF = p.rvfrc[c + SX]; const double pZT = p.Zt_avg1[c + SX]; const double pDU1 = p.DU_avg1[c + SX]; const double pDV1 = p.DV_avg1[c + SX]; const double pDU2 = p.DU_avg2[c + SX]; const double pDV2 = p.DV_avg2[c + SX];
    const double qH = p.h[c + 2 * SX]; const double qZS = s.zs[c + 2 * SX]; const double qZB = s.zb[c + 2 * SX]; const double qZO = s.zo[c + 2 * SX]; const double qVS = s.vs[c + 2 * SX]; const double qVB = s.vb[c + 2 * SX]; const double qVO = s.vo[c + 2 * SX]; const double qDMV = p.dm_v[c + 2 * SX];
    // ---- zeta_new(i,j) (step2d_FB.F:151-216)
    const double Drhs_n = nH + s.fwd * nZS + s.fwd1 * nZB + s.fwd2 * nZO;
    const double vrhs_n = s.fwd * nVS + s.fwd1 * nVB + s.fwd2 * nVO;
    const double DVom_n = 0.5 * (Drhs_n + Drhs_c) * nDMV * (vrhs_n);
    const double Drhs_w = p.h[c - 1] + s.fwd * s.zs[c - 1] + s.fwd1 * s.zb[c - 1] + s.fwd2 * s.zo[c - 1];
    const double Drhs_e = p.h[c + 1] + s.fwd * s.zs[c + 1] + s.fwd1 * s.zb[c + 1] + s.fwd2 * s.zo[c + 1];
    const double urhs_c = s.fwd * cUS + s.fwd1 * cUB + s.fwd2 * cUO;
    const double urhs_e = s.fwd * s.us[c + 1] + s.fwd1 * s.ub[c + 1] + s.fwd2 * s.uo[c + 1];
    const double DUon_c = 0.5 * (Drhs_c + Drhs_w) * cDNU * (urhs_c);
    const double DUon_e = 0.5 * (Drhs_e + Drhs_c) * p.dn_u[c + 1] * (urhs_e);
    const double zs = cZS, pm0 = cPM, pn0 = cPN, h0 = cH;
    double zn = zs + p.dtfast * pm0 * pn0 * (DUon_c - DUon_e + DVom_c - DVom_n) + p.dtfast * cSWF;
    if (p.pipes && i <= p.iend && p.pipe_idx[c] > 0.) zn = zn + p.dtfast * pm0 * pn0 * p.pipe_flx[c];   // step2d_FB.F:155-159
    zn = zn * cRM;
    znx[(j & 1) * TX + tx] = zn;
    __syncthreads();
    const double rhoA0 = cRHOA, rhoS0 = cRHOS;
    const double zb0 = cZB, zo0 = cZO;
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
    if (j >= ja && tx >= 1 && i <= p.iend) {
      // ---- own point (k_step2d phase 2)
      const double znw = znx[(j & 1) * TX + tx - 1];
      const bool swr = p.swall && j == p.jstr, nwr = p.nwall && j == p.jend;
      double zns = zn_s;
      if (swr) zns = zn * p.rmask[c - SX];       // zetabc.F: gradient condition on the row below the southern wall (s2d_zn_at)
      double DU1, DV1, Zt;
      s2d_store(p, s.zn, i, j, c, zn);
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
      const double h_w = p.h[c - 1], rhoA_w = p.rhoA[c - 1], rhoS_w = p.rhoS[c - 1], zs_w = s.zs[c - 1];
      S2Pt aw = zero, bw = zero;
      if (dou) {
        aw.UFx = s.bkw_new * znw + s.bkw * zs_w + s.bkw1 * s.zb[c - 1] + s.bkw2 * s.zo[c - 1];
        aw.UFe = (1.0 + rhoS_w) * aw.UFx;
        aw.VFx = aw.UFx * (rhoS_w - rhoA_w);
        aw.VFe = aw.UFe * aw.UFx;
      }
      S2Pt asv;      // component-wise selects: a struct select may be lowered through local memory (64 B of stack, -45 %)
      asv.UFx = sv ? as.UFx : 0.0; asv.UFe = sv ? as.UFe : 0.0; asv.VFx = sv ? as.VFx : 0.0; asv.VFe = sv ? as.VFe : 0.0;
      const double dnu = cDNU, dmv = cDMV;
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
        const double DUnew = ((D0 + Dw) * cUS + cff * (pm0 + p.pm[c - 1]) * (pn0 + p.pn[c - 1]) * (rubar + rufrc)) * cUM;
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
      s2d_store(p, s.un, i, j, c, un);
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
        s2d_store(p, s.vn, i, j, c, vn);
        p.DV_avg1[c] = DVa;
      } else {   // row jstr on a southern wall: v2dbc (no normal flow) and its DV_avg1 term (step2d_FB.F:449-529)
        const double vn = 0.0;
        s2d_store(p, s.vn, i, j, c, vn);
        p.DV_avg1[c] = DV1 + cff1 * (Dn0 + Dns) * (vn)*dmv;
      }
      if (s.iif == s.nfast) s2d_store(p, s.zn, i, j, c, Zt);
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
          s2d_store(p, s.un, i, jw, w, unw);
          if (dou) p.DU_avg1[w] = DU1w + cff1 * ((p.h[w] + znb) + (p.h[w - 1] + znbw)) * (unw)*p.dn_u[w];
          else if (s.iif == 1) p.DU_avg1[w] = DU1w;
          if (side == 1) {
            const double vnw = 0.0;
            s2d_store(p, s.vn, i, jw, w, vnw);
            p.DV_avg1[w] = DV1w + cff1 * ((znb + p.h[w]) + (zn + p.h[c])) * (vnw)*p.dm_v[w];
          } else if (s.iif == 1) p.DV_avg1[w] = DV1w;
          if (s.iif == s.nfast) s2d_store(p, s.zn, i, jw, w, Ztw);
        }
      }
    }
    // ---- this row becomes the southern neighbour
    zn_s = zn; h_s = h0; zs_s = zs; rhoA_s = rhoA0; pm_s = pm0; pn_s = pn0;
    as = a0; bs = b0;
    Drhs_c = Drhs_n; DVom_c = DVom_n;
    cH = nH; cZS = nZS; cZB = nZB; cZO = nZO; cVS = nVS; cDMV = nDMV;
    cUS = pUS; cUB = pUB; cUO = pUO; cDNU = pDNU; cPM = pPM; cPN = pPN; cRM = pRM; cSWF = pSWF; cRHOA = pRHOA; cRHOS = pRHOS; cUM = pUM; cVM = pVM; cRUF = pRUF; cRVF = pRVF; cZT = pZT; cDU1 = pDU1; cDV1 = pDV1; cDU2 = pDU2; cDV2 = pDV2;
    nH = qH; nZS = qZS; nZB = qZB; nZO = qZO; nVS = qVS; nVB = qVB; nVO = qVO; nDMV = qDMV;
  }
}
