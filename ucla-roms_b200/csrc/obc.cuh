// obc.cuh -- lateral boundary conditions on the non-periodic edges of a subdomain when at least one of them is OPEN
// (SURVEY.md 8(f) row f3): the combination every open-boundary case of the reference is built with, OBC_M2FLATHER (zetabc.F:28-36,
// u2dbc_im.F:28-56, v2dbc_im.F), the upstream-advection / binding rule for the tangential barotropic component (u2dbc_im.F:276-330),
// OBC_M3ORLANSKI (u3dbc_im.F:47-290, v3dbc_im.F) and OBC_TORLANSKI (t3dbc_im.F:36-277), with Z/M2/M3/T_FRC_BRY boundary data
// (boundary.F:21-60).  Closed edges of the same subdomain keep their wall rules (gradient zeta and tracers, no normal flow, gamma2 slip).
//
// Design: every boundary value is a PURE FUNCTION of (a) fields of the older time levels and (b) new-level values at interior
// points, which the interior kernels have already written.  ob_zn / ob_un / ob_vn (2-D) and ob3_u / ob3_v / ob3_t (3-D) evaluate that
// function at any point of the extended range, recursing from a corner to its two edges and from an edge to the interior, so one
// thread per boundary point can finish its point -- value, fast-time averages, flux accumulations (step2d_FB.F:449-529) -- without
// reading anything another thread of the same launch writes, in whatever order the reference sweeps its edges.  The arithmetic of
// each edge keeps the reference's operation order (the edges are not mirror images: t3dbc takes the interior neighbour at nstp on the
// west/south and at nnew on the east/north).
#pragma once

#define OB_B1(a, q) (a)[(q)]
#define OB_B2(a, q, k, L) (a)[(q) + ((k)-1) * (L)]
__device__ __forceinline__ bool ob_open(const KP& p, int side) { return (p.obc >> side) & 1; }
__device__ __forceinline__ double ob_max0(double a) { return (a < 0.) ? 0. : a; }     // max(a,0.) / min(a,0.) with the value returned for a = -0.
__device__ __forceinline__ double ob_min0(double a) { return (0. < a) ? 0. : a; }     // spelled out, so that both sides of the parity test agree

// ---- 2-D: zeta_new on the extended range (zetabc.F) -------------------------------------------------------------------------
__device__ __noinline__ double ob_zn_edge(const KP& p, const S2D& s, int i, int j) {
  const bool w = p.wwall && i < p.istr, e = p.ewall && i > p.iend, so = p.swall && j < p.jstr, n = p.nwall && j > p.jend;
  if (!(w | e | so | n)) return s2d_zeta_new(p, s, CIDX(i, j));
  const int ia = w ? p.istr : (e ? p.iend : i), ja = so ? p.jstr : (n ? p.jend : j);
  const ix2_t c = CIDX(i, j), ca = CIDX(ia, ja);
  const int side = w ? 0 : (e ? 1 : (so ? 2 : 3));
  if (ob_open(p, side) && p.obc_m2 == 0) {      // the Flather companion; OBC_M2SPECIFIED / no switch: gradient, like a wall (zetabc.F:54-61)
    const double cx = p.dtfast * (side < 2 ? p.pm[ca] : p.pn[ca]) * sqrt(p.g * p.h[ca]);
    return ((1. - cx) * s.zs[c] + cx * s.zs[ca]) * p.rmask[c];
  }
  return s2d_zeta_new(p, s, ca) * p.rmask[c];
}
__device__ __noinline__ double ob_zn(const KP& p, const S2D& s, int i, int j) {
  const bool we = (p.wwall && i < p.istr) || (p.ewall && i > p.iend), sn = (p.swall && j < p.jstr) || (p.nwall && j > p.jend);
  if (we && sn) {     // zetabc.F:206-222: mean of the two neighbouring edge values
    const int ia = i < p.istr ? p.istr : p.iend, ja = j < p.jstr ? p.jstr : p.jend;
    return 0.5 * (ob_zn_edge(p, s, ia, j) + ob_zn_edge(p, s, i, ja));
  }
  return ob_zn_edge(p, s, i, j);
}
// the characteristic value of the velocity normal to an open edge (u2dbc_im.F:29-56): zin/zout = zeta(kstp) on the interior/exterior side of
// the face, zin_new = zeta(knew) inside, vb = the face's own velocity, vin = the next face inwards (kstp); low: western/southern edge
__device__ __forceinline__ double ob_flather(double dtfast, double g, double hout, double hin, double pout, double pin, double zin, double zout, double zin_new,
                                             double vb, double vin, double vbry, double zbry, bool low) {
  const double cff = 0.5 * (hout + hin);
  const double hx = sqrt(g / cff);
  const double cx = dtfast * cff * hx * 0.5 * (pout + pin);
  double zx = (0.5 + cx) * zin + (0.5 - cx) * zout;
  if (cx > 0.292893218813452) {
    const double q = 1. - 0.292893218813452 / cx;
    zx = zx + (zin_new + cx * zout - (1. + cx) * zin) * (q * q);
  }
  return low ? 0.5 * ((1. - cx) * vb + cx * vin + vbry - hx * (zx - zbry)) : 0.5 * ((1. - cx) * vb + cx * vin + vbry + hx * (zx - zbry));
}
// upstream advection of the component tangential to an open edge plus binding to the data on inflow (u2dbc_im.F:293-330 and its mirrors):
// qb, qa = the point's own value and its interior neighbour's (old level), gb0/gb1, ga0/ga1 = the along-edge differences either side of them
__device__ __forceinline__ double ob_tang(double cx, double cy, double qb, double qa, double gb0, double gb1, double ga0, double ga1, double bry) {
  double cext;
  if (cx > 0.) cext = 0.; else { cext = -cx; cx = 0.; }
  double r = (1. - cx) * (qb - ob_max0(cy) * gb0 - ob_min0(cy) * gb1) + cx * (qa - ob_max0(cy) * ga0 - ob_min0(cy) * ga1);
  r = (1. - cext) * r + cext * bry;
  return r;
}
__device__ __noinline__ double ob_vn(const KP& p, const S2D& s, int i, int j);
// ubar(knew) at any u-point of the extended range (u2dbc_im.F)
__device__ __noinline__ double ob_un(const KP& p, const S2D& s, int i, int j) {
  const ix2_t c = CIDX(i, j), SX = p.SX;
  const bool srow = p.swall && j == p.jstr - 1, nrow = p.nwall && j == p.jend + 1;
  const bool wface = p.wwall && i == p.istr, eface = p.ewall && i == p.iend + 1;
  if (srow | nrow) {
    const int side = srow ? 2 : 3, ja = srow ? p.jstr : p.jend;
    if (wface | eface) {                       // the u-points in the corners of the basin
      if (ob_open(p, side)) {
        if (!ob_open(p, wface ? 0 : 1)) return s.un[c];          // open row beside a closed column: never set by the reference
        return 0.5 * (ob_un(p, s, wface ? i + 1 : i - 1, j) + ob_un(p, s, i, ja));   // u2dbc_im.F:453-476
      }
      return p.gamma2 * ob_un(p, s, i, ja) * p.umask[c];         // closed row: its loop runs over istr:iendR after the W/E edges are set
    }
    if (!ob_open(p, side)) return p.gamma2 * s.un[CIDX(i, ja)] * p.umask[c];
    if (p.obc_m2) return (p.obc_m2 == 1 ? OB_B1(p.bry_ubar[side], i) : s.un[CIDX(i, ja)]) * p.umask[c];   // specified / gradient (u2dbc_im.F:337-349)
    const ix2_t ca = CIDX(i, ja), cv = srow ? ca : c;            // the v-row on the edge: jstr (south), jend+1 (north)
    const double psn = srow ? (p.pn[c] + p.pn[c - 1] + p.pn[ca] + p.pn[ca - 1]) : (p.pn[ca] + p.pn[ca - 1] + p.pn[c] + p.pn[c - 1]);
    const double psm = srow ? (p.pm[c] + p.pm[c - 1] + p.pm[ca] + p.pm[ca - 1]) : (p.pm[ca] + p.pm[ca - 1] + p.pm[c] + p.pm[c - 1]);
    const double cx = (srow ? -0.125 : 0.125) * p.dtfast * (s.vs[cv] + s.vs[cv - 1]) * psn;
    const double cy = 0.125 * p.dtfast * (srow ? (s.us[c] + s.us[ca]) : (s.us[ca] + s.us[c])) * psm;
    const double r = ob_tang(cx, cy, s.us[c], s.us[ca], s.us[c] - s.us[c - 1], s.us[c + 1] - s.us[c], s.us[ca] - s.us[ca - 1], s.us[ca + 1] - s.us[ca],
                             OB_B1(p.bry_ubar[side], i));
    return r * p.umask[c];
  }
  if (wface) {
    if (!ob_open(p, 0)) return 0.0;
    if (p.obc_m2) return (p.obc_m2 == 1 ? OB_B1(p.bry_ubar[0], j) : s.un[c + 1]) * p.umask[c];            // u2dbc_im.F:130-148
    return ob_flather(p.dtfast, p.g, p.h[c - 1], p.h[c], p.pm[c - 1], p.pm[c], s.zs[c], s.zs[c - 1], s2d_zeta_new(p, s, c), s.us[c], s.us[c + 1],
                      OB_B1(p.bry_ubar[0], j), OB_B1(p.bry_zeta[0], j), true) * p.umask[c];
  }
  if (eface) {
    if (!ob_open(p, 1)) return 0.0;
    if (p.obc_m2) return (p.obc_m2 == 1 ? OB_B1(p.bry_ubar[1], j) : s.un[c - 1]) * p.umask[c];
    return ob_flather(p.dtfast, p.g, p.h[c], p.h[c - 1], p.pm[c], p.pm[c - 1], s.zs[c - 1], s.zs[c], s2d_zeta_new(p, s, c - 1), s.us[c], s.us[c - 1],
                      OB_B1(p.bry_ubar[1], j), OB_B1(p.bry_zeta[1], j), false) * p.umask[c];
  }
  (void)SX;
  return s.un[c];
}
// vbar(knew) at any v-point of the extended range (v2dbc_im.F)
__device__ __noinline__ double ob_vn(const KP& p, const S2D& s, int i, int j) {
  const ix2_t c = CIDX(i, j), SX = p.SX;
  const bool wcol = p.wwall && i == p.istr - 1, ecol = p.ewall && i == p.iend + 1;
  const bool sface = p.swall && j == p.jstr, nface = p.nwall && j == p.jend + 1;
  if (wcol | ecol) {
    const int side = wcol ? 0 : 1, ia = wcol ? p.istr : p.iend;
    if (sface | nface) {
      if (ob_open(p, side)) {
        if (!ob_open(p, sface ? 2 : 3)) return s.vn[c];
        return 0.5 * (ob_vn(p, s, i, sface ? j + 1 : j - 1) + ob_vn(p, s, ia, j));   // v2dbc_im.F corners
      }
      return p.gamma2 * ob_vn(p, s, ia, j) * p.vmask[c];
    }
    if (!ob_open(p, side)) return p.gamma2 * s.vn[CIDX(ia, j)] * p.vmask[c];
    if (p.obc_m2) return (p.obc_m2 == 1 ? OB_B1(p.bry_vbar[side], j) : s.vn[CIDX(ia, j)]) * p.vmask[c];
    const ix2_t ca = CIDX(ia, j), cu = wcol ? ca : c;            // the u-column on the edge: istr (west), iend+1 (east)
    const double psm = wcol ? (p.pm[c] + p.pm[c - SX] + p.pm[ca] + p.pm[ca - SX]) : (p.pm[ca] + p.pm[ca - SX] + p.pm[c] + p.pm[c - SX]);
    const double psn = wcol ? (p.pn[c] + p.pn[c - SX] + p.pn[ca] + p.pn[ca - SX]) : (p.pn[ca] + p.pn[ca - SX] + p.pn[c] + p.pn[c - SX]);
    const double cx = (wcol ? -0.125 : 0.125) * p.dtfast * (s.us[cu] + s.us[cu - SX]) * psm;
    const double cy = 0.125 * p.dtfast * (wcol ? (s.vs[c] + s.vs[ca]) : (s.vs[ca] + s.vs[c])) * psn;
    const double r = ob_tang(cx, cy, s.vs[c], s.vs[ca], s.vs[c] - s.vs[c - SX], s.vs[c + SX] - s.vs[c], s.vs[ca] - s.vs[ca - SX], s.vs[ca + SX] - s.vs[ca],
                             OB_B1(p.bry_vbar[side], j));
    return r * p.vmask[c];
  }
  if (sface) {
    if (!ob_open(p, 2)) return 0.0;
    if (p.obc_m2) return (p.obc_m2 == 1 ? OB_B1(p.bry_vbar[2], i) : s.vn[c + SX]) * p.vmask[c];
    return ob_flather(p.dtfast, p.g, p.h[c - SX], p.h[c], p.pn[c - SX], p.pn[c], s.zs[c], s.zs[c - SX], s2d_zeta_new(p, s, c), s.vs[c], s.vs[c + SX],
                      OB_B1(p.bry_vbar[2], i), OB_B1(p.bry_zeta[2], i), true) * p.vmask[c];
  }
  if (nface) {
    if (!ob_open(p, 3)) return 0.0;
    if (p.obc_m2) return (p.obc_m2 == 1 ? OB_B1(p.bry_vbar[3], i) : s.vn[c - SX]) * p.vmask[c];
    return ob_flather(p.dtfast, p.g, p.h[c], p.h[c - SX], p.pn[c], p.pn[c - SX], s.zs[c - SX], s.zs[c], s2d_zeta_new(p, s, c - SX), s.vs[c], s.vs[c - SX],
                      OB_B1(p.bry_vbar[3], i), OB_B1(p.bry_zeta[3], i), false) * p.vmask[c];
  }
  return s.vn[c];
}
// One thread per boundary point of the barotropic substep, launched after the interior kernel (which in this mode leaves every
// boundary point alone).  blockIdx.y: 0 southern ring row, 1 northern ring row (rho index, i = istrR:iendR), 2 western ring column,
// 3 eastern ring column (j = jstr:jend), 4 the u-faces on the western edge (istr, j = jstr:jend), 5 the v-faces on the southern
// edge (i = istr:iend, jstr).  A ring point owns zeta(knew), Zt_avg1, DU_avg2, DV_avg2 (and the first-substep resets) at its index
// plus the u- and v-values stored at that index where the reference defines them, with their DU_avg1 / DV_avg1 terms.
__global__ void k_step2d_bry(const __grid_constant__ KP p, const __grid_constant__ S2D s) {
  const int list = blockIdx.y, q = blockIdx.x * blockDim.x + threadIdx.x;
  const double cff1 = 0.5 * s.w1;
  const ix2_t SX = p.SX;
  int i, j;
  if (list < 2) {
    if (list == 0 ? !p.swall : !p.nwall) return;
    i = p.istrR + q; j = list == 0 ? p.jstr - 1 : p.jend + 1;
    if (i > p.iendR) return;
  } else if (list < 4) {
    if (list == 2 ? !p.wwall : !p.ewall) return;
    j = p.jstr + q; i = list == 2 ? p.istr - 1 : p.iend + 1;
    if (j > p.jend) return;
  } else if (list == 4) {
    if (!p.wwall) return;
    i = p.istr; j = p.jstr + q;
    if (j > p.jend) return;
    const ix2_t c = CIDX(i, j);
    const double un = ob_un(p, s, i, j);
    s2d_store(p, s.un, i, j, c, un);
    const double Dn = p.h[c] + ob_zn(p, s, i, j), Dnw = p.h[c - 1] + ob_zn(p, s, i - 1, j);
    const double base = s.iif == 1 ? 0.0 : p.DU_avg1[c];
    p.DU_avg1[c] = base + cff1 * (Dn + Dnw) * (un)*p.dn_u[c];
    return;
  } else {
    if (!p.swall) return;
    j = p.jstr; i = p.istr + q;
    if (i > p.iend) return;
    const ix2_t c = CIDX(i, j);
    const double vn = ob_vn(p, s, i, j);
    s2d_store(p, s.vn, i, j, c, vn);
    const double Dn = p.h[c] + ob_zn(p, s, i, j), Dns = p.h[c - SX] + ob_zn(p, s, i, j - 1);
    const double base = s.iif == 1 ? 0.0 : p.DV_avg1[c];
    p.DV_avg1[c] = base + cff1 * (Dn + Dns) * (vn)*p.dm_v[c];
    return;
  }
  // ---- ring point
  const ix2_t c = CIDX(i, j);
  const double zn = ob_zn(p, s, i, j);
  // DUon at istr-1 beside a western edge and DVom at jstr-1 beside a southern edge are never filled by the reference (zero scratch)
  const double DUon = (p.wwall && i < p.istr) ? 0.0 : s2d_DUon(p, s, c);
  const double DVom = (p.swall && j < p.jstr) ? 0.0 : s2d_DVom(p, s, c);
  double DU1, DV1, Zt;
  s2d_avg_out(p, s, i, j, c, zn, DUon, DVom, DU1, DV1, Zt);
  const bool has_u = (list < 2) ? (i >= p.istr) : (list == 3);      // rows: istr:iendR; the eastern column is the u-face iend+1
  const bool has_v = (list < 2) ? (list == 1) : true;               // the northern row is the v-face jend+1; both columns carry the tangential v
  if (has_u) {
    const double un = ob_un(p, s, i, j);
    s2d_store(p, s.un, i, j, c, un);
    const double Dn = p.h[c] + zn, Dnw = p.h[c - 1] + ob_zn(p, s, i - 1, j);
    p.DU_avg1[c] = DU1 + cff1 * (Dn + Dnw) * (un)*p.dn_u[c];
  } else if (s.iif == 1) p.DU_avg1[c] = DU1;
  if (has_v) {
    const double vn = ob_vn(p, s, i, j);
    s2d_store(p, s.vn, i, j, c, vn);
    const double Dn = p.h[c] + zn, Dns = p.h[c - SX] + ob_zn(p, s, i, j - 1);
    p.DV_avg1[c] = DV1 + cff1 * (Dn + Dns) * (vn)*p.dm_v[c];
  } else if (s.iif == 1) p.DV_avg1[c] = DV1;
  if (s.iif == s.nfast) s2d_store(p, s.zn, i, j, c, Zt);
}

// ---- 3-D ------------------------------------------------------------------------------------------------------------------
// fields of one 3-D boundary pass: a = the array being finished (nnew), as = the same quantity at nstp, ur/vr = u,v at nrhs
struct OB3 {
  double* a;
  const double *as, *ur, *vr;
  double dtfwd;      // dt/2 in the predictor stage (nnew == 3), dt in the corrector (u3dbc_im.F:36-40)
  int itrc;          // tracer index (0-based) of a BC_T pass
};
// Orlanski-type radiation of the component normal to an open edge with binding to the data on inflow (u3dbc_im.F:47-119 and mirrors)
__device__ __forceinline__ double ob_orlanski(double qb_s, double q1_s, double q1_n, double q2_n, double g1a, double g1b, double gba, double gbb, double bry,
                                              bool bry_inward, double ubind, double dtfwd, double pscale) {
  const double eps = 1.E-33;
  const double dft = q1_s - q1_n, dfx = q1_n - q2_n;
  const double dfy = (dft * (g1a + g1b) > 0.) ? g1a : g1b;
  const double sq = dfx * dfx + dfy * dfy;
  const double cff = (sq < eps) ? eps : sq;                     // max(dfx*dfx+dfy*dfy, eps)
  const double t1 = dft * dfy;
  const double t2 = (t1 < -cff) ? -cff : t1;                    // max(dft*dfy, -cff)
  double cy = (t2 < cff) ? t2 : cff;                            // min(cff, .)
  double cx = dft * dfx, cext;
  if (cx < 0.) {
    cx = 0.; cy = 0.;
    cext = bry_inward ? fabs(bry) : ubind;
    cext = cext * dtfwd * pscale;
  } else cext = 0.;
  double r = (cff * qb_s + cx * q1_n - ob_max0(cy) * gba - ob_min0(cy) * gbb) / (cff + cx);
  r = (1. - cext) * r + cext * bry;
  return r;
}
__device__ __noinline__ double ob3_u(const KP& p, const OB3& o, int i, int j, int k) {
  const ix_t c = CIDX(i, j), SX = p.SX;
  const bool srow = p.swall && j == p.jstr - 1, nrow = p.nwall && j == p.jend + 1;
  const bool wface = p.wwall && i == p.istr, eface = p.ewall && i == p.iend + 1;
  if (srow | nrow) {
    const int side = srow ? 2 : 3, ja = srow ? p.jstr : p.jend;
    if (wface | eface) {
      if (ob_open(p, side)) {
        if (!ob_open(p, wface ? 0 : 1)) return R3(o.a, c, k);
        return 0.5 * (ob3_u(p, o, wface ? i + 1 : i - 1, j, k) + ob3_u(p, o, i, ja, k));      // u3dbc_im.F:397-421
      }
      return p.gamma2 * ob3_u(p, o, i, ja, k) * p.umask[c];
    }
    if (!ob_open(p, side)) return p.gamma2 * R3(o.a, CIDX(i, ja), k) * p.umask[c];
    if (p.obc_m3) return (p.obc_m3 == 1 ? OB_B2(p.bry_u[side], i, k, BRYL(p, side)) : R3(o.a, CIDX(i, ja), k)) * p.umask[c];   // specified / gradient
    const ix_t ca = CIDX(i, ja), cv = srow ? ca : c;
    const double psn = p.pn[c] + p.pn[c - 1] + p.pn[ca] + p.pn[ca - 1];                        // u3dbc_im.F:212-213,259-260: boundary row first on both edges
    const double psm = p.pm[c] + p.pm[c - 1] + p.pm[ca] + p.pm[ca - 1];
    const double cx = (srow ? -0.125 : 0.125) * o.dtfwd * (R3(o.vr, cv, k) + R3(o.vr, cv - 1, k)) * psn;
    const double cy = 0.125 * o.dtfwd * (srow ? (R3(o.ur, c, k) + R3(o.ur, ca, k)) : (R3(o.ur, ca, k) + R3(o.ur, c, k))) * psm;
    const double r = ob_tang(cx, cy, R3(o.as, c, k), R3(o.as, ca, k), R3(o.as, c, k) - R3(o.as, c - 1, k), R3(o.as, c + 1, k) - R3(o.as, c, k),
                             R3(o.as, ca, k) - R3(o.as, ca - 1, k), R3(o.as, ca + 1, k) - R3(o.as, ca, k), OB_B2(p.bry_u[side], i, k, BRYL(p, side)));
    return r * p.umask[c];
  }
  if (wface | eface) {
    if (!ob_open(p, wface ? 0 : 1)) return 0.0;
    if (p.obc_m3) return (p.obc_m3 == 1 ? OB_B2(p.bry_u[wface ? 0 : 1], j, k, BRYL(p, 0)) : R3(o.a, c + (wface ? 1 : -1), k)) * p.umask[c];   // u3dbc_im.F:121-131
    const int d = wface ? 1 : -1;                // towards the interior
    const ix_t c1 = c + d, c2 = c + 2 * d;
    // along-edge differences at nstp, masked at the psi-points (u3dbc_im.F:50-57): grad(i,j) = (u(i,j)-u(i,j-1))*pmask(i,j)
    const double g1a = (R3(o.as, c1, k) - R3(o.as, c1 - SX, k)) * p.pmask[c1], g1b = (R3(o.as, c1 + SX, k) - R3(o.as, c1, k)) * p.pmask[c1 + SX];
    const double gba = (R3(o.as, c, k) - R3(o.as, c - SX, k)) * p.pmask[c], gbb = (R3(o.as, c + SX, k) - R3(o.as, c, k)) * p.pmask[c + SX];
    const double b = OB_B2(p.bry_u[wface ? 0 : 1], j, k, BRYL(p, 0));
    const double r = ob_orlanski(R3(o.as, c, k), R3(o.as, c1, k), R3(o.a, c1, k), R3(o.a, c2, k), g1a, g1b, gba, gbb, b, wface ? (b > 0.) : (b < 0.), p.ubind, o.dtfwd,
                                 0.5 * (p.pm[c - 1] + p.pm[c]));
    return r * p.umask[c];
  }
  return R3(o.a, c, k);
}
__device__ __noinline__ double ob3_v(const KP& p, const OB3& o, int i, int j, int k) {
  const ix_t c = CIDX(i, j), SX = p.SX;
  const bool wcol = p.wwall && i == p.istr - 1, ecol = p.ewall && i == p.iend + 1;
  const bool sface = p.swall && j == p.jstr, nface = p.nwall && j == p.jend + 1;
  if (wcol | ecol) {
    const int side = wcol ? 0 : 1, ia = wcol ? p.istr : p.iend;
    if (sface | nface) {
      if (ob_open(p, side)) {
        if (!ob_open(p, sface ? 2 : 3)) return R3(o.a, c, k);
        return 0.5 * (ob3_v(p, o, i, sface ? j + 1 : j - 1, k) + ob3_v(p, o, ia, j, k));      // v3dbc_im.F:406-430
      }
      return p.gamma2 * ob3_v(p, o, ia, j, k) * p.vmask[c];
    }
    if (!ob_open(p, side)) return p.gamma2 * R3(o.a, CIDX(ia, j), k) * p.vmask[c];
    if (p.obc_m3) return (p.obc_m3 == 1 ? OB_B2(p.bry_v[side], j, k, BRYL(p, side)) : R3(o.a, CIDX(ia, j), k)) * p.vmask[c];
    const ix_t ca = CIDX(ia, j), cu = wcol ? ca : c;
    const double psm = p.pm[c] + p.pm[c - SX] + p.pm[ca] + p.pm[ca - SX];                      // v3dbc_im.F: boundary column first on both edges
    const double psn = p.pn[c] + p.pn[c - SX] + p.pn[ca] + p.pn[ca - SX];
    const double cx = (wcol ? -0.125 : 0.125) * o.dtfwd * (R3(o.ur, cu, k) + R3(o.ur, cu - SX, k)) * psm;
    const double cy = 0.125 * o.dtfwd * (wcol ? (R3(o.vr, c, k) + R3(o.vr, ca, k)) : (R3(o.vr, ca, k) + R3(o.vr, c, k))) * psn;
    const double r = ob_tang(cx, cy, R3(o.as, c, k), R3(o.as, ca, k), R3(o.as, c, k) - R3(o.as, c - SX, k), R3(o.as, c + SX, k) - R3(o.as, c, k),
                             R3(o.as, ca, k) - R3(o.as, ca - SX, k), R3(o.as, ca + SX, k) - R3(o.as, ca, k), OB_B2(p.bry_v[side], j, k, BRYL(p, side)));
    return r * p.vmask[c];
  }
  if (sface | nface) {
    if (!ob_open(p, sface ? 2 : 3)) return 0.0;
    if (p.obc_m3) return (p.obc_m3 == 1 ? OB_B2(p.bry_v[sface ? 2 : 3], i, k, BRYL(p, 2)) : R3(o.a, c + (sface ? SX : -SX), k)) * p.vmask[c];
    const ix_t d = sface ? SX : -SX;
    const ix_t c1 = c + d, c2 = c + 2 * d;
    // grad(i,j) = (v(i,j)-v(i-1,j))*pmask(i,j) (v3dbc_im.F:46-51)
    const double g1a = (R3(o.as, c1, k) - R3(o.as, c1 - 1, k)) * p.pmask[c1], g1b = (R3(o.as, c1 + 1, k) - R3(o.as, c1, k)) * p.pmask[c1 + 1];
    const double gba = (R3(o.as, c, k) - R3(o.as, c - 1, k)) * p.pmask[c], gbb = (R3(o.as, c + 1, k) - R3(o.as, c, k)) * p.pmask[c + 1];
    const double b = OB_B2(p.bry_v[sface ? 2 : 3], i, k, BRYL(p, 2));
    const double r = ob_orlanski(R3(o.as, c, k), R3(o.as, c1, k), R3(o.a, c1, k), R3(o.a, c2, k), g1a, g1b, gba, gbb, b, sface ? (b > 0.) : (b < 0.), p.ubind, o.dtfwd,
                                 0.5 * (p.pn[c - SX] + p.pn[c]));
    return r * p.vmask[c];
  }
  return R3(o.a, c, k);
}
// tracer on an edge line (not a corner): t3dbc_im.F:36-277
__device__ __noinline__ double ob3_t_edge(const KP& p, const OB3& o, int i, int j, int k) {
  const ix_t c = CIDX(i, j), SX = p.SX;
  const bool w = p.wwall && i < p.istr, e = p.ewall && i > p.iend, so = p.swall && j < p.jstr, n = p.nwall && j > p.jend;
  if (!(w | e | so | n)) return R3(o.a, c, k);
  const int side = w ? 0 : (e ? 1 : (so ? 2 : 3));
  const ix_t ca = w ? c + 1 : (e ? c - 1 : (so ? c + SX : c - SX));
  if (!ob_open(p, side) || p.obc_t == 2) return R3(o.a, ca, k) * p.rmask[c];     // wall, or an open edge without a switch: gradient
  const double b = OB_B2(p.bry_t[side] + (size_t)o.itrc * p.N * BRYL(p, side), side < 2 ? j : i, k, BRYL(p, side));
  if (p.obc_t == 1) return b * p.rmask[c];                                         // OBC_TSPECIFIED (t3dbc_im.F:85-96)
  double cx, cy, gb0, gb1, ga0, ga1;
  if (side < 2) {
    const ix_t cu = w ? ca : c;            // the u-face on the edge: istr (west), iend+1 (east)
    cx = (w ? -o.dtfwd : o.dtfwd) * R3(o.ur, cu, k) * p.pm[c];
    cy = 0.5 * o.dtfwd * (R3(o.vr, c, k) + R3(o.vr, c + SX, k)) * p.pn[c];
    gb0 = (R3(o.as, c, k) - R3(o.as, c - SX, k)) * p.vmask[c]; gb1 = (R3(o.as, c + SX, k) - R3(o.as, c, k)) * p.vmask[c + SX];
    ga0 = (R3(o.as, ca, k) - R3(o.as, ca - SX, k)) * p.vmask[ca]; ga1 = (R3(o.as, ca + SX, k) - R3(o.as, ca, k)) * p.vmask[ca + SX];
  } else {
    const ix_t cv = so ? ca : c;           // the v-face on the edge: jstr (south), jend+1 (north)
    cx = (so ? -o.dtfwd : o.dtfwd) * R3(o.vr, cv, k) * p.pn[c];
    cy = 0.5 * o.dtfwd * (R3(o.ur, c, k) + R3(o.ur, c + 1, k)) * p.pm[c];
    gb0 = (R3(o.as, c, k) - R3(o.as, c - 1, k)) * p.umask[c]; gb1 = (R3(o.as, c + 1, k) - R3(o.as, c, k)) * p.umask[c + 1];
    ga0 = (R3(o.as, ca, k) - R3(o.as, ca - 1, k)) * p.umask[ca]; ga1 = (R3(o.as, ca + 1, k) - R3(o.as, ca, k)) * p.umask[ca + 1];
  }
  // interior neighbour: nstp on the west and south, nnew on the east and north (t3dbc_im.F:66 vs :131, :195 vs :260)
  const double qa = (w | so) ? R3(o.as, ca, k) : R3(o.a, ca, k);
  return ob_tang(cx, cy, R3(o.as, c, k), qa, gb0, gb1, ga0, ga1, b) * p.rmask[c];
}
__device__ __noinline__ double ob3_t(const KP& p, const OB3& o, int i, int j, int k) {
  const bool we = (p.wwall && i < p.istr) || (p.ewall && i > p.iend), sn = (p.swall && j < p.jstr) || (p.nwall && j > p.jend);
  if (we && sn) {     // t3dbc_im.F:281-344: mask-weighted mean of the two neighbouring edge values
    const int ia = i < p.istr ? p.istr : p.iend, ja = j < p.jstr ? p.jstr : p.jend;
    const double ma = p.rmask[CIDX(ia, j)], mb = p.rmask[CIDX(i, ja)];
    double cff = ma + mb;
    if (cff > 0.0) {
      cff = 1.0 / cff;
      return cff * (ma * ob3_t_edge(p, o, ia, j, k) + mb * ob3_t_edge(p, o, i, ja, k));
    }
    return 0.0;
  }
  return ob3_t_edge(p, o, i, j, k);
}
// One thread per boundary point and level.  blockIdx.y: 0 southern row, 1 northern row, 2 western column, 3 eastern column of the
// field's own boundary points; blockIdx.z: level.  KIND as in k_bc3d.
template <int KIND>
__global__ void k_bry3d(const __grid_constant__ KP p, const __grid_constant__ OB3 o) {
  const int list = blockIdx.y, q = blockIdx.x * blockDim.x + threadIdx.x, k = blockIdx.z + 1;
  int i, j;
  if (KIND == BC_U) {
    if (list < 2) {            // tangential rows jstr-1 / jend+1: i = istr:iendR (with a western edge the first one is the corner point)
      if (list == 0 ? !p.swall : !p.nwall) return;
      i = p.istr + q; j = list == 0 ? p.jstr - 1 : p.jend + 1;
      if (i > p.iendR) return;
    } else {                   // the faces on the western / eastern edge, j = jstr:jend
      if (list == 2 ? !p.wwall : !p.ewall) return;
      j = p.jstr + q; i = list == 2 ? p.istr : p.iend + 1;
      if (j > p.jend) return;
    }
    R3(o.a, CIDX(i, j), k) = ob3_u(p, o, i, j, k);
  } else if (KIND == BC_V) {
    if (list < 2) {            // the faces on the southern / northern edge, i = istr:iend
      if (list == 0 ? !p.swall : !p.nwall) return;
      i = p.istr + q; j = list == 0 ? p.jstr : p.jend + 1;
      if (i > p.iend) return;
    } else {                   // tangential columns istr-1 / iend+1: j = jstr:jendR
      if (list == 2 ? !p.wwall : !p.ewall) return;
      j = p.jstr + q; i = list == 2 ? p.istr - 1 : p.iend + 1;
      if (j > p.jendR) return;
    }
    R3(o.a, CIDX(i, j), k) = ob3_v(p, o, i, j, k);
  } else {
    if (list < 2) {            // rows incl. the corners
      if (list == 0 ? !p.swall : !p.nwall) return;
      i = p.istrR + q; j = list == 0 ? p.jstr - 1 : p.jend + 1;
      if (i > p.iendR) return;
    } else {
      if (list == 2 ? !p.wwall : !p.ewall) return;
      j = p.jstr + q; i = list == 2 ? p.istr - 1 : p.iend + 1;
      if (j > p.jend) return;
    }
    R3(o.a, CIDX(i, j), k) = ob3_t(p, o, i, j, k);
  }
}
