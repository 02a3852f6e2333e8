// romsb200.cu -- C ABI (include/roms_b200.h) of the B200-native UCLA-ROMS time step.
//
// Host side of the drop-in boundary: owns the device arena, moves fields between the Fortran
// module arrays' layout and the padded HBM layout, keeps the loop-control scalars of scalars.F,
// and exposes one export per reference tile subroutine (main.F:397-479 call order in roms_step).
// There is no CPU fallback: every numerical export launches CUDA kernels or fails.
#include <cuda_runtime.h>
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "../../include/roms_b200.h"
#include "kernels.cuh"
#include "comm.h"

static thread_local std::string g_err;
#define CK(call)                                                                              \
  do {                                                                                        \
    cudaError_t e_ = (call);                                                                  \
    if (e_ != cudaSuccess) {                                                                  \
      g_err = std::string(#call) + ": " + cudaGetErrorString(e_);                             \
      return -(int)e_;                                                                        \
    }                                                                                         \
  } while (0)

struct KTimer {
  const char* name;
  cudaEvent_t a, b;
};
struct FieldSlot {
  std::vector<double*> ptr;   // [lvl*ntrc + itrc]
  int nlvl = 1, ntrc = 1, nk = 1;
};

struct roms_ctx {
  roms_cfg c;
  int device;
  KP kp;
  cudaStream_t st;
  cudaStream_t st2 = nullptr;                      // high-priority stream of the ring kernels and halo swaps of the barotropic loop
  cudaEvent_t ev_int[2] = {nullptr, nullptr}, ev_edge[2] = {nullptr, nullptr}, ev_join = nullptr;
  int t_tma = 1;      // horizontal tracer kernels with TMA-staged tiles (tracer_tma.cuh); 0: the scalar-load kernels
  TmaMaps tmaps;      // three tensor maps over the whole arena (one per box shape)
  int tmaps_ok = 0;
  int overlap = 1;    // multi-rank: barotropic halo swaps overlapped with the interior launch of the next substep (ROMS_OVERLAP=0: serial)
  FieldSlot f[ROMS_NFIELDS];
  double* arena = nullptr;
  size_t arena_doubles = 0;
  // scratch planes for diag
  double *d_ub, *d_vb, *d_o[5];
  double* d_rows = nullptr;   // [7][Mm] row partials of diag + 8 results
  // scalars.F
  int iic = 0, iif = 1, nstp = 1, nrhs = 1, nnew = 1, kstp = 1, knew = 1, ntstart = 1, forw_start = 1;
  int nfast = 0;
  std::vector<double> w1, w2, Cs_w, Cs_r;
  long launches = 0;
  // optional per-kernel CUDA-event timing (bench.py roofline pass)
  int ktiming = 0;
  std::vector<KTimer> ktimers;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  int col_bx = 32, col_by = 2, pl_bx = 32, pl_by = 8;   // plane tiles 32x8: fewest halo-row re-reads through L2 (profiles/r1_tuning_log.txt)
  double* bry_buf = nullptr;   // boundary data of the open edges (ROMS_ZETA_WEST ... ROMS_T_NORTH), allocated when an edge is open
  size_t bry_doubles = 0;
  double* kpp_buf = nullptr;   // LMD/KPP arrays (roms_set_kpp)
  double *kpp_Rig = nullptr, *kpp_RigS = nullptr, *kpp_FC = nullptr, *kpp_hblS = nullptr, *kpp_bblS = nullptr, *kpp_alpha = nullptr, *kpp_beta = nullptr;
  int kpp_convec = 1;     // cppdefs LMD_CONVEC
  double kpp_Akv_bak = 0.0, kpp_Akt_bak[2] = {0.0, 0.0};   // vertical_mixing keyword
  int fuse_rho = 1;       // a rho_eos call is deferred; if prsgrd follows (main.F:397-413, 430-445) both run as one fused column kernel
  int rho_pending = 0;    // time index of the deferred rho_eos (0: none)
  int uv2_sm = 1;     // k_uv2_pass1 with the column parked in shared memory (1.0 -> 0.78 ms on B200); off when the column is too tall
  int s2d_tma = 12;   // barotropic kernel: 12 k_step2d_hyb<64,4> (shipped), 0 tile kernel k_step2d, 1-5 k_step2d_tma shapes, 6-11 k_step2d_march, 13-15 other hyb shapes
  int s2d_rows = 0;   // rows per block of the marching kernels; 0: one-wave rule for k_step2d_hyb, 16 for the others
  int n_sm = 148;
  int uv2_fused = 1;  // step3d_uv2 passes 0 and 1 in one shared-memory column kernel (k_uv2_sm)
  int tb_max = 4;     // most tracers per launch of the tracer kernels (ROMS_TB = 1, 2 or 4)
  int use_smem = 0;   // measured on B200 (profiles/r1_scratch_policy.txt): occupancy beats shared-memory scratch
  Comm comm;
  double* u(int l) { return f[ROMS_U].ptr[l - 1]; }
  double* v(int l) { return f[ROMS_V].ptr[l - 1]; }
  double* t(int l, int itrc) { return f[ROMS_T].ptr[(l - 1) * c.NT + (itrc - 1)]; }
  double* zeta(int l) { return f[ROMS_ZETA].ptr[l - 1]; }
  double* ubar(int l) { return f[ROMS_UBAR].ptr[l - 1]; }
  double* vbar(int l) { return f[ROMS_VBAR].ptr[l - 1]; }
};

// ------------------------------------------------------------------------------------------------
// launch geometry
static dim3 grid_plane(const roms_ctx* h, int i1, int j0, int j1, int nk, dim3 b) {
  return dim3((i1 + IOFF + 1 + b.x - 1) / b.x, (j1 - j0 + 1 + b.y - 1) / b.y, nk);
}
static void kt_begin(roms_ctx* h, const char* name, cudaStream_t st = nullptr);
static void kt_end(roms_ctx* h, cudaStream_t st = nullptr);
static int flush_rho(roms_ctx* h);
#define FLUSH_RHO(h) do { int rc_ = flush_rho(h); if (rc_) return rc_; } while (0)
// plane kernel whose threads each compute JB consecutive rows
#define PLANE_JB(kern, JB, i1, j0, j1, nk, ...)                                                  \
  do {                                                                                        \
    dim3 b_(h->pl_bx, h->pl_by, 1);                                                           \
    if (h->ktiming) kt_begin(h, #kern);                                                       \
    kern<<<grid_plane(h, i1, 0, ((j1) - (j0) + JB) / JB - 1, nk, b_), b_, 0, h->st>>>(__VA_ARGS__); \
    if (h->ktiming) kt_end(h);                                                                \
    h->launches++;                                                                            \
  } while (0)
#define PLANE(kern, i1, j0, j1, nk, ...)                                                      \
  do {                                                                                        \
    dim3 b_(h->pl_bx, h->pl_by, 1);                                                           \
    if (h->ktiming) kt_begin(h, #kern);                                                       \
    kern<<<grid_plane(h, i1, j0, j1, nk, b_), b_, 0, h->st>>>(__VA_ARGS__);                   \
    if (h->ktiming) kt_end(h);                                                                \
    h->launches++;                                                                            \
  } while (0)
static size_t col_smem(const roms_ctx* h, int narr) {
  return h->use_smem ? (size_t)narr * (h->c.N + 1) * h->col_bx * h->col_by * sizeof(double) : 0;
}
#define COLUMN(kern, narr, i1, j0, j1, ...)                                                   \
  do {                                                                                        \
    dim3 b_(h->col_bx, h->col_by, 1);                                                         \
    size_t sm_ = col_smem(h, narr);                                                           \
    if (sm_ > 48 * 1024) CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_)); \
    if (h->ktiming) kt_begin(h, #kern);                                                       \
    kern<<<grid_plane(h, i1, j0, j1, 1, b_), b_, sm_, h->st>>>(__VA_ARGS__);                  \
    if (h->ktiming) kt_end(h);                                                                \
    h->launches++;                                                                            \
  } while (0)

// column kernel that parks its column in shared memory: one warp-row per block, two arrays [k][lane]
#define COLUMN_SM(kern, i1, j0, j1, ...)                                                      \
  do {                                                                                        \
    dim3 b_(32, 1, 1);                                                                        \
    size_t sm_ = (size_t)2 * (h->c.N + 1) * 32 * sizeof(double);                              \
    static size_t attr_ = 0;                                                                  \
    if (attr_ != sm_) { CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_)); attr_ = sm_; } \
    if (h->ktiming) kt_begin(h, #kern);                                                       \
    kern<<<grid_plane(h, i1, j0, j1, 1, b_), b_, sm_, h->st>>>(__VA_ARGS__);                  \
    if (h->ktiming) kt_end(h);                                                                \
    h->launches++;                                                                            \
  } while (0)

static void kt_begin(roms_ctx* h, const char* name, cudaStream_t st) {
  KTimer t;
  t.name = name;
  cudaEventCreate(&t.a);
  cudaEventCreate(&t.b);
  cudaEventRecord(t.a, st ? st : h->st);
  h->ktimers.push_back(t);
}
static void kt_end(roms_ctx* h, cudaStream_t st) { cudaEventRecord(h->ktimers.back().b, st ? st : h->st); }

// ------------------------------------------------------------------------------------------------
// set_scoord.F:4-66 and set_weights.F:7-246, evaluated on the host in the reference's operation order
static double cs_curve(double sc, double theta_s, double theta_b) {
  const double csrf = theta_s > 0.0 ? (1.0 - cosh(theta_s * sc)) / (cosh(theta_s) - 1.0) : -sc * sc;
  return theta_b > 0.0 ? (exp(theta_b * csrf) - 1.0) / (1.0 - exp(-theta_b)) : csrf;
}
static void host_scoord(roms_ctx* h) {
  const int N = h->c.N;
  const double ds = 1.0 / double(N);
  h->Cs_w.assign(N + 1, 0.0);
  h->Cs_r.assign(N + 1, 0.0);
  h->Cs_w[0] = -1.0;
  for (int k = 1; k < N; ++k) h->Cs_w[k] = cs_curve(ds * double(k - N), h->c.theta_s, h->c.theta_b);
  for (int k = 1; k <= N; ++k) h->Cs_r[k] = cs_curve(ds * (double(k - N) - 0.5), h->c.theta_s, h->c.theta_b);
}
static void host_weights(roms_ctx* h) {
  const int nd = h->c.ndtfast;
  std::vector<double>& a = h->w1;
  std::vector<double>& b = h->w2;
  a.assign(2 * nd + 2, 0.0);
  b.assign(2 * nd + 2, 0.0);
  const double p = 2.0, q = 4.0, r = 0.25;
  double scale = (p + 1.0) * (p + q + 1.0) / ((p + 2.0) * (p + q + 2.0) * double(nd));
  int nfast = 0;
  auto moments = [&](double& sum, double& shft, bool iFirst) {
    sum = 0.0; shft = 0.0;
    for (int i = 1; i <= nfast; ++i) {
      sum = sum + a[i];
      shft = shft + (iFirst ? double(i) * a[i] : a[i] * double(i));
    }
  };
  for (int iter = 0; iter < 16; ++iter) {            // shape function, 16 centring iterations
    nfast = 0;
    for (int i = 1; i <= 2 * nd; ++i) {
      const double cff = scale * double(i);
      a[i] = pow(cff, p) - pow(cff, p + q) - r * cff;
      if (a[i] > 0.0) nfast = i;
      if (nfast > 0 && a[i] < 0.0) a[i] = 0.0;
    }
    double sum, shft;
    moments(sum, shft, false);
    scale = scale * shft / (sum * double(nd));
  }
  for (int iter = 0; iter < nd; ++iter) {            // shift so that the centroid sits exactly at ndtfast
    double sum, shft;
    moments(sum, shft, true);
    shft = shft / sum;
    const double cff = double(nd) - shft;
    if (cff > 1.0) {
      nfast++;
      for (int i = nfast; i >= 2; --i) a[i] = a[i - 1];
      a[1] = 0.0;
    } else if (cff > 0.0) {
      const double s1 = 1.0 - cff;
      for (int i = nfast; i >= 2; --i) a[i] = s1 * a[i] + cff * a[i - 1];
      a[1] = s1 * a[1];
    } else if (cff < -1.0) {
      nfast--;
      for (int i = 1; i <= nfast; ++i) a[i] = a[i + 1];
      a[nfast + 1] = 0.0;
    } else if (cff < 0.0) {
      const double s1 = 1.0 + cff;
      for (int i = 1; i <= nfast - 1; ++i) a[i] = s1 * a[i] - cff * a[i + 1];
      a[nfast] = s1 * a[nfast];
    }
  }
  for (int j = 1; j <= nfast; ++j)                   // secondary weights
    for (int i = 1; i <= j; ++i) b[i] = b[i] + a[j];
  double sa = 0.0, sb = 0.0;
  for (int i = 1; i <= nfast; ++i) { sa = sa + a[i]; sb = sb + b[i]; }
  sa = 1.0 / sa; sb = 1.0 / sb;
  for (int i = 1; i <= nfast; ++i) { a[i] = sa * a[i]; b[i] = sb * b[i]; }
  h->nfast = nfast;
}

// ------------------------------------------------------------------------------------------------
static void slot(roms_ctx* h, int id, int nlvl, int ntrc, int nk) {
  h->f[id].nlvl = nlvl; h->f[id].ntrc = ntrc; h->f[id].nk = nk;
  h->f[id].ptr.assign((size_t)nlvl * ntrc, nullptr);
}

extern "C" {

const char* roms_last_error(void) { return g_err.c_str(); }

int roms_create(const roms_cfg* cfg, int device, roms_ctx** out) {
  if (!cfg || !out) { g_err = "roms_create: null argument"; return -1; }
  if (cfg->NT > MAXNT || cfg->N < 3 || cfg->Lm < 2 || cfg->Mm < 2) { g_err = "roms_create: unsupported sizes"; return -1; }
  CK(cudaSetDevice(device));
  roms_ctx* h = new roms_ctx();
  h->c = *cfg;
  h->device = device;
  CK(cudaStreamCreateWithFlags(&h->st, cudaStreamNonBlocking));
  {
    int lo = 0, hi = 0;
    CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    CK(cudaStreamCreateWithPriority(&h->st2, cudaStreamNonBlocking, hi));
    for (int q = 0; q < 2; ++q) { CK(cudaEventCreateWithFlags(&h->ev_int[q], cudaEventDisableTiming)); CK(cudaEventCreateWithFlags(&h->ev_edge[q], cudaEventDisableTiming)); }
    CK(cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming));
  }
  const int N = cfg->N, NT = cfg->NT;
  KP& p = h->kp;
  memset(&p, 0, sizeof(p));
  p.Lm = cfg->Lm; p.Mm = cfg->Mm; p.N = N; p.NT = NT;
  p.SX = ((cfg->Lm + 18 + 15) / 16) * 16;
  if ((double)p.SX * (cfg->Mm + 4) * (N + 2) >= 2147483647.0) { g_err = "roms_create: subdomain too large for one GPU (a 3-D field must stay below 2^31 elements); split it over more GPUs"; return -1; }
  p.SXY = (ix_t)(p.SX * (cfg->Mm + 4));
  p.istr = 1; p.iend = cfg->Lm; p.jstr = 1; p.jend = cfg->Mm;
  {
    const int npe = cfg->np_eta > 0 ? cfg->np_eta : 1;
    p.swall = (!cfg->ns_periodic && cfg->jnode == 0) ? 1 : 0;          // SOUTHERN_EDGE of set_global_definitions.h
    p.nwall = (!cfg->ns_periodic && cfg->jnode == npe - 1) ? 1 : 0;    // NORTHERN_EDGE
    p.jstrR = p.swall ? p.jstr - 1 : p.jstr; p.jendR = p.nwall ? p.jend + 1 : p.jend; p.jstrV = p.swall ? p.jstr + 1 : p.jstr;
    const int npx = cfg->np_xi > 0 ? cfg->np_xi : 1;
    p.wwall = (!cfg->ew_periodic && cfg->inode == 0) ? 1 : 0;          // WESTERN_EDGE
    p.ewall = (!cfg->ew_periodic && cfg->inode == npx - 1) ? 1 : 0;    // EASTERN_EDGE
    p.istrR = p.wwall ? p.istr - 1 : p.istr; p.iendR = p.ewall ? p.iend + 1 : p.iend; p.istrU = p.wwall ? p.istr + 1 : p.istr;
  }
  p.nonlin = cfg->nonlin_eos; p.salinity = cfg->salinity;
  p.dt = cfg->dt; p.dtfast = cfg->dt / double(cfg->ndtfast);
  p.g = cfg->g; p.rho0 = cfg->rho0; p.Tcoef = cfg->Tcoef; p.T0 = cfg->T0; p.Scoef = cfg->Scoef; p.S0 = cfg->S0;
  p.rdrg = cfg->rdrg; p.rdrg2 = cfg->rdrg2; p.Zob = cfg->Zob; p.gamma2 = cfg->gamma2; p.hc = cfg->hc;
  if (const char* e = getenv("ROMS_COL_BX")) h->col_bx = atoi(e);
  if (const char* e = getenv("ROMS_COL_BY")) h->col_by = atoi(e);
  if (const char* e = getenv("ROMS_PL_BX")) h->pl_bx = atoi(e);
  if (const char* e = getenv("ROMS_PL_BY")) h->pl_by = atoi(e);
  const bool single = cfg->np_xi * cfg->np_eta == 1;
  p.wrap_x = (single && cfg->ew_periodic) ? 1 : 0;
  p.wrap_y = (single && cfg->ns_periodic) ? 1 : 0;
  if (const char* e = getenv("ROMS_COL_SMEM")) h->use_smem = atoi(e);
  if (const char* e = getenv("ROMS_UV2_SM")) h->uv2_sm = atoi(e);
  if (const char* e = getenv("ROMS_TB")) h->tb_max = atoi(e);
  if (const char* e = getenv("ROMS_UV2_FUSED")) h->uv2_fused = atoi(e);
  if (const char* e = getenv("ROMS_OVERLAP")) h->overlap = atoi(e);
  if (const char* e = getenv("ROMS_T_TMA")) h->t_tma = atoi(e);
  if (const char* e = getenv("ROMS_S2D_TMA")) h->s2d_tma = atoi(e);
  if (const char* e = getenv("ROMS_S2D_ROWS")) h->s2d_rows = atoi(e) > 0 ? atoi(e) : 0;
  CK(cudaDeviceGetAttribute(&h->n_sm, cudaDevAttrMultiProcessorCount, device));
  if (const char* e = getenv("ROMS_FUSE_RHO")) h->fuse_rho = atoi(e);
  if ((size_t)2 * (N + 1) * 32 * sizeof(double) > 64 * 1024) h->uv2_sm = 0;   // fewer than 3 warps per SM would fit: global re-read version
  if ((size_t)2 * (N + 1) * h->col_bx * h->col_by * sizeof(double) > 200 * 1024) { h->col_by = 1; }
  if ((size_t)2 * (N + 1) * h->col_bx * h->col_by * sizeof(double) > 200 * 1024) h->use_smem = 0;
  p.use_smem = h->use_smem;

  // field table
  for (int id = 0; id < ROMS_NFIELDS; ++id) slot(h, id, 1, 1, 1);
  for (int id = ROMS_ZETA_WEST; id <= ROMS_T_NORTH; ++id) slot(h, id, 0, 0, 0);     // boundary lines live in bry_buf, not in the plane arena
  slot(h, ROMS_ZETA, 4, 1, 1); slot(h, ROMS_UBAR, 4, 1, 1); slot(h, ROMS_VBAR, 4, 1, 1);
  slot(h, ROMS_STFLX, 1, NT, 1);
  slot(h, ROMS_DIFF2, 1, NT, 1);
  slot(h, ROMS_U, 3, 1, N); slot(h, ROMS_V, 3, 1, N); slot(h, ROMS_T, 3, NT, N);
  for (int id : {ROMS_HZ, ROMS_Z_R, ROMS_FLXU, ROMS_FLXV, ROMS_RHO1, ROMS_QP1, ROMS_RU, ROMS_RV}) slot(h, id, 1, 1, N);
  for (int id : {ROMS_Z_W, ROMS_WE, ROMS_WI, ROMS_AKV}) slot(h, id, 1, 1, N + 1);
  slot(h, ROMS_AKT, 1, 2, N + 1);
  // one arena: all fields + device-owned scratch
  size_t planes = 0;
  for (int id = 0; id < ROMS_NFIELDS; ++id) planes += (size_t)h->f[id].nlvl * h->f[id].ntrc * h->f[id].nk;
  if (h->use_smem) h->tb_max = 1;
  if (h->tb_max != 1 && h->tb_max != 2 && h->tb_max != 4) { g_err = "ROMS_TB must be 1, 2 or 4"; return -1; }
  // column scratch slab: spline CF + one array per tracer of the largest batch (k_t_vert), 2 arrays for k_uv_vert / k_omega
  const int scr_arrays = (NT - 2 >= 4 && h->tb_max >= 4) ? 5 : ((NT >= 2 && h->tb_max >= 2) ? 3 : 2);
  const size_t scratch_planes = (size_t)3 * N            /* FlxDiv, P, rho */
                                + (h->use_smem ? 0 : (size_t)scr_arrays * (N + 1)) /* column scratch slab */
                                + 1 /* zeta_new */ + 2 /* ub, vb */ + 5 /* diag planes */;
  const size_t small = 2 * (size_t)(N + 1) + 64 + 7 * (size_t)cfg->Mm + 8 + (size_t)MAXRIV * RIVROW + (size_t)MAXPIP * PIPROW(N);
  h->arena_doubles = (planes + scratch_planes) * (size_t)p.SXY + small;
  CK(cudaMalloc(&h->arena, h->arena_doubles * sizeof(double)));
  CK(cudaMemset(h->arena, 0, h->arena_doubles * sizeof(double)));   // init=0 (init_scalars.F:31)
  double* cur = h->arena;
  auto take = [&](size_t npl) { double* r = cur; cur += npl * (size_t)p.SXY; return r; };
  for (int id = 0; id < ROMS_NFIELDS; ++id)
    for (size_t s = 0; s < h->f[id].ptr.size(); ++s) h->f[id].ptr[s] = take(h->f[id].nk);
  p.FlxDiv = take(N); p.P = take(N); p.rho = take(N);
  p.scr = h->use_smem ? nullptr : take((size_t)scr_arrays * (N + 1));
  p.zeta_new = take(1);
  h->d_ub = take(1); h->d_vb = take(1);
  for (int q = 0; q < 5; ++q) h->d_o[q] = take(1);
  p.Cs_w = cur; cur += (N + 1);
  p.Cs_r = cur; cur += (N + 1);
  p.riv_tab = cur; cur += (size_t)MAXRIV * RIVROW;
  p.rivers = 0;
  p.pipe_tab = cur; cur += (size_t)MAXPIP * PIPROW(N);
  h->d_rows = cur; cur += 7 * (size_t)cfg->Mm + 8;
  p.pipes = 0;
  p.pot_tides = 0;
  p.lmd_kpp = 0;
#define P1(name, id) p.name = h->f[id].ptr[0]
  P1(h, ROMS_H); P1(pm, ROMS_PM); P1(pn, ROMS_PN); P1(f, ROMS_F); P1(rmask, ROMS_RMASK); P1(umask, ROMS_UMASK); P1(vmask, ROMS_VMASK);
  P1(dm_u, ROMS_DM_U); P1(dn_u, ROMS_DN_U); P1(dm_v, ROMS_DM_V); P1(dn_v, ROMS_DN_V); P1(dm_r, ROMS_DM_R); P1(dn_r, ROMS_DN_R);
  P1(fomn, ROMS_FOMN); P1(hinv, ROMS_HINV); P1(rhoA, ROMS_RHOA); P1(rhoS, ROMS_RHOS); P1(rufrc, ROMS_RUFRC); P1(rvfrc, ROMS_RVFRC);
  P1(Zt_avg1, ROMS_ZT_AVG1); P1(DU_avg1, ROMS_DU_AVG1); P1(DV_avg1, ROMS_DV_AVG1); P1(DU_avg2, ROMS_DU_AVG2); P1(DV_avg2, ROMS_DV_AVG2);
  P1(DU_avg_bak, ROMS_DU_AVG_BAK); P1(DV_avg_bak, ROMS_DV_AVG_BAK); P1(r_D, ROMS_R_D); P1(sustr, ROMS_SUSTR); P1(svstr, ROMS_SVSTR);
  P1(swflx, ROMS_SWFLX); P1(Hz, ROMS_HZ); P1(z_r, ROMS_Z_R); P1(z_w, ROMS_Z_W); P1(FlxU, ROMS_FLXU); P1(FlxV, ROMS_FLXV); P1(We, ROMS_WE);
  P1(Wi, ROMS_WI); P1(rho1, ROMS_RHO1); P1(qp1, ROMS_QP1); P1(Akv, ROMS_AKV); P1(ru, ROMS_RU); P1(rv, ROMS_RV);
  P1(pmask, ROMS_PMASK); P1(dm_p, ROMS_DM_P); P1(dn_p, ROMS_DN_P); P1(pmon_u, ROMS_PMON_U); P1(pnom_v, ROMS_PNOM_V);
  P1(visc2_r, ROMS_VISC2_R); P1(visc2_p, ROMS_VISC2_P);
  P1(riv_uflx, ROMS_RIV_UFLX); P1(riv_vflx, ROMS_RIV_VFLX); P1(pipe_idx, ROMS_PIPE_IDX); P1(pipe_flx, ROMS_PIPE_FLX); P1(ptide, ROMS_PTIDE); P1(srflx, ROMS_SRFLX);
#undef P1
  p.Akt[0] = h->f[ROMS_AKT].ptr[0]; p.Akt[1] = h->f[ROMS_AKT].ptr[1];
  if (h->t_tma) {
    // the arena as ONE 2-D tensor: SX columns x every row of every plane (all fields share the pitch and the plane height)
    typedef CUresult (*encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fp = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &qr) == cudaSuccess && fp && qr == cudaDriverEntryPointSuccess) {
      encode_fn enc = (encode_fn)fp;
      const cuuint64_t gdim[2] = {(cuuint64_t)p.SX, (cuuint64_t)(h->arena_doubles / (size_t)p.SX)};
      const cuuint64_t gstr[1] = {(cuuint64_t)p.SX * sizeof(double)};
      const cuuint32_t estr[2] = {1, 1};
      struct { CUtensorMap* m; cuuint32_t bx, by; } mk[3] = {{&h->tmaps.t36x12, TT_W, TT_H}, {&h->tmaps.f34x9, TF_W, TF_H}, {&h->tmaps.o32x8, 32, 8}};
      int ok = 1;
      for (auto& q : mk) {
        const cuuint32_t box[2] = {q.bx, q.by};
        if (enc(q.m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, h->arena, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) ok = 0;
      }
      h->tmaps_ok = ok;
    } else cudaGetLastError();
    if (!h->tmaps_ok) { g_err = "roms_create: cuTensorMapEncodeTiled failed (the TMA tracer kernels need it); set ROMS_T_TMA=0 to use the scalar-load kernels"; return -1; }
  }
  host_scoord(h);
  host_weights(h);
  CK(cudaMemcpy(p.Cs_w, h->Cs_w.data(), (N + 1) * sizeof(double), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(p.Cs_r, h->Cs_r.data(), (N + 1) * sizeof(double), cudaMemcpyHostToDevice));
  // open boundaries (obc.cuh)
  p.obc = (cfg->obc_west ? 1 : 0) | (cfg->obc_east ? 2 : 0) | (cfg->obc_south ? 4 : 0) | (cfg->obc_north ? 8 : 0);
  if (cfg->ew_periodic) p.obc &= ~3;
  if (cfg->ns_periodic) p.obc &= ~12;
  p.ubind = cfg->ubind;
  p.obc_m2 = cfg->obc_m2; p.obc_m3 = cfg->obc_m3; p.obc_t = cfg->obc_t;
  int bry = p.obc != 0;
  if (const char* e = getenv("ROMS_BRY")) bry = bry || (atoi(e) != 0 && (!cfg->ew_periodic || !cfg->ns_periodic));
  p.bry2d = p.bry3d = bry;
  for (int sd = 0; sd < 4; ++sd) { p.bry_zeta[sd] = p.bry_ubar[sd] = p.bry_vbar[sd] = p.bry_u[sd] = p.bry_v[sd] = p.bry_t[sd] = nullptr; }
  if (bry) {
    size_t n = 0;
    for (int sd = 0; sd < 4; ++sd) n += (size_t)BRYL(p, sd) * (3 + (size_t)(2 + NT) * N);
    h->bry_doubles = n;
    CK(cudaMalloc(&h->bry_buf, n * sizeof(double)));
    CK(cudaMemset(h->bry_buf, 0, n * sizeof(double)));
    double* cb = h->bry_buf;
    for (int sd = 0; sd < 4; ++sd) {
      const size_t L = BRYL(p, sd);
      p.bry_zeta[sd] = cb; cb += L; p.bry_ubar[sd] = cb; cb += L; p.bry_vbar[sd] = cb; cb += L;
      p.bry_u[sd] = cb; cb += L * N; p.bry_v[sd] = cb; cb += L * N; p.bry_t[sd] = cb; cb += L * N * NT;
    }
  }
  comm_init(&h->comm, cfg);
  *out = h;
  return 0;
}

int roms_destroy(roms_ctx* h) {
  if (!h) return 0;
  cudaSetDevice(h->device);
  cudaStreamSynchronize(h->st);
  comm_destroy(&h->comm);
  cudaFree(h->arena);
  cudaFree(h->bry_buf);
  cudaFree(h->kpp_buf);
  for (int q = 0; q < 2; ++q) { if (h->ev_int[q]) cudaEventDestroy(h->ev_int[q]); if (h->ev_edge[q]) cudaEventDestroy(h->ev_edge[q]); }
  if (h->ev_join) cudaEventDestroy(h->ev_join);
  if (h->st2) cudaStreamDestroy(h->st2);
  cudaStreamDestroy(h->st);
  delete h;
  return 0;
}

// work enqueued on the second stream (ring kernels and halo swaps of an unfinished barotropic loop) becomes visible to the main stream
static int join_st2(roms_ctx* h) {
  if (!h->st2) return 0;
  CK(cudaEventRecord(h->ev_join, h->st2));
  CK(cudaStreamWaitEvent(h->st, h->ev_join, 0));
  return 0;
}
static int find_slot(roms_ctx* h, int field, int lvl, int itrc, double** ptr, int* nk) {
  if (field < 0 || field >= ROMS_NFIELDS) { g_err = "bad field id"; return -1; }
  FieldSlot& s = h->f[field];
  const int l = s.nlvl > 1 ? lvl : 1, t = s.ntrc > 1 ? itrc : 1;
  if (l < 1 || l > s.nlvl || t < 1 || t > s.ntrc) { g_err = "bad level/tracer selector"; return -1; }
  *ptr = s.ptr[(size_t)(l - 1) * s.ntrc + (t - 1)];
  *nk = s.nk;
  return 0;
}
// boundary data of the open edges: device line and its length for a ROMS_*_WEST ... ROMS_*_NORTH id (nullptr when no edge is open: the
// data would never be read, pushes are accepted and dropped)
static bool bry_field(roms_ctx* h, int field, int itrc, double** d, size_t* n) {
  if (field < ROMS_ZETA_WEST || field > ROMS_T_NORTH) return false;
  const KP& p = h->kp;
  const int var = (field - ROMS_ZETA_WEST) / 4, sd = (field - ROMS_ZETA_WEST) % 4;
  const size_t L = BRYL(p, sd);
  *n = L * (var < 3 ? 1 : (size_t)p.N);
  double* base = var == 0 ? p.bry_zeta[sd] : var == 1 ? p.bry_ubar[sd] : var == 2 ? p.bry_vbar[sd] : var == 3 ? p.bry_u[sd] : var == 4 ? p.bry_v[sd] : p.bry_t[sd];
  if (var == 5 && base) {
    const int t = (itrc >= 1 && itrc <= h->c.NT) ? itrc : 1;
    base += (size_t)(t - 1) * L * p.N;
  }
  *d = base;
  return true;
}
static int bry_xfer(roms_ctx* h, int field, int itrc, double* host, bool push, bool wait) {
  double* d; size_t n;
  bry_field(h, field, itrc, &d, &n);
  if (!d) { if (!push) memset(host, 0, n * sizeof(double)); return 0; }
  CK(cudaSetDevice(h->device));
  if (push) CK(cudaMemcpyAsync(d, host, n * sizeof(double), cudaMemcpyHostToDevice, h->st));
  else CK(cudaMemcpyAsync(host, d, n * sizeof(double), cudaMemcpyDeviceToHost, h->st));
  if (wait) CK(cudaStreamSynchronize(h->st));
  return 0;
}
#define IS_BRY(field) ((field) >= ROMS_ZETA_WEST && (field) <= ROMS_T_NORTH)
long roms_field_count(roms_ctx* h, int field) {
  if (field < 0 || field >= ROMS_NFIELDS) return 0;
  if (IS_BRY(field)) { double* d; size_t n; bry_field(h, field, 1, &d, &n); return (long)n; }
  return (long)(h->c.Lm + 4) * (h->c.Mm + 4) * h->f[field].nk;
}
int roms_push(roms_ctx* h, int field, int lvl, int itrc, const double* host) {
  FLUSH_RHO(h);
  if (join_st2(h)) return -1;
  if (IS_BRY(field)) return bry_xfer(h, field, itrc, const_cast<double*>(host), true, true);
  double* d; int nk;
  if (find_slot(h, field, lvl, itrc, &d, &nk)) return -1;
  CK(cudaSetDevice(h->device));
  const size_t w = (size_t)(h->c.Lm + 4) * sizeof(double);
  CK(cudaMemcpy2DAsync(d + (IOFF - 1), (size_t)h->kp.SX * sizeof(double), host, w, w, (size_t)(h->c.Mm + 4) * nk, cudaMemcpyHostToDevice, h->st));
  CK(cudaStreamSynchronize(h->st));
  return 0;
}
int roms_pull(roms_ctx* h, int field, int lvl, int itrc, double* host) {
  FLUSH_RHO(h);
  if (join_st2(h)) return -1;
  if (IS_BRY(field)) return bry_xfer(h, field, itrc, host, false, true);
  double* d; int nk;
  if (find_slot(h, field, lvl, itrc, &d, &nk)) return -1;
  CK(cudaSetDevice(h->device));
  const size_t w = (size_t)(h->c.Lm + 4) * sizeof(double);
  CK(cudaMemcpy2DAsync(host, w, d + (IOFF - 1), (size_t)h->kp.SX * sizeof(double), w, (size_t)(h->c.Mm + 4) * nk, cudaMemcpyDeviceToHost, h->st));
  CK(cudaStreamSynchronize(h->st));
  return 0;
}
int roms_push_async(roms_ctx* h, int field, int lvl, int itrc, const double* host) {
  FLUSH_RHO(h);
  if (join_st2(h)) return -1;
  if (IS_BRY(field)) return bry_xfer(h, field, itrc, const_cast<double*>(host), true, false);
  double* d; int nk;
  if (find_slot(h, field, lvl, itrc, &d, &nk)) return -1;
  CK(cudaSetDevice(h->device));
  const size_t w = (size_t)(h->c.Lm + 4) * sizeof(double);
  CK(cudaMemcpy2DAsync(d + (IOFF - 1), (size_t)h->kp.SX * sizeof(double), host, w, w, (size_t)(h->c.Mm + 4) * nk, cudaMemcpyHostToDevice, h->st));
  return 0;
}
int roms_pull_async(roms_ctx* h, int field, int lvl, int itrc, double* host) {
  FLUSH_RHO(h);
  if (join_st2(h)) return -1;
  if (IS_BRY(field)) return bry_xfer(h, field, itrc, host, false, false);
  double* d; int nk;
  if (find_slot(h, field, lvl, itrc, &d, &nk)) return -1;
  CK(cudaSetDevice(h->device));
  const size_t w = (size_t)(h->c.Lm + 4) * sizeof(double);
  CK(cudaMemcpy2DAsync(host, w, d + (IOFF - 1), (size_t)h->kp.SX * sizeof(double), w, (size_t)(h->c.Mm + 4) * nk, cudaMemcpyDeviceToHost, h->st));
  return 0;
}
int roms_sync(roms_ctx* h) {
  FLUSH_RHO(h);
  CK(cudaSetDevice(h->device));
  if (join_st2(h)) return -1;
  CK(cudaStreamSynchronize(h->st));
  CK(cudaGetLastError());
  return comm_check_failed(&h->comm);
}
int roms_kernel_launches(roms_ctx* h, long* n) { *n = h->launches; return 0; }
// bytes of HBM this context keeps resident (state arena + KPP arrays + boundary data of open edges; halo buffers are not counted)
int roms_resident_bytes(roms_ctx* h, long long* n) {
  const size_t kpp = h->kpp_buf ? ((size_t)6 * h->kp.SXY * (h->kp.N + 1) + (size_t)13 * h->kp.SXY) * sizeof(double) : 0;
  *n = (long long)(h->arena_doubles * sizeof(double) + kpp + h->bry_doubles * sizeof(double));
  return 0;
}
// device-side timing on the library's own stream (bench.py): events bracket whatever is enqueued between
int roms_timer_start(roms_ctx* h) {
  FLUSH_RHO(h);
  CK(cudaSetDevice(h->device));
  if (!h->ev0) { CK(cudaEventCreate(&h->ev0)); CK(cudaEventCreate(&h->ev1)); }
  CK(cudaStreamSynchronize(h->st));
  CK(cudaEventRecord(h->ev0, h->st));
  return 0;
}
int roms_timer_stop(roms_ctx* h, double* ms) {
  FLUSH_RHO(h);
  if (join_st2(h)) return -1;
  CK(cudaEventRecord(h->ev1, h->st));
  CK(cudaEventSynchronize(h->ev1));
  float f = 0;
  CK(cudaEventElapsedTime(&f, h->ev0, h->ev1));
  *ms = f;
  return 0;
}
// per-kernel timing: while on, every launch is bracketed by its own pair of events
int roms_kernel_timing(roms_ctx* h, int on) {
  CK(cudaStreamSynchronize(h->st));
  for (auto& t : h->ktimers) { cudaEventDestroy(t.a); cudaEventDestroy(t.b); }
  h->ktimers.clear();
  h->ktiming = on;
  return 0;
}
// writes "name count total_ms\n" lines for every kernel timed since roms_kernel_timing(h,1)
int roms_kernel_times(roms_ctx* h, char* buf, int buflen) {
  CK(cudaStreamSynchronize(h->st));
  std::vector<std::string> names;
  std::vector<double> tot;
  std::vector<long> cnt;
  for (auto& t : h->ktimers) {
    float f = 0;
    cudaEventElapsedTime(&f, t.a, t.b);
    size_t q = 0;
    for (; q < names.size(); ++q) if (names[q] == t.name) break;
    if (q == names.size()) { names.push_back(t.name); tot.push_back(0); cnt.push_back(0); }
    tot[q] += f; cnt[q]++;
  }
  std::string out;
  char line[256];
  for (size_t q = 0; q < names.size(); ++q) {
    snprintf(line, sizeof line, "%s %ld %.6f\n", names[q].c_str(), cnt[q], tot[q]);
    out += line;
  }
  if ((int)out.size() + 1 > buflen) { g_err = "roms_kernel_times: buffer too small"; return -1; }
  memcpy(buf, out.c_str(), out.size() + 1);
  return 0;
}
int roms_get_nfast(roms_ctx* h) { return h->nfast; }
int roms_get_weights(roms_ctx* h, double* w1, double* w2, int n) {
  for (int i = 1; i <= n; ++i) { w1[i - 1] = h->w1[i]; w2[i - 1] = h->w2[i]; }
  return 0;
}
int roms_get_scoord(roms_ctx* h, double* Cs_w, double* Cs_r) {
  for (int k = 0; k <= h->c.N; ++k) Cs_w[k] = h->Cs_w[k];
  for (int k = 1; k <= h->c.N; ++k) Cs_r[k - 1] = h->Cs_r[k];
  return 0;
}
int roms_set_time_indices(roms_ctx* h, int iic, int iif, int nstp, int nrhs, int nnew, int kstp, int knew) {
  h->iic = iic; h->iif = iif; h->nstp = nstp; h->nrhs = nrhs; h->nnew = nnew; h->kstp = kstp; h->knew = knew;
  return 0;
}
int roms_get_time_indices(roms_ctx* h, int* iic, int* iif, int* nstp, int* nrhs, int* nnew, int* kstp, int* knew) {
  *iic = h->iic; *iif = h->iif; *nstp = h->nstp; *nrhs = h->nrhs; *nnew = h->nnew; *kstp = h->kstp; *knew = h->knew;
  return 0;
}
int roms_compute_tile_bounds(roms_ctx* h, int tile, int* istr, int* iend, int* jstr, int* jend) {
  if (tile != 0) { g_err = "only tile 0 exists (NSUB_X=NSUB_E=1, one tile per GPU subdomain)"; return -1; }
  *istr = 1; *iend = h->c.Lm; *jstr = 1; *jend = h->c.Mm;
  return 0;
}

// ---- exchange_xxx -------------------------------------------------------------------------------
static int exchange(roms_ctx* h, std::initializer_list<XF> fl, cudaStream_t st = nullptr) {
  if (h->ktiming) kt_begin(h, "exchange_xxx", st);
  int rc = comm_exchange(&h->comm, h->kp, st ? st : h->st, fl.begin(), (int)fl.size(), &h->launches);
  if (h->ktiming) kt_end(h, st);
  return rc;
}
int roms_exchange(roms_ctx* h, int nfields, const int* fields, const int* lvls, const int* itrcs) {
  FLUSH_RHO(h);
  CK(cudaSetDevice(h->device));
  std::vector<XF> v;
  for (int n = 0; n < nfields; ++n) {
    double* d; int nk;
    if (find_slot(h, fields[n], lvls ? lvls[n] : 1, itrcs ? itrcs[n] : 1, &d, &nk)) return -1;
    v.push_back({d, nk});
  }
  return comm_exchange(&h->comm, h->kp, h->st, v.data(), (int)v.size(), &h->launches);
}

// boundary rows of the 3-D fields (u3dbc/v3dbc/t3dbc/omega copies): no-ops unless this subdomain touches a non-periodic edge.
// itrc: tracer (1-based) of a BC_T pass -- the open-edge rule needs t(nstp,itrc) and the tracer's boundary data.
static int bc3d(roms_ctx* h, double* a, int nk, int kind, int itrc = 1) {
  const KP& p = h->kp;
  if (!p.swall && !p.nwall && !p.wwall && !p.ewall) return 0;
  const int len = (p.Lm > p.Mm ? p.Lm : p.Mm) + 3;
  dim3 b(128, 1, 1), g((len + 127) / 128, 4, nk);
  if (p.bry3d && kind != BC_W) {      // at least one open edge: the general boundary kernels (obc.cuh) take every non-periodic edge
    OB3 o;
    o.a = a; o.ur = h->u(h->nrhs); o.vr = h->v(h->nrhs);
    o.as = kind == BC_U ? h->u(h->nstp) : (kind == BC_V ? h->v(h->nstp) : h->t(h->nstp, itrc));
    o.dtfwd = (h->nnew == 3) ? 0.5 * h->c.dt : h->c.dt;
    o.itrc = itrc - 1;
    if (h->ktiming) kt_begin(h, "k_bry3d");
    if (kind == BC_U) k_bry3d<BC_U><<<g, b, 0, h->st>>>(p, o);
    else if (kind == BC_V) k_bry3d<BC_V><<<g, b, 0, h->st>>>(p, o);
    else k_bry3d<BC_T><<<g, b, 0, h->st>>>(p, o);
    if (h->ktiming) kt_end(h);
    h->launches++;
    return 0;
  }
  if (h->ktiming) kt_begin(h, "k_bc3d");
  const double g2 = h->c.gamma2;
  if (kind == BC_U) k_bc3d<BC_U><<<g, b, 0, h->st>>>(p, a, g2);
  else if (kind == BC_V) k_bc3d<BC_V><<<g, b, 0, h->st>>>(p, a, g2);
  else if (kind == BC_T) k_bc3d<BC_T><<<g, b, 0, h->st>>>(p, a, g2);
  else k_bc3d<BC_W><<<g, b, 0, h->st>>>(p, a, g2);
  if (h->ktiming) kt_end(h);
  h->launches++;
  return 0;
}
#define WALLED(h) ((h)->kp.swall || (h)->kp.nwall || (h)->kp.wwall || (h)->kp.ewall)

// ---- setup_grid1.F:13-210 (metrics, masks, fomn): host arithmetic on pulled planes, pushed back --
int roms_setup_grid(roms_ctx* h) {
  const int Lm = h->c.Lm, Mm = h->c.Mm, SXh = Lm + 4;
  const size_t n2 = (size_t)SXh * (Mm + 4);
  std::vector<double> pm(n2), pn(n2), f(n2), rm(n2);
  if (roms_pull(h, ROMS_PM, 1, 1, pm.data()) || roms_pull(h, ROMS_PN, 1, 1, pn.data()) || roms_pull(h, ROMS_F, 1, 1, f.data()) ||
      roms_pull(h, ROMS_RMASK, 1, 1, rm.data()))
    return -1;
  auto at = [&](int i, int j) { return (size_t)(i + 1) + (size_t)(j + 1) * SXh; };
  std::vector<double> fomn(n2, 0.0), dm_r(n2, 0.0), dn_r(n2, 0.0), dm_u(n2, 0.0), dn_u(n2, 0.0), dm_v(n2, 0.0), dn_v(n2, 0.0), um(n2, 0.0), vm(n2, 0.0);
  const KP& pp = h->kp;
  for (int j = (pp.swall ? 0 : -1); j <= (pp.nwall ? Mm + 1 : Mm + 2); ++j)     // extended bounds stop one row outside a wall
    for (int i = (pp.wwall ? 0 : -1); i <= (pp.ewall ? Lm + 1 : Lm + 2); ++i) fomn[at(i, j)] = f[at(i, j)] / (pm[at(i, j)] * pn[at(i, j)]);
  for (int j = pp.jstrR; j <= pp.jendR; ++j)
    for (int i = pp.istrR; i <= pp.iendR; ++i) {
      const size_t q = at(i, j);
      dm_r[q] = 1.0 / pm[q]; dn_r[q] = 1.0 / pn[q];
      if (i >= 1) {
        const size_t w = at(i - 1, j);
        dm_u[q] = 2.0 / (pm[q] + pm[w]); dn_u[q] = 2.0 / (pn[q] + pn[w]); um[q] = rm[q] * rm[w];
      }
      if (j >= 1) {
        const size_t s = at(i, j - 1);
        dm_v[q] = 2.0 / (pm[q] + pm[s]); dn_v[q] = 2.0 / (pn[q] + pn[s]); vm[q] = rm[q] * rm[s];
      }
    }
  // psi-point metrics and slipperiness mask, pm/pn ratios at u and v points (setup_grid1.F:107-191)
  std::vector<double> pmon_u(n2, 0.0), pnom_v(n2, 0.0), dm_p(n2, 0.0), dn_p(n2, 0.0), pmk(n2, 0.0);
  for (int j = pp.jstrR; j <= pp.jendR; ++j)
    for (int i = pp.istrR; i <= pp.iendR; ++i) {
      const size_t q = at(i, j);
      if (i >= 1) pmon_u[q] = (pm[q] + pm[at(i - 1, j)]) / (pn[q] + pn[at(i - 1, j)]);
      if (j >= 1) pnom_v[q] = (pn[q] + pn[at(i, j - 1)]) / (pm[q] + pm[at(i, j - 1)]);
      if (i >= 1 && j >= 1) {
        const size_t w = at(i - 1, j), s = at(i, j - 1), sw = at(i - 1, j - 1);
        dm_p[q] = 4. / (pm[q] + pm[s] + pm[w] + pm[sw]);
        dn_p[q] = 4. / (pn[q] + pn[s] + pn[w] + pn[sw]);
        const int land = (rm[w] < 0.5) + (rm[q] < 0.5) + (rm[sw] < 0.5) + (rm[s] < 0.5);
        // no-slip on land-mask boundaries: one land corner -> 1, two ADJACENT land corners -> 2, otherwise (diagonal pair, 3, 4) -> 0
        const bool diagonal = (rm[w] < 0.5 && rm[s] < 0.5 && !(rm[q] < 0.5) && !(rm[sw] < 0.5)) || (rm[q] < 0.5 && rm[sw] < 0.5 && !(rm[w] < 0.5) && !(rm[s] < 0.5));
        pmk[q] = land == 0 ? 1. : land == 1 ? 1. : (land == 2 && !diagonal) ? 2. : 0.;
      }
    }
  const size_t ncoef = n2;
  std::vector<double> coef(ncoef);
  struct { int id; std::vector<double>* a; } outv[] = {{ROMS_FOMN, &fomn}, {ROMS_DM_R, &dm_r}, {ROMS_DN_R, &dn_r}, {ROMS_DM_U, &dm_u}, {ROMS_DN_U, &dn_u},
                                                       {ROMS_DM_V, &dm_v}, {ROMS_DN_V, &dn_v}, {ROMS_UMASK, &um}, {ROMS_VMASK, &vm},
                                                       {ROMS_PMON_U, &pmon_u}, {ROMS_PNOM_V, &pnom_v}, {ROMS_DM_P, &dm_p}, {ROMS_DN_P, &dn_p}, {ROMS_PMASK, &pmk}};
  for (auto& o : outv)
    if (roms_push(h, o.id, 1, 1, o.a->data())) return -1;
  // background mixing coefficients: whole-array assignment (mixing.F:156-163)
  std::fill(coef.begin(), coef.end(), h->c.visc2);
  if (roms_push(h, ROMS_VISC2_R, 1, 1, coef.data()) || roms_push(h, ROMS_VISC2_P, 1, 1, coef.data())) return -1;
  for (int it = 1; it <= h->c.NT; ++it) {
    std::fill(coef.begin(), coef.end(), h->c.tnu2[it - 1]);
    if (roms_push(h, ROMS_DIFF2, 1, it, coef.data())) return -1;
  }
  const KP& p = h->kp;
  int rc = exchange(h, {{p.dm_r, 1}, {p.dn_r, 1}, {p.dm_u, 1}, {p.dn_u, 1}});
  if (!rc) rc = exchange(h, {{p.dm_v, 1}, {p.dn_v, 1}, {p.rmask, 1}, {p.umask, 1}});
  if (!rc) rc = exchange(h, {{p.vmask, 1}, {p.pmask, 1}, {p.dm_p, 1}, {p.dn_p, 1}});
  if (!rc) rc = exchange(h, {{p.pmon_u, 1}, {p.pnom_v, 1}});
  return rc;
}

// ---- the tile-subroutine surface ----------------------------------------------------------------
int roms_set_depth(roms_ctx* h, int tile) {
  FLUSH_RHO(h);
  (void)tile;
  CK(cudaSetDevice(h->device));
  const KP& p = h->kp;
  if (h->iic == 0) PLANE(k_set_depth_init, p.iendR, p.jstrR, p.jendR, 1, p, h->zeta(1), h->ubar(1), h->vbar(1));
  PLANE(k_set_depth, p.iendR, p.jstrR, p.jendR, p.N + 1, p, h->zeta(h->knew));
  if (h->iic == 0) { int rc = exchange(h, {{p.hinv, 1}}); if (rc) return rc; }
  return exchange(h, {{p.z_w, p.N + 1}, {p.z_r, p.N}, {p.Hz, p.N}});
}
int roms_set_huv(roms_ctx* h) {
  CK(cudaSetDevice(h->device));
  const KP& p = h->kp;
  PLANE(k_set_huv, p.iendR, p.jstrR, p.jendR, p.N, p, h->u(h->nrhs), h->v(h->nrhs));
  return exchange(h, {{p.FlxU, p.N}, {p.FlxV, p.N}});
}
int roms_set_huv1(roms_ctx* h, int tile) {
  FLUSH_RHO(h);
  (void)tile;
  CK(cudaSetDevice(h->device));
  const KP& p = h->kp;
  const int first = (h->iic == h->forw_start);
  COLUMN(k_set_huv1<0>, 0, p.iendR, p.jstrR, p.jendR, p, h->u(h->nnew), first);
  COLUMN(k_set_huv1<1>, 0, p.iendR, p.jstr, p.jendR, p, h->v(h->nnew), first);
  return exchange(h, {{p.FlxU, p.N}, {p.FlxV, p.N}, {h->u(h->nnew), p.N}, {h->v(h->nnew), p.N}});
}
int roms_omega(roms_ctx* h) {
  CK(cudaSetDevice(h->device));
  const KP& p = h->kp;
  double dtau;
  if (h->nrhs == 3) dtau = h->c.dt;
  else if (h->iic == h->forw_start) dtau = 0.5 * h->c.dt;
  else dtau = 0.6 * h->c.dt;
  COLUMN(k_omega, 2, p.iend, p.jstr, p.jend, p, dtau);
  bc3d(h, p.We, p.N + 1, BC_W);      // omega.F:171-231 edge copies
  bc3d(h, p.Wi, p.N + 1, BC_W);
  return exchange(h, {{p.We, p.N + 1}, {p.Wi, p.N + 1}});
}
}  // extern "C"
static int rho_eos_launch(roms_ctx* h, int tidx, bool with_prs_vert) {
  const KP& p = h->kp;
  const double* T = h->t(tidx, 1);
  const double* S = h->c.NT >= 2 ? h->t(tidx, 2) : nullptr;
  const int j0 = p.swall ? p.jstr - 1 : p.jstr - 2;
  if (h->c.nonlin_eos) {
    if (!S) { g_err = "NONLIN_EOS needs salinity (NT>=2)"; return -1; }
    if (with_prs_vert) COLUMN(k_rho_prs_vert<true>, 0, p.iend + 2, j0, p.jend + 2, p, T, S);
    else COLUMN(k_rho_eos<true>, 0, p.iend + 2, j0, p.jend + 2, p, T, S);
  } else {
    if (with_prs_vert) COLUMN(k_rho_prs_vert<false>, 0, p.iend + 2, j0, p.jend + 2, p, T, S);
    else COLUMN(k_rho_eos<false>, 0, p.iend + 2, j0, p.jend + 2, p, T, S);
  }
  return 0;
}
// a deferred rho_eos runs on its own as soon as anything but set_HUV / omega / prsgrd is called
static int flush_rho(roms_ctx* h) {
  if (!h->rho_pending) return 0;
  const int tidx = h->rho_pending;
  h->rho_pending = 0;
  return rho_eos_launch(h, tidx, false);
}
// ---- tracer batches: tracers handled together by one launch of the tracer kernels (see TrcBatch in kernels.cuh) ----
struct TPlan { int n; int it[4]; bool share; };
static std::vector<TPlan> plan_batches(const roms_ctx* h) {
  std::vector<TPlan> v;
  const int NT = h->c.NT, tmax = h->tb_max;
  int it = 1;
  if (NT >= 2 && tmax >= 2) { v.push_back({2, {1, 2, 0, 0}, false}); it = 3; }      // T and S: different diffusivities (Akt slots 1, 2)
  else { v.push_back({1, {1, 0, 0, 0}, true}); it = 2; }
  while (it <= NT) {                                                                // the rest share Akt(:,:,:,isalt)
    const int left = NT - it + 1;
    const int n = (left >= 4 && tmax >= 4) ? 4 : ((left >= 2 && tmax >= 2) ? 2 : 1);
    TPlan b = {n, {0, 0, 0, 0}, true};
    for (int q = 0; q < n; ++q) b.it[q] = it + q;
    v.push_back(b);
    it += n;
  }
  return v;
}
template <int TB>
static TrcBatch<TB> make_batch(roms_ctx* h, const TPlan& pl, int nrhs, int nnew, int nstp, int indx) {
  TrcBatch<TB> b;
  for (int q = 0; q < TB; ++q) {
    const int it = pl.it[q];
    b.rhs[q] = h->t(nrhs, it); b.tnew[q] = h->t(nnew, it); b.stp[q] = h->t(nstp, it); b.indx[q] = h->t(indx, it);
    b.akt[q] = h->kp.Akt[it >= 2 ? 1 : 0]; b.stflx[q] = h->f[ROMS_STFLX].ptr[it - 1]; b.itrc[q] = it;
  }
  return b;
}
// the three build variants of the horizontal tracer kernels: rivers (implies walls), walls, open periodic
#define T_HORIZ_LAUNCH(kern, TBn, ...)                                                                                          \
  do {                                                                                                                          \
    if (p.rivers) PLANE((kern<true, KB_T, true, TBn>), p.iend, p.jstr, p.jend, (p.N + KB_T - 1) / KB_T, p, __VA_ARGS__);       \
    else if (WALLED(h)) PLANE((kern<true, KB_T, false, TBn>), p.iend, p.jstr, p.jend, (p.N + KB_T - 1) / KB_T, p, __VA_ARGS__); \
    else PLANE((kern<false, KB_T, false, TBn>), p.iend, p.jstr, p.jend, (p.N + KB_T - 1) / KB_T, p, __VA_ARGS__);               \
  } while (0)
#define TTMA_KB 8
// ---- TMA-staged horizontal tracer kernels (tracer_tma.cuh)
template <int TB>
static TrcRows<TB> make_rows(roms_ctx* h, const TrcBatch<TB>& b) {
  const KP& p = h->kp;
  const int R = h->c.Mm + 4;
  auto row = [&](const double* a) { return (int)(((size_t)(a - h->arena) / (size_t)p.SXY) * (size_t)R); };
  TrcRows<TB> r;
  r.flxu = row(p.FlxU); r.flxv = row(p.FlxV); r.hz = row(p.Hz); r.fd = row(p.FlxDiv);
  for (int q = 0; q < TB; ++q) { r.rhs[q] = row(b.rhs[q]); r.tnew[q] = row(b.tnew[q]); r.stp[q] = row(b.stp[q]); r.indx[q] = row(b.indx[q]); }
  return r;
}
template <bool PRE, bool WALL, bool RIV, int TB, int NS>
static int t_tma_launch(roms_ctx* h, const TrcBatch<TB>& b, double dtau, double cf_stp, double cf_bak) {
  const KP& p = h->kp;
  typedef TtmaLayout<PRE, TB> L;
  const size_t sm_ = (size_t)NS * L::stage * sizeof(double) + NS * sizeof(unsigned long long);
  static bool attr_ = false;
  if (!attr_) { CK(cudaFuncSetAttribute((k_t_horiz_tma<PRE, WALL, RIV, TB, TTMA_KB, NS>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_)); attr_ = true; }
  dim3 b_(288, 1, 1), g_((p.iend + IOFF + 1 + 31) / 32, (p.jend - p.jstr + 1 + 7) / 8, (p.N + TTMA_KB - 1) / TTMA_KB);
  if (h->ktiming) kt_begin(h, PRE ? "k_pre_t_horiz_tma" : "k_t_horiz_tma");
  k_t_horiz_tma<PRE, WALL, RIV, TB, TTMA_KB, NS><<<g_, b_, sm_, h->st>>>(p, b, h->tmaps, make_rows<TB>(h, b), dtau, cf_stp, cf_bak);
  if (h->ktiming) kt_end(h);
  h->launches++;
  return 0;
}
template <bool PRE, int TB, int NS>
static int t_tma_variant(roms_ctx* h, const TrcBatch<TB>& b, double dtau, double cf_stp, double cf_bak) {
  if (h->kp.rivers) return t_tma_launch<PRE, true, true, TB, NS>(h, b, dtau, cf_stp, cf_bak);
  if (WALLED(h)) return t_tma_launch<PRE, true, false, TB, NS>(h, b, dtau, cf_stp, cf_bak);
  return t_tma_launch<PRE, false, false, TB, NS>(h, b, dtau, cf_stp, cf_bak);
}
// TMA-staged horizontal momentum RHS (k_rhs_uv_tma)
template <bool UP, bool WALL>
static int rhs_uv_tma_launch(roms_ctx* h, const double* u, const double* v) {
  const KP& p = h->kp;
  constexpr int NS = 3;
  const int R = h->c.Mm + 4;
  auto row = [&](const double* a) { return (int)(((size_t)(a - h->arena) / (size_t)p.SXY) * (size_t)R); };
  RhsRows r;
  r.u = row(u); r.v = row(v); r.flxu = row(p.FlxU); r.flxv = row(p.FlxV); r.hz = row(p.Hz); r.ru = row(p.ru); r.rv = row(p.rv);
  const size_t sm_ = (size_t)NS * RhsLayout::stage * sizeof(double) + NS * sizeof(unsigned long long);
  static bool attr_ = false;
  if (!attr_) { CK(cudaFuncSetAttribute((k_rhs_uv_tma<UP, WALL, TTMA_KB, NS>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_)); attr_ = true; }
  dim3 b_(288, 1, 1), g_((p.iend + IOFF + 1 + 31) / 32, (p.jend - p.jstr + 1 + 7) / 8, (p.N + TTMA_KB - 1) / TTMA_KB);
  if (h->ktiming) kt_begin(h, UP ? "k_rhs_uv_tma<UP3>" : "k_rhs_uv_tma<C4>");
  k_rhs_uv_tma<UP, WALL, TTMA_KB, NS><<<g_, b_, sm_, h->st>>>(p, h->tmaps, r, u, v);
  if (h->ktiming) kt_end(h);
  h->launches++;
  return 0;
}
// TMA-staged horizontal pressure gradient (k_prsgrd_tma, nonlinear split EOS)
template <bool WALL, int MINB>
static int prsgrd_tma_launch(roms_ctx* h) {
  const KP& p = h->kp;
  constexpr int NS = 3;
  const int R = h->c.Mm + 4;
  auto row = [&](const double* a) { return (int)(((size_t)(a - h->arena) / (size_t)p.SXY) * (size_t)R); };
  PrsRows r;
  r.zr = row(p.z_r); r.r1 = row(p.rho1); r.q1 = row(p.qp1); r.P = row(p.P); r.rho = row(p.rho); r.hz = row(p.Hz);
  const size_t sm_ = (size_t)NS * PrsLayout::stage * sizeof(double) + NS * sizeof(unsigned long long);
  static bool attr_ = false;
  if (!attr_) { CK(cudaFuncSetAttribute((k_prsgrd_tma<WALL, TTMA_KB, NS, MINB>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_)); attr_ = true; }
  dim3 b_(288, 1, 1), g_((p.iend + IOFF + 1 + 31) / 32, (p.jend - p.jstr + 1 + 7) / 8, (p.N + TTMA_KB - 1) / TTMA_KB);
  if (h->ktiming) kt_begin(h, "k_prsgrd_tma");
  k_prsgrd_tma<WALL, TTMA_KB, NS, MINB><<<g_, b_, sm_, h->st>>>(p, h->tmaps, r);
  if (h->ktiming) kt_end(h);
  h->launches++;
  return 0;
}
static int tracer_pre_horiz(roms_ctx* h, const TPlan& pl, int nrhs, int nnew, int nstp, int indx, double dtau, double cf_stp, double cf_bak) {
  const KP& p = h->kp;
  if (h->t_tma) {
    if (pl.n == 4) {      // two half-batches: a stage of four predictor tracers would leave one block per SM
      for (int half = 0; half < 2; ++half) {
        TPlan q = {2, {pl.it[2 * half], pl.it[2 * half + 1], 0, 0}, pl.share};
        int rc = t_tma_variant<true, 2, 3>(h, make_batch<2>(h, q, nrhs, nnew, nstp, indx), dtau, cf_stp, cf_bak);
        if (rc) return rc;
      }
      return 0;
    }
    if (pl.n == 2) return t_tma_variant<true, 2, 3>(h, make_batch<2>(h, pl, nrhs, nnew, nstp, indx), dtau, cf_stp, cf_bak);
    return t_tma_variant<true, 1, 3>(h, make_batch<1>(h, pl, nrhs, nnew, nstp, indx), dtau, cf_stp, cf_bak);
  }
  if (pl.n == 4) T_HORIZ_LAUNCH(k_pre_t_horiz, 4, make_batch<4>(h, pl, nrhs, nnew, nstp, indx), dtau, cf_stp, cf_bak);
  else if (pl.n == 2) T_HORIZ_LAUNCH(k_pre_t_horiz, 2, make_batch<2>(h, pl, nrhs, nnew, nstp, indx), dtau, cf_stp, cf_bak);
  else T_HORIZ_LAUNCH(k_pre_t_horiz, 1, make_batch<1>(h, pl, nrhs, nnew, nstp, indx), dtau, cf_stp, cf_bak);
  return 0;
}
static int tracer_cor_horiz(roms_ctx* h, const TPlan& pl, int nrhs, int nnew) {
  const KP& p = h->kp;
  if (h->t_tma) {
    if (pl.n == 4) {
      for (int half = 0; half < 2; ++half) {
        TPlan q = {2, {pl.it[2 * half], pl.it[2 * half + 1], 0, 0}, pl.share};
        int rc = t_tma_variant<false, 2, 3>(h, make_batch<2>(h, q, nrhs, nnew, nrhs, nnew), 0.0, 0.0, 0.0);
        if (rc) return rc;
      }
      return 0;
    }
    if (pl.n == 2) return t_tma_variant<false, 2, 3>(h, make_batch<2>(h, pl, nrhs, nnew, nrhs, nnew), 0.0, 0.0, 0.0);
    return t_tma_variant<false, 1, 3>(h, make_batch<1>(h, pl, nrhs, nnew, nrhs, nnew), 0.0, 0.0, 0.0);
  }
  if (pl.n == 4) {
    // measured on B200 (profiles/r2_kernels_tracers_v0.txt): the UP3 corrector is faster per tracer in batches of 2 (1.87 vs 2.54 ms)
    for (int half = 0; half < 2; ++half) {
      TPlan q = {2, {pl.it[2 * half], pl.it[2 * half + 1], 0, 0}, pl.share};
      T_HORIZ_LAUNCH(k_t_horiz, 2, make_batch<2>(h, q, nrhs, nnew, nrhs, nnew));
    }
  } else if (pl.n == 2) T_HORIZ_LAUNCH(k_t_horiz, 2, make_batch<2>(h, pl, nrhs, nnew, nrhs, nnew));
  else T_HORIZ_LAUNCH(k_t_horiz, 1, make_batch<1>(h, pl, nrhs, nnew, nrhs, nnew));
  return 0;
}
#define T_VERT_LAUNCH(MODEn)                                                                                                                            \
  do {                                                                                                                                                  \
    if (pl.n == 4) COLUMN((k_t_vert<MODEn, 4, true>), 5, p.iend, p.jstr, p.jend, p, make_batch<4>(h, pl, nrhs, nnew, nrhs, nnew), dtau);                 \
    else if (pl.n == 2 && pl.share) COLUMN((k_t_vert<MODEn, 2, true>), 3, p.iend, p.jstr, p.jend, p, make_batch<2>(h, pl, nrhs, nnew, nrhs, nnew), dtau); \
    else if (pl.n == 2) COLUMN((k_t_vert<MODEn, 2, false>), 3, p.iend, p.jstr, p.jend, p, make_batch<2>(h, pl, nrhs, nnew, nrhs, nnew), dtau);            \
    else COLUMN((k_t_vert<MODEn, 1, true>), 2, p.iend, p.jstr, p.jend, p, make_batch<1>(h, pl, nrhs, nnew, nrhs, nnew), dtau);                           \
  } while (0)
static int tracer_vert(roms_ctx* h, int mode, const TPlan& pl, int nrhs, int nnew, double dtau) {
  const KP& p = h->kp;
  if (mode == 0) T_VERT_LAUNCH(0);
  else T_VERT_LAUNCH(1);
  return 0;
}
extern "C" {
int roms_rho_eos(roms_ctx* h, int tidx) {
  CK(cudaSetDevice(h->device));
  // A deferred call for the SAME time level is still pending and nothing has touched the state since (every other export
  // flushes first): this call would recompute bit-identical values from identical inputs -- rho_eos(nnew) at the end of a
  // step (main.F:477) followed by rho_eos(nrhs) at the start of the next (main.F:397, nrhs = the previous nnew).  Keep one.
  if (h->fuse_rho && h->rho_pending == tidx) return 0;
  FLUSH_RHO(h);
  if (h->fuse_rho) { h->rho_pending = tidx; return 0; }
  return rho_eos_launch(h, tidx, false);
}
int roms_prsgrd(roms_ctx* h) {
  CK(cudaSetDevice(h->device));
  const KP& p = h->kp;
  bool fused = false;
  if (h->rho_pending) {      // rho_eos of this time level is still due: EOS + vertical pressure integration in one sweep
    const int tidx = h->rho_pending;
    h->rho_pending = 0;
    int rc = rho_eos_launch(h, tidx, true);
    if (rc) return rc;
    fused = true;
  }
  if (h->c.nonlin_eos) {
    if (!fused) COLUMN(k_prsgrd_vert<true>, 0, p.Lm, 0, p.Mm, p);
    if (h->t_tma) return WALLED(h) ? prsgrd_tma_launch<true, 3>(h) : prsgrd_tma_launch<false, 3>(h);   // bound 3 (72 registers): 1.42 ms; bound 2 (96, no spill): 1.57 ms
    if (WALLED(h)) PLANE((k_prsgrd_horiz<true, true, KB_PRS>), p.iend, p.jstr, p.jend, (p.N + KB_PRS - 1) / KB_PRS, p);
    else PLANE((k_prsgrd_horiz<true, false, KB_PRS>), p.iend, p.jstr, p.jend, (p.N + KB_PRS - 1) / KB_PRS, p);
  } else {
    if (!fused) COLUMN(k_prsgrd_vert<false>, 0, p.Lm, 0, p.Mm, p);
    if (WALLED(h)) PLANE((k_prsgrd_horiz<false, true, KB_PRS>), p.iend, p.jstr, p.jend, (p.N + KB_PRS - 1) / KB_PRS, p);
    else PLANE((k_prsgrd_horiz<false, false, KB_PRS>), p.iend, p.jstr, p.jend, (p.N + KB_PRS - 1) / KB_PRS, p);
  }
  return 0;
}
int roms_pre_step3d(roms_ctx* h, int tile) {
  FLUSH_RHO(h);
  (void)tile;
  CK(cudaSetDevice(h->device));
  const KP& p = h->kp;
  const int nstp = h->nstp, nrhs = h->nrhs, nnew = h->nnew, indx = 3 - nstp;
  const double dt = h->c.dt, AM3_crv = 1.0 / 6.0;
  double dtau, cf_stp, cf_bak;
  if (h->iic == h->forw_start) { dtau = 0.5 * dt; cf_stp = 1.0; cf_bak = 0.0; }
  else { dtau = dt * (1.0 - AM3_crv); cf_stp = 0.5 + AM3_crv; cf_bak = 0.5 - AM3_crv; }
  const bool wl = WALLED(h);
  PLANE(k_pre_flxdiv, p.iend, p.jstrV - 1, p.jend, p.N, p, dtau);
  const std::vector<TPlan> plan = plan_batches(h);
  for (const TPlan& pl : plan) { int rc = tracer_pre_horiz(h, pl, nrhs, nnew, nstp, indx, dtau, cf_stp, cf_bak); if (rc) return rc; }
  if (h->t_tma) { int rc = wl ? rhs_uv_tma_launch<false, true>(h, h->u(nrhs), h->v(nrhs)) : rhs_uv_tma_launch<false, false>(h, h->u(nrhs), h->v(nrhs)); if (rc) return rc; }
  else if (wl) PLANE_JB((k_rhs_uv_horiz<false, true, 4>), 4, p.iend, p.jstr, p.jend, p.N, p, h->u(nrhs), h->v(nrhs));
  else PLANE_JB((k_rhs_uv_horiz<false, false, 4>), 4, p.iend, p.jstr, p.jend, p.N, p, h->u(nrhs), h->v(nrhs));
  PLANE(k_rd_bott_drag, p.iend, p.jstrV - 1, p.jend, 1, p, h->u(nstp), h->v(nstp));
  for (const TPlan& pl : plan) { int rc = tracer_vert(h, 0, pl, nrhs, nnew, dtau); if (rc) return rc; }
  COLUMN((k_uv_vert<0, 0>), 2, p.iend, p.jstr, p.jend, p, h->u(nrhs), h->u(nnew), h->u(nstp), h->u(indx), dtau, cf_stp, cf_bak);
  COLUMN((k_uv_vert<0, 1>), 2, p.iend, p.jstrV, p.jend, p, h->v(nrhs), h->v(nnew), h->v(nstp), h->v(indx), dtau, cf_stp, cf_bak);
  if (p.rivers) PLANE(k_river_uv, p.iend, p.jstr, p.jend, 1, p, h->u(nnew), h->v(nnew), p.istrU, p.jstrV);   // pre_step3d4S.F:493-522
  // u3dbc/v3dbc/t3dbc (pre_step3d4S.F:533-545), then one exchange per tracer (:548)
  bc3d(h, h->u(nnew), p.N, BC_U);
  bc3d(h, h->v(nnew), p.N, BC_V);
  for (int it = 1; it <= h->c.NT; ++it) {
    bc3d(h, h->t(nnew, it), p.N, BC_T, it);
    int rc = exchange(h, {{h->t(nnew, it), p.N}});
    if (rc) return rc;
  }
  return 0;
}
int roms_step3d_uv1(roms_ctx* h, int tile) {
  FLUSH_RHO(h);
  (void)tile;
  CK(cudaSetDevice(h->device));
  const KP& p = h->kp;
  const int nrhs = h->nrhs, nnew = h->nnew;
  if (h->t_tma) { int rc = WALLED(h) ? rhs_uv_tma_launch<true, true>(h, h->u(nrhs), h->v(nrhs)) : rhs_uv_tma_launch<true, false>(h, h->u(nrhs), h->v(nrhs)); if (rc) return rc; }
  else if (WALLED(h)) PLANE_JB((k_rhs_uv_horiz<true, true, 1>), 1, p.iend, p.jstr, p.jend, p.N, p, h->u(nrhs), h->v(nrhs));
  else PLANE_JB((k_rhs_uv_horiz<true, false, 1>), 1, p.iend, p.jstr, p.jend, p.N, p, h->u(nrhs), h->v(nrhs));
  COLUMN((k_uv_vert<1, 0>), 2, p.iend, p.jstr, p.jend, p, h->u(nrhs), h->u(nnew), (const double*)nullptr, (double*)nullptr, h->c.dt, 0.0, 0.0);
  COLUMN((k_uv_vert<1, 1>), 2, p.iend, p.jstrV, p.jend, p, h->v(nrhs), h->v(nnew), (const double*)nullptr, (double*)nullptr, h->c.dt, 0.0, 0.0);
  return 0;
}
}  // extern "C"
// k_step2d_hyb<64,4> (step2d_hyb.cuh): the shipped barotropic kernel.  rows per block: as few as keep the whole grid resident in ONE wave
// (occupancy x SMs block slots) -- a block's run time is (rows+1) row iterations of ~2 us each, so a second wave doubles it
// (measured, profiles/r2_step2d_variants.txt: 1024x1024 on 148 SMs -> 17 strips x 52 row blocks of 20 rows, 58 us)
static int s2d_hyb_launch(roms_ctx* h, const S2D& s, const char* name) {
  const KP& p = h->kp;
  S2PlanesH ap;
  ap.a[HA_H] = p.h; ap.a[HA_ZS] = s.zs; ap.a[HA_ZB] = s.zb; ap.a[HA_ZO] = s.zo; ap.a[HA_US] = s.us; ap.a[HA_UB] = s.ub; ap.a[HA_UO] = s.uo;
  ap.a[HA_DNU] = p.dn_u; ap.a[HA_PM] = p.pm; ap.a[HA_PN] = p.pn; ap.a[HA_RHOA] = p.rhoA; ap.a[HA_RHOS] = p.rhoS;
  constexpr int TXn = 64, Sn = 4, MINBn = 5;     // bound 5, not 6: ptxas keeps 168 registers (6 blocks fit) but must not spill -- a 64-byte spill costs 45 %
  const int strips = (p.iend - p.istr + 1 + TXn - 2) / (TXn - 1), Mrows = p.jend - p.jstr + 1;
  const size_t sm_ = ((size_t)Sn * HA_N * (TXn + 4) + 2 * TXn) * sizeof(double) + Sn * sizeof(unsigned long long);
  static int occ_ = 0;
  if (!occ_) {
    CK(cudaFuncSetAttribute((k_step2d_hyb<TXn, Sn, MINBn>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_));
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_, (k_step2d_hyb<TXn, Sn, MINBn>), TXn, sm_));
    if (occ_ < 1) occ_ = 1;
  }
  int rows = h->s2d_rows;
  if (rows <= 0) {
    const int slots = h->n_sm * occ_, by = slots / strips > 0 ? slots / strips : 1;
    rows = (Mrows + by - 1) / by;
    if (rows < 8) rows = 8;
  }
  dim3 b_(TXn, 1, 1), g_(strips, (Mrows + rows - 1) / rows, 1);
  if (h->ktiming) kt_begin(h, name);
  k_step2d_hyb<TXn, Sn, MINBn><<<g_, b_, sm_, h->st>>>(p, s, ap, rows);
  if (h->ktiming) kt_end(h);
  h->launches++;
  return 0;
}
// operand planes of the TMA marching kernel, in the order of step2d_tma.cuh
static void s2d_planes(const KP& p, const S2D& s, S2Planes& ap) {
  ap.a[A_H] = p.h; ap.a[A_ZS] = s.zs; ap.a[A_ZB] = s.zb; ap.a[A_ZO] = s.zo; ap.a[A_VS] = s.vs; ap.a[A_VB] = s.vb; ap.a[A_VO] = s.vo; ap.a[A_DMV] = p.dm_v;
  ap.a[A_US] = s.us; ap.a[A_UB] = s.ub; ap.a[A_UO] = s.uo; ap.a[A_DNU] = p.dn_u; ap.a[A_PM] = p.pm; ap.a[A_PN] = p.pn; ap.a[A_RM] = p.rmask;
  ap.a[A_SWF] = p.swflx; ap.a[A_RHOA] = p.rhoA; ap.a[A_RHOS] = p.rhoS; ap.a[A_UM] = p.umask; ap.a[A_VM] = p.vmask; ap.a[A_RUF] = p.rufrc;
  ap.a[A_RVF] = p.rvfrc; ap.a[A_ZT] = p.Zt_avg1; ap.a[A_DU1] = p.DU_avg1; ap.a[A_DV1] = p.DV_avg1; ap.a[A_DU2] = p.DU_avg2; ap.a[A_DV2] = p.DV_avg2;
}
extern "C" {
int roms_step2d(roms_ctx* h) {
  FLUSH_RHO(h);
  CK(cudaSetDevice(h->device));
  const KP& p = h->kp;
  const int iif = h->iif, kstp = h->kstp, knew = h->knew;
  int kbak, kold;
  S2D s;
  if (iif == 1) {
    kbak = kstp; kold = kstp;
    s.fwd = 1.0; s.fwd1 = 0.0; s.fwd2 = 0.0; s.bkw_new = 0.0; s.bkw = 1.0; s.bkw1 = 0.0; s.bkw2 = 0.0;
  } else if (iif == 2) {
    kbak = kstp - 1; if (kbak < 1) kbak = 4;
    kold = kbak;
    s.fwd = 1.0; s.fwd1 = 0.0; s.fwd2 = 0.0;
    s.bkw_new = 1.0833333333333; s.bkw = -0.1666666666666; s.bkw1 = 0.0833333333333; s.bkw2 = 0.0;
  } else {
    kbak = kstp - 1; if (kbak < 1) kbak = 4;
    kold = kbak - 1; if (kold < 1) kold = 4;
    s.fwd = 1.781105; s.fwd1 = -1.06221; s.fwd2 = 0.281105;
    s.bkw_new = 0.614; s.bkw = 0.285; s.bkw1 = 0.088; s.bkw2 = 0.013;
  }
  s.zs = h->zeta(kstp); s.zb = h->zeta(kbak); s.zo = h->zeta(kold);
  s.us = h->ubar(kstp); s.ub = h->ubar(kbak); s.uo = h->ubar(kold);
  s.vs = h->vbar(kstp); s.vb = h->vbar(kbak); s.vo = h->vbar(kold);
  s.zn = h->zeta(knew); s.un = h->ubar(knew); s.vn = h->vbar(knew);
  s.w1 = h->w1[iif]; s.w2 = h->w2[iif];
  s.iif = iif; s.nfast = h->nfast;
  s.region = 0;
  const bool multi = h->c.np_xi * h->c.np_eta > 1;
  if (multi && h->overlap && p.Lm >= 8 && p.Mm >= 8) {
    // ---- halo swaps of the barotropic loop hidden behind compute (north_star; the reference blocks in exchange_xxx, mpi_exchanges.F:672-800).
    // Substep n is launched as INTERIOR (all points at least two cells from the subdomain's edges: they read no halo) on the main
    // stream and RING (the two outermost cells, which read the halos of substeps n-1..n-3 and fill the messages of n) followed by
    // the pack/flag/unpack pair on a second, high-priority stream.  interior(n) waits for ring(n-1), ring(n) waits for
    // interior(n-1) and (stream order) for unpack(n-1): the swap of substep n runs while interior(n) and interior(n+1) compute.
    const int e = iif & 1, o = e ^ 1;
    dim3 b_(S2_TX, S2_TY, 1), g_((p.iend + IOFF + 1 + S2_TX - 1) / S2_TX, (p.jend - p.jstr + 1 + S2_TY - 1) / S2_TY, 1);
    if (iif == 1) { CK(cudaEventRecord(h->ev_join, h->st)); CK(cudaStreamWaitEvent(h->st2, h->ev_join, 0)); }
    else CK(cudaStreamWaitEvent(h->st, h->ev_edge[o], 0));
    S2D si = s, sr = s;
    si.region = 1; sr.region = 2;
    const int s2mode_ = (p.bry2d && h->s2d_tma != 12) ? 0 : h->s2d_tma;   // the marching kernels carry the closed-wall code; as INTERIOR kernel (no boundary point in its mask) the shipped one is fine
    if (s2mode_ == 12) {     // interior by the shipped marching kernel (its emit mask skips the ring)
      int rc = s2d_hyb_launch(h, si, "k_step2d_hyb(interior)");
      if (rc) return rc;
      h->launches--;
    } else if (s2mode_ >= 1 && s2mode_ <= 5) {
      S2Planes ap;
      s2d_planes(p, s, ap);
      const int rows = h->s2d_rows > 0 ? h->s2d_rows : 16;
      const size_t sm_ = ((size_t)4 * S2M_NA * (64 + 4) + 2 * 64) * sizeof(double) + 4 * sizeof(unsigned long long);
      static bool attr_ = false;
      if (!attr_) { CK(cudaFuncSetAttribute((k_step2d_tma<64, 4>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_)); attr_ = true; }
      dim3 bt_(64 + 32, 1, 1), gt_((p.iend - p.istr + 1 + 64 - 2) / (64 - 1), (p.jend - p.jstr + 1 + rows - 1) / rows, 1);
      if (h->ktiming) kt_begin(h, "k_step2d_tma(interior)");
      k_step2d_tma<64, 4><<<gt_, bt_, sm_, h->st>>>(p, si, ap, rows);
      if (h->ktiming) kt_end(h);
    } else {
      if (h->ktiming) kt_begin(h, "k_step2d(interior)");
      k_step2d<<<g_, b_, 0, h->st>>>(p, si);
      if (h->ktiming) kt_end(h);
    }
    CK(cudaEventRecord(h->ev_int[e], h->st));
    if (iif > 1) CK(cudaStreamWaitEvent(h->st2, h->ev_int[o], 0));
    if (h->ktiming) kt_begin(h, "k_step2d(ring)", h->st2);
    k_step2d<<<g_, b_, 0, h->st2>>>(p, sr);
    if (h->ktiming) kt_end(h, h->st2);
    h->launches += 2;
    if (p.bry2d) {
      const int len_ = (p.Lm > p.Mm ? p.Lm : p.Mm) + 3;
      dim3 bw_(128, 1, 1), gw_((len_ + 127) / 128, 6, 1);
      k_step2d_bry<<<gw_, bw_, 0, h->st2>>>(p, s);
      h->launches++;
    } else if (p.wwall || p.ewall) {
      dim3 bw_(128, 1, 1), gw_((p.jendR - p.jstrR + 1 + 127) / 128, 2, 1);
      k_step2d_ew<<<gw_, bw_, 0, h->st2>>>(p, s);
      h->launches++;
    }
    CK(cudaEventRecord(h->ev_edge[e], h->st2));
    int rc = exchange(h, {{s.zn, 1}, {s.un, 1}, {s.vn, 1}}, h->st2);
    if (rc) return rc;
    if (iif == h->nfast) {       // the 3-D part that follows needs the whole level and its halos
      CK(cudaEventRecord(h->ev_join, h->st2));
      CK(cudaStreamWaitEvent(h->st, h->ev_join, 0));
      return roms_set_depth(h, 0);
    }
    return 0;
  }
  if (p.bry2d) {
    // open edges: interior by the tile kernel (it leaves every boundary point alone in this mode), then one thread per boundary point
    dim3 b_(S2_TX, S2_TY, 1), g_((p.iend + IOFF + 1 + S2_TX - 1) / S2_TX, (p.jend - p.jstr + 1 + S2_TY - 1) / S2_TY, 1);
    if (h->s2d_tma == 12 && p.Lm >= 8 && p.Mm >= 8) {
      // the marching kernel for everything but the ring of width 2 (its wall code never runs there), the tile kernel for the ring
      S2D si = s, sr = s;
      si.region = 1; sr.region = 2;
      int rc = s2d_hyb_launch(h, si, "k_step2d_hyb(interior)");
      if (rc) return rc;
      if (h->ktiming) kt_begin(h, "k_step2d(ring)");
      k_step2d<<<g_, b_, 0, h->st>>>(p, sr);
      if (h->ktiming) kt_end(h);
    } else {
      if (h->ktiming) kt_begin(h, "k_step2d");
      k_step2d<<<g_, b_, 0, h->st>>>(p, s);
      if (h->ktiming) kt_end(h);
    }
    const int len_ = (p.Lm > p.Mm ? p.Lm : p.Mm) + 3;
    dim3 bw_(128, 1, 1), gw_((len_ + 127) / 128, 6, 1);
    if (h->ktiming) kt_begin(h, "k_step2d_bry");
    k_step2d_bry<<<gw_, bw_, 0, h->st>>>(p, s);
    if (h->ktiming) kt_end(h);
    h->launches += 2;
  } else if (h->s2d_tma == 12) {
    int rc = s2d_hyb_launch(h, s, "k_step2d_hyb");
    if (rc) return rc;
  } else if (h->s2d_tma > 12) {
    // other shapes of k_step2d_hyb (tuning)
    S2PlanesH ap;
    ap.a[HA_H] = p.h; ap.a[HA_ZS] = s.zs; ap.a[HA_ZB] = s.zb; ap.a[HA_ZO] = s.zo; ap.a[HA_US] = s.us; ap.a[HA_UB] = s.ub; ap.a[HA_UO] = s.uo;
    ap.a[HA_DNU] = p.dn_u; ap.a[HA_PM] = p.pm; ap.a[HA_PN] = p.pn; ap.a[HA_RHOA] = p.rhoA; ap.a[HA_RHOS] = p.rhoS;
    const int rows = h->s2d_rows > 0 ? h->s2d_rows : 20;
#define S2H_LAUNCH(TXn, Sn, MINBn)                                                                                                                      \
  do {                                                                                                                                             \
    const size_t sm_ = ((size_t)Sn * HA_N * (TXn + 4) + 2 * TXn) * sizeof(double) + Sn * sizeof(unsigned long long);                                \
    static bool attr_ = false;                                                                                                                     \
    if (!attr_) { CK(cudaFuncSetAttribute((k_step2d_hyb<TXn, Sn, MINBn>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_)); attr_ = true; }        \
    dim3 b_(TXn, 1, 1), g_((p.iend - p.istr + 1 + TXn - 2) / (TXn - 1), (p.jend - p.jstr + 1 + rows - 1) / rows, 1);                                \
    if (h->ktiming) kt_begin(h, "k_step2d_hyb");                                                                                                   \
    k_step2d_hyb<TXn, Sn, MINBn><<<g_, b_, sm_, h->st>>>(p, s, ap, rows);                                                                                 \
    if (h->ktiming) kt_end(h);                                                                                                                     \
    h->launches++;                                                                                                                                 \
  } while (0)
    if (h->s2d_tma == 13) S2H_LAUNCH(64, 4, 5);
    else if (h->s2d_tma == 14) S2H_LAUNCH(64, 4, 4);
    else if (h->s2d_tma == 15) S2H_LAUNCH(32, 4, 8);
    else S2H_LAUNCH(32, 4, 10);
#undef S2H_LAUNCH
  } else if (h->s2d_tma >= 6) {
    // register-marching kernel (step2d_march.cuh): strips of TX-1 columns x s2d_rows rows per block
    const int rows = h->s2d_rows > 0 ? h->s2d_rows : 16;
#define S2R_LAUNCH(TXn, MB)                                                                                                   \
  do {                                                                                                                      \
    dim3 b_(TXn, 1, 1), g_((p.iend - p.istr + 1 + TXn - 2) / (TXn - 1), (p.jend - p.jstr + 1 + rows - 1) / rows, 1);         \
    if (h->ktiming) kt_begin(h, "k_step2d_march");                                                                          \
    k_step2d_march<TXn, MB><<<g_, b_, 0, h->st>>>(p, s, rows);                                                                \
    if (h->ktiming) kt_end(h);                                                                                              \
    h->launches++;                                                                                                          \
  } while (0)
    if (h->s2d_tma == 7) S2R_LAUNCH(128, 1);
    else if (h->s2d_tma == 8) S2R_LAUNCH(32, 1);
    else if (h->s2d_tma == 9) S2R_LAUNCH(64, 8);
    else if (h->s2d_tma == 10) S2R_LAUNCH(128, 4);
    else if (h->s2d_tma == 11) S2R_LAUNCH(64, 6);
    else S2R_LAUNCH(64, 1);
#undef S2R_LAUNCH
  } else if (h->s2d_tma) {
    // TMA-fed marching kernel (step2d_tma.cuh): strips of S2M_TX-1 columns x s2d_rows rows per block
    S2Planes ap;
    s2d_planes(p, s, ap);
    const int rows = h->s2d_rows > 0 ? h->s2d_rows : 16;
#define S2M_LAUNCH(TXn, Sn)                                                                                                                        \
  do {                                                                                                                                             \
    const size_t sm_ = ((size_t)Sn * S2M_NA * (TXn + 4) + 2 * TXn) * sizeof(double) + Sn * sizeof(unsigned long long);                              \
    static bool attr_ = false;                                                                                                                     \
    if (!attr_) { CK(cudaFuncSetAttribute((k_step2d_tma<TXn, Sn>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm_)); attr_ = true; }        \
    dim3 b_(TXn + 32, 1, 1), g_((p.iend - p.istr + 1 + TXn - 2) / (TXn - 1), (p.jend - p.jstr + 1 + rows - 1) / rows, 1);                           \
    if (h->ktiming) kt_begin(h, "k_step2d_tma");                                                                                                   \
    k_step2d_tma<TXn, Sn><<<g_, b_, sm_, h->st>>>(p, s, ap, rows);                                                                                 \
    if (h->ktiming) kt_end(h);                                                                                                                     \
    h->launches++;                                                                                                                                 \
  } while (0)
    switch (h->s2d_tma) {
      case 2: S2M_LAUNCH(128, 3); break;
      case 3: S2M_LAUNCH(128, 4); break;
      case 4: S2M_LAUNCH(256, 3); break;
      case 5: S2M_LAUNCH(64, 3); break;
      default: S2M_LAUNCH(64, 4); break;
    }
#undef S2M_LAUNCH
  } else {
    dim3 b_(S2_TX, S2_TY, 1), g_((p.iend + IOFF + 1 + S2_TX - 1) / S2_TX, (p.jend - p.jstr + 1 + S2_TY - 1) / S2_TY, 1);
    if (h->ktiming) kt_begin(h, "k_step2d");
    k_step2d<<<g_, b_, 0, h->st>>>(p, s);
    if (h->ktiming) kt_end(h);
    h->launches++;
  }
  if (!p.bry2d && (p.wwall || p.ewall)) {
    dim3 b_(128, 1, 1), g_((p.jendR - p.jstrR + 1 + 127) / 128, 2, 1);
    if (h->ktiming) kt_begin(h, "k_step2d_ew");
    k_step2d_ew<<<g_, b_, 0, h->st>>>(p, s);
    if (h->ktiming) kt_end(h);
    h->launches++;
  }
  if (iif == h->nfast) {
    int rc = roms_set_depth(h, 0);
    if (rc) return rc;
  }
  if (h->c.np_xi * h->c.np_eta == 1 && h->c.ew_periodic) return 0;      // periodic images already stored by the kernel
  return exchange(h, {{s.zn, 1}, {s.un, 1}, {s.vn, 1}});
}
int roms_step3d_uv2(roms_ctx* h, int tile) {
  FLUSH_RHO(h);
  (void)tile;
  CK(cudaSetDevice(h->device));
  const KP& p = h->kp;
  const int nnew = h->nnew, nstp = h->nstp, knew = h->knew;
  if (h->uv2_sm && h->uv2_fused && !p.bry3d) {
    // both passes in one kernel per direction; the wall rows / columns (u3dbc, v3dbc between the passes) first, from the original u(nnew)
    const double g2 = h->c.gamma2;
#define UV2_EDGE(DIRn, arr, sarr, barr, a0, a1, b0, b1) COLUMN_SM((k_uv2_sm<DIRn, true>), a1, b0, b1, p, arr, sarr, barr, g2, a0, a1, b0, b1)
    if (p.swall) UV2_EDGE(0, h->u(nnew), h->u(nstp), h->ubar(knew), p.istr, p.iendR, p.jstr - 1, p.jstr - 1);
    if (p.nwall) UV2_EDGE(0, h->u(nnew), h->u(nstp), h->ubar(knew), p.istr, p.iendR, p.jend + 1, p.jend + 1);
    if (p.wwall) UV2_EDGE(0, h->u(nnew), h->u(nstp), h->ubar(knew), p.istr, p.istr, p.jstr, p.jend);
    if (p.ewall) UV2_EDGE(0, h->u(nnew), h->u(nstp), h->ubar(knew), p.iend + 1, p.iend + 1, p.jstr, p.jend);
    if (p.wwall) UV2_EDGE(1, h->v(nnew), h->v(nstp), h->vbar(knew), p.istr - 1, p.istr - 1, p.jstr, p.jendR);
    if (p.ewall) UV2_EDGE(1, h->v(nnew), h->v(nstp), h->vbar(knew), p.iend + 1, p.iend + 1, p.jstr, p.jendR);
    if (p.swall) UV2_EDGE(1, h->v(nnew), h->v(nstp), h->vbar(knew), p.istr, p.iend, p.jstr, p.jstr);
    if (p.nwall) UV2_EDGE(1, h->v(nnew), h->v(nstp), h->vbar(knew), p.istr, p.iend, p.jend + 1, p.jend + 1);
#undef UV2_EDGE
    COLUMN_SM((k_uv2_sm<0, false>), p.iend, p.jstr, p.jend, p, h->u(nnew), h->u(nstp), h->ubar(knew), g2, p.istrU, p.iend, p.jstr, p.jend);
    COLUMN_SM((k_uv2_sm<1, false>), p.iend, p.jstrV, p.jend, p, h->v(nnew), h->v(nstp), h->vbar(knew), g2, p.istr, p.iend, p.jstrV, p.jend);
    if (p.rivers) PLANE(k_river_uv, p.iend, p.jstr, p.jend, 1, p, h->u(nnew), h->v(nnew), p.istr, p.jstr);   // step3d_uv2.F:689-717
    int rc = exchange(h, {{p.FlxU, p.N}, {h->u(nnew), p.N}, {h->ubar(knew), 1}});
    if (rc) return rc;
    return exchange(h, {{p.FlxV, p.N}, {h->v(nnew), p.N}, {h->vbar(knew), 1}});
  }
  COLUMN(k_uv2_pass0<0>, 0, p.iend, p.jstr, p.jend, p, h->u(nnew));
  COLUMN(k_uv2_pass0<1>, 0, p.iend, p.jstrV, p.jend, p, h->v(nnew));
  bc3d(h, h->u(nnew), p.N, BC_U);     // u3dbc_tile / v3dbc_tile (step3d_uv2.F:471-472)
  bc3d(h, h->v(nnew), p.N, BC_V);
  if (h->uv2_sm) {
    COLUMN_SM(k_uv2_pass1_sm<0>, p.iendR, p.jstrR, p.jendR, p, h->u(nnew), h->u(nstp), h->ubar(knew));
    COLUMN_SM(k_uv2_pass1_sm<1>, p.iendR, p.jstr, p.jendR, p, h->v(nnew), h->v(nstp), h->vbar(knew));
  } else {
    COLUMN(k_uv2_pass1<0>, 0, p.iendR, p.jstrR, p.jendR, p, h->u(nnew), h->u(nstp), h->ubar(knew));
    COLUMN(k_uv2_pass1<1>, 0, p.iendR, p.jstr, p.jendR, p, h->v(nnew), h->v(nstp), h->vbar(knew));
  }
  if (p.rivers) PLANE(k_river_uv, p.iend, p.jstr, p.jend, 1, p, h->u(nnew), h->v(nnew), p.istr, p.jstr);   // step3d_uv2.F:689-717
  int rc = exchange(h, {{p.FlxU, p.N}, {h->u(nnew), p.N}, {h->ubar(knew), 1}});
  if (rc) return rc;
  return exchange(h, {{p.FlxV, p.N}, {h->v(nnew), p.N}, {h->vbar(knew), 1}});
}
int roms_step3d_t(roms_ctx* h, int tile) {
  FLUSH_RHO(h);
  (void)tile;
  CK(cudaSetDevice(h->device));
  const KP& p = h->kp;
  const int nrhs = h->nrhs, nnew = h->nnew, NT = h->c.NT;
  for (const TPlan& pl : plan_batches(h)) {
    int rc = tracer_cor_horiz(h, pl, nrhs, nnew);
    if (!rc) rc = tracer_vert(h, 1, pl, nrhs, nnew, h->c.dt);
    if (rc) return rc;
  }
  for (int it = 1; it <= NT; ++it) bc3d(h, h->t(nnew, it), p.N, BC_T, it);   // t3dbc_tile (step3d_t_ISO.F:1157-1159)
  for (int it = 1; it + 1 <= NT; it += 2) {
    int rc = exchange(h, {{h->t(nnew, it), p.N}, {h->t(nnew, it + 1), p.N}});
    if (rc) return rc;
  }
  if (NT % 2 == 1) return exchange(h, {{h->t(nnew, NT), p.N}});
  return 0;
}
// visc3d_S.F:4-134 (UV_VIS2).  Called between step3d_uv1 and the barotropic loop (main.F:448-450): u,v(3-nstp) hold Hz*u.
int roms_visc3d(roms_ctx* h) {
  FLUSH_RHO(h);
  CK(cudaSetDevice(h->device));
  const KP& p = h->kp;
  const int nstp = h->nstp, indx = 3 - nstp;
  PLANE(k_visc3d, p.iend, p.jstr, p.jend, 1, p, h->u(nstp), h->v(nstp), h->u(indx), h->v(indx));
  return 0;
}
// t3dmix_S.F:4-258 (TS_DIF2 branch), every tracer, each followed by its exchange (:254-256)
int roms_t3dmix(roms_ctx* h) {
  FLUSH_RHO(h);
  CK(cudaSetDevice(h->device));
  const KP& p = h->kp;
  const int nrhs = h->nrhs, nnew = h->nnew;
  for (int it = 1; it <= h->c.NT; ++it) {
    PLANE((k_t3dmix<KB_T>), p.iend, p.jstr, p.jend, (p.N + KB_T - 1) / KB_T, p, h->t(nrhs, it), h->t(nnew, it), h->f[ROMS_DIFF2].ptr[it - 1]);
    int rc = exchange(h, {{h->t(nnew, it), p.N}});
    if (rc) return rc;
  }
  return 0;
}
int roms_alfabeta(roms_ctx* h, double* alpha_host, double* beta_host) {
  FLUSH_RHO(h);
  CK(cudaSetDevice(h->device));
  const KP& p = h->kp;
  const double* T = h->t(h->nstp, 1);
  const double* S = h->c.NT >= 2 ? h->t(h->nstp, 2) : T;
  COLUMN(k_alfabeta, 0, p.iend + 2, p.jstr - 2, p.jend + 2, p, T, S, h->d_o[0], h->d_o[1]);
  const size_t w = (size_t)(h->c.Lm + 4) * sizeof(double);
  CK(cudaMemcpy2DAsync(alpha_host, w, h->d_o[0] + (IOFF - 1), (size_t)p.SX * sizeof(double), w, h->c.Mm + 4, cudaMemcpyDeviceToHost, h->st));
  CK(cudaMemcpy2DAsync(beta_host, w, h->d_o[1] + (IOFF - 1), (size_t)p.SX * sizeof(double), w, h->c.Mm + 4, cudaMemcpyDeviceToHost, h->st));
  CK(cudaStreamSynchronize(h->st));
  return 0;
}

// diag.F:50-658.  Column values on the device; the sums are formed on the host in a fixed order
// (row-wise pairwise) so that the result is deterministic.  Multi-GPU: each rank reports its own
// partial sums; roms_diag combines them through the communicator.
static double pairsum(const double* a, size_t n) {
  if (n <= 8) { double s = 0; for (size_t i = 0; i < n; ++i) s += a[i]; return s; }
  return pairsum(a, n / 2) + pairsum(a + n / 2, n - n / 2);
}
int roms_set_pot_tides(roms_ctx* h, int on) { h->kp.pot_tides = on ? 1 : 0; return 0; }
// cppdefs LMD_MIXING + LMD_KPP + LMD_BKPP + LMD_RIMIX + LMD_NONLOCAL: allocates the closure's arrays and makes roms_step call
// roms_lmd_vmix where main.F:408-410 and :434-436 do.  Call before roms_init_step0.
int roms_set_kpp(roms_ctx* h, int on) {
  CK(cudaSetDevice(h->device));
  KP& p = h->kp;
  if (!on) { p.lmd_kpp = 0; return 0; }
  if (!h->c.nonlin_eos) { g_err = "roms_set_kpp: the Brunt-Vaisala frequency is implemented for the nonlinear split EOS only"; return -1; }
  if (!h->kpp_buf) {
    const size_t n2 = (size_t)p.SXY, nw = n2 * (p.N + 1), tot = 6 * nw + 13 * n2;
    CK(cudaMalloc(&h->kpp_buf, tot * sizeof(double)));
    CK(cudaMemset(h->kpp_buf, 0, tot * sizeof(double)));
    double* cur = h->kpp_buf;
    auto take = [&](size_t n) { double* r = cur; cur += n; return r; };
    p.bvf = take(nw); p.swr_frac = take(nw); p.ghat = take(nw);
    h->kpp_Rig = take(nw); h->kpp_RigS = take(nw); h->kpp_FC = take(nw);
    p.hbls = take(n2); p.hbbl = take(n2); take(n2);   /* (srflx is a field of the arena: ROMS_SRFLX) */
    p.k_ustar = take(n2); p.k_Bo = take(n2); p.k_Bosol = take(n2); p.k_hbl = take(n2); p.k_bbl = take(n2);
    h->kpp_hblS = take(n2); h->kpp_bblS = take(n2); h->kpp_alpha = take(n2); h->kpp_beta = take(n2);
  }
  p.lmd_kpp = 1;
  return 0;
}
int roms_set_kpp_opts(roms_ctx* h, int lmd_convec, double Akv_bak, double Akt_bak_T, double Akt_bak_S) {
  h->kpp_convec = lmd_convec ? 1 : 0; h->kpp_Akv_bak = Akv_bak; h->kpp_Akt_bak[0] = Akt_bak_T; h->kpp_Akt_bak[1] = Akt_bak_S;
  return 0;
}
int roms_get_hbl(roms_ctx* h, double* hbls, double* hbbl) {
  CK(cudaSetDevice(h->device));
  if (!h->kp.lmd_kpp) { g_err = "roms_get_hbl: KPP is off"; return -1; }
  const KP& p = h->kp;
  const size_t w = (size_t)(h->c.Lm + 4) * sizeof(double);
  CK(cudaMemcpy2DAsync(hbls, w, p.hbls + (IOFF - 1), (size_t)p.SX * sizeof(double), w, h->c.Mm + 4, cudaMemcpyDeviceToHost, h->st));
  CK(cudaMemcpy2DAsync(hbbl, w, p.hbbl + (IOFF - 1), (size_t)p.SX * sizeof(double), w, h->c.Mm + 4, cudaMemcpyDeviceToHost, h->st));
  CK(cudaStreamSynchronize(h->st));
  return 0;
}
// lmd_vmix(tind) (lmd_vmix.F:4-29): lmd_vmix_tile then lmd_kpp_tile; result in Akv, Akt(:,:,:,1:2), hbls, hbbl, ghat
int roms_lmd_vmix(roms_ctx* h, int tind) {
  FLUSH_RHO(h);
  CK(cudaSetDevice(h->device));
  const KP& p = h->kp;
  if (!p.lmd_kpp) { g_err = "roms_lmd_vmix: call roms_set_kpp first"; return -1; }
  const int N = p.N;
  const int imin = p.wwall ? p.istr : p.istr - 1, imax = p.ewall ? p.iend : p.iend + 1, jmin = p.swall ? p.jstr : p.jstr - 1, jmax = p.nwall ? p.jend : p.jend + 1;
  const int len = (p.Lm > p.Mm ? p.Lm : p.Mm) + 3;
  const dim3 eb(128, 1, 1);
  PLANE(k_lmd_rig, imax, jmin, jmax, N - 1, p, h->u(tind), h->v(tind), h->kpp_Rig, imin, imax, jmin, jmax);
  k_lmd_edges<<<dim3((len + 127) / 128, 5, N - 1), eb, 0, h->st>>>(p, h->kpp_Rig + p.SXY, imin, imax, jmin, jmax);
  h->launches++;
  PLANE(k_lmd_smooth, p.iend, p.jstr, p.jend, N - 1, p, h->kpp_Rig + p.SXY, h->kpp_RigS + p.SXY, 0);
  COLUMN(k_lmd_vmix_col, 0, p.iend, p.jstr, p.jend, p, h->kpp_RigS, h->kpp_Akv_bak, h->kpp_Akt_bak[0], h->kpp_Akt_bak[1], h->kpp_convec);
  COLUMN(k_alfabeta, 0, p.iend + 2, p.jstr - 2, p.jend + 2, p, h->t(h->nstp, 1), h->c.NT >= 2 ? h->t(h->nstp, 2) : nullptr, h->kpp_alpha, h->kpp_beta);
  COLUMN(k_lmd_kpp1, 0, imax, jmin, jmax, p, h->u(tind), h->v(tind), h->kpp_alpha, h->kpp_beta, h->f[ROMS_STFLX].ptr[0],
         h->c.NT >= 2 ? h->f[ROMS_STFLX].ptr[1] : (double*)nullptr, h->kpp_FC, imin, imax, jmin, jmax);
  k_lmd_edges<<<dim3((len + 127) / 128, 5, 1), eb, 0, h->st>>>(p, p.k_hbl, imin, imax, jmin, jmax);
  k_lmd_edges<<<dim3((len + 127) / 128, 5, 1), eb, 0, h->st>>>(p, p.k_bbl, imin, imax, jmin, jmax);
  h->launches += 2;
  PLANE(k_lmd_smooth, p.iend, p.jstr, p.jend, 1, p, p.k_hbl, h->kpp_hblS, 1);
  PLANE(k_lmd_smooth, p.iend, p.jstr, p.jend, 1, p, p.k_bbl, h->kpp_bblS, 1);
  COLUMN(k_lmd_kpp2, 0, p.iend, p.jstr, p.jend, p, h->kpp_hblS, h->kpp_bblS, h->u(h->nstp), h->v(h->nstp), h->iic == h->forw_start ? 1 : 0);
  k_lmd_edges<<<dim3((len + 127) / 128, 5, 1), eb, 0, h->st>>>(p, p.hbls, p.istr, p.iend, p.jstr, p.jend);
  k_lmd_edges<<<dim3((len + 127) / 128, 5, 1), eb, 0, h->st>>>(p, p.hbbl, p.istr, p.iend, p.jstr, p.jend);
  h->launches += 2;
  int rc = exchange(h, {{p.Akv, N + 1}, {p.hbls, 1}, {p.hbbl, 1}});
  if (rc) return rc;
  return exchange(h, {{p.Akt[0], N + 1}, {p.Akt[1], N + 1}});
}
// pipe_frc.F (init_arrays_pipes, set_pipe_frc): release profile and tracer values of one pipe
int roms_set_pipe(roms_ctx* h, int ipipe, const double* prf, const double* trc) {
  CK(cudaSetDevice(h->device));
  if (ipipe < 1 || ipipe > MAXPIP || !prf || !trc) { g_err = "roms_set_pipe: pipe index out of range (1..8)"; return -1; }
  const int N = h->c.N;
  std::vector<double> row(PIPROW(N), 0.0);
  for (int k = 0; k < N; ++k) row[k] = prf[k];
  for (int it = 0; it < h->c.NT; ++it) row[N + it] = trc[it];
  CK(cudaMemcpyAsync(h->kp.pipe_tab + (size_t)(ipipe - 1) * PIPROW(N), row.data(), row.size() * sizeof(double), cudaMemcpyHostToDevice, h->st));
  CK(cudaStreamSynchronize(h->st));
  h->kp.pipes = 1;
  return 0;
}
// river_frc.F set_river_frc: volume flux and tracer values of one river; the table lives next to the fields in the arena
int roms_set_river(roms_ctx* h, int iriver, double vol, const double* trc) {
  CK(cudaSetDevice(h->device));
  if (iriver < 1 || iriver > MAXRIV || !trc) { g_err = "roms_set_river: river index out of range (1..8)"; return -1; }
  double row[RIVROW] = {0.0};
  row[0] = vol;
  for (int it = 0; it < h->c.NT; ++it) row[1 + it] = trc[it];
  CK(cudaMemcpyAsync(h->kp.riv_tab + (size_t)(iriver - 1) * RIVROW, row, sizeof(row), cudaMemcpyHostToDevice, h->st));
  CK(cudaStreamSynchronize(h->st));
  h->kp.rivers = 1;
  return 0;
}
// setup_grid2.F:64-100 for this subdomain (the reference prints the global values once at start-up): extrema over the water
// points of the rho-range incl. boundary rows on a physical edge, unmasked area and volume.  A one-off start-up report
// like the reference's: the five planes are pulled and scanned on the host in index order.
int roms_grid_stats(roms_ctx* h, double out[9]) {
  CK(cudaSetDevice(h->device));
  const KP& p = h->kp;
  const int Lm = h->c.Lm, Mm = h->c.Mm, SXh = Lm + 4;
  const size_t n2 = (size_t)SXh * (Mm + 4);
  std::vector<double> pmv(n2), pnv(n2), hv(n2), rmv(n2), fv(n2);
  if (roms_pull(h, ROMS_PM, 1, 1, pmv.data()) || roms_pull(h, ROMS_PN, 1, 1, pnv.data()) || roms_pull(h, ROMS_H, 1, 1, hv.data()) ||
      roms_pull(h, ROMS_RMASK, 1, 1, rmv.data()) || roms_pull(h, ROMS_F, 1, 1, fv.data()))
    return -1;
  double hmin = 1e20, hmax = -1e20, gmin = 1e20, gmax = -1e20, cmin = 1e20, cmax = -1e20, ccor = -1e20, area = 0, vol = 0;
  for (int j = p.jstrR; j <= p.jendR; ++j)
    for (int i = p.istrR; i <= p.iendR; ++i) {
      const size_t o = (size_t)(j + 1) * SXh + (i + 1);
      if (rmv[o] > 0.5) {
        hmin = std::min(hmin, hv[o]); hmax = std::max(hmax, hv[o]);
        double cff = 1. / sqrt(pmv[o] * pnv[o]);
        gmin = std::min(gmin, cff); gmax = std::max(gmax, cff);
        cff = p.dtfast * sqrt(p.g * hv[o] * (pmv[o] * pmv[o] + pnv[o] * pnv[o]));
        cmin = std::min(cmin, cff); cmax = std::max(cmax, cff);
        cff = p.dt * fabs(fv[o]);
        ccor = std::max(ccor, cff);
      }
    }
  for (int j = p.jstr; j <= p.jend; ++j)
    for (int i = p.istr; i <= p.iend; ++i) {
      const size_t o = (size_t)(j + 1) * SXh + (i + 1);
      const double dA = rmv[o] / (pmv[o] * pnv[o]);
      area += dA; vol += dA * hv[o];
    }
  out[0] = hmin; out[1] = hmax; out[2] = gmin; out[3] = gmax; out[4] = cmin; out[5] = cmax; out[6] = ccor; out[7] = area; out[8] = vol;
  return 0;
}
int roms_diag(roms_ctx* h, double out[4]) {
  FLUSH_RHO(h);
  CK(cudaSetDevice(h->device));
  const KP& p = h->kp;
  const int Mm = h->c.Mm;
  COLUMN(k_diag_bar, 0, p.iend + 1, p.jstr, p.jend + 1, p, h->u(h->nstp), h->v(h->nstp), h->d_ub, h->d_vb);
  COLUMN(k_diag_col, 0, p.iend, p.jstr, p.jend, p, h->u(h->nstp), h->v(h->nstp), h->d_ub, h->d_vb, h->d_o[0], h->d_o[1], h->d_o[2], h->d_o[3], h->d_o[4]);
  {
    dim3 b_(128, 1, 1), g_((Mm + 127) / 128, 1, 1);
    if (h->ktiming) kt_begin(h, "k_diag_rows");
    k_diag_rows<<<g_, b_, 0, h->st>>>(p, h->d_o[0], h->d_o[1], h->d_o[2], h->d_o[3], h->d_o[4], h->d_rows);
    if (h->ktiming) kt_end(h);
    k_diag_final<<<1, 32, 0, h->st>>>(h->d_rows, Mm, h->d_rows + (size_t)7 * Mm);
    h->launches += 2;
  }
  double red[7];
  CK(cudaMemcpyAsync(red, h->d_rows + (size_t)7 * Mm, 7 * sizeof(double), cudaMemcpyDeviceToHost, h->st));
  CK(cudaStreamSynchronize(h->st));
  int rc = comm_diag_reduce(&h->comm, red);
  if (rc) return rc;
  const double avke = red[0], avke2b = red[1], avzeta = red[2], volume = red[4];
  out[0] = avke / (volume + avzeta);
  out[1] = avke2b / (volume + avzeta);
  out[2] = red[5];
  out[3] = red[6];
  for (int q = 0; q < 4; ++q)
    if (!std::isfinite(out[q])) { g_err = "BLOWUP: non-finite diagnostics (diag.F:622-634)"; return 1; }
  return 0;
}

// ---- driver: main.F:86-321 (tail of roms_init) and main.F:333-520 (roms_step) -------------------
int roms_init_step0(roms_ctx* h) {
  FLUSH_RHO(h);
  h->iic = 0; h->kstp = 1; h->knew = 1; h->iif = 1; h->nstp = 1; h->nnew = 1; h->nrhs = 1;
  h->forw_start = h->ntstart;
  int rc;
  if ((rc = roms_set_depth(h, 0))) return rc;
  if (h->kp.lmd_kpp) {       // main.F:219 swr_frac
    const KP& p = h->kp;
    COLUMN(k_swr_frac, 0, p.iend, p.jstr, p.jend, p);
    if ((rc = exchange(h, {{p.swr_frac, p.N + 1}}))) return rc;
  }
  if ((rc = roms_set_huv(h))) return rc;
  if ((rc = roms_omega(h))) return rc;
  if ((rc = roms_rho_eos(h, h->nrhs))) return rc;
  return 0;
}
int roms_step(roms_ctx* h) {
  int rc;
  h->iic = (h->iic == 0) ? h->ntstart : h->iic + 1;
  h->nstp = 1 + (h->iic - h->ntstart) % 2;
  h->nrhs = h->nstp; h->nnew = 3;
#define RUN(x) if ((rc = (x))) return rc
  RUN(roms_rho_eos(h, h->nrhs));
  RUN(roms_set_huv(h));
  RUN(roms_omega(h));
  RUN(roms_prsgrd(h));
  // main.F:408-410 calls lmd_vmix before prsgrd; the two share no data, and after prsgrd the deferred rho_eos (which also
  // produces bvf) has run fused with the pressure integration
  if (h->kp.lmd_kpp) RUN(roms_lmd_vmix(h, h->nstp));
  RUN(roms_pre_step3d(h, 0));
  RUN(roms_set_huv1(h, 0));
  h->nrhs = 3; h->nnew = 3 - h->nstp;
  RUN(roms_omega(h));
  RUN(roms_rho_eos(h, h->nrhs));
  RUN(roms_prsgrd(h));
  if (h->kp.lmd_kpp) RUN(roms_lmd_vmix(h, h->nrhs));      // main.F:434-436 (before prsgrd there; independent of it)
  RUN(roms_step3d_uv1(h, 0));
  if (h->c.uv_vis2) RUN(roms_visc3d(h));
  for (h->iif = 1; h->iif <= h->nfast; ++h->iif) {
    h->kstp = h->knew;
    h->knew = h->kstp + 1;
    if (h->knew > 4) h->knew = 1;
    RUN(roms_step2d(h));
  }
  RUN(roms_step3d_uv2(h, 0));
  RUN(roms_omega(h));
  RUN(roms_step3d_t(h, 0));
  if (h->c.ts_dif2) RUN(roms_t3dmix(h));
  RUN(roms_rho_eos(h, h->nnew));
#undef RUN
  return 0;
}
int roms_run(roms_ctx* h, int nsteps) {
  for (int n = 0; n < nsteps; ++n) {
    int rc = roms_step(h);
    if (rc) return rc;
  }
  return 0;
}

}  // extern "C"

#include "comm_impl.cuh"

// ================================================================================================
// roms_omega_noalloc -- an allocation-free variant of the omega kernel (omega.F:17-164 arithmetic of k_omega).
// It has NO counterpart in the reference; it was added for callers that cannot
// call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid allocation).  The two
// per-level column coefficients that k_omega keeps in the context's arena live here in ONE file-scope device array
// of fixed size, held by every GPU that loads the library until the process exits:
//     2 * (N+1) * Lm * Mm  <=  ROMS_NOALLOC_MAX_DOUBLES  (4 Mi doubles = 32 MiB)
// The entry point calls no cudaMalloc/cudaFree and no allocating library routine.  Every call shares the array, which
// keeps the previous call's contents: nothing in it has to start at zero (each column writes its levels 1..N in the
// up-sweep before the down-sweep reads them), but two calls that overlap on the GPU (different streams, threads or
// processes' contexts on one device do not share it; streams and threads of ONE process do) race on it.
#define ROMS_NOALLOC_MAX_DOUBLES (4u << 20)
static __device__ __align__(256) double g_noalloc_scratch[ROMS_NOALLOC_MAX_DOUBLES];

struct OmegaNA {
  int Lm, Mm, N, SXh;
  long SXYh, ncell;
  double dtau;
  const double *FlxU, *FlxV, *z_w, *Hz, *pm, *pn, *swflx;
  double *We, *Wi;
  double* scr;
};
// one thread per (i,j) column of the interior; fields are DEVICE arrays in the reference's unpadded layout
// (-1:Lm+2, -1:Mm+2 [, k]) -- what roms_pull returns, uploaded as is
__global__ void __launch_bounds__(64) k_omega_noalloc(const __grid_constant__ OmegaNA a) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x + 1, j = blockIdx.y * blockDim.y + threadIdx.y + 1;
  if (i > a.Lm || j > a.Mm) return;
  const int N = a.N;
  const long SX = a.SXh, SXY = a.SXYh;
  const long c = (i + 1) + (long)(j + 1) * SX;
  double* __restrict__ s0 = a.scr + ((long)(j - 1) * a.Lm + (i - 1));
  double* __restrict__ s1 = s0 + (long)(N + 1) * a.ncell;
#define NA3(f, x, k) (f)[(x) + (long)((k)-1) * SXY]
#define NAW(f, x, k) (f)[(x) + (long)(k)*SXY]
  const double cu_min = 0.6, cu_max = 1.0, cmnx_ratio = cu_min / cu_max, cutoff = 2.0 - cmnx_ratio, r4cmx = 0.25 / (1.0 - cmnx_ratio);
  double W = 0.0;
  NAW(a.Wi, c, 0) = 0.0;
  for (int k = 1; k <= N; ++k) {
    const double fu1 = NA3(a.FlxU, c + 1, k), fu0 = NA3(a.FlxU, c, k), fv1 = NA3(a.FlxV, c + SX, k), fv0 = NA3(a.FlxV, c, k);
    W = W - fu1 + fu0 - fv1 + fv0;
    s0[(long)k * a.ncell] = DMAX(fu1, 0.0) - DMIN(fu0, 0.0) + DMAX(fv1, 0.0) - DMIN(fv0, 0.0);
    s1[(long)k * a.ncell] = W;
  }
  const double dm_r = 1.0 / a.pm[c], dn_r = 1.0 / a.pn[c];      // setup_grid1.F: dm_r = 1/pm, dn_r = 1/pn
  W = W + a.swflx[c] * dm_r * dn_r;
  const double zw0 = NAW(a.z_w, c, 0);
  const double wrk = W / (NAW(a.z_w, c, N) - zw0);
  NAW(a.Wi, c, N) = 0.0;
  NAW(a.We, c, N) = 0.0;
  NAW(a.We, c, 0) = 0.0;
  const double CX0 = a.dtau * a.pm[c] * a.pn[c];
  double CXkp = s0[(long)N * a.ncell], Hzkp = NA3(a.Hz, c, N);
  for (int k = N - 1; k >= 1; --k) {
    double Wik = s1[(long)k * a.ncell] - wrk * (NAW(a.z_w, c, k) - zw0);
    const double CXk = s0[(long)k * a.ncell], Hzk = NA3(a.Hz, c, k);
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
    NAW(a.We, c, k) = Wek;
    NAW(a.Wi, c, k) = Wik;
    CXkp = CXk; Hzkp = Hzk;
  }
#undef NA3
#undef NAW
}

#include <mutex>
static std::mutex g_noalloc_mu;
static double* g_noalloc_ptr[64] = {nullptr};     // device address of g_noalloc_scratch, per device ordinal

extern "C" int roms_omega_noalloc(int Lm, int Mm, int N, double dtau, const double* FlxU, const double* FlxV, const double* z_w, const double* Hz,
                                  const double* pm, const double* pn, const double* swflx, double* We, double* Wi, void* stream) {
  if (Lm < 1 || Mm < 1 || N < 2 || !FlxU || !FlxV || !z_w || !Hz || !pm || !pn || !swflx || !We || !Wi) { g_err = "roms_omega_noalloc: bad argument"; return -1; }
  const unsigned long long need = 2ull * (unsigned long long)(N + 1) * (unsigned long long)Lm * (unsigned long long)Mm;
  if (need > ROMS_NOALLOC_MAX_DOUBLES) {
    char msg[256];
    snprintf(msg, sizeof msg, "roms_omega_noalloc: 2*(N+1)*Lm*Mm = %llu doubles exceeds the static scratch of %u doubles (32 MiB); use roms_omega on a context",
             need, (unsigned)ROMS_NOALLOC_MAX_DOUBLES);
    g_err = msg;
    return -3;
  }
  int dev = -1;
  CK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) { g_err = "roms_omega_noalloc: device ordinal out of range"; return -1; }
  double* scr;
  {
    // once per device, with that device current: in host code the array's NAME is only a host-side stand-in for the symbol,
    // not a device address -- the address the kernel needs comes from cudaGetSymbolAddress
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
      void* ptr = nullptr;
      CK(cudaGetSymbolAddress(&ptr, g_noalloc_scratch));
      g_noalloc_ptr[dev] = (double*)ptr;
    }
    scr = g_noalloc_ptr[dev];
  }
  OmegaNA a;
  a.Lm = Lm; a.Mm = Mm; a.N = N; a.SXh = Lm + 4;
  a.SXYh = (long)(Lm + 4) * (Mm + 4); a.ncell = (long)Lm * Mm;
  a.dtau = dtau;
  a.FlxU = FlxU; a.FlxV = FlxV; a.z_w = z_w; a.Hz = Hz; a.pm = pm; a.pn = pn; a.swflx = swflx; a.We = We; a.Wi = Wi;
  a.scr = scr;
  dim3 b(32, 2, 1), g((Lm + 31) / 32, (Mm + 1) / 2, 1);
  k_omega_noalloc<<<g, b, 0, (cudaStream_t)stream>>>(a);
  CK(cudaGetLastError());
  return 0;
}
