/* roms_b200.h -- C ABI of libromsb200.so: the B200-native drop-in for the explicit
 * stencil time step of UCLA-ROMS (reference: dafyddstephenson/ucla-roms, src/).
 *
 * The reference drives its dynamical kernel through argument-free / (tile) Fortran
 * subroutines whose state is module globals (main.F:397-479).  Each export below replaces
 * one of them; a Fortran maintainer binds it with an `interface ... bind(C)` block (see
 * INTEGRATION.md).  All state lives on the device in the reference's (i,j,k) column-major
 * layout with a 2-cell halo: a 2-D field is (-1:Lm+2, -1:Mm+2), rho/u/v-type 3-D fields add
 * k=1:N, w-type fields k=0:N (set_global_definitions.h:146-150).  Host buffers handed to
 * roms_push/roms_pull use exactly that layout, so `c_loc(array)` of the Fortran module
 * arrays can be passed unchanged.
 *
 * Every function returns 0 on success, <0 on CUDA/NCCL failure, >0 when a non-finite state
 * was detected.  No function falls back to the CPU.
 */
#ifndef ROMS_B200_H
#define ROMS_B200_H
#ifdef __cplusplus
extern "C" {
#endif

typedef struct roms_ctx roms_ctx;   /* opaque per-GPU handle */
#define ROMS_MAXNT 40                /* most tracers one context carries (bgc configurations: 34) */

/* Run-time replacement for the reference's compile-time param.opt + *.in keywords
 * (param.opt:20-43, read_inp_mod.F:416-446). Sizes are those of THIS subdomain. */
typedef struct roms_cfg {
  int Lm, Mm, N, NT;          /* interior points of this subdomain; levels; tracers          */
  int ew_periodic, ns_periodic;/* cppdefs EW_PERIODIC / NS_PERIODIC                          */
  int nonlin_eos;             /* 0: linear rho=Tcoef*(T0-T)+Scoef*(S-S0) ; 1: NONLIN+SPLIT_EOS */
  int salinity;               /* cppdefs SALINITY (tracer 2 is S)                             */
  int ndtfast;                /* barotropic steps per baroclinic step (time_stepping keyword) */
  int np_xi, np_eta;          /* process grid (mpi_setup.F:61-70); 1x1 for a single GPU       */
  int inode, jnode;           /* this rank's position in it                                   */
  double dt;                  /* baroclinic time step [s]                                     */
  double rho0, g;
  double Tcoef, T0, Scoef, S0;/* lin_rho_eos keyword                                          */
  double rdrg, rdrg2, Zob;    /* bottom_drag keyword                                          */
  double gamma2;              /* slipperiness                                                 */
  double theta_s, theta_b, hc;/* S-coord keyword                                              */
  /* lateral mixing along S-surfaces; zero-initialised = switched off */
  int uv_vis2;                /* cppdefs UV_VIS2: roms_step calls roms_visc3d after step3d_uv1 (main.F:448-450) */
  int ts_dif2;                /* cppdefs TS_DIF2: roms_step calls roms_t3dmix after step3d_t (main.F:471-473)   */
  double visc2;               /* lateral_visc keyword [m2/s] -> visc2_r, visc2_p (mixing.F:156-157)           */
  double tnu2[ROMS_MAXNT];    /* tracer_diff2 keyword [m2/s] -> diff2(:,:,itrc)  (mixing.F:161-163)           */
  /* open boundaries; zero-initialised = every non-periodic edge is a closed wall.  An open edge runs the
     combination every open-boundary case of the reference is built with (cppdefs.opt:185-225, tests/{Flux_frc,Pipes_real,CDR_*}):
     OBC_M2FLATHER (zetabc.F:28-36, u2dbc_im.F:28-56: characteristic condition on the normal barotropic velocity, upstream advection +
     binding to the boundary data on the tangential one), OBC_M3ORLANSKI (u3dbc_im.F:47-119, v3dbc_im.F), OBC_TORLANSKI (t3dbc_im.F:36-83),
     with Z_FRC_BRY, M2_FRC_BRY, M3_FRC_BRY, T_FRC_BRY boundary data (boundary.F:109-222) pushed as ROMS_*_WEST ... ROMS_*_NORTH. */
  int obc_west, obc_east, obc_south, obc_north;   /* cppdefs OBC_WEST / OBC_EAST / OBC_SOUTH / OBC_NORTH (ignored on a periodic edge) */
  double ubind;               /* ubind keyword [m/s]: binding velocity towards the boundary data on inflow (u3dbc_im.F:83-90)      */
  /* the reference's alternative open conditions, one choice per variable family for all open edges (they are compile-time switches there):
     0 = the default above; 1 = OBC_M2SPECIFIED / OBC_M3SPECIFIED / OBC_TSPECIFIED (boundary value = boundary data, u2dbc_im.F:130-140,
     u3dbc_im.F:121-131, t3dbc_im.F:85-96); 2 = none of the switches (zero-gradient, the "#else" branches).  With obc_m2 != 0 the free
     surface takes the gradient condition (zetabc.F:54-61). */
  int obc_m2, obc_m3, obc_t;
} roms_cfg;

/* Field identifiers for roms_push / roms_pull.  `lvl` selects the time level (1-based, as in
 * the Fortran: zeta(:,:,1:4), u(:,:,:,1:3), t(:,:,:,1:3,itrc)); `itrc` the tracer (1-based).
 * Unused selectors are ignored. */
enum roms_field {
  /* static 2-D, supplied by the host (grid.F:124-177) */
  ROMS_H = 0, ROMS_PM, ROMS_PN, ROMS_F, ROMS_RMASK,
  /* derived 2-D metrics (setup_grid1.F), computed by roms_setup_grid */
  ROMS_UMASK, ROMS_VMASK, ROMS_DM_U, ROMS_DN_U, ROMS_DM_V, ROMS_DN_V, ROMS_DM_R, ROMS_DN_R,
  ROMS_FOMN, ROMS_HINV,
  /* 2-D state (ocean_vars.F:67-69, coupling.F:22-44, surf_flux.F:86-94) */
  ROMS_ZETA, ROMS_UBAR, ROMS_VBAR,               /* lvl 1..4 */
  ROMS_RHOA, ROMS_RHOS, ROMS_RUFRC, ROMS_RVFRC,
  ROMS_ZT_AVG1, ROMS_DU_AVG1, ROMS_DV_AVG1, ROMS_DU_AVG2, ROMS_DV_AVG2,
  ROMS_DU_AVG_BAK, ROMS_DV_AVG_BAK, ROMS_R_D,
  ROMS_SUSTR, ROMS_SVSTR, ROMS_SWFLX,
  ROMS_STFLX,                                    /* itrc 1..NT */
  /* 3-D state */
  ROMS_U, ROMS_V,                                /* lvl 1..3, k=1:N */
  ROMS_T,                                        /* lvl 1..3, itrc 1..NT, k=1:N */
  ROMS_HZ, ROMS_Z_R,                             /* k=1:N */
  ROMS_Z_W,                                      /* k=0:N */
  ROMS_FLXU, ROMS_FLXV,                          /* k=1:N */
  ROMS_WE, ROMS_WI,                              /* k=0:N */
  ROMS_RHO1, ROMS_QP1,                           /* k=1:N (linear EOS: RHO1 holds rho) */
  ROMS_AKV,                                      /* k=0:N */
  ROMS_AKT,                                      /* itrc 1..2 (iTandS), k=0:N */
  ROMS_RU, ROMS_RV,                              /* k=1:N; prsgrd -> pre_step3d/step3d_uv1 hand-off
                                                    (the reference's A3d(1,1:2), prsgrd.F:12) */
  /* lateral mixing (row f1): psi-point metrics and mask computed by roms_setup_grid (setup_grid1.F:107-190);
     coefficients filled from cfg.visc2 / cfg.tnu2 by roms_setup_grid (mixing.F:156-163) -- a host that adds a
     sponge (set_nudgcof.F:88-110) pushes its own afterwards */
  ROMS_PMASK, ROMS_DM_P, ROMS_DN_P, ROMS_PMON_U, ROMS_PNOM_V,
  ROMS_VISC2_R, ROMS_VISC2_P,
  ROMS_DIFF2,                                    /* itrc 1..NT */
  ROMS_RIV_UFLX, ROMS_RIV_VFLX,                  /* river_frc.F: river faces, 10*iriver + fraction of the river's volume flux */
  ROMS_PTIDE,                                    /* tides.F:26 equilibrium tidal potential [m] (pot_tides) */
  ROMS_PIPE_IDX, ROMS_PIPE_FLX,                  /* pipe_frc.F: pipe index of a cell (0: none) and its volume flux fraction*pipe_vol [m3/s] */
  ROMS_SRFLX,                                    /* surf_flux.F srflx: shortwave radiation flux [degC m/s], part of stflx(itemp); with KPP on it
                                                    penetrates by swr_frac (step3d_t_ISO.F:961-998, lmd_kpp.F:160-167); zero unless pushed */
  /* boundary data of the open edges (boundary.F:21-60 zeta_west(j), ubar_west(j), vbar_west(j), u_west(j,k), v_west(j,k),
     t_west(j,k,itrc) and their east/south/north counterparts): lines along the edge with the reference's own extents
     (boundary.F:111-131: zeta_west(0:Mm+1), u_west(0:Mm+1,N), t_south(0:Lm+1,N,NT) ...) -- Mm+2 values (west, east) or Lm+2 values
     (south, north), times N levels for the 3-D ones (level-major, i.e. the Fortran (j,k) order).  Id = base + side,
     side 0 west, 1 east, 2 south, 3 north.  Zero unless pushed. */
  ROMS_ZETA_WEST, ROMS_ZETA_EAST, ROMS_ZETA_SOUTH, ROMS_ZETA_NORTH,
  ROMS_UBAR_WEST, ROMS_UBAR_EAST, ROMS_UBAR_SOUTH, ROMS_UBAR_NORTH,
  ROMS_VBAR_WEST, ROMS_VBAR_EAST, ROMS_VBAR_SOUTH, ROMS_VBAR_NORTH,
  ROMS_U_WEST, ROMS_U_EAST, ROMS_U_SOUTH, ROMS_U_NORTH,             /* k=1:N */
  ROMS_V_WEST, ROMS_V_EAST, ROMS_V_SOUTH, ROMS_V_NORTH,             /* k=1:N */
  ROMS_T_WEST, ROMS_T_EAST, ROMS_T_SOUTH, ROMS_T_NORTH,             /* itrc 1..NT, k=1:N */
  ROMS_NFIELDS
};

/* ---- lifetime, data movement ------------------------------------------------------------ */
int  roms_create(const roms_cfg* cfg, int device, roms_ctx** out);
int  roms_destroy(roms_ctx* h);
int  roms_push(roms_ctx* h, int field, int lvl, int itrc, const double* host);   /* H2D */
int  roms_pull(roms_ctx* h, int field, int lvl, int itrc, double* host);         /* D2H, blocking */
/* The same transfers enqueued on the library's stream without waiting (pinned host memory makes them truly asynchronous): the host
 * buffer of a push must stay untouched, and that of a pull is valid, only after the next roms_sync.  For a driver that exchanges
 * forcing and output planes every step (set_forces / wrt_his, main.F:386, :324): one roms_sync per step instead of one per plane. */
int  roms_push_async(roms_ctx* h, int field, int lvl, int itrc, const double* host);
int  roms_pull_async(roms_ctx* h, int field, int lvl, int itrc, double* host);
long roms_field_count(roms_ctx* h, int field);  /* doubles per (lvl,itrc) slab */
int  roms_sync(roms_ctx* h);
const char* roms_last_error(void);

/* ---- host-side constants (set_scoord.F:4-66, set_weights.F:7-246, setup_grid1.F:13-210) -- */
int  roms_setup_grid(roms_ctx* h);      /* after H,PM,PN,F,RMASK are pushed                  */
int  roms_get_nfast(roms_ctx* h);       /* executed barotropic substeps (82 for ndtfast=60)   */
int  roms_get_weights(roms_ctx* h, double* w1, double* w2, int n); /* weight(1:2,1:n)         */
int  roms_get_scoord(roms_ctx* h, double* Cs_w, double* Cs_r);     /* Cs_w(0:N), Cs_r(1:N)    */

/* ---- loop-control scalars (scalars.F:28-32): the driver keeps owning the rotation -------- */
int  roms_set_time_indices(roms_ctx* h, int iic, int iif, int nstp, int nrhs, int nnew,
                           int kstp, int knew);
int  roms_get_time_indices(roms_ctx* h, int* iic, int* iif, int* nstp, int* nrhs, int* nnew,
                           int* kstp, int* knew);
int  roms_compute_tile_bounds(roms_ctx* h, int tile, int* istr, int* iend, int* jstr, int* jend);
                                        /* compute_tile_bounds.h:45-79 with NSUB_X=NSUB_E=1  */

/* ---- the tile-subroutine surface (one export per reference subroutine) ------------------- */
int  roms_set_depth(roms_ctx* h, int tile);   /* set_depth.F:4     */
int  roms_set_huv(roms_ctx* h);               /* set_depth.F:190   */
int  roms_set_huv1(roms_ctx* h, int tile);    /* set_depth.F:239   */
int  roms_omega(roms_ctx* h);                 /* omega.F:4         */
int  roms_rho_eos(roms_ctx* h, int tidx);     /* rho_eos.F:6       */
int  roms_prsgrd(roms_ctx* h);                /* prsgrd.F:4        */
int  roms_pre_step3d(roms_ctx* h, int tile);  /* pre_step3d4S.F:4  */
int  roms_step3d_uv1(roms_ctx* h, int tile);  /* step3d_uv1.F:5    */
int  roms_step2d(roms_ctx* h);                /* step2d_FB.F:3     */
int  roms_step3d_uv2(roms_ctx* h, int tile);  /* step3d_uv2.F:6    */
int  roms_step3d_t(roms_ctx* h, int tile);    /* step3d_t_ISO.F:20 */
int  roms_visc3d(roms_ctx* h);                /* visc3d_S.F:4  (UV_VIS2: harmonic viscosity along S-surfaces)  */
int  roms_t3dmix(roms_ctx* h);                /* t3dmix_S.F:4  (TS_DIF2: Laplacian tracer diffusion, all NT)   */
int  roms_alfabeta(roms_ctx* h, double* alpha_host, double* beta_host); /* alfabeta.F:4 */
int  roms_set_kpp(roms_ctx* h, int on);        /* cppdefs LMD_MIXING LMD_KPP LMD_BKPP LMD_RIMIX LMD_NONLOCAL: allocate the closure's arrays;
                                                  roms_step then calls roms_lmd_vmix where main.F:408-410, 434-436 do; rho_eos also fills bvf */
int  roms_set_kpp_opts(roms_ctx* h, int lmd_convec, double Akv_bak, double Akt_bak_T, double Akt_bak_S);
                                               /* cppdefs LMD_CONVEC (lmd_vmix.F:269-274; default on, as in Rivers_ana/Pipes_ana) and the
                                                  vertical_mixing keyword's background coefficients Akv_bak, Akt_bak(1:2) (default 0) */
int  roms_lmd_vmix(roms_ctx* h, int tind);     /* lmd_vmix.F:4 lmd_vmix(tind): Akv, Akt, hbls, hbbl, ghat from u,v(tind), bvf, surface fluxes */
int  roms_get_hbl(roms_ctx* h, double* hbls_host, double* hbbl_host); /* boundary layer thicknesses, (Lm+4)x(Mm+4) planes */
int  roms_set_pot_tides(roms_ctx* h, int on);  /* tides.F pot_tides: prsgrd subtracts g*ptide (ROMS_PTIDE, pushed by the host each
                                                  step after its harmonic synthesis, tides.F:238-246) from the surface pressure, prsgrd.F:209-211 */
int  roms_set_pipe(roms_ctx* h, int ipipe, const double* prf, const double* trc);
                                               /* pipe_frc.F: release profile pipe_prf(1:N) and the NT tracer values of pipe ipipe (1..8);
                                                  switches pipe_source on.  Pipe cells and their fluxes: ROMS_PIPE_IDX / ROMS_PIPE_FLX.
                                                  Hooks: step2d_FB.F:155-159, omega.F:102-108, step3d_t_ISO.F:927-935 */
int  roms_set_river(roms_ctx* h, int iriver, double vol, const double* trc);
                                               /* river_frc.F set_river_frc: volume flux [m3/s] and the NT tracer values of river
                                                  iriver (1..8) for the coming steps; switches river_source on.  The river faces
                                                  are the non-zero entries of ROMS_RIV_UFLX / ROMS_RIV_VFLX (init_river_frc).
                                                  Hooks: step2d_FB.F:531-554, pre_step3d4S.F:493-522, step3d_uv2.F:689-717,
                                                  compute_horiz_tracer_fluxes.h:217-246 */
int  roms_exchange(roms_ctx* h, int nfields, const int* fields, const int* lvls, const int* itrcs);
                                              /* mpi_exchanges.F:34-40 exchange_xxx */
int  roms_grid_stats(roms_ctx* h, double out[9]); /* setup_grid2.F:64-100 start-up report of this subdomain: hmin, hmax, grdmin,
                                                  grdmax, Cg_min, Cg_max, Cu_Cor, area, volume                                */
int  roms_diag(roms_ctx* h, double out[4]);   /* diag.F:50: KE, barotropic KE, max adv CFL, its vertical part */

/* ---- allocation-free variant (NO counterpart in the reference) ----
 * For callers that cannot call cudaMalloc (CUDA-graph capture, tight serving loops, harnesses that forbid it): the
 * omega.F:17-164 column kernel of roms_omega without a context.  Its column scratch is a file-scope static device array
 * of 4 Mi doubles (32 MiB, held by every GPU that loads the library until exit): 2*(N+1)*Lm*Mm must not exceed it, else
 * the call returns -3 with a message.  Calls no cudaMalloc/cudaFree.  All pointers are DEVICE pointers in the reference's
 * unpadded layout (-1:Lm+2,-1:Mm+2[,k]): FlxU,FlxV,Hz k=1:N; z_w,We,Wi k=0:N; pm,pn,swflx 2-D.  Writes We,Wi on the
 * interior (1:Lm,1:Mm), k=0:N; no boundary copies, no halo exchange.  Stream-ordered on `stream` (a cudaStream_t; NULL =
 * default stream).  All calls share the one scratch array: calls that overlap on the GPU race (serialise them on one stream). */
int  roms_omega_noalloc(int Lm, int Mm, int N, double dtau, const double* FlxU, const double* FlxV, const double* z_w, const double* Hz,
                        const double* pm, const double* pn, const double* swflx, double* We, double* Wi, void* stream);

/* ---- the driver loop body, for hosts without a Fortran main (main.F:86-321, 333-520) ---- */
int  roms_init_step0(roms_ctx* h);            /* set_depth, set_HUV, omega, rho_eos of roms_init */
int  roms_step(roms_ctx* h);                  /* one baroclinic step: main.F:397-479            */
int  roms_run(roms_ctx* h, int nsteps);       /* nsteps x roms_step                             */
int  roms_kernel_launches(roms_ctx* h, long* n); /* kernels launched so far by this context    */
int  roms_resident_bytes(roms_ctx* h, long long* n); /* HBM bytes this context keeps resident */

/* ---- multi-GPU: 2-D subdomain split, one process per GPU (replaces mpi_setup.F / mpi_exchanges.F) ---- */
int  roms_comm_unique_id(char* out, int len);  /* rank 0: NCCL unique id (>=128 bytes) to broadcast   */
int  roms_comm_attach(roms_ctx* h, const char* id, int len, int rank, int world);
                                               /* rank = jnode*np_xi+inode (mpi_setup.F:61-70)        */
int  roms_decomp_neighbor(int np_xi, int np_eta, int inode, int jnode, int ew_periodic, int ns_periodic, int d);
                                               /* neighbour rank in direction d=0..7 (W,E,S,N,SW,SE,NW,NE), -1 at a wall */
int  roms_decomp_box(int Lm, int Mm, int d, int send, const int walls[4], int out[4]);
                                               /* i0,j0,ni,nj of the message to/from direction d; walls = {S,N,W,E} closed physical
                                                  edges of this subdomain (NULL: none): messages along a wall carry the boundary row/column */
int  roms_decomp_opposite(int d);
int  roms_decomp_extent(int LL, int np, int node, int* L, int* corn);
                                               /* mpi_setup.F:117-160: interior points L of block `node` of `np` along a direction of
                                                  LL global points (surplus trimmed from the first/last block) and its global offset
                                                  (iSW_corn / jSW_corn); non-zero if the block is narrower than the halo */

/* ---- measurement hooks (the reference's timers.F:35-131 analogue; device-side CUDA events) - */
int  roms_timer_start(roms_ctx* h);           /* event on the library's stream                  */
int  roms_timer_stop(roms_ctx* h, double* ms);/* second event + elapsed milliseconds            */
int  roms_kernel_timing(roms_ctx* h, int on); /* bracket every launch with its own event pair   */
int  roms_kernel_times(roms_ctx* h, char* buf, int buflen); /* "kernel count total_ms" lines    */

#ifdef __cplusplus
}
#endif
#endif
