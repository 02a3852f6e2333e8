"""Analytic (synthetic, deterministic) cases for the explicit time step: host-side numpy only.

Every case returns (cfg, state) where `state` maps field names of include/roms_b200.h to arrays in
the reference layout (j, i) / (k, j, i) incl. the 2-cell halo, ready for `push_state`.  The same
state is pushed into libromsb200 and, in the tests, into the CPU oracle.

  channel(...)   BASELINE.json configs[1]: analytic baroclinic channel, T+S, nonlinear split EOS
                 (SURVEY.md section 8(d) input (2)).  `ns_walls=False` gives the doubly periodic variant.
  filament(...)  configs[0]: tests/Filament/{ana_grid.h,ana_init.h,param.opt,benchmark.in}
"""
import numpy as np
from .capi import RomsCfg


def _scoord(N, theta_s, theta_b):
    """set_scoord.F:4-66 (host copy used only to place the analytic T,S profiles)."""
    def csf(sc):
        csrf = (1.0 - np.cosh(theta_s * sc)) / (np.cosh(theta_s) - 1.0) if theta_s > 0 else -sc ** 2
        return (np.exp(theta_b * csrf) - 1.0) / (1.0 - np.exp(-theta_b)) if theta_b > 0 else csrf
    ds = 1.0 / N
    k = np.arange(1, N + 1)
    sc_r = ds * (k - N - 0.5)
    sc_w = ds * (np.arange(0, N + 1) - N)
    cs_w = csf(sc_w); cs_w[0] = -1.0; cs_w[-1] = 0.0
    return sc_r, csf(sc_r), sc_w, cs_w


def _depths(h, zeta, N, theta_s, theta_b, hc):
    sc_r, cs_r, sc_w, cs_w = _scoord(N, theta_s, theta_b)
    hinv = 1.0 / (h + hc)
    z_r = zeta[None] + (zeta + h)[None] * (hc * sc_r[:, None, None] + cs_r[:, None, None] * h[None]) * hinv[None]
    z_w = zeta[None] + (zeta + h)[None] * (hc * sc_w[:, None, None] + cs_w[:, None, None] * h[None]) * hinv[None]
    z_w[0] = -h
    return z_r, z_w


def _wrap(a, ew, ns):
    """periodic halo fill of the last two axes (j, i), 2 cells"""
    if ew:
        a[..., :, :2] = a[..., :, -4:-2]
        a[..., :, -2:] = a[..., :, 2:4]
    if ns:
        a[..., :2, :] = a[..., -4:-2, :]
        a[..., -2:, :] = a[..., 2:4, :]
    return a


def rank_grid(world):
    """NP_XI x NP_ETA for `world` GPUs of one box (1x1, 2x1, 2x2, 4x2)."""
    return {1: (1, 1), 2: (2, 1), 4: (2, 2), 8: (4, 2)}.get(world, (world, 1))


def block_extent(LL, np_, node):
    """mpi_setup.F:117-160 (mirror of roms_decomp_extent): interior points of block `node` of `np_` along a direction of LL global
    points and its global offset.  Blocks get ceil(LL/np_) points; the surplus is trimmed from the first and the last block."""
    L0 = (LL + np_ - 1) // np_
    off = np_ * L0 - LL
    L = L0 - (off // 2 if node == 0 else 0) - ((off + 1) // 2 if node == np_ - 1 else 0)
    return L, (0 if node == 0 else node * L0 - off // 2)


def river_faces(rmask, ridx, rfrc):
    """river_frc.F:228-277 calc_river_flux: a river cell (a masked cell with index ridx and fraction rfrc of its river's
    volume flux) feeds every unmasked neighbour through the face between them; faces are coded 10*iriver +- fraction/faces,
    the sign being the direction of the flow along the axis."""
    ru = np.zeros_like(rmask); rv = np.zeros_like(rmask)
    for j, i in zip(*np.nonzero(rfrc[1:-1, 1:-1] > 0)):
        j += 1; i += 1
        faces = rmask[j, i - 1] + rmask[j, i + 1] + rmask[j - 1, i] + rmask[j + 1, i]
        assert faces > 0 and rmask[j, i] == 0, "river grid position error"
        if rmask[j, i - 1] > 0:
            ru[j, i] = -rfrc[j, i] / faces + 10 * ridx[j, i]
        if rmask[j, i + 1] > 0:
            ru[j, i + 1] = rfrc[j, i] / faces + 10 * ridx[j, i]
        if rmask[j - 1, i] > 0:
            rv[j, i] = -rfrc[j, i] / faces + 10 * ridx[j, i]
        if rmask[j + 1, i] > 0:
            rv[j + 1, i] = rfrc[j, i] / faces + 10 * ridx[j, i]
    return ru, rv


def _noise(ig, jg, k, salt):
    """deterministic pseudo-random field in [-1,1) as a function of GLOBAL indices (so that every rank of a
    decomposed run generates exactly the values a single-domain run has)"""
    x = (ig.astype(np.uint64) * np.uint64(73856093)) ^ (jg.astype(np.uint64) * np.uint64(19349663)) ^ \
        (k.astype(np.uint64) * np.uint64(83492791)) ^ np.uint64(salt * 2654435761 % (1 << 32))
    x = (x ^ (x >> np.uint64(15))) * np.uint64(2246822519) % np.uint64(1 << 32)
    x = (x ^ (x >> np.uint64(13))) * np.uint64(3266489917) % np.uint64(1 << 32)
    x = x ^ (x >> np.uint64(16))
    return x.astype(np.float64) / float(1 << 31) - 1.0


def channel(Lm=64, Mm=64, N=16, NT=2, ns_walls=False, nonlin_eos=True, dt=120.0, ndtfast=60, dx=1.0e3, dy=1.0e3,
            depth=4000.0, seed=20260101, perturb=1e-6, amp=None, sub=(0, 0, 1, 1), seamount=False, island=False, ew_walls=False,
            visc2=None, tnu2=None, sponge=False, ishift=0, river=False, pipe=False, tide=False, open_edges=""):
    """Lm, Mm are the GLOBAL interior sizes; sub=(inode, jnode, np_xi, np_eta) selects this rank's block
    (mpi_setup.F:61-70; uneven splits as mpi_setup.F:117-160, see block_extent).  Returns the cfg and state of that block.
    seamount: Gaussian seamount h = depth*(1 - 0.5*exp(-r^2/(L/5)^2)) (SURVEY.md 8(d) input (3), form of
    NHMG/tests/mg_setup_tests.f90:150); island: a circular land mask (rmask=0) north-east of it, so that the
    rmask/umask/vmask factors of every routine are exercised.  ew_walls: closed western/eastern walls as well (with
    ns_walls: a fully closed basin, BASELINE configs[2] geometry).
    visc2 / tnu2 [m2/s]: switch on UV_VIS2 / TS_DIF2 lateral mixing (visc3d_S.F, t3dmix_S.F) with these background
    coefficients; sponge: add a smooth spatially varying part to visc2_r, visc2_p, diff2 (what set_nudgcof.F:88-110 does
    near open boundaries) so that the coefficient averaging of the flux formulas is exercised.
    river: the southern tenth of the domain becomes land cut by a river channel over the middle fifth in xi, the southernmost
    2 % is all land (the geometry of tests/Rivers_ana/ana_grid.h); the land cells at the head of the channel are the river cells of
    river 1 (river_frc.F:125-137), its faces are coded by `river_faces` and its volume flux and tracer values are those of
    tests/Rivers_ana/ana_frc_river.h scaled to the cell size, with a milder T,S contrast (state key "RIVERS").
    pipe: two neighbouring water cells share the outfall of pipe 1 (pipe_frc.F: index + fraction, pipe_flx = fraction*pipe_vol),
    released over the two bottom levels (the profile of pipe_frc.F:232-234) with its own tracer values (state key "PIPES").
    tide: an equilibrium tidal potential ptide (tides.F) of 0.3 m amplitude, one wave across the domain (state key "PTIDE";
    `initialize` switches pot_tides on).
    open_edges: letters out of "WESN" -- those (non-periodic) edges are open boundaries (cfg.obc_*, ubind = 0.1 m/s) instead of walls; the
    boundary data (boundary.F: zeta_west ... t_north) are the initial fields on the boundary points, i.e. the exterior is clamped to
    the initial state (state keys "ZETA_WEST" ... "T_NORTH"), and the initial flow through an open edge is kept.
    ishift: translate the whole E-W periodic problem by this many cells in xi (the state is the roll of the unshifted one, so
    the solution must be the roll of the unshifted solution, bit for bit -- a size-independent parity property)."""
    inode, jnode, np_xi, np_eta = sub
    LmG, MmG = Lm, Mm
    (Lm, i0), (Mm, j0) = block_extent(LmG, np_xi, inode), block_extent(MmG, np_eta, jnode)
    cfg = RomsCfg(Lm=Lm, Mm=Mm, N=N, NT=NT, ew_periodic=0 if ew_walls else 1, ns_periodic=0 if ns_walls else 1,
                  nonlin_eos=1 if nonlin_eos else 0, salinity=1, ndtfast=ndtfast, np_xi=np_xi, np_eta=np_eta, inode=inode, jnode=jnode,
                  dt=dt, rho0=1027.0, g=9.81, Tcoef=0.2, T0=10.0, Scoef=0.8, S0=35.0,
                  rdrg=3.0e-4, rdrg2=0.0, Zob=0.0, gamma2=1.0, theta_s=6.0, theta_b=2.0, hc=25.0)
    if visc2 is not None:
        cfg.uv_vis2, cfg.visc2 = 1, visc2
    if tnu2 is not None:
        cfg.ts_dif2 = 1
        for it in range(NT):
            cfg.tnu2[it] = tnu2 * (1.0 + 0.25 * it)
    # work on a 3-cell halo so that u/v-point averages exist on the whole 2-cell halo, trim at the end
    sy, sx = Mm + 6, Lm + 6
    ig = (np.arange(-2, Lm + 4) + i0)[None, :] + np.zeros((sy, 1), dtype=np.int64)
    jg = (np.arange(-2, Mm + 4) + j0)[:, None] + np.zeros((1, sx), dtype=np.int64)
    if not ew_walls:
        ig = (ig - 1 + ishift) % LmG + 1         # E-W periodic images
    if not ns_walls:
        jg = (jg - 1) % MmG + 1
    x = dx * (ig - 0.5); y = dy * (jg - 0.5)
    Lx, Ly = dx * LmG, dy * MmG
    if amp is None:
        amp = 0.02 * min(Lx, Ly) / 64.0e3 + 0.01   # keeps the thermal-wind shear independent of the grid size
    st = {}
    h = np.full((sy, sx), depth)
    if seamount:
        r2 = (x - 0.5 * Lx) ** 2 + (y - 0.5 * Ly) ** 2
        h = depth * (1.0 - 0.5 * np.exp(-r2 / (min(Lx, Ly) / 5.0) ** 2))
    rmask = np.ones((sy, sx))
    if island:
        rmask[(x - 0.72 * Lx) ** 2 + (y - 0.70 * Ly) ** 2 < (0.09 * min(Lx, Ly)) ** 2] = 0.0
    if river:
        rmask[(y < 0.1 * Ly) & ((x < 0.4 * Lx) | (x > 0.6 * Lx))] = 0.0
        rmask[y < 0.02 * Ly + 0.75 * dy] = 0.0
    zeta = 0.01 * np.sin(2 * np.pi * x / Lx) * np.cos(2 * np.pi * y / Ly)
    z_r, z_w = _depths(h, zeta, N, cfg.theta_s, cfg.theta_b, cfg.hc)
    Hz = z_w[1:] - z_w[:-1]
    if ns_walls:
        front = amp * np.tanh((y - 0.5 * Ly) / (0.15 * Ly))
    else:
        front = amp * np.sin(2 * np.pi * y / Ly)
    kk = np.arange(1, N + 1)[:, None, None]
    nz = lambda salt: _noise(ig[None] + 0 * kk, jg[None] + 0 * kk, kk + 0 * ig[None], salt + seed % 1000)
    T = 4.0 + 10.0 * np.exp(z_r / 50.0) + (front + 0.2 * amp * np.sin(2 * np.pi * x / Lx) * np.cos(2 * np.pi * y / Ly))[None] * np.exp(z_r / 400.0)
    S = 35.0 + 0.1 * amp * np.cos(2 * np.pi * x / Lx)[None] * np.exp(z_r / 300.0) - 0.5 * np.exp(z_r / 100.0)
    u = 0.02 * (np.sin(2 * np.pi * y / Ly) * (1 + 0.3 * np.cos(2 * np.pi * x / Lx)))[None] * np.exp(z_r / 500.0)
    v = 0.01 * (np.sin(2 * np.pi * x / Lx) * np.cos(2 * np.pi * y / Ly))[None] * np.exp(z_r / 500.0)
    T = T * (1 + perturb * nz(1)); S = S * (1 + perturb * nz(2))
    u = u + perturb * 0.1 * nz(3); v = v + perturb * 0.1 * nz(4)
    if open_edges:
        assert np_xi == 1 and np_eta == 1, "open-boundary cases are single-domain"
        assert all(e in "WESN" for e in open_edges) and (ew_walls or not set(open_edges) & set("WE")) and (ns_walls or not set(open_edges) & set("SN"))
        cfg.obc_west, cfg.obc_east, cfg.obc_south, cfg.obc_north = ("W" in open_edges), ("E" in open_edges), ("S" in open_edges), ("N" in open_edges)
        cfg.ubind = 0.1
    if ns_walls:
        jl = (np.arange(-2, Mm + 4) + j0)
        if "S" not in open_edges:
            v[:, jl <= 1, :] = 0.0                     # closed south wall (j=1)
        if "N" not in open_edges:
            v[:, jl >= MmG + 1, :] = 0.0               # closed north wall (j=Mm+1)
    if ew_walls:
        il = (np.arange(-2, Lm + 4) + i0)
        if "W" not in open_edges:
            u[:, :, il <= 1] = 0.0                     # closed west wall (i=1)
        if "E" not in open_edges:
            u[:, :, il >= LmG + 1] = 0.0               # closed east wall (i=Lm+1)
    if island or river:
        um = np.ones_like(rmask); vm = np.ones_like(rmask)
        um[:, 1:] = rmask[:, 1:] * rmask[:, :-1]; vm[1:, :] = rmask[1:, :] * rmask[:-1, :]
        u = u * um[None]; v = v * vm[None]; zeta = zeta * rmask
    Hzu = np.empty_like(Hz); Hzv = np.empty_like(Hz)
    Hzu[:, :, 1:] = 0.5 * (Hz[:, :, 1:] + Hz[:, :, :-1]); Hzu[:, :, 0] = Hz[:, :, 0]
    Hzv[:, 1:, :] = 0.5 * (Hz[:, 1:, :] + Hz[:, :-1, :]); Hzv[:, 0, :] = Hz[:, 0, :]
    ubar = (Hzu * u).sum(0) / Hzu.sum(0); vbar = (Hzv * v).sum(0) / Hzv.sum(0)
    t2 = lambda a: np.ascontiguousarray(a[..., 1:-1, 1:-1])
    st["H"] = t2(h)
    st["PM"] = np.full((Mm + 4, Lm + 4), 1.0 / dx); st["PN"] = np.full((Mm + 4, Lm + 4), 1.0 / dy)
    st["F"] = np.full((Mm + 4, Lm + 4), 1.0e-4)
    st["RMASK"] = t2(rmask)
    zeta, ubar, vbar, u, v, T, S = (t2(a) for a in (zeta, ubar, vbar, u, v, T, S))
    st["ZETA"] = {(1, 1): zeta, (2, 1): zeta}
    st["UBAR"] = {(1, 1): ubar, (2, 1): ubar}
    st["VBAR"] = {(1, 1): vbar, (2, 1): vbar}
    st["U"] = {(1, 1): u, (2, 1): u}
    st["V"] = {(1, 1): v, (2, 1): v}
    tr = {(1, 1): T, (2, 1): T, (1, 2): S, (2, 2): S}
    for it in range(3, NT + 1):   # passive tracers (SURVEY.md 8(d) input (4))
        pz = t2(np.sin(2 * np.pi * (it - 2) * x / Lx)[None] * np.cos(2 * np.pi * y / Ly)[None] * np.exp(z_r / 100.0))
        tr[(1, it)] = pz; tr[(2, it)] = pz
    st["T"] = tr
    akv = np.full((N + 1, Mm + 4, Lm + 4), 1.0e-4)
    st["AKV"] = akv
    st["AKT"] = {(1, 1): akv, (1, 2): akv}
    if open_edges:
        # python index = Fortran index + 1: rho/v columns istr-1 -> 1, iend+1 -> Lm+2; the u-face ON the western edge is istr -> 2
        line = {"W": lambda a, face: a[..., 2 if face else 1], "E": lambda a, face: a[..., Lm + 2],
                "S": lambda a, face: a[..., 2 if face else 1, :], "N": lambda a, face: a[..., Mm + 2, :]}
        name = {"W": "WEST", "E": "EAST", "S": "SOUTH", "N": "NORTH"}
        for e in open_edges:
            cut = lambda a, face=False: np.ascontiguousarray(line[e](a, face)[..., 1:-1])      # (0:Mm+1) / (0:Lm+1)
            st["ZETA_" + name[e]] = cut(zeta)
            st["UBAR_" + name[e]] = cut(ubar, e == "W"); st["VBAR_" + name[e]] = cut(vbar, e == "S")
            st["U_" + name[e]] = cut(u, e == "W"); st["V_" + name[e]] = cut(v, e == "S")
            st["T_" + name[e]] = {(1, it): cut(tr[(1, it)]) for it in range(1, NT + 1)}
    if river:
        ridx = np.zeros_like(rmask); rfrc = np.zeros_like(rmask)
        head = np.zeros_like(rmask, dtype=bool)
        head[:-1, :] = (rmask[:-1, :] == 0) & (rmask[1:, :] == 1) & (x[:-1, :] > 0.4 * Lx) & (x[:-1, :] < 0.6 * Lx)
        ncell = int(round(0.2 * LmG))
        ridx[head] = 1.0; rfrc[head] = 1.0 / ncell
        ru, rv = river_faces(rmask, ridx, rfrc)
        st["RIV_UFLX"] = t2(ru); st["RIV_VFLX"] = t2(rv)
        st["RIVERS"] = [(5.0e2 * dx * dy / 1.0e4, [15.0, 34.5] + [0.5] * (NT - 2))]   # Rivers_ana volume flux per unit area; mild T,S contrast (deep basin)
    if pipe:
        pidx = np.zeros_like(rmask); pflx = np.zeros_like(rmask)
        ip, jp = int(0.3 * LmG) + 1, int(0.6 * MmG) + 1
        sel = ((ig == ip) | (ig == ip + 1)) & (jg == jp)
        pidx[sel] = 1.0; pflx[sel] = 0.5 * 2.0e2 * dx * dy / 1.0e4
        st["PIPE_IDX"] = t2(pidx); st["PIPE_FLX"] = t2(pflx)
        st["PIPES"] = [([0.5, 0.5] + [0.0] * (N - 2), [4.5, 34.95] + [1.0] * (NT - 2))]   # close to the ambient bottom water: a gentle plume
    if tide:
        st["PTIDE"] = t2(0.3 * np.cos(2 * np.pi * x / Lx) * np.cos(np.pi * y / Ly))
    if sponge:
        wr = 0.5 + 0.5 * np.cos(2 * np.pi * x / Lx) * np.sin(2 * np.pi * y / Ly)          # at rho points, 3-cell halo
        wp = np.zeros_like(wr)
        wp[1:, 1:] = 0.25 * (wr[1:, 1:] + wr[1:, :-1] + wr[:-1, 1:] + wr[:-1, :-1])      # psi points (set_nudgcof.F:97-98)
        if visc2 is not None:
            st["VISC2_R"] = t2(visc2 * (1.0 + wr)); st["VISC2_P"] = t2(visc2 * (1.0 + wp))
        if tnu2 is not None:
            st["DIFF2"] = {(1, it + 1): t2(cfg.tnu2[it] * (1.0 + wr)) for it in range(NT)}
    return cfg, st


def pulse_box(Lm=64, Mm=64, N=4, open_edges="WESN", dx=1.0e3, depth=100.0, dt=60.0, ndtfast=30):
    """Radiation test of the open boundaries: a homogeneous box at rest (f = 0, no drag, no mixing) with a Gaussian free-surface
    bump in its centre.  The gravity-wave ring it sheds has to leave through the open edges (boundary data: zero = a resting
    exterior); behind closed walls the same energy stays in the box.  Returns cfg, state."""
    cfg = RomsCfg(Lm=Lm, Mm=Mm, N=N, NT=2, ew_periodic=0, ns_periodic=0, nonlin_eos=0, salinity=1, ndtfast=ndtfast, np_xi=1, np_eta=1, inode=0, jnode=0,
                  dt=dt, rho0=1027.0, g=9.81, Tcoef=0.0, T0=10.0, Scoef=0.0, S0=35.0, rdrg=0.0, rdrg2=0.0, Zob=0.0, gamma2=1.0,
                  theta_s=3.0, theta_b=0.0, hc=25.0)
    cfg.obc_west, cfg.obc_east, cfg.obc_south, cfg.obc_north = ("W" in open_edges), ("E" in open_edges), ("S" in open_edges), ("N" in open_edges)
    cfg.ubind = 0.1
    sy, sx = Mm + 4, Lm + 4
    x = dx * (np.arange(-1, Lm + 3)[None, :] - 0.5) + np.zeros((sy, 1))
    y = dx * (np.arange(-1, Mm + 3)[:, None] - 0.5) + np.zeros((1, sx))
    zeta = 0.1 * np.exp(-((x - 0.5 * dx * Lm) ** 2 + (y - 0.5 * dx * Mm) ** 2) / (5.0 * dx) ** 2)
    zero2, zero3 = np.zeros((sy, sx)), np.zeros((N, sy, sx))
    st = {"H": np.full((sy, sx), depth), "PM": np.full((sy, sx), 1.0 / dx), "PN": np.full((sy, sx), 1.0 / dx), "F": zero2.copy(), "RMASK": np.ones((sy, sx)),
          "ZETA": {(1, 1): zeta, (2, 1): zeta}, "UBAR": {(1, 1): zero2, (2, 1): zero2}, "VBAR": {(1, 1): zero2, (2, 1): zero2},
          "U": {(1, 1): zero3, (2, 1): zero3}, "V": {(1, 1): zero3, (2, 1): zero3},
          "T": {(1, 1): zero3 + 10.0, (2, 1): zero3 + 10.0, (1, 2): zero3 + 35.0, (2, 2): zero3 + 35.0},
          "AKV": np.zeros((N + 1, sy, sx)), "AKT": {(1, 1): np.zeros((N + 1, sy, sx)), (1, 2): np.zeros((N + 1, sy, sx))}}
    for e, nm in (("W", "WEST"), ("E", "EAST"), ("S", "SOUTH"), ("N", "NORTH")):
        if e in open_edges:
            L = (sy if e in "WE" else sx) - 2
            st["T_" + nm] = {(1, 1): np.full((N, L), 10.0), (1, 2): np.full((N, L), 35.0)}     # the other boundary data stay zero
    return cfg, st


def strat_box(open_edges="WESN", Lm=32, Mm=32, N=8, dx=500.0, depth=100.0):
    """Baroclinic radiation test of the open boundaries: `pulse_box` with a linear stratification (10 K over the depth, linear EOS) and a
    mode-1 Gaussian displacement of the isotherms (5 m) in the centre instead of the free-surface bump; the internal-wave ring it sheds
    (c1 = 0.44 m/s) has to leave through the 3-D conditions (u3dbc/v3dbc radiation, t3dbc upstream advection towards the resting profile).
    Returns cfg, state and the background profile Tb(k,j,i)."""
    cfg, st = pulse_box(Lm=Lm, Mm=Mm, N=N, open_edges=open_edges, dx=dx, depth=depth, dt=60.0, ndtfast=30)
    cfg.Tcoef = 0.2
    sy, sx = Mm + 4, Lm + 4
    z_r, _ = _depths(st["H"], np.zeros((sy, sx)), N, cfg.theta_s, cfg.theta_b, cfg.hc)
    x = dx * (np.arange(-1, Lm + 3)[None, :] - 0.5) + np.zeros((sy, 1))
    y = dx * (np.arange(-1, Mm + 3)[:, None] - 0.5) + np.zeros((1, sx))
    r2 = (x - 0.5 * dx * Lm) ** 2 + (y - 0.5 * dx * Mm) ** 2
    Tb = 10.0 + 10.0 * (z_r + depth) / depth
    disp = 5.0 * np.exp(-r2 / (3.0 * dx) ** 2)[None] * np.sin(np.pi * (z_r + depth) / depth)
    T = Tb - (10.0 / depth) * disp
    S = st["T"][(1, 2)]
    st["ZETA"] = {(1, 1): np.zeros((sy, sx)), (2, 1): np.zeros((sy, sx))}
    st["T"] = {(1, 1): T, (2, 1): T, (1, 2): S, (2, 2): S}
    for e, nm in (("W", "WEST"), ("E", "EAST"), ("S", "SOUTH"), ("N", "NORTH")):
        if e in open_edges:
            L = (sy if e in "WE" else sx) - 2
            st["T_" + nm] = {(1, 1): np.repeat(Tb[:, 2, 2][:, None], L, 1), (1, 2): np.full((N, L), 35.0)}
    return cfg, st, Tb


def wave_channel(Lm=80, Mm=8, N=3, dx=1.0e3, depth=100.0, dt=60.0, ndtfast=30):
    """Forcing test of the open boundaries: a homogeneous flat channel at rest, periodic in eta, open to the west and east.  The caller
    pushes an incoming gravity wave as western boundary data step by step (zeta_west = a sin(wt), ubar_west = u_west = sqrt(g/h) zeta_west:
    the incoming characteristic); the eastern data stay zero.  Returns cfg, state."""
    cfg = RomsCfg(Lm=Lm, Mm=Mm, N=N, NT=2, ew_periodic=0, ns_periodic=1, nonlin_eos=0, salinity=1, ndtfast=ndtfast, np_xi=1, np_eta=1, inode=0, jnode=0,
                  dt=dt, rho0=1027.0, g=9.81, Tcoef=0.0, T0=10.0, Scoef=0.0, S0=35.0, rdrg=0.0, rdrg2=0.0, Zob=0.0, gamma2=1.0,
                  theta_s=3.0, theta_b=0.0, hc=25.0)
    cfg.obc_west = cfg.obc_east = 1
    cfg.ubind = 0.1
    sy, sx = Mm + 4, Lm + 4
    z2, z3 = np.zeros((sy, sx)), np.zeros((N, sy, sx))
    st = {"H": np.full((sy, sx), depth), "PM": np.full((sy, sx), 1.0 / dx), "PN": np.full((sy, sx), 1.0 / dx), "F": z2.copy(), "RMASK": np.ones((sy, sx)),
          "ZETA": {(1, 1): z2, (2, 1): z2}, "UBAR": {(1, 1): z2, (2, 1): z2}, "VBAR": {(1, 1): z2, (2, 1): z2},
          "U": {(1, 1): z3, (2, 1): z3}, "V": {(1, 1): z3, (2, 1): z3},
          "T": {(1, 1): z3 + 10.0, (2, 1): z3 + 10.0, (1, 2): z3 + 35.0, (2, 2): z3 + 35.0},
          "AKV": np.zeros((N + 1, sy, sx)), "AKT": {(1, 1): np.zeros((N + 1, sy, sx)), (1, 2): np.zeros((N + 1, sy, sx))}}
    for nm in ("WEST", "EAST"):
        st["T_" + nm] = {(1, 1): np.full((N, Mm + 2), 10.0), (1, 2): np.full((N, Mm + 2), 35.0)}
    return cfg, st


def channel_strips(Lm, Mm, N, strips=8, sub=(0, 0, 1, 1), **kw):
    """`channel(...)` for a large block without its dozen full-size temporaries: the rank's block is assembled from `strips`
    horizontal strips, each produced by `channel` itself as the block of a finer process grid (every formula is a function of the
    GLOBAL indices, so the values are those of the direct call, bit for bit).  Needs Mm divisible by np_eta*strips."""
    inode, jnode, np_xi, np_eta = sub
    if strips <= 1 or Mm % (np_eta * strips) != 0 or (Mm // (np_eta * strips)) < 4:
        return channel(Lm, Mm, N, sub=sub, **kw)
    cfg = st = None
    rows = Mm // (np_eta * strips)
    for s_ in range(strips):
        c, part = channel(Lm, Mm, N, sub=(inode, jnode * strips + s_, np_xi, np_eta * strips), **kw)
        if st is None:
            cfg = c
            cfg.Mm, cfg.np_eta, cfg.jnode = rows * strips, np_eta, jnode
            st = {}
        # rows of this strip that land in the assembled block: its interior, plus the 2 halo rows at the block's own ends
        lo = 0 if s_ == 0 else 2
        hi = rows + 4 if s_ == strips - 1 else rows + 2
        dst = slice(s_ * rows + lo, s_ * rows + hi)

        def put(key, sel, a):
            tgt = st.setdefault(key, {}) if sel is not None else st
            k = sel if sel is not None else key
            if k not in tgt:
                tgt[k] = np.empty(a.shape[:-2] + (rows * strips + 4, a.shape[-1]), dtype=a.dtype)
            tgt[k][..., dst, :] = a[..., lo:hi, :]
        for key, v in part.items():
            if isinstance(v, dict):
                seen = {}
                for sel, a in v.items():
                    if id(a) in seen:                      # (1,it) and (2,it) share one array in `channel`: keep them shared
                        st[key][sel] = st[key][seen[id(a)]]
                        continue
                    seen[id(a)] = sel
                    put(key, sel, a)
            elif isinstance(v, np.ndarray):
                put(key, None, v)
            else:
                st[key] = v
        del part
    return cfg, st


def filament(Lm=64, Mm=64, N=32):
    """tests/Filament analytical case (ana_grid.h, ana_init.h) evaluated with numpy."""
    cfg = RomsCfg(Lm=Lm, Mm=Mm, N=N, NT=1, ew_periodic=1, ns_periodic=1, nonlin_eos=0, salinity=0, ndtfast=60,
                  np_xi=1, np_eta=1, inode=0, jnode=0, dt=5.0, rho0=1000.0, g=9.81, Tcoef=0.20, T0=1.0, Scoef=0.822, S0=1.0,
                  rdrg=0.0, rdrg2=1.0e-3, Zob=1.0e-2, gamma2=1.0, theta_s=6.0, theta_b=2.0, hc=25.0)
    sy, sx = Mm + 4, Lm + 4
    SizeX, SizeY, HD = 12.8e3, 3.2e3, 1000.0
    dx, dy = SizeX / Lm, SizeY / Mm
    i = np.arange(-1, Lm + 3)[None, :] + np.zeros((sy, 1))
    xr = dx * (i - 0.5) - SizeX / 2.0
    st = {"H": np.full((sy, sx), HD), "PM": np.full((sy, sx), 1.0 / dx), "PN": np.full((sy, sx), 1.0 / dy),
          "F": np.full((sy, sx), 2 * 7.81e-5), "RMASK": np.ones((sy, sx))}
    h = st["H"]
    z_r, z_w = _depths(h, np.zeros((sy, sx)), N, cfg.theta_s, cfg.theta_b, cfg.hc)
    Hz = z_w[1:] - z_w[:-1]
    g, alpha = cfg.g, cfg.Tcoef / cfg.rho0
    b0, B, lam, Nb, N0, h0, dh0, L = 5.0e-2, 0.025, 8.0, 1.0e-7, 3.0e-5, 60.0, 15.0, 2000.0
    hs = h0 + dh0 * np.exp(-(xr / L) ** 2)
    bfun = lambda z, hsbl: b0 + Nb * (z + HD) + 0.5 * N0 * ((1 + B) * z - (1 - B) * (hsbl + lam * np.log(np.cosh((1.0 / lam) * (z + hsbl)))))
    T = bfun(z_r, hs[None]) / (g * alpha)
    bf_int = (Hz[:, 2, 2] * bfun(z_r[:, 2, 2], h0) / g).sum()
    zeta = (T * alpha * Hz).sum(0) - bf_int
    f = st["F"]; pm = st["PM"]
    v = np.zeros_like(T)
    dzdx = np.zeros((sy, sx)); dzdx[:, 1:-1] = 0.5 * pm[:, 1:-1] * (zeta[:, 2:] - zeta[:, :-2])
    v[N - 1] = g * dzdx / f
    for k in range(N - 2, -1, -1):
        dbdx = np.zeros((sy, sx))
        dbdx[:, 1:-1] = 0.25 * pm[:, 1:-1] * g * alpha * (T[k, :, 2:] - T[k, :, :-2] + T[k + 1, :, 2:] - T[k + 1, :, :-2])
        v[k] = v[k + 1] - dbdx * (z_r[k + 1] - z_r[k]) / f
    vbar = (v[:N - 1] * Hz[:N - 1]).sum(0) / HD
    u = np.zeros_like(v); ubar = np.zeros((sy, sx))
    for a in (zeta, vbar, v, T):
        _wrap(a, True, True)
    st["ZETA"] = {(1, 1): zeta, (2, 1): zeta}
    st["UBAR"] = {(1, 1): ubar}
    st["VBAR"] = {(1, 1): vbar, (2, 1): vbar}
    st["U"] = {(1, 1): u, (2, 1): u}
    st["V"] = {(1, 1): v, (2, 1): v}
    st["T"] = {(1, 1): T, (2, 1): T}
    return cfg, st


def rivers_ana_grid():
    """Grid, mask and run parameters of the reference's tests/Rivers_ana (ana_grid.h, param.opt, benchmark.in): 100x100x10,
    dx=dy=100 m, closed walls, flat 5 m shelf over the southern fifth then a uniform slope to 100 m, land strip with a river
    channel.  Only the start-up quantities of this case are pinned by the reference's log (nfast, Cs_w table, hmin/hmax,
    area/volume, Cg_min/Cg_max: benchmark.result_github_gnu:58,101-118): its time stepping needs KPP and river forcing."""
    Lm = Mm = 100
    N = 10
    cfg = RomsCfg(Lm=Lm, Mm=Mm, N=N, NT=2, ew_periodic=0, ns_periodic=0, nonlin_eos=1, salinity=1, ndtfast=30,
                  np_xi=1, np_eta=1, inode=0, jnode=0, dt=20.0, rho0=1027.5, g=9.81, Tcoef=0.0, T0=0.0, Scoef=0.0, S0=0.0,
                  rdrg=0.0, rdrg2=1.0e-3, Zob=1.0e-2, gamma2=1.0, theta_s=6.0, theta_b=6.0, hc=25.0)
    sy, sx = Mm + 4, Lm + 4
    size_xi = size_eta = 1.0e4
    depth, max_depth = 5.0, 100.0
    dx, dy = size_xi / Lm, size_eta / Mm
    i = np.arange(-1, Lm + 3)[None, :] + np.zeros((sy, 1))
    j = np.arange(-1, Mm + 3)[:, None] + np.zeros((1, sx))
    xr, yr = dx * (i - 0.5), dy * (j - 0.5)
    shelf = size_eta / 5
    slope = (max_depth - depth) / (size_eta * 4 / 5)
    h = np.where(yr < shelf, depth, depth + (yr - shelf) * slope)
    land, coast, riv_west, riv_east = size_eta * 0.1, size_eta * 0.02, size_xi * 0.4, size_xi * 0.6
    rmask = np.ones((sy, sx))
    rmask[(yr < land) & ((xr < riv_west) | (xr > riv_east))] = 0.0
    rmask[yr < coast] = 0.0
    st = {"H": h, "PM": np.full((sy, sx), 1.0 / dx), "PN": np.full((sy, sx), 1.0 / dy), "F": np.zeros((sy, sx)), "RMASK": rmask}
    return cfg, st


def rivers_ana(akv=0.0, akt=0.0, kpp=False):
    """The reference's tests/Rivers_ana case without its KPP closure: grid and mask of `rivers_ana_grid`, ana_init.h (rest,
    T = 4 + 10 exp(z/50), S = 36), zero surface forcing (analytical.F:216-305), the analytic river of river_frc.F:125-137 /
    ana_frc_river.h (20 coast cells at the head of the channel share 500 m3/s of 24 degC, 1 PSU water).  Akv, Akt are uniform
    stand-ins for what lmd_vmix/lmd_kpp would compute; kpp=True switches the closure itself on (state key "KPP") and
    the UV_VIS2/TS_DIF2 switches of the case's cppdefs.opt (with its zero coefficients)."""
    cfg, st = rivers_ana_grid()
    if kpp:
        cfg.uv_vis2, cfg.ts_dif2 = 1, 1
        st["KPP"] = True
    N = cfg.N
    sy, sx = cfg.Mm + 4, cfg.Lm + 4
    h, rmask = st["H"], st["RMASK"]
    z_r, z_w = _depths(h, np.zeros_like(h), N, cfg.theta_s, cfg.theta_b, cfg.hc)
    T = 4.0 + 10.0 * np.exp(z_r / 50.0)
    S = np.full_like(T, 36.0)
    zero2, zero3 = np.zeros((sy, sx)), np.zeros_like(T)
    xr = 100.0 * (np.arange(-1, cfg.Lm + 3)[None, :] + np.zeros((sy, 1)) - 0.5)
    ridx = np.zeros_like(rmask); rfrc = np.zeros_like(rmask)
    head = np.zeros_like(rmask, dtype=bool)
    head[:-1, :] = (rmask[:-1, :] == 0) & (rmask[1:, :] == 1) & (xr[:-1, :] > 4000.0) & (xr[:-1, :] < 6000.0)
    ridx[head] = 1.0; rfrc[head] = 1.0 / 20.0
    ru, rv = river_faces(rmask, ridx, rfrc)
    st.update({"ZETA": {(1, 1): zero2}, "UBAR": {(1, 1): zero2}, "VBAR": {(1, 1): zero2}, "U": {(1, 1): zero3, (2, 1): zero3},
               "V": {(1, 1): zero3, (2, 1): zero3}, "T": {(1, 1): T, (2, 1): T, (1, 2): S, (2, 2): S},
               "AKV": np.full((N + 1, sy, sx), akv), "AKT": {(1, 1): np.full((N + 1, sy, sx), akt), (1, 2): np.full((N + 1, sy, sx), akt)},
               "RIV_UFLX": ru, "RIV_VFLX": rv, "RIVERS": [(5.0e2, [24.0, 1.0])]})
    return cfg, st


def take_block(cfg, state, sub):
    """Cuts the block sub=(inode, jnode, np_xi, np_eta) (mpi_setup.F:61-70, extents as `block_extent`) out of a single-domain
    cfg/state: every field keeps its 2-cell halo, which at an interior block edge holds the neighbour's values."""
    import copy
    inode, jnode, np_xi, np_eta = sub
    (Lm, i0), (Mm, j0) = block_extent(cfg.Lm, np_xi, inode), block_extent(cfg.Mm, np_eta, jnode)
    c = copy.copy(cfg)
    c.Lm, c.Mm, c.np_xi, c.np_eta, c.inode, c.jnode = Lm, Mm, np_xi, np_eta, inode, jnode
    cut = lambda a: np.ascontiguousarray(a[..., j0:j0 + Mm + 4, i0:i0 + Lm + 4])
    cut_j = lambda a: np.ascontiguousarray(a[..., j0:j0 + Mm + 2])      # boundary data of the western / eastern edge: lines over j = 0:Mm+1
    cut_i = lambda a: np.ascontiguousarray(a[..., i0:i0 + Lm + 2])      # southern / northern edge: lines over i = 0:Lm+1
    st = {}
    for k, v in state.items():
        side = k.rsplit("_", 1)[-1]
        f = cut_j if side in ("WEST", "EAST") else cut_i if side in ("SOUTH", "NORTH") else cut
        if isinstance(v, dict):
            st[k] = {kk: f(a) for kk, a in v.items()}
        elif isinstance(v, np.ndarray):
            st[k] = f(v)
        else:
            st[k] = v
    return c, st


def pipes_ana(kpp=True):
    """The reference's tests/Pipes_ana case (ana_grid.h, ana_init.h, ana_pipe_frc.h, benchmark.in): 100x100x10, dx=dy=300 m,
    f=1e-4, depth 10 m shelf then a slope to 100 m, the land strip of Rivers_ana, closed walls, dt=60 s, ndtfast=30; a 2x2-cell
    pipe in the middle of the domain releases 500 m3/s of 24 degC, 1 PSU water over the two bottom levels."""
    Lm = Mm = 100
    N = 10
    cfg = RomsCfg(Lm=Lm, Mm=Mm, N=N, NT=2, ew_periodic=0, ns_periodic=0, nonlin_eos=1, salinity=1, ndtfast=30,
                  np_xi=1, np_eta=1, inode=0, jnode=0, dt=60.0, rho0=1027.5, g=9.81, Tcoef=0.0, T0=0.0, Scoef=0.0, S0=0.0,
                  rdrg=0.0, rdrg2=1.0e-3, Zob=1.0e-2, gamma2=1.0, theta_s=6.0, theta_b=6.0, hc=25.0)
    cfg.uv_vis2, cfg.ts_dif2 = 1, 1
    sy, sx = Mm + 4, Lm + 4
    SizeX = SizeY = 30.0e3
    dx, dy = SizeX / Lm, SizeY / Mm
    i = np.arange(-1, Lm + 3)[None, :] + np.zeros((sy, 1))
    j = np.arange(-1, Mm + 3)[:, None] + np.zeros((1, sx))
    xr, yr = dx * (i - 0.5), dy * (j - 0.5)
    depth, max_depth = 10.0, 100.0
    shelf = SizeY / 5
    slope = (max_depth - depth) / (SizeY * 4 / 5)
    h = np.where(yr < shelf, depth, depth + (yr - shelf) * slope)
    land, coast, riv_west, riv_east = SizeY * 0.1, SizeY * 0.02, SizeX * 0.4, SizeX * 0.6
    rmask = np.ones((sy, sx))
    rmask[(yr < land) & ((xr < riv_west) | (xr > riv_east))] = 0.0
    rmask[yr < coast] = 0.0
    psz, px, py = SizeX * 0.02, SizeX * .5, SizeY * .5
    pipe_cells = float(round(psz / dx)) ** 2
    sel = (xr > px - 0.5 * psz) & (xr < px + 0.5 * psz) & (yr > py - 0.5 * psz) & (yr < py + 0.5 * psz)
    pidx = np.where(sel, 1.0, 0.0)
    pflx = np.where(sel, 1.0 / pipe_cells, 0.0) * 5.0e2
    z_r, z_w = _depths(h, np.zeros_like(h), N, cfg.theta_s, cfg.theta_b, cfg.hc)
    T = 4.0 + 10.0 * np.exp(z_r / 50.0)
    S = np.full_like(T, 36.0)
    zero2, zero3 = np.zeros((sy, sx)), np.zeros_like(T)
    st = {"H": h, "PM": np.full((sy, sx), 1.0 / dx), "PN": np.full((sy, sx), 1.0 / dy), "F": np.full((sy, sx), 1.0e-4), "RMASK": rmask,
          "ZETA": {(1, 1): zero2}, "UBAR": {(1, 1): zero2}, "VBAR": {(1, 1): zero2}, "U": {(1, 1): zero3, (2, 1): zero3},
          "V": {(1, 1): zero3, (2, 1): zero3}, "T": {(1, 1): T, (2, 1): T, (1, 2): S, (2, 2): S},
          "AKV": np.zeros((N + 1, sy, sx)), "AKT": {(1, 1): np.zeros((N + 1, sy, sx)), (1, 2): np.zeros((N + 1, sy, sx))},
          "PIPE_IDX": pidx, "PIPE_FLX": pflx, "PIPES": [([0.5, 0.5] + [0.0] * (N - 2), [24.0, 1.0])]}
    if kpp:
        st["KPP"] = True
    return cfg, st


def initialize(model, state):
    """roms_init (main.F:86-321) for a host without Fortran: push the analytic fields, derive the
    grid metrics, then set_depth / set_HUV / omega / rho_eos."""
    grid = {k: state[k] for k in ("H", "PM", "PN", "F", "RMASK")}
    model.push_state(grid)
    model.call("setup_grid")
    model.push_state({k: v for k, v in state.items() if k not in grid and k not in ("RIVERS", "PIPES", "KPP")})
    if state.get("KPP"):
        model.call("set_kpp", 1)     # lmd_vmix/lmd_kpp
    for n, (vol, trc) in enumerate(state.get("RIVERS", []), 1):
        model.set_river(n, vol, trc)
    for n, (prf, trc) in enumerate(state.get("PIPES", []), 1):
        model.set_pipe(n, prf, trc)
    if "PTIDE" in state:
        model.call("set_pot_tides", 1)
    model.call("init_step0")
