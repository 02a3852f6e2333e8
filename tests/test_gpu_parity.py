"""GPU parity tests: libromsb200.so (through its C ABI) against the CPU oracle on the same analytic
inputs.  Tolerances are north_star's: relative L2 <= 1e-12 after one baroclinic step and <= 1e-8
after 100 steps on zeta, ubar/vbar, u, v, T, S, rho."""
import json
import os
import numpy as np
import pytest
from orclib import Oracle, cases
from romsb200 import Roms
import parity_walk

pytestmark = pytest.mark.gpu
TOL1, TOL100 = 1e-12, 1e-8
GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "filament_diag.json")))


def prognostic(m, NT):
    iic, iif, nstp, nrhs, nnew, kstp, knew = m.get_idx()
    out = {"zeta": m.pull("ZETA", knew), "ubar": m.pull("UBAR", knew), "vbar": m.pull("VBAR", knew),
           "u": m.pull("U", nnew), "v": m.pull("V", nnew), "rho": m.pull("RHO1")}
    for it in range(1, NT + 1):
        out["T" if it == 1 else "S" if it == 2 else f"t{it}"] = m.pull("T", nnew, it)
    return out


def compare(g, o, NT, tol, cfg=None):
    a, b = prognostic(g, NT), prognostic(o, NT)
    res = {}
    for k in b:
        x, y = (parity_walk.view(a[k], cfg), parity_walk.view(b[k], cfg)) if cfg is not None else (a[k], b[k])
        assert np.isfinite(x).all(), k
        res[k] = parity_walk.rel(x, y)
    assert max(res.values()) <= tol, res
    return res


def test_push_pull_roundtrip():
    cfg, st = cases.channel(20, 12, 6)
    g = Roms(cfg)
    rng = np.random.default_rng(1)
    for f, lvl, it in [("ZETA", 3, 1), ("U", 2, 1), ("T", 3, 2), ("Z_W", 1, 1), ("AKT", 1, 2)]:
        a = rng.standard_normal(g.shape(f))
        g.push(f, a, lvl, it)
        assert np.array_equal(g.pull(f, lvl, it), a)
    g.close()


def test_async_push_pull_roundtrip():
    """roms_push_async / roms_pull_async: same planes as the blocking calls once roms_sync has returned."""
    import ctypes as C
    from romsb200 import FID
    cfg, st = cases.channel(20, 12, 6)
    g = Roms(cfg)
    rng = np.random.default_rng(2)
    dptr = C.POINTER(C.c_double)
    for f, lvl, it in [("ZETA", 2, 1), ("V", 1, 1), ("T", 2, 2)]:
        a = np.ascontiguousarray(rng.standard_normal(g.shape(f)))
        b = np.empty_like(a)
        assert g.lib.roms_push_async(g.h, FID[f], lvl, it, a.ctypes.data_as(dptr)) == 0
        assert g.lib.roms_pull_async(g.h, FID[f], lvl, it, b.ctypes.data_as(dptr)) == 0
        g.sync()
        assert np.array_equal(a, b)
        assert np.array_equal(g.pull(f, lvl, it), a)
    g.close()


def test_constants_match_oracle():
    cfg, _ = cases.filament()
    g, o = Roms(cfg), Oracle(cfg)
    assert g.nfast() == o.nfast() == 82
    for a, b in zip(g.weights(82), o.weights(82)):
        assert np.array_equal(a, b)
    for a, b in zip(g.scoord(), o.scoord()):
        assert np.array_equal(a, b)
    g.close(); o.close()


def test_rivers_ana_startup_on_gpu():
    """nfast, weights, Cs tables and the setup_grid2 report of the reference's Rivers_ana grid (masked, sloping bottom, closed
    walls): library == oracle, and both == the reference's log (tests/golden/rivers_ana_startup.json)."""
    G = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "rivers_ana_startup.json")))
    cfg, st = cases.rivers_ana_grid()
    g, o = Roms(cfg), Oracle(cfg)
    assert g.nfast() == o.nfast() == G["nfast"]
    for a, b in zip(g.weights(41), o.weights(41)):
        assert np.array_equal(a, b)
    for a, b in zip(g.scoord(), o.scoord()):
        assert np.array_equal(a, b)
    for m in (g, o):
        m.push_state(st)
        m.call("setup_grid")
    sg, so = g.grid_stats(), o.grid_stats()
    assert np.array_equal(sg[:7], so[:7]) and np.allclose(sg[7:], so[7:], rtol=1e-14, atol=0)
    assert sg[0] == G["hmin"] and sg[1] == G["hmax"] and abs(sg[7] / G["area"] - 1) < 1e-13 and abs(sg[8] / G["volume"] - 1) < 1e-13
    for f in ("UMASK", "VMASK", "PMASK", "DM_U", "DN_V", "PMON_U", "FOMN"):
        assert np.array_equal(parity_walk.view(g.pull(f), cfg), parity_walk.view(o.pull(f), cfg)), f
    g.close(); o.close()


@pytest.mark.parametrize("case", ["channel", "filament", "channel_linear_odd", "channel_walls", "channel_walls_linear_odd",
                                  "seamount_island_walls", "seamount_island_periodic", "closed_basin", "closed_basin_seamount_odd",
                                  "ew_walls_ns_periodic", "channel_mixing", "closed_basin_seamount_mixing", "filament_mixing_switches",
                                  "closed_basin_river", "walled_channel_river_mixing", "walled_channel_pipe", "closed_basin_river_pipe", "closed_basin_tide", "rivers_ana_without_kpp",
                                  "channel_nt8", "closed_basin_nt7_river", "channel_walls_n64", "closed_basin_seamount_n100", "periodic_nt5_mixing"])
def test_routine_by_routine(case):
    if case == "channel":
        cfg, st = cases.channel(48, 40, 12)
    elif case == "channel_walls":
        cfg, st = cases.channel(48, 40, 12, ns_walls=True)             # closed N/S walls (configs[1] geometry)
    elif case == "seamount_island_walls":                              # bottom topography + land mask + closed N/S walls
        cfg, st = cases.channel(48, 40, 12, ns_walls=True, seamount=True, island=True)
    elif case == "seamount_island_periodic":
        cfg, st = cases.channel(45, 38, 11, seamount=True, island=True)
    elif case == "closed_basin":                                        # walls on all four sides (corner rules)
        cfg, st = cases.channel(48, 40, 12, ns_walls=True, ew_walls=True)
    elif case == "closed_basin_seamount_odd":                           # BASELINE configs[2] geometry, ragged sizes, land mask
        cfg, st = cases.channel(37, 29, 9, ns_walls=True, ew_walls=True, seamount=True, island=True)
    elif case == "ew_walls_ns_periodic":
        cfg, st = cases.channel(40, 36, 10, ew_walls=True)
    elif case == "channel_walls_linear_odd":
        cfg, st = cases.channel(37, 29, 9, NT=3, nonlin_eos=False, ns_walls=True)
    elif case == "channel_mixing":                                      # UV_VIS2 + TS_DIF2, spatially varying coefficients, 3 tracers
        cfg, st = cases.channel(45, 38, 11, NT=3, visc2=80.0, tnu2=40.0, sponge=True)
    elif case == "closed_basin_seamount_mixing":                        # + walls on all sides, land mask (pmask = 0/1/2)
        cfg, st = cases.channel(37, 29, 9, ns_walls=True, ew_walls=True, seamount=True, island=True, visc2=120.0, tnu2=60.0, sponge=True)
    elif case == "filament_mixing_switches":                            # tests/Filament/cppdefs.opt:34-35 with visc2 = tnu2 = 0
        cfg, st = cases.filament(64, 64, 32)
        cfg.uv_vis2, cfg.ts_dif2 = 1, 1
    elif case == "closed_basin_river":                                  # river source through a land strip (Rivers_ana geometry)
        cfg, st = cases.channel(40, 36, 8, ns_walls=True, ew_walls=True, river=True, dt=60.0)
    elif case == "walled_channel_river_mixing":                         # E-W periodic, 3 tracers, seamount, lateral mixing + river
        cfg, st = cases.channel(45, 38, 9, NT=3, ns_walls=True, seamount=True, river=True, dt=60.0, visc2=60.0, tnu2=30.0)
    elif case == "walled_channel_pipe":                                 # pipe outfall (pipe_frc.F) in an E-W periodic channel
        cfg, st = cases.channel(40, 36, 8, ns_walls=True, pipe=True, dt=60.0)
    elif case == "closed_basin_river_pipe":                             # both sources, 3 tracers, ragged sizes
        cfg, st = cases.channel(37, 31, 9, NT=3, ns_walls=True, ew_walls=True, river=True, pipe=True, dt=60.0)
    elif case == "closed_basin_tide":                                   # tidal potential in the surface pressure (prsgrd.F:209-211)
        cfg, st = cases.channel(40, 36, 8, ns_walls=True, ew_walls=True, seamount=True, tide=True)
    elif case == "rivers_ana_without_kpp":                              # the reference's tests/Rivers_ana, uniform Akv/Akt instead of KPP
        cfg, st = cases.rivers_ana(1.0e-4, 1.0e-5)
    elif case == "channel_nt8":                                         # tracer batches (T,S) + 4 + 2 (TrcBatch in kernels.cuh)
        cfg, st = cases.channel(40, 33, 10, NT=8, ns_walls=True)
    elif case == "closed_basin_nt7_river":                              # batches 2 + 4 + 1, river tracer values per tracer, land mask
        cfg, st = cases.channel(37, 31, 9, NT=7, ns_walls=True, ew_walls=True, river=True, pipe=True, dt=60.0)
    elif case == "periodic_nt5_mixing":                                 # batches 2 + 2 + 1, doubly periodic, per-tracer diffusivities
        cfg, st = cases.channel(33, 30, 8, NT=5, visc2=60.0, tnu2=30.0, sponge=True)
    elif case == "channel_walls_n64":                                   # the bench column height (configs[1]: N=64): KB_T=8 blocks, KC=4 chunks
        cfg, st = cases.channel(70, 44, 64, ns_walls=True)
    elif case == "closed_basin_seamount_n100":                          # configs[2] column height (N=100 is not a multiple of 8)
        cfg, st = cases.channel(44, 36, 100, ns_walls=True, ew_walls=True, seamount=True)
    elif case == "filament":
        cfg, st = cases.filament(64, 64, 32)
    else:
        cfg, st = cases.channel(37, 29, 9, NT=3, nonlin_eos=False)     # ragged sizes, passive tracer
    w = parity_walk.walk(cfg, st, nsteps=2)
    bad = {k: v for k, v in w.items() if not v <= 1e-11}
    assert not bad, bad


@pytest.mark.parametrize("walls,topo,ew,mix", [(False, False, False, False), (True, False, False, False), (True, True, False, False),
                                               (True, True, True, False), (True, True, True, True)])
def test_one_step_and_100_steps_channel(walls, topo, ew, mix):
    cfg, st = cases.channel(64, 48, 16, ns_walls=walls, seamount=topo, island=topo, ew_walls=ew, **(dict(visc2=100.0, tnu2=50.0, sponge=True) if mix else {}))
    g, o = Roms(cfg), Oracle(cfg)
    cases.initialize(g, st); cases.initialize(o, st)
    g.call("step"); o.call("step")
    r1 = compare(g, o, cfg.NT, TOL1, cfg)
    g.call("run", 99); o.call("run", 99)
    r100 = compare(g, o, cfg.NT, TOL100, cfg)
    print("1 step:", r1, "\n100 steps:", r100)
    d1, d2 = g.diag(), o.diag()
    assert np.allclose(d1, d2, rtol=1e-9)
    g.close(); o.close()


@pytest.mark.parametrize("env", [{"ROMS_S2D_TMA": "0"}, {"ROMS_S2D_TMA": "1"}, {"ROMS_S2D_TMA": "2", "ROMS_S2D_ROWS": "5"}, {"ROMS_S2D_TMA": "6", "ROMS_S2D_ROWS": "7"},
                                 {"ROMS_S2D_TMA": "12", "ROMS_S2D_ROWS": "9"}, {"ROMS_UV2_FUSED": "0"}, {"ROMS_T_TMA": "0"}])
@pytest.mark.parametrize("closed", [False, True])
def test_kernel_variants_are_bit_identical(env, closed, monkeypatch):
    """The shipped alternatives of the barotropic substep (tile kernel, full TMA ring, another ring shape, register marching, the
    default hybrid with ragged row blocks), the unfused step3d_uv2 passes and the scalar-load plane kernels: every variant must
    reproduce the oracle exactly like the defaults do."""
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    cfg, st = cases.channel(70, 45, 12, ns_walls=True, seamount=closed, island=closed, ew_walls=closed)
    g, o = Roms(cfg), Oracle(cfg)
    cases.initialize(g, st); cases.initialize(o, st)
    g.call("run", 3); o.call("run", 3)
    r = compare(g, o, cfg.NT, TOL1, cfg)
    assert max(r.values()) == 0.0, r
    g.close(); o.close()


def test_filament_golden_on_gpu():
    """The CUDA path itself against the reference's golden log (tests/Filament/benchmark.result_github_gnu:146-167)."""
    cfg, _ = cases.filament()
    cfg.uv_vis2, cfg.ts_dif2 = 1, 1     # tests/Filament/cppdefs.opt:34-35 (UV_VIS2, TS_DIF2 compiled in; visc2 = tnu2 = 0 in benchmark.in)
    o = Oracle(cfg)
    o.call("init_filament")            # exact restatement of ana_grid.h / ana_init.h
    g = Roms(cfg)
    for f in ("H", "PM", "PN", "F", "RMASK"):
        g.push(f, o.pull(f))
    g.call("setup_grid")
    for f, lv in (("ZETA", (1, 2)), ("UBAR", (1,)), ("VBAR", (1, 2)), ("U", (1, 2)), ("V", (1, 2)), ("T", (1, 2))):
        for l in lv:
            g.push(f, o.pull(f, l), l)
    g.call("init_step0")
    gold = np.array([[float(x) for x in r] for r in GOLD["github_gnu"]])
    for s in range(21):
        if s:
            g.call("step")
        d = g.diag()
        rel = np.abs(d[:3] - gold[s, :3]) / np.abs(gold[s, :3])
        assert rel.max() < 1e-11, (s, d, gold[s])
        assert d[3] == 0.0
    g.close(); o.close()


def test_alfabeta():
    cfg, st = cases.channel(24, 20, 8)
    g, o = Roms(cfg), Oracle(cfg)
    cases.initialize(g, st); cases.initialize(o, st)
    a, b = g.alfabeta()
    import ctypes as C
    ao = np.empty(o.shape2); bo = np.empty(o.shape2)
    o.call("alfabeta", ao.ctypes.data_as(C.POINTER(C.c_double)), bo.ctypes.data_as(C.POINTER(C.c_double)))
    assert parity_walk.rel(a, ao) < 1e-14 and parity_walk.rel(b, bo) < 1e-14
    g.close(); o.close()


def test_river_source_budgets_and_100_steps():
    """River sources (river_frc.F; hooks step2d_FB.F:531-554, pre_step3d4S.F:493-522, step3d_uv2.F:689-717,
    compute_horiz_tracer_fluxes.h:217-246): no golden log of the reference pins them without KPP, so besides GPU-vs-oracle
    parity the volume, heat and salt budgets must close: per baroclinic step the basin gains riv_vol*dt of water and
    riv_vol*dt*riv_trc of each tracer."""
    cfg, st = cases.channel(40, 36, 8, ns_walls=True, ew_walls=True, river=True, dt=60.0)
    g, o = Roms(cfg), Oracle(cfg)
    cases.initialize(g, st); cases.initialize(o, st)
    vol, trc = st["RIVERS"][0]
    rm = st["RMASK"][2:-2, 2:-2]
    dA = 1.0e6

    def budget(m):
        iic, iif, nstp, nrhs, nnew, kstp, knew = m.get_idx()
        Hz = m.pull("HZ")[:, 2:-2, 2:-2]
        return np.array([(m.pull("ZETA", knew)[2:-2, 2:-2] * rm).sum() * dA] +
                        [(Hz * m.pull("T", nnew, it)[:, 2:-2, 2:-2] * rm).sum() * dA for it in (1, 2)])
    g.call("step"); o.call("step")
    compare(g, o, cfg.NT, TOL1, cfg)
    b0 = budget(g)
    for n in range(3):
        g.call("step"); o.call("step")
        b = budget(g)
        gain = (b - b0) / (vol * cfg.dt * np.array([1.0, trc[0], trc[1]]))
        assert np.all(np.abs(gain - 1.0) < 1e-6), gain
        b0 = b
    g.call("run", 96); o.call("run", 96)
    r100 = compare(g, o, cfg.NT, TOL100, cfg)
    print("river case, 100 steps:", r100)
    g.close(); o.close()


def test_rivers_ana_without_kpp_on_gpu():
    """tests/Rivers_ana of the reference (closed walls, land mask, sloping bottom, nonlinear EOS, river) with uniform Akv/Akt in
    place of KPP: 20 steps on the GPU are bit-identical to the oracle, and step 2 of the reference's log is met to 5e-9
    (see tests/test_oracle_golden.py::test_rivers_ana_golden_without_kpp for what that pins)."""
    G = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "rivers_ana_diag.json")))["diag"]
    cfg, st = cases.rivers_ana(1.0e-4, 1.0e-5)
    g, o = Roms(cfg), Oracle(cfg)
    cases.initialize(g, st); cases.initialize(o, st)
    g.call("run", 2); o.call("run", 2)
    d = g.diag()
    assert abs(d[0] / G[2]["KE"] - 1) < 5e-9 and abs(d[1] / G[2]["BAROTR_KE"] - 1) < 5e-9, d
    g.call("run", 18); o.call("run", 18)
    r = compare(g, o, cfg.NT, 1e-11, cfg)
    print("Rivers_ana without KPP, 20 steps:", r)
    g.close(); o.close()


def test_kpp_closure_on_gpu():
    """LMD/KPP vertical mixing (lmd_vmix.F, lmd_kpp.F, lmd_swr_frac.F, bvf of rho_eos.F) on the reference's
    Rivers_ana case: one call of lmd_vmix against the oracle field by field (Akv, Akt, hbls, hbbl), then 20 steps of the whole
    model with the closure on, and the reference's own log (Courant columns to round-off, see test_rivers_ana_golden_with_kpp).
    sin/log/exp of the closure come from the CUDA math library here and from libm in the oracle, hence 1e-13 instead of 0."""
    G = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "rivers_ana_diag.json")))["diag"]
    cfg, st = cases.rivers_ana(kpp=True)
    g, o = Roms(cfg), Oracle(cfg)
    cases.initialize(g, st); cases.initialize(o, st)
    for m in (g, o):
        m.set_idx(1, 1, 1, 1, 3, 1, 1)
        m.call("lmd_vmix", 1)
    for f, it in (("AKV", 1), ("AKT", 1), ("AKT", 2)):
        a, b = parity_walk.view(g.pull(f, 1, it), cfg), parity_walk.view(o.pull(f, 1, it), cfg)
        assert parity_walk.rel(a, b) <= 1e-13, (f, it, parity_walk.rel(a, b))
    for a, b in zip(g.hbl(), o.hbl()):
        assert parity_walk.rel(parity_walk.view(a, cfg), parity_walk.view(b, cfg)) <= 1e-13
    g.close(); o.close()
    g, o = Roms(cfg), Oracle(cfg)
    cases.initialize(g, st); cases.initialize(o, st)
    for n in range(1, 21):
        g.call("step"); o.call("step")
        d = g.diag()
        if n <= 16:
            assert abs(d[2] / G[n]["MAX_ADV_CFL"] - 1) <= 1e-11 and abs(d[3] / G[n]["MAX_VERT_CFL"] - 1) <= 1e-11, (n, d)
    r = compare(g, o, cfg.NT, 1e-9, cfg)
    print("Rivers_ana with KPP, 20 steps:", r)
    g.close(); o.close()


def kpp_flux_case(Lm=40, Mm=36, N=12):
    """closed basin 200 m deep with the KPP closure on and NON-ZERO surface forcing: net cooling with shortwave heating
    (Bfsfc < 0 under the surface -> ghat != 0: the LMD_NONLOCAL terms of step3d_t_ISO.F:961-998), a salt flux and a
    non-uniform wind stress (ustar of lmd_kpp.F:176-180)."""
    cfg, st = cases.channel(Lm, Mm, N, ns_walls=True, ew_walls=True, seamount=True, depth=200.0, dt=60.0)
    cfg.Zob = 1.0e-2
    st["KPP"] = True
    sy, sx = Mm + 4, Lm + 4
    x = (np.arange(sx)[None, :] - 1.5) / Lm + np.zeros((sy, 1)); y = (np.arange(sy)[:, None] - 1.5) / Mm + np.zeros((1, sx))
    srflx = 3.7e-5 * (1.0 + 0.3 * np.cos(2 * np.pi * x))
    st["SRFLX"] = srflx
    st["STFLX"] = {(1, 1): srflx - 4.9e-5 * (1.0 + 0.5 * np.sin(2 * np.pi * y)), (1, 2): 2.0e-6 * (1.0 + 0.5 * np.cos(2 * np.pi * (x + y)))}
    st["SUSTR"] = 1.0e-4 * (1.0 + 0.8 * np.sin(2 * np.pi * y) * np.cos(np.pi * x))
    st["SVSTR"] = 0.6e-4 * np.cos(2 * np.pi * x) * (1.0 + 0.5 * y)
    return cfg, st


def test_kpp_surface_fluxes_on_gpu():
    """the consumer half of the closure (VERDICT f2): srflx*swr_frac and ghat in the tracer corrector, ustar from a
    non-uniform wind stress, LMD_CONVEC off and non-zero background coefficients (roms_set_kpp_opts)."""
    for opts in (None, (0, 1.0e-4, 1.0e-5, 2.0e-5)):
        cfg, st = kpp_flux_case()
        g, o = Roms(cfg), Oracle(cfg)
        for m in (g, o):
            cases.initialize(m, st)
            if opts:
                m.call("set_kpp_opts", opts[0], *[__import__("ctypes").c_double(v) for v in opts[1:]])
        g.call("step"); o.call("step")
        r1 = compare(g, o, cfg.NT, 1e-11, cfg)
        g.call("run", 19); o.call("run", 19)
        r20 = compare(g, o, cfg.NT, 1e-9, cfg)
        for f, it in (("AKV", 1), ("AKT", 1), ("AKT", 2)):
            a, b = parity_walk.view(g.pull(f, 1, it), cfg), parity_walk.view(o.pull(f, 1, it), cfg)
            assert parity_walk.rel(a, b) <= 1e-9, (f, it, parity_walk.rel(a, b))
        print("KPP with surface fluxes", opts, r1, r20)
        g.close(); o.close()


def test_pipes_ana_golden_on_gpu():
    """The CUDA path against the reference's Pipes_ana log directly: 21 lines x 4 columns (walls, land mask, slope, nonlinear EOS,
    Coriolis, KPP with convective adjustment, pipe source) to <= 1e-10 relative (the closure's sin/log come from the CUDA math
    library), and against the oracle field by field after the 20 steps."""
    G = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "pipes_ana_diag.json")))["diag"]
    cfg, st = cases.pipes_ana()
    g, o = Roms(cfg), Oracle(cfg)
    cases.initialize(g, st); cases.initialize(o, st)
    worst = 0.0
    for n in range(0, 21):
        if n:
            g.call("step"); o.call("step")
        d = g.diag()
        ref = np.array([G[n]["KE"], G[n]["BAROTR_KE"], G[n]["MAX_ADV_CFL"], G[n]["MAX_VERT_CFL"]])
        dev = np.where(ref != 0, np.abs(d - ref) / np.where(ref != 0, np.abs(ref), 1.0), np.abs(d))
        worst = max(worst, dev.max())
        assert dev.max() <= 1e-10, (n, d, ref)
    r = compare(g, o, cfg.NT, 1e-9, cfg)
    print("Pipes_ana golden on the GPU: worst deviation", worst, "; vs oracle after 20 steps:", r)
    g.close(); o.close()


def test_pipe_source_budgets_and_100_steps():
    """Pipe outfalls (pipe_frc.F; hooks step2d_FB.F:155-159, omega.F:102-108, step3d_t_ISO.F:927-935): GPU-vs-oracle parity
    and closed budgets: per baroclinic step the domain gains sum(pipe_flx)*dt of water and sum(pipe_flx)*dt*pipe_trc of each tracer."""
    cfg, st = cases.channel(40, 36, 8, ns_walls=True, pipe=True, dt=60.0)
    g, o = Roms(cfg), Oracle(cfg)
    cases.initialize(g, st); cases.initialize(o, st)
    vol, trc = st["PIPE_FLX"][2:-2, 2:-2].sum(), st["PIPES"][0][1]

    def budget(m):
        iic, iif, nstp, nrhs, nnew, kstp, knew = m.get_idx()
        Hz = m.pull("HZ")[:, 2:-2, 2:-2]
        return np.array([m.pull("ZETA", knew)[2:-2, 2:-2].sum() * 1.0e6] + [(Hz * m.pull("T", nnew, it)[:, 2:-2, 2:-2]).sum() * 1.0e6 for it in (1, 2)])
    g.call("step"); o.call("step")
    compare(g, o, cfg.NT, TOL1, cfg)
    b0 = budget(g)
    for n in range(3):
        g.call("step"); o.call("step")
        b = budget(g)
        gain = (b - b0) / (vol * cfg.dt * np.array([1.0, trc[0], trc[1]]))
        assert np.all(np.abs(gain - 1.0) < 1e-6), gain
        b0 = b
    g.call("run", 96); o.call("run", 96)
    r100 = compare(g, o, cfg.NT, TOL100, cfg)
    print("pipe case, 100 steps:", r100)
    g.close(); o.close()


def test_full_size_properties():
    """BASELINE configs[1] at FULL size (1024x1024x64, the bench workload), where the CPU oracle is out of reach: parity through
    size-independent properties.  (1) Translation: the E-W periodic problem shifted by 37 cells is the roll of the unshifted
    one, so after two baroclinic steps every prognostic field must be the roll of the unshifted result BIT FOR BIT (any
    tile-, block- or alignment-dependent arithmetic in a kernel would break this: 37 is coprime to every tile size).
    (2) Conservation: with closed N/S walls and no surface fluxes the volume (sum zeta*dA) and the heat and salt content
    (sum Hz*T*dA) are conserved to round-off."""
    Lm, Mm, N, shift = 1024, 1024, 64, 37

    def run(ishift):
        cfg, st = cases.channel(Lm, Mm, N, ns_walls=True, ishift=ishift)
        g = Roms(cfg)
        cases.initialize(g, st)
        del st
        dA = 1.0e6
        Hz0 = g.pull("HZ")[:, 2:-2, 2:-2]
        c0 = [float((Hz0 * g.pull("T", 1, it)[:, 2:-2, 2:-2]).sum()) * dA for it in (1, 2)]
        v0 = float(g.pull("ZETA", 1)[2:-2, 2:-2].sum()) * dA
        hsum = float(g.pull("H")[2:-2, 2:-2].sum()) * dA
        del Hz0
        g.call("run", 2)
        iic, iif, nstp, nrhs, nnew, kstp, knew = g.get_idx()
        out = {"zeta": g.pull("ZETA", knew), "ubar": g.pull("UBAR", knew), "vbar": g.pull("VBAR", knew), "u": g.pull("U", nnew),
               "v": g.pull("V", nnew), "T": g.pull("T", nnew, 1), "S": g.pull("T", nnew, 2), "rho": g.pull("RHO1")}
        Hz1 = g.pull("HZ")[:, 2:-2, 2:-2]
        c1 = [float((Hz1 * out[k][:, 2:-2, 2:-2]).sum()) * dA for k in ("T", "S")]
        v1 = float(out["zeta"][2:-2, 2:-2].sum()) * dA
        d = g.diag()
        g.close()
        cons = {"volume": abs(v1 - v0) / hsum, "heat": abs(c1[0] - c0[0]) / abs(c0[0]), "salt": abs(c1[1] - c0[1]) / abs(c0[1])}
        return {k: np.ascontiguousarray(a[..., 2:-2, 2:-2]) for k, a in out.items()}, cons, d

    a, cons, da = run(0)
    print("full-size conservation after 2 steps:", cons)
    assert cons["volume"] <= 1e-15 and cons["heat"] <= 1e-12 and cons["salt"] <= 1e-12, cons
    assert all(np.isfinite(x).all() for x in a.values())
    b, cons_b, db = run(shift)
    for k in a:
        assert np.array_equal(np.roll(a[k], -shift, axis=-1), b[k]), k
    assert np.allclose(da, db, rtol=1e-12)


def test_two_gpu_halo_exchange_parity():
    """NCCL halo swaps: 2 ranks (2x1 split) against the single-domain oracle; skipped on a 1-GPU box."""
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    script = os.path.join(os.path.dirname(__file__), "mgpu_parity.py")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29731", script, "3"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29732", script, "3", "walls"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29733", script, "3", "seamount"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29734", script, "3", "closed"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29735", script, "3", "closedmix"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    for port, mode in ((29736, "uneven"), (29737, "unevenclosed"), (29738, "river"), (29739, "kpp"), (29740, "obc")):      # blocks of unequal size (mpi_setup.F:117-160); river source; KPP; open edges
        r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                            "--master-port", str(port), script, "3", mode], capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]


def test_six_gpu_run_reproduces_the_rivers_ana_log():
    """The CUDA path decomposed like the run that wrote the reference's Rivers_ana log (3x2 blocks) and given the bathymetry as that run
    had it (h = 0 in the second halo line of every block: tests/Rivers_ana/ana_grid.h fills 0:nx+1 only, analytic grids are never
    exchanged) reproduces the log itself on all 21 lines x 4 columns (observed <= 4.3e-16 on the energies, <= 4.3e-13 on the Courant
    numbers: profiles/r2_rivers_ana_reflog_6gpu.txt).  Needs 6 GPUs; skipped on smaller boxes."""
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 6:
        pytest.skip("needs 6 GPUs")
    script = os.path.join(os.path.dirname(__file__), "mgpu_parity.py")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=6", "--master-addr", "127.0.0.1",
                        "--master-port", "29750", script, "20", "reflog"], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
