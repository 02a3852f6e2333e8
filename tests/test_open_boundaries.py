"""Open boundaries: the OBC_M2FLATHER / OBC_M3ORLANSKI / OBC_TORLANSKI combination of the reference
(zetabc.F, u2dbc_im.F, v2dbc_im.F, u3dbc_im.F, v3dbc_im.F, t3dbc_im.F) with boundary data (boundary.F).

PARITY UNPINNED: every reference case with an open edge (tests/Flux_frc, Pipes_real, Rivers_real, CDR_*) needs netCDF input that
is not part of the reference tree, so no golden log covers these branches.  The oracle restates them statement by statement;
what stands in for the pin are the physical properties below (CPU), and the GPU path is compared with the oracle bit for bit."""
import numpy as np
import pytest
from orclib import Oracle, cases


def _energy(m, depth):
    knew = m.get_idx()[6]
    z = m.pull("ZETA", knew)[2:-2, 2:-2]
    ub, vb = m.pull("UBAR", knew), m.pull("VBAR", knew)
    ke = 0.5 * depth * (0.5 * (ub[2:-2, 2:-2] ** 2 + ub[2:-2, 3:-1] ** 2) + 0.5 * (vb[2:-2, 2:-2] ** 2 + vb[3:-1, 2:-2] ** 2))
    return (0.5 * 9.81 * z ** 2).sum() + ke.sum()


def test_gravity_wave_pulse_leaves_through_open_edges():
    """A free-surface bump in a resting homogeneous box sheds a gravity-wave ring (c = 31 m/s, 32 km to the edges).  With all four
    edges open less than 1 % of the energy is left after 30 steps (1800 s) and less than 1e-5 after 90; behind walls more than half
    stays; with only the eastern edge open the box drains more slowly."""
    left = {}
    for edges in ("", "WESN", "E"):
        cfg, st = cases.pulse_box(open_edges=edges)
        o = Oracle(cfg)
        cases.initialize(o, st)
        e0 = _energy(o, 100.0)
        r = []
        for n in range(90):
            o.call("step")
            if n in (29, 89):
                r.append(_energy(o, 100.0) / e0)
        o.close()
        left[edges] = r
    assert left["WESN"][0] < 1e-2 and left["WESN"][1] < 1e-5, left
    assert left[""][1] > 0.5, left
    assert left["WESN"][1] < left["E"][1] < left[""][1], left


def test_internal_wave_pulse_leaves_through_open_edges():
    """The 3-D conditions: a mode-1 displacement of the isotherms in a linearly stratified resting box sheds an internal-wave ring
    (0.44 m/s, 8 km to the edges).  After 1000 steps (60 000 s) the box with four open edges holds < 15 % of the temperature variance and
    < 20 % of the kinetic energy that the same box keeps behind walls (observed 6 % and 12 %: radiation plus upstream advection towards
    the resting profile is a good, not a perfect, absorber of internal waves -- as in the reference)."""
    left = {}
    for edges in ("", "WESN"):
        cfg, st, Tb = cases.strat_box(open_edges=edges)
        o = Oracle(cfg)
        cases.initialize(o, st)
        o.call("run", 1000)
        nnew = o.get_idx()[4]
        T, u, v = o.pull("T", nnew, 1)[:, 2:-2, 2:-2], o.pull("U", nnew)[:, 2:-2, 2:-2], o.pull("V", nnew)[:, 2:-2, 2:-2]
        assert np.isfinite(T).all()
        left[edges] = (((T - Tb[:, 2:-2, 2:-2]) ** 2).sum(), (u ** 2 + v ** 2).sum())
        o.close()
    assert left["WESN"][0] < 0.15 * left[""][0] and left["WESN"][1] < 0.20 * left[""][1], left


def _drive_wave(m, cfg, nsteps, a=0.01, period=1800.0, depth=100.0):
    """push the incoming wave as western boundary data before every step (what set_bry_all does on the host); returns zeta(x) of row 2 per step"""
    cg = np.sqrt(cfg.g / depth)
    ser = []
    for n in range(1, nsteps + 1):
        zw = a * np.sin(2 * np.pi / period * (n - 0.5) * cfg.dt)
        m.push("ZETA_WEST", np.full(cfg.Mm + 2, zw))
        m.push("UBAR_WEST", np.full(cfg.Mm + 2, cg * zw))
        m.push("U_WEST", np.full((cfg.N, cfg.Mm + 2), cg * zw))
        m.call("step")
        ser.append(m.pull("ZETA", m.get_idx()[6])[4, 2:-2].copy())
    return np.array(ser)


def test_incoming_wave_enters_through_the_boundary_data():
    """The forcing side of the Flather condition: a gravity wave prescribed as western boundary data (zeta_west = a sin(wt) with the
    matching ubar_west, period 1800 s, c = 31 m/s) must travel down the channel as a sin(w(t - x/c)) -- amplitude within 2 %, correlation
    > 0.998 at x = 10, 40 and 70 km once it has filled the 80 km -- and leave through the open eastern edge without building up."""
    cfg, st = cases.wave_channel()
    o = Oracle(cfg)
    cases.initialize(o, st)
    ser = _drive_wave(o, cfg, 120)
    o.close()
    a, period, c = 0.01, 1800.0, np.sqrt(9.81 * 100.0)
    x = 1.0e3 * (np.arange(1, cfg.Lm + 1) - 0.5)
    t = np.arange(1, 121) * cfg.dt
    late = t > 4000.0
    for i in (10, 40, 70):
        ref = a * np.sin(2 * np.pi / period * (t - x[i] / c))
        assert abs(ser[late, i].std() / ref[late].std() - 1.0) < 0.02
        assert np.corrcoef(ser[late, i], ref[late])[0, 1] > 0.998
    assert np.abs(ser[-30:]).max() < 1.02 * a


@pytest.mark.parametrize("kind", [0, 1, 2])
@pytest.mark.parametrize("U0,V0", [(0.05, 0.0), (0.0, -0.03), (0.04, 0.02)])
def test_uniform_throughflow_is_transparent(U0, V0, kind):
    """A uniform current over a flat bottom (f = 0, no drag) with the same current as boundary data is a steady state of the
    equations; the open edges -- characteristic condition, upstream advection of the tangential component, 3-D radiation,
    tracer advection, corner means -- must leave it untouched (observed: exactly).  kind 1 / 2: the reference's alternative conditions
    (OBC_M2SPECIFIED + OBC_M3SPECIFIED + OBC_TSPECIFIED; no switch = zero gradient) for all three variable families."""
    cfg, st = cases.pulse_box(Lm=32, Mm=24, N=4, open_edges="WESN")
    cfg.obc_m2 = cfg.obc_m3 = cfg.obc_t = kind
    sy, sx, N = cfg.Mm + 4, cfg.Lm + 4, cfg.N
    st["ZETA"] = {(1, 1): np.zeros((sy, sx)), (2, 1): np.zeros((sy, sx))}
    for nm, val in (("U", U0), ("V", V0)):
        st[nm + "BAR"] = {(1, 1): np.full((sy, sx), val), (2, 1): np.full((sy, sx), val)}
        st[nm] = {(1, 1): np.full((N, sy, sx), val), (2, 1): np.full((N, sy, sx), val)}
        for side, L in (("WEST", sy - 2), ("EAST", sy - 2), ("SOUTH", sx - 2), ("NORTH", sx - 2)):
            st[nm + "BAR_" + side] = np.full(L, val)
            st[nm + "_" + side] = np.full((N, L), val)
    o = Oracle(cfg)
    cases.initialize(o, st)
    for n in range(20):
        o.call("step")
    idx = o.get_idx()
    nnew, knew = idx[4], idx[6]
    ring = lambda a: a[..., 1:-1, 1:-1]
    assert np.abs(ring(o.pull("ZETA", knew))).max() <= 1e-13
    assert np.abs(ring(o.pull("UBAR", knew)) - U0).max() <= 1e-13 and np.abs(ring(o.pull("VBAR", knew)) - V0).max() <= 1e-13
    assert np.abs(ring(o.pull("U", nnew)) - U0).max() <= 1e-13 and np.abs(ring(o.pull("V", nnew)) - V0).max() <= 1e-13
    assert np.abs(ring(o.pull("T", nnew, 1)) - 10.0).max() <= 1e-12 and np.abs(ring(o.pull("T", nnew, 2)) - 35.0).max() <= 1e-12
    o.close()


def test_closed_flags_off_is_the_old_path():
    """obc_* = 0 must not change a closed basin: 5 steps of the walled seamount case with and without the (unused) boundary data."""
    res = []
    for push_bry in (False, True):
        cfg, st = cases.channel(Lm=24, Mm=20, N=6, ns_walls=True, ew_walls=True, seamount=True)
        if push_bry:
            st["ZETA_WEST"] = np.full(cfg.Mm + 2, 0.3)
            st["U_EAST"] = np.full((cfg.N, cfg.Mm + 2), 0.2)
        o = Oracle(cfg)
        cases.initialize(o, st)
        for n in range(5):
            o.call("step")
        res.append([o.pull("ZETA", o.get_idx()[6]), o.pull("U", o.get_idx()[4]), o.pull("T", o.get_idx()[4], 1)])
        o.close()
    for a, b in zip(*res):
        assert np.array_equal(a, b)


def test_baroclinic_open_box_runs_and_stays_bounded():
    """The stratified sheared test flow over a seamount in a box with four open edges clamped to the initial state: 60 steps stay
    finite and the kinetic energy stays within a factor of two of its initial value (no boundary instability)."""
    cfg, st = cases.channel(Lm=40, Mm=32, N=8, ns_walls=True, ew_walls=True, seamount=True, open_edges="WESN", dt=60.0, ndtfast=30, depth=500.0)
    o = Oracle(cfg)
    cases.initialize(o, st)
    ke0 = None
    for n in range(60):
        o.call("step")
        d = o.diag()
        assert np.all(np.isfinite(d))
        if n == 1:
            ke0 = d[0]
    assert 0.5 * ke0 < o.diag()[0] < 2.0 * ke0
    o.close()


# ---- GPU: libromsb200.so against the oracle, bit for bit ------------------------------------------------------------------------
def _same(g, o, cfg, steps_done):
    import parity_walk
    from test_gpu_parity import prognostic
    a, b = prognostic(g, cfg.NT), prognostic(o, cfg.NT)
    for f in ("ZT_AVG1", "DU_AVG1", "DV_AVG1", "DU_AVG2", "DV_AVG2", "DU_AVG_BAK", "DV_AVG_BAK", "FLXU", "FLXV", "HZ"):
        a[f], b[f] = g.pull(f), o.pull(f)
    worst = {}
    for k in b:
        x, y = parity_walk.view(a[k], cfg), parity_walk.view(b[k], cfg)
        assert np.isfinite(x).all(), (k, steps_done)
        worst[k] = parity_walk.rel(x, y)
    assert max(worst.values()) == 0.0, (steps_done, worst)


@pytest.mark.gpu
@pytest.mark.parametrize("ns,ew,topo", [(True, False, False), (False, True, True), (True, True, True)])
def test_boundary_kernels_reproduce_the_wall_code(ns, ew, topo, monkeypatch):
    """ROMS_BRY=1 routes the CLOSED edges of a basin through the general boundary kernels (k_step2d_bry, k_bry3d: obc.cuh) instead of
    the wall code inside the interior kernels; the results must not move -- this pins the recursion, the point lists, the zero-scratch
    rules and the flux accumulations of those kernels against code that is itself pinned by the reference's Pipes_ana log."""
    from romsb200 import Roms
    monkeypatch.setenv("ROMS_BRY", "1")
    cfg, st = cases.channel(70, 45, 12, ns_walls=ns, ew_walls=ew, seamount=topo, island=topo)
    g, o = Roms(cfg), Oracle(cfg)
    cases.initialize(g, st); cases.initialize(o, st)
    for n in (1, 2, 5):
        g.call("run", n); o.call("run", n)
        _same(g, o, cfg, n)
    g.close(); o.close()


@pytest.mark.gpu
@pytest.mark.parametrize("edges,ns,ew", [("E", False, True), ("WE", True, True), ("SN", True, False), ("WESN", True, True), ("WS", True, True), ("EN", True, True)])
def test_open_edges_gpu_equals_oracle(edges, ns, ew):
    """Open edges in every combination with periodic and closed ones (incl. the mixed corners: an open row beside a closed column),
    stratified sheared flow over a seamount with an island, exterior clamped to the initial state: GPU == oracle exactly after 1, 3 and
    43 steps, on the prognostic fields and on the fast-time averages the barotropic mode hands to the 3-D part."""
    from romsb200 import Roms
    cfg, st = cases.channel(56, 44, 10, ns_walls=ns, ew_walls=ew, seamount=True, island=True, open_edges=edges, dt=60.0, ndtfast=30, depth=500.0)
    g, o = Roms(cfg), Oracle(cfg)
    cases.initialize(g, st); cases.initialize(o, st)
    done = 0
    for n in (1, 2, 40):
        g.call("run", n); o.call("run", n)
        done += n
        _same(g, o, cfg, done)
    g.close(); o.close()


@pytest.mark.gpu
@pytest.mark.parametrize("m2,m3,mt", [(1, 1, 1), (2, 2, 2), (0, 1, 2), (1, 0, 0)])
def test_alternative_open_conditions_gpu_equals_oracle(m2, m3, mt):
    """OBC_M2SPECIFIED / OBC_M3SPECIFIED / OBC_TSPECIFIED and the zero-gradient defaults, mixed per variable family, on four open edges and
    on a west+south pair beside closed edges: GPU == oracle exactly."""
    from romsb200 import Roms
    for edges in ("WESN", "WS"):
        cfg, st = cases.channel(48, 40, 8, ns_walls=True, ew_walls=True, seamount=True, island=True, open_edges=edges, dt=60.0, ndtfast=30, depth=500.0)
        cfg.obc_m2, cfg.obc_m3, cfg.obc_t = m2, m3, mt
        g, o = Roms(cfg), Oracle(cfg)
        cases.initialize(g, st); cases.initialize(o, st)
        done = 0
        for n in (1, 9):
            g.call("run", n); o.call("run", n)
            done += n
            _same(g, o, cfg, done)
        g.close(); o.close()


@pytest.mark.gpu
def test_incoming_wave_on_gpu():
    """The boundary data pushed step by step (roms_push of the lines between steps) drive the CUDA path exactly like the oracle."""
    from romsb200 import Roms
    cfg, st = cases.wave_channel()
    g, o = Roms(cfg), Oracle(cfg)
    cases.initialize(g, st); cases.initialize(o, st)
    sg, so = _drive_wave(g, cfg, 40), _drive_wave(o, cfg, 40)
    assert np.array_equal(sg, so)
    _same(g, o, cfg, 40)
    g.close(); o.close()


@pytest.mark.gpu
def test_gravity_wave_pulse_leaves_on_gpu():
    """The radiation test through the CUDA path: same energy history as the oracle, < 1 % left after 30 steps."""
    from romsb200 import Roms
    cfg, st = cases.pulse_box(open_edges="WESN")
    g, o = Roms(cfg), Oracle(cfg)
    cases.initialize(g, st); cases.initialize(o, st)
    e0 = _energy(g, 100.0)
    g.call("run", 30); o.call("run", 30)
    assert _energy(g, 100.0) / e0 < 1e-2
    _same(g, o, cfg, 30)
    g.close(); o.close()


def test_boundary_lines_are_cut_like_the_planes():
    """take_block hands every rank the slice of the boundary data that belongs to its block: index q of a local line is the same point
    as index q + offset of the global one (lines run 0:Mm+1 / 0:Lm+1 like the reference's zeta_west(0:Mm+1), boundary.F:111-131)."""
    cfgG, stG = cases.channel(46, 37, 5, ns_walls=True, ew_walls=True, open_edges="WESN")
    for npx, npe in ((2, 1), (3, 2)):
        for rank in range(npx * npe):
            inode, jnode = rank % npx, rank // npx
            cfg, st = cases.take_block(cfgG, stG, (inode, jnode, npx, npe))
            (Lm, i0), (Mm, j0) = cases.block_extent(46, npx, inode), cases.block_extent(37, npe, jnode)
            assert (cfg.Lm, cfg.Mm, cfg.obc_west, cfg.obc_north, cfg.ubind) == (Lm, Mm, 1, 1, 0.1)
            assert np.array_equal(st["ZETA_WEST"], stG["ZETA_WEST"][j0:j0 + Mm + 2]) and st["ZETA_WEST"].shape == (Mm + 2,)
            assert np.array_equal(st["U_SOUTH"], stG["U_SOUTH"][:, i0:i0 + Lm + 2]) and st["U_SOUTH"].shape == (5, Lm + 2)
            assert np.array_equal(st["T_EAST"][(1, 2)], stG["T_EAST"][(1, 2)][:, j0:j0 + Mm + 2])
            # the line value at local q is the plane value at the boundary point of the same local index
            if inode == 0:
                assert np.array_equal(st["ZETA_WEST"], st["ZETA"][(1, 1)][1:-1, 1])
            if jnode == npe - 1:
                assert np.array_equal(st["V_NORTH"], st["V"][(1, 1)][:, Mm + 2, 1:-1])
