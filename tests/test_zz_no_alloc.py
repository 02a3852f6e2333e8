"""roms_omega_noalloc (include/roms_b200.h): the allocation-free variant of the omega kernel, whose column scratch is a static
device array.  It has no counterpart in the reference; the check is against the
existing entry point roms_omega on the same inputs, through the C ABI, on two calls in a row (the second on another state, so
that the scratch left behind by the first call is seen not to matter).  Collected last on purpose (file name)."""
import ctypes as C
import numpy as np
import pytest
from orclib import cases
from romsb200 import Roms

pytestmark = pytest.mark.gpu
TOL = 1e-12     # the one-step tolerance of tests/test_gpu_parity.py; the arithmetic is the same, so 0.0 is expected


def _rel(a, b):
    n = np.linalg.norm(b.ravel())
    d = np.linalg.norm((a - b).ravel())
    return d / n if n > 0 else d


def test_omega_noalloc_matches_roms_omega_twice():
    import torch
    cfg, st = cases.channel(48, 40, 12, ns_walls=True, seamount=True)
    Lm, Mm, N = cfg.Lm, cfg.Mm, cfg.N
    g = Roms(cfg, 0)
    cases.initialize(g, st)
    sy, sx = Mm + 4, Lm + 4
    x = np.arange(sx)[None, :] / Lm + np.zeros((sy, 1))
    g.push("SWFLX", 2.0e-7 * np.cos(2 * np.pi * x))          # so that the surface-flux term of omega.F is exercised
    fn = g.lib.roms_omega_noalloc
    dp = C.POINTER(C.c_double)
    fn.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double] + [dp] * 9 + [C.c_void_p]
    fn.restype = C.c_int
    ptr = lambda t: C.cast(t.data_ptr(), dp)
    worst = []
    for call in range(2):
        g.call("step")                                         # another state for each call
        g.call("omega")                                        # existing entry point: nrhs = 3 after a step -> dtau = dt (omega.F:50-58)
        g.sync()
        ref = {f: g.pull(f) for f in ("WE", "WI")}
        dev = {f: torch.from_numpy(g.pull(f)).cuda() for f in ("FLXU", "FLXV", "Z_W", "HZ", "PM", "PN", "SWFLX")}
        We = torch.full((N + 1, sy, sx), float("nan"), dtype=torch.float64, device="cuda")
        Wi = torch.full_like(We, float("nan"))
        torch.cuda.synchronize()
        rc = fn(Lm, Mm, N, cfg.dt, ptr(dev["FLXU"]), ptr(dev["FLXV"]), ptr(dev["Z_W"]), ptr(dev["HZ"]), ptr(dev["PM"]), ptr(dev["PN"]),
                ptr(dev["SWFLX"]), ptr(We), ptr(Wi), None)
        assert rc == 0, rc
        torch.cuda.synchronize()
        for name, out in (("WE", We), ("WI", Wi)):
            a, b = out.cpu().numpy()[:, 2:-2, 2:-2], ref[name][:, 2:-2, 2:-2]
            assert np.isfinite(a).all(), name
            assert np.abs(b).max() > 0.0, name
            worst.append(_rel(a, b))
            assert worst[-1] <= TOL, (call, name, worst[-1])
    print("roms_omega_noalloc vs roms_omega, relative L2 (We, Wi) x 2 calls:", worst)
    # beyond the static scratch: a clear error, nothing launched
    g.lib.roms_last_error.restype = C.c_char_p
    rc = fn(2048, 2048, 64, cfg.dt, ptr(dev["FLXU"]), ptr(dev["FLXV"]), ptr(dev["Z_W"]), ptr(dev["HZ"]), ptr(dev["PM"]), ptr(dev["PN"]),
            ptr(dev["SWFLX"]), ptr(We), ptr(Wi), None)
    assert rc == -3 and b"exceeds the static scratch" in g.lib.roms_last_error()
    g.close()
