"""Walks one baroclinic step routine by routine (main.F:397-479 order) through the CUDA library and
the CPU oracle on the same state and reports, after every routine, the relative L2 / max deviation of
every field that routine writes.  Used by test_gpu_parity.py and runnable on its own for debugging:
    python tests/parity_walk.py [Lm Mm N]
"""
import sys
import numpy as np
from orclib import Oracle, cases
from romsb200 import Roms

STATE2 = ["ZETA", "UBAR", "VBAR"]
CHECK = {
    "rho_eos": ["RHO1", "QP1", "RHOA", "RHOS"],
    "set_huv": ["FLXU", "FLXV"],
    "omega": ["WE", "WI"],
    "prsgrd": ["RU", "RV"],
    "pre_step3d": ["T3", "U3", "V3", "Tindx", "Uindx", "Vindx", "R_D"],
    "set_huv1": ["U3", "V3", "FLXU", "FLXV"],
    "step3d_uv1": ["Unew", "Vnew", "RUFRC", "RVFRC"],
    "step2d": ["ZETAk", "UBARk", "VBARk", "ZT_AVG1", "DU_AVG1", "DV_AVG1", "DU_AVG2", "DV_AVG2", "DU_AVG_BAK", "DV_AVG_BAK", "RUFRC", "RVFRC"],
    "set_depth": ["HZ", "Z_R", "Z_W"],
    "step3d_uv2": ["Unew", "Vnew", "FLXU", "FLXV", "UBARk", "VBARk"],
    "step3d_t": ["Tnew"],
    "visc3d": ["Unew", "Vnew", "RUFRC", "RVFRC"],
    "t3dmix": ["Tnew"],
}


def rel(a, b):
    d = np.linalg.norm((a - b).ravel())
    n = np.linalg.norm(b.ravel())
    return d / n if n > 0 else d


def view(a, cfg):
    """rows/columns that exist in the model: beyond a closed wall the outermost halo line is allocated but never defined"""
    a = a if cfg.ns_periodic else a[..., 1:-1, :]
    return a if cfg.ew_periodic else a[..., :, 1:-1]


def fetch(m, name, NT):
    iic, iif, nstp, nrhs, nnew, kstp, knew = m.get_idx()
    indx = 3 - nstp
    if name == "T3":
        return np.stack([m.pull("T", 3, it) for it in range(1, NT + 1)])
    if name == "Tindx":
        return np.stack([m.pull("T", indx, it) for it in range(1, NT + 1)])
    if name == "Tnew":
        return np.stack([m.pull("T", nnew, it) for it in range(1, NT + 1)])
    if name in ("U3", "V3"):
        return m.pull(name[0], 3)
    if name in ("Uindx", "Vindx"):
        return m.pull(name[0], indx)
    if name in ("Unew", "Vnew"):
        return m.pull(name[0], nnew)
    if name in ("ZETAk", "UBARk", "VBARk"):
        return m.pull(name[:-1], knew)
    return m.pull(name)


def walk(cfg, st, nsteps=1, verbose=False, device=0):
    """returns dict routine -> worst relative L2 deviation over its outputs (max over steps/calls)"""
    o = Oracle(cfg)
    g = Roms(cfg, device)
    cases.initialize(o, st)
    cases.initialize(g, st)
    NT = cfg.NT
    worst = {}

    def both(name, *args, check=None):
        o.call(name, *args)
        g.call(name, *args)
        for f in CHECK[check or name]:
            a, b = view(fetch(g, f, NT), cfg), view(fetch(o, f, NT), cfg)
            assert np.isfinite(b).all(), ("oracle non-finite", name, f)
            r = rel(a, b) if np.isfinite(a).all() else np.inf
            worst[(check or name, f)] = max(worst.get((check or name, f), 0.0), r)
            if verbose:
                print(f"  {name:12s} {f:10s} relL2={r:.3e} max|d|={np.abs(a - b).max():.3e}")

    def setidx(**kw):
        for m in (o, g):
            i = dict(zip(["iic", "iif", "nstp", "nrhs", "nnew", "kstp", "knew"], m.get_idx()))
            i.update(kw)
            m.set_idx(i["iic"], i["iif"], i["nstp"], i["nrhs"], i["nnew"], i["kstp"], i["knew"])

    # init_step0 outputs
    for f in ["HZ", "Z_R", "Z_W", "FLXU", "FLXV", "WE", "WI", "RHO1", "RHOA", "RHOS", "DU_AVG1", "DV_AVG1", "HINV", "PMASK", "DM_P", "DN_P", "PMON_U", "PNOM_V",
              "VISC2_R", "VISC2_P"]:
        a, b = view(g.pull(f), cfg), view(o.pull(f), cfg)
        worst[("init_step0", f)] = rel(a, b)
        if verbose:
            print(f"  init_step0   {f:10s} relL2={worst[('init_step0', f)]:.3e}")
    nfast = o.nfast()
    assert g.nfast() == nfast
    knew = 1
    for step in range(1, nsteps + 1):
        nstp = 1 + (step - 1) % 2
        setidx(iic=step, nstp=nstp, nrhs=nstp, nnew=3)
        both("rho_eos", nstp)
        both("set_huv")
        both("omega")
        both("prsgrd")
        both("pre_step3d", 0)
        both("set_huv1", 0)
        setidx(nrhs=3, nnew=3 - nstp)
        both("omega")
        both("rho_eos", 3)
        both("prsgrd")
        both("step3d_uv1", 0)
        if cfg.uv_vis2:
            both("visc3d")
        for iif in range(1, nfast + 1):
            kstp = knew
            knew = kstp + 1 if kstp < 4 else 1
            setidx(iif=iif, kstp=kstp, knew=knew)
            both("step2d")
        both("set_depth", 0) if False else None
        for f in CHECK["set_depth"]:
            a, b = view(fetch(g, f, NT), cfg), view(fetch(o, f, NT), cfg)
            worst[("set_depth", f)] = max(worst.get(("set_depth", f), 0.0), rel(a, b))
        setidx(iif=nfast + 1)
        both("step3d_uv2", 0)
        both("omega")
        both("step3d_t", 0)
        if cfg.ts_dif2:
            both("t3dmix")
        both("rho_eos", 3 - nstp)
    o.close()
    g.close()
    return worst


if __name__ == "__main__":
    dims = [int(x) for x in sys.argv[1:4]] or [48, 40, 12]
    for tag, (cfg, st) in {"channel(nonlin EOS, T+S)": cases.channel(*dims), "channel(N/S walls)": cases.channel(*dims, ns_walls=True),
                             "closed basin": cases.channel(*dims, ns_walls=True, ew_walls=True), "filament(linear EOS, T)": cases.filament(*dims)}.items():
        print(tag)
        w = walk(cfg, st, nsteps=2, verbose=True)
        print(" worst:", max(w.values()), max(w, key=w.get))
