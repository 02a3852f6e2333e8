#!/usr/bin/env python
"""bench.py -- baroclinic grid-point updates/s of the B200-native UCLA-ROMS time step.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--size Lm Mm N] [--nt NT]

One "step" = one full baroclinic step (main.F:397-479: rho_eos, set_HUV, omega, prsgrd, pre_step3d,
set_HUV1, omega, rho_eos, prsgrd, step3d_uv1, nfast x step2d, step3d_uv2, omega, step3d_t, rho_eos)
of the analytic baroclinic channel (BASELINE.json configs[1]: 1024x1024x64 per GPU, T+S, nonlinear
split EOS, 82 executed barotropic substeps).  Inputs are far larger than L2 (15 GB of state per GPU),
so no L2 flush is needed between timed iterations.  Timing: CUDA events on the library's stream,
barrier + synchronize on both sides, max over ranks.  Prints ONE JSON line on rank 0.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "ucla-roms_b200"))

METRIC = "baroclinic grid-point updates/s"
UNIT = "grid-point updates/s"

# Algorithmic fp64 words per point per launch (SURVEY.md section 8(a)/(d) rule: every 3-D operand of a
# kernel read once, every 3-D result written once, 2-D operands ignored in 3-D kernels, scratch = 0).
# (words, dimensionality)
KERNEL_WORDS = {
    "k_rho_eos": lambda t: (6 if t[0] == "true" else 4, 3), "k_set_huv": (5, 3), "k_omega": (6, 3),
    "k_prsgrd_vert": lambda t: (5 if t[0] == "true" else 3, 3),
    "k_rho_prs_vert": lambda t: (8 if t[0] == "true" else 6, 3),   # rho_eos fused with the vertical part of prsgrd: T,S,Hz,z_r read; rho1,qp1,P,rho written
    "k_prsgrd_horiz": lambda t: (8 if t[0] == "true" else 6, 3), "k_prsgrd_tma": (8, 3),
    "k_pre_flxdiv": (5, 3), "k_rhs_uv_horiz": (9, 3), "k_rhs_uv_tma": (9, 3),
    "k_pre_t_horiz_tma": lambda t: (13, 3), "k_t_horiz_tma": lambda t: (8, 3),     # TB = 2 in every launch of the TMA kernels (3+5*2, 2+3*2); TB = 1 launches over-count
    # tracer kernels, TB tracers per launch: operands shared by the batch count once
    "k_pre_t_horiz": lambda t: (3 + 5 * int(t[3]), 3),    # FlxU,FlxV,FlxDiv + per tracer t(nrhs),t(nstp),t(indx) R, t(nnew),t(indx) W
    "k_t_horiz": lambda t: (2 + 3 * int(t[3]), 3),        # FlxU,FlxV + per tracer t(nrhs) R, t(nnew) RW
    "k_t_vert": lambda t: ((4 if t[0] == "0" else 3) + (1 if t[2] == "true" else int(t[1])) + 3 * int(t[1]), 3),   # Hz,We,Wi[,FlxDiv] + Akt slots + per tracer t(nrhs) R, t(nnew) RW
    "k_uv_vert": lambda t: (10 if t[0] == "0" else 9, 3),
    "k_set_huv1": (4, 3), "k_uv2_pass0": (3, 3), "k_uv2_pass1": (6, 3), "k_uv2_pass1_sm": (6, 3),
    "k_uv2_sm": (6, 3),      # step3d_uv2 passes 0+1 fused: Hz, u(nnew), u(nstp), Flx read; u(nnew), Flx written
    # KPP closure (row f2): 3-D operands read once, results written once
    "k_lmd_rig": (5, 3), "k_lmd_smooth": (2, 3), "k_lmd_vmix_col": (5, 3), "k_lmd_kpp1": (8, 3), "k_lmd_kpp2": (9, 3),
    "k_set_depth": (3, 3), "k_step2d_zeta": (24, 2), "k_step2d_uv": (23, 2), "k_rd_bott_drag": (3, 2),
    "k_visc3d": (7, 3),      # row f1: u,v(nstp),Hz read; u,v(3-nstp) read+written
    "k_t3dmix": (4, 3),      # row f1: t(nrhs),Hz read; t(nnew) read+written
    "k_step2d_tma": (34, 2), "k_step2d_march": (34, 2), "k_step2d_hyb": (34, 2),
    "k_step2d": (34, 2),     # 9 state R + 3 W + 12 static R + 10 averages RW per 2-D point and substep
}


def kernel_words(name):
    """(words, dim) of a kernel from its stringified launch name, e.g. '(k_prsgrd_horiz<true, false>)'; None if not a compute kernel."""
    n = name.replace("(", "").replace(")", "").replace(" ", "")
    base, _, targs = n.partition("<")
    w = KERNEL_WORDS.get(base)
    if w is None:
        return None
    return w(targs.rstrip(">").split(",")) if callable(w) else w


def step_bytes_per_update(NT, nfast, N):
    """B = 8*(113 + 7*NT + 34*nfast/N) bytes per baroclinic grid-point update."""
    return 8.0 * (113 + 7 * NT + 34.0 * nfast / N)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel` from the committed `ncu --set full` capture of the channel
    workload (profiles/r2_ncu_traffic.json, made by tools/ncu_summary.py from the raw pages of that capture); None if the kernel is not in it."""
    p = os.path.join(ROOT, "profiles", "r2_ncu_traffic.json")
    if not os.path.exists(p):
        return None
    norm = lambda n: n.replace("(", "").replace(")", "").replace(" ", "").replace("true", "1").replace("false", "0")
    want = norm(kernel)
    for k, v in json.load(open(p)).items():
        if norm(k) == want or norm(k).split("<")[0] == want.split("<")[0]:
            return v["dram_bytes"]
    return None


class ClockSampler:
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu):
        self.gpu, self.f, self.p = gpu, None, None

    def start(self):
        try:
            self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "50"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.p.terminate()
        self.p.wait()
        self.f.flush()
        rows = [l.strip().split(", ") for l in open(self.f.name) if l.strip()]
        os.unlink(self.f.name)
        sm = sorted(int(float(r[1])) for r in rows if len(r) >= 9)
        reasons = set()
        for r in rows:
            if len(r) < 9:
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[5:9]):
                if v.strip().lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": int(float(rows[0][2])) if rows else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# Workloads = BASELINE.json configs[1..4] as per-GPU blocks (weak scaling: the block is fixed, the domain grows with the rank grid
# 1x1, 2x1, 2x2, 4x2).  `total_at` = the rank count at which the assembled domain is the configuration BASELINE.json names.
WORKLOADS = {
    "channel": dict(size=(1024, 1024, 64), nt=2, kw=dict(ns_walls=True), cpu_sample=(256, 256, 64), total_at=1,
                    what="analytic baroclinic channel, T+S, nonlinear split EOS, E-W periodic, closed free-slip N/S walls"),
    "seamount": dict(size=(1024, 2048, 100), nt=2, kw=dict(ns_walls=True, ew_walls=True, seamount=True), cpu_sample=(128, 256, 100), total_at=8,
                     what="analytic seamount h=4500(1-0.5exp(-r^2/(L/5)^2)) in a closed basin, T+S, nonlinear split EOS (4096x4096x100 on 8 GPUs)"),
    "tracers": dict(size=(1024, 2048, 64), nt=32, kw=dict(ns_walls=True, ew_walls=True, river=True, dt=60.0), cpu_sample=(96, 128, 64), total_at=2,
                    what="Rivers_ana geometry (land strip, river channel and river source) in a closed basin, T+S+30 passive tracers (2048x2048x64 on 2 GPUs)"),
    "regional": dict(size=(1024, 1024, 64), nt=2, kw=dict(ns_walls=True, ew_walls=True, seamount=True, open_edges="WESN"), cpu_sample=(256, 256, 64), total_at=1,
                     what="regional box with four OPEN edges (Flather / Orlanski radiation, exterior clamped to the initial state) over the analytic seamount, T+S, nonlinear split EOS; "
                          "not a BASELINE.json config -- the cost of SURVEY.md 8(f) row f3 (single GPU)"),
    "weak": dict(size=(2304, 2304, 100), nt=2, kw=dict(ns_walls=True, ew_walls=True, seamount=True), cpu_sample=(128, 256, 100), total_at=8,
                 what="largest block that fits 180 GB: analytic seamount in a closed basin, T+S (9216x4608x100 on 8 GPUs; 8192x8192x100 needs 1.8 TB)"),
}


def rel_l2(a, b):
    import numpy as np
    n = float(np.linalg.norm(b.ravel()))
    d = float(np.linalg.norm((a - b).ravel()))
    return d / n if n > 0 else d


def parity_check(case_kw, NT, N, rank, world, local, dist):
    """Run before the timed region, so that every bench line carries its own parity evidence:
    (a) rank 0: two baroclinic steps of a 96x64xN block of THIS workload's physics (same N, same NT up to 8) through the CUDA
        library against the CPU oracle, worst relative L2 over zeta, ubar, vbar, u, v, every tracer and rho;
    (b) N > 1 ranks: a small domain cut into the rank grid, stepped with the halo swaps, against the single-domain oracle."""
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from orclib import Oracle
    from romsb200 import Roms, cases
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
    out = {"tolerance": "relative L2 <= 1e-12 after one step (north_star)"}
    nt = min(NT, 8)
    kw = dict(case_kw, NT=nt)

    def fields(m):
        idx = m.get_idx()
        nnew, knew = idx[4], idx[6]
        return [("zeta", m.pull("ZETA", knew)), ("ubar", m.pull("UBAR", knew)), ("vbar", m.pull("VBAR", knew)), ("u", m.pull("U", nnew)),
                ("v", m.pull("V", nnew)), ("rho", m.pull("RHO1"))] + [(f"t{it}", m.pull("T", nnew, it)) for it in range(1, nt + 1)]

    def trim(a, cfg, r=None):
        ns = slice(1 if (not cfg.ns_periodic and cfg.jnode == 0) else 0, -1 if (not cfg.ns_periodic and cfg.jnode == cfg.np_eta - 1) else None)
        ew = slice(1 if (not cfg.ew_periodic and cfg.inode == 0) else 0, -1 if (not cfg.ew_periodic and cfg.inode == cfg.np_xi - 1) else None)
        return a[..., ns, ew]
    if rank == 0:
        cfg, st = cases.channel(96, 64, N, **kw)
        g, o = Roms(cfg, local), Oracle(cfg)
        cases.initialize(g, st); cases.initialize(o, st)
        g.call("run", 2); o.call("run", 2)
        out["single_domain"] = {"sample": f"96x64x{N}, {nt} tracers, 2 steps", "worst_rel_l2": max(rel_l2(trim(a, cfg), trim(b, cfg)) for (_, a), (_, b) in zip(fields(g), fields(o)))}
        g.close(); o.close()
    if world > 1:
        npx, npe = cases.rank_grid(world)
        LG, MG, Ns = 32 * npx, 28 * npe, min(N, 16)
        sub = (rank % npx, rank // npx, npx, npe)
        cfg, st = cases.channel(LG, MG, Ns, sub=sub, **kw)
        g = Roms(cfg, local)
        g.attach_torch_comm(dist)
        cases.initialize(g, st)
        g.call("run", 2)
        cfgG, stG = cases.channel(LG, MG, Ns, **kw)
        o = Oracle(cfgG)
        cases.initialize(o, stG)
        o.call("run", 2)
        i0, j0 = cases.block_extent(LG, npx, sub[0])[1], cases.block_extent(MG, npe, sub[1])[1]
        w = max(rel_l2(trim(a, cfg), trim(b[..., j0:j0 + cfg.Mm + 4, i0:i0 + cfg.Lm + 4], cfg)) for (_, a), (_, b) in zip(fields(g), fields(o)))
        g.close(); o.close()
        import torch
        t = torch.tensor([w], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out["decomposed_vs_single_domain"] = {"sample": f"{LG}x{MG}x{Ns} over {npx}x{npe} ranks, 2 steps", "worst_rel_l2": float(t.item())}
    return out


def cpu_oracle_rate(cfg_kw, steps, warmup, sample):
    """Times the CPU restatement of the path (oracle/, OpenMP over the reference's independent j/k loops)
    on all host cores for a bounded sample of the workload; returns (updates/s, cores, seconds, sample str)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from orclib import Oracle
    from romsb200 import cases
    cores = os.cpu_count() or 1
    os.environ["OMP_NUM_THREADS"] = str(cores)   # torchrun presets 1; the baseline uses every host core
    Lm, Mm, N = sample
    cfg, st = cases.channel(Lm, Mm, N, **cfg_kw)
    o = Oracle(cfg, omp=True)
    cases.initialize(o, st)
    for _ in range(warmup):
        o.call("step")
    t0 = time.perf_counter()
    for _ in range(steps):
        o.call("step")
    dt = time.perf_counter() - t0
    o.close()
    return Lm * Mm * N * steps / dt, cores, dt, f"{Lm}x{Mm}x{N} block of the same workload ({cfg.NT} tracers), {steps} baroclinic steps after {warmup} warm-up, {cores} OpenMP threads"


_STDOUT_FD = None


def _quiet_stdout():
    """Libraries (NCCL's version banner) write to fd 1; the contract is ONE JSON line there.  Everything but the final
    line goes to stderr."""
    global _STDOUT_FD
    sys.stdout.flush()
    _STDOUT_FD = os.dup(1)
    os.dup2(2, 1)


def _emit(line):
    sys.stdout.flush()
    if _STDOUT_FD is not None:
        os.dup2(_STDOUT_FD, 1)
    print(json.dumps(line), flush=True)


def main():
    _quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="channel", choices=sorted(WORKLOADS), help="BASELINE.json configs[1..4] as per-GPU blocks; the default is the configuration the metric is quoted on")
    ap.add_argument("--size", type=int, nargs=3, default=None, metavar=("Lm", "Mm", "N"), help="per-GPU block (default: the workload's)")
    ap.add_argument("--nt", type=int, default=None)
    ap.add_argument("--cpu-sample", type=int, nargs=3, default=None)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the GPU-vs-oracle check that precedes the timed region")
    ap.add_argument("--periodic", action="store_true", help="doubly periodic variant instead of the closed N/S walls of BASELINE configs[1]")
    ap.add_argument("--mixing", action="store_true", help="also run visc3d / t3dmix (UV_VIS2, TS_DIF2; SURVEY.md 8(f) row f1) inside every step")
    ap.add_argument("--kpp", action="store_true", help="also run the LMD/KPP vertical mixing closure (lmd_vmix twice per step; SURVEY.md 8(f) row f2)")
    ap.add_argument("--kernel-table", default=None, help="write the per-kernel timing table to this file")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    wl = WORKLOADS[a.workload]
    if a.workload == "regional" and world > 1:
        raise SystemExit("bench.py: the 'regional' (open-boundary) workload is built single-domain; decomposed open boundaries are covered by tests/mgpu_parity.py obc")
    Lm, Mm, N = a.size or wl["size"]
    NT = a.nt or wl["nt"]
    cpu_sample = tuple(a.cpu_sample or wl["cpu_sample"])
    case_kw = dict(wl["kw"], NT=NT, nonlin_eos=True)
    if a.periodic:
        case_kw.update(ns_walls=False, ew_walls=False)
    if a.mixing:
        case_kw.update(visc2=100.0, tnu2=50.0)
    workload = f"{wl['what']}; {Lm}x{Mm}x{N} per GPU, {NT} tracers" + (", doubly periodic variant" if a.periodic else "") + \
               ", ndtfast=60 (82 executed barotropic substeps per step)" + \
               (", + harmonic lateral viscosity/diffusion (visc3d, t3dmix)" if a.mixing else "") + (", + KPP vertical mixing (lmd_vmix x2 per step)" if a.kpp else "")
    config = {"workload": workload, "name": a.workload, "grid_per_gpu": [Lm, Mm, N], "tracers": NT,
              "l2_policy": "inputs (the resident state, see resident_gb_per_gpu) far larger than the 126 MB L2; no flush",
              "timing": "CUDA events on the library stream, barrier+sync both sides, max over ranks"}

    if a.impl == "reference":
        # The reference Fortran cannot be built in this image (no gfortran/MPI/netCDF): the arm times the CPU
        # restatement of its algorithm (oracle/) with all host threads on a bounded sample of the workload.
        if rank != 0:
            return 0
        steps = max(1, min(a.steps, 3))
        rate, cores, secs, sample = cpu_oracle_rate(case_kw, steps, min(a.warmup, 1), cpu_sample)
        config["cpu_sample"] = sample
        line = {"impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": a.gpus, "steps": steps, "warmup": min(a.warmup, 1),
                "ms_per_step": 1e3 * secs / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": config,
                "cpu_baseline": {"value": rate, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
                "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        _emit(line)
        return 0

    import numpy as np
    import torch
    import torch.distributed as dist
    from romsb200 import Roms, cases, FID
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    torch.cuda.set_device(local)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    parity = None if a.no_parity else parity_check(case_kw, NT, N, rank, world, local, dist)
    np_xi, np_eta = cases.rank_grid(world)
    cfg, st = cases.channel_strips(Lm * np_xi, Mm * np_eta, N, strips=(8 if Lm * Mm * N > (1 << 26) else 1), sub=(rank % np_xi, rank // np_xi, np_xi, np_eta), **case_kw)
    config["domain"] = [Lm * np_xi, Mm * np_eta, N]
    config["rank_grid"] = [np_xi, np_eta]
    if a.kpp:
        cfg.Zob = 1.0e-2          # the bottom boundary layer of the closure needs a roughness length (log-layer drag then, too)
    g = Roms(cfg, local)
    if world > 1:
        g.attach_torch_comm(dist)
    if a.kpp:
        st["KPP"] = True
    cases.initialize(g, st)
    del st
    lib = g.lib
    config["resident_gb_per_gpu"] = round(g.resident_bytes() / 1e9, 2)
    nfast = g.nfast()

    def timed_steps(k):
        barrier()
        lib.roms_timer_start(g.h)
        g.call("run", k)
        ms = C.c_double()
        lib.roms_timer_stop(g.h, C.byref(ms))
        barrier()
        t = torch.tensor([ms.value], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    g.call("run", a.warmup)
    g.sync()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = g.launches()
    ms = timed_steps(a.steps)
    launches = g.launches() - l0
    clocks = sampler.stop() if rank == 0 else None
    d = g.diag()
    assert np.isfinite(d).all(), d
    updates = float(Lm) * Mm * N * world * a.steps
    value = updates / (ms * 1e-3)

    # ---- e2e: the same steps driven the way the Fortran host drives them: every step the surface forcing comes
    # from pinned HOST memory (set_forces, main.F:386) and the free surface / barotropic velocities go back to the host.
    sy, sx = Mm + 4, Lm + 4
    pin_in = [torch.zeros((sy, sx), dtype=torch.float64).pin_memory() for _ in range(3 + NT)]
    pin_out = [torch.empty((sy, sx), dtype=torch.float64).pin_memory() for _ in range(3)]
    in_fields = [("SUSTR", 1, 1), ("SVSTR", 1, 1), ("SWFLX", 1, 1)] + [("STFLX", 1, it) for it in range(1, NT + 1)]
    dptr = C.POINTER(C.c_double)

    def e2e_step():
        for t_, (f, l, it) in zip(pin_in, in_fields):
            lib.roms_push_async(g.h, FID[f], l, it, C.cast(t_.data_ptr(), dptr))
        g.call("step")
        knew = g.get_idx()[6]
        for t_, f in zip(pin_out, ("ZETA", "UBAR", "VBAR")):
            lib.roms_pull_async(g.h, FID[f], knew, 1, C.cast(t_.data_ptr(), dptr))
        g.sync()      # one synchronisation per step: the forcing buffers are free again, the output planes are valid
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    lib.roms_timer_start(g.h)
    for _ in range(a.steps):
        e2e_step()
    ems = C.c_double()
    lib.roms_timer_stop(g.h, C.byref(ems))
    barrier()
    e2e_wall = time.perf_counter() - t0
    te = torch.tensor([max(ems.value, 0.0), e2e_wall * 1e3], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_ms = float(te.max().item())
    plane = sy * sx * 8
    e2e = {"value": updates / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": len(pin_in) * plane, "d2h_bytes_per_step": len(pin_out) * plane,
           "what": "per step: H2D of sustr,svstr,swflx,stflx(1:NT) from pinned host (roms_push_async), roms_step, D2H of zeta,ubar,vbar(knew) (roms_pull_async), one roms_sync"}

    # ---- roofline of the dominant kernel: a second pass of the same K steps with every launch bracketed by its own events
    peak, peak_src = peaks()
    roof = None
    table = []
    # (every rank runs the pass -- the halo swaps are collective -- rank 0 reports)
    lib.roms_kernel_timing(g.h, 1)
    g.call("run", a.steps)
    g.sync()
    if rank != 0:
        lib.roms_kernel_timing(g.h, 0)
    if rank == 0:
        buf = C.create_string_buffer(1 << 16)
        lib.roms_kernel_times(g.h, buf, len(buf))
        lib.roms_kernel_timing(g.h, 0)
        tot = 0.0
        for ln in buf.value.decode().splitlines():
            name, cnt, tms = ln.rsplit(" ", 2)
            table.append((name, int(cnt), float(tms)))
            tot += float(tms)
        table.sort(key=lambda r: -r[2])
        top = next((r for r in table if kernel_words(r[0])), table[0])   # dominant compute kernel (exchange_xxx is reported in the table)
        words, dim = kernel_words(top[0]) or (None, 3)
        npts = Lm * Mm * (N if dim == 3 else 1)
        avg_ms = top[2] / top[1]
        if words:
            ach = words * 8.0 * npts / (avg_ms * 1e-3) / 1e9
            roof = {"bound": "hbm", "kernel": top[0], "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                    "traffic": ncu_traffic(top[0]) if (Lm, Mm, N) == (1024, 1024, 64) else None,
                    "peak_source": peak_src, "algorithmic_bytes_per_launch": words * 8.0 * npts, "avg_launch_ms": avg_ms,
                    "launches_per_step": top[1] / a.steps, "share_of_step": top[2] / tot,
                    "traffic_source": "profiles/r2_ncu_traffic.json (ncu --set full, dram read+write bytes per launch, this workload)"}
        if a.kernel_table:
            with open(a.kernel_table, "w") as f:
                f.write(f"# per-kernel CUDA-event times, {a.steps} steps, {workload}\n# kernel launches total_ms avg_ms share algorithmic_GB/s\n")
                for name, cnt, tms in table:
                    w, dm = kernel_words(name) or (0, 3)
                    gbs = w * 8.0 * Lm * Mm * (N if dm == 3 else 1) / (tms / cnt * 1e-3) / 1e9 if w else 0.0
                    f.write(f"{name.replace(' ', '')} {cnt} {tms:.3f} {tms / cnt:.4f} {tms / tot:.4f} {gbs:.1f}\n")
    B = step_bytes_per_update(NT, nfast, N)
    step_roof = {"bytes_per_update": B, "achieved": value * B / 1e9 / world, "peak": peak, "unit": "GB/s per GPU", "frac": value * B / 1e9 / world / peak}

    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        rate, cores, secs, sample = cpu_oracle_rate(case_kw, 2, 1, cpu_sample)
        cpu = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
    g.close()
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms / a.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
                "roofline": roof, "step_roofline": step_roof, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks, "parity": parity,
                "diag": {"KE": d[0], "BAROTR_KE": d[1], "MAX_ADV_CFL": d[2], "MAX_VERT_CFL": d[3]}}
        _emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
