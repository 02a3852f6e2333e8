"""Multi-GPU parity: run under torch.distributed.run with WORLD_SIZE in {2,4,8}.  Every rank steps its
block of the analytic channel through libromsb200 (NCCL halo swaps) and compares the block with the
single-domain CPU oracle solution of the whole domain.  Exits non-zero on mismatch."""
import os
import sys
import numpy as np
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from orclib import Oracle, cases  # noqa: E402
from romsb200 import Roms  # noqa: E402


def main():
    nsteps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
    mode = sys.argv[2] if len(sys.argv) > 2 else "periodic"
    walls = mode in ("walls", "seamount", "closed", "closedmix", "unevenclosed", "river")
    topo = mode in ("seamount", "closed", "closedmix", "unevenclosed")           # + bottom topography and a land mask
    ew = mode in ("closed", "closedmix", "unevenclosed", "river")                # closed basin: walls on all four sides
    mix = dict(visc2=100.0, tnu2=50.0, sponge=True) if mode == "closedmix" else {}   # + visc3d / t3dmix
    if mode == "river":                                                              # + a river source whose faces straddle the block boundary
        mix = dict(river=True, dt=60.0)
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    npx, npe = cases.rank_grid(world)
    if mode == "kpp":        # the reference's Rivers_ana case with the KPP closure on, cut into blocks
        return main_kpp(nsteps, rank, world, local, npx, npe)
    if mode == "reflog":     # the reference's own Rivers_ana run: 3x2 blocks, h = 0 in the second halo line of every block
        return main_reflog(nsteps, rank, world, local)
    if mode == "obc":        # four open edges (Flather / Orlanski, exterior clamped to the initial state), cut into blocks
        return main_obc(nsteps, rank, world, local, npx, npe)
    LG, MG, N = 24 * npx, 20 * npe, 10
    if mode.startswith("uneven"):       # blocks of unequal size (mpi_setup.F:117-160): 3 points short in xi, 1 in eta
        LG, MG = LG - (3 if npx > 1 else 0), MG - (1 if npe > 1 else 0)
    cfg, st = cases.channel(LG, MG, N, ns_walls=walls, ew_walls=ew, seamount=topo, island=topo, sub=(rank % npx, rank // npx, npx, npe), **mix)
    g = Roms(cfg, local)
    g.attach_torch_comm(dist)
    cases.initialize(g, st)
    g.call("run", nsteps)
    dg = g.diag()
    cfgG, stG = cases.channel(LG, MG, N, ns_walls=walls, ew_walls=ew, seamount=topo, island=topo, **mix)
    o = Oracle(cfgG)
    cases.initialize(o, stG)
    o.call("run", nsteps)
    do = o.diag()
    iic, iif, nstp, nrhs, nnew, kstp, knew = g.get_idx()
    Lm, Mm = cfg.Lm, cfg.Mm
    i0, j0 = cases.block_extent(LG, npx, rank % npx)[1], cases.block_extent(MG, npe, rank // npx)[1]
    worst = 0.0
    for f, l, it in [("ZETA", knew, 1), ("UBAR", knew, 1), ("VBAR", knew, 1), ("U", nnew, 1), ("V", nnew, 1), ("T", nnew, 1), ("T", nnew, 2),
                     ("RHO1", 1, 1), ("HZ", 1, 1), ("WE", 1, 1), ("FLXU", 1, 1)]:
        a = g.pull(f, l, it)
        b = o.pull(f, l, it)[..., j0:j0 + Mm + 4, i0:i0 + Lm + 4]
        if walls:      # rows beyond a closed wall are allocated but undefined
            ja, jb = (1 if rank // npx == 0 else 0), (-1 if rank // npx == npe - 1 else None)
            a, b = a[..., ja:jb, :], b[..., ja:jb, :]
        if ew:
            ia, ib = (1 if rank % npx == 0 else 0), (-1 if rank % npx == npx - 1 else None)
            a, b = a[..., :, ia:ib], b[..., :, ia:ib]
        r = np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-300)
        worst = max(worst, r)
    ok = worst <= 1e-12 and np.allclose(dg, do, rtol=1e-11, atol=1e-18)
    print(f"rank {rank}/{world}: worst relative L2 vs single-domain oracle after {nsteps} steps = {worst:.2e}; diag gpu {dg} oracle {do}", flush=True)
    t = torch.tensor([0 if ok else 1], device="cuda")
    dist.all_reduce(t)
    g.close()
    dist.destroy_process_group()
    sys.exit(int(t.item() != 0))


def main_kpp(nsteps, rank, world, local, npx, npe):
    cfgG, stG = cases.rivers_ana(kpp=True)
    cfg, st = cases.take_block(cfgG, stG, (rank % npx, rank // npx, npx, npe))
    g = Roms(cfg, local)
    g.attach_torch_comm(dist)
    cases.initialize(g, st)
    g.call("run", nsteps)
    dg = g.diag()
    o = Oracle(cfgG)
    cases.initialize(o, stG)
    o.call("run", nsteps)
    do = o.diag()
    iic, iif, nstp, nrhs, nnew, kstp, knew = g.get_idx()
    Lm, Mm = cfg.Lm, cfg.Mm
    i0, j0 = cases.block_extent(cfgG.Lm, npx, rank % npx)[1], cases.block_extent(cfgG.Mm, npe, rank // npx)[1]
    worst = 0.0
    for f, l, it in [("ZETA", knew, 1), ("UBAR", knew, 1), ("VBAR", knew, 1), ("U", nnew, 1), ("V", nnew, 1), ("T", nnew, 1), ("T", nnew, 2), ("AKV", 1, 1), ("AKT", 1, 2)]:
        a = g.pull(f, l, it)[..., 2:-2, 2:-2]
        b = o.pull(f, l, it)[..., j0 + 2:j0 + Mm + 2, i0 + 2:i0 + Lm + 2]
        worst = max(worst, np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-300))
    ok = worst <= 1e-11 and np.allclose(dg, do, rtol=1e-10, atol=1e-18)
    print(f"rank {rank}/{world}: KPP case, worst relative L2 vs single-domain oracle after {nsteps} steps = {worst:.2e}; diag gpu {dg} oracle {do}", flush=True)
    t = torch.tensor([0 if ok else 1], device="cuda")
    dist.all_reduce(t)
    g.close()
    dist.destroy_process_group()
    sys.exit(int(t.item() != 0))


def main_reflog(nsteps, rank, world, local):
    """Six ranks decomposed exactly like the run that wrote tests/Rivers_ana/benchmark.result_github_gnu (NP_XI=3, NP_ETA=2,
    tests/Rivers_ana/param.opt), each block given the bathymetry the way that run had it: tests/Rivers_ana/ana_grid.h fills 0:nx+1 x
    0:ny+1, the second halo line keeps init = 0 and an analytic grid is never exchanged (grid.F:444-446).  With that input the CUDA
    path must reproduce the reference's log itself -- all 21 lines x 4 columns -- not just the oracle."""
    import json
    assert world == 6, "the reference's Rivers_ana log was written by 6 ranks"
    npx, npe = 3, 2
    cfgG, stG = cases.rivers_ana(kpp=True)
    cfg, st = cases.take_block(cfgG, stG, (rank % npx, rank // npx, npx, npe))
    h = st["H"].copy()
    h[0, :] = 0.0; h[-1, :] = 0.0; h[:, 0] = 0.0; h[:, -1] = 0.0
    st["H"] = h
    g = Roms(cfg, local)
    g.attach_torch_comm(dist)
    cases.initialize(g, st)
    G = json.load(open(os.path.join(HERE, "golden", "rivers_ana_diag.json")))["diag"]
    worst = np.zeros(4)
    rows = []
    for n in range(0, min(nsteps, 20) + 1):
        if n:
            g.call("step")
        d = g.diag()
        ref = np.array([G[n]["KE"], G[n]["BAROTR_KE"], G[n]["MAX_ADV_CFL"], G[n]["MAX_VERT_CFL"]])
        dev = np.where(ref != 0, np.abs(d - ref) / np.where(ref != 0, np.abs(ref), 1.0), np.abs(d))
        worst = np.maximum(worst, dev)
        rows.append((n, d.tolist(), dev.tolist()))
    ok = worst.max() <= 1e-11
    if rank == 0:
        for n, d, dev in rows:
            print(f"line {n:2d}: " + " ".join(f"{x:.16e}" for x in d) + "   rel.dev " + " ".join(f"{x:.1e}" for x in dev), flush=True)
        print(f"Rivers_ana log on 3x2 GPU blocks with the reference's halo of h: worst relative deviation per column {worst}  -> {'OK' if ok else 'MISMATCH'}", flush=True)
    t = torch.tensor([0 if ok else 1], device="cuda")
    dist.all_reduce(t)
    g.close()
    dist.destroy_process_group()
    sys.exit(int(t.item() != 0))


def main_obc(nsteps, rank, world, local, npx, npe):
    cfgG, stG = cases.channel(28 * npx, 22 * npe, 10, ns_walls=True, ew_walls=True, seamount=True, island=True, open_edges="WESN", dt=60.0, ndtfast=30, depth=500.0)
    cfg, st = cases.take_block(cfgG, stG, (rank % npx, rank // npx, npx, npe))
    g = Roms(cfg, local)
    g.attach_torch_comm(dist)
    cases.initialize(g, st)
    g.call("run", nsteps)
    dg = g.diag()
    o = Oracle(cfgG)
    cases.initialize(o, stG)
    o.call("run", nsteps)
    do = o.diag()
    iic, iif, nstp, nrhs, nnew, kstp, knew = g.get_idx()
    Lm, Mm = cfg.Lm, cfg.Mm
    i0, j0 = cases.block_extent(cfgG.Lm, npx, rank % npx)[1], cases.block_extent(cfgG.Mm, npe, rank // npx)[1]
    worst = 0.0
    for f, l, it in [("ZETA", knew, 1), ("UBAR", knew, 1), ("VBAR", knew, 1), ("U", nnew, 1), ("V", nnew, 1), ("T", nnew, 1), ("T", nnew, 2), ("DU_AVG1", 1, 1), ("DV_AVG2", 1, 1),
                     ("FLXU", 1, 1)]:
        # the block's interior plus the boundary points of the open edges it touches (halo values of the averages are not exchanged)
        ia, ib = (1 if rank % npx == 0 else 2), (Lm + 3 if rank % npx == npx - 1 else Lm + 2)
        ja, jb = (1 if rank // npx == 0 else 2), (Mm + 3 if rank // npx == npe - 1 else Mm + 2)
        a = g.pull(f, l, it)[..., ja:jb, ia:ib]
        b = o.pull(f, l, it)[..., j0 + ja:j0 + jb, i0 + ia:i0 + ib]
        worst = max(worst, np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-300))
    ok = worst <= 1e-12 and np.allclose(dg, do, rtol=1e-11, atol=1e-18)
    print(f"rank {rank}/{world}: open-boundary case, worst relative L2 vs single-domain oracle after {nsteps} steps = {worst:.2e}; diag gpu {dg} oracle {do}", flush=True)
    t = torch.tensor([0 if ok else 1], device="cuda")
    dist.all_reduce(t)
    g.close()
    dist.destroy_process_group()
    sys.exit(int(t.item() != 0))


if __name__ == "__main__":
    main()
