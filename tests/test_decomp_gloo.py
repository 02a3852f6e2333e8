"""world_size-2 `gloo` test of the multi-GPU host logic: two CPU processes split a
periodic global field 2x1, build the eight halo messages from the library's geometry functions
(the same boxes and neighbour ranks the CUDA pack kernel and the ncclSend/ncclRecv group use),
swap them with torch.distributed, and must reproduce the global periodic halo exactly."""
import os
import socket
import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _worker(rank, world, port, npx, npe, uneven=False):
    import ctypes as C
    from romsb200 import load_library, cases
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lib = load_library()
    LG, MG = 12 * npx - (3 if uneven else 0), 10 * npe - (1 if uneven else 0)
    i_n, jn = rank % npx, rank // npx
    # mpi_setup.F:117-160: with `uneven` the first/last block of a direction is narrower than the others
    Lc, ic, Mc, jc = C.c_int(), C.c_int(), C.c_int(), C.c_int()
    assert lib.roms_decomp_extent(LG, npx, i_n, C.byref(Lc), C.byref(ic)) == 0 and lib.roms_decomp_extent(MG, npe, jn, C.byref(Mc), C.byref(jc)) == 0
    Lm, Mm, i0, j0 = Lc.value, Mc.value, ic.value, jc.value
    assert (Lm, i0) == cases.block_extent(LG, npx, i_n) and (Mm, j0) == cases.block_extent(MG, npe, jn)
    glob = np.arange((MG) * (LG), dtype=np.float64).reshape(MG, LG) + 0.5       # interior only, (j,i)
    loc = np.full((Mm + 4, Lm + 4), np.nan)
    loc[2:-2, 2:-2] = glob[j0:j0 + Mm, i0:i0 + Lm]
    reqs, recvs = [], []
    for d in range(8):
        o = lib.roms_decomp_opposite(d)
        nb = lib.roms_decomp_neighbor(npx, npe, i_n, jn, 1, 1, d)
        sb = (C.c_int * 4)(); lib.roms_decomp_box(Lm, Mm, d, 1, None, sb)
        msg = torch.from_numpy(np.ascontiguousarray(loc[sb[1] + 1:sb[1] + 1 + sb[3], sb[0] + 1:sb[0] + 1 + sb[2]]))
        src = lib.roms_decomp_neighbor(npx, npe, i_n, jn, 1, 1, o)
        rb = (C.c_int * 4)(); lib.roms_decomp_box(Lm, Mm, o, 0, None, rb)
        buf = torch.empty((rb[3], rb[2]), dtype=torch.float64)
        if nb == rank:
            buf.copy_(msg)
        else:
            reqs.append(dist.isend(msg, nb, tag=d))
            reqs.append(dist.irecv(buf, src, tag=d))
        recvs.append((rb, buf))
    for r in reqs:
        r.wait()
    for rb, buf in recvs:
        loc[rb[1] + 1:rb[1] + 1 + rb[3], rb[0] + 1:rb[0] + 1 + rb[2]] = buf.numpy()
    jj = (np.arange(-2, Mm + 2) + j0) % MG
    ii = (np.arange(-2, Lm + 2) + i0) % LG
    ref = glob[np.ix_(jj, ii)]
    assert np.array_equal(loc, ref), rank
    dist.destroy_process_group()


def _worker_closed(rank, world, port, npx, npe):
    """closed basin: no wrap, and messages along a wall also carry the boundary row / column (mpi_exchanges.F:538-543)"""
    import ctypes as C
    from romsb200 import load_library
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lib = load_library()
    Lm, Mm = 9, 7
    LG, MG = Lm * npx, Mm * npe
    i_n, jn = rank % npx, rank // npx
    G = np.arange((MG + 2) * (LG + 2), dtype=np.float64).reshape(MG + 2, LG + 2) + 0.25     # global rho-range 0..MG+1 x 0..LG+1
    walls = (C.c_int * 4)(int(jn == 0), int(jn == npe - 1), int(i_n == 0), int(i_n == npx - 1))     # S, N, W, E
    gi = np.arange(-1, Lm + 3) + i_n * Lm       # global index of local i = -1..Lm+2
    gj = np.arange(-1, Mm + 3) + jn * Mm
    own_i = (np.arange(-1, Lm + 3) >= (0 if walls[2] else 1)) & (np.arange(-1, Lm + 3) <= (Lm + 1 if walls[3] else Lm))
    own_j = (np.arange(-1, Mm + 3) >= (0 if walls[0] else 1)) & (np.arange(-1, Mm + 3) <= (Mm + 1 if walls[1] else Mm))
    val_i = (gi >= 0) & (gi <= LG + 1)
    val_j = (gj >= 0) & (gj <= MG + 1)
    ref = np.full((Mm + 4, Lm + 4), np.nan)
    ref[np.ix_(val_j, val_i)] = G[np.ix_(gj[val_j], gi[val_i])]
    loc = np.full((Mm + 4, Lm + 4), np.nan)
    loc[np.ix_(own_j, own_i)] = ref[np.ix_(own_j, own_i)]
    reqs, recvs = [], []
    expect = np.zeros((Mm + 4, Lm + 4), dtype=bool)
    expect[np.ix_(own_j, own_i)] = True
    for d in range(8):
        o = lib.roms_decomp_opposite(d)
        nb = lib.roms_decomp_neighbor(npx, npe, i_n, jn, 0, 0, d)
        src = lib.roms_decomp_neighbor(npx, npe, i_n, jn, 0, 0, o)
        if nb >= 0:
            sb = (C.c_int * 4)(); lib.roms_decomp_box(Lm, Mm, d, 1, walls, sb)
            msg = torch.from_numpy(np.ascontiguousarray(loc[sb[1] + 1:sb[1] + 1 + sb[3], sb[0] + 1:sb[0] + 1 + sb[2]]))
            assert not np.isnan(msg.numpy()).any()
            reqs.append(dist.isend(msg, nb, tag=d))
        if src >= 0:
            rb = (C.c_int * 4)(); lib.roms_decomp_box(Lm, Mm, o, 0, walls, rb)
            buf = torch.empty((rb[3], rb[2]), dtype=torch.float64)
            reqs.append(dist.irecv(buf, src, tag=d))
            recvs.append((rb, buf))
    for r in reqs:
        r.wait()
    for rb, buf in recvs:
        loc[rb[1] + 1:rb[1] + 1 + rb[3], rb[0] + 1:rb[0] + 1 + rb[2]] = buf.numpy()
        expect[rb[1] + 1:rb[1] + 1 + rb[3], rb[0] + 1:rb[0] + 1 + rb[2]] = True
    assert np.array_equal(loc[expect], ref[expect]), rank
    assert np.isnan(loc[~expect]).all()
    # every valid halo cell that has a neighbour on that side must have been filled
    need = np.zeros_like(expect)
    need[np.ix_(val_j, val_i)] = True
    need[0, :] &= not walls[0]; need[-1, :] &= not walls[1]; need[:, 0] &= not walls[2]; need[:, -1] &= not walls[3]
    assert (expect | ~need).all(), rank
    dist.destroy_process_group()


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def test_halo_messages_world2_gloo():
    mp.spawn(_worker, args=(2, _free_port(), 2, 1), nprocs=2, join=True)


def test_halo_messages_world4_gloo():
    mp.spawn(_worker, args=(4, _free_port(), 2, 2), nprocs=4, join=True)


def test_halo_messages_uneven_blocks_world4_gloo():
    mp.spawn(_worker, args=(4, _free_port(), 2, 2, True), nprocs=4, join=True)


def test_halo_messages_closed_basin_world4_gloo():
    mp.spawn(_worker_closed, args=(4, _free_port(), 2, 2), nprocs=4, join=True)
