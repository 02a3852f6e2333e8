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


def _worker(rank, world, port, npx, npe):
    import ctypes as C
    from romsb200 import load_library
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lib = load_library()
    Lm, Mm = 12, 10
    LG, MG = Lm * npx, Mm * npe
    i_n, jn = rank % npx, rank // npx
    glob = np.arange((MG) * (LG), dtype=np.float64).reshape(MG, LG) + 0.5       # interior only, (j,i)
    loc = np.full((Mm + 4, Lm + 4), np.nan)
    loc[2:-2, 2:-2] = glob[jn * Mm:(jn + 1) * Mm, i_n * Lm:(i_n + 1) * Lm]
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
    jj = (np.arange(-2, Mm + 2) + jn * Mm) % MG
    ii = (np.arange(-2, Lm + 2) + i_n * Lm) % LG
    ref = glob[np.ix_(jj, ii)]
    assert np.array_equal(loc, ref), rank
    dist.destroy_process_group()


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def test_halo_messages_world2_gloo():
    mp.spawn(_worker, args=(2, _free_port(), 2, 1), nprocs=2, join=True)


def test_halo_messages_world4_gloo():
    mp.spawn(_worker, args=(4, _free_port(), 2, 2), nprocs=4, join=True)
