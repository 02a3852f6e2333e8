"""CPU-side checks of the drop-in boundary: libromsb200.so loads and exports every symbol that
include/roms_b200.h declares; the ctypes mirror of roms_cfg matches the C struct; the host-only
decomposition geometry (mpi_setup.F neighbours, mpi_exchanges.F message boxes) is self-consistent.
No compute calls are made (there is no GPU here)."""
import ctypes as C
import os
import re
import numpy as np
from orclib import ROOT
from romsb200 import load_library, RomsCfg, FIELDS


def header():
    return open(os.path.join(ROOT, "include", "roms_b200.h")).read()


def test_exports_every_declared_symbol():
    lib = load_library()
    names = sorted(set(re.findall(r"\b(roms_[a-z0-9_]+)\s*\(", header())))
    assert len(names) >= 35
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_field_enum_matches_python_mirror():
    body = re.search(r"enum roms_field \{(.*?)ROMS_NFIELDS", header(), re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    ids = [x.strip().split("=")[0].strip()[5:] for x in body.split(",") if x.strip()]
    assert ids == FIELDS


def test_cfg_struct_layout():
    # 13 ints (padded to 56) + 14 doubles + 2 ints + visc2 + tnu2[ROMS_MAXNT=40] + 4 ints (obc_*) + ubind + 3 ints, natural alignment
    assert C.sizeof(RomsCfg) == 56 + 14 * 8 + 8 + 8 + 40 * 8 + 16 + 8 + 16          # + obc_m2, obc_m3, obc_t (+4 padding)
    assert RomsCfg.obc_west.offset == 56 + 14 * 8 + 8 + 8 + 40 * 8 and RomsCfg.ubind.offset == RomsCfg.obc_west.offset + 16
    assert RomsCfg.uv_vis2.offset == 56 + 14 * 8 and RomsCfg.visc2.offset == 56 + 14 * 8 + 8


def test_no_cpu_fallback_error_path():
    lib = load_library()
    cfg = RomsCfg(Lm=8, Mm=8, N=4, NT=1, ndtfast=10, np_xi=1, np_eta=1, dt=1.0)
    h = C.c_void_p()
    rc = lib.roms_create(C.byref(cfg), 0, C.byref(h))
    import torch
    if not torch.cuda.is_available():
        assert rc != 0          # fails loudly without a device
        lib.roms_last_error.restype = C.c_char_p
        assert lib.roms_last_error()


def test_decomp_geometry():
    lib = load_library()
    for (npx, npe, ew, ns) in [(1, 1, 1, 1), (2, 1, 1, 1), (2, 2, 1, 0), (4, 2, 0, 0), (4, 2, 1, 1)]:
        for jn in range(npe):
            for i_n in range(npx):
                me = jn * npx + i_n
                for d in range(8):
                    nb = lib.roms_decomp_neighbor(npx, npe, i_n, jn, ew, ns, d)
                    o = lib.roms_decomp_opposite(d)
                    if nb >= 0:
                        # my neighbour in direction d sees me in the opposite direction
                        assert lib.roms_decomp_neighbor(npx, npe, nb % npx, nb // npx, ew, ns, o) == me
                    sb = (C.c_int * 4)(); rb = (C.c_int * 4)()
                    lib.roms_decomp_box(16, 12, d, 1, None, sb); lib.roms_decomp_box(16, 12, o, 0, None, rb)
                    assert sb[2] == rb[2] and sb[3] == rb[3]      # message extents agree
    # the 8 receive boxes tile the halo exactly once
    Lm, Mm = 10, 7
    cover = np.zeros((Mm + 4, Lm + 4), int)
    for d in range(8):
        b = (C.c_int * 4)(); lib.roms_decomp_box(Lm, Mm, d, 0, None, b)
        cover[b[1] + 1:b[1] + 1 + b[3], b[0] + 1:b[0] + 1 + b[2]] += 1
    assert (cover[2:-2, 2:-2] == 0).all() and cover.sum() == cover.size - Lm * Mm and cover.max() == 1


def test_uneven_block_extents():
    """mpi_setup.F:117-160: blocks of ceil(LL/np) points, surplus trimmed from the first and last block; the C export and the
    Python mirror used by the analytic cases agree, blocks tile the global range without gaps."""
    import ctypes as C
    from romsb200 import load_library, cases
    lib = load_library()
    for LL in (7, 45, 64, 100, 1023, 4096):
        for np_ in (1, 2, 3, 4, 8):
            nxt = 0
            for node in range(np_):
                L, c = C.c_int(), C.c_int()
                rc = lib.roms_decomp_extent(LL, np_, node, C.byref(L), C.byref(c))
                assert (L.value, c.value) == cases.block_extent(LL, np_, node)
                assert rc == (0 if L.value >= 2 else 1)
                assert c.value == nxt
                nxt += L.value
            assert nxt == LL
    # the reference's example geometry: LLm=45 on 2 ranks -> Lm=23 per block, off=1: first keeps 23, last gets 22
    assert cases.block_extent(45, 2, 0) == (23, 0) and cases.block_extent(45, 2, 1) == (22, 23)
    assert lib.roms_decomp_extent(45, 2, 2, None, None) != 0
