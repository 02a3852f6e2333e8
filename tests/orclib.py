"""ctypes binding of the CPU parity oracle (oracle/liboracle.so). Test infrastructure only."""
import ctypes as C
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "ucla-roms_b200"))
from romsb200.capi import RomsCfg, FIELDS, FID, CApi  # noqa: E402,F401
from romsb200 import cases  # noqa: E402,F401


def filament_cfg(Lm=64, Mm=64, N=32):
    """tests/Filament/{param.opt,benchmark.in} of the reference."""
    return cases.filament(Lm, Mm, N)[0]


def build_oracle(omp=False):
    name = "liboracle_omp.so" if omp else "liboracle.so"
    so = os.path.join(ROOT, "oracle", name)
    src = os.path.join(ROOT, "oracle", "roms_oracle.cpp")
    if (not os.path.exists(so)) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), name])
    return so


class Oracle(CApi):
    def __init__(self, cfg, omp=False):
        lib = C.CDLL(build_oracle(omp))
        h = C.c_void_p()
        assert lib.orc_create(C.byref(cfg), C.byref(h)) == 0
        super().__init__(lib, "orc_", cfg, h)

    def close(self):
        if self.h:
            self.lib.orc_destroy(self.h)
            self.h = None

    def set_block_halo_artifact(self, np_xi, np_eta):
        """Emulate the set-up artifact of the reference's OWN decomposed Rivers_ana run (oracle step2d): that run leaves
        h = 0 in the second halo line of each of its np_xi x np_eta MPI blocks, which halves Drhs(istr-2,j) / Drhs(i,jstr-2) of every block
        with a western / southern neighbour.  Columns and rows are the first owned ones of those blocks (mpi_setup.F:117-160)."""
        bi = [cases.block_extent(self.cfg.Lm, np_xi, n)[1] + 1 for n in range(1, np_xi)]
        bj = [cases.block_extent(self.cfg.Mm, np_eta, n)[1] + 1 for n in range(1, np_eta)]
        self._check(self.lib.orc_set_block_halo_artifact(self.h, len(bi), (C.c_int * max(len(bi), 1))(*bi), len(bj), (C.c_int * max(len(bj), 1))(*bj)),
                    "set_block_halo_artifact")
        return bi, bj

    def exchange(self, field, lvl=1, itrc=1):
        self.call("exchange", FID[field], lvl, itrc)
