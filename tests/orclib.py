"""ctypes binding of the CPU parity oracle (oracle/liboracle.so). Test infrastructure only."""
import ctypes as C
import os
import subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class RomsCfg(C.Structure):
    """Mirror of `roms_cfg` in include/roms_b200.h."""
    _fields_ = [("Lm", C.c_int), ("Mm", C.c_int), ("N", C.c_int), ("NT", C.c_int),
                ("ew_periodic", C.c_int), ("ns_periodic", C.c_int), ("nonlin_eos", C.c_int),
                ("salinity", C.c_int), ("ndtfast", C.c_int), ("np_xi", C.c_int), ("np_eta", C.c_int),
                ("inode", C.c_int), ("jnode", C.c_int),
                ("dt", C.c_double), ("rho0", C.c_double), ("g", C.c_double),
                ("Tcoef", C.c_double), ("T0", C.c_double), ("Scoef", C.c_double), ("S0", C.c_double),
                ("rdrg", C.c_double), ("rdrg2", C.c_double), ("Zob", C.c_double), ("gamma2", C.c_double),
                ("theta_s", C.c_double), ("theta_b", C.c_double), ("hc", C.c_double)]


# enum roms_field, in the order of include/roms_b200.h
FIELDS = ["H", "PM", "PN", "F", "RMASK", "UMASK", "VMASK", "DM_U", "DN_U", "DM_V", "DN_V", "DM_R", "DN_R",
          "FOMN", "HINV", "ZETA", "UBAR", "VBAR", "RHOA", "RHOS", "RUFRC", "RVFRC", "ZT_AVG1", "DU_AVG1",
          "DV_AVG1", "DU_AVG2", "DV_AVG2", "DU_AVG_BAK", "DV_AVG_BAK", "R_D", "SUSTR", "SVSTR", "SWFLX",
          "STFLX", "U", "V", "T", "HZ", "Z_R", "Z_W", "FLXU", "FLXV", "WE", "WI", "RHO1", "QP1", "AKV", "AKT",
          "RU", "RV"]
FID = {n: i for i, n in enumerate(FIELDS)}


def filament_cfg(Lm=64, Mm=64, N=32):
    """tests/Filament/{param.opt,benchmark.in} of the reference."""
    return RomsCfg(Lm=Lm, Mm=Mm, N=N, NT=1, ew_periodic=1, ns_periodic=1, nonlin_eos=0, salinity=0,
                   ndtfast=60, np_xi=1, np_eta=1, inode=0, jnode=0, dt=5.0, rho0=1000.0, g=9.81,
                   Tcoef=0.20, T0=1.0, Scoef=0.822, S0=1.0, rdrg=0.0, rdrg2=1.0e-3, Zob=1.0e-2,
                   gamma2=1.0, theta_s=6.0, theta_b=2.0, hc=25.0)


def build_oracle():
    so = os.path.join(ROOT, "oracle", "liboracle.so")
    src = os.path.join(ROOT, "oracle", "roms_oracle.cpp")
    if (not os.path.exists(so)) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "liboracle.so"])
    return so


class _Lib:
    """Common ctypes surface of liboracle.so (prefix orc_) and libromsb200.so (prefix roms_)."""

    def __init__(self, lib, prefix, cfg, handle):
        self.lib, self.p, self.cfg, self.h = lib, prefix, cfg, handle
        self.shape2 = (cfg.Mm + 4, cfg.Lm + 4)

    def _f(self, name):
        return getattr(self.lib, self.p + name)

    def count(self, field):
        fn = self._f("field_count"); fn.restype = C.c_long
        return fn(self.h, FID[field])

    def shape(self, field):
        n = self.count(field) // (self.shape2[0] * self.shape2[1])
        return self.shape2 if n == 1 else (n,) + self.shape2

    def push(self, field, arr, lvl=1, itrc=1):
        a = np.ascontiguousarray(arr, dtype=np.float64)
        assert a.size == self.count(field), (field, a.shape, self.count(field))
        rc = self._f("push")(self.h, FID[field], lvl, itrc, a.ctypes.data_as(C.POINTER(C.c_double)))
        assert rc == 0, (field, rc)

    def pull(self, field, lvl=1, itrc=1):
        a = np.empty(self.shape(field), dtype=np.float64)
        rc = self._f("pull")(self.h, FID[field], lvl, itrc, a.ctypes.data_as(C.POINTER(C.c_double)))
        assert rc == 0, (field, rc)
        return a

    def call(self, name, *args):
        rc = self._f(name)(self.h, *args)
        assert rc == 0, (name, rc)

    def set_idx(self, iic, iif, nstp, nrhs, nnew, kstp, knew):
        self.call("set_time_indices", iic, iif, nstp, nrhs, nnew, kstp, knew)

    def get_idx(self):
        v = [C.c_int() for _ in range(7)]
        self._f("get_time_indices")(self.h, *[C.byref(x) for x in v])
        return tuple(x.value for x in v)

    def diag(self):
        out = (C.c_double * 4)()
        rc = self._f("diag")(self.h, out)
        assert rc == 0
        return np.array(out[:])

    def nfast(self):
        return self._f("get_nfast")(self.h)

    def weights(self, n):
        w1 = (C.c_double * n)(); w2 = (C.c_double * n)()
        self._f("get_weights")(self.h, w1, w2, n)
        return np.array(w1[:]), np.array(w2[:])

    def scoord(self):
        N = self.cfg.N
        cw = (C.c_double * (N + 1))(); cr = (C.c_double * N)()
        self._f("get_scoord")(self.h, cw, cr)
        return np.array(cw[:]), np.array(cr[:])


class Oracle(_Lib):
    def __init__(self, cfg, omp=False):
        so = build_oracle()
        if omp:
            so = os.path.join(ROOT, "oracle", "liboracle_omp.so")
        lib = C.CDLL(so)
        h = C.c_void_p()
        assert lib.orc_create(C.byref(cfg), C.byref(h)) == 0
        super().__init__(lib, "orc_", cfg, h)

    def close(self):
        if self.h:
            self.lib.orc_destroy(self.h); self.h = None

    def exchange(self, field, lvl=1, itrc=1):
        self.call("exchange", FID[field], lvl, itrc)
