import ctypes as C
import os
import numpy as np

PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class RomsCfg(C.Structure):
    """Mirror of `roms_cfg` (include/roms_b200.h)."""
    _fields_ = [("Lm", C.c_int), ("Mm", C.c_int), ("N", C.c_int), ("NT", C.c_int),
                ("ew_periodic", C.c_int), ("ns_periodic", C.c_int), ("nonlin_eos", C.c_int),
                ("salinity", C.c_int), ("ndtfast", C.c_int), ("np_xi", C.c_int), ("np_eta", C.c_int),
                ("inode", C.c_int), ("jnode", C.c_int),
                ("dt", C.c_double), ("rho0", C.c_double), ("g", C.c_double),
                ("Tcoef", C.c_double), ("T0", C.c_double), ("Scoef", C.c_double), ("S0", C.c_double),
                ("rdrg", C.c_double), ("rdrg2", C.c_double), ("Zob", C.c_double), ("gamma2", C.c_double),
                ("theta_s", C.c_double), ("theta_b", C.c_double), ("hc", C.c_double),
                ("uv_vis2", C.c_int), ("ts_dif2", C.c_int), ("visc2", C.c_double), ("tnu2", C.c_double * 40),
                ("obc_west", C.c_int), ("obc_east", C.c_int), ("obc_south", C.c_int), ("obc_north", C.c_int), ("ubind", C.c_double),
                ("obc_m2", C.c_int), ("obc_m3", C.c_int), ("obc_t", C.c_int)]


# enum roms_field, in the order of include/roms_b200.h
FIELDS = ["H", "PM", "PN", "F", "RMASK", "UMASK", "VMASK", "DM_U", "DN_U", "DM_V", "DN_V", "DM_R", "DN_R",
          "FOMN", "HINV", "ZETA", "UBAR", "VBAR", "RHOA", "RHOS", "RUFRC", "RVFRC", "ZT_AVG1", "DU_AVG1",
          "DV_AVG1", "DU_AVG2", "DV_AVG2", "DU_AVG_BAK", "DV_AVG_BAK", "R_D", "SUSTR", "SVSTR", "SWFLX",
          "STFLX", "U", "V", "T", "HZ", "Z_R", "Z_W", "FLXU", "FLXV", "WE", "WI", "RHO1", "QP1", "AKV", "AKT",
          "RU", "RV", "PMASK", "DM_P", "DN_P", "PMON_U", "PNOM_V", "VISC2_R", "VISC2_P", "DIFF2", "RIV_UFLX", "RIV_VFLX",
          "PTIDE", "PIPE_IDX", "PIPE_FLX", "SRFLX"] + \
         [v + "_" + s for v in ("ZETA", "UBAR", "VBAR", "U", "V", "T") for s in ("WEST", "EAST", "SOUTH", "NORTH")]
FID = {n: i for i, n in enumerate(FIELDS)}


def lib_path():
    # ROMS_B200_LIB: developer hook to A/B a differently tuned build of the same library
    return os.environ.get("ROMS_B200_LIB") or os.path.join(PKG, "libromsb200.so")


def load_library():
    p = lib_path()
    if not os.path.exists(p):
        raise RuntimeError(f"{p} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)")
    return C.CDLL(p)


class CApi:
    """Common ctypes surface: libromsb200.so exports `roms_*`; the test oracle exports the same
    functions with another prefix, so the parity tests drive both through this class."""

    def __init__(self, lib, prefix, cfg, handle):
        self.lib, self.p, self.cfg, self.h = lib, prefix, cfg, handle
        self.shape2 = (cfg.Mm + 4, cfg.Lm + 4)

    def _f(self, name):
        return getattr(self.lib, self.p + name)

    def count(self, field):
        fn = self._f("field_count")
        fn.restype = C.c_long
        return fn(self.h, FID[field])

    def shape(self, field):
        side = field.rsplit("_", 1)[-1]
        if side in ("WEST", "EAST", "SOUTH", "NORTH"):       # boundary data: a line along the edge (times N levels)
            L = (self.shape2[0] if side in ("WEST", "EAST") else self.shape2[1]) - 2     # zeta_west(0:Mm+1), ubar_south(0:Lm+1)
            n = self.count(field) // L
            return (L,) if n == 1 else (n, L)
        n = self.count(field) // (self.shape2[0] * self.shape2[1])
        return self.shape2 if n == 1 else (n,) + self.shape2

    def push(self, field, arr, lvl=1, itrc=1):
        a = np.ascontiguousarray(arr, dtype=np.float64)
        assert a.size == self.count(field), (field, a.shape, self.count(field))
        rc = self._f("push")(self.h, FID[field], lvl, itrc, a.ctypes.data_as(C.POINTER(C.c_double)))
        self._check(rc, "push " + field)

    def pull(self, field, lvl=1, itrc=1, out=None):
        a = np.empty(self.shape(field), dtype=np.float64) if out is None else out
        rc = self._f("pull")(self.h, FID[field], lvl, itrc, a.ctypes.data_as(C.POINTER(C.c_double)))
        self._check(rc, "pull " + field)
        return a

    def _check(self, rc, what):
        if rc != 0:
            msg = ""
            if self.p == "roms_":
                self.lib.roms_last_error.restype = C.c_char_p
                msg = self.lib.roms_last_error().decode()
            raise RuntimeError(f"{self.p}{what} failed with code {rc}: {msg}")

    def call(self, name, *args):
        self._check(self._f(name)(self.h, *args), name)

    def set_idx(self, iic, iif, nstp, nrhs, nnew, kstp, knew):
        self.call("set_time_indices", iic, iif, nstp, nrhs, nnew, kstp, knew)

    def get_idx(self):
        v = [C.c_int() for _ in range(7)]
        self._f("get_time_indices")(self.h, *[C.byref(x) for x in v])
        return tuple(x.value for x in v)

    def diag(self):
        out = (C.c_double * 4)()
        self._check(self._f("diag")(self.h, out), "diag")
        return np.array(out[:])

    def grid_stats(self):
        """setup_grid2.F:64-100: hmin, hmax, grdmin, grdmax, Cg_min, Cg_max, Cu_Cor, area, volume"""
        out = (C.c_double * 9)()
        self._check(self._f("grid_stats")(self.h, out), "grid_stats")
        return np.array(out[:])

    def hbl(self):
        """surface and bottom boundary layer thicknesses of the KPP closure (hbls, hbbl)"""
        a = np.empty(self.shape2); b = np.empty(self.shape2)
        self._check(self._f("get_hbl")(self.h, a.ctypes.data_as(C.POINTER(C.c_double)), b.ctypes.data_as(C.POINTER(C.c_double))), "get_hbl")
        return a, b

    def set_river(self, iriver, vol, trc):
        """river_frc.F set_river_frc: volume flux [m3/s] and tracer values of river `iriver` (1-based)"""
        a = (C.c_double * len(trc))(*trc)
        self._check(self._f("set_river")(self.h, iriver, C.c_double(vol), a), "set_river")

    def set_pipe(self, ipipe, prf, trc):
        """pipe_frc.F: vertical release profile (N values, bottom first) and tracer values of pipe `ipipe` (1-based)"""
        a = (C.c_double * len(prf))(*prf)
        b = (C.c_double * len(trc))(*trc)
        self._check(self._f("set_pipe")(self.h, ipipe, a, b), "set_pipe")

    def nfast(self):
        return self._f("get_nfast")(self.h)

    def weights(self, n):
        w1 = (C.c_double * n)()
        w2 = (C.c_double * n)()
        self._f("get_weights")(self.h, w1, w2, n)
        return np.array(w1[:]), np.array(w2[:])

    def scoord(self):
        N = self.cfg.N
        cw = (C.c_double * (N + 1))()
        cr = (C.c_double * N)()
        self._f("get_scoord")(self.h, cw, cr)
        return np.array(cw[:]), np.array(cr[:])

    # state transfer helpers ---------------------------------------------------------------
    def push_state(self, st):
        """st: dict name -> array, or name -> {(lvl,itrc): array}"""
        for name, v in st.items():
            if isinstance(v, dict):
                for (lvl, itrc), a in v.items():
                    self.push(name, a, lvl, itrc)
            else:
                self.push(name, v)


class Roms(CApi):
    """One GPU subdomain of the B200 time step (opaque `roms_ctx`)."""

    def __init__(self, cfg, device=0):
        lib = load_library()
        h = C.c_void_p()
        rc = lib.roms_create(C.byref(cfg), device, C.byref(h))
        if rc != 0:
            lib.roms_last_error.restype = C.c_char_p
            raise RuntimeError("roms_create failed: " + lib.roms_last_error().decode())
        super().__init__(lib, "roms_", cfg, h)

    def close(self):
        if self.h:
            self.lib.roms_destroy(self.h)
            self.h = None

    def sync(self):
        self.call("sync")

    def launches(self):
        n = C.c_long()
        self.lib.roms_kernel_launches(self.h, C.byref(n))
        return n.value

    def resident_bytes(self):
        n = C.c_longlong()
        self.lib.roms_resident_bytes(self.h, C.byref(n))
        return n.value

    def exchange(self, field, lvl=1, itrc=1):
        f = (C.c_int * 1)(FID[field]); l = (C.c_int * 1)(lvl); t = (C.c_int * 1)(itrc)
        self._check(self.lib.roms_exchange(self.h, 1, f, l, t), "exchange")

    def attach_torch_comm(self, dist):
        """multi-rank runs: hand the NCCL unique id (broadcast through torch.distributed) to the library"""
        uid = (C.c_char * 128)()
        if dist.get_rank() == 0:
            self._check(self.lib.roms_comm_unique_id(uid, 128), "comm_unique_id")
        import torch
        t = torch.frombuffer(bytearray(uid.raw), dtype=torch.uint8).cuda()
        dist.broadcast(t, 0)
        raw = bytes(t.cpu().numpy().tobytes())
        self._check(self.lib.roms_comm_attach(self.h, raw, 128, dist.get_rank(), dist.get_world_size()), "comm_attach")

    def alfabeta(self):
        a = np.empty(self.shape2); b = np.empty(self.shape2)
        self._check(self.lib.roms_alfabeta(self.h, a.ctypes.data_as(C.POINTER(C.c_double)), b.ctypes.data_as(C.POINTER(C.c_double))), "alfabeta")
        return a, b
