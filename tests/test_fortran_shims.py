"""The committed Fortran binding (fortran/roms_b200_mod.F90, roms_b200_shims.F90) cannot be compiled here (no Fortran compiler in
the image): check it textually against the C header and the built library -- every bind(C) name is declared and exported, the
argument counts agree, the field enumeration has the header's order, and every reference subroutine main.F:397-479 calls has a shim."""
import ctypes as C
import os
import re
from orclib import ROOT
from romsb200 import load_library, FIELDS

MOD = open(os.path.join(ROOT, "fortran", "roms_b200_mod.F90")).read()
SHIMS = open(os.path.join(ROOT, "fortran", "roms_b200_shims.F90")).read()
HDR = open(os.path.join(ROOT, "include", "roms_b200.h")).read()


def test_bind_names_exist_with_matching_argument_counts():
    lib = load_library()
    decl = {m.group(1): m.group(2) for m in re.finditer(r"\b(roms_\w+)\s*\(([^;{]*?)\)\s*;", HDR)}
    binds = re.findall(r"function\s+\w+\s*\(([^)]*)\)\s*bind\(C,\s*name='(\w+)'\)", MOD)
    assert len(binds) >= 30
    for args, name in binds:
        assert name in decl, name
        getattr(lib, name)                                   # exported
        nf = len([a for a in args.split(",") if a.strip()])
        nc = len([a for a in decl[name].split(",") if a.strip() and a.strip() != "void"])
        assert nf == nc, (name, nf, nc)


def test_field_enum_order():
    names = re.findall(r"ROMS_[A-Z0-9_]+", MOD[MOD.index("enum, bind(C)"):MOD.index("end enum")])
    assert names == ["ROMS_" + f for f in FIELDS] + ["ROMS_NFIELDS"]


def test_every_driver_call_has_a_shim():
    subs = {s.lower() for s in re.findall(r"^subroutine\s+(\w+)", SHIMS, flags=re.M)}
    for s in ("step2d", "pre_step3d", "prsgrd", "rho_eos", "step3d_uv1", "step3d_uv2", "step3d_t", "omega", "set_depth", "set_huv", "set_huv1"):
        assert s in subs, s
    used = set(re.findall(r"\b(roms_\w+_c)\b", SHIMS))
    declared = set(re.findall(r"function\s+(roms_\w+_c)\b", MOD))
    assert used <= declared, used - declared


def test_boundary_data_shim_pushes_every_line():
    """roms_b200_push_bry hands all 4 x 6 boundary arrays of boundary.F to the library under the same OBC_* switches, with the field ids
    of the header and the reference's own lower bound 0."""
    body = SHIMS[SHIMS.index("subroutine roms_b200_push_bry"):SHIMS.index("end subroutine roms_b200_push_bry")]
    for side in ("WEST", "EAST", "SOUTH", "NORTH"):
        blk = body[body.index("#ifdef OBC_" + side):]
        blk = blk[:blk.index("#endif")]
        for var in ("ZETA", "UBAR", "VBAR", "U", "V", "T"):
            fid = f"ROMS_{var}_{side}"
            assert fid[5:] in FIELDS and fid in blk, fid
            assert f"{var.lower()}_{side.lower()}(0" in blk
    assert "roms_push_async" in MOD

