"""No Fortran compiler exists in this image, so the bind(C) interfaces of
fishman_b200/fortran/fishman_gpu_c.F90 are checked mechanically against
include/fishman_b200.h: same names, same argument counts, scalars by value where the C
prototype takes them by value, pointers by reference."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def c_protos():
    src = open(os.path.join(ROOT, "include", "fishman_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    out = {}
    for m in re.finditer(r"\b(fm_[a-z0-9_]+)\s*\(([^)]*)\)\s*;", src):
        args = [a.strip() for a in m.group(2).split(",") if a.strip() and a.strip() != "void"]
        out[m.group(1)] = ["*" in a or "(" in a for a in args]  # True = pointer
    return out


def f_interfaces():
    src = open(os.path.join(ROOT, "fishman_b200", "fortran", "fishman_gpu_c.F90")).read()
    src = re.sub(r"&\s*\n\s*", " ", src)
    out = {}
    pat = re.compile(r"(?:function|subroutine)\s+(\w+)\s*\(([^)]*)\)\s*bind\(C,\s*name='(\w+)'\)(.*?)end (?:function|subroutine)",
                     re.S | re.I)
    for m in pat.finditer(src):
        name, args, cname, body = m.group(1), [a.strip() for a in m.group(2).split(",")], m.group(3), m.group(4)
        assert name == cname
        byval = set()
        for line in body.splitlines():
            if "value" in line.split("::")[0]:
                byval |= {v.strip() for v in line.split("::")[1].split(",")}
        out[cname] = [a not in byval for a in args]  # True = by reference (pointer)
    return out


def test_fortran_interfaces_match_header():
    c, f = c_protos(), f_interfaces()
    assert len(f) >= 18
    for name, fargs in f.items():
        assert name in c, f"{name} bound in Fortran but not declared in the header"
        cargs = c[name]
        assert len(cargs) == len(fargs), f"{name}: {len(fargs)} Fortran args vs {len(cargs)} C args"
        for k, (cp, fp) in enumerate(zip(cargs, fargs)):
            # a C pointer-typed handle passed by value (type(c_ptr), value) is fine: skip arg 0 handles
            if k == 0 and not fp and cp:
                continue
            assert cp == fp, f"{name} arg {k}: C pointer={cp}, Fortran by-reference={fp}"
    core = {"fm_create", "fm_create_lonlat", "fm_destroy", "fm_last_error", "fm_triangulate", "fm_counts",
            "fm_get_triangles", "fm_voronoi", "fm_get_circumcentres", "fm_get_rings", "fm_get_edges", "fm_lloyd",
            "fm_get_generators"}
    assert core <= set(f)
