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


def _split_args(text):
    out, depth, cur = [], 0, ""
    for ch in text:
        if ch == "(":
            depth += 1
        if ch == ")":
            depth -= 1
        if ch == "," and depth == 0:
            out.append(cur.strip())
            cur = ""
        else:
            cur += ch
    if cur.strip():
        out.append(cur.strip())
    return out


def _module_source():
    src = open(os.path.join(ROOT, "fishman_b200", "fortran", "delaunay_voronoi_gpu.F90")).read()
    src = "\n".join(ln.split("!")[0] if not ln.lstrip().startswith("#") else "" for ln in src.splitlines())
    return re.sub(r"&\s*\n\s*&?", " ", src)


def test_replacement_module_calls_match_the_header():
    """fishman_b200/fortran/delaunay_voronoi_gpu.F90 (the replacement delaunay_voronoi_mod) cannot be
    compiled here: check mechanically that every fm_* it calls is bound in fishman_gpu_c.F90 and
    declared in the header with that many arguments."""
    src = _module_source()
    c, f = c_protos(), f_interfaces()
    calls = {}
    for m in re.finditer(r"\b(fm_[a-z0-9_]+)\s*\(", src):
        name, i, depth = m.group(1), m.end(), 1
        j = i
        while depth:
            depth += {"(": 1, ")": -1}.get(src[j], 0)
            j += 1
        calls.setdefault(name, set()).add(len(_split_args(src[i:j - 1])))
    helpers = {"fm_error_message": 1}          # Fortran helper of fishman_gpu_c, not a C entry point
    used = set(calls) - set(helpers) - {"fm_density", "fm_quality_stats"}
    assert {"fm_create", "fm_create_lonlat", "fm_triangulate", "fm_voronoi", "fm_counts", "fm_get_triangles", "fm_get_rings",
            "fm_get_circumcentres", "fm_get_edges", "fm_mpas", "fm_get_mpas_cells", "fm_get_mpas_vertices", "fm_lloyd",
            "fm_get_generators", "fm_retriangulate", "fm_destroy"} <= used
    for name in used:
        assert name in c and name in f, f"{name} is called but not declared / bound"
        for nargs in calls[name]:
            assert nargs == len(c[name]), f"{name} called with {nargs} arguments, the header has {len(c[name])}"
    for name, n in helpers.items():
        assert calls.get(name, {n}) == {n}


def test_replacement_module_keeps_the_reference_interface():
    """The public names of DV:15-30 and the argument lists of the four procedures (DV:188-195, :226,
    :358-360, :587-589) and of output_interface (DV:169-181), hard-coded here from the reference (and
    cross-checked against it when /root/reference is mounted)."""
    src = _module_source()
    assert re.search(r"^\s*module\s+delaunay_voronoi_mod\b", src, re.M)
    public = set(re.findall(r"^\s*public\s+(\w+)", src, re.M))
    ref_public = {"delaunay_voronoi_init", "delaunay_triangulation", "voronoi_diagram", "delaunay_voronoi_output",
                  "delaunay_vertex_type", "delaunay_triangle_type", "delaunay_edge_type", "voronoi_vertex_type",
                  "voronoi_cell_type", "voronoi_edge_type", "get_DVT", "get_DT", "get_DE", "get_VVT", "get_VC", "get_VE"}
    assert ref_public <= public
    want = {"delaunay_voronoi_init": ["num_point", "lon", "lat", "x", "y", "z"], "delaunay_triangulation": [],
            "voronoi_diagram": ["all"], "delaunay_voronoi_output": ["output"],
            "output_interface": ["DVT_array", "DT_list", "DE_array", "VVT_array", "VC_array", "VE_array", "tag"]}

    def arglists(text):
        return {m.group(1): [a.strip() for a in m.group(2).split(",") if a.strip()]
                for m in re.finditer(r"^\s*subroutine\s+(\w+)\s*\(([^)]*)\)", text, re.M)}
    got = arglists(src)
    for name, args in want.items():
        assert got.get(name) == args, (name, got.get(name))
    # the components a plug-in reads (consumer: src/mpas_mesh_mod.F90:205-294, :311-337)
    for tname, comps in {"delaunay_vertex_type": ["id", "incDT", "linkDVT"], "delaunay_triangle_type": ["id", "DVT", "adjDT", "center"],
                         "delaunay_edge_type": ["DVT1", "DVT2", "DT1", "DT2"], "voronoi_vertex_type": ["id", "DT", "VE", "VC"],
                         "voronoi_edge_type": ["id", "VVT1", "VVT2", "VC1", "VC2"],
                         "voronoi_cell_type": ["id", "center", "VVT", "VE", "adjVC"]}.items():
        body = re.search(r"type\b[^\n]*::\s*" + tname + r"\b(.*?)end type", src, re.S) or \
            re.search(r"^\s*type\s+" + tname + r"\b(.*?)end type", src, re.S | re.M)
        assert body, tname
        for comp in comps:
            assert re.search(r"\b" + comp + r"\b", body.group(1)), (tname, comp)
    ref = "/root/reference/src/delaunay_voronoi_mod.F90"
    if os.path.exists(ref):
        rsrc = "\n".join(ln.split("!")[0] for ln in open(ref).read().splitlines())
        rgot = arglists(rsrc)
        for name, args in want.items():
            assert rgot.get(name) == args, ("reference", name, rgot.get(name))
        assert set(re.findall(r"^\s*public\s+(\w+)", rsrc, re.M)) == ref_public
