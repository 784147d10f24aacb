#!/usr/bin/env python
"""Counts the fp64 instructions of the centroid / quadrature kernel in the SASS of the built
library and writes profiles/sass_counts.json (read by bench.py for roofline.fp64) plus a
tracked excerpt profiles/<tag>_sass_centroid_uniform.txt.

  python tools/sass_dp_count.py [tag]

Method: cuobjdump -sass of fm_centroid_update<true>; every backward branch closes a loop; the
quadrature loops are the ones with eight MUFU.RSQ64H per trip (seven node normalisations + the
sub-triangle area).  The hot loop is the ring-regrouped form (quad_uniform, taken for rings of
<= 12 triangles): the one with fewer DP instructions; the other is the long-ring fallback.
DP instructions = DFMA, DMUL, DADD (flops: 2, 1, 1) + DSETP / DMNMX (pipe slots, no flops).
"""
import hashlib
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "fishman_b200", "libfishman_b200.so")
FUNC = "_ZN2fm18fm_centroid_updateILb1ELb0EEEvNS_5VdCtxEi"


def csrc_sha16():
    h = hashlib.sha256()
    d = os.path.join(ROOT, "fishman_b200", "csrc")
    for f in sorted(os.listdir(d)):
        with open(os.path.join(d, f), "rb") as fh:
            h.update(f.encode() + b"\0" + fh.read())
    return h.hexdigest()[:16]


def function_sass(func):
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    lines, on = [], False
    for ln in out.splitlines():
        if "Function :" in ln:
            on = func in ln
            continue
        if on and re.search(r"/\*[0-9a-f]{4,}\*/\s+\S", ln):
            lines.append(ln)
    ins = []
    for ln in lines:
        m = re.search(r"/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            ins.append((int(m.group(1), 16), m.group(2).strip()))
    return ins


def opcode(text):
    t = text.split()
    if t and t[0].startswith("@"):
        t = t[1:]
    return t[0] if t else ""


def count(ins):
    c = {"DFMA": 0, "DMUL": 0, "DADD": 0, "DSETP": 0, "DMNMX": 0, "MUFU.RSQ64H": 0, "MUFU.RCP64H": 0, "total": len(ins)}
    for _, t in ins:
        op = opcode(t)
        base = op.split(".")[0]
        if base in ("DFMA", "DMUL", "DADD", "DSETP", "DMNMX"):
            c[base] += 1
        elif op.startswith("MUFU.RSQ64H"):
            c["MUFU.RSQ64H"] += 1
        elif op.startswith("MUFU.RCP64H"):
            c["MUFU.RCP64H"] += 1
    c["dp_instr"] = c["DFMA"] + c["DMUL"] + c["DADD"] + c["DSETP"] + c["DMNMX"]
    c["flop"] = 2 * c["DFMA"] + c["DMUL"] + c["DADD"]
    return c


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "r02"
    ins = function_sass(FUNC)
    if not ins:
        sys.exit("function not found in the SASS")
    loops = []
    for a, t in ins:
        m = re.search(r"\bBRA\b.*?(0x[0-9a-f]+)", t)
        if m and int(m.group(1), 16) < a:
            lo = int(m.group(1), 16)
            body = [(x, y) for x, y in ins if lo <= x <= a]
            loops.append((lo, a, count(body)))
    quad = [L for L in loops if L[2]["MUFU.RSQ64H"] >= 8 and L[2]["MUFU.RSQ64H"] % 8 == 0]
    if not quad:
        sys.exit("no quadrature loop found")
    quad.sort(key=lambda L: L[2]["dp_instr"] / (L[2]["MUFU.RSQ64H"] // 8))
    lo, hi, c = quad[0]
    unroll = c["MUFU.RSQ64H"] // 8
    whole = count(ins)
    in_all_quad = {k: sum(L[2][k] for L in quad) for k in ("dp_instr", "flop")}
    outside = {k: whole[k] - in_all_quad[k] for k in ("dp_instr", "flop")}
    res = {
        "csrc_sha16": csrc_sha16(),
        "source": f"profiles/{tag}_sass_centroid_uniform.txt (cuobjdump -sass of fm_centroid_update<true>, tools/sass_dp_count.py)",
        "centroid_update_uniform": {
            "loop_addresses": [hex(lo), hex(hi)], "unroll": unroll,
            "per_trip": c,
            "dp_instr_per_subtriangle": c["dp_instr"] / unroll,
            "flop_per_subtriangle": c["flop"] / unroll,
            "mufu_per_subtriangle": (c["MUFU.RSQ64H"] + c["MUFU.RCP64H"]) / unroll,
            "instr_per_subtriangle": c["total"] / unroll,
            "dp_instr_per_cell_outside_loop": outside["dp_instr"],
            "flop_per_cell_outside_loop": outside["flop"],
            "other_quadrature_loops": [{"addresses": [hex(L[0]), hex(L[1])], "per_trip": L[2]} for L in quad[1:]],
        },
    }
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    with open(os.path.join(ROOT, "profiles", "sass_counts.json"), "w") as f:
        json.dump(res, f, indent=1)
    with open(os.path.join(ROOT, "profiles", f"{tag}_sass_centroid_uniform.txt"), "w") as f:
        f.write(f"# {FUNC}: the hot quadrature loop {hex(lo)}..{hex(hi)} (one trip = {unroll} sub-triangle(s))\n")
        f.write(f"# per trip: {json.dumps(c)}\n# csrc_sha16 {res['csrc_sha16']}\n")
        for a, t in ins:
            if lo <= a <= hi:
                f.write(f"/*{a:04x}*/ {t} ;\n")
    print(json.dumps(res["centroid_update_uniform"], indent=1))


if __name__ == "__main__":
    main()
