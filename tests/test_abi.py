"""The drop-in boundary: libfishman_b200.so must load without a GPU and export exactly
the entry points include/fishman_b200.h declares; without a CUDA device the product
fails loudly (no CPU fallback); the product package never touches oracle/."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "fishman_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(fm_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from fishman_b200 import _lib, build
    build.build()
    L = C.CDLL(build.LIB)
    syms = header_symbols()
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(L, s), f"{s} declared in include/fishman_b200.h but not exported"
    assert sorted(_lib.SYMBOLS) == syms, "ctypes binding and header disagree"
    out = subprocess.check_output(["nm", "-D", "--defined-only", build.LIB], text=True)
    exported = sorted(set(re.findall(r" T (fm_[a-z0-9_]+)$", out, flags=re.M)))
    assert exported == syms, "library exports entry points the header does not declare"


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import numpy as np
    import fishman_b200 as fb
    with pytest.raises(fb.FishmanError, match="no CPU path"):
        fb.Mesh.from_xyz(np.eye(4, 3))


def test_product_never_references_the_oracle():
    pkg = os.path.join(ROOT, "fishman_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".inc", ".F90")):
                txt = open(os.path.join(d, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt and "libfishman_oracle" not in txt, f
                assert "tests/emu" not in txt.replace("tests/emu (CPU logic tests", "").replace("in tests/emu", "").replace("(tests/emu)", ""), f


def test_sass_is_sm100a_only():
    from fishman_b200 import build
    out = subprocess.check_output(["cuobjdump", "--list-elf", build.LIB], text=True)
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_build_staleness_check_covers_every_source():
    """The in-tree library is rebuilt when ANY file under csrc/ changes (a list by name once missed a new header
    and the GPU runs silently used a stale library)."""
    from fishman_b200 import build
    csrc = os.path.join(ROOT, "fishman_b200", "csrc")
    deps = {os.path.realpath(d) for d in build.DEPS}
    for f in os.listdir(csrc):
        assert os.path.realpath(os.path.join(csrc, f)) in deps, f
