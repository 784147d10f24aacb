"""Builds and binds tests/emu/fm_emu.cpp: the kernel bodies of fishman_b200/csrc run
thread by thread on the host.  TEST INFRASTRUCTURE ONLY (CPU logic tests); the
product library has no host path."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "emu", "fm_emu.cpp")
LIB = os.path.join(HERE, "emu", "libfm_emu.so")
CSRC = os.path.join(HERE, "..", "fishman_b200", "csrc")
_lib = None


def build():
    deps = [SRC] + [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    if (not os.path.exists(LIB)) or os.path.getmtime(LIB) < max(os.path.getmtime(d) for d in deps):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared",
                               "-o", LIB, SRC, "-lm"])
    return LIB


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB)
        L.fme_last_error.restype = C.c_char_p
        _lib = L
    return _lib


def triangulate(xyz, sort=True):
    """Returns (tri, adj, stats) with 1-based ids, through the emulated kernels."""
    p = np.ascontiguousarray(xyz, dtype=np.float64)
    n = len(p)
    T = 2 * n - 4
    tri = np.empty((T, 3), np.int32)
    adj = np.empty((T, 3), np.int32)
    st = np.zeros(4, np.int32)
    ip = C.POINTER(C.c_int)
    rc = lib().fme_triangulate(n, p.ctypes.data_as(C.POINTER(C.c_double)), tri.ctypes.data_as(ip),
                               adj.ctypes.data_as(ip), st.ctypes.data_as(ip), 1 if sort else 0)
    if rc:
        raise RuntimeError(lib().fme_last_error().decode())
    return tri, adj, dict(rounds=int(st[0]), subrounds=int(st[1]), flips=int(st[2]), max_active=int(st[3]))


def repair(xyz_old, xyz_new, nranks=1, reverse=False):
    """Warm repair of the triangulation of xyz_old after the points moved to xyz_new, as `nranks`
    ranks of a sharded mesh do it (nranks = 1: the unsharded sweep).  Returns (tri, adj, turned,
    n_requests); tri/adj are slot-for-slot comparable between runs."""
    a = np.ascontiguousarray(xyz_old, dtype=np.float64)
    b = np.ascontiguousarray(xyz_new, dtype=np.float64)
    n = len(a)
    T = 2 * n - 4
    tri = np.empty((T, 3), np.int32)
    adj = np.empty((T, 3), np.int32)
    turned, nreq = C.c_int(0), C.c_int(0)
    ip, dp = C.POINTER(C.c_int), C.POINTER(C.c_double)
    rc = lib().fme_repair(n, a.ctypes.data_as(dp), b.ctypes.data_as(dp), int(nranks), 1 if reverse else 0,
                          tri.ctypes.data_as(ip), adj.ctypes.data_as(ip), C.byref(turned), C.byref(nreq))
    if rc:
        raise RuntimeError(lib().fme_last_error().decode())
    return tri, adj, turned.value, nreq.value


def triangulate_sharded(xyz, nranks):
    """Cold build with every level's point location done in nranks slices (the sharded form)."""
    p = np.ascontiguousarray(xyz, dtype=np.float64)
    n = len(p)
    T = 2 * n - 4
    tri = np.empty((T, 3), np.int32)
    adj = np.empty((T, 3), np.int32)
    ip = C.POINTER(C.c_int)
    rc = lib().fme_triangulate_sharded(n, p.ctypes.data_as(C.POINTER(C.c_double)), int(nranks),
                                       tri.ctypes.data_as(ip), adj.ctypes.data_as(ip))
    if rc:
        raise RuntimeError(lib().fme_last_error().decode())
    return tri, adj


def rebuild_after_failed_repair(xyz_old, xyz_new, skip_bump=False):
    """build(xyz_old), move the points, warm sweep (finds turned-over triangles and leaves), then
    the in-place rebuild.  Returns (tri, adj, turned).  skip_bump: start the rebuild on the stamp the
    sweep used (the defect fm_retriangulate(h, warm=1) had)."""
    a = np.ascontiguousarray(xyz_old, dtype=np.float64)
    b = np.ascontiguousarray(xyz_new, dtype=np.float64)
    n = len(a)
    T = 2 * n - 4
    tri = np.empty((T, 3), np.int32)
    adj = np.empty((T, 3), np.int32)
    turned = C.c_int(0)
    ip, dp = C.POINTER(C.c_int), C.POINTER(C.c_double)
    rc = lib().fme_rebuild_after_failed_repair(n, a.ctypes.data_as(dp), b.ctypes.data_as(dp), 1 if skip_bump else 0,
                                               tri.ctypes.data_as(ip), adj.ctypes.data_as(ip), C.byref(turned))
    if rc:
        raise RuntimeError(lib().fme_last_error().decode())
    return tri, adj, turned.value


def triangulate_star(xyz, kappa=2.0, wide=False):
    """Cold build by stars (csrc/star_kernels.cuh) on the host.  Returns (rc, tri, adj, stats): rc 0 = complete,
    2 = the product would fall back to the insertion build (stats say why)."""
    p = np.ascontiguousarray(xyz, dtype=np.float64)
    n = len(p)
    T = 2 * n - 4
    tri = np.zeros((T, 3), np.int32)
    adj = np.zeros((T, 3), np.int32)
    st = np.zeros(12, np.int32)
    ip = C.POINTER(C.c_int)
    rc = lib().fme_triangulate_star(n, p.ctypes.data_as(C.POINTER(C.c_double)), C.c_double(kappa), 1 if wide else 0,
                                    tri.ctypes.data_as(ip), adj.ctypes.data_as(ip), st.ctypes.data_as(ip))
    if rc == 1:
        raise RuntimeError(lib().fme_last_error().decode())
    names = ("retried", "exact", "failed", "broken", "owned", "degree_sum", "cells_per_axis", "slow", "full_wrap")
    return rc, tri, adj, dict(zip(names, (int(v) for v in st[:9])))
