"""CPU tests of the two-mode winding-number walk (kGridFwn6) without a GPU: the brick-level tests of the product header
meshlib_b200/csrc/b2fwn_brick_math.h compiled for the host, and a host replay of the kernel's walk (tests/brick_walk_host.cpp)
held against the oracle.  What must hold: a uniform answer (ALL_NEAR / ALL_FAR) never contradicts the exactly rounded per-voxel
decision of Dipole::addIfGoodApprox (/root/reference/source/MRMesh/MRDipole.h:18-27), so every voxel accepts exactly the
reference's set of dipoles and its winding number differs only by the rounding of the accepted terms."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from meshlib_b200 import make_basis_xf, meshgen
from oracle import checker

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "brick_walk_host.cpp")
HDR = os.path.join(HERE, "..", "meshlib_b200", "csrc", "b2fwn_brick_math.h")


@pytest.fixture(scope="module")
def bw(tmp_path_factory):
    so = os.path.join(tmp_path_factory.mktemp("bw"), "libbrickwalk.so")
    flags = ["-O2", "-ffp-contract=off", "-shared", "-fPIC"]
    if "fma" in open("/proc/cpuinfo").read().split():
        flags.append("-mfma")  # fmaf as one instruction; without it libm's (equally exact) software fmaf is used
    subprocess.run(["g++", *flags, "-o", so, SRC], check=True)
    lib = C.CDLL(so)
    lib.bw_audit_brick_tests.restype = C.c_uint64
    lib.bw_audit_brick_tests.argtypes = [C.c_uint64, C.c_uint64, C.c_float, C.c_float, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    lib.bw_calc_from_grid.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float,
                                      C.c_void_p, C.c_void_p, C.c_void_p]
    return lib


@pytest.mark.parametrize("coord_scale,voxel", [(2.0, 1e-3), (2.0, 0.05), (4000.0, 1e-3), (1e-2, 1e-6), (3e5, 7.0)])
def test_uniform_answers_never_contradict_a_voxel(bw, coord_scale, voxel):
    near, far = C.c_uint64(0), C.c_uint64(0)
    bad = bw.bw_audit_brick_tests(12345, 150000, coord_scale, voxel, C.byref(near), C.byref(far))
    assert bad == 0
    assert near.value > 1000 and far.value > 1000  # both uniform answers were actually given


def _replay(bw, orc, pts, tris, dims, xf, beta=2.0, valid=None):
    nodes, dip = orc.nodes(), orc.dipoles()
    pts = np.ascontiguousarray(pts, np.float32)
    tris = np.ascontiguousarray(orc.tris if hasattr(orc, "tris") else tris, np.int32)
    d = np.asarray(dims, np.int32)
    xf = np.ascontiguousarray(xf, np.float32)
    out = np.full(int(d[0]) * int(d[1]) * int(d[2]), np.nan, np.float32)
    cnt = np.zeros(8, np.uint64)
    vp = None if valid is None else np.ascontiguousarray(valid, np.uint8).ctypes.data
    bw.bw_calc_from_grid(nodes.ctypes.data, dip.ctypes.data, len(nodes), pts.ctypes.data, tris.ctypes.data, d.ctypes.data,
                         xf.ctypes.data, beta, vp, out.ctypes.data, cnt.ctypes.data)
    return out, cnt


@pytest.mark.parametrize("mesh,dims", [("torus", (29, 22, 13)), ("sphere", (40, 40, 40)), ("noisy", (67, 41, 23))])
def test_two_mode_walk_matches_the_oracle(bw, mesh, dims):
    if mesh == "torus":
        pts, tris = meshgen.make_torus(2, 1, 96, 48)
    else:
        pts, tris = meshgen.make_uv_sphere(1, 80, 40)
        if mesh == "noisy":
            pts = meshgen.add_noise(pts, 4e-3)
    orc = checker.best_oracle(pts, tris)
    lo, hi = pts.min(0) - 0.3, pts.max(0) + 0.3
    voxel = ((hi - lo) / np.asarray(dims)).astype(np.float32)
    xf = make_basis_xf(lo.astype(np.float32), voxel)
    want = orc.calc_from_grid(dims, xf)
    got, cnt = _replay(bw, orc, pts, tris, dims, xf)
    assert cnt[4] == 0 and cnt[5] == 0, cnt
    assert cnt[1] > 0 and cnt[2] > 0 and cnt[3] > 0  # uniform ALL_FAR answers and diverged sections both happened
    rep = checker.compare_w(got, want, atol=checker.W_GRID_ATOL)
    assert rep["ok"], rep
    # voxels that take no part (NaN distances in mode 1) shrink the ball; the others keep their value bit for bit
    rng = np.random.default_rng(7)
    valid = (rng.random(got.size) < 0.3).astype(np.uint8)
    got2, cnt2 = _replay(bw, orc, pts, tris, dims, xf, valid=valid)
    assert cnt2[4] == 0 and cnt2[5] == 0
    sel = valid.astype(bool)
    assert np.array_equal(got2[sel].view(np.uint32), got[sel].view(np.uint32))
    assert np.isnan(got2[~sel]).all()
