"""oracle/checker.py -- TEST INFRASTRUCTURE ONLY.  Parity rules of BASELINE.json's north_star made executable
(SURVEY.md 8(c) "Parity rules"): sign identical outside |w_ref - threshold| < band, NaN masks identical, distances within
rel_tol relative; winding numbers compared by absolute difference; self-intersection bits outside the bands at 0 and 1."""
import numpy as np

from . import portlib, reflib

W_BAND = 1e-3
DIST_REL_TOL = 1e-5
# grid kernels: far-field DECISIONS are exact, accepted dipole terms use MUFU rsqrt (<= 1e-6 relative each); the contract
# is the sign outside |w - threshold| < W_BAND, this bound just documents how far w itself may drift
W_GRID_ATOL = 1e-5


def best_oracle(points, tris):
    """the unmodified reference (oracle/_ref) where it was built, else the C restatement"""
    if reflib.available():
        m = reflib.RefMesh.from_arrays(points, tris)
        m.kind = "reference"
        return m
    return portlib.PortMesh(points, tris)


def compare_w(got, want, atol=2e-6):
    got, want = np.asarray(got, np.float32), np.asarray(want, np.float32)
    diff = np.abs(got.astype(np.float64) - want.astype(np.float64))
    return {"ok": bool(got.shape == want.shape and (diff <= atol).all()), "max_abs_diff": float(diff.max()) if diff.size else 0.0,
            "bit_exact_frac": float((got.view(np.uint32) == want.view(np.uint32)).mean()) if diff.size else 1.0, "n": int(diff.size)}


def compare_sdf(got, want, w_ref=None, threshold=0.5, band=W_BAND, rel_tol=DIST_REL_TOL):
    """got/want: signed distances with NaN; w_ref: reference winding numbers of the same voxels (None = no band excuse)"""
    got, want = np.asarray(got, np.float32), np.asarray(want, np.float32)
    rep = {"n": int(want.size)}
    if got.shape != want.shape:
        return {"ok": False, "why": f"shape {got.shape} vs {want.shape}"}
    gn, wn = np.isnan(got), np.isnan(want)
    rep["nan_mismatch"] = int((gn != wn).sum())
    both = ~gn & ~wn
    g, w = got[both], want[both]
    sign_diff = np.signbit(g) != np.signbit(w)
    if w_ref is not None:
        inband = np.abs(np.asarray(w_ref, np.float32)[both].astype(np.float64) - threshold) < band
        rep["in_band"] = int(inband.sum())
        rep["sign_mismatch_in_band"] = int((sign_diff & inband).sum())
        sign_diff = sign_diff & ~inband
    rep["sign_mismatch"] = int(sign_diff.sum())
    ad, aw = np.abs(g).astype(np.float64), np.abs(w).astype(np.float64)
    err = np.abs(ad - aw)
    bad = err > rel_tol * aw
    rep["dist_mismatch"] = int(bad.sum())
    rep["max_rel_err"] = float((err[aw > 0] / aw[aw > 0]).max()) if (aw > 0).any() else 0.0
    rep["dist_bit_exact_frac"] = float((np.abs(g).view(np.uint32) == np.abs(w).view(np.uint32)).mean()) if g.size else 1.0
    rep["ok"] = rep["nan_mismatch"] == 0 and rep["sign_mismatch"] == 0 and rep["dist_mismatch"] == 0
    return rep


def compare_bits(got_words, want_words, w_ref, n_faces, band=W_BAND):
    """self-intersection bitsets: equal outside the bands |w| < band, |w-1| < band"""
    got = np.unpackbits(np.asarray(got_words, np.uint64).view(np.uint8), bitorder="little")[:n_faces].astype(bool)
    want = np.unpackbits(np.asarray(want_words, np.uint64).view(np.uint8), bitorder="little")[:n_faces].astype(bool)
    w = np.asarray(w_ref, np.float64)
    inband = (np.abs(w) < band) | (np.abs(w - 1) < band)
    mism = (got != want)
    return {"ok": bool(not (mism & ~inband).any()), "mismatch": int((mism & ~inband).sum()), "mismatch_in_band": int((mism & inband).sum()),
            "set": int(want.sum()), "n": int(n_faces)}


# ---- callers of the interface, restated over the oracle (checker only) ------------------------------------------------
def ref_mesh_to_distance_volume(orc, origin, voxel_size, dims, **opts):
    """MR::meshToDistanceVolume, HoleWindingRule branch (source/MRVoxels/MRMeshToDistanceVolume.cpp:43-62):
    basis = (scale(voxelSize), origin + 0.5f * voxelSize); min/max ignore NaN (MRParallelMinMax.h:14-42)"""
    o = np.asarray(origin, np.float32)
    v = np.asarray(voxel_size, np.float32) * np.ones(3, np.float32)
    b = o + np.float32(0.5) * v
    xf = np.array([v[0], 0, 0, 0, v[1], 0, 0, 0, v[2], b[0], b[1], b[2]], np.float32)
    data = orc.calc_from_grid_with_distances(dims, xf, **opts)
    ok = ~np.isnan(data)
    fmax = np.finfo(np.float32).max
    return data, (float(data[ok].min()) if ok.any() else float(fmax)), (float(data[ok].max()) if ok.any() else float(-fmax))


def ref_make_signed_by_winding_number(orc, volume, dims, xf, threshold=0.5, beta=2.0):
    """MR::makeSignedByWindingNumber on a dense grid (source/MRVoxels/MRVDBConversions.cpp:491-568): winding numbers of all
    voxels from the oracle's calcFromGrid (:552-556), then `val = -val` for every voxel with windVal > threshold (:527-533).
    The OpenVDB plumbing around that loop (dense fill, accessors, prune) cannot be built here and has no arithmetic.
    Returns (signed copy of the volume, winding numbers)."""
    w = np.asarray(orc.calc_from_grid(dims, xf, beta=beta), np.float32)
    out = np.array(volume, np.float32, copy=True).reshape(-1)
    flip = w > np.float32(threshold)
    out[flip] = -out[flip]
    return out, w


def ref_offset_grid(points, tris, offset, voxel_size):
    """grid of mcOffsetMesh: bbox.expanded(|offset|) -> calcOriginAndDimensions (source/MRVoxels/MROffset.cpp:163-165,
    MRCalcDims.h:18-27), in float32 like the reference"""
    f = np.float32
    t = np.asarray(tris)
    used = np.unique(t[(t >= 0).all(1)])
    p = np.asarray(points, f)[used]
    a = f(abs(offset))
    bmin, bmax = p.min(0) - a, p.max(0) + a
    exp = f(2) * f(voxel_size)
    origin = (bmin - exp).astype(f)
    inv = f(1) / f(voxel_size)
    dims = ((((bmax + exp).astype(f) - origin).astype(f) * inv).astype(f)).astype(np.int32) + 1
    return origin, tuple(int(d) for d in dims)


def ref_mc_offset_mesh(orc, points, tris, offset, voxel_size, close_holes=True, threshold=0.5, beta=2.0):
    """MR::mcOffsetMesh, HoleWindingRule (source/MRVoxels/MROffset.cpp:117-246) over the oracle's SDF and marching cubes:
    returns (verts, tris, sdf, origin, dims)"""
    from . import portlib, reflib
    f = np.float32
    origin, dims = ref_offset_grid(points, tris, offset, voxel_size)
    v, a = f(voxel_size), f(abs(offset))
    lo = max(f(a - f(f(1.001) * v)), f(0))
    hi = f(a + f(f(1.001) * v))
    b = origin + f(0.5) * v
    xf = np.array([v, 0, 0, 0, v, 0, 0, 0, v, b[0], b[1], b[2]], f)
    sdf = orc.calc_from_grid_with_distances(dims, xf, min_dist_sq=float(f(lo * lo)), max_dist_sq=float(f(hi * hi)),
                                            null_outside=not close_holes, threshold=threshold, beta=beta)
    mc = reflib.marching_cubes if reflib.available() else portlib.marching_cubes
    verts, mtris = mc(sdf, dims, float(v), origin, float(offset), True)
    return verts, mtris, sdf, origin, dims
