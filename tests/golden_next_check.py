"""Shared checker for tests/golden/ref_golden_next.npz (generated from the unmodified reference by
tests/golden/make_golden_next.py).  An implementation under test is an object with the methods of `Impl` below; the same
checks run over the reference itself (sanity of the fixture), over the CPU restatements, and over the CUDA library."""
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_golden_next.npz")
NAMES = ("torus", "noisy_sphere")
W_BAND = 1e-3       # sign contract of the winding number: |w - threshold| below this is excused
DOT_TIE = 1e-5      # ProjectionNormal: |dot( proj - pt, pseudonormal )| <= DOT_TIE * |d| is a rounding tie
FLT_MAX = float(np.finfo(np.float32).max)


class Impl:
    """what check() calls; every method may return None = "not available in this implementation" (the check is skipped)"""

    def find_projections(self, q, up, lo):          # structured array with face, point, a, b, dist_sq
        return None

    def signed_distance_pn(self, q, min_dist_sq, max_dist_sq, null_outside):
        return None

    def make_signed(self, unsigned, dims, xf, threshold, beta):
        return None

    def refit(self, new_points):                    # -> (nodes, dipoles, winding numbers of q after the refit) via callables
        return None


def load():
    return np.load(GOLDEN)


def _same_bits(a, b):
    a, b = np.ascontiguousarray(a), np.ascontiguousarray(b)
    return a.shape == b.shape and a.tobytes() == b.tobytes()


def check(make_impl, g, name, w_atol=2e-6):
    """make_impl(points, tris) -> Impl.  Returns the list of checks that ran."""
    pts, tris, q = g[f"{name}_points"], g[f"{name}_tris"], g[f"{name}_q"]
    impl = make_impl(pts, tris)
    ran = []
    # N1: records bit for bit
    for key, up, lo in ((f"{name}_proj", FLT_MAX, 0.0), (f"{name}_proj_limits", 0.25, 0.01)):
        got = impl.find_projections(q, up, lo)
        if got is not None:
            want = g[key]
            for f in ("face", "point", "a", "b", "dist_sq"):
                assert _same_bits(got[f].view(np.uint32), want[f].view(np.uint32)), (name, key, f)
            ran.append(key)
    # N2 (b): NaN masks and magnitudes bit for bit, signs outside rounding ties
    for key, kw in ((f"{name}_pn", dict(min_dist_sq=0.0, max_dist_sq=FLT_MAX, null_outside=False)),
                    (f"{name}_pn_band", dict(min_dist_sq=0.0, max_dist_sq=0.25, null_outside=True))):
        got = impl.signed_distance_pn(q, **kw)
        if got is not None:
            want, dots = g[key], g[key + "_dots"]
            assert np.array_equal(np.isnan(got), np.isnan(want)), (name, key)
            ok = ~np.isnan(want)
            assert _same_bits(np.abs(got[ok]), np.abs(want[ok])), (name, key)
            clear = ok & (np.abs(dots) > DOT_TIE * np.abs(want))  # the query points ON the mesh (distance 0) have no sign
            assert clear.sum() > 0.98 * (ok & (want != 0)).sum() and np.array_equal(np.signbit(got[clear]), np.signbit(want[clear])), (name, key)
            ran.append(key)
    # N2 (a): magnitudes untouched, signs outside the winding-number band
    dims, xf = tuple(int(d) for d in g[f"{name}_dims"]), g[f"{name}_xf"]
    got = impl.make_signed(g[f"{name}_unsigned"].copy(), dims, xf, 0.4, 2.5)
    if got is not None:
        want, w = g[f"{name}_signed"], g[f"{name}_sign_w"]
        assert _same_bits(np.abs(got), np.abs(want)), name
        outside = np.abs(w.astype(np.float64) - 0.4) >= W_BAND
        assert np.array_equal(np.signbit(got)[outside], np.signbit(want)[outside]), name
        assert np.signbit(want).any() and not np.signbit(want).all()
        ran.append(f"{name}_signed")
    # N3: refitted tree and dipoles bit for bit, winding numbers on it
    r = impl.refit(g[f"{name}_moved_points"])
    if r is not None:
        nodes, dipoles, w = r
        assert _same_bits(nodes, g[f"{name}_refit_nodes"]) and _same_bits(dipoles, g[f"{name}_refit_dipoles"]), name
        assert np.abs(np.asarray(w, np.float64) - g[f"{name}_refit_w"]).max() <= w_atol, name
        ran.append(f"{name}_refit")
    return ran
