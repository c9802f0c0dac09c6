"""GPU: the CUDA library against tests/golden/ref_golden_next.npz (vectors of the unmodified reference for the SURVEY 8f rows
N1-N3) -- the same checker the CPU restatements pass in tests/test_oracle.py; needs no reference library on the box."""
import numpy as np
import pytest

import golden_next_check as gc
import meshlib_b200 as mb

pytestmark = pytest.mark.gpu


class _Ours(gc.Impl):
    def __init__(self, pts, tris, q):
        self.fwn, self.q = mb.FastWindingNumber(pts, tris), q

    def find_projections(self, q, up, lo):
        return self.fwn.find_projections(q, None, None, up, lo)

    def signed_distance_pn(self, q, min_dist_sq, max_dist_sq, null_outside):
        return self.fwn.signed_distances_by_projection_normal(q, mb.DistOptions(min_dist_sq, max_dist_sq, null_outside))

    def make_signed(self, unsigned, dims, xf, threshold, beta):
        return self.fwn.make_signed_by_winding_number(unsigned, dims, xf, threshold, beta)

    def refit(self, new_points):
        self.fwn.update_points(new_points)
        nodes, dipoles = self.fwn.get_tree()
        return nodes, dipoles, self.fwn.calc_from_vector(self.q)


@pytest.mark.parametrize("name", gc.NAMES)
def test_cuda_library_against_golden_next_fixture(name):
    g = gc.load()
    made = []

    def make(pts, tris):
        made.append(_Ours(pts, tris, g[f"{name}_q"]))
        return made[-1]
    ran = gc.check(make, g, name)
    made[0].fwn.close()
    assert len(ran) == 6, ran
