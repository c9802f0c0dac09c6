"""CPU test of bench.py's parity block (the comparison every bench line carries): fed with the reference's own values it reports
all-zero; a flipped sign away from the threshold, a moved distance and a NaN that should not be there are each counted."""
import numpy as np
import pytest

import bench
from oracle import reflib

pytestmark = pytest.mark.skipif(not reflib.available(), reason="needs oracle/_ref/libmrref.so")


def test_parity_block_counts_what_it_should():
    wl = bench.build_workload("c1")
    m, kind, _cores, _s = bench.cpu_reference_setup(wl)
    assert kind == "reference"
    layers = bench.sample_layers(wl["dims"][2], 16)[5:8]
    ref = {}
    bench.cpu_run_layers(m, wl, layers, ref)
    same = {z: v.copy() for z, v in ref.items()}
    rep = bench.parity_block(m, wl, ref, same)
    assert rep["ok"] and rep["layers"] == 3 and rep["dist_bit_exact_frac"] == 1.0
    assert (rep["nan_mismatch"], rep["sign_mismatch"], rep["dist_mismatch"]) == (0, 0, 0)
    bad = {z: v.copy() for z, v in ref.items()}
    z0 = layers[0]
    far = int(np.argmax(np.abs(ref[z0])))          # far from the surface: its winding number is nowhere near the threshold
    bad[z0][far] = -bad[z0][far]
    bad[layers[1]][100] *= np.float32(1.001)
    bad[layers[2]][7] = np.float32("nan")
    rep = bench.parity_block(m, wl, ref, bad)
    assert not rep["ok"]
    assert (rep["nan_mismatch"], rep["sign_mismatch"], rep["dist_mismatch"]) == (1, 1, 1)
    assert rep["dist_bit_exact_frac"] < 1.0


def test_sample_layers_are_spread_and_in_range():
    for dz in (1, 7, 16, 128, 2048):
        ls = bench.sample_layers(dz, min(dz, 16))
        assert ls == sorted(set(ls)) and ls[0] >= 0 and ls[-1] < dz and len(ls) == min(dz, 16)
