"""CPU tests of the drop-in boundary: libb2fwn.so loads, exports every symbol include/b2fwn.h declares, and fails loudly
(no CPU fallback) when there is no CUDA device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import meshlib_b200 as mb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "b2fwn.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b2fwn_[a-z_0-9]+)\s*\(", src)))


def test_header_symbols_exported():
    L = mb.load_library()
    names = _declared()
    assert len(names) >= 18
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/b2fwn.h but not exported"
    assert sorted(mb.EXPORTS) == names


def test_struct_layouts_match_header():
    assert C.sizeof(mb.DistOptions) == 20
    assert C.sizeof(mb.api.Config) == 96
    assert C.sizeof(mb.api.Stats) == 64
    assert C.sizeof(mb.api.McParams) == 48
    assert mb.PROJECTION_DTYPE.itemsize == 32
    assert C.sizeof(mb.VolumeParams) == 48 and C.sizeof(mb.OffsetParams) == 32 and C.sizeof(mb.OffsetResult) == 56
    o = mb.DistOptions()
    assert (o.min_dist_sq, o.null_outside_min_max, o.winding_number_threshold, o.winding_number_beta) == (0.0, 1, 0.5, 2.0)
    assert o.max_dist_sq == np.finfo(np.float32).max


def test_struct_sizes_against_the_c_compiler(tmp_path):
    # the header compiled as plain C: its sizeof values are what the ctypes mirrors must have
    import subprocess
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "b2fwn.h"\nint main(void){printf("%zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(b2fwn_config),'
                   'sizeof(b2fwn_dist_options), sizeof(b2fwn_stats), sizeof(b2fwn_projection), sizeof(b2fwn_mc_params),'
                   'sizeof(b2fwn_volume_params), sizeof(b2fwn_offset_params), sizeof(b2fwn_offset_result));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", f"-I{ROOT}/include", str(src), "-o", str(exe)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    want = [C.sizeof(mb.api.Config), C.sizeof(mb.DistOptions), C.sizeof(mb.api.Stats), mb.PROJECTION_DTYPE.itemsize, C.sizeof(mb.api.McParams),
            C.sizeof(mb.VolumeParams), C.sizeof(mb.OffsetParams), C.sizeof(mb.OffsetResult)]
    assert got == want


def test_no_cpu_fallback_without_device():
    L = mb.load_library()
    if L.b2fwn_device_count() > 0:
        pytest.skip("a CUDA device is present")
    from meshlib_b200 import meshgen
    pts, tris = meshgen.make_uv_sphere(1, 8, 8)
    with pytest.raises(mb.B2Error) as e:
        mb.FastWindingNumber(pts, tris)
    assert "NVIDIA GPU error" in str(e.value)


def test_product_does_not_touch_oracle():
    # the product path may not import, link, load or execute anything under oracle/
    bad = re.compile(r"^\s*(from|import)\s+oracle|liboracle|libmrref|oracle/|from \.\. ?import oracle", re.M)
    for dirpath, _d, files in os.walk(os.path.join(ROOT, "meshlib_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")) or f == "Makefile":
                s = open(os.path.join(dirpath, f)).read()
                assert not bad.search(s), f"{f} references the oracle"


def test_meshgen_shapes():
    from meshlib_b200 import meshgen
    p, t = meshgen.make_torus(2, 1, 256, 128)
    assert p.shape == (32768, 3) and t.shape == (65536, 3) and t.max() == 32767
    p, t = meshgen.make_uv_sphere(1, 1000, 500)
    assert t.shape == (1000000, 3) and p.shape == (500002, 3)
    q = meshgen.query_points(1000)
    assert q.shape == (1000, 3) and q.min() >= -1.5 and q.max() <= 1.5
    assert np.array_equal(q, meshgen.query_points(1000))


@pytest.mark.skipif(not os.path.isdir("/root/reference/source/MRMesh"), reason="reference headers not present")
def test_cpp_adapter_compiles_against_reference_headers():
    # the adapter derives from MR::IFastWindingNumber / IFastWindingNumberByParts (source/MRMesh/MRFastWindingNumber.h)
    import subprocess
    ref = "/root/reference"
    src = os.path.join(ROOT, "meshlib_b200", "host", "b2fwn_meshlib_adapter.hpp")
    cmd = ["g++", "-std=c++23", "-fsyntax-only", "-w", "-x", "c++", f"-I{ref}/source", f"-I{ref}/thirdparty/onetbb/include",
           f"-I{ref}/thirdparty/parallel-hashmap", f"-I{ref}/thirdparty/expected/include", f"-I{ROOT}/include", src]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]


def test_part_plan_matches_split_by_chunks():
    """b2fwn_plan_parts (the partition runChunked uses; no device needed) against the restated splitByChunks of the oracle and the
    reference's own known answers (/root/reference/source/MRMesh/MRChunkIterator.cpp:52-80), plus the round-robin deal over devices"""
    from oracle import portlib
    assert [(b, s) for b, s, _d in mb.plan_parts(8, 4, 0)] == [(0, 4), (4, 4)]
    assert [(b, s) for b, s, _d in mb.plan_parts(9, 4, 0)] == [(0, 4), (4, 4), (8, 1)]
    assert [(b, s) for b, s, _d in mb.plan_parts(10, 4, 1)] == [(0, 4), (3, 4), (6, 4)]
    assert [(b, s) for b, s, _d in mb.plan_parts(11, 4, 1)] == [(0, 4), (3, 4), (6, 4), (9, 2)]
    assert mb.plan_parts(0, 4, 0) == []
    for total in (1, 2, 7, 19, 64, 2048):
        for part in (1, 2, 3, 5, 8, 64, 5000):
            for overlap in (0, 1, 2):
                if part <= overlap:
                    with pytest.raises(mb.B2Error):
                        mb.plan_parts(total, part, overlap)
                    continue
                want = portlib.split_by_chunks(total, min(part, total), overlap)
                if total <= overlap:
                    continue  # the reference yields no chunk there; the pipeline computes the few layers as one part
                got = mb.plan_parts(total, part, overlap, n_devices=3)
                assert [(b, s) for b, s, _d in got] == want, (total, part, overlap)
                nd = min(3, len(got))
                assert [d for _b, _s, d in got] == [k % nd for k in range(len(got))]
