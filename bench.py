#!/usr/bin/env python
"""bench.py -- headline benchmark of the mesh -> signed-distance-volume hot path (BASELINE.json metric:
SDF voxels/s of calcFromGridWithDistances = fast-winding-number sign + closest-point distance).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

A "step" is one pass of the hot path over the whole voxel grid of the workload.
  value  : voxels/s, result resident in HBM (the _device form of the C ABI), device time by CUDA events on the
           launching stream (b2fwn_stats.last_kernel_ms), max over ranks.
  e2e    : voxels/s through the reference-facing call with HOST buffers: mesh arrays on the host -> b2fwn_create
           (H2D + device tree/dipole build) -> b2fwn_calc_from_grid_with_distances into a PAGEABLE host volume (chunked,
           staged D2H inside the timed region) -> b2fwn_destroy; one process drives all N GPUs (multi-device handle).
  --impl reference : the reference's own CPU/TBB implementation (oracle/_ref when it was built, else the C
           restatement) on a bounded sample of the same workload, all host threads.
N > 1 (torchrun): the grid is sharded by z (interleaved chunks of a few layers per rank, meshlib_b200/slabs.py), mesh and
tree replicated per GPU, the slabs gathered into the full volume on every GPU by NCCL all-gather (north_star), stage by
stage behind the kernels of the next stage; total work is fixed => "scaling": "strong".
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from meshlib_b200 import meshgen, slabs  # noqa: E402

METRIC = "sdf_voxels_per_sec"
UNIT = "voxels/s"

# name -> (mesh spec, grid n, options); C-numbers refer to BASELINE.json configs / SURVEY.md 8(d)
WORKLOADS = {
    # C5: noisy UV sphere (1, 3162, 1581) -> 9,998,244 faces, 2048^3, narrow band (maxDist = 3 voxels, no NaN)
    "c5_band": dict(mesh=("noisy_sphere", 3162, 1581, 1e-3), grid=2048, band_voxels=3.0, null_outside=False),
    # same mesh, exact distance everywhere (default options)
    "c5_full": dict(mesh=("noisy_sphere", 3162, 1581, 1e-3), grid=2048, band_voxels=None, null_outside=True),
    # C1: torus 65,536 faces on 128^3, default options (the reference's own CPU-runnable case)
    "c1": dict(mesh=("torus", 256, 128), grid=128, band_voxels=None, null_outside=True),
    # C3-shaped: 2M-face noisy sphere at 512^3, band
    "c3_band": dict(mesh=("noisy_sphere", 1414, 707, 1e-3), grid=512, band_voxels=3.0, null_outside=False),
    "c3_full": dict(mesh=("noisy_sphere", 1414, 707, 1e-3), grid=512, band_voxels=None, null_outside=True),
}
DEFAULT_WORKLOAD = "c5_band"

# C3 as BASELINE.json states it: generalOffsetMesh (Standard mode, HoleWindingRule sign) of the 2 M-face noisy sphere at ~512^3 voxels
# INCLUDING marching cubes: b2fwn_mc_offset_mesh (distance band + marching cubes, both in HBM).  Diagnostic workload, not the headline.
OFFSET_WORKLOADS = {
    "c3_offset": dict(mesh=("noisy_sphere", 1414, 707, 1e-3), voxel=2.04 / 508.0, offset=0.02, close_holes=True),
    "c5_offset": dict(mesh=("noisy_sphere", 3162, 1581, 1e-3), voxel=2.04 / 2040.0, offset=0.005, close_holes=True),
}

# DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) of ONE launch of the dominant kernel (kGridFwn6, whole grid on one
# GPU): not measurable inside a timed run, so READ from the committed raw page of the `ncu --set full` capture on B200
# (scripts/gpu_r3.sh stages fwn6 / c3; c5_band by application replay so that all counter passes complete)
NCU_CAPTURES = {
    "c5_band": ("profiles/r03/ncu_full_kGridFwn6_c5_band.csv", "kGridFwn6"),
    "c3_band": ("profiles/r03/ncu_full_c3_band.csv", "kGridFwn6"),
}
# flops per voxel by SURVEY 8(d)'s weights over the reference algorithm's counters, as the N = 1 run measures them on its sample
# layers (roofline.flops_per_voxel of that line): carried to the lines that have no CPU leg (N > 1, --no-cpu)
FLOPS_PER_VOXEL = {
    "c5_band": (2645.5708909032824, "BENCH_r02.json parsed.roofline.flops_per_voxel (N = 1, 200336-voxel sample of 16 layers)"),
}
_NCU_UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def ncu_traffic(workload):
    """(bytes per launch, source) of the workload's dominant kernel from its committed capture; (None, reason) without one"""
    if workload not in NCU_CAPTURES:
        return None, "no capture of this workload"
    rel, kernel = NCU_CAPTURES[workload]
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), rel)
    if not os.path.exists(path):
        return None, f"{rel} is missing"
    import csv
    with open(path, newline="") as f:
        rows = list(csv.reader(f))
    head, units = rows[0], rows[1]
    for r in rows[2:]:
        if kernel in r[head.index("Kernel Name")]:
            total = 0.0
            for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                c = head.index(k)
                total += float(r[c].replace(",", "")) * _NCU_UNIT[units[c]]
            return total, f"{rel} ({kernel}: dram__bytes_read.sum + dram__bytes_write.sum)"
    return None, f"{rel} holds no {kernel}"


def build_workload(name, grid_override=None):
    w = dict(WORKLOADS[name])
    spec = w["mesh"]
    if spec[0] == "torus":
        pts, tris = meshgen.make_torus(2, 1, spec[1], spec[2])
        n = grid_override or w["grid"]
        dims, origin, voxel = (n, n, n), np.array([-3.2, -3.2, -3.2], np.float32), np.float32(6.4 / n)
    else:
        pts, tris = meshgen.make_uv_sphere(1, spec[1], spec[2])
        pts = meshgen.add_noise(pts, spec[3], seed=42)
        n = grid_override or w["grid"]
        dims, origin, voxel = meshgen.grid_for_bbox(pts, n, 2.0)
    from meshlib_b200 import make_basis_xf
    xf = make_basis_xf(origin, voxel)
    fmax = float(np.finfo(np.float32).max)
    max_sq = fmax if w["band_voxels"] is None else float(np.float32((w["band_voxels"] * float(voxel)) ** 2))
    opts = dict(min_dist_sq=0.0, max_dist_sq=max_sq, null_outside=bool(w["null_outside"]), threshold=0.5, beta=2.0)
    return dict(name=name, points=pts, tris=tris, dims=dims, xf=xf, voxel=float(voxel), opts=opts,
                desc=f"{name}: {len(tris)} faces, grid {dims[0]}x{dims[1]}x{dims[2]}, "
                     + ("exact distance everywhere" if w["band_voxels"] is None else f"band maxDist={w['band_voxels']} voxels")
                     + f", nullOutsideMinMax={w['null_outside']}")


class ClockSampler:
    """samples nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)"""

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for k, n in enumerate(names):
                if len(r) > 3 + k and r[3 + k].lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def flops_per_voxel(c, n):
    """8*Vf + 10*Af + 80*Lf + 12*Bd + 60*Ld, counts of the REFERENCE algorithm on the reference tree"""
    return (8 * c["fwn_visits"] + 10 * c["fwn_accepts"] + 80 * c["fwn_leaves"] + 12 * c["box_tests"] + 60 * c["tri_tests"]) / max(n, 1)


def sample_layers(dz, k):
    return sorted(set(int(round((i + 0.5) * dz / k - 0.5)) for i in range(k)))


# ---------------------------------------------------------------------------------------------------------------------
# CPU side: the reference's own implementation (checker code; only this leg of bench.py may execute oracle/)
# ---------------------------------------------------------------------------------------------------------------------

def cpu_reference_setup(wl):
    from oracle import portlib, reflib
    t0 = time.perf_counter()
    if reflib.available():
        m = reflib.RefMesh.from_arrays(wl["points"], wl["tris"])
        kind, cores = "reference", reflib.hardware_threads()
        m.nodes()  # builds the AABB tree once (Mesh::getAABBTree), as the first call would
        m.dipoles()
    else:
        m = portlib.PortMesh(wl["points"], wl["tris"])
        kind, cores = "port", os.cpu_count()
    return m, kind, cores, time.perf_counter() - t0


def cpu_run_layers(m, wl, layers, keep=None):
    """times the reference on the given z-layers; keep (a dict) receives {z: the reference's values of that layer}"""
    o = wl["opts"]
    t0 = time.perf_counter()
    for z in layers:
        v = m.calc_grid_with_distances_layers(wl["dims"], wl["xf"], z, z + 1, min_dist_sq=o["min_dist_sq"], max_dist_sq=o["max_dist_sq"],
                                              null_outside=o["null_outside"], threshold=o["threshold"], beta=o["beta"])
        if keep is not None:
            keep[z] = v
    return time.perf_counter() - t0


PARITY_LAYERS = 16  # candidates: the GPU copies of these layers are kept for the parity block


def choose_sample(m, wl, budget_s):
    """evenly spaced z-layers worth about budget_s seconds of CPU work (probe: the middle layer); always a subset of
    sample_layers(dz, PARITY_LAYERS), whose GPU values the bench keeps"""
    dz = wl["dims"][2]
    cand = sample_layers(dz, min(dz, PARITY_LAYERS))
    t1 = cpu_run_layers(m, wl, [cand[len(cand) // 2]])
    k = int(max(1, min(len(cand), budget_s / max(t1, 1e-6))))
    return [cand[int((i + 0.5) * len(cand) / k)] for i in range(k)], t1


def voxel_centres(wl, z, idx):
    """float32 voxel centres xf(x, y, z) of the voxels idx (flat indices inside layer z)"""
    dx = wl["dims"][0]
    xf = np.asarray(wl["xf"], np.float32)
    pos = np.stack([(idx % dx), (idx // dx), np.full(len(idx), z)], 1).astype(np.float32)
    return (pos @ xf[:9].reshape(3, 3).T + xf[9:]).astype(np.float32)


def parity_block(m, wl, ref_layers, gpu_layers):
    """north_star's parity rules (oracle/checker.compare_sdf) on the sampled layers: the reference's values (computed for the
    cpu_baseline anyway) against the same layers of the GPU volume.  Sign mismatches are excused only where the reference's
    own winding number of that voxel lies within 1e-3 of the threshold."""
    from oracle import checker
    tot = {"layers": 0, "n": 0, "nan_mismatch": 0, "sign_mismatch": 0, "sign_mismatch_in_band": 0, "dist_mismatch": 0,
           "dist_bit_exact": 0, "finite": 0, "max_rel_err": 0.0}
    o = wl["opts"]
    for z, want in ref_layers.items():
        got = gpu_layers[z]
        w_ref = None
        both = ~np.isnan(got) & ~np.isnan(want)
        diff = np.flatnonzero(both & (np.signbit(got) != np.signbit(want)))
        if len(diff):
            w_ref = np.full(want.shape, 10.0, np.float32)  # far from any threshold: only the differing voxels can be excused
            w_ref[diff] = m.calc_from_vector(voxel_centres(wl, z, diff), beta=o["beta"])
        rep = checker.compare_sdf(got, want, w_ref, threshold=o["threshold"])
        tot["layers"] += 1
        tot["n"] += rep["n"]
        for k in ("nan_mismatch", "sign_mismatch", "dist_mismatch"):
            tot[k] += rep[k]
        tot["sign_mismatch_in_band"] += rep.get("sign_mismatch_in_band", 0)
        tot["finite"] += int(both.sum())
        tot["dist_bit_exact"] += int((np.abs(got[both]).view(np.uint32) == np.abs(want[both]).view(np.uint32)).sum())
        tot["max_rel_err"] = max(tot["max_rel_err"], rep["max_rel_err"])
    tot["dist_bit_exact_frac"] = tot.pop("dist_bit_exact") / max(tot["finite"], 1)
    tot["ok"] = tot["nan_mismatch"] == 0 and tot["sign_mismatch"] == 0 and tot["dist_mismatch"] == 0
    tot["rule"] = "NaN masks equal; sign equal outside |w_ref - threshold| < 1e-3; |d| within 1e-5 relative (oracle/checker.py)"
    return tot


def cpu_baseline(wl, budget_s=15.0, gpu_layers=None):
    m, kind, cores, setup_s = cpu_reference_setup(wl)
    layers, _ = choose_sample(m, wl, budget_s)
    keep = {} if gpu_layers is not None else None
    dt = cpu_run_layers(m, wl, layers, keep)
    nvox = len(layers) * wl["dims"][0] * wl["dims"][1]
    parity = parity_block(m, wl, keep, gpu_layers) if keep else None
    return {"value": nvox / dt, "unit": UNIT, "cores": cores, "kind": kind, "parity": parity,
            "sample": f"{len(layers)} evenly spaced z-layers of {wl['dims'][0]}x{wl['dims'][1]} ({nvox} voxels, {dt:.2f} s); "
                      f"mesh+tree setup {setup_s:.1f} s excluded", "layers": layers}


def ref_cuda_baseline(wl, layers):
    """SURVEY.md 8(d) second baseline: the reference's own CUDA kernels (oracle/_ref/libmrrefcuda.so, unmodified
    MRCudaFastWindingNumber.cu for sm_100, -use_fast_math) on this GPU, fed the reference's CPU-built tree, on the same
    bounded sample of z-layers as the CPU baseline; kernel time by CUDA events"""
    from oracle import refcudalib, reflib
    if not (refcudalib.available() and reflib.available()):
        return None
    m = reflib.RefMesh.from_arrays(wl["points"], wl["tris"])
    ref = refcudalib.RefCudaFwn(m.nodes(), m.dipoles(), m.points, m.tris)  # what prepareData_ uploads
    o = wl["opts"]
    kw = dict(min_dist_sq=o["min_dist_sq"], max_dist_sq=o["max_dist_sq"], null_outside=o["null_outside"], threshold=o["threshold"],
              beta=o["beta"], fetch=False)
    ref.signed_distance_layers(wl["dims"], wl["xf"], layers[0], layers[0] + 1, **kw)  # warm-up, not counted
    ms = sum(ref.signed_distance_layers(wl["dims"], wl["xf"], z, z + 1, **kw)[0] for z in layers)
    ref.close()
    nvox = len(layers) * wl["dims"][0] * wl["dims"][1]
    return {"value": nvox / (ms * 1e-3), "unit": UNIT, "kind": "reference CUDA kernels (MRCudaFastWindingNumber.cu, sm_100, -use_fast_math), kernel time only",
            "sample": f"{len(layers)} z-layers ({nvox} voxels, {ms:.1f} ms)"}


def algorithmic_flops(wl, layers, max_samples=200000):
    """operation counts of the reference algorithm (C restatement, counters) on a stride sample of the given layers"""
    from oracle import portlib
    m = portlib.PortMesh(wl["points"], wl["tris"])
    o = wl["opts"]
    per_layer = wl["dims"][0] * wl["dims"][1]
    stride = max(1, (per_layer * len(layers)) // max_samples) | 1
    n = 0
    m.reset_counters()
    for z in layers:
        m.calc_from_grid_with_distances(wl["dims"], wl["xf"], min_dist_sq=o["min_dist_sq"], max_dist_sq=o["max_dist_sq"],
                                        null_outside=o["null_outside"], threshold=o["threshold"], beta=o["beta"], z0=z, z1=z + 1,
                                        stride=stride)
        n += (per_layer + stride - 1) // stride
    c = m.counter_dict()
    return flops_per_voxel(c, n), {k: v / n for k, v in c.items()}, n


def run_reference_arm(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    m, kind, cores, setup_s = cpu_reference_setup(wl)
    layers, _ = choose_sample(m, wl, 6.0)
    nvox = len(layers) * wl["dims"][0] * wl["dims"][1]
    for _ in range(args.warmup):
        cpu_run_layers(m, wl, layers[:1])
    t = 0.0
    for _ in range(args.steps):
        t += cpu_run_layers(m, wl, layers)
    v = nvox * args.steps / t
    sample = f"each step = {len(layers)} evenly spaced z-layers ({nvox} voxels) of the workload; mesh+tree setup {setup_s:.1f} s excluded"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["desc"], "timing": "host wall clock, all host threads (TBB)"},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


# ---------------------------------------------------------------------------------------------------------------------
# query workloads: winding numbers of point vectors (C2) and of face centres (C4); the metric's second half, single GPU
# ---------------------------------------------------------------------------------------------------------------------
QUERY_WORKLOADS = {
    # C2: calcFromVector, 10 M random points in [-1.5, 1.5]^3 against the 1 M-triangle UV sphere
    "c2": dict(kind="vector", mesh=("uv_sphere", 1000, 500), n=10_000_000),
    # C4: calcSelfIntersections of two interpenetrating tori, 5.12 M triangles (one query per face)
    "c4": dict(kind="faces", mesh=("two_tori", 1600, 800)),
}


def run_queries(args, name):
    import torch
    import meshlib_b200 as mb
    w = QUERY_WORKLOADS[name]
    if w["mesh"][0] == "uv_sphere":
        pts, tris = meshgen.make_uv_sphere(1, w["mesh"][1], w["mesh"][2])
    else:
        pts, tris = meshgen.two_tori(2.0, 0.7, w["mesh"][1], w["mesh"][2])
    torch.cuda.set_device(0)
    fwn = mb.FastWindingNumber(pts, tris)
    vector = w["kind"] == "vector"
    if vector:
        q = meshgen.query_points(w["n"])
        n = len(q)
        dq = torch.from_numpy(q).cuda()
        dout = torch.empty(n, dtype=torch.float32, device="cuda")
        step = lambda: fwn.calc_from_vector_device(dq.data_ptr(), n, dout.data_ptr())
    else:
        n = len(tris)
        dbits = torch.zeros((n + 63) // 64, dtype=torch.int64, device="cuda")
        dout = torch.empty(n, dtype=torch.float32, device="cuda")
        step = lambda: fwn.calc_self_intersections_device(dbits.data_ptr(), dout.data_ptr())
    for _ in range(args.warmup):
        step()
    sampler = ClockSampler(0)
    sampler.start()
    l0 = fwn.stats().kernel_launches
    ms = 0.0
    for _ in range(args.steps):
        step()
        ms += fwn.stats().last_kernel_ms
    clocks = sampler.stop()
    launches = fwn.stats().kernel_launches - l0
    ms /= args.steps
    got_w = dout.cpu().numpy()
    got_bits = None if vector else dbits.cpu().numpy().view(np.uint64)
    st = fwn.stats()
    fwn.close()
    # e2e: host mesh + host queries -> create -> call with host buffers (pageable) -> destroy
    e_steps = max(1, min(args.steps, 3))
    for it in range(e_steps + 1):  # the first round is a warm-up (staging ring, allocator state)
        if it == 1:
            t0 = time.perf_counter()
        f2 = mb.FastWindingNumber(pts, tris)
        res = f2.calc_from_vector(q) if vector else f2.calc_self_intersections()
        f2.close()
    e_ms = 1e3 * (time.perf_counter() - t0) / e_steps
    fp32_peak = mb.measure_fp32_peak(0)
    desc = (f"{name}: calcFromVector, {n} random points vs {len(tris)} triangles (UV sphere)" if vector else
            f"{name}: calcSelfIntersections, {len(tris)} triangles (two interpenetrating tori), one query per face")
    out = {"metric": "winding_number_queries_per_sec", "value": n / (ms * 1e-3), "unit": "queries/s", "n_gpus": 1, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
           "data": "synthetic",
           "config": {"workload": desc, "timing": "CUDA events on the launching stream (incl. the Morton sort of the queries)",
                      "l2": "tree+triangles %.0f MB, queries %.0f MB per step" % ((st.num_nodes * 64 + st.num_leaves * 48) / 1e6, (12 if vector else 0) * n / 1e6 + 4 * n / 1e6)},
           "clocks": clocks, "gpu_launches": int(launches), "tree_build_ms": st.tree_build_ms,
           "e2e": {"value": n / (e_ms * 1e-3), "unit": "queries/s", "ms_per_step": e_ms, "steps": e_steps,
                   "h2d_bytes_per_step": int(pts.nbytes + tris.nbytes + (q.nbytes if vector else 0)),
                   "d2h_bytes_per_step": int(4 * n if vector else 8 * ((n + 63) // 64)),
                   "includes": "b2fwn_create (H2D + device tree build) + the host-buffer call (pageable numpy) + b2fwn_destroy; host wall clock"},
           "roofline": {"bound": "fp32", "achieved": None, "peak": fp32_peak, "unit": "TFLOP/s", "frac": None, "traffic": None,
                        "peak_source": "FFMA micro-benchmark run on this GPU just now (b2fwn_measure_fp32_peak)"}}
    if not args.no_cpu:
        from oracle import checker, portlib, reflib
        orc = checker.best_oracle(pts, tris)
        # bounded sample of the same queries on all host threads
        if vector:
            k = 200_000
            orc.calc_from_vector(q[:k])
            t0 = time.perf_counter()
            want = orc.calc_from_vector(q[:k])
            t1 = time.perf_counter() - t0
            k2 = int(min(n, max(k, k * args.cpu_budget / max(t1, 1e-6))))
            idx = np.arange(0, n, max(1, n // k2))[:k2]
            t0 = time.perf_counter()
            want = orc.calc_from_vector(q[idx])
            c_s = time.perf_counter() - t0
            rep = checker.compare_w(got_w[idx], want, atol=2e-6)
            out["parity"] = {"n": int(len(idx)), "max_abs_dw": rep["max_abs_diff"], "bit_exact_frac": rep["bit_exact_frac"], "ok": rep["ok"],
                             "rule": "|w - w_ref| <= 2e-6 on the sampled queries"}
            sample, nq = f"{len(idx)} of the {n} queries (every {max(1, n // k2)}th), {c_s:.2f} s", len(idx)
        else:
            t0 = time.perf_counter()
            want = orc.calc_self_intersections()
            c_s = time.perf_counter() - t0
            rep = checker.compare_bits(got_bits, want, got_w, len(tris))
            out["parity"] = {"n": int(len(tris)), "mismatch_outside_bands": rep["mismatch"], "mismatch_in_bands": rep["mismatch_in_band"],
                             "set_bits": rep["set"], "ok": rep["ok"], "rule": "bit f equal wherever w is not within 1e-3 of 0 or 1"}
            sample, nq = f"all {len(tris)} faces once, {c_s:.2f} s", len(tris)
        out["cpu_baseline"] = {"value": nq / c_s, "unit": "queries/s", "cores": reflib.hardware_threads() if reflib.available() else 1,
                               "kind": getattr(orc, "kind", "port"), "sample": sample}
        # reference-algorithm operation counts on a sample (C restatement with counters): 8 Vf + 10 Af + 80 Lf per query
        m = portlib.PortMesh(pts, tris)
        m.reset_counters()
        ks = 50_000
        if vector:
            m.calc_from_vector(q[:: max(1, n // ks)][:ks])
            cnt_n = len(q[:: max(1, n // ks)][:ks])
        else:
            P, T = m_points(pts), tris[:: max(1, len(tris) // ks)][:ks]
            centres = ((P[T[:, 0]] + P[T[:, 1]] + P[T[:, 2]]) * np.float32(1 / 3.0)).astype(np.float32)
            m.calc_from_vector(centres)
            cnt_n = len(centres)
        c = m.counter_dict()
        fpq = (8 * c["fwn_visits"] + 10 * c["fwn_accepts"] + 80 * c["fwn_leaves"]) / max(cnt_n, 1)
        ach = fpq * n / (ms * 1e-3) / 1e12
        out["roofline"].update({"achieved": ach, "frac": ach / fp32_peak, "flops_per_query": fpq,
                                "ops_per_query": {k2_: v / cnt_n for k2_, v in c.items() if k2_.startswith("fwn")},
                                "ops_sample": f"{cnt_n} queries, reference algorithm counters (face-centre sample without the skipped face for c4)"})
    print(json.dumps(out))


def m_points(pts):
    return np.asarray(pts, np.float32)


# ---------------------------------------------------------------------------------------------------------------------
# offset workloads: distance band + marching cubes (north_star subsystem 4), single GPU
# ---------------------------------------------------------------------------------------------------------------------

def run_offset(args, name):
    import torch
    import meshlib_b200 as mb
    w = OFFSET_WORKLOADS[name]
    spec = w["mesh"]
    pts, tris = meshgen.make_uv_sphere(1, spec[1], spec[2])
    pts = meshgen.add_noise(pts, spec[3], seed=42)
    torch.cuda.set_device(0)
    fwn = mb.FastWindingNumber(pts, tris)
    for _ in range(args.warmup):
        fwn.mc_offset_mesh(w["offset"], w["voxel"], close_holes=w["close_holes"], fetch=False)
    sampler = ClockSampler(0)
    sampler.start()
    l0 = fwn.stats().kernel_launches
    sdf = mc = 0.0
    for _ in range(args.steps):
        _v, _t, info = fwn.mc_offset_mesh(w["offset"], w["voxel"], close_holes=w["close_holes"], fetch=False)
        sdf += info["sdf_ms"]
        mc += info["mc_ms"]
    clocks = sampler.stop()
    launches = fwn.stats().kernel_launches - l0
    sdf, mc = sdf / args.steps, mc / args.steps
    dims = info["dims"]
    total = dims[0] * dims[1] * dims[2]
    # e2e: host mesh -> create -> offset -> TriMesh on the host (pageable numpy), create + fetch + destroy inside
    fwn.close()
    e_steps = max(1, min(args.steps, 3))
    for it in range(e_steps + 1):  # the first round is a warm-up
        if it == 1:
            t0 = time.perf_counter()
        f2 = mb.FastWindingNumber(pts, tris)
        v, t, info2 = f2.mc_offset_mesh(w["offset"], w["voxel"], close_holes=w["close_holes"], fetch=True)
        f2.close()
    e_ms = 1e3 * (time.perf_counter() - t0) / e_steps
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak, hbm_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (6650.0, "fallback")
    out_bytes = 12 * info["n_verts"] + 12 * info["n_tris"]
    mc_bytes = 10.5 * total + out_bytes  # value read twice, code written + read, segment totals; + the compacted output
    desc = (f"{name}: mcOffsetMesh of {len(tris)} faces, voxel {w['voxel']:.6g}, offset {w['offset']}, grid {dims[0]}x{dims[1]}x{dims[2]}, "
            f"HoleWindingRule, closeHoles={w['close_holes']}, incl. marching cubes")
    out = {"metric": METRIC, "value": total / ((sdf + mc) * 1e-3), "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": sdf + mc, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": desc, "timing": "CUDA events on the launching stream (b2fwn_offset_result.sdf_ms + mc_ms)",
                      "l2": "volume %.0f MB per step, larger than L2" % (4 * total / 1e6)},
           "sdf_ms": sdf, "mc_ms": mc, "n_verts": int(info["n_verts"]), "n_tris": int(info["n_tris"]), "clocks": clocks, "gpu_launches": int(launches),
           "e2e": {"value": total / (e_ms * 1e-3), "unit": UNIT, "ms_per_step": e_ms, "steps": e_steps,
                   "h2d_bytes_per_step": int(pts.nbytes + tris.nbytes), "d2h_bytes_per_step": int(out_bytes),
                   "includes": "b2fwn_create + b2fwn_mc_offset_mesh + b2fwn_marching_cubes_fetch into pageable numpy arrays + b2fwn_destroy; host wall clock"},
           "roofline": {"bound": "hbm", "kernel": "kMcClassify + scans + kMcEmit (marching cubes stage)", "achieved": mc_bytes / (mc * 1e-3) / 1e9,
                        "peak": hbm_peak, "peak_source": hbm_src, "unit": "GB/s", "frac": mc_bytes / (mc * 1e-3) / 1e9 / hbm_peak,
                        "algorithmic_bytes": mc_bytes, "traffic": None}}
    if not args.no_cpu:
        from oracle import checker
        orc = checker.best_oracle(pts, tris)
        t0 = time.perf_counter()
        wv, wt, _sdf, worigin, wdims = checker.ref_mc_offset_mesh(orc, pts, orc.tris if hasattr(orc, "tris") else tris, w["offset"], w["voxel"],
                                                                 close_holes=w["close_holes"])
        c_s = time.perf_counter() - t0
        from oracle import reflib
        out["cpu_baseline"] = {"value": total / c_s, "unit": UNIT, "cores": reflib.hardware_threads() if reflib.available() else 1,
                               "kind": getattr(orc, "kind", "port"), "sample": f"the whole pipeline once: reference SDF band + reference marching cubes, {c_s:.1f} s"}
        same_tris = (len(t) == len(wt)) and bool(np.array_equal(t, wt)) if hasattr(orc, "tris") and np.array_equal(orc.tris, tris) else None
        out["parity"] = {"dims_equal": tuple(wdims) == tuple(info2["dims"]), "origin_equal": bool(np.array_equal(worigin, info2["origin"])),
                         "n_verts": [int(len(v)), int(len(wv))], "n_tris": [int(len(t)), int(len(wt))], "triangles_identical": same_tris,
                         "max_vertex_diff": float(np.abs(v - wv).max()) if len(v) == len(wv) else None}
    print(json.dumps(out))


# ---------------------------------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------------------------------

def run_ours(args, wl):
    import torch
    import torch.distributed as dist
    import meshlib_b200 as mb

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun --nproc-per-node {args.gpus}")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)

    dims, xf = wl["dims"], wl["xf"]
    o = wl["opts"]
    opts = mb.DistOptions(o["min_dist_sq"], o["max_dist_sq"], o["null_outside"], o["threshold"], o["beta"])
    layer = dims[0] * dims[1]
    total = layer * dims[2]
    z0, z1 = slabs.slab_range(rank, world, dims[2])

    fwn = mb.FastWindingNumber(wl["points"], wl["tris"], device=local)
    st0 = fwn.stats()
    # full volume on every GPU (north_star: the slabs of all ranks gathered on every GPU); rank r writes its share in place
    p2p = world > 1 and args.gather in ("p2p", "ce")  # both need the peer-mapped volumes
    ce = world > 1 and args.gather == "ce"
    peer_ptrs, my_ptr = [], None
    if p2p:
        # peer-visible volumes: every rank maps the volumes of the others; the kernels store each result into all of them.
        # Where CUDA IPC / peer access is not available (some container set-ups) every rank falls back to the NCCL gather.
        ok, err, token = 1, "", b""
        try:
            my_ptr, token = fwn.shared_alloc(4 * total)
        except mb.B2Error as e:
            ok, err = 0, str(e)
        tokens = [None] * world
        dist.all_gather_object(tokens, (ok, token))
        if all(t[0] for t in tokens):
            try:
                peer_ptrs = [fwn.shared_open(tokens[r][1]) for r in range(world) if r != rank]
            except mb.B2Error as e:
                ok, err = 0, str(e)
        flag = torch.tensor([ok], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag[0]) == 0:
            if rank == 0:
                print(f"bench.py: peer-mapped volumes unavailable ({err or 'another rank failed'}); using --gather nccl", file=sys.stderr)
            for q in peer_ptrs:
                fwn.shared_close(q)
            if my_ptr:
                fwn.shared_free(my_ptr)
            p2p, ce, peer_ptrs, my_ptr = False, False, [], None
    if p2p:
        volume = torch.as_tensor(mb.DeviceBuffer(my_ptr, total), device=dev)
    else:
        volume = torch.empty(total, dtype=torch.float32, device=dev)
    slab = volume[z0 * layer:z1 * layer]

    host_group = dist.new_group(backend="gloo") if world > 1 else None

    def barrier(host_only=False):
        torch.cuda.synchronize()
        if world > 1:
            if host_only:
                # the ranks that wait must not occupy their GPU: an NCCL barrier spins in a kernel, and rank 0 drives ALL GPUs in
                # the e2e leg (two processes on one GPU are time-sliced: r02 measured 1322 ms instead of ~750 ms at N = 2)
                dist.barrier(group=host_group)
            else:
                dist.barrier()
            torch.cuda.synchronize()


    # N > 1: z-interleaved chunks (equal cost per rank).  --gather p2p (default): one call per step, the gather fused into the
    # kernel (stores into every rank's volume over NVLink), then a barrier.  --gather nccl: the z-range cut into stages, the
    # all-gather of one stage issued on NCCL's stream while the next stage is computed.  N = 1: one plain call.
    plan = slabs.ChunkPlan(world, dims[2], chunk=args.z_chunk, stages=1 if (p2p and not ce) else args.stages) if world > 1 else None
    if ce:
        # --gather ce: the kernels write the rank's own volume only; every finished stage's chunks (contiguous blocks of z_chunk
        # layers) are pushed into the peers' volumes by the COPY ENGINES (cudaMemcpyAsync on the peer-mapped pointers, two copy
        # streams) while the next stage is computed -- no SM issues a remote store; exposed: the last stage's copies + barrier
        from cuda.bindings import runtime as cudart
        copy_streams = [torch.cuda.Stream(device=dev) for _ in range(max(1, min(args.copy_streams, world - 1)))]
        D2D = cudart.cudaMemcpyKind.cudaMemcpyDeviceToDevice
    if p2p and plan.call_args(rank, 0) is None:
        raise SystemExit("grid too small for this many ranks")

    def step():
        """returns (kernel ms on the launching stream, ms between the last kernel's end and the last gather's end)"""
        if world == 1:
            fwn.calc_from_grid_with_distances_device(dims, xf, slab.data_ptr(), opts, z_begin=z0, z_end=z1)
            return fwn.stats().last_kernel_ms, 0.0
        if ce:
            k_ms, n_copy = 0.0, 0
            for s in range(plan.stages):
                a = plan.call_args(rank, s)
                if not a:
                    continue
                zb, ze, zc, zs = a
                fwn.calc_from_grid_with_distances_device_chunks(dims, xf, my_ptr + 4 * zb * layer, opts, z_begin=zb, z_end=ze,
                                                                z_chunk=zc, z_stride=zs)  # returns when the kernels are done
                k_ms += fwn.stats().last_kernel_ms
                for z in range(zb, ze, zs):
                    off, nbytes = 4 * z * layer, 4 * layer * min(zc, ze - z)
                    for i in range(len(peer_ptrs)):
                        q = peer_ptrs[(i + rank) % len(peer_ptrs)]  # the ranks start with different peers
                        err, = cudart.cudaMemcpyAsync(q + off, my_ptr + off, nbytes, D2D, copy_streams[n_copy % len(copy_streams)].cuda_stream)
                        if err != cudart.cudaError_t.cudaSuccess:
                            raise RuntimeError(f"cudaMemcpyAsync to a peer volume: {err}")
                        n_copy += 1
            t = time.perf_counter()
            for cs in copy_streams:
                cs.synchronize()
            dist.barrier()  # every rank's copies have landed: the volume is complete everywhere
            torch.cuda.synchronize()
            return k_ms, 1e3 * (time.perf_counter() - t)
        if p2p:
            zb, ze, zc, zs = plan.call_args(rank, 0)
            off = 4 * zb * layer
            fwn.calc_from_grid_with_distances_device_chunks_p2p(dims, xf, my_ptr + off, [q + off for q in peer_ptrs], opts, z_begin=zb,
                                                                z_end=ze, z_chunk=zc, z_stride=zs)  # returns when the kernels are done
            t = time.perf_counter()
            dist.barrier()  # every rank's stores have landed: the volume is complete everywhere
            torch.cuda.synchronize()
            return fwn.stats().last_kernel_ms, 1e3 * (time.perf_counter() - t)
        k_ms, works = 0.0, []
        for s in range(plan.stages):
            a = plan.call_args(rank, s)
            if a:
                zb, ze, zc, zs = a
                fwn.calc_from_grid_with_distances_device_chunks(dims, xf, volume.data_ptr() + 4 * zb * layer, opts, z_begin=zb,
                                                                z_end=ze, z_chunk=zc, z_stride=zs)  # returns when the kernels are done
                k_ms += fwn.stats().last_kernel_ms
            works += plan.gather_stage(volume, rank, s, dims, dist, async_op=True)
        t = time.perf_counter()
        for w in works:
            w.wait()
        torch.cuda.synchronize()
        return k_ms, 1e3 * (time.perf_counter() - t)

    for _ in range(args.warmup):
        step()
    launches0 = fwn.stats().kernel_launches
    sampler = ClockSampler(local)
    barrier()
    if rank == 0:
        sampler.start()
    t0 = time.perf_counter()
    k_sum = g_sum = 0.0
    for _ in range(args.steps):
        k, g = step()
        k_sum += k
        g_sum += g
    barrier()
    wall_ms = 1e3 * (time.perf_counter() - t0)
    clocks = sampler.stop() if rank == 0 else None
    launches = fwn.stats().kernel_launches - launches0
    per_rank_kernel_ms = None
    if world > 1:  # the slowest rank must be visible
        g = [torch.zeros(1, dtype=torch.float64, device=dev) for _ in range(world)]
        dist.all_gather(g, torch.tensor([k_sum / args.steps], dtype=torch.float64, device=dev))
        per_rank_kernel_ms = [float(x[0]) for x in g]
    t = torch.tensor([k_sum + g_sum, k_sum, g_sum, wall_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, k_ms, g_ms, wall_ms = (float(x) / args.steps for x in t.tolist())
    if world > 1:
        dev_ms = wall_ms  # gathers overlap the kernels: the step is the bracketed region (device synchronised on both sides)
    checksum = float(torch.nan_to_num(volume[:: max(1, total // (1 << 22))]).double().abs().sum())
    # parity block: the GPU's values of the candidate sample layers, compared below with the reference's values of the same layers
    gpu_layers = None
    if rank == 0 and not args.no_cpu:
        gpu_layers = {z: volume[z * layer:(z + 1) * layer].cpu().numpy() for z in sample_layers(dims[2], min(dims[2], PARITY_LAYERS))}
    if world > 1:  # every rank must hold the same, complete volume
        c = torch.tensor([checksum, -checksum], dtype=torch.float64, device=dev)
        dist.all_reduce(c, op=dist.ReduceOp.MAX)
        if float(c[0]) != -float(c[1]):
            raise SystemExit(f"volumes differ between ranks: checksum range [{-float(c[1])}, {float(c[0])}]")

    # ---- e2e: host mesh -> create (H2D + device tree build) -> host volume, all inside the timed region --------------
    # ONE process drives all N GPUs through the library's own multi-device handle (b2fwn_config.n_devices): what a MeshLib caller
    # gets from MR::B200::FastWindingNumber( mesh, devices ).  Under torchrun the other ranks release their GPUs and wait.
    # The output buffer is PAGEABLE host memory that the caller has sized and touched (a std::vector<float> after resize);
    # the same call into pinned memory is reported beside it.
    e2e = None
    if not args.no_e2e:
        del volume, slab
        if p2p:
            barrier()
            for q in peer_ptrs:
                fwn.shared_close(q)
            barrier()
            fwn.shared_free(my_ptr)
        fwn.close()
        torch.cuda.empty_cache()
        barrier()
        barrier(host_only=True)
        if rank == 0:
            devices = list(range(world))
            # every round is timed on its own and the MEDIAN is the e2e figure: one stalled cudaMalloc inside b2fwn_create (2 s, once,
            # in the r02 driver run) moved a three-round mean by 39 %; max and all rounds are reported beside it
            e_steps = max(1, min(args.steps, 5))
            e_rounds = {}

            def run(out, steps, key=None):
                ms = []
                for _ in range(steps):
                    t0 = time.perf_counter()
                    f2 = mb.FastWindingNumber(wl["points"], wl["tris"], devices=devices)
                    f2.calc_from_grid_with_distances(dims, xf, opts, out=out)
                    st = f2.stats()
                    f2.close()
                    ms.append(1e3 * (time.perf_counter() - t0))
                if key:
                    e_rounds[key] = ms
                return float(np.median(ms)), st

            host_np = np.empty(total, np.float32)
            host_np.fill(0)  # touched, like the vector a caller resized
            run(host_np, 1)  # warm-up: contexts on the other devices, staging rings
            e_ms, st2 = run(host_np, e_steps, "pageable")
            # the host-output call must deliver what the device-resident call left in HBM (same kernels, chunked): bit for bit
            e2e_same = None
            if gpu_layers is not None:
                e2e_same = all(np.array_equal(host_np[z * layer:(z + 1) * layer].view(np.uint32), g.view(np.uint32)) for z, g in gpu_layers.items())
            e2e_checksum = float(np.nan_to_num(host_np[:: max(1, total // (1 << 22))]).astype(np.float64).__abs__().sum())
            del host_np
            pinned_ms = None
            if not args.no_pinned and (world == 1 or args.pinned):  # N > 1: only on request (--pinned), it pins 34 GB once more
                host = torch.empty(total, dtype=torch.float32, pin_memory=True)
                run(host.numpy(), 1)
                pinned_ms, _ = run(host.numpy(), min(e_steps, 3), "pinned")
                del host
            e2e = {"value": total / (e_ms * 1e-3), "unit": UNIT, "ms_per_step": e_ms, "steps": e_steps,
                   "statistic": "median over the rounds, each timed on its own (host wall clock)",
                   "ms_max": max(e_rounds["pageable"]), "ms_min": min(e_rounds["pageable"]),
                   "ms_mean": float(np.mean(e_rounds["pageable"])), "ms_rounds": [round(v, 1) for v in e_rounds["pageable"]],
                   "h2d_bytes_per_step": int(wl["points"].nbytes + wl["tris"].nbytes + 12 * 4 + 20),
                   "d2h_bytes_per_step": int(4 * total), "equals_resident_volume_on_sample_layers": e2e_same, "checksum": e2e_checksum,
                   "host_buffer": "pageable (numpy, sized and touched by the caller): pinned staging ring + copier threads inside the library",
                   "pinned_buffer": None if pinned_ms is None else {"value": total / (pinned_ms * 1e-3), "ms_per_step": pinned_ms,
                                                                    "ms_rounds": [round(v, 1) for v in e_rounds["pinned"]]},
                   "kernel_ms_of_the_call": st2.last_kernel_ms, "tree_build_ms": st2.tree_build_ms,
                   "includes": f"one process, {world} GPU(s): b2fwn_create (mesh H2D + device tree/dipole build"
                               + (", on every GPU at once" if world > 1 else "")
                               + ") + b2fwn_calc_from_grid_with_distances into the host buffer + b2fwn_destroy; host wall clock"}
        barrier(host_only=True)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    value = total / (dev_ms * 1e-3)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak, hbm_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (6650.0, "fallback")
    fp32_peak = mb.measure_fp32_peak(local)
    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dev_ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": wl["desc"], "sharding": "single GPU" if world == 1 else
                   (f"z-interleaved chunks of {args.z_chunk} layers x{world} ranks in {plan.stages} stages, mesh+tree replicated, every "
                    "finished stage pushed into the peers' volumes by the copy engines (cudaMemcpyAsync over NVLink) behind the next "
                    "stage's kernels + barrier") if ce else
                   (f"z-interleaved chunks of {args.z_chunk} layers x{world} ranks, mesh+tree replicated, gather fused into the kernel "
                    "(peer stores over NVLink into every rank's volume) + barrier") if p2p else
                   (f"z-interleaved chunks of {args.z_chunk} layers x{world} ranks in {plan.stages} stages, mesh+tree replicated, NCCL "
                    "all-gather of each stage issued behind the next stage's kernels"),
                   "l2": "inputs larger than L2 (tree+triangles %.0f MB, output %.0f MB per step)" % (
                       (st0.num_nodes * 64 + st0.num_leaves * 64) / 1e6, 4 * total / 1e6),
                   "timing": ("CUDA events on the launching stream" if world == 1 else
                              "host clock around the K steps, device synchronised and ranks barriered on both sides") + ", max over ranks"},
        "kernel_ms_per_step": k_ms, "kernel_ms_per_rank": per_rank_kernel_ms, "gather_exposed_ms_per_step": g_ms, "wall_ms_per_step": wall_ms,
        "tree_build_ms": st0.tree_build_ms, "num_faces": int(st0.num_leaves), "num_nodes": int(st0.num_nodes),
        "clocks": clocks, "gpu_launches": int(launches), "checksum": checksum,
    }
    if e2e:
        out["e2e"] = e2e
    # roofline of the step (kGridDist + kGridFwn6, the latter ~84 % of it): FP32 pipe, per north_star; HBM figures beside it
    out["roofline"] = {"bound": "fp32", "achieved": None, "peak": fp32_peak, "unit": "TFLOP/s", "frac": None, "traffic": None,
                       "peak_source": "FFMA micro-benchmark run on this GPU just now (b2fwn_measure_fp32_peak)",
                       "hbm": {"algorithmic_bytes_per_voxel": 4.0 + (st0.num_nodes * 64 + st0.num_leaves * 64) * world / total,
                               "peak": hbm_peak, "peak_source": hbm_src, "unit": "GB/s"}}
    if world == 1 and not args.grid:
        traffic, source = ncu_traffic(wl["name"])
        if traffic is not None:
            out["roofline"]["traffic"], out["roofline"]["traffic_source"] = traffic, source
        out["roofline"]["traffic_note"] = ("dominant kernel (the winding walk) alone; its algorithmic bytes are 8 B/voxel (|d| read, signed d written) "
                                           "+ one read of the tree")
    nslab0 = (dims[2] // world) * layer  # voxels of one launch (rank 0's slab)
    out["roofline"]["hbm"]["achieved"] = (4.0 * nslab0 + st0.num_nodes * 64 + st0.num_leaves * 64) / (k_ms * 1e-3) / 1e9
    out["roofline"]["hbm"]["frac"] = out["roofline"]["hbm"]["achieved"] / hbm_peak
    if world == 1 and not args.no_cpu:
        cb = cpu_baseline(wl, args.cpu_budget, gpu_layers)
        out["parity"] = cb.pop("parity")
        fpv, per_voxel, n = algorithmic_flops(wl, cb["layers"])
        layers_for_gpu_ref = cb["layers"]
        del cb["layers"]
        out["cpu_baseline"] = cb
        try:
            rc = ref_cuda_baseline(wl, layers_for_gpu_ref)
            if rc:
                out["ref_cuda_baseline"] = rc
        except Exception as e:  # a baseline, never a reason to lose the bench line
            out["ref_cuda_baseline"] = {"error": str(e)[:200]}
        ach = fpv * nslab0 / (k_ms * 1e-3) / 1e12
        out["roofline"].update({"achieved": ach, "frac": ach / fp32_peak, "flops_per_voxel": fpv, "ops_per_voxel": per_voxel,
                                "ops_sample": f"{n} voxels (stride sample of the cpu_baseline layers), reference algorithm counters"})
    if out["roofline"]["achieved"] is None and wl["name"] in FLOPS_PER_VOXEL and not args.grid:
        # N > 1 (or --no-cpu): the flop count per voxel is a property of the workload, not of the run -- the N = 1 figure with its
        # source; one rank's launch does 1/N of the grid in the slowest rank's kernel time, against ONE GPU's peak
        fpv, src = FLOPS_PER_VOXEL[wl["name"]]
        t_ms = max(per_rank_kernel_ms) if per_rank_kernel_ms else k_ms
        ach = fpv * (total / world) / (t_ms * 1e-3) / 1e12
        out["roofline"].update({"achieved": ach, "frac": ach / fp32_peak, "flops_per_voxel": fpv, "flops_per_voxel_source": src,
                                "per": "one rank's launch (1/N of the grid) in the slowest rank's kernel time, one GPU's peak"})
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS) + sorted(OFFSET_WORKLOADS) + sorted(QUERY_WORKLOADS))
    ap.add_argument("--grid", type=int, default=None, help="override the grid size (diagnostics only; not the headline)")
    ap.add_argument("--z-chunk", type=int, default=8, help="N > 1: layers per interleaved chunk (multiple of 4)")
    ap.add_argument("--gather", default="p2p", choices=["p2p", "ce", "nccl"],
                    help="N > 1: fused peer stores; copy-engine pushes of finished stages into the peers' volumes; or NCCL all-gather")
    ap.add_argument("--copy-streams", type=int, default=7, help="--gather ce: copy streams the pushes are dealt over (at most one per peer)")
    ap.add_argument("--stages", type=int, default=4, help="N > 1: compute/gather pipeline stages over z")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-pinned", action="store_true", help="skip the pinned-buffer variant of the e2e leg")
    ap.add_argument("--pinned", action="store_true", help="N > 1: run the pinned-buffer variant of the e2e leg as well")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-budget", type=float, default=15.0)
    args = ap.parse_args()
    if args.workload in QUERY_WORKLOADS:
        if args.impl == "reference" or args.gpus != 1:
            raise SystemExit("query workloads are single-GPU diagnostics of our arm")
        return run_queries(args, args.workload)
    if args.workload in OFFSET_WORKLOADS:
        if args.impl == "reference" or args.gpus != 1:
            raise SystemExit("offset workloads are single-GPU diagnostics of our arm")
        return run_offset(args, args.workload)
    wl = build_workload(args.workload, args.grid)
    if args.impl == "reference":
        run_reference_arm(args, wl)
    else:
        run_ours(args, wl)


if __name__ == "__main__":
    main()
