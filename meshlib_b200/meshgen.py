"""meshlib_b200.meshgen -- deterministic synthetic inputs (triangle soups) shared by tests and bench.py.

Restates the reference's primitive generators as numpy (same vertex and face numbering):
  make_torus      MR::makeTorus      /root/reference/source/MRMesh/MRTorus.cpp:9-59
  make_uv_sphere  MR::makeUVSphere   /root/reference/source/MRMesh/MRMakeSphereMesh.cpp:35-102
  add_noise       distribution of MR::addNoise (source/MRMesh/MRAddNoise.cpp): Gaussian displacement; our own
                  counter-based bit stream (both the oracle and the GPU read the SAME generated arrays).
  query_points    splitmix64 counter generator of SURVEY.md 8(d) C2.
Only float32 arrays leave this module: points (V,3) float32, tris (F,3) int32.
"""
import numpy as np

PI_F = np.float32(3.14159265358979323846)


def make_torus(R=2.0, r=1.0, p_res=256, s_res=128):
    R, r = np.float32(R), np.float32(r)
    i = np.arange(s_res, dtype=np.int64)[:, None]
    j = np.arange(p_res, dtype=np.int64)[None, :]
    a = (np.float32(2) * i.astype(np.float32) * PI_F / np.float32(s_res)).astype(np.float32)
    b = (np.float32(2) * j.astype(np.float32) * PI_F / np.float32(p_res)).astype(np.float32)
    rad = (R - r * np.cos(a)).astype(np.float32)
    x = (rad * np.cos(b)).astype(np.float32)
    y = (rad * np.sin(b)).astype(np.float32)
    z = np.broadcast_to((r * np.sin(a)).astype(np.float32), x.shape)
    pts = np.stack([x, y, z], -1).reshape(-1, 3).astype(np.float32)
    ii, jj = np.broadcast_arrays(i, j)
    t0 = np.stack([ii * p_res + jj, (ii + 1) % s_res * p_res + jj, ii * p_res + (jj + 1) % p_res], -1)
    t1 = np.stack([ii * p_res + jj, (ii + s_res - 1) % s_res * p_res + jj, ii * p_res + (jj + p_res - 1) % p_res], -1)
    tris = np.stack([t0, t1], 2).reshape(-1, 3).astype(np.int32)
    return np.ascontiguousarray(pts), np.ascontiguousarray(tris)


def make_uv_sphere(radius=1.0, h_res=16, v_res=16):
    radius = np.float32(radius)
    j = np.arange(v_res, dtype=np.int64)[:, None]
    i = np.arange(h_res, dtype=np.int64)[None, :]
    a = (PI_F * ((j + 1).astype(np.float32) / np.float32(v_res + 1) - np.float32(0.5))).astype(np.float32)
    b = (np.float32(2) * i.astype(np.float32) * PI_F / np.float32(h_res)).astype(np.float32)
    ad, bd = a.astype(np.float64), b.astype(np.float64)
    x = (float(radius) * np.cos(ad) * np.sin(bd)).astype(np.float32)
    y = (float(radius) * np.cos(ad) * np.cos(bd)).astype(np.float32)
    z = np.broadcast_to((float(radius) * np.sin(ad)).astype(np.float32), x.shape)
    body = np.stack([x, y, z], -1).reshape(-1, 3)
    pts = np.concatenate([body, [[0, 0, -radius]], [[0, 0, radius]]]).astype(np.float32)
    bottom, top = h_res * v_res, h_res * v_res + 1
    jj, ii = np.broadcast_arrays(np.arange(v_res - 1, dtype=np.int64)[:, None], i)
    ip = (ii + 1) % h_res
    t0 = np.stack([(jj + 1) * h_res + ii, jj * h_res + ip, jj * h_res + ii], -1)
    t1 = np.stack([(jj + 1) * h_res + ii, (jj + 1) * h_res + ip, jj * h_res + ip], -1)
    side = np.stack([t0, t1], 2).reshape(-1, 3)
    i1 = np.arange(h_res, dtype=np.int64)
    i1p = (i1 + 1) % h_res
    c0 = np.stack([i1, i1p, np.full_like(i1, bottom)], -1)
    c1 = np.stack([(v_res - 1) * h_res + i1p, (v_res - 1) * h_res + i1, np.full_like(i1, top)], -1)
    caps = np.stack([c0, c1], 1).reshape(-1, 3)
    tris = np.concatenate([side, caps]).astype(np.int32)
    return np.ascontiguousarray(pts), np.ascontiguousarray(tris)


def _splitmix64(x):
    x = (x + np.uint64(0x9E3779B97F4A7C15)).astype(np.uint64)
    z = x
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def _uniform01(n, seed, stream):
    """n float64 in (0,1), counter based: splitmix64(seed + stream*2^40 + i), top 53 bits"""
    with np.errstate(over="ignore"):
        ctr = np.arange(n, dtype=np.uint64) + np.uint64(seed) + (np.uint64(stream) << np.uint64(40))
        bits = _splitmix64(ctr)
    return ((bits >> np.uint64(11)).astype(np.float64) + 0.5) * (1.0 / 9007199254740992.0)


def add_noise(points, sigma, seed=42):
    """independent Gaussian displacement of every coordinate (Box-Muller on the counter stream)"""
    n = points.size
    u1, u2 = _uniform01(n, seed, 1), _uniform01(n, seed, 2)
    g = np.sqrt(-2.0 * np.log(u1)) * np.cos(2.0 * np.pi * u2)
    return (points.astype(np.float64) + sigma * g.reshape(points.shape)).astype(np.float32)


def query_points(n, lo=-1.5, hi=1.5, seed=0xB200):
    """n points uniform in [lo,hi]^3: 24-bit mantissas from splitmix64(seed + i)"""
    with np.errstate(over="ignore"):
        bits = _splitmix64(np.arange(n, dtype=np.uint64) + np.uint64(seed))
    out = np.empty((n, 3), np.float32)
    for k in range(3):
        m = ((bits >> np.uint64(20 * k)) & np.uint64(0xFFFFF)).astype(np.float64) / float(1 << 20)
        out[:, k] = (lo + (hi - lo) * m).astype(np.float32)
    return out


def two_tori(R=2.0, r=0.7, p_res=128, s_res=64, shift=(1.0, 0.0, 0.0)):
    """C4 input: two interpenetrating tori concatenated as one triangle soup"""
    p, t = make_torus(R, r, p_res, s_res)
    p2 = (p + np.asarray(shift, np.float32)).astype(np.float32)
    return np.concatenate([p, p2]), np.concatenate([t, t + len(p)]).astype(np.int32)


def grid_for_bbox(points, n, margin_voxels=2.0):
    """cubic n^3 grid over the bbox +- margin voxels; returns (dims, origin, voxel_size) with
    voxel centres at origin + (i+0.5)*voxel (meshToDistanceVolume convention)"""
    lo = points.min(0).astype(np.float64)
    hi = points.max(0).astype(np.float64)
    ext = float((hi - lo).max())
    voxel = ext / (n - 2.0 * margin_voxels)
    ctr = 0.5 * (lo + hi)
    origin = ctr - 0.5 * n * voxel
    return (n, n, n), origin.astype(np.float32), np.float32(voxel)
