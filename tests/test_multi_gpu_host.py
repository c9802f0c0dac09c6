"""Host-side multi-GPU logic on CPU: z-slab partition + slab gather over a world-size-2 gloo group.
The slabs here are filled by the oracle (the checker), not by the product kernels: what is under test is the partition
and the collective plumbing bench.py uses, i.e. that rank-local slabs concatenate into exactly the single-rank volume."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from meshlib_b200 import meshgen, slabs


def test_slab_ranges_cover_the_grid():
    for dz in (1, 7, 64, 2048):
        for world in (1, 2, 3, 4, 8):
            r = [slabs.slab_range(k, world, dz) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == dz
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            assert max(b - a for a, b in r) - min(b - a for a, b in r) <= 1


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, dims, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import portlib
    pts, tris = meshgen.make_torus(2, 1, 24, 12)
    m = portlib.PortMesh(pts, tris)
    xf = np.array([0.4, 0, 0, 0, 0.4, 0, 0, 0, 0.4, -3.0, -3.0, -1.4], np.float32)
    total = dims[0] * dims[1] * dims[2]
    volume = torch.full((total,), float("nan"), dtype=torch.float32)
    z0, z1 = slabs.slab_range(rank, world, dims[2])
    a, b = slabs.slab_elems(rank, world, dims)
    if z1 > z0:
        volume[a:b] = torch.from_numpy(m.calc_grid_with_distances_layers(dims, xf, z0, z1))
    slabs.gather_slabs(volume, rank, world, dims, dist)
    np.save(os.path.join(out_dir, f"rank{rank}.npy"), volume.numpy())
    if rank == 0:
        np.save(os.path.join(out_dir, "full.npy"), m.calc_from_grid_with_distances(dims, xf))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("dims", [(15, 16, 8), (15, 16, 7)], ids=["even", "ragged"])
def test_world2_gloo_slab_gather_equals_single_rank(tmp_path, dims):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, dims, str(tmp_path)), nprocs=2, join=True)
    full = np.load(tmp_path / "full.npy")
    for r in range(2):
        got = np.load(tmp_path / f"rank{r}.npy")
        assert np.array_equal(got.view(np.uint32), full.view(np.uint32))


# ---- z-interleaved chunks in stages (what bench.py runs for N > 1) ------------------------------------------------------------

def test_chunk_plan_covers_the_grid_once():
    for dz in (1, 7, 8, 64, 100, 2048):
        for world in (1, 2, 3, 8):
            for stages in (1, 4):
                plan = slabs.ChunkPlan(world, dz, 8, stages)
                seen = []
                for r in range(world):
                    mine = []
                    for s in range(plan.stages):
                        a = plan.call_args(r, s)
                        if a:
                            zb, ze, c, sd = a
                            while zb < ze:
                                mine += list(range(zb, min(zb + c, ze)))
                                zb += sd
                    assert mine == plan.layers_of(r)
                    seen += mine
                assert sorted(seen) == list(range(dz))
    with pytest.raises(ValueError):
        slabs.ChunkPlan(2, 64, 6, 1)


def _chunk_worker(rank, world, port, dims, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import portlib
    pts, tris = meshgen.make_torus(2, 1, 24, 12)
    m = portlib.PortMesh(pts, tris)
    xf = np.array([0.4, 0, 0, 0, 0.4, 0, 0, 0, 0.2, -3.0, -3.0, -1.4], np.float32)
    layer = dims[0] * dims[1]
    volume = torch.full((layer * dims[2],), float("nan"), dtype=torch.float32)
    plan = slabs.ChunkPlan(world, dims[2], 4, 2)
    works = []
    for s in range(plan.stages):
        for z in plan.layers_of(rank):
            z0, z1 = plan.stage_range(s)
            if z0 <= z < z1:
                volume[z * layer:(z + 1) * layer] = torch.from_numpy(m.calc_grid_with_distances_layers(dims, xf, z, z + 1))
        works += plan.gather_stage(volume, rank, s, dims, dist, async_op=True)
    for w in works:
        w.wait()
    np.save(os.path.join(out_dir, f"rank{rank}.npy"), volume.numpy())
    if rank == 0:
        np.save(os.path.join(out_dir, "full.npy"), m.calc_from_grid_with_distances(dims, xf))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("dims", [(9, 8, 16), (9, 8, 21)], ids=["even", "ragged"])
def test_world2_gloo_staged_chunk_gather_equals_single_rank(tmp_path, dims):
    port = _free_port()
    mp.spawn(_chunk_worker, args=(2, port, dims, str(tmp_path)), nprocs=2, join=True)
    full = np.load(tmp_path / "full.npy")
    for r in range(2):
        got = np.load(tmp_path / f"rank{r}.npy")
        assert np.array_equal(got.view(np.uint32), full.view(np.uint32))
