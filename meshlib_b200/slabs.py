"""meshlib_b200.slabs -- host-side z-slab partition of a voxel grid over ranks.

The reference slices grids into whole-z-layer chunks for memory reasons (maxBufferSizeAlignedByBlock,
/root/reference/source/MRCuda/MRCudaBasic.cpp:103-107); the same axis shards the grid across GPUs: because
id = x + dx*(y + dy*z) (source/MRMesh/MRVolumeIndexer.h:132-140) every slab is one contiguous range of the volume."""


def slab_range(rank, world, dz):
    """layers [z0, z1) of rank `rank`: z_k = floor(k * dz / world)"""
    return rank * dz // world, (rank + 1) * dz // world


def slab_elems(rank, world, dims):
    z0, z1 = slab_range(rank, world, dims[2])
    layer = dims[0] * dims[1]
    return z0 * layer, z1 * layer


def even_split(world, dz):
    """True when every rank owns the same number of layers (the in-place all-gather needs equal slabs)"""
    return dz % world == 0


def gather_slabs(volume, rank, world, dims, dist=None):
    """gathers the ranks' slabs into `volume` (a 1-D torch tensor of the whole grid on every rank).  Equal slabs: one
    in-place all_gather_into_tensor; ragged slabs: one broadcast per rank."""
    if world == 1:
        return volume
    if dist is None:
        import torch.distributed as dist
    if even_split(world, dims[2]):
        a, b = slab_elems(rank, world, dims)
        dist.all_gather_into_tensor(volume, volume[a:b])
    else:
        for r in range(world):
            a, b = slab_elems(r, world, dims)
            if b > a:
                dist.broadcast(volume[a:b], src=r)
    return volume


# ---- z-interleaved partition (equal cost per rank) --------------------------------------------------------------------------
# Contiguous slabs cost different amounts (c5_band on 8 GPUs: slowest slab 192 ms, mean 152 ms -- the pole and surface
# bricks are not spread evenly over z), and the all-gather can only start when the slowest rank is done.  Interleaving
# chunks of a few layers gives every rank an equal share of every part of the grid; splitting the z-range into stages lets
# the gather of one stage run while the next stage is computed.  Kernel side: b2fwn_calc_from_grid_with_distances_device_chunks.

class ChunkPlan:
    """rank r owns the layers  z = p * period + r * chunk + k  (0 <= k < chunk, z < dz),  period = world * chunk.
    The periods are grouped into `stages` consecutive z-ranges, each computed by one call and gathered afterwards."""

    def __init__(self, world, dz, chunk=8, stages=4):
        if chunk <= 0 or chunk % 4:
            raise ValueError("chunk must be a positive multiple of 4 (one brick is 4 layers)")
        self.world, self.dz, self.chunk = world, dz, chunk
        self.period = world * chunk
        self.full_periods = dz // self.period
        n = max(1, min(stages, max(self.full_periods, 1)))
        # stage s = periods [b[s], b[s+1]); the ragged tail (dz % period layers) belongs to the last stage
        self.bounds = [s * self.full_periods // n for s in range(n + 1)]

    @property
    def stages(self):
        return len(self.bounds) - 1

    def stage_range(self, s):
        """z-range [z0, z1) of stage s (all ranks together)"""
        z0 = self.bounds[s] * self.period
        z1 = self.dz if s == self.stages - 1 else self.bounds[s + 1] * self.period
        return z0, z1

    def call_args(self, rank, s):
        """(z_begin, z_end, z_chunk, z_stride) of rank `rank`'s call for stage s, or None when it owns nothing there"""
        z0, z1 = self.stage_range(s)
        zb = z0 + rank * self.chunk
        return (zb, z1, self.chunk, self.period) if zb < z1 else None

    def layers_of(self, rank):
        return [z for z in range(self.dz) if (z % self.period) // self.chunk == rank]

    def gather_stage(self, volume, rank, s, dims, dist, async_op=False):
        """gathers stage s into `volume` on every rank: one in-place all_gather_into_tensor per full period, one broadcast
        per rank for the ragged tail.  Returns the work handles when async_op."""
        layer = dims[0] * dims[1]
        works = []
        if self.world == 1:
            return works
        for p in range(self.bounds[s], self.bounds[s + 1]):
            a = p * self.period * layer
            mine = a + rank * self.chunk * layer
            w = dist.all_gather_into_tensor(volume[a:a + self.period * layer], volume[mine:mine + self.chunk * layer], async_op=async_op)
            works.append(w)
        if s == self.stages - 1 and self.dz % self.period:
            t0 = self.full_periods * self.period
            for r in range(self.world):
                za, zb = t0 + r * self.chunk, min(t0 + (r + 1) * self.chunk, self.dz)
                if zb > za:
                    works.append(dist.broadcast(volume[za * layer:zb * layer], src=r, async_op=async_op))
        return [w for w in works if w is not None]
