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
