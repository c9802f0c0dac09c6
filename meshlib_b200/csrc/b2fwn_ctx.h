// b2fwn_ctx.h -- the handle behind b2fwn_handle and the host-side chunk pipeline of the host-output / by-parts calls.
// Private to libb2fwn.so (b2fwn_api.cu, b2fwn_pipeline.cu).
#pragma once
#include "b2fwn_internal.h"
#include <functional>
#include <vector>

struct b2fwn_ctx
{
    int device = 0;
    cudaStream_t stream = nullptr;     // compute
    cudaStream_t stream2 = nullptr;    // compute of every second chunk of the host-output calls (overlaps the tail of the previous one)
    cudaStream_t copyStream = nullptr; // D2H of finished parts
    cudaEvent_t evA = nullptr, evB = nullptr;
    b2::DeviceTree tree;
    b2fwn_stats stats{};
    b2::McResult mc;
    b2::McWorkspace mcWork;            // scratch of the mesher, grow-only; released by b2fwn_destroy, b2fwn_trim, and by reclaimCaches
                                       // when an allocation of this handle fails for lack of memory
    b2::TopoResult topo;
    // scratch for the host-pointer entry points
    float* dScratch[2] = { nullptr, nullptr };
    size_t scratchFloats[2] = { 0, 0 };
    float* hPinned[2] = { nullptr, nullptr };
    size_t pinnedFloats[2] = { 0, 0 };
    cudaEvent_t partDone[2] = { nullptr, nullptr };
    cudaEvent_t computeDone[2] = { nullptr, nullptr };
    unsigned* dMinMax = nullptr;       // running min / max of meshToDistanceVolume (ordered-uint encoding)
    // multi-device handle (b2fwn_config.n_devices > 1): the same tree on the other devices; the host-output grid calls deal
    // their z-chunks round-robin over [this] + replicas, one host thread per device
    std::vector<b2fwn_ctx*> replicas;
};

namespace b2
{

// The caches a handle keeps between calls (the two scratch slots, the mesher's workspace, the pinned part buffers, the unused
// blocks of its memory pool) are given back in three ways: b2fwn_destroy, b2fwn_trim, and -- the parts named in `what` -- by the
// allocation sites of the handle themselves when cudaMalloc reports cudaErrorMemoryAllocation, before they try once more.
// A site names only what the running call does not use.  The device of h must be current; the device is synchronised first.
enum : unsigned { RECLAIM_SCRATCH0 = 1, RECLAIM_SCRATCH1 = 2, RECLAIM_MC_WORK = 4, RECLAIM_PINNED = 8, RECLAIM_POOL = 16, RECLAIM_ALL = 31 };
void reclaimCaches( b2fwn_ctx* h, unsigned what );
// device bytes reclaimCaches( h, what ) would release: part of the memory a call of this handle can count on
size_t reclaimableBytes( const b2fwn_ctx* h, unsigned what );
// fn() once more after reclaimCaches( h, what ) when it failed with cudaErrorMemoryAllocation
template <class Fn>
int retryAfterReclaim( b2fwn_ctx* h, unsigned what, Fn fn )
{
    int rc = fn();
    if ( rc == B2FWN_ERR_CUDA && lastCudaError() == cudaErrorMemoryAllocation && reclaimableBytes( h, what ) > 0 )
    {
        reclaimCaches( h, what );
        rc = fn();
    }
    return rc;
}

int ensureScratch( b2fwn_ctx* h, int slot, size_t floats );
int ensurePinned( b2fwn_ctx* h, int slot, size_t floats );

// number of parts of nLayers layers in parts of layersPerPart with `overlap` shared layers (0 when the part is not larger than the overlap)
int partCount( int nLayers, int layersPerPart, int overlap );

// layers per part runChunked uses for nz layers of a dims grid on this handle right now: the caller's wish (<= 0: chosen from the grid),
// cut to what fits twice into the memory its devices can count on, at least overlap + 1, at most nz
int layersPerPartOf( b2fwn_ctx* h, const int32_t dims[3], int nz, int overlap, int wanted );

// getCudaSafeMemoryLimit (source/MRCuda/MRCudaBasic.cpp:75-90): 80 % of ( free - 128 MB ) of the device, in bytes
size_t safeMemoryLimit( int device );

// enqueues the kernels of the layers [z0, z1) of the call's grid on stream st of device ctx dev, result at dOut (device memory
// of that device, first float = voxel (0,0,z0)); called on the host thread that drives dev
using ChunkLaunch = std::function<int( b2fwn_ctx* dev, int z0, int z1, float* dOut, cudaStream_t st )>;

struct ChunkedCall
{
    const int32_t* dims = nullptr;
    int zBegin = 0, zEnd = 0;
    int overlap = 0;            // by-parts: layers shared by consecutive parts
    int layersPerPart = 0;      // <= 0: chosen from the grid and the free device memory
    float* hostOut = nullptr;   // contiguous host output (layer zBegin first), or nullptr with parts
    b2fwn_part_cb part = nullptr;            // by-parts consumer (calling thread, increasing z)
    b2fwn_part_buffer_cb partBuffer = nullptr; // optional: destination of each part, asked on the calling thread
    void* partUser = nullptr;
    b2fwn_progress progress = nullptr;
    void* progressUser = nullptr;
};

// z-chunked pipeline of every host-output grid call: chunks dealt round-robin over the devices of the handle, one driver thread
// per device, kernels of chunk k+1 enqueued before the copy of chunk k is driven, D2H straight into pinned destinations and
// through a pinned staging ring + copier threads into pageable ones; consumer and progress callbacks on the calling thread
int runChunked( b2fwn_ctx* h, const ChunkedCall& call, const ChunkLaunch& launch );

} // namespace b2
