// b2fwn_kernels.cu -- traversal kernels of the path (hand-written for sm_100a; no tensor cores: the work is tree
// traversal + FP32/FP64 FMA-pipe arithmetic).
//
// Replaces the reference's one-thread-per-element kernels (source/MRCuda/MRCudaFastWindingNumber.cu:135-243:
// block of 32 threads, private 32-entry stacks in local memory, TWO traversals per voxel) by
//   * a STACKLESS pre-order walk: the tree is stored in DFS pre-order, so "push r, push l, pop" of
//     calcFastWindingNumber (source/MRMesh/MRDipole.cpp:104-121) is "i+1 to descend, skip[i] to leave a subtree";
//     the per-query order of additions is exactly the reference's.
// The grid kernels (warp-cooperative brick walk, fused sign + distance) live in b2fwn_grid.cu; this file holds the
// one-thread-per-query kernels for point vectors and face centroids.
#include "b2fwn_ctx.h"
#include "b2fwn_math.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <cfloat>
#include <climits>
#include <cstdlib>
#include <mutex>

namespace b2
{

namespace
{

constexpr unsigned FULL = 0xFFFFFFFFu;

__device__ __forceinline__ int skipOf( int link, int i ) { return link < 0 ? i + 1 : link; }

// ---- per-thread stackless winding number (vectors of arbitrary points, face centroids) --------------------------
__device__ __forceinline__ float fwnSingle( const TravNode* __restrict__ nodes, const LeafTri* __restrict__ tris, int numNodes,
                                            const float3 q, const float betaSq, const int skipFace )
{
    float res = 0.f;
    int i = 0;
    while ( i < numNodes )
    {
        const float4* np = reinterpret_cast<const float4*>( nodes + i );
        const float4 q2 = __ldg( np + 2 ), q3 = __ldg( np + 3 );
        const int link = __float_as_int( q3.w );
        if ( dipoleAddIfGoodApprox( q, betaSq, q2, q3, res ) ) { i = skipOf( link, i ); continue; }
        if ( link >= 0 ) { ++i; continue; }
        const float4* tp = reinterpret_cast<const float4*>( tris + ( -link - 1 ) );
        const float4 ta = __ldg( tp ), tb = __ldg( tp + 1 ), tc = __ldg( tp + 2 );
        if ( __float_as_int( ta.w ) != skipFace )
            res = FA( res, triangleSolidAngle( q, ta, tb, tc ) );
        ++i;
    }
    constexpr float INV_4PI = 1.0f / ( 4 * 3.14159265358979323846f ); // MRDipole.cpp:126 (PI_F: MRConstants.h)
    return FM( INV_4PI, res );
}

// order[t] = index of the query thread t handles (Morton order of the points, so that the lanes of a warp walk nearly the same
// nodes); nullptr = input order
__global__ void __launch_bounds__( 128 ) kFwnVector( const TravNode* __restrict__ nodes, const LeafTri* __restrict__ tris, int numNodes,
    const float* __restrict__ pts, size_t n, float betaSq, int skipFace, float* __restrict__ out, const uint32_t* __restrict__ order )
{
    const size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if ( t >= n ) return;
    const size_t i = order ? order[t] : t;
    const float3 q = make_float3( pts[3 * i], pts[3 * i + 1], pts[3 * i + 2] );
    out[i] = numNodes > 0 ? fwnSingle( nodes, tris, numNodes, q, betaSq, skipFace ) : 0.f;
}

__device__ __forceinline__ uint32_t spread10( uint32_t v ) // 10 bits -> every third bit
{
    v = ( v | ( v << 16 ) ) & 0x030000FFu;
    v = ( v | ( v << 8 ) ) & 0x0300F00Fu;
    v = ( v | ( v << 4 ) ) & 0x030C30C3u;
    v = ( v | ( v << 2 ) ) & 0x09249249u;
    return v;
}

// 30-bit Morton code of every point inside the box of the mesh grown by half its size (points beyond are clamped: they are far
// from every node and walk the same few top nodes anyway)
__global__ void __launch_bounds__( 256 ) kMortonKeys( const float* __restrict__ pts, size_t n, float3 lo, float3 inv, uint32_t* __restrict__ keys,
                                                      uint32_t* __restrict__ vals )
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if ( i >= n ) return;
    const float fx = fminf( fmaxf( ( pts[3 * i] - lo.x ) * inv.x, 0.f ), 1023.f );
    const float fy = fminf( fmaxf( ( pts[3 * i + 1] - lo.y ) * inv.y, 0.f ), 1023.f );
    const float fz = fminf( fmaxf( ( pts[3 * i + 2] - lo.z ) * inv.z, 0.f ), 1023.f );
    // NaN coordinates: fmaxf( NaN, 0 ) = 0 -- any cell will do, the walk itself is order-independent
    keys[i] = spread10( (uint32_t)fx ) | ( spread10( (uint32_t)fy ) << 1 ) | ( spread10( (uint32_t)fz ) << 2 );
    vals[i] = (uint32_t)i;
}

// calcSelfIntersections (MRFastWindingNumber.cpp:40-52): query = triCenter(f) = (1/3.0f)*(a+b+c), skipFace = f,
// bit set iff w < 0 || w > 1; thresholding on the device (the reference GPU does it on the host after D2H,
// MRCudaFastWindingNumber.cpp:163-169)
__global__ void __launch_bounds__( 128 ) kFwnFaces( const TravNode* __restrict__ nodes, const LeafTri* __restrict__ tris, int numNodes,
    const float* __restrict__ pts, const int32_t* __restrict__ faces, size_t nLeaves, float betaSq,
    float* __restrict__ outW, uint32_t* __restrict__ outBits32 )
{
    // thread t takes the face of leaf slot t: leaves are stored in the tree's pre-order, so the faces of a warp are neighbours in
    // space and their walks share most nodes (in face-id order a warp's walks have nothing to do with each other unless the
    // mesh happens to be numbered coherently).  Deleted faces are no leaves: their bits and values stay zero (cleared before).
    const size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if ( t >= nLeaves ) return;
    const size_t f = (size_t)__float_as_int( __ldg( reinterpret_cast<const float4*>( tris + t ) ).w );
    bool flag = false;
    {
        const int32_t ia = faces[3 * f], ib = faces[3 * f + 1], ic = faces[3 * f + 2];
        float w = 0.f;
        {
            const float third = 1 / 3.0f;
            float3 q;
            q.x = FM( third, FA( FA( pts[3 * (size_t)ia], pts[3 * (size_t)ib] ), pts[3 * (size_t)ic] ) );
            q.y = FM( third, FA( FA( pts[3 * (size_t)ia + 1], pts[3 * (size_t)ib + 1] ), pts[3 * (size_t)ic + 1] ) );
            q.z = FM( third, FA( FA( pts[3 * (size_t)ia + 2], pts[3 * (size_t)ib + 2] ), pts[3 * (size_t)ic + 2] ) );
            w = fwnSingle( nodes, tris, numNodes, q, betaSq, (int)f );
            flag = w < 0 || w > 1;
        }
        if ( outW ) outW[f] = w;
    }
    if ( flag && outBits32 ) atomicOr( outBits32 + ( f >> 5 ), 1u << ( f & 31 ) );
}

} // anonymous namespace

int launchFwnVector( const DeviceTree& t, const float* dPts, size_t n, float beta, int32_t skipFace, float* dOut,
                     cudaStream_t st, uint64_t* launches )
{
    if ( n == 0 ) return B2FWN_OK;
    const size_t blocks = ( n + 127 ) / 128;
    // Morton order for vectors large enough to pay for the sort (r02, C2: 10 M random points against 1 M triangles)
    static const bool noSort = devSwitch( "B2FWN_NO_QUERY_SORT" ) != nullptr; // dev switch
    uint32_t* order = nullptr;
    uint8_t* scratch = nullptr;
    if ( !noSort && n >= ( 1u << 15 ) && n <= (size_t)INT_MAX && t.numLeaves > 0 && t.pool ) // cub counts the items in an int; longer vectors are walked in input order
    {
        size_t tmpBytes = 0;
        cub::DeviceRadixSort::SortPairs( nullptr, tmpBytes, (uint32_t*)nullptr, (uint32_t*)nullptr, (uint32_t*)nullptr, (uint32_t*)nullptr, (int)n, 0, 30, st );
        const size_t nPad = ( n + 63 ) & ~(size_t)63;
        B2_CUDA( cudaMallocFromPoolAsync( &scratch, 4 * nPad * sizeof( uint32_t ) + tmpBytes, t.pool, st ) );
        uint32_t* keys = reinterpret_cast<uint32_t*>( scratch );
        uint32_t *keys2 = keys + nPad, *vals = keys2 + nPad, *vals2 = vals + nPad;
        float3 lo, inv;
        const float ext[3] = { t.bbox[3] - t.bbox[0], t.bbox[4] - t.bbox[1], t.bbox[5] - t.bbox[2] };
        lo = make_float3( t.bbox[0] - 0.5f * ext[0], t.bbox[1] - 0.5f * ext[1], t.bbox[2] - 0.5f * ext[2] );
        inv = make_float3( ext[0] > 0 ? 1024.f / ( 2 * ext[0] ) : 0.f, ext[1] > 0 ? 1024.f / ( 2 * ext[1] ) : 0.f, ext[2] > 0 ? 1024.f / ( 2 * ext[2] ) : 0.f );
        kMortonKeys<<<(unsigned)( ( n + 255 ) / 256 ), 256, 0, st>>>( dPts, n, lo, inv, keys, vals );
        B2_CUDA( cub::DeviceRadixSort::SortPairs( vals2 + nPad, tmpBytes, keys, keys2, vals, vals2, (int)n, 0, 30, st ) );
        order = vals2;
        if ( launches ) *launches += 5; // keys + 4 radix passes
    }
    // One thread per query, in Morton order.  The warp-cooperative walk of the grid kernels (32 x 4 neighbours per warp,
    // launchFwnQueriesWarp) is kept for the faces of calcSelfIntersections, where it wins (C4: 16.7 -> 10.9 ms); for scattered
    // points it loses (C2, 10 M random points: 5.69 ms against 4.29 ms -- the ball of 128 random neighbours is ten triangles
    // wide, few nodes are near or far for all of them).  B2FWN_QUERY_WARPS=1 selects it here as well.
    static const bool warps = devSwitch( "B2FWN_QUERY_WARPS" ) && atoi( devSwitch( "B2FWN_QUERY_WARPS" ) );
    if ( order && warps )
    {
        if ( int e = launchFwnQueriesWarp( t, dPts, order, n, false, beta, skipFace, dOut, nullptr, st, launches ) ) return e;
    }
    else
    {
        kFwnVector<<<(unsigned)blocks, 128, 0, st>>>( t.nodes, t.leafTris, (int)t.numNodes, dPts, n, beta * beta,
                                                      skipFace < 0 ? -1 : skipFace, dOut, order );
        if ( launches ) ++*launches;
    }
    if ( scratch ) B2_CUDA( cudaFreeAsync( scratch, st ) );
    B2_CUDA( cudaGetLastError() );
    return B2FWN_OK;
}

int launchFwnFaces( const DeviceTree& t, float beta, float* dOutW, uint64_t* dBits, cudaStream_t st, uint64_t* launches )
{
    const size_t n = t.numFaces;
    if ( n == 0 ) return B2FWN_OK;
    if ( dBits ) B2_CUDA( cudaMemsetAsync( dBits, 0, ( ( n + 63 ) / 64 ) * sizeof( uint64_t ), st ) );
    if ( dOutW ) B2_CUDA( cudaMemsetAsync( dOutW, 0, n * sizeof( float ), st ) );
    const size_t L = t.numLeaves;
    if ( L == 0 ) return B2FWN_OK;
    static const bool perThread = devSwitch( "B2FWN_QUERY_THREADS" ) && atoi( devSwitch( "B2FWN_QUERY_THREADS" ) );
    if ( L >= ( 1u << 15 ) && !perThread )
        return launchFwnQueriesWarp( t, nullptr, nullptr, L, true, beta, -1, dOutW, reinterpret_cast<uint32_t*>( dBits ), st, launches );
    kFwnFaces<<<(unsigned)( ( L + 127 ) / 128 ), 128, 0, st>>>( t.nodes, t.leafTris, (int)t.numNodes, t.points, t.tris, L, beta * beta, dOutW,
                                                              reinterpret_cast<uint32_t*>( dBits ) );
    if ( launches ) ++*launches;
    B2_CUDA( cudaGetLastError() );
    return B2FWN_OK;
}

// ---- b2fwn_calc_from_vector_device_noalloc: the Morton-ordered query walk with its scratch inside the library image ----------
//
// launchFwnVector above borrows 4 index arrays + the radix sort's workspace from the handle's memory pool for the length of the
// call.  Some hosts of this library cannot let it do that: an application that has handed (nearly) all of the GPU's memory to its own
// arena at start-up, so that any later cudaMalloc / pool growth inside a plug-in fails or evicts it; and a frame- or
// latency-bound loop, where an allocation is a device-wide synchronisation point that stalls the caller's other streams.  For those
// the same walk exists once more with the scratch as a file-scope array in device memory: it is part of the loaded module, so
// every GPU that loads libb2fwn.so carries these 17 MiB from the first CUDA call of the process to its exit, and no call ever
// asks the allocator for anything.  The price is the fixed upper limit on the number of queries below.
//
// gNoAllocScratch is a HOST-side handle of the symbol in this translation unit's host code -- not a device address.  The
// address a kernel can use comes from cudaGetSymbolAddress with the device current, once per device (noAllocScratchOf).
constexpr size_t kNoAllocMaxQueries = B2FWN_NOALLOC_MAX_QUERIES;
constexpr size_t kNoAllocSortBytes = (size_t)1 << 20; // workspace of cub's radix sort; what cub really asks for is checked per call
constexpr size_t kNoAllocScratchBytes = 4 * kNoAllocMaxQueries * sizeof( uint32_t ) + kNoAllocSortBytes;
static __device__ __align__( 256 ) unsigned char gNoAllocScratch[kNoAllocScratchBytes];

namespace
{

constexpr int kNoAllocMaxDeviceOrdinal = 64;
std::mutex gNoAllocMutex[kNoAllocMaxDeviceOrdinal];              // one call at a time per device: every call uses the same array
unsigned char* gNoAllocScratchAddr[kNoAllocMaxDeviceOrdinal] = {}; // device address of gNoAllocScratch on that device, once resolved

// device address of the scratch array on `device`, which must be the current device; caller holds gNoAllocMutex[device]
int noAllocScratchOf( int device, unsigned char** out )
{
    if ( !gNoAllocScratchAddr[device] )
    {
        unsigned char* p = nullptr;
        B2_CUDA( cudaGetSymbolAddress( (void**)&p, gNoAllocScratch ) );
        if ( !p ) { setError( "no device address for the static scratch array" ); return B2FWN_ERR_CUDA; }
        gNoAllocScratchAddr[device] = p; // a failed look-up is not remembered: the next call tries again
    }
    *out = gNoAllocScratchAddr[device];
    return B2FWN_OK;
}

} // anonymous namespace

} // namespace b2

#pragma GCC visibility push( default )
extern "C" int b2fwn_calc_from_vector_device_noalloc( b2fwn_handle h, const float* dPts, size_t n, float beta, int32_t skipFace, float* dOut )
{
    using namespace b2;
    if ( !h || ( n && ( !dPts || !dOut ) ) ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    if ( n > kNoAllocMaxQueries )
    {
        setError( "b2fwn_calc_from_vector_device_noalloc: " + std::to_string( n ) + " queries exceed the static scratch (B2FWN_NOALLOC_MAX_QUERIES = "
                  + std::to_string( kNoAllocMaxQueries ) + "); split the vector or use b2fwn_calc_from_vector_device" );
        return B2FWN_ERR_INVALID;
    }
    if ( h->device < 0 || h->device >= kNoAllocMaxDeviceOrdinal ) { setError( "device ordinal beyond the static scratch table" ); return B2FWN_ERR_INVALID; }
    if ( n == 0 ) return B2FWN_OK;
    B2_CUDA( cudaSetDevice( h->device ) );
    // held until the stream has drained: a second call on this device (another thread, another handle) would sort into the
    // same array while this one still reads its order from it
    std::lock_guard<std::mutex> lock( gNoAllocMutex[h->device] );
    const DeviceTree& t = h->tree;
    const cudaStream_t st = h->stream;
    uint32_t* order = nullptr;
    if ( n >= ( 1u << 15 ) && t.numLeaves > 0 ) // same threshold as launchFwnVector: below it the sort costs more than it saves
    {
        unsigned char* scratch = nullptr;
        if ( int e = noAllocScratchOf( h->device, &scratch ) ) return e;
        // [keys | keys2 | vals | vals2 | sort workspace]; every word read below is written first in this call (the sort clears
        // its own counters), so nothing of the previous call's contents has to be zeroed.  The sort ping-pongs between the two
        // halves (cub::DoubleBuffer): the plain-pointer form would want a second copy of keys and values in its workspace.
        uint32_t* keys = reinterpret_cast<uint32_t*>( scratch );
        uint32_t *keys2 = keys + kNoAllocMaxQueries, *vals = keys2 + kNoAllocMaxQueries, *vals2 = vals + kNoAllocMaxQueries;
        void* sortTmp = vals2 + kNoAllocMaxQueries;
        cub::DoubleBuffer<uint32_t> dbKeys( keys, keys2 ), dbVals( vals, vals2 );
        size_t tmpBytes = 0;
        B2_CUDA( cub::DeviceRadixSort::SortPairs( nullptr, tmpBytes, dbKeys, dbVals, (int)n, 0, 30, st ) );
        if ( tmpBytes > kNoAllocSortBytes ) { setError( "radix sort workspace exceeds its share of the static scratch" ); return B2FWN_ERR_INVALID; }
        const float ext[3] = { t.bbox[3] - t.bbox[0], t.bbox[4] - t.bbox[1], t.bbox[5] - t.bbox[2] };
        const float3 lo = make_float3( t.bbox[0] - 0.5f * ext[0], t.bbox[1] - 0.5f * ext[1], t.bbox[2] - 0.5f * ext[2] );
        const float3 inv = make_float3( ext[0] > 0 ? 1024.f / ( 2 * ext[0] ) : 0.f, ext[1] > 0 ? 1024.f / ( 2 * ext[1] ) : 0.f, ext[2] > 0 ? 1024.f / ( 2 * ext[2] ) : 0.f );
        B2_CUDA( cudaEventRecord( h->evA, st ) );
        kMortonKeys<<<(unsigned)( ( n + 255 ) / 256 ), 256, 0, st>>>( dPts, n, lo, inv, keys, vals );
        B2_CUDA( cub::DeviceRadixSort::SortPairs( sortTmp, tmpBytes, dbKeys, dbVals, (int)n, 0, 30, st ) );
        order = dbVals.Current();
        h->stats.kernel_launches += 5;
    }
    else
        B2_CUDA( cudaEventRecord( h->evA, st ) );
    kFwnVector<<<(unsigned)( ( n + 127 ) / 128 ), 128, 0, st>>>( t.nodes, t.leafTris, (int)t.numNodes, dPts, n, beta * beta, skipFace < 0 ? -1 : skipFace, dOut, order );
    ++h->stats.kernel_launches;
    B2_CUDA( cudaGetLastError() );
    B2_CUDA( cudaEventRecord( h->evB, st ) );
    B2_CUDA( cudaStreamSynchronize( st ) );
    float ms = 0;
    if ( cudaEventElapsedTime( &ms, h->evA, h->evB ) == cudaSuccess ) h->stats.last_kernel_ms = ms;
    return B2FWN_OK;
}
#pragma GCC visibility pop
