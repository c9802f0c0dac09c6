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
#include "b2fwn_internal.h"
#include "b2fwn_math.cuh"
#include <cfloat>
#include <climits>

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

__global__ void __launch_bounds__( 128 ) kFwnVector( const TravNode* __restrict__ nodes, const LeafTri* __restrict__ tris, int numNodes,
    const float* __restrict__ pts, size_t n, float betaSq, int skipFace, float* __restrict__ out )
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if ( i >= n ) return;
    const float3 q = make_float3( pts[3 * i], pts[3 * i + 1], pts[3 * i + 2] );
    out[i] = numNodes > 0 ? fwnSingle( nodes, tris, numNodes, q, betaSq, skipFace ) : 0.f;
}

// calcSelfIntersections (MRFastWindingNumber.cpp:40-52): query = triCenter(f) = (1/3.0f)*(a+b+c), skipFace = f,
// bit set iff w < 0 || w > 1; thresholding on the device (the reference GPU does it on the host after D2H,
// MRCudaFastWindingNumber.cpp:163-169)
__global__ void __launch_bounds__( 128 ) kFwnFaces( const TravNode* __restrict__ nodes, const LeafTri* __restrict__ tris, int numNodes,
    const float* __restrict__ pts, const int32_t* __restrict__ faces, size_t nFaces, float betaSq,
    float* __restrict__ outW, uint32_t* __restrict__ outBits32 )
{
    const size_t f = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    bool flag = false;
    if ( f < nFaces )
    {
        const int32_t ia = faces[3 * f], ib = faces[3 * f + 1], ic = faces[3 * f + 2];
        float w = 0.f;
        if ( ia >= 0 && ib >= 0 && ic >= 0 && numNodes > 0 )
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
    const unsigned bal = __ballot_sync( FULL, flag );
    if ( ( threadIdx.x & 31 ) == 0 && f < nFaces && outBits32 )
        outBits32[f >> 5] = bal;
}

} // anonymous namespace

int launchFwnVector( const DeviceTree& t, const float* dPts, size_t n, float beta, int32_t skipFace, float* dOut,
                     cudaStream_t st, uint64_t* launches )
{
    if ( n == 0 ) return B2FWN_OK;
    const size_t blocks = ( n + 127 ) / 128;
    kFwnVector<<<(unsigned)blocks, 128, 0, st>>>( t.nodes, t.leafTris, (int)t.numNodes, dPts, n, beta * beta,
                                                  skipFace < 0 ? -1 : skipFace, dOut );
    if ( launches ) ++*launches;
    B2_CUDA( cudaGetLastError() );
    return B2FWN_OK;
}

int launchFwnFaces( const DeviceTree& t, float beta, float* dOutW, uint64_t* dBits, cudaStream_t st, uint64_t* launches )
{
    const size_t n = t.numFaces;
    if ( n == 0 ) return B2FWN_OK;
    if ( dBits ) B2_CUDA( cudaMemsetAsync( dBits, 0, ( ( n + 63 ) / 64 ) * sizeof( uint64_t ), st ) );
    const size_t blocks = ( n + 127 ) / 128;
    kFwnFaces<<<(unsigned)blocks, 128, 0, st>>>( t.nodes, t.leafTris, (int)t.numNodes, t.points, t.tris, n, beta * beta, dOutW,
                                                 reinterpret_cast<uint32_t*>( dBits ) );
    if ( launches ) ++*launches;
    B2_CUDA( cudaGetLastError() );
    return B2FWN_OK;
}

} // namespace b2
