// b2fwn_project.cu -- closest point on the mesh for a vector of points, WITH face id and barycentric coordinates:
// the GPU back-end of MR::IPointsToMeshProjector (source/MRMesh/MRPointsToMeshProjector.h:12-31,
// CPU: MRPointsToMeshProjector.cpp:17-33 -> findProjectionSubtree, MRMeshProject.cpp:13-103; reference GPU:
// source/MRCuda/MRCudaPointsToMeshProjector.cu:8-97).
//
// One thread per point, private stack of (node, box distance) pairs, exactly the reference's visiting order: both children
// are box-tested at the parent, the farther one is pushed first, a popped task is dropped when its distance is no longer
// below the running best, a candidate replaces the best only when STRICTLY closer.  The face that wins a tie therefore is
// the reference's, and with the closest point evaluated in double like the reference (MRMeshProject.cpp:67-75) the whole
// record -- face, point, barycentrics, distSq -- is the reference's bit for bit.
// Two things keep that exact and make it fast: (1) the points are WALKED in Morton order (30-bit keys, cub radix sort of an
// index permutation; results go back to the caller's positions), so the threads of a warp descend through the same nodes;
// (2) before the double-precision closest point a candidate passes the FP32 frame-space estimate of the distance kernels
// (b2fwn_math.cuh: rejects only candidates that provably cannot be STRICTLY closer than the running best, so neither the
// result nor the winner of a tie changes).  10 M random points against a 1 M-triangle sphere: 1159 ms -> see profiles/.
// Optional transforms (createProjectionTransforms, MRMeshProject.cpp:104-121): pointXf moves the query point, treeXf (the
// non-rigid case) moves node boxes (hull of the 8 transformed corners, MRBox.h:336-345) and triangle vertices.
#include "b2fwn_internal.h"
#include "b2fwn_math.cuh"
#include <cub/device/device_radix_sort.cuh>
#include <cfloat>

namespace b2
{

namespace
{

struct Xf12
{
    float A[9];
    float b[3];
};

template <bool TXF>
__device__ __forceinline__ float nodeBoxDistSq( const TravNode* __restrict__ nodes, const int n, const Xf12& treeXf, const float3 pt, int& right, int& link )
{
    const float4* np = reinterpret_cast<const float4*>( nodes + n );
    float4 q0 = __ldg( np ), q1 = __ldg( np + 1 );
    right = __float_as_int( q1.z );
    link = __float_as_int( q1.w );
    if ( TXF )
    {
        float lo[3] = { FLT_MAX, FLT_MAX, FLT_MAX }, hi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
#pragma unroll
        for ( int c = 0; c < 8; ++c )
        {
            const float3 q = applyXf( treeXf.A, treeXf.b, ( c & 1 ) ? q0.w : q0.x, ( c & 2 ) ? q1.x : q0.y, ( c & 4 ) ? q1.y : q0.z );
            if ( q.x < lo[0] ) lo[0] = q.x; if ( q.x > hi[0] ) hi[0] = q.x; // Box::include, MRBox.h
            if ( q.y < lo[1] ) lo[1] = q.y; if ( q.y > hi[1] ) hi[1] = q.y;
            if ( q.z < lo[2] ) lo[2] = q.z; if ( q.z > hi[2] ) hi[2] = q.z;
        }
        q0 = make_float4( lo[0], lo[1], lo[2], hi[0] );
        q1.x = hi[1]; q1.y = hi[2];
    }
    return boxDistSq( pt, q0, q1 );
}

template <bool TXF>
__global__ void __launch_bounds__( 128 ) kFindProjections( const TravNode* __restrict__ nodes, const LeafTri* __restrict__ tris,
                                                           const LeafFilt* __restrict__ filt, const int numNodes,
                                                           const float* __restrict__ pts, const uint32_t* __restrict__ order,
                                                           const size_t n, const int hasPointXf, const Xf12 pointXf,
                                                           const Xf12 treeXf, const float upDistLimitSq, const float loDistLimitSq,
                                                           b2fwn_projection* __restrict__ out )
{
    const size_t slot = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if ( slot >= n ) return;
    const size_t idx = order ? order[slot] : slot;
    float3 pt = make_float3( pts[3 * idx], pts[3 * idx + 1], pts[3 * idx + 2] );
    if ( hasPointXf ) pt = applyXf( pointXf.A, pointXf.b, pt.x, pt.y, pt.z );
    const float pmag = 1e-6f * ( fabsf( pt.x ) + fabsf( pt.y ) + fabsf( pt.z ) );

    int face = -1;
    float best = upDistLimitSq, sbest = sqrtf( upDistLimitSq ), bpx = 0.f, bpy = 0.f, bpz = 0.f, ba = 0.f, bb = 0.f;
    int stN[32];
    float stD[32];
    int sp = 0;
    if ( numNodes > 0 )
    {
        int right, link;
        const float d = nodeBoxDistSq<TXF>( nodes, 0, treeXf, pt, right, link );
        if ( d < best ) { stN[0] = 0; stD[0] = d; sp = 1; }
    }
    while ( sp > 0 )
    {
        --sp;
        const int node = stN[sp];
        if ( stD[sp] >= best ) continue;
        const float4 q1 = __ldg( reinterpret_cast<const float4*>( nodes + node ) + 1 );
        const int link = __float_as_int( q1.w );
        if ( link < 0 )
        {
            if ( !TXF )
            {
                // FP32 estimate in the triangle's own frame: a candidate that cannot be strictly closer is not evaluated
                const float4* fp = reinterpret_cast<const float4*>( filt + ( -link - 1 ) );
                const float4 f0 = __ldg( fp ), f1 = __ldg( fp + 1 ), f2 = __ldg( fp + 2 ), f3 = __ldg( fp + 3 );
                const float3 vv = make_float3( fmaf( f2.y, f1.z, -f2.z * f1.y ), fmaf( f2.z, f1.x, -f2.x * f1.z ), fmaf( f2.x, f1.y, -f2.y * f1.x ) );
                if ( estimateRejects( pt, pmag, sbest, f0, f1, f2, f3, vv ) ) continue;
            }
            const float4* tp = reinterpret_cast<const float4*>( tris + ( -link - 1 ) );
            const float4 ta = __ldg( tp ), tb = __ldg( tp + 1 ), tc = __ldg( tp + 2 );
            float3 a = make_float3( ta.x, ta.y, ta.z ), b = make_float3( tb.x, tb.y, tb.z ), c = make_float3( tc.x, tc.y, tc.z );
            if ( TXF )
            {
                a = applyXf( treeXf.A, treeXf.b, a.x, a.y, a.z );
                b = applyXf( treeXf.A, treeXf.b, b.x, b.y, b.z );
                c = applyXf( treeXf.A, treeXf.b, c.x, c.y, c.z );
            }
            double rx, ry, rz, va, vb;
            closestPointBary( pt, a, b, c, rx, ry, rz, va, vb );
            const float fx = __double2float_rn( rx ), fy = __double2float_rn( ry ), fz = __double2float_rn( rz );
            const float ex = FS( fx, pt.x ), ey = FS( fy, pt.y ), ez = FS( fz, pt.z );
            const float cand = dot3( ex, ey, ez, ex, ey, ez );
            if ( cand < best ) // MRMeshProject.cpp:78
            {
                best = cand; sbest = sqrtf( cand ); face = __float_as_int( ta.w ); bpx = fx; bpy = fy; bpz = fz;
                ba = __double2float_rn( va ); bb = __double2float_rn( vb );
                if ( best <= loDistLimitSq ) break; // MRMeshProject.cpp:81-82
            }
            continue;
        }
        const int l = node + 1, r = __float_as_int( q1.z );
        int dummyR, dummyL;
        const float d1 = nodeBoxDistSq<TXF>( nodes, l, treeXf, pt, dummyR, dummyL );
        const float d2 = nodeBoxDistSq<TXF>( nodes, r, treeXf, pt, dummyR, dummyL );
        // the task with the smaller distance is pushed last, to be entered first (MRMeshProject.cpp:90-99)
        if ( d1 < d2 )
        {
            if ( d2 < best ) { stN[sp] = r; stD[sp] = d2; ++sp; }
            if ( d1 < best ) { stN[sp] = l; stD[sp] = d1; ++sp; }
        }
        else
        {
            if ( d1 < best ) { stN[sp] = l; stD[sp] = d1; ++sp; }
            if ( d2 < best ) { stN[sp] = r; stD[sp] = d2; ++sp; }
        }
    }
    float4* o = reinterpret_cast<float4*>( out + idx );
    o[0] = make_float4( __int_as_float( face ), bpx, bpy, bpz );
    o[1] = make_float4( ba, bb, best, 0.f );
}

Xf12 toXf12( const float* x )
{
    Xf12 r{};
    if ( x )
    {
        for ( int k = 0; k < 9; ++k ) r.A[k] = x[k];
        for ( int k = 0; k < 3; ++k ) r.b[k] = x[9 + k];
    }
    return r;
}

__device__ __forceinline__ uint32_t spread10( uint32_t v )
{
    v = ( v | ( v << 16 ) ) & 0x030000FFu;
    v = ( v | ( v << 8 ) ) & 0x0300F00Fu;
    v = ( v | ( v << 4 ) ) & 0x030C30C3u;
    v = ( v | ( v << 2 ) ) & 0x09249249u;
    return v;
}
// 30-bit Morton key of every (transformed) point inside the box [lo, lo + 1024 / scale); points outside clamp to its faces
__global__ void __launch_bounds__( 256 ) kMortonKeys( const float* __restrict__ pts, const size_t n, const int hasPointXf, const Xf12 pointXf,
                                                      const float3 lo, const float3 scale, uint32_t* __restrict__ keys, uint32_t* __restrict__ idx )
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if ( i >= n ) return;
    float3 p = make_float3( pts[3 * i], pts[3 * i + 1], pts[3 * i + 2] );
    if ( hasPointXf ) p = applyXf( pointXf.A, pointXf.b, p.x, p.y, p.z );
    const uint32_t x = (uint32_t)fminf( fmaxf( ( p.x - lo.x ) * scale.x, 0.f ), 1023.f );
    const uint32_t y = (uint32_t)fminf( fmaxf( ( p.y - lo.y ) * scale.y, 0.f ), 1023.f );
    const uint32_t z = (uint32_t)fminf( fmaxf( ( p.z - lo.z ) * scale.z, 0.f ), 1023.f );
    keys[i] = spread10( x ) | ( spread10( y ) << 1 ) | ( spread10( z ) << 2 );
    idx[i] = (uint32_t)i;
}

} // anonymous namespace

static_assert( sizeof( b2fwn_projection ) == 32, "b2fwn_projection is two float4" );

int launchFindProjections( const DeviceTree& t, const float* dPts, size_t n, const float* pointXf12, const float* treeXf12,
                           float upDistLimitSq, float loDistLimitSq, b2fwn_projection* dOut, cudaStream_t st, uint64_t* launches )
{
    if ( n == 0 ) return B2FWN_OK;
    if ( t.depth > 30 ) { setError( "tree too deep for the projection stack" ); return B2FWN_ERR_INVALID; }
    if ( n > 0xFFFFFFFFull ) { setError( "too many points for one call" ); return B2FWN_ERR_INVALID; }
    const unsigned blocks = (unsigned)( ( n + 127 ) / 128 );
    const Xf12 pxf = toXf12( pointXf12 ), txf = toXf12( treeXf12 );

    // walk order: Morton order of the points over twice the mesh box (worth it from a few thousand points on)
    uint32_t* scratch = nullptr;
    const uint32_t* order = nullptr;
    if ( n >= 4096 && t.numNodes > 0 )
    {
        size_t tempBytes = 0;
        cub::DeviceRadixSort::SortPairs( nullptr, tempBytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const uint32_t*)nullptr, (uint32_t*)nullptr,
                                         (int)n, 0, 30, st );
        const size_t words = ( n + 3 ) & ~(size_t)3;
        B2_CUDA( cudaMallocFromPoolAsync( &scratch, 4 * words * sizeof( uint32_t ) + tempBytes, t.pool, st ) );
        uint32_t *keysIn = scratch, *keysOut = scratch + words, *idxIn = scratch + 2 * words, *idxOut = scratch + 3 * words;
        float3 lo, scale;
        {
            const float ex = t.bbox[3] - t.bbox[0], ey = t.bbox[4] - t.bbox[1], ez = t.bbox[5] - t.bbox[2];
            lo = make_float3( t.bbox[0] - 0.5f * ex, t.bbox[1] - 0.5f * ey, t.bbox[2] - 0.5f * ez );
            scale = make_float3( 1024.f / fmaxf( 2.f * ex, FLT_MIN ), 1024.f / fmaxf( 2.f * ey, FLT_MIN ), 1024.f / fmaxf( 2.f * ez, FLT_MIN ) );
        }
        kMortonKeys<<<(unsigned)( ( n + 255 ) / 256 ), 256, 0, st>>>( dPts, n, pointXf12 != nullptr, pxf, lo, scale, keysIn, idxIn );
        B2_CUDA( cub::DeviceRadixSort::SortPairs( scratch + 4 * words, tempBytes, keysIn, keysOut, idxIn, idxOut, (int)n, 0, 30, st ) );
        order = idxOut;
        if ( launches ) *launches += 2;
    }
    if ( treeXf12 )
        kFindProjections<true><<<blocks, 128, 0, st>>>( t.nodes, t.leafTris, t.leafFilt, (int)t.numNodes, dPts, order, n, pointXf12 != nullptr, pxf, txf,
                                                        upDistLimitSq, loDistLimitSq, dOut );
    else
        kFindProjections<false><<<blocks, 128, 0, st>>>( t.nodes, t.leafTris, t.leafFilt, (int)t.numNodes, dPts, order, n, pointXf12 != nullptr, pxf, txf,
                                                         upDistLimitSq, loDistLimitSq, dOut );
    if ( launches ) ++*launches;
    if ( scratch ) B2_CUDA( cudaFreeAsync( scratch, st ) );
    B2_CUDA( cudaGetLastError() );
    return B2FWN_OK;
}

} // namespace b2
