// b2fwn_kernels.cu -- traversal kernels of the path (hand-written for sm_100a; no tensor cores: the work is tree
// traversal + FP32/FP64 FMA-pipe arithmetic).
//
// Replaces the reference's one-thread-per-element kernels (source/MRCuda/MRCudaFastWindingNumber.cu:135-243:
// block of 32 threads, private 32-entry stacks in local memory, TWO traversals per voxel) by
//   * a STACKLESS pre-order walk: the tree is stored in DFS pre-order, so "push r, push l, pop" of
//     calcFastWindingNumber (source/MRMesh/MRDipole.cpp:104-121) is "i+1 to descend, skip[i] to leave a subtree";
//     the per-query order of additions is exactly the reference's.
//   * WARP-LOCKSTEP traversal for grids: the 32 lanes of a warp own an 8x2x2 brick of voxels and walk the tree
//     with ONE warp-uniform node index.  Every lane keeps its own "resume" index per half (winding number /
//     distance): a lane that accepted a dipole (or pruned a box) at node i sleeps until the walk reaches skip[i].
//     The warp descends when any lane needs to (warp vote); otherwise it jumps to the minimum resume index
//     (redux.sync).  Node fetches are warp-uniform 16-B loads (one transaction, broadcast), so there is no memory
//     divergence, and each lane performs exactly the additions the reference performs for its voxel, in its order.
//   * ONE fused walk for sign + distance (calcWithDistances, source/MRMesh/MRFastWindingNumber.cpp:69-76): both halves
//     share the node index; the 64-B node keeps each half in its own 32-B sector.
#include "b2fwn_internal.h"
#include "b2fwn_math.cuh"
#include <cfloat>
#include <climits>

namespace b2
{

namespace
{

constexpr unsigned FULL = 0xFFFFFFFFu;
constexpr int BRICK_X = 8, BRICK_Y = 2, BRICK_Z = 2;
constexpr int GRID_TB = 256; // 8 warps = 8 bricks per block

struct GridParams
{
    int dx, dy;          // grid dims in x, y
    int zBegin, zEnd;    // layers computed by this launch
    int bricksX, bricksY, bricksZ;
    float A[9];
    float b[3];
    float betaSq;
    float minDistSq, maxDistSq, threshold;
    int nullOutside;
    int numNodes;
};

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

// ---- warp-lockstep grid kernel -------------------------------------------------------------------------------------
// MODE 0: winding number only (calcFromGrid); MODE 1: signed distance (calcFromGridWithDistances)
template <int MODE>
__global__ void __launch_bounds__( GRID_TB ) kGridLockstep( const TravNode* __restrict__ nodes, const LeafTri* __restrict__ tris,
                                                            const GridParams P, float* __restrict__ out )
{
    const unsigned lane = threadIdx.x & 31;
    const size_t brick = blockIdx.x * (size_t)( GRID_TB / 32 ) + ( threadIdx.x >> 5 );
    const size_t numBricks = (size_t)P.bricksX * P.bricksY * P.bricksZ;
    if ( brick >= numBricks ) return; // whole warp
    const int bx = (int)( brick % P.bricksX );
    const int by = (int)( ( brick / P.bricksX ) % P.bricksY );
    const int bz = (int)( brick / ( (size_t)P.bricksX * P.bricksY ) );
    const int x = bx * BRICK_X + ( lane & 7 );
    const int y = by * BRICK_Y + ( ( lane >> 3 ) & 1 );
    const int z = P.zBegin + bz * BRICK_Z + ( lane >> 4 );
    const bool valid = x < P.dx && y < P.dy && z < P.zEnd;
    const float3 p = applyXf( P.A, P.b, (float)x, (float)y, (float)z );
    const int N = P.numNodes;

    float sum = 0.f;
    float best = P.maxDistSq;
    int fRes = valid ? 0 : INT_MAX; // winding-number half sleeps until the walk reaches this node index
    int dRes = ( MODE == 1 && valid ) ? 0 : INT_MAX;
    bool isNaN = false;

    if ( MODE == 1 && N > 0 )
    {
        // greedy descent towards the brick centre: one near triangle gives every lane a finite upper bound, which
        // makes the order-independent pre-order walk prune like the reference's nearest-first walk
        const float3 pc = applyXf( P.A, P.b, (float)( bx * BRICK_X ) + 3.5f, (float)( by * BRICK_Y ) + 0.5f,
                                   (float)( P.zBegin + bz * BRICK_Z ) + 0.5f );
        int j = 0;
        int link = __float_as_int( __ldg( reinterpret_cast<const float4*>( nodes ) + 1 ).w );
        while ( link >= 0 )
        {
            const int l = j + 1;
            const float4* lp = reinterpret_cast<const float4*>( nodes + l );
            const float4 l0 = __ldg( lp ), l1 = __ldg( lp + 1 );
            const int llink = __float_as_int( l1.w );
            const int r = skipOf( llink, l );
            const float4* rp = reinterpret_cast<const float4*>( nodes + r );
            const float4 r0 = __ldg( rp ), r1 = __ldg( rp + 1 );
            if ( boxDistSq( pc, l0, l1 ) <= boxDistSq( pc, r0, r1 ) ) { j = l; link = llink; }
            else { j = r; link = __float_as_int( r1.w ); }
        }
        const float4* tp = reinterpret_cast<const float4*>( tris + ( -link - 1 ) );
        const float4 ta = __ldg( tp ), tb = __ldg( tp + 1 ), tc = __ldg( tp + 2 );
        if ( valid )
        {
            const float cand = triDistSqExact( p, ta, tb, tc );
            if ( cand < best )
            {
                best = cand;
                if ( best <= P.minDistSq ) dRes = INT_MAX; // MRMeshProject.cpp:81-82
            }
        }
    }

    // narrow band + NaN outside: the reference skips the winding number of NaN voxels (MRFastWindingNumber.cpp:72-74);
    // run the distance half alone first, then the winding-number half for the surviving lanes
    const bool twoPhase = MODE == 1 && P.nullOutside && P.maxDistSq < FLT_MAX;
    const int passes = twoPhase ? 2 : 1;
    for ( int pass = 0; pass < passes; ++pass )
    {
        int fR = fRes, dR = dRes;
        if ( twoPhase )
        {
            if ( pass == 0 ) fR = INT_MAX;
            else
            {
                dR = INT_MAX;
                isNaN = best < P.minDistSq || best >= P.maxDistSq;
                if ( isNaN ) fR = INT_MAX;
            }
        }
        int i = __reduce_min_sync( FULL, min( fR, dR ) );
        while ( i < N )
        {
            const float4* np = reinterpret_cast<const float4*>( nodes + i );
            bool descend = false;
            if ( MODE == 1 && __any_sync( FULL, i >= dR ) )
            {
                const float4 q0 = __ldg( np ), q1 = __ldg( np + 1 );
                const int link = __float_as_int( q1.w );
                if ( i >= dR )
                {
                    if ( boxDistSq( p, q0, q1 ) >= best ) dR = skipOf( link, i ); // MRMeshProject.cpp:30, 48
                    else if ( link >= 0 ) descend = true;
                    else
                    {
                        const float4* tp = reinterpret_cast<const float4*>( tris + ( -link - 1 ) );
                        const float4 ta = __ldg( tp ), tb = __ldg( tp + 1 ), tc = __ldg( tp + 2 );
                        const float cand = triDistSqExact( p, ta, tb, tc );
                        dR = i + 1;
                        if ( cand < best )
                        {
                            best = cand;
                            if ( best <= P.minDistSq )
                            {
                                dR = INT_MAX;
                                if ( !twoPhase && P.nullOutside && best < P.minDistSq ) { isNaN = true; fR = INT_MAX; }
                            }
                        }
                    }
                }
            }
            if ( __any_sync( FULL, i >= fR ) )
            {
                const float4 q2 = __ldg( np + 2 ), q3 = __ldg( np + 3 );
                const int link = __float_as_int( q3.w );
                if ( i >= fR )
                {
                    if ( dipoleAddIfGoodApprox( p, P.betaSq, q2, q3, sum ) ) fR = skipOf( link, i );
                    else if ( link >= 0 ) descend = true;
                    else
                    {
                        const float4* tp = reinterpret_cast<const float4*>( tris + ( -link - 1 ) );
                        const float4 ta = __ldg( tp ), tb = __ldg( tp + 1 ), tc = __ldg( tp + 2 );
                        sum = FA( sum, triangleSolidAngle( p, ta, tb, tc ) );
                        fR = i + 1;
                    }
                }
            }
            if ( __any_sync( FULL, descend ) ) ++i;
            else i = __reduce_min_sync( FULL, min( fR, dR ) );
        }
    }

    if ( !valid ) return;
    constexpr float INV_4PI = 1.0f / ( 4 * 3.14159265358979323846f );
    const float w = FM( INV_4PI, sum );
    float res;
    if ( MODE == 0 ) res = w;
    else
    {
        if ( P.nullOutside && ( best < P.minDistSq || best >= P.maxDistSq ) ) res = __int_as_float( 0x7fc00000 ); // cQuietNan
        else res = ( w > P.threshold ) ? -__fsqrt_rn( best ) : __fsqrt_rn( best );
    }
    const size_t id = (size_t)x + (size_t)P.dx * ( (size_t)y + (size_t)P.dy * (size_t)( z - P.zBegin ) );
    out[id] = res;
}

inline int checkDims( const int32_t dims[3], int zBegin, int zEnd )
{
    if ( dims[0] < 0 || dims[1] < 0 || dims[2] < 0 || zBegin < 0 || zEnd > dims[2] || zBegin > zEnd )
    {
        setError( "invalid grid dimensions" );
        return B2FWN_ERR_INVALID;
    }
    return B2FWN_OK;
}

template <int MODE>
int launchGrid( const DeviceTree& t, const int32_t dims[3], const GridXf& xf, float beta, const b2fwn_dist_options* opt,
                int zBegin, int zEnd, float* dOut, cudaStream_t st, uint64_t* launches )
{
    if ( int e = checkDims( dims, zBegin, zEnd ) ) return e;
    if ( dims[0] == 0 || dims[1] == 0 || zEnd == zBegin ) return B2FWN_OK;
    GridParams P{};
    P.dx = dims[0]; P.dy = dims[1]; P.zBegin = zBegin; P.zEnd = zEnd;
    P.bricksX = ( dims[0] + BRICK_X - 1 ) / BRICK_X;
    P.bricksY = ( dims[1] + BRICK_Y - 1 ) / BRICK_Y;
    P.bricksZ = ( zEnd - zBegin + BRICK_Z - 1 ) / BRICK_Z;
    for ( int k = 0; k < 9; ++k ) P.A[k] = xf.A[k];
    for ( int k = 0; k < 3; ++k ) P.b[k] = xf.b[k];
    P.betaSq = beta * beta; // sqr(beta), MRDipole.cpp:103
    if ( opt )
    {
        P.minDistSq = opt->min_dist_sq; P.maxDistSq = opt->max_dist_sq; P.threshold = opt->winding_number_threshold;
        P.nullOutside = opt->null_outside_min_max;
    }
    P.numNodes = (int)t.numNodes;
    const size_t numBricks = (size_t)P.bricksX * P.bricksY * P.bricksZ;
    const size_t blocks = ( numBricks + GRID_TB / 32 - 1 ) / ( GRID_TB / 32 );
    if ( blocks > 0x7FFFFFFFull ) { setError( "grid too large for one launch" ); return B2FWN_ERR_INVALID; }
    kGridLockstep<MODE><<<(unsigned)blocks, GRID_TB, 0, st>>>( t.nodes, t.leafTris, P, dOut );
    if ( launches ) ++*launches;
    B2_CUDA( cudaGetLastError() );
    return B2FWN_OK;
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

int launchFwnGrid( const DeviceTree& t, const int32_t dims[3], const GridXf& xf, float beta, int zBegin, int zEnd,
                   float* dOut, cudaStream_t st, uint64_t* launches )
{
    return launchGrid<0>( t, dims, xf, beta, nullptr, zBegin, zEnd, dOut, st, launches );
}

int launchSdfGrid( const DeviceTree& t, const int32_t dims[3], const GridXf& xf, const b2fwn_dist_options& opt,
                   int zBegin, int zEnd, float* dOut, cudaStream_t st, uint64_t* launches )
{
    return launchGrid<1>( t, dims, xf, opt.winding_number_beta, &opt, zBegin, zEnd, dOut, st, launches );
}

} // namespace b2
