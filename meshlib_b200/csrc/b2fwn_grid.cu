// b2fwn_grid.cu -- the grid kernels of the path: calcFromGrid and calcFromGridWithDistances
// (source/MRMesh/MRFastWindingNumber.cpp:54-94; reference GPU kernels source/MRCuda/MRCudaFastWindingNumber.cu:175-243).
// Hand-written for sm_100a; FP32/FP64 FMA-pipe work fed from L1/L2, no tensor cores (tree traversal, not a contraction).
//
// One warp owns a BRICK of 8 x 4 x 4 voxels; every lane owns 4 consecutive voxels in x (register blocking) and writes
// them with one float4 store.  The warp walks the pre-order tree with ONE warp-uniform node index (stackless: "i+1"
// descends, the node's skip link leaves its subtree -- the order of calcFastWindingNumber's stack, MRDipole.cpp:104-121).
// Every voxel keeps a "resume" index per half (winding number / distance): a voxel that accepted a dipole or pruned a box
// at node i sleeps until the walk reaches skip[i].  The warp descends when any voxel needs to (warp vote), otherwise it
// jumps to the minimum resume index (redux.sync).  Node fetches are warp-uniform 16-B loads (one transaction, broadcast
// to 128 voxels), the per-voxel work is branch-free and 4-way independent per lane (ILP), and the two halves share the
// walk (ONE traversal for sign + distance; the reference GPU does two, .cu:230,238).
//
// Arithmetic contract (b2fwn_math.cuh): far-field DECISIONS and box / closest-point distances round exactly like the
// reference's scalar code, so every voxel accepts exactly the reference's set of dipoles and the distance is the
// reference's value bit for bit; the accepted dipole TERMS use rsqrt (1e-7 relative).  Distance candidates are first
// filtered by a conservative plane-distance lower bound in FP32; survivors get the reference's double closest point.
#include "b2fwn_internal.h"
#include "b2fwn_math.cuh"
#include <cfloat>
#include <climits>
#include <cstdio>
#include <cstdlib>

namespace b2
{

namespace
{

constexpr unsigned FULL = 0xFFFFFFFFu;
constexpr int VX = 4;                  // voxels per lane, consecutive in x
constexpr int BX = 8, BY = 4, BZ = 4;  // brick of one warp: 2 x 4 x 4 lanes
constexpr int GRID_TB = 128;           // 4 warps = 4 bricks (adjacent in x) per block

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
    int vecStore;        // rows are 16-B aligned: float4 stores
};

__device__ __forceinline__ int skipOf( int link, int i ) { return link < 0 ? i + 1 : link; }

__device__ __forceinline__ int warpMinResume( const int ( &fR )[VX], const int ( &dR )[VX] )
{
    int m = INT_MAX;
#pragma unroll
    for ( int v = 0; v < VX; ++v ) m = min( m, min( fR[v], dR[v] ) );
    return __reduce_min_sync( FULL, m );
}

// MODE 0: winding number only (calcFromGrid); MODE 1: signed distance (calcFromGridWithDistances)
// STATS (development only, B2FWN_STATS=1): operation counters of the walk, summed into `stats`
//   [0] distance-half node visits (warp level)  [1] box tests by awake voxels   [2] leaf visits that loaded the triangle (warp)
//   [3] exact closest-point evaluations (voxel)  [4] candidates rejected by the plane bound (voxel)
//   [5] winding-half node visits (warp)          [6] far-field tests by awake voxels [7] exact solid angles (voxel)
template <int MODE, bool STATS>
__global__ void __launch_bounds__( GRID_TB ) kGridBrick( const TravNode* __restrict__ nodes, const LeafTri* __restrict__ tris,
                                                         const GridParams P, float* __restrict__ out, unsigned long long* __restrict__ stats )
{
    unsigned long long cnt[8] = { 0, 0, 0, 0, 0, 0, 0, 0 };
#define B2_COUNT( k, n ) do { if ( STATS ) cnt[k] += ( n ); } while ( 0 )
    const unsigned lane = threadIdx.x & 31;
    const size_t brick = blockIdx.x * (size_t)( GRID_TB / 32 ) + ( threadIdx.x >> 5 );
    const size_t numBricks = (size_t)P.bricksX * P.bricksY * P.bricksZ;
    if ( brick >= numBricks ) return; // whole warp
    const int bx = (int)( brick % P.bricksX );
    const int by = (int)( ( brick / P.bricksX ) % P.bricksY );
    const int bz = (int)( brick / ( (size_t)P.bricksX * P.bricksY ) );
    const int x0 = bx * BX + (int)( lane & 1 ) * VX;
    const int y = by * BY + (int)( ( lane >> 1 ) & 3 );
    const int z = P.zBegin + bz * BZ + (int)( lane >> 3 );
    const bool rowValid = y < P.dy && z < P.zEnd;
    const int N = P.numNodes;

    float3 p[VX];
    float pmag[VX];
    bool valid[VX];
    float sum[VX], best[VX];
    int fR[VX], dR[VX]; // resume indices of the winding-number / distance half of every voxel
#pragma unroll
    for ( int v = 0; v < VX; ++v )
    {
        valid[v] = rowValid && x0 + v < P.dx;
        p[v] = applyXf( P.A, P.b, (float)( x0 + v ), (float)y, (float)z );
        pmag[v] = 1e-6f * ( fabsf( p[v].x ) + fabsf( p[v].y ) + fabsf( p[v].z ) );
        sum[v] = 0.f;
        best[v] = P.maxDistSq;
        fR[v] = valid[v] ? 0 : INT_MAX;
        dR[v] = ( MODE == 1 && valid[v] ) ? 0 : INT_MAX;
    }
    bool isNaN[VX] = { false, false, false, false };

    if ( MODE == 1 && N > 0 )
    {
        // greedy descent towards the brick centre: one near triangle gives every voxel a finite upper bound, which
        // makes the order-independent pre-order walk prune like the reference's nearest-first walk
        const float3 pc = applyXf( P.A, P.b, (float)( bx * BX ) + 3.5f, (float)( by * BY ) + 1.5f, (float)( P.zBegin + bz * BZ ) + 1.5f );
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
            if ( boxDistSqBF( pc, l0, l1 ) <= boxDistSqBF( pc, r0, r1 ) ) { j = l; link = llink; }
            else { j = r; link = __float_as_int( r1.w ); }
        }
        const float4* tp = reinterpret_cast<const float4*>( tris + ( -link - 1 ) );
        const float4 ta = __ldg( tp ), tb = __ldg( tp + 1 ), tc = __ldg( tp + 2 );
#pragma unroll
        for ( int v = 0; v < VX; ++v )
            if ( valid[v] )
            {
                const float cand = triDistSqExact( p[v], ta, tb, tc );
                if ( cand < best[v] )
                {
                    best[v] = cand;
                    if ( best[v] <= P.minDistSq ) dR[v] = INT_MAX; // MRMeshProject.cpp:81-82
                }
            }
    }

    // narrow band + NaN outside: the reference skips the winding number of NaN voxels (MRFastWindingNumber.cpp:72-74);
    // run the distance half alone first, then the winding-number half for the surviving voxels
    const bool twoPhase = MODE == 1 && P.nullOutside && P.maxDistSq < FLT_MAX;
    if ( twoPhase )
    {
#pragma unroll
        for ( int v = 0; v < VX; ++v ) fR[v] = INT_MAX;
    }
    for ( int pass = 0; pass < ( twoPhase ? 2 : 1 ); ++pass )
    {
        if ( pass == 1 )
        {
#pragma unroll
            for ( int v = 0; v < VX; ++v )
            {
                dR[v] = INT_MAX;
                isNaN[v] = best[v] < P.minDistSq || best[v] >= P.maxDistSq;
                fR[v] = ( valid[v] && !isNaN[v] ) ? 0 : INT_MAX;
            }
        }
        int i = warpMinResume( fR, dR );
        while ( i < N )
        {
            const float4* np = reinterpret_cast<const float4*>( nodes + i );
            bool descend = false;
            if ( MODE == 1 )
            {
                const int dmin = min( min( dR[0], dR[1] ), min( dR[2], dR[3] ) );
                if ( __any_sync( FULL, i >= dmin ) )
                {
                    const float4 q0 = __ldg( np ), q1 = __ldg( np + 1 );
                    const int link = __float_as_int( q1.w );
                    if ( lane == 0 ) B2_COUNT( 0, 1 );
                    if ( link >= 0 )
                    {
#pragma unroll
                        for ( int v = 0; v < VX; ++v )
                        {
                            const bool a = i >= dR[v];
                            B2_COUNT( 1, a );
                            const bool pr = boxDistSqBF( p[v], q0, q1 ) >= best[v]; // MRMeshProject.cpp:30, 48
                            dR[v] = ( a && pr ) ? link : dR[v];
                            descend |= a && !pr;
                        }
                    }
                    else
                    {
                        bool go[VX], need = false;
#pragma unroll
                        for ( int v = 0; v < VX; ++v )
                        {
                            const bool a = i >= dR[v];
                            go[v] = a && !( boxDistSqBF( p[v], q0, q1 ) >= best[v] );
                            B2_COUNT( 1, a );
                            dR[v] = a ? i + 1 : dR[v];
                            need |= go[v];
                        }
                        if ( __any_sync( FULL, need ) )
                        {
                            const float4* tp = reinterpret_cast<const float4*>( tris + ( -link - 1 ) );
                            const float4 ta = __ldg( tp ), tb = __ldg( tp + 1 ), tc = __ldg( tp + 2 ), tn = __ldg( tp + 3 );
                            if ( lane == 0 ) B2_COUNT( 2, 1 );
#pragma unroll
                            for ( int v = 0; v < VX; ++v )
                            {
                                if ( !go[v] ) continue;
                                if ( planeBoundRejects( p[v], pmag[v], ta, tn, best[v] ) ) { B2_COUNT( 4, 1 ); continue; }
                                B2_COUNT( 3, 1 );
                                const float cand = triDistSqExact( p[v], ta, tb, tc );
                                if ( cand < best[v] )
                                {
                                    best[v] = cand;
                                    if ( best[v] <= P.minDistSq )
                                    {
                                        dR[v] = INT_MAX;
                                        if ( !twoPhase && P.nullOutside && best[v] < P.minDistSq ) { isNaN[v] = true; fR[v] = INT_MAX; }
                                    }
                                }
                            }
                        }
                    }
                }
            }
            const int fmin = min( min( fR[0], fR[1] ), min( fR[2], fR[3] ) );
            if ( __any_sync( FULL, i >= fmin ) )
            {
                const float4 q2 = __ldg( np + 2 ), q3 = __ldg( np + 3 );
                const int link = __float_as_int( q3.w );
                const float thr = FM( P.betaSq, q2.w );
                if ( lane == 0 ) B2_COUNT( 5, 1 );
                if ( link >= 0 )
                {
#pragma unroll
                    for ( int v = 0; v < VX; ++v )
                    {
                        const bool a = i >= fR[v];
                        float ddx, ddy, ddz, dd;
                        const bool far = dipoleFar( p[v], thr, q2, ddx, ddy, ddz, dd );
                        const float term = dipoleTermFast( ddx, ddy, ddz, dd, q3 );
                        const bool acc = a && far;
                        B2_COUNT( 6, a );
                        sum[v] += acc ? term : 0.f;
                        fR[v] = acc ? link : fR[v];
                        descend |= a && !far;
                    }
                }
                else
                {
                    bool go[VX], need = false;
#pragma unroll
                    for ( int v = 0; v < VX; ++v )
                    {
                        const bool a = i >= fR[v];
                        float ddx, ddy, ddz, dd;
                        const bool far = dipoleFar( p[v], thr, q2, ddx, ddy, ddz, dd );
                        const float term = dipoleTermFast( ddx, ddy, ddz, dd, q3 );
                        sum[v] += ( a && far ) ? term : 0.f;
                        B2_COUNT( 6, a );
                        go[v] = a && !far;
                        fR[v] = a ? i + 1 : fR[v];
                        need |= go[v];
                    }
                    if ( __any_sync( FULL, need ) )
                    {
                        const float4* tp = reinterpret_cast<const float4*>( tris + ( -link - 1 ) );
                        const float4 ta = __ldg( tp ), tb = __ldg( tp + 1 ), tc = __ldg( tp + 2 );
#pragma unroll
                        for ( int v = 0; v < VX; ++v )
                            if ( go[v] ) { B2_COUNT( 7, 1 ); sum[v] += triangleSolidAngle( p[v], ta, tb, tc ); }
                    }
                }
            }
            if ( __any_sync( FULL, descend ) ) ++i;
            else i = warpMinResume( fR, dR );
        }
    }

    if ( STATS )
    {
#pragma unroll
        for ( int k = 0; k < 8; ++k )
        {
            const unsigned long long tot = __reduce_add_sync( FULL, (unsigned)cnt[k] ); // per-warp counts fit 32 bits
            if ( lane == 0 ) atomicAdd( stats + k, tot );
        }
    }
#undef B2_COUNT
    if ( !rowValid ) return;
    constexpr float INV_4PI = 1.0f / ( 4 * 3.14159265358979323846f ); // MRDipole.cpp:126
    float res[VX];
#pragma unroll
    for ( int v = 0; v < VX; ++v )
    {
        const float w = FM( INV_4PI, sum[v] );
        if ( MODE == 0 ) res[v] = w;
        else if ( P.nullOutside && ( best[v] < P.minDistSq || best[v] >= P.maxDistSq ) ) res[v] = __int_as_float( 0x7fc00000 ); // cQuietNan
        else res[v] = ( w > P.threshold ) ? -__fsqrt_rn( best[v] ) : __fsqrt_rn( best[v] );
    }
    const size_t id = (size_t)x0 + (size_t)P.dx * ( (size_t)y + (size_t)P.dy * (size_t)( z - P.zBegin ) );
    if ( P.vecStore && x0 + VX <= P.dx )
        *reinterpret_cast<float4*>( out + id ) = make_float4( res[0], res[1], res[2], res[3] );
    else
    {
#pragma unroll
        for ( int v = 0; v < VX; ++v )
            if ( valid[v] ) out[id + v] = res[v];
    }
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
    P.bricksX = ( dims[0] + BX - 1 ) / BX;
    P.bricksY = ( dims[1] + BY - 1 ) / BY;
    P.bricksZ = ( zEnd - zBegin + BZ - 1 ) / BZ;
    for ( int k = 0; k < 9; ++k ) P.A[k] = xf.A[k];
    for ( int k = 0; k < 3; ++k ) P.b[k] = xf.b[k];
    P.betaSq = beta * beta; // sqr(beta), MRDipole.cpp:103
    if ( opt )
    {
        P.minDistSq = opt->min_dist_sq; P.maxDistSq = opt->max_dist_sq; P.threshold = opt->winding_number_threshold;
        P.nullOutside = opt->null_outside_min_max;
    }
    P.numNodes = (int)t.numNodes;
    P.vecStore = ( dims[0] % VX == 0 ) && ( reinterpret_cast<uintptr_t>( dOut ) % 16 == 0 );
    const size_t numBricks = (size_t)P.bricksX * P.bricksY * P.bricksZ;
    const size_t blocks = ( numBricks + GRID_TB / 32 - 1 ) / ( GRID_TB / 32 );
    if ( blocks > 0x7FFFFFFFull ) { setError( "grid too large for one launch" ); return B2FWN_ERR_INVALID; }
    static const bool wantStats = getenv( "B2FWN_STATS" ) != nullptr;
    if ( wantStats )
    {
        unsigned long long* dStats = nullptr; unsigned long long h[8];
        B2_CUDA( cudaMalloc( &dStats, sizeof( h ) ) );
        B2_CUDA( cudaMemsetAsync( dStats, 0, sizeof( h ), st ) );
        kGridBrick<MODE, true><<<(unsigned)blocks, GRID_TB, 0, st>>>( t.nodes, t.leafTris, P, dOut, dStats );
        B2_CUDA( cudaMemcpyAsync( h, dStats, sizeof( h ), cudaMemcpyDeviceToHost, st ) );
        B2_CUDA( cudaStreamSynchronize( st ) );
        cudaFree( dStats );
        const double nv = (double)dims[0] * dims[1] * ( zEnd - zBegin ), nb = (double)numBricks;
        fprintf( stderr, "[b2fwn stats] voxels %.0f bricks %.0f | dist: visits/brick %.1f boxtests/voxel %.1f trileaf/brick %.1f exact/voxel %.2f "
                 "filtered/voxel %.2f | fwn: visits/brick %.1f tests/voxel %.1f exact/voxel %.2f\n", nv, nb, h[0] / nb, h[1] / nv, h[2] / nb,
                 h[3] / nv, h[4] / nv, h[5] / nb, h[6] / nv, h[7] / nv );
    }
    else
        kGridBrick<MODE, false><<<(unsigned)blocks, GRID_TB, 0, st>>>( t.nodes, t.leafTris, P, dOut, nullptr );
    if ( launches ) ++*launches;
    B2_CUDA( cudaGetLastError() );
    return B2FWN_OK;
}

} // anonymous namespace

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
