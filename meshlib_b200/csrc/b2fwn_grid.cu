// b2fwn_grid.cu -- the grid kernels of the path: calcFromGrid and calcFromGridWithDistances
// (source/MRMesh/MRFastWindingNumber.cpp:54-94; reference GPU kernels source/MRCuda/MRCudaFastWindingNumber.cu:175-243).
// Hand-written for sm_100a; FP32/FP64 FMA-pipe work fed from L1/L2, no tensor cores (tree traversal, not a contraction).
//
// One warp owns a BRICK of 8 x 4 x 4 voxels; every lane owns 4 consecutive voxels in x (register blocking) and moves
// them with one float4 access.  The warp walks the pre-order tree with ONE warp-uniform node index (stackless: "i+1"
// descends, the node's skip link leaves its subtree -- the order of calcFastWindingNumber's stack, MRDipole.cpp:104-121).
// Every voxel keeps a "resume" index: a voxel that accepted a dipole (or pruned a box) at node i sleeps until the walk
// reaches skip[i].  The warp descends when any voxel needs to (warp vote), otherwise it jumps to the minimum resume index
// (redux.sync).  Node fetches are warp-uniform 16-B loads (one transaction, broadcast to 128 voxels) and the per-voxel
// work is branch-free and 4-way independent per lane (ILP).
//
//   kGridDist  closest-point distance of every voxel (findProjection semantics: upper/lower limits, MRMeshProject.cpp:13-103)
//              -> writes |d| (or NaN outside [minDist, maxDist) when nullOutsideMinMax)
//   kGridFwn   winding number of every voxel; MODE 1 reads |d|, skips NaN voxels exactly like the reference
//              (MRFastWindingNumber.cpp:72-74) and writes the signed distance in place
// Two walks, not one shared walk: the two halves visit different node sets (band mode: 52 vs 217 nodes per brick, full
// mode: 5014 vs 217), and a shared walk measured 30 % slower (c3_band 83.8 ms vs 64.3 ms).  The reference GPU
// also walks twice, but per THREAD with private stacks (.cu:230,238).
//
// Arithmetic contract (b2fwn_math.cuh): far-field DECISIONS and box / closest-point distances round exactly like the
// reference's scalar code, so every voxel accepts exactly the reference's set of dipoles and the distance is the
// reference's value bit for bit; the accepted dipole TERMS use MUFU rsqrt (<= 1e-6 relative).  Distance candidates are
// first filtered by an FP32 distance estimate in the triangle's own frame with a stated allowance (b2fwn_math.cuh);
// survivors get the reference's double closest point.
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
    int vecStore;        // rows are 16-B aligned: float4 accesses
};

__device__ __forceinline__ int skipOf( int link, int i ) { return link < 0 ? i + 1 : link; }
__device__ __forceinline__ int min4( const int ( &a )[VX] ) { return min( min( a[0], a[1] ), min( a[2], a[3] ) ); }

struct Brick
{
    int bx, by, bz, x0, y, z;
    bool rowValid;
    size_t id; // output index of the lane's first voxel
};

__device__ __forceinline__ bool brickOfWarp( const GridParams& P, Brick& B )
{
    const unsigned lane = threadIdx.x & 31;
    const size_t brick = blockIdx.x * (size_t)( GRID_TB / 32 ) + ( threadIdx.x >> 5 );
    const size_t numBricks = (size_t)P.bricksX * P.bricksY * P.bricksZ;
    if ( brick >= numBricks ) return false; // whole warp
    B.bx = (int)( brick % P.bricksX );
    B.by = (int)( ( brick / P.bricksX ) % P.bricksY );
    B.bz = (int)( brick / ( (size_t)P.bricksX * P.bricksY ) );
    B.x0 = B.bx * BX + (int)( lane & 1 ) * VX;
    B.y = B.by * BY + (int)( ( lane >> 1 ) & 3 );
    B.z = P.zBegin + B.bz * BZ + (int)( lane >> 3 );
    B.rowValid = B.y < P.dy && B.z < P.zEnd;
    B.id = (size_t)B.x0 + (size_t)P.dx * ( (size_t)B.y + (size_t)P.dy * (size_t)( B.z - P.zBegin ) );
    return true;
}

// development counters (B2FWN_STATS=1): [0] node visits (warp) [1] tests by awake voxels [2] leaf visits that loaded the
// triangle (warp) [3] exact evaluations (voxel) [4] candidates rejected by the plane bound (voxel)
#define B2_COUNT( k, n ) do { if ( STATS ) cnt[k] += ( n ); } while ( 0 )
template <bool STATS>
__device__ __forceinline__ void flushCounters( const unsigned ( &cnt )[5], unsigned long long* stats )
{
    if ( !STATS ) return;
#pragma unroll
    for ( int k = 0; k < 5; ++k )
    {
        const unsigned tot = __reduce_add_sync( FULL, cnt[k] );
        if ( ( threadIdx.x & 31 ) == 0 ) atomicAdd( stats + k, (unsigned long long)tot );
    }
}

// ---- brick cull for band-limited distances ---------------------------------------------------------------------------
// With an upper limit (maxDistSq < FLT_MAX, the offset flows) most bricks lie farther than maxDist from every triangle and
// their voxels keep the initial value maxDistSq.  One THREAD per brick proves that with a box-box walk (brick AABB in mesh
// space against the node boxes): about 20 instructions per visit on one thread instead of a 128-voxel warp walk.
// Conservative: a brick is culled only if every leaf BOX is farther than maxDist * (1 + 1e-4) from the padded brick box, so
// no triangle can produce a candidate < maxDistSq for any of its voxels (candidates must be strictly smaller,
// MRMeshProject.cpp:78) -- the culled result is exactly what the walk would return.
__global__ void __launch_bounds__( 256 ) kBrickCull( const TravNode* __restrict__ nodes, const GridParams P, uint8_t* __restrict__ active )
{
    const size_t brick = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    const size_t numBricks = (size_t)P.bricksX * P.bricksY * P.bricksZ;
    if ( brick >= numBricks ) return;
    const int bx = (int)( brick % P.bricksX );
    const int by = (int)( ( brick / P.bricksX ) % P.bricksY );
    const int bz = (int)( brick / ( (size_t)P.bricksX * P.bricksY ) );
    float lo[3] = { FLT_MAX, FLT_MAX, FLT_MAX }, hi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
#pragma unroll
    for ( int c = 0; c < 8; ++c )
    {
        const float3 q = applyXf( P.A, P.b, (float)( bx * BX + ( ( c & 1 ) ? BX - 1 : 0 ) ), (float)( by * BY + ( ( c & 2 ) ? BY - 1 : 0 ) ),
                                  (float)( P.zBegin + bz * BZ + ( ( c & 4 ) ? BZ - 1 : 0 ) ) );
        lo[0] = fminf( lo[0], q.x ); hi[0] = fmaxf( hi[0], q.x );
        lo[1] = fminf( lo[1], q.y ); hi[1] = fmaxf( hi[1], q.y );
        lo[2] = fminf( lo[2], q.z ); hi[2] = fmaxf( hi[2], q.z );
    }
#pragma unroll
    for ( int k = 0; k < 3; ++k )
    {
        const float pad = 1e-5f * fmaxf( fabsf( lo[k] ), fabsf( hi[k] ) ); // interior voxels round a few ulp off the corner hull
        lo[k] -= pad; hi[k] += pad;
    }
    const float lim = P.maxDistSq * 1.0001f;
    int i = 0, hit = 0;
    while ( i < P.numNodes )
    {
        const float4* np = reinterpret_cast<const float4*>( nodes + i );
        const float4 q0 = __ldg( np ), q1 = __ldg( np + 1 );
        const int link = __float_as_int( q1.w );
        const float ddx = fmaxf( fmaxf( q0.x - hi[0], lo[0] - q0.w ), 0.f );
        const float ddy = fmaxf( fmaxf( q0.y - hi[1], lo[1] - q1.x ), 0.f );
        const float ddz = fmaxf( fmaxf( q0.z - hi[2], lo[2] - q1.y ), 0.f );
        const float d = ddx * ddx + ddy * ddy + ddz * ddz;
        if ( d > lim ) i = skipOf( link, i );
        else if ( link < 0 ) { hit = 1; break; }
        else ++i;
    }
    active[brick] = (uint8_t)hit;
}

// ---- distance walk -----------------------------------------------------------------------------------------------------
// Nearest-first depth-first walk with ONE warp-uniform stack (the order of findProjectionSubtree, MRMeshProject.cpp:43-100:
// both children are box-tested at the parent, the nearer one is entered first, the other is pushed).  The stack lives in
// registers, one entry per lane (depth <= log2 F < 32), push = predicated move, pop = one shuffle.  A child is entered when
// ANY of the brick's 128 voxels still needs it (box distance < that voxel's running best); near/far is decided for the
// brick centre.  No per-voxel traversal state at all: a popped subtree is simply re-tested against the (tighter) bests.
// The result is the exact minimum over all triangles of the reference's per-triangle value, hence independent of the order.
template <bool AA, bool STATS>
__global__ void __launch_bounds__( GRID_TB ) kGridDist( const TravNode* __restrict__ nodes, const LeafTri* __restrict__ tris,
                                                        const LeafFilt* __restrict__ filt, const GridParams P, float* __restrict__ out,
                                                        const uint8_t* __restrict__ active, unsigned long long* __restrict__ stats )
{
    unsigned cnt[5] = { 0, 0, 0, 0, 0 };
    const unsigned lane = threadIdx.x & 31;
    Brick B;
    if ( !brickOfWarp( P, B ) ) return;
    // a culled brick keeps maxDistSq everywhere: no walk
    const bool walk = P.numNodes > 0 && !( active && !active[blockIdx.x * (size_t)( GRID_TB / 32 ) + ( threadIdx.x >> 5 )] );

    float3 p[VX];
    float pmag[VX], best[VX], lim[VX], sbest[VX]; // lim = best while the voxel still searches, -1 once it stopped; sbest = sqrt(best)
#pragma unroll
    for ( int v = 0; v < VX; ++v )
    {
        const bool valid = B.rowValid && B.x0 + v < P.dx;
        p[v] = applyXf( P.A, P.b, (float)( B.x0 + v ), (float)B.y, (float)B.z );
        if ( AA ) { p[v].y = p[0].y; p[v].z = p[0].z; } // axis-aligned grid: the lane's voxels share y, z (see kGridFwn)
        pmag[v] = 1e-6f * ( fabsf( p[v].x ) + fabsf( p[v].y ) + fabsf( p[v].z ) );
        best[v] = P.maxDistSq;
        sbest[v] = sqrtf( P.maxDistSq );
        lim[v] = valid ? P.maxDistSq : -1.f;
    }
    const float3 pc = applyXf( P.A, P.b, (float)( B.bx * BX ) + 3.5f, (float)( B.by * BY ) + 1.5f, (float)( P.zBegin + B.bz * BZ ) + 1.5f );

    if ( walk )
    {
        int stN = 0, stI = 0, sp = 0; // lane k holds stack entry k
        int n = 0;
        // info of a node: leaf -> its (negative) link, internal -> index of its right child
        const float4 r1 = __ldg( reinterpret_cast<const float4*>( nodes ) + 1 );
        int info = __float_as_int( r1.w ) < 0 ? __float_as_int( r1.w ) : __float_as_int( r1.z );
        for ( ;; )
        {
            bool pop = true;
            if ( info < 0 )
            {
                // leaf: FP32 estimate first, the reference's double closest point for the survivors
                const int slot = -info - 1;
                const float4* fp = reinterpret_cast<const float4*>( filt + slot );
                const float4 f0 = __ldg( fp ), f1 = __ldg( fp + 1 ), f2 = __ldg( fp + 2 ), f3 = __ldg( fp + 3 );
                const float3 vv = make_float3( fmaf( f2.y, f1.z, -f2.z * f1.y ), fmaf( f2.z, f1.x, -f2.x * f1.z ), fmaf( f2.x, f1.y, -f2.y * f1.x ) ); // v = n x u
                if ( lane == 0 ) B2_COUNT( 2, 1 );
                bool go[VX], ex = false;
#pragma unroll
                for ( int v = 0; v < VX; ++v )
                {
                    const bool rej = estimateRejects( p[v], pmag[v], sbest[v], f0, f1, f2, f3, vv );
                    go[v] = lim[v] >= 0.f && !rej;
                    B2_COUNT( 4, lim[v] >= 0.f && rej );
                    ex |= go[v];
                }
                if ( __any_sync( FULL, ex ) )
                {
                    const float4* tp = reinterpret_cast<const float4*>( tris + slot );
                    const float4 ta = __ldg( tp ), tb = __ldg( tp + 1 ), tc = __ldg( tp + 2 );
#pragma unroll
                    for ( int v = 0; v < VX; ++v )
                    {
                        if ( !go[v] ) continue;
                        B2_COUNT( 3, 1 );
                        const float cand = triDistSqExact( p[v], ta, tb, tc );
                        if ( cand < best[v] ) // MRMeshProject.cpp:78
                        {
                            best[v] = cand;
                            sbest[v] = sqrtf( cand );
                            lim[v] = cand <= P.minDistSq ? -1.f : cand; // MRMeshProject.cpp:81-82: this voxel stops searching
                        }
                    }
                }
            }
            else
            {
                const int l = n + 1, r = info;
                const float4* lp = reinterpret_cast<const float4*>( nodes + l );
                const float4* rp = reinterpret_cast<const float4*>( nodes + r );
                const float4 l0 = __ldg( lp ), l1 = __ldg( lp + 1 ), r0 = __ldg( rp ), r1q = __ldg( rp + 1 );
                if ( lane == 0 ) B2_COUNT( 0, 2 );
                bool wl = false, wr = false;
#pragma unroll
                for ( int v = 0; v < VX; ++v )
                {
                    B2_COUNT( 1, lim[v] >= 0.f ? 2 : 0 );
                    wl |= boxDistSqBF( p[v], l0, l1 ) < lim[v]; // prune on boxDistSq >= best, MRMeshProject.cpp:30, 48
                    wr |= boxDistSqBF( p[v], r0, r1q ) < lim[v];
                }
                const bool WL = __any_sync( FULL, wl ), WR = __any_sync( FULL, wr );
                const int infoL = __float_as_int( l1.w ) < 0 ? __float_as_int( l1.w ) : __float_as_int( l1.z );
                const int infoR = __float_as_int( r1q.w ) < 0 ? __float_as_int( r1q.w ) : __float_as_int( r1q.z );
                if ( WL || WR )
                {
                    pop = false;
                    const bool leftFirst = WL && ( !WR || boxDistSqBF( pc, l0, l1 ) <= boxDistSqBF( pc, r0, r1q ) );
                    if ( WL && WR )
                    {
                        if ( (int)lane == sp ) { stN = leftFirst ? r : l; stI = leftFirst ? infoR : infoL; }
                        ++sp;
                    }
                    n = leftFirst ? l : r;
                    info = leftFirst ? infoL : infoR;
                }
            }
            if ( pop )
            {
                if ( sp == 0 ) break;
                --sp;
                n = __shfl_sync( FULL, stN, sp );
                info = __shfl_sync( FULL, stI, sp );
            }
        }
    }
    flushCounters<STATS>( cnt, stats );

    if ( !B.rowValid ) return;
    float res[VX];
#pragma unroll
    for ( int v = 0; v < VX; ++v )
    {
        // MRFastWindingNumber.cpp:71-73: NaN outside [minDist, maxDist) -- such voxels get no winding number at all
        if ( P.nullOutside && ( best[v] < P.minDistSq || best[v] >= P.maxDistSq ) ) res[v] = __int_as_float( 0x7fc00000 ); // cQuietNan
        else res[v] = __fsqrt_rn( best[v] );
    }
    if ( P.vecStore && B.x0 + VX <= P.dx )
        *reinterpret_cast<float4*>( out + B.id ) = make_float4( res[0], res[1], res[2], res[3] );
    else
    {
#pragma unroll
        for ( int v = 0; v < VX; ++v )
            if ( B.x0 + v < P.dx ) out[B.id + v] = res[v];
    }
}

// ---- winding-number walk -------------------------------------------------------------------------------------------------
// MODE 0: out = w (calcFromGrid).  MODE 1: out holds |d| or NaN from kGridDist; NaN voxels are skipped, the others become
// w > threshold ? -|d| : |d| (calcWithDistances, MRFastWindingNumber.cpp:69-76)
// AA: the grid is axis-aligned in mesh space (A diagonal, every MeshLib caller): the 4 voxels of a lane share y and z, so
// the y/z differences and squares of the far-field test are computed once per lane instead of four times (same rounding).
template <int MODE, bool AA, bool STATS>
__global__ void __launch_bounds__( GRID_TB ) kGridFwn( const TravNode* __restrict__ nodes, const LeafTri* __restrict__ tris,
                                                       const GridParams P, float* __restrict__ out, unsigned long long* __restrict__ stats )
{
    unsigned cnt[5] = { 0, 0, 0, 0, 0 };
    const unsigned lane = threadIdx.x & 31;
    Brick B;
    if ( !brickOfWarp( P, B ) ) return;
    const int N = P.numNodes;

    float dist[VX] = { 0.f, 0.f, 0.f, 0.f };
    const bool vec = P.vecStore && B.x0 + VX <= P.dx;
    if ( MODE == 1 && B.rowValid )
    {
        if ( vec )
        {
            const float4 d4 = *reinterpret_cast<const float4*>( out + B.id );
            dist[0] = d4.x; dist[1] = d4.y; dist[2] = d4.z; dist[3] = d4.w;
        }
        else
        {
#pragma unroll
            for ( int v = 0; v < VX; ++v )
                if ( B.x0 + v < P.dx ) dist[v] = out[B.id + v];
        }
    }
    float3 p[VX];
    float sum[VX];
    int fR[VX];
#pragma unroll
    for ( int v = 0; v < VX; ++v )
    {
        const bool valid = B.rowValid && B.x0 + v < P.dx && !( dist[v] != dist[v] );
        p[v] = applyXf( P.A, P.b, (float)( B.x0 + v ), (float)B.y, (float)B.z );
        if ( AA ) { p[v].y = p[0].y; p[v].z = p[0].z; } // bitwise equal anyway; lets the compiler share the y/z arithmetic
        sum[v] = 0.f;
        fR[v] = valid ? 0 : INT_MAX;
    }
    // brick centre and radius (distance of the farthest voxel centre, padded): one warp-uniform test per node tells when
    // EVERY voxel of the brick is inside beta * r of the node, i.e. all of them must descend -- no per-voxel work then.
    // Conservative by 1e-5 relative against the exactly rounded per-voxel decision, so decisions are unchanged.
    const float3 pc = applyXf( P.A, P.b, (float)( B.bx * BX ) + 3.5f, (float)( B.by * BY ) + 1.5f, (float)( P.zBegin + B.bz * BZ ) + 1.5f );
    float rad = 0.f;
#pragma unroll
    for ( int v = 0; v < VX; ++v )
    {
        const float ex = p[v].x - pc.x, ey = p[v].y - pc.y, ez = p[v].z - pc.z;
        rad = fmaxf( rad, ex * ex + ey * ey + ez * ez );
    }
    rad = sqrtf( __uint_as_float( __reduce_max_sync( FULL, __float_as_uint( rad ) ) ) ) * 1.00001f;

    int i = __reduce_min_sync( FULL, min4( fR ) );
    while ( i < N )
    {
        const float4* np = reinterpret_cast<const float4*>( nodes + i );
        const float4 q2 = __ldg( np + 2 ), q3 = __ldg( np + 3 );
        const int link = __float_as_int( q3.w );
        const float thr = FM( P.betaSq, q2.w );
        if ( lane == 0 ) B2_COUNT( 0, 1 );
        bool descend = false;
        // (|pos - centre| + radius)^2 <= thr * (1 - 1e-5)  =>  dd <= thr for every voxel of the brick
        const float cx = q2.x - pc.x, cy = q2.y - pc.y, cz = q2.z - pc.z;
        const float dcr = sqrtf( fmaf( cx, cx, fmaf( cy, cy, cz * cz ) ) ) * 1.00001f + rad;
        const bool allNear = dcr * dcr <= thr * 0.99999f;
        if ( link >= 0 && allNear )
        {
            if ( lane == 0 ) B2_COUNT( 4, 1 );
            descend = true; // the walk is here because some voxel is awake, and every awake voxel descends
        }
        else if ( link >= 0 )
        {
#pragma unroll
            for ( int v = 0; v < VX; ++v )
            {
                const bool a = i >= fR[v];
                B2_COUNT( 1, a );
                float ddx, ddy, ddz, dd;
                const bool far = dipoleFar( p[v], thr, q2, ddx, ddy, ddz, dd );
                const float term = dipoleTermFast( ddx, ddy, ddz, dd, q3 );
                const bool acc = a && far;
                sum[v] += ( acc && dd >= FLT_MIN ) ? term : 0.f;
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
                B2_COUNT( 1, a );
                float ddx, ddy, ddz, dd;
                const bool far = dipoleFar( p[v], thr, q2, ddx, ddy, ddz, dd );
                const float term = dipoleTermFast( ddx, ddy, ddz, dd, q3 );
                sum[v] += ( a && far && dd >= FLT_MIN ) ? term : 0.f;
                go[v] = a && !far;
                fR[v] = a ? i + 1 : fR[v];
                need |= go[v];
            }
            if ( __any_sync( FULL, need ) )
            {
                const float4* tp = reinterpret_cast<const float4*>( tris + ( -link - 1 ) );
                const float4 ta = __ldg( tp ), tb = __ldg( tp + 1 ), tc = __ldg( tp + 2 );
                if ( lane == 0 ) B2_COUNT( 2, 1 );
#pragma unroll
                for ( int v = 0; v < VX; ++v )
                    if ( go[v] ) { B2_COUNT( 3, 1 ); sum[v] += triangleSolidAngle( p[v], ta, tb, tc ); }
            }
        }
        if ( __any_sync( FULL, descend ) ) ++i;
        else i = __reduce_min_sync( FULL, min4( fR ) );
    }
    flushCounters<STATS>( cnt, stats );

    if ( !B.rowValid ) return;
    constexpr float INV_4PI = 1.0f / ( 4 * 3.14159265358979323846f ); // MRDipole.cpp:126
    float res[VX];
#pragma unroll
    for ( int v = 0; v < VX; ++v )
    {
        const float w = FM( INV_4PI, sum[v] );
        if ( MODE == 0 ) res[v] = w;
        else res[v] = ( dist[v] != dist[v] ) ? dist[v] : ( ( w > P.threshold ) ? -dist[v] : dist[v] );
    }
    if ( vec )
        *reinterpret_cast<float4*>( out + B.id ) = make_float4( res[0], res[1], res[2], res[3] );
    else
    {
#pragma unroll
        for ( int v = 0; v < VX; ++v )
            if ( B.x0 + v < P.dx ) out[B.id + v] = res[v];
    }
}
#undef B2_COUNT

inline int checkDims( const int32_t dims[3], int zBegin, int zEnd )
{
    if ( dims[0] < 0 || dims[1] < 0 || dims[2] < 0 || zBegin < 0 || zEnd > dims[2] || zBegin > zEnd )
    {
        setError( "invalid grid dimensions" );
        return B2FWN_ERR_INVALID;
    }
    return B2FWN_OK;
}

void printStats( const char* what, const unsigned long long* h, double voxels, double bricks )
{
    fprintf( stderr, "[b2fwn stats] %s: voxels %.0f bricks %.0f | visits/brick %.1f tests/voxel %.1f leaf-loads/brick %.1f exact/voxel %.2f "
             "plane-rejected/voxel %.2f\n", what, voxels, bricks, h[0] / bricks, h[1] / voxels, h[2] / bricks, h[3] / voxels, h[4] / voxels );
}

// MODE 0: winding numbers only; MODE 1: distance walk, then winding-number walk over the same slab
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
    const bool aa = xf.A[1] == 0 && xf.A[2] == 0 && xf.A[3] == 0 && xf.A[5] == 0 && xf.A[6] == 0 && xf.A[7] == 0;
    uint8_t* dActive = nullptr;
    if ( MODE == 1 && P.maxDistSq < FLT_MAX && P.numNodes > 0 )
    {
        B2_CUDA( cudaMallocAsync( &dActive, numBricks, st ) );
        kBrickCull<<<(unsigned)( ( numBricks + 255 ) / 256 ), 256, 0, st>>>( t.nodes, P, dActive );
        if ( launches ) ++*launches;
    }
    static const bool wantStats = getenv( "B2FWN_STATS" ) != nullptr;
    if ( wantStats )
    {
        unsigned long long* dStats = nullptr; unsigned long long h[10];
        B2_CUDA( cudaMalloc( &dStats, sizeof( h ) ) );
        B2_CUDA( cudaMemsetAsync( dStats, 0, sizeof( h ), st ) );
        if ( MODE == 1 )
        {
            if ( aa ) kGridDist<true, true><<<(unsigned)blocks, GRID_TB, 0, st>>>( t.nodes, t.leafTris, t.leafFilt, P, dOut, dActive, dStats );
            else kGridDist<false, true><<<(unsigned)blocks, GRID_TB, 0, st>>>( t.nodes, t.leafTris, t.leafFilt, P, dOut, dActive, dStats );
        }
        {
            if ( aa ) kGridFwn<MODE, true, true><<<(unsigned)blocks, GRID_TB, 0, st>>>( t.nodes, t.leafTris, P, dOut, dStats + 5 );
            else kGridFwn<MODE, false, true><<<(unsigned)blocks, GRID_TB, 0, st>>>( t.nodes, t.leafTris, P, dOut, dStats + 5 );
        }
        B2_CUDA( cudaMemcpyAsync( h, dStats, sizeof( h ), cudaMemcpyDeviceToHost, st ) );
        B2_CUDA( cudaStreamSynchronize( st ) );
        cudaFree( dStats );
        const double nv = (double)dims[0] * dims[1] * ( zEnd - zBegin ), nb = (double)numBricks;
        if ( MODE == 1 ) printStats( "dist", h, nv, nb );
        printStats( "fwn ", h + 5, nv, nb );
    }
    else
    {
        if ( MODE == 1 )
        {
            if ( aa ) kGridDist<true, false><<<(unsigned)blocks, GRID_TB, 0, st>>>( t.nodes, t.leafTris, t.leafFilt, P, dOut, dActive, nullptr );
            else kGridDist<false, false><<<(unsigned)blocks, GRID_TB, 0, st>>>( t.nodes, t.leafTris, t.leafFilt, P, dOut, dActive, nullptr );
        }
        {
            if ( aa ) kGridFwn<MODE, true, false><<<(unsigned)blocks, GRID_TB, 0, st>>>( t.nodes, t.leafTris, P, dOut, nullptr );
            else kGridFwn<MODE, false, false><<<(unsigned)blocks, GRID_TB, 0, st>>>( t.nodes, t.leafTris, P, dOut, nullptr );
        }
    }
    if ( launches ) *launches += ( MODE == 1 ? 2 : 1 );
    if ( dActive ) B2_CUDA( cudaFreeAsync( dActive, st ) );
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
