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
#include "b2fwn_brick_math.h"
#include <cfloat>
#include <climits>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <type_traits>

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
    int zBegin, nz;      // first layer and number of layers computed by this launch ...
    int zChunk, zSkip;   // ... in chunks of zChunk consecutive layers with zSkip layers left out after each chunk (z-interleaved
                         // multi-GPU partition; a contiguous slab has zChunk >= nz, zSkip = 0).  zChunk is a multiple of BZ.
    int bricksX, bricksY, bricksZ;
    float A[9];
    float b[3];
    float betaSq;
    float minDistSq, maxDistSq, threshold;
    int nullOutside;
    int numNodes;
    int vecStore;        // rows are 16-B aligned: float4 accesses
};

// grid layer of local layer lz
__device__ __forceinline__ int globalZ( const GridParams& P, const int lz ) { return P.zBegin + lz + ( lz / P.zChunk ) * P.zSkip; }
__device__ __forceinline__ int skipOf( int link, int i ) { return link < 0 ? i + 1 : link; }
__device__ __forceinline__ int min4( const int ( &a )[VX] ) { return min( min( a[0], a[1] ), min( a[2], a[3] ) ); }

struct Brick
{
    int bx, by, bz, x0, y, z;
    bool rowValid;
    size_t id;    // output index of the lane's first voxel
    size_t brick; // linear brick index
};

// Heavy-first dispatch.  Walk lengths are very uneven (c3_band: mean 217 node visits per brick, longest 17,845 -- the pole
// of the UV sphere, thousands of slivers the far field cannot approximate) and the long walks used to sit wherever the
// linear brick order put them: with the pole in the LAST z-layer one SM worked alone for the final 30 % of the kernel
// (ncu r01: sm__cycles_active avg 45.6 M of 64.7 M elapsed).  A one-thread-per-brick estimate of the walk length picks the
// bricks far above the mean; they take the first `cap` warp slots of the grid (started first, hidden behind the bulk), the
// linear pass skips them.  Scheduling only: every brick is computed by the same code, results do not depend on the order.
struct BrickOrder
{
    const uint32_t* heavy = nullptr; // brick indices of the heavy bricks
    const unsigned* count = nullptr; // number of appended entries (may exceed cap: the surplus stayed in the linear pass)
    const uint8_t* cls = nullptr;    // 1 = brick is in the heavy list
    unsigned cap = 0;                // warp slots reserved at the front of the grid; 0 = plain linear order
};

__device__ __forceinline__ void brickFromIndex( const GridParams& P, const size_t brick, Brick& B )
{
    const unsigned lane = threadIdx.x & 31;
    B.brick = brick;
    B.bx = (int)( brick % P.bricksX );
    B.by = (int)( ( brick / P.bricksX ) % P.bricksY );
    B.bz = (int)( brick / ( (size_t)P.bricksX * P.bricksY ) );
    B.x0 = B.bx * BX + (int)( lane & 1 ) * VX;
    B.y = B.by * BY + (int)( ( lane >> 1 ) & 3 );
    B.z = globalZ( P, B.bz * BZ ) + (int)( lane >> 3 ); // a brick never straddles two chunks
    B.rowValid = B.y < P.dy && B.bz * BZ + (int)( lane >> 3 ) < P.nz;
    B.id = (size_t)B.x0 + (size_t)P.dx * ( (size_t)B.y + (size_t)P.dy * (size_t)( B.z - P.zBegin ) );
}

// warp slot -> brick.  The first `cap` slots belong to the heavy list, the rest walk the bricks in linear order and skip the
// listed ones.  false: nothing to do for this warp.
__device__ __forceinline__ bool brickIndexOfWarp( const GridParams& P, const BrickOrder& O, size_t& brick )
{
    const size_t slot = blockIdx.x * (size_t)( GRID_TB / 32 ) + ( threadIdx.x >> 5 );
    const size_t numBricks = (size_t)P.bricksX * P.bricksY * P.bricksZ;
    if ( slot < O.cap )
    {
        if ( slot >= *O.count ) return false; // whole warp
        brick = O.heavy[slot];
    }
    else
    {
        brick = slot - O.cap;
        if ( brick >= numBricks ) return false;
        if ( O.cap && O.cls[brick] ) return false;
    }
    return true;
}
__device__ __forceinline__ bool brickOfWarp( const GridParams& P, const BrickOrder& O, Brick& B )
{
    size_t brick;
    if ( !brickIndexOfWarp( P, O, brick ) ) return false;
    brickFromIndex( P, brick, B );
    return true;
}

// development counters (B2FWN_STATS=1): [0] node visits (warp) [1] tests by awake voxels [2] leaf visits that loaded the
// triangle (warp) [3] exact evaluations (voxel) [4] candidates rejected by the plane bound (voxel)
#define B2_COUNT( k, n ) do { if ( STATS ) cnt[k] += ( n ); } while ( 0 )
constexpr int STAT_HIST = 5, STAT_MAX = 38, STAT_WORDS = 40; // per kernel
template <bool STATS>
__device__ __forceinline__ void flushCounters( const unsigned ( &cnt )[5], unsigned long long* stats, const size_t brickIdx )
{
    if ( !STATS ) return;
#pragma unroll
    for ( int k = 0; k < 5; ++k )
    {
        const unsigned tot = __reduce_add_sync( FULL, cnt[k] );
        if ( ( threadIdx.x & 31 ) == 0 ) atomicAdd( stats + k, (unsigned long long)tot );
    }
    // distribution of the walk length over bricks: histogram of log2(node visits), and the longest walk with its brick
    if ( ( threadIdx.x & 31 ) == 0 )
    {
        const unsigned long long brick = brickIdx;
        atomicAdd( stats + STAT_HIST + ( cnt[0] ? 32 - __clz( cnt[0] ) : 0 ), 1ull );
        atomicMax( stats + STAT_MAX, ( (unsigned long long)cnt[0] << 32 ) | ( brick & 0xFFFFFFFFull ) );
    }
}

// ---- brick cull for band-limited distances ---------------------------------------------------------------------------
// With an upper limit (maxDistSq < FLT_MAX, the offset flows) most bricks lie farther than maxDist from every triangle and
// their voxels keep the initial value maxDistSq.  One THREAD per brick proves that with a box-box walk (brick AABB in mesh
// space against the node boxes): about 20 instructions per visit on one thread instead of a 128-voxel warp walk.
// Conservative: a brick is culled only if every leaf BOX is farther than maxDist * (1 + 1e-4) from the padded brick box, so
// no triangle can produce a candidate < maxDistSq for any of its voxels (candidates must be strictly smaller,
// MRMeshProject.cpp:78) -- the culled result is exactly what the walk would return.
// The walk does not stop at the first leaf in range: it counts the node boxes in range (leaves weighted), which is the
// length of the distance walk the brick's warp is going to do.  cost = 0 <=> culled; the counts feed the heavy-first order.
constexpr unsigned COST_MAX = 4096u; // estimates stop here (one thread walks them): anything longer is heavy anyway
__global__ void __launch_bounds__( 256 ) kBrickCull( const TravNode* __restrict__ nodes, const GridParams P, uint16_t* __restrict__ cost,
                                                     unsigned long long* __restrict__ totals )
{
    const size_t brick = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    const size_t numBricks = (size_t)P.bricksX * P.bricksY * P.bricksZ;
    unsigned c = 0;
    if ( brick < numBricks )
    {
        const int bx = (int)( brick % P.bricksX );
        const int by = (int)( ( brick / P.bricksX ) % P.bricksY );
        const int bz = (int)( brick / ( (size_t)P.bricksX * P.bricksY ) );
        float lo[3] = { FLT_MAX, FLT_MAX, FLT_MAX }, hi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
#pragma unroll
        for ( int k = 0; k < 8; ++k )
        {
            const float3 q = applyXf( P.A, P.b, (float)( bx * BX + ( ( k & 1 ) ? BX - 1 : 0 ) ), (float)( by * BY + ( ( k & 2 ) ? BY - 1 : 0 ) ),
                                      (float)( globalZ( P, bz * BZ ) + ( ( k & 4 ) ? BZ - 1 : 0 ) ) );
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
        int i = 0;
        unsigned visits = 0, leaves = 0;
        while ( i < P.numNodes && visits < COST_MAX )
        {
            const float4* np = reinterpret_cast<const float4*>( nodes + i );
            const float4 q0 = __ldg( np ), q1 = __ldg( np + 1 );
            const int link = __float_as_int( q1.w );
            const float ddx = fmaxf( fmaxf( q0.x - hi[0], lo[0] - q0.w ), 0.f );
            const float ddy = fmaxf( fmaxf( q0.y - hi[1], lo[1] - q1.x ), 0.f );
            const float ddz = fmaxf( fmaxf( q0.z - hi[2], lo[2] - q1.y ), 0.f );
            const float d = ddx * ddx + ddy * ddy + ddz * ddz;
            ++visits;
            if ( d > lim ) i = skipOf( link, i );
            else { leaves += link < 0; ++i; }
        }
        // a walk cut short by the cap has not seen every leaf: such a brick stays active whatever it found
        c = ( leaves || visits >= COST_MAX ) ? min( visits + 8 * leaves, COST_MAX ) : 0;
        cost[brick] = (uint16_t)c;
    }
    // totals[0] = sum of the costs, totals[1] = number of bricks with a walk
    const unsigned sum = __reduce_add_sync( FULL, c ), cntw = __popc( __ballot_sync( FULL, c != 0 ) );
    if ( ( threadIdx.x & 31 ) == 0 && sum ) { atomicAdd( totals, (unsigned long long)sum ); atomicAdd( totals + 1, (unsigned long long)cntw ); }
}

// Length of a winding-number walk over a ball of voxels (centre pc, radius rad), estimated by ONE thread with the kernel's
// own brick-level tests: a node is entered when some voxel of the ball may be inside beta * r of it (centre distance minus
// radius), otherwise every voxel accepts its dipole and the walk leaves the subtree.  Leaves that need the triangle are
// weighted.  The estimate is monotone in the ball: a ball inside another never walks more.
__device__ __forceinline__ unsigned fwnCostWalk( const TravNode* __restrict__ nodes, const int numNodes, const float betaSq, const float3 pc,
                                                 const float rad )
{
    unsigned c = 0;
    int i = 0;
    while ( i < numNodes && c < COST_MAX )
    {
        const float4* np = reinterpret_cast<const float4*>( nodes + i );
        const float4 q2 = __ldg( np + 2 );
        const int link = __float_as_int( __ldg( reinterpret_cast<const float*>( np + 3 ) + 3 ) );
        const float cx = q2.x - pc.x, cy = q2.y - pc.y, cz = q2.z - pc.z;
        const float dn = fmaxf( sqrtf( cx * cx + cy * cy + cz * cz ) - rad, 0.f );
        const bool near = dn * dn <= betaSq * q2.w;
        if ( link < 0 ) { c += near ? 8 : 1; ++i; }
        else { ++c; i = near ? i + 1 : link; }
    }
    return min( c, COST_MAX );
}
// radius in mesh space of a box of voxel centres with half extents (hx, hy, hz) in voxel units
__device__ __forceinline__ float boxRadius( const GridParams& P, const float hx, const float hy, const float hz )
{
    float rad = 0.f;
#pragma unroll
    for ( int k = 0; k < 4; ++k )
    {
        const float sx = hx, sy = ( k & 1 ) ? -hy : hy, sz = ( k & 2 ) ? -hz : hz;
        const float ex = P.A[0] * sx + P.A[1] * sy + P.A[2] * sz, ey = P.A[3] * sx + P.A[4] * sy + P.A[5] * sz,
                    ez = P.A[6] * sx + P.A[7] * sy + P.A[8] * sz;
        rad = fmaxf( rad, ex * ex + ey * ey + ez * ez );
    }
    return sqrtf( rad ) * 1.0001f;
}

// first level: one estimate per GROUP of 2 x 2 x 2 bricks (per brick the estimate cost 54 ms at c5, 5 % of the step); the
// ball of a group contains the balls of its bricks, so a brick can only be heavy inside a heavy group
// TX x TY x TZ = brick of the winding kernel the estimates are for (v5: 8 x 4 x 4, v6: 8 x 8 x 4)
constexpr int GRP = 2;
template <int TX, int TY, int TZ>
__global__ void __launch_bounds__( 256 ) kGroupCostFwn( const TravNode* __restrict__ nodes, const GridParams P, const int gX, const int gY,
                                                        const int gZ, uint16_t* __restrict__ cost, unsigned long long* __restrict__ totals )
{
    const size_t group = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    const size_t numGroups = (size_t)gX * gY * gZ;
    unsigned c = 0;
    if ( group < numGroups )
    {
        const int gx = (int)( group % gX ), gy = (int)( ( group / gX ) % gY ), gz = (int)( group / ( (size_t)gX * gY ) );
        const float hx = 0.5f * ( GRP * TX - 1 ), hy = 0.5f * ( GRP * TY - 1 ), hz = 0.5f * ( GRP * TZ - 1 );
        const float3 pc = applyXf( P.A, P.b, (float)( gx * GRP * TX ) + hx, (float)( gy * GRP * TY ) + hy, (float)globalZ( P, gz * GRP * TZ ) + hz );
        c = fwnCostWalk( nodes, P.numNodes, P.betaSq, pc, boxRadius( P, hx, hy, hz ) );
        cost[group] = (uint16_t)c;
    }
    const unsigned sum = __reduce_add_sync( FULL, c ), cntw = __popc( __ballot_sync( FULL, c != 0 ) );
    if ( ( threadIdx.x & 31 ) == 0 && sum ) { atomicAdd( totals, (unsigned long long)sum ); atomicAdd( totals + 1, (unsigned long long)cntw ); }
}

// Estimates far above the mean go to the heavy list (BrickOrder).  The threshold is a multiple of the mean estimate of the
// walking bricks / groups, kept inside the range the capped estimates can express.
constexpr unsigned HEAVY_FACTOR = 8, HEAVY_MIN = COST_MAX / 2;
__device__ __forceinline__ unsigned heavyThreshold( const unsigned long long* __restrict__ totals )
{
    const unsigned long long mean = totals[1] ? totals[0] / totals[1] : 0;
    return (unsigned)min( (unsigned long long)COST_MAX - 1, max( (unsigned long long)HEAVY_MIN, HEAVY_FACTOR * mean ) );
}
__global__ void __launch_bounds__( 256 ) kBrickHeavy( const uint16_t* __restrict__ cost, const size_t numBricks,
                                                      const unsigned long long* __restrict__ totals, uint32_t* __restrict__ heavy,
                                                      unsigned* __restrict__ count, uint8_t* __restrict__ cls, const unsigned cap )
{
    const size_t brick = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if ( brick >= numBricks ) return;
    uint8_t h = 0;
    if ( cost[brick] > heavyThreshold( totals ) )
    {
        const unsigned k = atomicAdd( count, 1u );
        if ( k < cap ) { heavy[k] = (uint32_t)brick; h = 1; }
    }
    cls[brick] = h;
}
// second level: the bricks of a heavy group get their own estimate; those above the same threshold are listed
template <int TX, int TY, int TZ>
__global__ void __launch_bounds__( 256 ) kGroupHeavy( const TravNode* __restrict__ nodes, const uint16_t* __restrict__ cost, const GridParams P,
                                                      const int gX, const int gY, const unsigned long long* __restrict__ totals,
                                                      uint32_t* __restrict__ heavy, unsigned* __restrict__ count, uint8_t* __restrict__ cls,
                                                      const unsigned cap )
{
    const size_t brick = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    const size_t numBricks = (size_t)P.bricksX * P.bricksY * P.bricksZ;
    if ( brick >= numBricks ) return;
    const int bx = (int)( brick % P.bricksX ), by = (int)( ( brick / P.bricksX ) % P.bricksY ), bz = (int)( brick / ( (size_t)P.bricksX * P.bricksY ) );
    const size_t group = (size_t)( bx / GRP ) + (size_t)gX * ( (size_t)( by / GRP ) + (size_t)gY * (size_t)( bz / GRP ) );
    const unsigned thr = heavyThreshold( totals );
    uint8_t h = 0;
    if ( cost[group] > thr )
    {
        const float hx = 0.5f * ( TX - 1 ), hy = 0.5f * ( TY - 1 ), hz = 0.5f * ( TZ - 1 );
        const float3 pc = applyXf( P.A, P.b, (float)( bx * TX ) + hx, (float)( by * TY ) + hy, (float)globalZ( P, bz * TZ ) + hz );
        if ( fwnCostWalk( nodes, P.numNodes, P.betaSq, pc, boxRadius( P, hx, hy, hz ) ) > thr )
        {
            const unsigned k = atomicAdd( count, 1u );
            if ( k < cap ) { heavy[k] = (uint32_t)brick; h = 1; }
        }
    }
    cls[brick] = h;
}

// ---- distance walk -----------------------------------------------------------------------------------------------------
// Nearest-first depth-first walk with ONE warp-uniform stack (the order of findProjectionSubtree, MRMeshProject.cpp:43-100:
// both children are box-tested at the parent, the nearer one is entered first, the other is pushed).  The stack lives in
// registers, one entry per lane (depth <= log2 F < 32), push = predicated move, pop = one shuffle.  A child is entered when
// ANY of the brick's 128 voxels still needs it (box distance < that voxel's running best); near/far is decided for the
// brick centre.  No per-voxel traversal state at all: a popped subtree is simply re-tested against the (tighter) bests.
// The result is the exact minimum over all triangles of the reference's per-triangle value, hence independent of the order.
// 8 blocks per SM (64 registers, ~200 B of spills) measured faster than the spill-free 118 registers at 4 blocks: c3_full
// 626 -> 535 ms, c5_band 1221 -> 1150 ms -- the walk is latency-bound (one dependent node fetch after another)
// PROJ (SignDetectionMode::ProjectionNormal over a volume): the voxel centres are origin + voxelSize * ( pos + 0.5 ), rounded
// like meshToDistanceFunctionVolume rounds them (P.b = origin, P.A = diag( voxelSize )), and the leaf slot of the closest
// triangle is stored beside the distance -- the sign kernel (b2fwn_normals.cu) recomputes the closest point on that one triangle
template <bool AA, bool STATS, int MINB = 8, bool PROJ = false>
__global__ void __launch_bounds__( GRID_TB, MINB ) kGridDist( const TravNode* __restrict__ nodes, const LeafTri* __restrict__ tris,
                                                        const LeafFilt* __restrict__ filt, const GridParams P, float* __restrict__ out,
                                                        const uint16_t* __restrict__ cullCost, const BrickOrder O,
                                                        unsigned long long* __restrict__ stats, int32_t* __restrict__ slotOut = nullptr )
{
    unsigned cnt[5] = { 0, 0, 0, 0, 0 };
    const unsigned lane = threadIdx.x & 31;
    Brick B;
    if ( !brickOfWarp( P, O, B ) ) return;
    // a culled brick (kBrickCull: cost 0) keeps maxDistSq everywhere: no walk
    const bool walk = P.numNodes > 0 && !( cullCost && !cullCost[B.brick] );

    float3 p[VX];
    float pmag[VX], best[VX], lim[VX], sbest[VX]; // lim = best while the voxel still searches, -1 once it stopped; sbest = sqrt(best)
    int bslot[PROJ ? VX : 1];                     // PROJ: leaf slot of the closest triangle so far
#pragma unroll
    for ( int v = 0; v < VX; ++v )
    {
        const bool valid = B.rowValid && B.x0 + v < P.dx;
        if ( PROJ )
            p[v] = make_float3( __fadd_rn( P.b[0], __fmul_rn( P.A[0], (float)( B.x0 + v ) + 0.5f ) ), __fadd_rn( P.b[1], __fmul_rn( P.A[4], (float)B.y + 0.5f ) ),
                                __fadd_rn( P.b[2], __fmul_rn( P.A[8], (float)B.z + 0.5f ) ) ); // MRMeshToDistanceVolume.cpp:83-84
        else
            p[v] = applyXf( P.A, P.b, (float)( B.x0 + v ), (float)B.y, (float)B.z );
        if ( AA ) { p[v].y = p[0].y; p[v].z = p[0].z; } // axis-aligned grid: the lane's voxels share y, z (see kGridFwn)
        pmag[v] = 1e-6f * ( fabsf( p[v].x ) + fabsf( p[v].y ) + fabsf( p[v].z ) );
        if ( PROJ ) bslot[v] = -1;
        best[v] = P.maxDistSq;
        sbest[v] = sqrtf( P.maxDistSq );
        lim[v] = valid ? P.maxDistSq : -1.f;
    }
    const float3 pc = applyXf( P.A, P.b, (float)( B.bx * BX ) + 3.5f, (float)( B.by * BY ) + 1.5f, (float)globalZ( P, B.bz * BZ ) + 1.5f );

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
                            if ( PROJ ) bslot[v] = slot;
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
    flushCounters<STATS>( cnt, stats, B.brick );

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
    if ( PROJ )
    {
#pragma unroll
        for ( int v = 0; v < VX; ++v )
            if ( B.x0 + v < P.dx ) slotOut[B.id + v] = bslot[v];
    }
}

// ---- winding-number walk -------------------------------------------------------------------------------------------------
// MODE 0: out = w (calcFromGrid).  MODE 1: out holds |d| or NaN from kGridDist; NaN voxels are skipped, the others become
// w > threshold ? -|d| : |d| (calcWithDistances, MRFastWindingNumber.cpp:69-76)
// AA: the grid is axis-aligned in mesh space (A diagonal, every MeshLib caller): the 4 voxels of a lane share y and z, so
// the y/z differences and squares of the far-field test are computed once per lane instead of four times (same rounding).
template <int MODE, bool AA, bool STATS>
__global__ void __launch_bounds__( GRID_TB ) kGridFwn( const TravNode* __restrict__ nodes, const LeafTri* __restrict__ tris,
                                                       const GridParams P, float* __restrict__ out, const BrickOrder O,
                                                       unsigned long long* __restrict__ stats )
{
    unsigned cnt[5] = { 0, 0, 0, 0, 0 };
    const unsigned lane = threadIdx.x & 31;
    Brick B;
    if ( !brickOfWarp( P, O, B ) ) return;
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
    const float3 pc = applyXf( P.A, P.b, (float)( B.bx * BX ) + 3.5f, (float)( B.by * BY ) + 1.5f, (float)globalZ( P, B.bz * BZ ) + 1.5f );
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
    flushCounters<STATS>( cnt, stats, B.brick );

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

// ---- winding-number walk, v5 -------------------------------------------------------------------------------------------
// Same walk and the same decisions as kGridFwn above with fewer issued instructions per node (ncu r01: 46 % of the issue
// slots were warp-uniform per-node overhead, the ALU pipe -- compares and selects -- was the busiest pipe at 46 %):
//  * the next node comes from ONE warp reduction: m = min over the warp of the resume indices AFTER the node was
//    processed; an awake voxel that neither accepted nor slept keeps resume <= i, so "descend" <=> m <= i and the next node
//    is max(m, i+1) -- no per-voxel descend predicates, no vote;
//  * the all-near test uses one MUFU square root (sqrt.approx, 2^-22 relative, inside the test's 1e-5 margins) instead of
//    the IEEE sqrt sequence;
//  * the denormal guard of the reference (`d > 0`, MRDipole.h:24) can only matter when beta^2 rr < FLT_MIN; such nodes take
//    a guarded copy of the loop body, all others carry no guard;
//  * accepted terms are added under a predicate (one FFMA) instead of multiply + select + add.
// (A second copy of the winding halves packed 32 B apart -- 4 nodes per line instead of 2 -- measured the same: c5_band 1151 ms.)
// (Fetching node i+1 ahead of time measured SLOWER: c5_band 1419 ms against 1182 ms -- 8 more registers, 2 more loads.
//  Block size 32 / 64 / 128 measured the same.)
// The walk is a function of NV voxels per lane, used with NV = 4 by the regular bricks and with NV = 1 by the slices of
// heavy bricks (below); every term is added with the same explicit operations, so a voxel gets the same bits either way.
__device__ __forceinline__ float sqrtApprox( const float x )
{
    float r;
    asm( "sqrt.approx.ftz.f32 %0, %1;" : "=f"( r ) : "f"( x ) );
    return r;
}
// the two factors of an accepted term, dot(dp, dirArea) and rsqrt(dd)^3, operation by operation; the caller adds them with
// one explicit fmaf( dot, r3, sum ) -- no compiler-chosen contraction anywhere
__device__ __forceinline__ void dipoleTermParts( const float dx, const float dy, const float dz, const float dd, const float4 q3, float& dot, float& r3 )
{
    const float r = rsqrtApprox( dd );
    dot = fmaf( dx, q3.x, fmaf( dy, q3.y, FM( dz, q3.z ) ) );
    r3 = FM( FM( r, r ), r );
}
template <int NV>
__device__ __forceinline__ int minOf( const int ( &a )[NV] )
{
    int m = a[0];
#pragma unroll
    for ( int v = 1; v < NV; ++v ) m = min( m, a[v] );
    return m;
}

template <bool GUARD, int NV>
__device__ __forceinline__ void fwnInner( const int i, const int link, const float thr, const float4 q2, const float4 q3,
                                          const float3 ( &p )[NV], float ( &sum )[NV], int ( &fR )[NV] )
{
#pragma unroll
    for ( int v = 0; v < NV; ++v )
    {
        float ddx, ddy, ddz, dd, dot, r3;
        const bool far = dipoleFar( p[v], thr, q2, ddx, ddy, ddz, dd );
        dipoleTermParts( ddx, ddy, ddz, dd, q3, dot, r3 );
        if ( i >= fR[v] && far )
        {
            if ( !GUARD || dd >= FLT_MIN ) sum[v] = fmaf( dot, r3, sum[v] );
            fR[v] = link;
        }
    }
}

// walks the tree for the NV voxels of every lane; pc, rad = centre and padded radius of all voxels of the warp
// SKIP: skip[v] = face whose exact solid angle voxel v leaves out (calcFromVector's skipFace, calcSelfIntersections' own face:
// MRDipole.cpp:113-116 -- the leaf's dipole test still runs, only the triangle term is dropped)
template <int NV, bool SKIP = false>
__device__ __forceinline__ void fwnWalk( const TravNode* __restrict__ nodes, const LeafTri* __restrict__ tris, const int N, const float betaSq,
                                         const float3 ( &p )[NV], const float3 pc, const float rad, float ( &sum )[NV], int ( &fR )[NV],
                                         const int* skip = nullptr )
{
    int i = __reduce_min_sync( FULL, minOf( fR ) );
    while ( i < N )
    {
        const float4* np = reinterpret_cast<const float4*>( nodes + i );
        const float4 q2 = __ldg( np + 2 ), q3 = __ldg( np + 3 );
        const int link = __float_as_int( q3.w );
        const float thr = FM( betaSq, q2.w );
        if ( link >= 0 )
        {
            // (|pos - centre| + radius)^2 <= thr * (1 - 1e-5)  =>  dd <= thr for every voxel of the warp: all descend
            const float cx = q2.x - pc.x, cy = q2.y - pc.y, cz = q2.z - pc.z;
            const float dcr = fmaf( sqrtApprox( fmaf( cx, cx, fmaf( cy, cy, cz * cz ) ) ), 1.00001f, rad );
            if ( dcr * dcr <= thr * 0.99999f ) { ++i; continue; }
            if ( thr >= FLT_MIN ) fwnInner<false>( i, link, thr, q2, q3, p, sum, fR );
            else fwnInner<true>( i, link, thr, q2, q3, p, sum, fR );
        }
        else
        {
            bool go[NV], need = false;
#pragma unroll
            for ( int v = 0; v < NV; ++v )
            {
                const bool a = i >= fR[v];
                float ddx, ddy, ddz, dd, dot, r3;
                const bool far = dipoleFar( p[v], thr, q2, ddx, ddy, ddz, dd );
                dipoleTermParts( ddx, ddy, ddz, dd, q3, dot, r3 );
                if ( a && far && dd >= FLT_MIN ) sum[v] = fmaf( dot, r3, sum[v] );
                go[v] = a && !far;
                fR[v] = a ? i + 1 : fR[v];
                need |= go[v];
            }
            if ( __any_sync( FULL, need ) )
            {
                const float4* tp = reinterpret_cast<const float4*>( tris + ( -link - 1 ) );
                const float4 ta = __ldg( tp ), tb = __ldg( tp + 1 ), tc = __ldg( tp + 2 );
#pragma unroll
                for ( int v = 0; v < NV; ++v )
                    if ( go[v] && !( SKIP && __float_as_int( ta.w ) == skip[v] ) )
                        sum[v] = FA( sum[v], triangleSolidAngle( p[v], ta, tb, tc ) ); // (all NV angles + predicated adds measured slower)
            }
        }
        i = max( __reduce_min_sync( FULL, minOf( fR ) ), i + 1 );
    }
}

__device__ __forceinline__ float fwnResult( const int MODE, const float sum, const float dist, const float threshold )
{
    constexpr float INV_4PI = 1.0f / ( 4 * 3.14159265358979323846f ); // MRDipole.cpp:126
    const float w = FM( INV_4PI, sum );
    if ( MODE == 0 ) return w;
    return ( dist != dist ) ? dist : ( ( w > threshold ) ? -dist : dist ); // MRFastWindingNumber.cpp:72-76
}

// Bricks of the heavy list (BrickOrder) are simply walked FIRST, by the same code.  Cutting them into 4 z-slices of one voxel
// per lane was tried (as a branch of this kernel: ptxas then guards every warp reduction of the loop against divergence,
// +10 % on the regular walk; as a kernel of its own on a side stream: the slices pay every node for 32 voxels instead of
// 128, 4,195 listed bricks of c3_band cost 9 ms beside a 20 ms regular kernel) and dropped.
// PEERS: the multi-GPU gather fused into the kernel.  Every rank computes its own bricks and stores each result, besides
// its own volume, straight into the volumes of the other ranks (peer memory over NVLink / NVSwitch, posted writes): when
// the kernels of all ranks have finished, every GPU holds the whole volume -- no collective kernel competes for SMs (the
// NCCL all-gather could not overlap this kernel: it fills every SM's register file) and nothing is left to do after the
// last brick.  Traffic: (P-1) x 4 B per voxel leave each GPU during its step, far below the NVLink rate.
// 9 blocks per SM (56 registers) measured 1.8 % faster than 8 (63 registers); 10 (48 registers, 60 B spills) slower again
template <int MODE, bool AA, bool PEERS, int MINB = 9>
__global__ void __launch_bounds__( GRID_TB, MINB ) kGridFwn5( const TravNode* __restrict__ nodes, const LeafTri* __restrict__ tris,
                                                        const GridParams P, float* __restrict__ out, const BrickOrder O, const __grid_constant__ PeerOut peers )
{
    Brick B;
    if ( !brickOfWarp( P, O, B ) ) return;
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
    const float3 pc = applyXf( P.A, P.b, (float)( B.bx * BX ) + 3.5f, (float)( B.by * BY ) + 1.5f, (float)globalZ( P, B.bz * BZ ) + 1.5f );
    float rad = 0.f;
#pragma unroll
    for ( int v = 0; v < VX; ++v )
    {
        const float ex = p[v].x - pc.x, ey = p[v].y - pc.y, ez = p[v].z - pc.z;
        rad = fmaxf( rad, ex * ex + ey * ey + ez * ez );
    }
    rad = sqrtf( __uint_as_float( __reduce_max_sync( FULL, __float_as_uint( rad ) ) ) ) * 1.00001f;

    fwnWalk<VX>( nodes, tris, N, P.betaSq, p, pc, rad, sum, fR );

    if ( !B.rowValid ) return;
    float res[VX];
#pragma unroll
    for ( int v = 0; v < VX; ++v ) res[v] = fwnResult( MODE, sum[v], dist[v], P.threshold );
    if ( vec )
        *reinterpret_cast<float4*>( out + B.id ) = make_float4( res[0], res[1], res[2], res[3] );
    else
    {
#pragma unroll
        for ( int v = 0; v < VX; ++v )
            if ( B.x0 + v < P.dx ) out[B.id + v] = res[v];
    }
    if ( PEERS )
    {
        for ( int k = 0; k < peers.n; ++k )
        {
            float* po = peers.p[k];
            if ( vec )
                *reinterpret_cast<float4*>( po + B.id ) = make_float4( res[0], res[1], res[2], res[3] );
            else
            {
#pragma unroll
                for ( int v = 0; v < VX; ++v )
                    if ( B.x0 + v < P.dx ) po[B.id + v] = res[v];
            }
        }
    }
}

// ---- winding numbers of query vectors, warp-cooperative ---------------------------------------------------------------------
// calcFromVector / calcSelfIntersections for large vectors: the queries arrive in an order in which neighbours in the vector are
// neighbours in space (Morton order of the points, b2fwn_kernels.cu; leaf order of the faces), a warp takes 32 x QV consecutive
// ones as its "brick" and walks the tree for them with the grid kernel's code: one fetch and one all-near test per node for 128
// queries, per-query resume indices, the reference's exactly rounded decisions.  One thread per query (kFwnVector / kFwnFaces)
// walked every node once per query, each lane its own way: C2 4.3 ms, C4 16.7 ms (r02).
constexpr int QV = 4;

// order: query index of slot t (nullptr: t itself); FACES: the queries are the centres of the leaf triangles (slot t = leaf t),
// each skipping its own face, bits set where w < 0 || w > 1 (MRFastWindingNumber.cpp:40-52)
template <bool FACES>
__global__ void __launch_bounds__( GRID_TB, 8 ) kFwnQueriesWarp( const TravNode* __restrict__ nodes, const LeafTri* __restrict__ tris, const int N,
    const float* __restrict__ pts, const int32_t* __restrict__ faces, const uint32_t* __restrict__ order, const size_t n, const float betaSq,
    const int skipFace, float* __restrict__ outW, uint32_t* __restrict__ outBits32 )
{
    const unsigned lane = threadIdx.x & 31;
    const size_t base = ( blockIdx.x * (size_t)( GRID_TB / 32 ) + ( threadIdx.x >> 5 ) ) * ( 32 * QV );
    if ( base >= n ) return;
    float3 p[QV];
    float sum[QV];
    int fR[QV], skip[QV];
    size_t idx[QV];
    float lo[3] = { FLT_MAX, FLT_MAX, FLT_MAX }, hi[3] = { -FLT_MAX, -FLT_MAX, -FLT_MAX };
#pragma unroll
    for ( int v = 0; v < QV; ++v )
    {
        const size_t t = base + (size_t)v * 32 + lane;
        const bool valid = t < n;
        idx[v] = 0; skip[v] = skipFace; p[v] = make_float3( 0.f, 0.f, 0.f );
        if ( valid )
        {
            if ( FACES )
            {
                const size_t f = (size_t)__float_as_int( __ldg( reinterpret_cast<const float4*>( tris + t ) ).w );
                const int32_t ia = faces[3 * f], ib = faces[3 * f + 1], ic = faces[3 * f + 2];
                const float third = 1 / 3.0f; // triCenter: (1/3.0f) * (a + b + c), MRMeshMath.cpp:119-124
                p[v].x = FM( third, FA( FA( pts[3 * (size_t)ia], pts[3 * (size_t)ib] ), pts[3 * (size_t)ic] ) );
                p[v].y = FM( third, FA( FA( pts[3 * (size_t)ia + 1], pts[3 * (size_t)ib + 1] ), pts[3 * (size_t)ic + 1] ) );
                p[v].z = FM( third, FA( FA( pts[3 * (size_t)ia + 2], pts[3 * (size_t)ib + 2] ), pts[3 * (size_t)ic + 2] ) );
                idx[v] = f; skip[v] = (int)f;
            }
            else
            {
                idx[v] = order ? order[t] : t;
                p[v] = make_float3( pts[3 * idx[v]], pts[3 * idx[v] + 1], pts[3 * idx[v] + 2] );
            }
            const float c[3] = { p[v].x, p[v].y, p[v].z };
#pragma unroll
            for ( int k = 0; k < 3; ++k )
                if ( fabsf( c[k] ) <= FLT_MAX ) { lo[k] = fminf( lo[k], c[k] ); hi[k] = fmaxf( hi[k], c[k] ); }
        }
        // a query with a non-finite coordinate gets NaN from the reference's arithmetic (every dd is NaN or inf: the terms are);
        // it is kept out of the walk and of the warp's ball, its sum starts and stays NaN
        const bool finite = fabsf( p[v].x ) <= FLT_MAX && fabsf( p[v].y ) <= FLT_MAX && fabsf( p[v].z ) <= FLT_MAX;
        sum[v] = finite ? 0.f : __int_as_float( 0x7fc00000 );
        fR[v] = valid && finite ? 0 : INT_MAX;
    }
    // ball of the warp's queries: centre of their box, radius to the farthest one (padded like the grid bricks')
    float3 pc;
    {
        float c[3];
#pragma unroll
        for ( int k = 0; k < 3; ++k )
        {
            float a = lo[k], b = hi[k];
#pragma unroll
            for ( int o = 16; o > 0; o >>= 1 ) { a = fminf( a, __shfl_xor_sync( FULL, a, o ) ); b = fmaxf( b, __shfl_xor_sync( FULL, b, o ) ); }
            c[k] = 0.5f * ( a + b );
        }
        pc = make_float3( c[0], c[1], c[2] );
    }
    float rad = 0.f;
#pragma unroll
    for ( int v = 0; v < QV; ++v )
        if ( fR[v] == 0 )
        {
            const float ex = p[v].x - pc.x, ey = p[v].y - pc.y, ez = p[v].z - pc.z;
            rad = fmaxf( rad, ex * ex + ey * ey + ez * ez );
        }
    rad = sqrtf( __uint_as_float( __reduce_max_sync( FULL, __float_as_uint( rad ) ) ) ) * 1.00001f;
    if ( !( pc.x == pc.x && pc.y == pc.y && pc.z == pc.z ) || !( rad < FLT_MAX ) ) { pc = make_float3( 0.f, 0.f, 0.f ); rad = FLT_MAX; } // no all-near shortcut then

    fwnWalk<QV, true>( nodes, tris, N, betaSq, p, pc, rad, sum, fR, skip );

    constexpr float INV_4PI = 1.0f / ( 4 * 3.14159265358979323846f ); // MRDipole.cpp:126
#pragma unroll
    for ( int v = 0; v < QV; ++v )
    {
        const size_t t = base + (size_t)v * 32 + lane;
        if ( t >= n ) continue;
        const float w = FM( INV_4PI, sum[v] );
        if ( outW ) outW[idx[v]] = w;
        if ( FACES && outBits32 && ( w < 0 || w > 1 ) ) atomicOr( outBits32 + ( idx[v] >> 5 ), 1u << ( idx[v] & 31 ) );
    }
}

// ---- winding-number walk, v6 (axis-aligned grids: every MeshLib caller) ------------------------------------------------------
// What v5 still paid per node, for every voxel: the exactly rounded far-field decision, the resume-index compare and select,
// and the warp reduction that finds the next node.  A CPU replay of the walk on the bench workloads (c5_band, c3_band) shows
// that most node visits need none of that: of ~200 visits per brick, ~85 are nodes EVERY voxel of the brick is near to
// (all descend) and 60-80 are nodes EVERY voxel is far from (all accept the dipole); only the nodes whose beta-sphere cuts
// the brick (28 of 193 visits at c5_band, 76 of 214 at c3_band) need per-voxel decisions.  So the walk has two modes:
//  * UNIFORM: all voxels of the brick are awake at the same node.  One conservative brick-level test (b2fwn_brick_math.h,
//    no square root) classifies the node: ALL_NEAR -> i+1, nothing else; ALL_FAR -> every voxel adds the dipole term, no
//    decision, no resume index, no reduction, next node = skip link; MIXED -> a diverged section.
//  * DIVERGED section = the subtree [i, skip(i)) of a MIXED node, walked like v5 with per-voxel resume indices and the
//    reference's exactly rounded decisions.  At skip(i) every voxel is awake again: back to UNIFORM.
//    (MIXED leaves are sections of length one: the leaf code lives in the diverged walk only.)
// Bricks are 8 x 8 x 4 voxels, 8 consecutive x-voxels per lane: the per-node work that does not depend on the voxel (two 16-B
// loads, the brick test, the y/z parts of dd and of the dot product -- the lane's voxels share y and z on an axis-aligned
// grid) is paid once per 256 voxels; replay: 200 visits per 256-voxel brick against 193 per 128-voxel brick.
// Arithmetic: Blackwell's packed FP32 (FADD2 / FMUL2 / FFMA2, two IEEE operations per issue slot) on voxel pairs.  ptxas 12.9
// contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2 (also through the __fmul2_rn / __fadd2_rn intrinsics), which would change
// the rounding of the DECISION, so the exactly rounded dx^2 is a scalar FMUL and only its two additions are packed.
// The accepted TERM is the same sequence of operations in both modes (contracted dd, MUFU rsqrt, FFMA), so a voxel gets the
// same bits whichever brick or mode it is computed in; terms are summed in pre-order of the accepted nodes, as in v5.
constexpr int FX = 8;                       // voxels per lane
constexpr int F6X = 8, F6Y = 8, F6Z = 4;    // brick of one warp: 1 x 8 x 4 lanes

__device__ __forceinline__ float2 bc2( const float s ) { return make_float2( s, s ); }

// what a lane needs of a node besides the voxel's own x: y/z parts of the squared distance and of dot(dp, dirArea)
struct NodeLane
{
    float dy2, dz2; // exactly rounded squares (decision)
    float syz;      // dy^2 + dz^2, contracted (term)
    float yz;       // dy * dirArea.y + dz * dirArea.z (term)
};
__device__ __forceinline__ NodeLane nodeLane( const float4 q2, const float4 q3, const float py, const float pz )
{
    const float dy = FS( q2.y, py ), dz = FS( q2.z, pz );
    NodeLane L;
    L.dy2 = FM( dy, dy );
    L.dz2 = FM( dz, dz );
    L.syz = fmaf( dy, dy, L.dz2 );
    L.yz = fmaf( dy, q3.y, FM( dz, q3.z ) );
    return L;
}
// term of a voxel pair: dot(dp, dirArea) and rsqrt(dd)^3, dd contracted
__device__ __forceinline__ void termPair( const float2 dx, const NodeLane& L, const float q3x, float2& dot, float2& r3, float2& ddc )
{
    ddc = __ffma2_rn( dx, dx, bc2( L.syz ) );
    const float2 r = make_float2( rsqrtApprox( ddc.x ), rsqrtApprox( ddc.y ) );
    r3 = __fmul2_rn( __fmul2_rn( r, r ), r );
    dot = __ffma2_rn( dx, bc2( q3x ), bc2( L.yz ) );
}
__device__ __forceinline__ float2 negate2( const float2 a ) { return make_float2( -a.x, -a.y ); }

// ALL_FAR node: every voxel of the lane adds the dipole term
__device__ __forceinline__ void fwn6AddAll( const float4 q2, const float4 q3, const float2 ( &px )[FX / 2], const float py, const float pz,
                                            float2 ( &sum )[FX / 2] )
{
    const NodeLane L = nodeLane( q2, q3, py, pz );
#pragma unroll
    for ( int k = 0; k < FX / 2; ++k )
    {
        const float2 dx = __fadd2_rn( bc2( q2.x ), negate2( px[k] ) );
        float2 dot, r3, ddc;
        termPair( dx, L, q3.x, dot, r3, ddc );
        sum[k] = __ffma2_rn( dot, r3, sum[k] );
    }
}

// exactly rounded dd of a voxel pair: ( dx*dx + dy*dy ) + dz*dz, every operation rounded (MRDipole.h:21)
__device__ __forceinline__ float2 exactDd( const float2 dx, const NodeLane& L )
{
    const float2 sq = make_float2( FM( dx.x, dx.x ), FM( dx.y, dx.y ) );
    return __fadd2_rn( __fadd2_rn( sq, bc2( L.dy2 ) ), bc2( L.dz2 ) );
}

template <bool GUARD>
__device__ __forceinline__ void fwn6Inner( const int i, const int link, const float thr, const float4 q2, const float4 q3,
                                           const float2 ( &px )[FX / 2], const float py, const float pz, float2 ( &sum )[FX / 2], int ( &fR )[FX] )
{
    const NodeLane L = nodeLane( q2, q3, py, pz );
#pragma unroll
    for ( int k = 0; k < FX / 2; ++k )
    {
        const float2 dx = __fadd2_rn( bc2( q2.x ), negate2( px[k] ) );
        const float2 dd = exactDd( dx, L );
        float2 dot, r3, ddc;
        termPair( dx, L, q3.x, dot, r3, ddc );
        if ( i >= fR[2 * k] && !( dd.x <= thr ) )
        {
            if ( !GUARD || ( dd.x >= FLT_MIN && ddc.x >= FLT_MIN ) ) sum[k].x = fmaf( dot.x, r3.x, sum[k].x );
            fR[2 * k] = link;
        }
        if ( i >= fR[2 * k + 1] && !( dd.y <= thr ) )
        {
            if ( !GUARD || ( dd.y >= FLT_MIN && ddc.y >= FLT_MIN ) ) sum[k].y = fmaf( dot.y, r3.y, sum[k].y );
            fR[2 * k + 1] = link;
        }
    }
}

// diverged section: nodes [i, end) with per-voxel resume indices; returns end
__device__ __forceinline__ int fwn6Section( const TravNode* __restrict__ nodes, const LeafTri* __restrict__ tris, int i, const int end,
                                            const float betaSq, const float2 ( &px )[FX / 2], const float py, const float pz,
                                            float2 ( &sum )[FX / 2], int ( &fR )[FX], const float pxLane, unsigned char* __restrict__ smList,
                                            float* __restrict__ smRes )
{
    const unsigned lane = threadIdx.x & 31;
    while ( i < end )
    {
        const float4* np = reinterpret_cast<const float4*>( nodes + i );
        const float4 q2 = __ldg( np + 2 ), q3 = __ldg( np + 3 );
        const int link = __float_as_int( q3.w );
        const float thr = FM( betaSq, q2.w );
        if ( link >= 0 )
        {
            if ( thr >= 2.f * FLT_MIN ) fwn6Inner<false>( i, link, thr, q2, q3, px, py, pz, sum, fR );
            else fwn6Inner<true>( i, link, thr, q2, q3, px, py, pz, sum, fR );
        }
        else
        {
            const NodeLane L = nodeLane( q2, q3, py, pz );
            unsigned go = 0;
#pragma unroll
            for ( int k = 0; k < FX / 2; ++k )
            {
                const float2 dx = __fadd2_rn( bc2( q2.x ), negate2( px[k] ) );
                const float2 dd = exactDd( dx, L );
                float2 dot, r3, ddc;
                termPair( dx, L, q3.x, dot, r3, ddc );
                const bool a0 = i >= fR[2 * k], a1 = i >= fR[2 * k + 1];
                const bool far0 = !( dd.x <= thr ), far1 = !( dd.y <= thr );
                if ( a0 && far0 && dd.x >= FLT_MIN && ddc.x >= FLT_MIN ) sum[k].x = fmaf( dot.x, r3.x, sum[k].x );
                if ( a1 && far1 && dd.y >= FLT_MIN && ddc.y >= FLT_MIN ) sum[k].y = fmaf( dot.y, r3.y, sum[k].y );
                go |= ( a0 && !far0 ? 1u : 0u ) << ( 2 * k );
                go |= ( a1 && !far1 ? 1u : 0u ) << ( 2 * k + 1 );
                fR[2 * k] = a0 ? i + 1 : fR[2 * k];
                fR[2 * k + 1] = a1 ? i + 1 : fR[2 * k + 1];
            }
            if ( __any_sync( FULL, go != 0 ) )
            {
                // The voxels that are near this triangle are few and scattered over lanes and voxel slots (ncu, c5_band: the exact
                // solid angle ran with 12 of 32 lanes active, slot by slot, and took 31 % of the kernel's time).  They are
                // compacted: every (lane, slot) pair becomes one work item in a per-warp list, the warp evaluates 32 items at a
                // time with all lanes busy and hands the results back through shared memory.  Each voxel still adds its own term
                // at the same place of its sum: the bits do not change.
                const float4* tp = reinterpret_cast<const float4*>( tris + ( -link - 1 ) );
                const float4 ta = __ldg( tp ), tb = __ldg( tp + 1 ), tc = __ldg( tp + 2 );
                const int cnt = __popc( go );
                int incl = cnt;
#pragma unroll
                for ( int d = 1; d < 32; d <<= 1 )
                {
                    const int t = __shfl_up_sync( FULL, incl, d );
                    if ( lane >= (unsigned)d ) incl += t;
                }
                const int total = __shfl_sync( FULL, incl, 31 );
                int pos = incl - cnt;
                for ( unsigned g = go; g; g &= g - 1 ) smList[pos++] = (unsigned char)( lane * FX + ( __ffs( g ) - 1 ) );
                __syncwarp();
                for ( int t0 = 0; t0 < total; t0 += 32 )
                {
                    const int t = t0 + (int)lane;
                    const unsigned code = t < total ? smList[t] : 0u;
                    const int L = (int)( code >> 3 ), v = (int)( code & 7u );
                    const float qx = __shfl_sync( FULL, pxLane, v ); // lane v holds the x of voxel slot v (the same for every lane)
                    const float qy = __shfl_sync( FULL, py, L ), qz = __shfl_sync( FULL, pz, L );
                    if ( t < total ) smRes[( v << 5 ) + L] = triangleSolidAngle( make_float3( qx, qy, qz ), ta, tb, tc );
                }
                __syncwarp();
#pragma unroll
                for ( int k = 0; k < FX / 2; ++k )
                {
                    if ( go & ( 1u << ( 2 * k ) ) ) sum[k].x = FA( sum[k].x, smRes[( ( 2 * k ) << 5 ) + lane] );
                    if ( go & ( 2u << ( 2 * k ) ) ) sum[k].y = FA( sum[k].y, smRes[( ( 2 * k + 1 ) << 5 ) + lane] );
                }
                __syncwarp();
            }
        }
        i = max( __reduce_min_sync( FULL, minOf( fR ) ), i + 1 );
    }
    return end;
}

template <int MODE, bool PEERS, int MINB = 7>
__global__ void __launch_bounds__( GRID_TB, MINB ) kGridFwn6( const TravNode* __restrict__ nodes, const LeafTri* __restrict__ tris,
                                                        const GridParams P, float* __restrict__ out, const BrickOrder O, const __grid_constant__ PeerOut peers )
{
    // work items of a leaf's exact solid angles and their results (fwn6Section), per warp
    __shared__ unsigned char sList[GRID_TB / 32][32 * FX];
    __shared__ float sRes[GRID_TB / 32][32 * FX];
    const unsigned lane = threadIdx.x & 31;
    size_t brick;
    if ( !brickIndexOfWarp( P, O, brick ) ) return;
    const int bx = (int)( brick % P.bricksX ), by = (int)( ( brick / P.bricksX ) % P.bricksY ), bz = (int)( brick / ( (size_t)P.bricksX * P.bricksY ) );
    const int x0 = bx * F6X, y = by * F6Y + (int)( lane & 7 ), lz = bz * F6Z + (int)( lane >> 3 );
    const int z = globalZ( P, bz * F6Z ) + (int)( lane >> 3 ); // a brick never straddles two chunks
    const bool rowValid = y < P.dy && lz < P.nz;
    const size_t id = (size_t)x0 + (size_t)P.dx * ( (size_t)y + (size_t)P.dy * (size_t)( z - P.zBegin ) );
    const int N = P.numNodes;

    // voxels that get a winding number: inside the grid and, MODE 1, not NaN (MRFastWindingNumber.cpp:72-74)
    unsigned valid = 0;
    if ( rowValid )
    {
#pragma unroll
        for ( int h = 0; h < 2; ++h )
        {
            if ( x0 + 4 * h >= P.dx ) continue;
            if ( P.vecStore ) // dx % 4 == 0: the four voxels are inside the grid
            {
                unsigned m = 15u;
                if ( MODE == 1 )
                {
                    const float4 d4 = *reinterpret_cast<const float4*>( out + id + 4 * h );
                    m = ( d4.x == d4.x ? 1u : 0u ) | ( d4.y == d4.y ? 2u : 0u ) | ( d4.z == d4.z ? 4u : 0u ) | ( d4.w == d4.w ? 8u : 0u );
                }
                valid |= m << ( 4 * h );
            }
            else
            {
#pragma unroll
                for ( int v = 4 * h; v < 4 * h + 4; ++v )
                    if ( x0 + v < P.dx )
                    {
                        bool ok = true;
                        if ( MODE == 1 ) { const float d = out[id + v]; ok = d == d; }
                        valid |= ( ok ? 1u : 0u ) << v;
                    }
            }
        }
    }
    if ( !__any_sync( FULL, valid != 0 ) ) return; // nothing to compute, the distances (NaN) stay as they are

    float2 px[FX / 2], sum[FX / 2];
    float py, pz;
    {
        const float3 p0 = applyXf( P.A, P.b, (float)x0, (float)y, (float)z );
        py = p0.y; pz = p0.z; // axis-aligned grid: y and z do not depend on x
#pragma unroll
        for ( int k = 0; k < FX / 2; ++k )
        {
            px[k] = make_float2( applyXf( P.A, P.b, (float)( x0 + 2 * k ), (float)y, (float)z ).x, applyXf( P.A, P.b, (float)( x0 + 2 * k + 1 ), (float)y, (float)z ).x );
            sum[k] = make_float2( 0.f, 0.f );
        }
    }
    // x of voxel slot (lane & 7), the same bits as px[] (same function, same arguments): lane v serves slot v in fwn6Section
    const float pxLane = applyXf( P.A, P.b, (float)( x0 + (int)( lane & 7 ) ), (float)y, (float)z ).x;
    // ball of the voxels that take part: centre of the brick, radius to the farthest such voxel
    const float3 pc = applyXf( P.A, P.b, (float)x0 + 3.5f, (float)( by * F6Y ) + 3.5f, (float)globalZ( P, bz * F6Z ) + 1.5f );
    float radSq = 0.f;
    {
        const float ey = py - pc.y, ez = pz - pc.z;
        const float eyz = fmaf( ey, ey, ez * ez );
#pragma unroll
        for ( int v = 0; v < FX; ++v )
        {
            const float ex = ( ( v & 1 ) ? px[v / 2].y : px[v / 2].x ) - pc.x;
            if ( valid & ( 1u << v ) ) radSq = fmaxf( radSq, fmaf( ex, ex, eyz ) );
        }
    }
    radSq = __uint_as_float( __reduce_max_sync( FULL, __float_as_uint( radSq ) ) );
    const BrickBall ball = makeBrickBall( pc.x, pc.y, pc.z, radSq );

    int fR[FX];
    int i = 0;
    // (r02: classifying 32 consecutive nodes at once -- lane l takes node i + l, two ballots, the warp then walks the window on
    // the bits -- was built and measured: bit-identical results, but c5_band 1069 ms against 857 ms for this node-by-node loop.
    // A window is thrown away whenever a skip link leaves it, which is what most ALL_FAR and MIXED nodes do, and the two
    // ballots plus the second fetch of the node sit on the critical path of every step.)
    while ( i < N )
    {
        const float4* np = reinterpret_cast<const float4*>( nodes + i );
        const float4 q2 = __ldg( np + 2 ), q3 = __ldg( np + 3 );
        const int link = __float_as_int( q3.w );
        const float thr = FM( P.betaSq, q2.w );
        const float s = brickDistSq( ball, q2.x, q2.y, q2.z ), rhs = s * ball.r4;
        if ( link >= 0 && brickAllNear( ball, s, rhs, thr ) ) { ++i; continue; }
        const int next = skipOf( link, i );
        if ( brickAllFar( ball, s, rhs, thr ) )
        {
            fwn6AddAll( q2, q3, px, py, pz, sum );
            i = next;
            continue;
        }
#pragma unroll
        for ( int v = 0; v < FX; ++v ) fR[v] = ( valid & ( 1u << v ) ) ? i : INT_MAX;
        i = fwn6Section( nodes, tris, i, next, P.betaSq, px, py, pz, sum, fR, pxLane, sList[threadIdx.x >> 5], sRes[threadIdx.x >> 5] );
    }

    if ( !rowValid ) return;
    float4 r8[2] = { make_float4( 0.f, 0.f, 0.f, 0.f ), make_float4( 0.f, 0.f, 0.f, 0.f ) };
#pragma unroll
    for ( int h = 0; h < 2; ++h )
    {
        if ( x0 + 4 * h >= P.dx ) continue;
        const float s4[4] = { sum[2 * h].x, sum[2 * h].y, sum[2 * h + 1].x, sum[2 * h + 1].y };
        if ( P.vecStore )
        {
            float4 d4 = make_float4( 0.f, 0.f, 0.f, 0.f );
            if ( MODE == 1 ) d4 = *reinterpret_cast<const float4*>( out + id + 4 * h );
            const float4 r4 = make_float4( fwnResult( MODE, s4[0], d4.x, P.threshold ), fwnResult( MODE, s4[1], d4.y, P.threshold ),
                                           fwnResult( MODE, s4[2], d4.z, P.threshold ), fwnResult( MODE, s4[3], d4.w, P.threshold ) );
            *reinterpret_cast<float4*>( out + id + 4 * h ) = r4;
            if ( PEERS )
            {
                r8[h] = r4;
                // rows of whole 32-byte sectors (dx % 8 == 0, all volumes 32-B aligned: P.vecStore == 2): the lane's 8 results go to
                // every peer as ONE 256-bit store after the loop.  Two 16-byte stores per row, issued by two different
                // instructions, reached the peers as two partial-sector writes each: r02, N = 8: 234 ms per rank for 107 ms of
                // work, 30 GB of posted writes per GPU at 128 GB/s (profiles/r02/scale/)
                if ( P.vecStore != 2 )
                    for ( int k = 0; k < peers.n; ++k ) *reinterpret_cast<float4*>( peers.p[k] + id + 4 * h ) = r4;
            }
        }
        else
        {
#pragma unroll
            for ( int v = 0; v < 4; ++v )
                if ( x0 + 4 * h + v < P.dx )
                {
                    const float d = MODE == 1 ? out[id + 4 * h + v] : 0.f;
                    const float r = fwnResult( MODE, s4[v], d, P.threshold );
                    out[id + 4 * h + v] = r;
                    if ( PEERS )
                        for ( int k = 0; k < peers.n; ++k ) peers.p[k][id + 4 * h + v] = r;
                }
        }
    }
    if ( PEERS && P.vecStore == 2 ) // dx % 8 == 0: both halves of the row lie inside the grid
        for ( int k = 0; k < peers.n; ++k )
            asm volatile( "st.global.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" :: "l"( peers.p[k] + id ), "f"( r8[0].x ), "f"( r8[0].y ),
                          "f"( r8[0].z ), "f"( r8[0].w ), "f"( r8[1].x ), "f"( r8[1].y ), "f"( r8[1].z ), "f"( r8[1].w ) : "memory" );
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

void printStats( const char* what, const unsigned long long* h, double voxels, double bricks )
{
    fprintf( stderr, "[b2fwn stats] %s: voxels %.0f bricks %.0f | visits/brick %.1f tests/voxel %.1f leaf-loads/brick %.1f exact/voxel %.2f "
             "plane-rejected/voxel %.2f\n", what, voxels, bricks, h[0] / bricks, h[1] / voxels, h[2] / bricks, h[3] / voxels, h[4] / voxels );
    fprintf( stderr, "[b2fwn stats] %s: longest walk %llu visits at brick %llu; bricks by log2(visits):", what, h[STAT_MAX] >> 32, h[STAT_MAX] & 0xFFFFFFFFull );
    for ( int k = 0; k < 33; ++k )
        if ( h[STAT_HIST + k] ) fprintf( stderr, " [<2^%d]=%llu", k, h[STAT_HIST + k] );
    fprintf( stderr, "\n" );
}

// MODE 0: winding numbers only; MODE 1: distance walk, then winding-number walk over the same slab.
// signOnly (MODE 1): the slab already holds (unsigned) distances -- no distance walk, the winding-number walk alone negates
// the values of the voxels with w > threshold (NaN voxels are left as they are and are not walked)
template <int MODE>
int launchGrid( const DeviceTree& t, const int32_t dims[3], const GridXf& xf, float beta, const b2fwn_dist_options* opt,
                int zBegin, int zEnd, const ZChunks* zc, const PeerOut* peers, float* dOut, cudaStream_t st, uint64_t* launches,
                const bool signOnly = false )
{
    const bool distPass = MODE == 1 && !signOnly;
    if ( int e = checkDims( dims, zBegin, zEnd ) ) return e;
    if ( peers && peers->n > 0 && devSwitch( "B2FWN_STATS" ) ) { setError( "peer output is not available in stats mode" ); return B2FWN_ERR_INVALID; }
    if ( dims[0] == 0 || dims[1] == 0 || zEnd == zBegin ) return B2FWN_OK;
    GridParams P{};
    P.dx = dims[0]; P.dy = dims[1]; P.zBegin = zBegin;
    if ( zc )
    {
        // layers zBegin + c * stride + k, k < chunk, below zEnd; dOut points at layer zBegin of a volume that has them all
        if ( zc->chunk <= 0 || zc->chunk % BZ != 0 || zc->stride < zc->chunk )
        {
            setError( "invalid z-chunk layout (chunk must be a positive multiple of 4, stride >= chunk)" );
            return B2FWN_ERR_INVALID;
        }
        const int span = zEnd - zBegin;
        P.nz = ( span / zc->stride ) * zc->chunk + std::min( span % zc->stride, zc->chunk );
        P.zChunk = zc->chunk; P.zSkip = zc->stride - zc->chunk;
    }
    else { P.nz = zEnd - zBegin; P.zChunk = INT_MAX / 2; P.zSkip = 0; }
    P.bricksX = ( dims[0] + BX - 1 ) / BX;
    P.bricksY = ( dims[1] + BY - 1 ) / BY;
    P.bricksZ = ( P.nz + BZ - 1 ) / BZ;
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
    static const bool wantStats = devSwitch( "B2FWN_STATS" ) != nullptr;
    static const bool heavyFirst = !( devSwitch( "B2FWN_LINEAR" ) && atoi( devSwitch( "B2FWN_LINEAR" ) ) ); // dev switch: plain linear order
    static const bool forceV5 = devSwitch( "B2FWN_FWN5" ) && atoi( devSwitch( "B2FWN_FWN5" ) ); // dev switch: v5 winding walk on axis-aligned grids too
    // winding walk: v6 (8 x 8 x 4 bricks) on axis-aligned grids, v5 (8 x 4 x 4) on general affine grids; the stats build counts v4's walk
    const bool v6 = aa && !forceV5 && !wantStats;
    GridParams PF = P; // brick layout of the winding kernel
    if ( v6 )
    {
        PF.bricksX = ( dims[0] + F6X - 1 ) / F6X;
        PF.bricksY = ( dims[1] + F6Y - 1 ) / F6Y;
        PF.bricksZ = ( P.nz + F6Z - 1 ) / F6Z;
    }
    const size_t numBricksF = (size_t)PF.bricksX * PF.bricksY * PF.bricksZ;
    const bool band = distPass && P.maxDistSq < FLT_MAX && P.numNodes > 0;
    // small grids (c1: 16,384 bricks, 6.4 ms) lose more to the estimate kernels than a tail can cost them
    const bool order = heavyFirst && !wantStats && P.numNodes > 0 && numBricks >= ( 1u << 16 ) && numBricks <= 0xFFFFFFFFull;

    // scratch of this call: per-brick walk-length estimates (2 x u16), heavy lists (2 x cap u32), class bytes (2 x u8),
    // totals (4 x u64) and list counters (2 x u32) in one stream-ordered allocation (sized for the distance bricks; the
    // winding kernel never has more)
    const unsigned cap = order ? (unsigned)std::min<size_t>( numBricks, 1u << 16 ) : 0;
    const unsigned capF = order ? (unsigned)std::min<size_t>( numBricksF, cap ) : 0;
    const size_t nbPad = ( numBricks + 15 ) & ~(size_t)15;
    uint8_t* scratch = nullptr;
    const size_t scratchBytes = 64 + 2 * nbPad * sizeof( uint16_t ) + 2 * nbPad + 2 * (size_t)cap * sizeof( uint32_t );
    if ( band || order )
    {
        B2_CUDA( cudaMallocFromPoolAsync( &scratch, scratchBytes, t.pool, st ) );
        B2_CUDA( cudaMemsetAsync( scratch, 0, 64, st ) );
    }
    unsigned long long* totals = reinterpret_cast<unsigned long long*>( scratch ); // [0,1] distance, [2,3] winding number
    unsigned* counts = reinterpret_cast<unsigned*>( scratch + 32 );               // [0] distance, [1] winding number
    uint16_t* costD = reinterpret_cast<uint16_t*>( scratch + 64 );
    uint16_t* costF = costD + nbPad;
    uint8_t* clsD = reinterpret_cast<uint8_t*>( costF + nbPad );
    uint8_t* clsF = clsD + nbPad;
    uint32_t* heavyD = reinterpret_cast<uint32_t*>( clsF + nbPad );
    uint32_t* heavyF = heavyD + cap;
    const unsigned nb256 = (unsigned)( ( numBricks + 255 ) / 256 );
    BrickOrder oD, oF;
    // The winding-number estimates do not depend on the distance kernels: with distances (MODE 1) they run on the tree's
    // high-priority side stream beside kBrickCull / kGridDist and are joined before kGridFwn5 -- their cost is latency (one
    // thread walks up to COST_MAX nodes), not throughput, and per launch: the host-output calls pay it once per chunk.
    const bool sideEst = order && distPass && t.sideStream;
    auto fwnEstimates = [&]( cudaStream_t es )
    {
        const int gX = ( PF.bricksX + GRP - 1 ) / GRP, gY = ( PF.bricksY + GRP - 1 ) / GRP, gZ = ( PF.bricksZ + GRP - 1 ) / GRP;
        const unsigned gBlocks = (unsigned)( ( (size_t)gX * gY * gZ + 255 ) / 256 ), bBlocks = (unsigned)( ( numBricksF + 255 ) / 256 );
        if ( v6 )
        {
            kGroupCostFwn<F6X, F6Y, F6Z><<<gBlocks, 256, 0, es>>>( t.nodes, PF, gX, gY, gZ, costF, totals + 2 );
            kGroupHeavy<F6X, F6Y, F6Z><<<bBlocks, 256, 0, es>>>( t.nodes, costF, PF, gX, gY, totals + 2, heavyF, counts + 1, clsF, capF );
        }
        else
        {
            kGroupCostFwn<BX, BY, BZ><<<gBlocks, 256, 0, es>>>( t.nodes, PF, gX, gY, gZ, costF, totals + 2 );
            kGroupHeavy<BX, BY, BZ><<<bBlocks, 256, 0, es>>>( t.nodes, costF, PF, gX, gY, totals + 2, heavyF, counts + 1, clsF, capF );
        }
        if ( launches ) *launches += 2;
        oF.heavy = heavyF; oF.count = counts + 1; oF.cls = clsF; oF.cap = capF;
    };
    if ( sideEst )
    {
        B2_CUDA( cudaEventRecord( t.evFork, st ) ); // after the scratch allocation and its memset
        B2_CUDA( cudaStreamWaitEvent( t.sideStream, t.evFork, 0 ) );
        fwnEstimates( t.sideStream );
        B2_CUDA( cudaEventRecord( t.evJoin, t.sideStream ) );
    }
    if ( band )
    {
        kBrickCull<<<nb256, 256, 0, st>>>( t.nodes, P, costD, totals );
        if ( launches ) ++*launches;
        if ( order )
        {
            kBrickHeavy<<<nb256, 256, 0, st>>>( costD, numBricks, totals, heavyD, counts, clsD, cap );
            if ( launches ) ++*launches;
            oD.heavy = heavyD; oD.count = counts; oD.cls = clsD; oD.cap = cap;
        }
    }
    if ( order && !sideEst ) fwnEstimates( st );
    const uint16_t* cullCost = band ? costD : nullptr;
    auto blocksFor = [&]( const BrickOrder& o ) { return (unsigned)( ( numBricks + o.cap + GRID_TB / 32 - 1 ) / ( GRID_TB / 32 ) ); };
    const unsigned blocksF = (unsigned)( ( numBricksF + oF.cap + GRID_TB / 32 - 1 ) / ( GRID_TB / 32 ) );
    if ( wantStats )
    {
        unsigned long long* dStats = nullptr; unsigned long long h[2 * STAT_WORDS];
        B2_CUDA( cudaMalloc( &dStats, sizeof( h ) ) );
        B2_CUDA( cudaMemsetAsync( dStats, 0, sizeof( h ), st ) );
        if ( distPass )
        {
            if ( aa ) kGridDist<true, true><<<blocksFor( oD ), GRID_TB, 0, st>>>( t.nodes, t.leafTris, t.leafFilt, P, dOut, cullCost, oD, dStats );
            else kGridDist<false, true><<<blocksFor( oD ), GRID_TB, 0, st>>>( t.nodes, t.leafTris, t.leafFilt, P, dOut, cullCost, oD, dStats );
        }
        if ( aa ) kGridFwn<MODE, true, true><<<blocksFor( oF ), GRID_TB, 0, st>>>( t.nodes, t.leafTris, P, dOut, oF, dStats + STAT_WORDS );
        else kGridFwn<MODE, false, true><<<blocksFor( oF ), GRID_TB, 0, st>>>( t.nodes, t.leafTris, P, dOut, oF, dStats + STAT_WORDS );
        B2_CUDA( cudaMemcpyAsync( h, dStats, sizeof( h ), cudaMemcpyDeviceToHost, st ) );
        B2_CUDA( cudaStreamSynchronize( st ) );
        cudaFree( dStats );
        const double nv = (double)dims[0] * dims[1] * P.nz, nb = (double)numBricks;
        if ( distPass ) printStats( "dist", h, nv, nb );
        printStats( "fwn ", h + STAT_WORDS, nv, nb );
    }
    else
    {
        if ( distPass )
        {
            // (r02: 6 / 5 / 4 blocks per SM -- 80 / 96 / 118 registers, fewer or no spills -- measured again on the full-range walk:
            //  c3_full 533 ms at 8 blocks, 571 at 5, 602 at 4; c1 6.1 / 6.0 / 6.2 / 6.5 ms; c3_band 34.5 -> 38.6 ms: 8 stays)
            if ( aa ) kGridDist<true, false><<<blocksFor( oD ), GRID_TB, 0, st>>>( t.nodes, t.leafTris, t.leafFilt, P, dOut, cullCost, oD, nullptr );
            else kGridDist<false, false><<<blocksFor( oD ), GRID_TB, 0, st>>>( t.nodes, t.leafTris, t.leafFilt, P, dOut, cullCost, oD, nullptr );
        }
        if ( sideEst ) B2_CUDA( cudaStreamWaitEvent( st, t.evJoin, 0 ) );
        PeerOut po{};
        const bool toPeers = peers && peers->n > 0;
        if ( toPeers )
        {
            po = *peers;
            for ( int k = 0; k < po.n; ++k )
                if ( reinterpret_cast<uintptr_t>( po.p[k] ) % 16 != 0 ) PF.vecStore = 0;
            if ( v6 && PF.vecStore && dims[0] % 8 == 0 )
            {
                bool a32 = true; // the peer volumes (the only targets of the 256-bit stores)
                for ( int k = 0; k < po.n; ++k ) a32 = a32 && reinterpret_cast<uintptr_t>( po.p[k] ) % 32 == 0;
                static const bool noV8 = devSwitch( "B2FWN_NO_V8" ) != nullptr; // dev switch: two 16-byte stores per row, as before
                if ( a32 && !noV8 ) PF.vecStore = 2;
            }
        }
        if ( v6 )
        {
            if ( toPeers ) kGridFwn6<MODE, true><<<blocksF, GRID_TB, 0, st>>>( t.nodes, t.leafTris, PF, dOut, oF, po );
            else kGridFwn6<MODE, false><<<blocksF, GRID_TB, 0, st>>>( t.nodes, t.leafTris, PF, dOut, oF, po );
        }
        else if ( toPeers )
        {
            if ( aa ) kGridFwn5<MODE, true, true><<<blocksF, GRID_TB, 0, st>>>( t.nodes, t.leafTris, PF, dOut, oF, po );
            else kGridFwn5<MODE, false, true><<<blocksF, GRID_TB, 0, st>>>( t.nodes, t.leafTris, PF, dOut, oF, po );
        }
        else if ( aa ) kGridFwn5<MODE, true, false><<<blocksF, GRID_TB, 0, st>>>( t.nodes, t.leafTris, PF, dOut, oF, po );
        else kGridFwn5<MODE, false, false><<<blocksF, GRID_TB, 0, st>>>( t.nodes, t.leafTris, PF, dOut, oF, po );
    }
    if ( launches ) *launches += ( distPass ? 2 : 1 );
    static const bool debugOrder = devSwitch( "B2FWN_DEBUG" ) != nullptr;
    if ( debugOrder && scratch )
    {
        unsigned long long h[8];
        B2_CUDA( cudaMemcpyAsync( h, scratch, sizeof( h ), cudaMemcpyDeviceToHost, st ) );
        B2_CUDA( cudaStreamSynchronize( st ) );
        const unsigned* hc = reinterpret_cast<const unsigned*>( h + 4 );
        fprintf( stderr, "[b2fwn order] bricks %zu cap %u | dist: cost sum %llu over %llu bricks, heavy %u | fwn: cost sum %llu over %llu groups, heavy %u\n",
                 numBricks, cap, h[0], h[1], hc[0], h[2], h[3], hc[1] );
    }
    if ( scratch ) B2_CUDA( cudaFreeAsync( scratch, st ) );
    B2_CUDA( cudaGetLastError() );
    return B2FWN_OK;
}

} // anonymous namespace

// Distance pass alone for SignDetectionMode::ProjectionNormal over the layers [zBegin, zEnd) of a volume: dOut = |d| (NaN outside
// [minDistSq, maxDistSq) when nullOutside), dSlot = leaf slot of the closest triangle per voxel (-1 = none).  Voxel centres are
// origin + voxelSize * ( pos + 0.5 ); minDistSq must be 0 (an early exit would depend on the visiting order).
int launchDistSlotGrid( const DeviceTree& t, const int32_t dims[3], const float origin[3], const float voxelSize[3], float maxDistSq,
                        int nullOutside, int zBegin, int zEnd, float* dOut, int32_t* dSlot, cudaStream_t st, uint64_t* launches )
{
    if ( int e = checkDims( dims, zBegin, zEnd ) ) return e;
    if ( dims[0] == 0 || dims[1] == 0 || zEnd == zBegin ) return B2FWN_OK;
    GridParams P{};
    P.dx = dims[0]; P.dy = dims[1]; P.zBegin = zBegin;
    P.nz = zEnd - zBegin; P.zChunk = INT_MAX / 2; P.zSkip = 0;
    P.bricksX = ( dims[0] + BX - 1 ) / BX;
    P.bricksY = ( dims[1] + BY - 1 ) / BY;
    P.bricksZ = ( P.nz + BZ - 1 ) / BZ;
    P.A[0] = voxelSize[0]; P.A[4] = voxelSize[1]; P.A[8] = voxelSize[2];
    P.minDistSq = 0.f; P.maxDistSq = maxDistSq; P.nullOutside = nullOutside;
    P.numNodes = (int)t.numNodes;
    P.vecStore = ( dims[0] % VX == 0 ) && ( reinterpret_cast<uintptr_t>( dOut ) % 16 == 0 );
    const size_t numBricks = (size_t)P.bricksX * P.bricksY * P.bricksZ;
    const size_t blocks = ( numBricks + GRID_TB / 32 - 1 ) / ( GRID_TB / 32 );
    if ( blocks > 0x7FFFFFFFull ) { setError( "grid too large for one launch" ); return B2FWN_ERR_INVALID; }
    // the brick cull works on boxes of voxel centres: the usual basis form b = origin + 0.5 * voxelSize is exact enough for it
    // (its boxes are padded); the distance kernel itself takes the origin form
    GridParams Pc = P;
    for ( int k = 0; k < 3; ++k ) { Pc.b[k] = origin[k] + 0.5f * voxelSize[k]; P.b[k] = origin[k]; }
    const bool band = maxDistSq < FLT_MAX && P.numNodes > 0;
    uint8_t* scratch = nullptr;
    const size_t nbPad = ( numBricks + 15 ) & ~(size_t)15;
    uint16_t* costD = nullptr;
    if ( band )
    {
        B2_CUDA( cudaMallocFromPoolAsync( &scratch, 64 + nbPad * sizeof( uint16_t ), t.pool, st ) );
        B2_CUDA( cudaMemsetAsync( scratch, 0, 64, st ) );
        costD = reinterpret_cast<uint16_t*>( scratch + 64 );
        kBrickCull<<<(unsigned)( ( numBricks + 255 ) / 256 ), 256, 0, st>>>( t.nodes, Pc, costD, reinterpret_cast<unsigned long long*>( scratch ) );
        if ( launches ) ++*launches;
    }
    kGridDist<true, false, 8, true><<<(unsigned)blocks, GRID_TB, 0, st>>>( t.nodes, t.leafTris, t.leafFilt, P, dOut, costD, BrickOrder{}, nullptr, dSlot );
    if ( launches ) ++*launches;
    if ( scratch ) B2_CUDA( cudaFreeAsync( scratch, st ) );
    B2_CUDA( cudaGetLastError() );
    return B2FWN_OK;
}

int launchFwnGrid( const DeviceTree& t, const int32_t dims[3], const GridXf& xf, float beta, int zBegin, int zEnd,
                   float* dOut, cudaStream_t st, uint64_t* launches, const ZChunks* zc )
{
    return launchGrid<0>( t, dims, xf, beta, nullptr, zBegin, zEnd, zc, nullptr, dOut, st, launches );
}

int launchSdfGrid( const DeviceTree& t, const int32_t dims[3], const GridXf& xf, const b2fwn_dist_options& opt,
                   int zBegin, int zEnd, float* dOut, cudaStream_t st, uint64_t* launches, const ZChunks* zc, const PeerOut* peers )
{
    return launchGrid<1>( t, dims, xf, opt.winding_number_beta, &opt, zBegin, zEnd, zc, peers, dOut, st, launches );
}

// warp-cooperative winding numbers of a query vector in a spatially coherent order (b2fwn_kernels.cu decides when to use it)
int launchFwnQueriesWarp( const DeviceTree& t, const float* dPts, const uint32_t* dOrder, size_t n, bool faces, float beta, int32_t skipFace,
                          float* dOutW, uint32_t* dBits32, cudaStream_t st, uint64_t* launches )
{
    if ( n == 0 || t.numNodes == 0 ) return B2FWN_OK;
    const size_t warps = ( n + 32 * QV - 1 ) / ( 32 * QV );
    const size_t blocks = ( warps + GRID_TB / 32 - 1 ) / ( GRID_TB / 32 );
    if ( blocks > 0x7FFFFFFFull ) { setError( "too many queries for one launch" ); return B2FWN_ERR_INVALID; }
    if ( faces )
        kFwnQueriesWarp<true><<<(unsigned)blocks, GRID_TB, 0, st>>>( t.nodes, t.leafTris, (int)t.numNodes, t.points, t.tris, nullptr, n, beta * beta, -1, dOutW, dBits32 );
    else
        kFwnQueriesWarp<false><<<(unsigned)blocks, GRID_TB, 0, st>>>( t.nodes, t.leafTris, (int)t.numNodes, dPts, nullptr, dOrder, n, beta * beta,
                                                                      skipFace < 0 ? -1 : skipFace, dOutW, nullptr );
    if ( launches ) ++*launches;
    B2_CUDA( cudaGetLastError() );
    return B2FWN_OK;
}

// In-place sign of a distance slab: value -> -value where the winding number of the voxel exceeds the threshold
// (the per-voxel update of makeSignedByWindingNumber, source/MRVoxels/MRVDBConversions.cpp:517-533)
int launchSignGrid( const DeviceTree& t, const int32_t dims[3], const GridXf& xf, float beta, float threshold, int zBegin, int zEnd,
                    float* dVolumeAtZBegin, cudaStream_t st, uint64_t* launches )
{
    b2fwn_dist_options o{};
    o.min_dist_sq = 0; o.max_dist_sq = FLT_MAX; o.null_outside_min_max = 0;
    o.winding_number_threshold = threshold; o.winding_number_beta = beta;
    return launchGrid<1>( t, dims, xf, beta, &o, zBegin, zEnd, nullptr, nullptr, dVolumeAtZBegin, st, launches, true );
}

} // namespace b2
