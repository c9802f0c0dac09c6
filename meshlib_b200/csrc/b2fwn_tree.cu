// b2fwn_tree.cu -- device build of the AABB tree and of the dipole (far-field) data.
//
// Compiled with -fmad=false: every float operation below rounds exactly like the reference's scalar CPU code
// (no FMA contraction on baseline x86-64), so that the device-built tree and dipoles are BIT-IDENTICAL to
//   MR::AABBTree::AABBTree(MeshPart)       source/MRMesh/MRAABBTree.cpp:15-60
//   AABBTreeMaker::construct / partitionLeaves_ / makeNode_   source/MRMesh/MRAABBTreeMaker.hpp:51-171
//   MR::calcDipoles                         source/MRMesh/MRDipole.cpp:11-71
//
// The reference splits every subtree at the median of a STRICT TOTAL ORDER (box centre along the box's
// dimensions from longest to shortest, then leaf id), so the leaf SET of every node is independent of thread
// count and of the selection algorithm.  We therefore build the identical tree breadth-first: one segmented
// sort per level (all segments of a level at once), node indices in DFS pre-order computed in closed form
// (left = i+1, right = i + 2*leftLeaves).  ceil(log2 F)+1 levels, everything stays in HBM.
#include "b2fwn_internal.h"
#include <cub/device/device_merge_sort.cuh>
#include <cub/device/device_select.cuh>
#include <thrust/iterator/counting_iterator.h>
#include <cfloat>
#include <cstring>
#include <cstdlib>
#include <algorithm>
#include <vector>

namespace b2
{

namespace
{

constexpr int TB = 256;

// per-leaf record kept during the build
struct LeafBox
{
    float mn[3];
    float mx[3];
    int32_t face;
};

// sort key of one leaf at one level: (segment, centre along longest, middle, shortest dim, leaf id)
struct LevelKey
{
    uint32_t seg;
    float k1, k2, k3;
    int32_t face;
    int32_t leaf; // index into LeafBox array
};

struct LevelKeyLess
{
    __device__ __forceinline__ bool operator()( const LevelKey& a, const LevelKey& b ) const
    {
        if ( a.seg != b.seg ) return a.seg < b.seg;
        // MRAABBTreeMaker.hpp:56-69
        if ( a.k1 != b.k1 ) return a.k1 < b.k1;
        if ( a.k2 != b.k2 ) return a.k2 < b.k2;
        if ( a.k3 != b.k3 ) return a.k3 < b.k3;
        return a.face < b.face;
    }
};

// segment (= subtree under construction) of one level
struct Seg
{
    int32_t first; // first leaf position
    int32_t n;     // number of leaves; 0 = empty slot
    int32_t node;  // pre-order node index of the subtree root
    int32_t done;  // leaf node already written
};

__device__ __forceinline__ uint32_t f2ord( float f )
{
    uint32_t u = __float_as_uint( f );
    return ( u & 0x80000000u ) ? ~u : ( u | 0x80000000u );
}
__device__ __forceinline__ float ord2f( uint32_t u )
{
    return __uint_as_float( ( u & 0x80000000u ) ? ( u & 0x7fffffffu ) : ~u );
}

// a face is valid when none of its ids is negative (MeshTopology::getTriangulation marks deleted faces that way); an id beyond
// the vertex array is a malformed triangulation: counted, so that the build fails with an error instead of reading out of bounds
__global__ void kValidFlags( const int32_t* __restrict__ tris, size_t nFaces, size_t nVerts, uint8_t* __restrict__ flags, int* __restrict__ outOfRange )
{
    size_t f = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if ( f >= nFaces ) return;
    const int a = tris[3 * f], b = tris[3 * f + 1], c = tris[3 * f + 2];
    const bool valid = a >= 0 && b >= 0 && c >= 0;
    if ( valid && ( (size_t)a >= nVerts || (size_t)b >= nVerts || (size_t)c >= nVerts ) ) { atomicAdd( outOfRange, 1 ); flags[f] = 0; return; }
    flags[f] = valid ? 1 : 0;
}

// computeFaceBox (MRAABBTree.cpp:15-29) + Box::insignificantlyExpanded (MRBox.h:260-270)
// also reduces the exact (unexpanded) bounding box of all valid faces into bbox6 (ordered-integer images)
__global__ void kFaceBoxes( const float* __restrict__ pts, const int32_t* __restrict__ tris,
                            const int32_t* __restrict__ validFaces, int nLeaves, LeafBox* __restrict__ out, uint32_t* __restrict__ bbox6 )
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t bb[6] = { 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u, 0u };
    if ( i < nLeaves )
    {
    const int f = validFaces[i];
    LeafBox lb;
    lb.face = f;
#pragma unroll
    for ( int d = 0; d < 3; ++d )
    {
        const float a = pts[3 * (size_t)tris[3 * (size_t)f] + d];
        const float b = pts[3 * (size_t)tris[3 * (size_t)f + 1] + d];
        const float c = pts[3 * (size_t)tris[3 * (size_t)f + 2] + d];
        // Box::include: if (pt < min) min = pt; if (pt > max) max = pt  (MRBox.h)
        float mn = FLT_MAX, mx = -FLT_MAX;
        if ( a < mn ) mn = a; if ( a > mx ) mx = a;
        if ( b < mn ) mn = b; if ( b > mx ) mx = b;
        if ( c < mn ) mn = c; if ( c > mx ) mx = c;
        lb.mn[d] = nextafterf( mn, -FLT_MAX );
        lb.mx[d] = nextafterf( mx, FLT_MAX );
        bb[d] = f2ord( mn ); bb[3 + d] = f2ord( mx );
    }
    out[i] = lb;
    }
#pragma unroll
    for ( int c = 0; c < 6; ++c )
    {
        const uint32_t r = c < 3 ? __reduce_min_sync( 0xFFFFFFFFu, bb[c] ) : __reduce_max_sync( 0xFFFFFFFFu, bb[c] );
        if ( ( threadIdx.x & 31 ) == 0 )
        {
            if ( c < 3 ) atomicMin( bbox6 + c, r ); else atomicMax( bbox6 + c, r );
        }
    }
}

__global__ void kInitPerm( int nLeaves, LevelKey* __restrict__ keys, const LeafBox* __restrict__ leaves )
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if ( i >= nLeaves ) return;
    LevelKey k;
    k.seg = 0; k.k1 = k.k2 = k.k3 = 0; k.face = leaves[i].face; k.leaf = i;
    keys[i] = k;
}

__global__ void kClearSegBoxes( uint32_t* __restrict__ segBox, size_t nSegs )
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if ( i >= nSegs * 6 ) return;
    // min slots start at +inf-ordered, max slots at -inf-ordered
    segBox[i] = ( i % 6 < 3 ) ? 0xFFFFFFFFu : 0u;
}

// union of the leaf boxes of every segment (makeNode_, MRAABBTreeMaker.hpp:80-83): segmented warp reduction over
// runs of equal segment id, then one atomic per run and component on order-preserving integer images
__global__ void kSegBoxes( int nLeaves, const LevelKey* __restrict__ keys, const LeafBox* __restrict__ leaves,
                           uint32_t* __restrict__ segBox )
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned lane = threadIdx.x & 31;
    const bool valid = i < nLeaves;
    uint32_t seg = 0xFFFFFFFFu;
    uint32_t v[6];
#pragma unroll
    for ( int c = 0; c < 6; ++c ) v[c] = c < 3 ? 0xFFFFFFFFu : 0u;
    if ( valid )
    {
        const LevelKey k = keys[i];
        seg = k.seg;
        const LeafBox lb = leaves[k.leaf];
#pragma unroll
        for ( int c = 0; c < 3; ++c ) { v[c] = f2ord( lb.mn[c] ); v[3 + c] = f2ord( lb.mx[c] ); }
    }
    // segments are contiguous runs; reduce towards the head lane of each run
#pragma unroll
    for ( int off = 1; off < 32; off <<= 1 )
    {
        const uint32_t oseg = __shfl_down_sync( 0xFFFFFFFFu, seg, off );
        uint32_t ov[6];
#pragma unroll
        for ( int c = 0; c < 6; ++c ) ov[c] = __shfl_down_sync( 0xFFFFFFFFu, v[c], off );
        if ( lane + off < 32 && oseg == seg )
        {
#pragma unroll
            for ( int c = 0; c < 3; ++c ) { v[c] = min( v[c], ov[c] ); v[3 + c] = max( v[3 + c], ov[3 + c] ); }
        }
    }
    const uint32_t pseg = __shfl_up_sync( 0xFFFFFFFFu, seg, 1 );
    const bool head = valid && ( lane == 0 || pseg != seg );
    if ( head )
    {
        uint32_t* dst = segBox + (size_t)seg * 6;
#pragma unroll
        for ( int c = 0; c < 3; ++c ) { atomicMin( dst + c, v[c] ); atomicMax( dst + 3 + c, v[3 + c] ); }
    }
}

// sort keys of a level: findSortedBoxDims (MRBox.h:376-396) + the comparator of partitionLeaves_
__global__ void kMakeKeys( int nLeaves, LevelKey* __restrict__ keys, const LeafBox* __restrict__ leaves,
                           const uint32_t* __restrict__ segBox, const Seg* __restrict__ segs )
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if ( i >= nLeaves ) return;
    LevelKey k = keys[i];
    if ( segs[k.seg].n <= 1 ) { k.k1 = k.k2 = k.k3 = 0; keys[i] = k; return; }
    const uint32_t* sb = segBox + (size_t)k.seg * 6;
    float ext[3]; int dim[3] = { 0, 1, 2 };
#pragma unroll
    for ( int c = 0; c < 3; ++c ) ext[c] = ord2f( sb[3 + c] ) - ord2f( sb[c] );
    // bubble sort of pairs (extent, axis), ascending, std::pair operator<
#pragma unroll
    for ( int a = 0; a + 1 < 3; ++a )
#pragma unroll
        for ( int b = a + 1; b < 3; ++b )
            if ( ext[b] < ext[a] || ( !( ext[a] < ext[b] ) && dim[b] < dim[a] ) )
            {
                float te = ext[a]; ext[a] = ext[b]; ext[b] = te;
                int td = dim[a]; dim[a] = dim[b]; dim[b] = td;
            }
    const LeafBox lb = leaves[k.leaf];
    float c2[3];
#pragma unroll
    for ( int c = 0; c < 3; ++c ) c2[c] = lb.mn[c] + lb.mx[c];
    k.k1 = dim[2] == 0 ? c2[0] : dim[2] == 1 ? c2[1] : c2[2];
    k.k2 = dim[1] == 0 ? c2[0] : dim[1] == 1 ? c2[1] : c2[2];
    k.k3 = dim[0] == 0 ? c2[0] : dim[0] == 1 ? c2[1] : c2[2];
    keys[i] = k;
}

// after the sort: move every leaf into the left or right child segment of the next level
__global__ void kDescend( int nLeaves, LevelKey* __restrict__ keys, const Seg* __restrict__ segs )
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if ( i >= nLeaves ) return;
    const uint32_t s = keys[i].seg;
    const Seg sg = segs[s];
    uint32_t ns = 2u * s;
    if ( sg.n > 1 && i - sg.first >= sg.n / 2 ) ns += 1;
    keys[i].seg = ns;
}

// writes the nodes of this level and the segment table of the next one
__global__ void kEmitLevel( uint32_t nSegs, int level, const Seg* __restrict__ segs, Seg* __restrict__ next,
                            const uint32_t* __restrict__ segBox, const LevelKey* __restrict__ keys,
                            const LeafBox* __restrict__ leaves, RefNode* __restrict__ nodes,
                            int32_t* __restrict__ skip, uint8_t* __restrict__ depth, int32_t* __restrict__ leafNodeOfPos )
{
    const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
    if ( s >= nSegs ) return;
    const Seg sg = segs[s];
    Seg l = { 0, 0, 0, 0 }, r = { 0, 0, 0, 0 };
    if ( sg.n == 1 )
    {
        if ( !sg.done )
        {
            const LevelKey k = keys[sg.first];
            const LeafBox lb = leaves[k.leaf];
            RefNode nd;
#pragma unroll
            for ( int c = 0; c < 3; ++c ) { nd.bmin[c] = lb.mn[c]; nd.bmax[c] = lb.mx[c]; }
            nd.l = lb.face; nd.r = -1; // setLeafId (MRAABBTreeNode.h:41)
            nodes[sg.node] = nd;
            skip[sg.node] = sg.node + 1;
            depth[sg.node] = (uint8_t)level;
            leafNodeOfPos[sg.first] = sg.node;
        }
        l = { sg.first, 1, sg.node, 1 };
    }
    else if ( sg.n > 1 )
    {
        const int nl = sg.n / 2; // midLeaf = first + n/2 (MRAABBTreeMaker.hpp:71)
        RefNode nd;
        const uint32_t* sb = segBox + (size_t)s * 6;
#pragma unroll
        for ( int c = 0; c < 3; ++c ) { nd.bmin[c] = ord2f( sb[c] ); nd.bmax[c] = ord2f( sb[3 + c] ); }
        nd.l = sg.node + 1;
        nd.r = sg.node + 2 * nl; // root + 1 + getNumNodes(leftNumLeaves) (MRAABBTreeMaker.hpp:88-89)
        nodes[sg.node] = nd;
        skip[sg.node] = sg.node + 2 * sg.n - 1;
        depth[sg.node] = (uint8_t)level;
        l = { sg.first, nl, sg.node + 1, 0 };
        r = { sg.first + nl, sg.n - nl, sg.node + 2 * nl, 0 };
    }
    next[2 * (size_t)s] = l;
    next[2 * (size_t)s + 1] = r;
}

// ---- the lower levels of the build, one thread block per subtree --------------------------------------------------------
// Once the segments of a level hold at most SUB_CAP leaves, a block takes one segment and finishes its whole subtree in shared
// memory: per level the same steps as above (segment boxes, keys, sort by (segment, k1, k2, k3, face), emit, descend), the sort
// being a bitonic network over the block's leaves.  The global passes sort ALL leaves once per level whatever the segment sizes
// (2.9 ms per level at 10 M leaves); for the last 10 of 24 levels this kernel replaces them.  Same comparator, same strict
// total order, same closed-form node numbering: the tree is bit-identical (test_tree_and_dipoles_bit_exact, golden fixtures).
constexpr int SUB_CAP = 1024;
constexpr int SUB_TB = 256;

struct SubSeg
{
    uint16_t first, n;
    int32_t node;
    uint32_t done;
};

struct SubShared
{
    float mn[3][SUB_CAP], mx[3][SUB_CAP];
    float k1[SUB_CAP], k2[SUB_CAP], k3[SUB_CAP];
    int32_t face[SUB_CAP], leaf[SUB_CAP];
    uint32_t box[SUB_CAP][6];
    SubSeg seg[2][SUB_CAP];
    uint16_t segOf[SUB_CAP]; // by element
    uint16_t ord[SUB_CAP];   // position -> element, 0xFFFF = padding
};

__device__ __forceinline__ bool subLess( const SubShared& S, unsigned a, unsigned b )
{
    if ( a == 0xFFFFu ) return false;
    if ( b == 0xFFFFu ) return true;
    if ( S.segOf[a] != S.segOf[b] ) return S.segOf[a] < S.segOf[b];
    if ( S.k1[a] != S.k1[b] ) return S.k1[a] < S.k1[b]; // MRAABBTreeMaker.hpp:56-69
    if ( S.k2[a] != S.k2[b] ) return S.k2[a] < S.k2[b];
    if ( S.k3[a] != S.k3[b] ) return S.k3[a] < S.k3[b];
    return S.face[a] < S.face[b];
}

__global__ void __launch_bounds__( SUB_TB ) kSubtrees( const Seg* __restrict__ segs, int level0, int levels, LevelKey* __restrict__ keys,
                                                       const LeafBox* __restrict__ leaves, RefNode* __restrict__ nodes, int32_t* __restrict__ skip,
                                                       uint8_t* __restrict__ depth, int32_t* __restrict__ leafNodeOfPos )
{
    extern __shared__ __align__( 16 ) unsigned char subRaw[];
    SubShared& S = *reinterpret_cast<SubShared*>( subRaw );
    const Seg root = segs[blockIdx.x];
    const int n = root.n;
    if ( n <= 0 || ( n == 1 && root.done ) ) return;
    const int tid = threadIdx.x;
    int P = 1;
    while ( P < n ) P <<= 1;
    for ( int i = tid; i < P; i += SUB_TB )
    {
        if ( i < n )
        {
            const LevelKey k = keys[root.first + i];
            const LeafBox lb = leaves[k.leaf];
#pragma unroll
            for ( int c = 0; c < 3; ++c ) { S.mn[c][i] = lb.mn[c]; S.mx[c][i] = lb.mx[c]; }
            S.face[i] = lb.face; S.leaf[i] = k.leaf;
            S.segOf[i] = 0;
            S.ord[i] = (uint16_t)i;
        }
        else S.ord[i] = 0xFFFFu;
    }
    if ( tid == 0 ) S.seg[0][0] = SubSeg{ 0, (uint16_t)n, root.node, (uint32_t)root.done };
    __syncthreads();
    int cur = 0;
    for ( int lev = level0; lev <= levels; ++lev )
    {
        const int nSegs = 1 << ( lev - level0 );
        const bool split = lev < levels;
        if ( split )
        {
            for ( int i = tid; i < nSegs * 6; i += SUB_TB ) S.box[i / 6][i % 6] = ( i % 6 < 3 ) ? 0xFFFFFFFFu : 0u;
            __syncthreads();
            // union of the leaf boxes of every segment that will be split (makeNode_, MRAABBTreeMaker.hpp:80-83)
            for ( int e = tid; e < n; e += SUB_TB )
            {
                const int sg = S.segOf[e];
                if ( S.seg[cur][sg].n > 1 )
                {
#pragma unroll
                    for ( int c = 0; c < 3; ++c ) { atomicMin( &S.box[sg][c], f2ord( S.mn[c][e] ) ); atomicMax( &S.box[sg][3 + c], f2ord( S.mx[c][e] ) ); }
                }
            }
            __syncthreads();
            // keys: findSortedBoxDims (MRBox.h:376-396) + the comparator of partitionLeaves_ (as kMakeKeys)
            for ( int e = tid; e < n; e += SUB_TB )
            {
                const int sg = S.segOf[e];
                float a1 = 0, a2 = 0, a3 = 0;
                if ( S.seg[cur][sg].n > 1 )
                {
                    float ext[3]; int dim[3] = { 0, 1, 2 };
#pragma unroll
                    for ( int c = 0; c < 3; ++c ) ext[c] = ord2f( S.box[sg][3 + c] ) - ord2f( S.box[sg][c] );
#pragma unroll
                    for ( int a = 0; a + 1 < 3; ++a )
#pragma unroll
                        for ( int b = a + 1; b < 3; ++b )
                            if ( ext[b] < ext[a] || ( !( ext[a] < ext[b] ) && dim[b] < dim[a] ) )
                            {
                                float te = ext[a]; ext[a] = ext[b]; ext[b] = te;
                                int td = dim[a]; dim[a] = dim[b]; dim[b] = td;
                            }
                    float c2[3];
#pragma unroll
                    for ( int c = 0; c < 3; ++c ) c2[c] = S.mn[c][e] + S.mx[c][e];
                    a1 = dim[2] == 0 ? c2[0] : dim[2] == 1 ? c2[1] : c2[2];
                    a2 = dim[1] == 0 ? c2[0] : dim[1] == 1 ? c2[1] : c2[2];
                    a3 = dim[0] == 0 ? c2[0] : dim[0] == 1 ? c2[1] : c2[2];
                }
                S.k1[e] = a1; S.k2[e] = a2; S.k3[e] = a3;
            }
            __syncthreads();
            // bitonic sort of the positions by (segment, k1, k2, k3, face)
            for ( int k = 2; k <= P; k <<= 1 )
                for ( int j = k >> 1; j > 0; j >>= 1 )
                {
                    for ( int i = tid; i < P; i += SUB_TB )
                    {
                        const int ixj = i ^ j;
                        if ( ixj > i )
                        {
                            const unsigned a = S.ord[i], b = S.ord[ixj];
                            const bool up = ( i & k ) == 0;
                            if ( subLess( S, b, a ) == up ) { S.ord[i] = (uint16_t)b; S.ord[ixj] = (uint16_t)a; }
                        }
                    }
                    __syncthreads();
                }
        }
        // nodes of this level and the segment table of the next (as kEmitLevel)
        for ( int sidx = tid; sidx < nSegs; sidx += SUB_TB )
        {
            const SubSeg sg = S.seg[cur][sidx];
            SubSeg l = { 0, 0, 0, 0 }, r = { 0, 0, 0, 0 };
            if ( sg.n == 1 )
            {
                if ( !sg.done )
                {
                    const int e = S.ord[sg.first];
                    RefNode nd;
#pragma unroll
                    for ( int c = 0; c < 3; ++c ) { nd.bmin[c] = S.mn[c][e]; nd.bmax[c] = S.mx[c][e]; }
                    nd.l = S.face[e]; nd.r = -1; // setLeafId (MRAABBTreeNode.h:41)
                    nodes[sg.node] = nd;
                    skip[sg.node] = sg.node + 1;
                    depth[sg.node] = (uint8_t)lev;
                    leafNodeOfPos[root.first + sg.first] = sg.node;
                }
                l = SubSeg{ sg.first, 1, sg.node, 1 };
            }
            else if ( sg.n > 1 )
            {
                const int nl = sg.n / 2; // midLeaf = first + n/2 (MRAABBTreeMaker.hpp:71)
                RefNode nd;
#pragma unroll
                for ( int c = 0; c < 3; ++c ) { nd.bmin[c] = ord2f( S.box[sidx][c] ); nd.bmax[c] = ord2f( S.box[sidx][3 + c] ); }
                nd.l = sg.node + 1;
                nd.r = sg.node + 2 * nl; // root + 1 + getNumNodes(leftNumLeaves) (MRAABBTreeMaker.hpp:88-89)
                nodes[sg.node] = nd;
                skip[sg.node] = sg.node + 2 * sg.n - 1;
                depth[sg.node] = (uint8_t)lev;
                l = SubSeg{ sg.first, (uint16_t)nl, sg.node + 1, 0 };
                r = SubSeg{ (uint16_t)( sg.first + nl ), (uint16_t)( sg.n - nl ), sg.node + 2 * nl, 0 };
            }
            if ( split ) { S.seg[cur ^ 1][2 * sidx] = l; S.seg[cur ^ 1][2 * sidx + 1] = r; }
        }
        __syncthreads();
        if ( !split ) break;
        // every leaf moves into the left or right child segment (as kDescend)
        for ( int pos = tid; pos < n; pos += SUB_TB )
        {
            const int e = S.ord[pos];
            const int sidx = S.segOf[e];
            const SubSeg sg = S.seg[cur][sidx];
            int ns = 2 * sidx;
            if ( sg.n > 1 && pos - sg.first >= sg.n / 2 ) ns += 1;
            S.segOf[e] = (uint16_t)ns;
        }
        __syncthreads();
        cur ^= 1;
    }
    // final leaf order of the subtree
    for ( int pos = tid; pos < n; pos += SUB_TB )
    {
        const int e = S.ord[pos];
        keys[root.first + pos].face = S.face[e];
        keys[root.first + pos].leaf = S.leaf[e];
    }
}

// ---- dipoles (MRDipole.cpp:23-71) ---------------------------------------------------------------------------------

// leaves: da = 0.5f*cross(b-a, c-a); area = |da|; pos = area * ((1/3.0f)*(a+b+c))   (MRDipole.cpp:29-45,
// MRMeshMath.cpp:119-124, 140-149)
__global__ void kLeafDipoles( size_t nNodes, const RefNode* __restrict__ nodes, const float* __restrict__ pts,
                              const int32_t* __restrict__ tris, RefDipole* __restrict__ dip )
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if ( i >= nNodes ) return;
    const RefNode nd = nodes[i];
    if ( nd.r >= 0 ) return;
    const size_t f = (size_t)nd.l;
    float a[3], b[3], c[3];
#pragma unroll
    for ( int d = 0; d < 3; ++d )
    {
        a[d] = pts[3 * (size_t)tris[3 * f] + d];
        b[d] = pts[3 * (size_t)tris[3 * f + 1] + d];
        c[d] = pts[3 * (size_t)tris[3 * f + 2] + d];
    }
    const float ux = b[0] - a[0], uy = b[1] - a[1], uz = b[2] - a[2];
    const float vx = c[0] - a[0], vy = c[1] - a[1], vz = c[2] - a[2];
    const float cx = uy * vz - uz * vy, cy = uz * vx - ux * vz, cz = ux * vy - uy * vx;
    const float dax = 0.5f * cx, day = 0.5f * cy, daz = 0.5f * cz;
    const float area = sqrtf( dax * dax + day * day + daz * daz );
    const float third = 1 / 3.0f;
    const float ctrx = third * ( ( a[0] + b[0] ) + c[0] ), ctry = third * ( ( a[1] + b[1] ) + c[1] ), ctrz = third * ( ( a[2] + b[2] ) + c[2] );
    RefDipole d;
    d.pos[0] = area * ctrx; d.pos[1] = area * ctry; d.pos[2] = area * ctrz;
    d.area = area;
    d.dirArea[0] = dax; d.dirArea[1] = day; d.dirArea[2] = daz;
    d.rr = 0;
    dip[i] = d;
}

// inner nodes of one depth: dipole = left + right (MRDipole.cpp:48-61); deeper levels are complete already
__global__ void kInnerDipoles( size_t nNodes, int level, const uint8_t* __restrict__ depth,
                               const RefNode* __restrict__ nodes, RefDipole* __restrict__ dip )
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if ( i >= nNodes || depth[i] != level ) return;
    const RefNode nd = nodes[i];
    if ( nd.r < 0 ) return;
    const RefDipole dl = dip[nd.l], dr = dip[nd.r];
    RefDipole d;
#pragma unroll
    for ( int c = 0; c < 3; ++c ) { d.pos[c] = dl.pos[c] + dr.pos[c]; d.dirArea[c] = dl.dirArea[c] + dr.dirArea[c]; }
    d.area = dl.area + dr.area;
    d.rr = 0;
    dip[i] = d;
}

// pos /= area (as pos *= 1/area, MRVector3.h:131-137) and rr = distToFarthestCornerSq (MRDipole.cpp:11-21, 63-70);
// also packs the 64-B traversal node
__global__ void kFinishDipoles( size_t nNodes, const RefNode* __restrict__ nodes, RefDipole* __restrict__ dip,
                                const int32_t* __restrict__ skip, const int32_t* __restrict__ leafSlotOfNode,
                                TravNode* __restrict__ trav )
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if ( i >= nNodes ) return;
    const RefNode nd = nodes[i];
    RefDipole d = dip[i];
    if ( d.area > 0 )
    {
        const float inv = 1 / d.area;
        d.pos[0] *= inv; d.pos[1] *= inv; d.pos[2] *= inv;
    }
    float rr = 0;
#pragma unroll
    for ( int c = 0; c < 3; ++c )
    {
        const float a = d.pos[c] - nd.bmin[c], b = d.pos[c] - nd.bmax[c];
        const float toMinSq = a * a, toMaxSq = b * b;
        rr += fmaxf( toMinSq, toMaxSq );
    }
    d.rr = rr;
    dip[i] = d;
    const int link = skip ? ( nd.r < 0 ? -( leafSlotOfNode[i] + 1 ) : skip[i] ) : __float_as_int( trav[i].q3.w ); // refit: links stay
    TravNode t;
    t.q0 = make_float4( nd.bmin[0], nd.bmin[1], nd.bmin[2], nd.bmax[0] );
    t.q1 = make_float4( nd.bmax[1], nd.bmax[2], __int_as_float( nd.r < 0 ? 0 : nd.r ), __int_as_float( link ) ); // .z = right child
    t.q2 = make_float4( d.pos[0], d.pos[1], d.pos[2], d.rr );
    t.q3 = make_float4( d.dirArea[0], d.dirArea[1], d.dirArea[2], __int_as_float( link ) );
    trav[i] = t;
}

// leaf slot = leaf position in the final sorted order (= pre-order rank of the leaf)
__global__ void kLeafSlots( int nLeaves, const int32_t* __restrict__ leafNodeOfPos, int32_t* __restrict__ leafSlotOfNode,
                            const LevelKey* __restrict__ keys, const float* __restrict__ pts,
                            const int32_t* __restrict__ tris, LeafTri* __restrict__ out, LeafFilt* __restrict__ filt )
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if ( p >= nLeaves ) return;
    size_t f;
    if ( leafNodeOfPos ) { leafSlotOfNode[leafNodeOfPos[p]] = p; f = (size_t)keys[p].face; }
    else f = (size_t)__float_as_int( out[p].a.w ); // refit: the same leaf order at new coordinates
    const int32_t ia = tris[3 * f], ib = tris[3 * f + 1], ic = tris[3 * f + 2];
    LeafTri t;
    t.a = make_float4( pts[3 * (size_t)ia], pts[3 * (size_t)ia + 1], pts[3 * (size_t)ia + 2], __int_as_float( (int)f ) );
    t.b = make_float4( pts[3 * (size_t)ib], pts[3 * (size_t)ib + 1], pts[3 * (size_t)ib + 2], 0.f );
    t.c = make_float4( pts[3 * (size_t)ic], pts[3 * (size_t)ic + 1], pts[3 * (size_t)ic + 2], 0.f );
    // the triangle in its own frame, in double, rounded once (LeafFilt, b2fwn_internal.h)
    const double ux = (double)t.b.x - t.a.x, uy = (double)t.b.y - t.a.y, uz = (double)t.b.z - t.a.z;
    const double vx = (double)t.c.x - t.a.x, vy = (double)t.c.y - t.a.y, vz = (double)t.c.z - t.a.z;
    double nx = uy * vz - uz * vy, ny = uz * vx - ux * vz, nz = ux * vy - uy * vx;
    const double nlen = sqrt( nx * nx + ny * ny + nz * nz ), bx = sqrt( ux * ux + uy * uy + uz * uz );
    const double lenCA2 = vx * vx + vy * vy + vz * vz;
    const double ex = vx - ux, ey = vy - uy, ez = vz - uz, lenBC2 = ex * ex + ey * ey + ez * ez;
    LeafFilt fr;
    // slivers (sine of the angle at a below 1e-3) and degenerate triangles always take the exact path
    const bool good = bx > 0 && lenCA2 > 0 && lenBC2 > 0 && nlen > 1e-3 * bx * sqrt( lenCA2 );
    if ( good )
    {
        nx /= nlen; ny /= nlen; nz /= nlen;
        const double e1x = ux / bx, e1y = uy / bx, e1z = uz / bx;
        const double e2x = ny * e1z - nz * e1y, e2y = nz * e1x - nx * e1z, e2z = nx * e1y - ny * e1x; // v = n x u
        const double cx = vx * e1x + vy * e1y + vz * e1z, cy = vx * e2x + vy * e2y + vz * e2z;
        const float mag = fmaxf( fmaxf( fabsf( t.a.x ), fabsf( t.b.x ) ), fabsf( t.c.x ) ) + fmaxf( fmaxf( fabsf( t.a.y ), fabsf( t.b.y ) ), fabsf( t.c.y ) )
                          + fmaxf( fmaxf( fabsf( t.a.z ), fabsf( t.b.z ) ), fabsf( t.c.z ) );
        fr.f0 = make_float4( t.a.x, t.a.y, t.a.z, (float)bx );
        fr.f1 = make_float4( (float)e1x, (float)e1y, (float)e1z, (float)cx );
        fr.f2 = make_float4( (float)nx, (float)ny, (float)nz, (float)cy );
        fr.f3 = make_float4( (float)( 1.0 / lenBC2 ), (float)( 1.0 / lenCA2 ), 3e-6f * (float)( bx + sqrt( lenCA2 ) ) + 1e-6f * mag, 0.f );
    }
    else
    {
        fr.f0 = make_float4( t.a.x, t.a.y, t.a.z, 0.f );
        fr.f1 = make_float4( 1.f, 0.f, 0.f, 0.f );
        fr.f2 = make_float4( 0.f, 0.f, 1.f, 1.f );
        fr.f3 = make_float4( 1.f, 1.f, __int_as_float( 0x7f800000 ), 0.f );
    }
    filt[p] = fr;
    out[p] = t;
}

// ---- refit (AABBTree::refit, MRAABBTree.cpp:62-94, with every vertex marked as changed) -----------------------------------

// leaf nodes: computeFaceBox at the new coordinates; also the exact bounding box of the valid faces (as kFaceBoxes)
__global__ void kRefitLeafNodes( size_t nNodes, RefNode* __restrict__ nodes, const float* __restrict__ pts, const int32_t* __restrict__ tris,
                                 uint32_t* __restrict__ bbox6 )
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    uint32_t bb[6] = { 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u, 0u };
    if ( i < nNodes && nodes[i].r < 0 )
    {
        const size_t f = (size_t)nodes[i].l;
#pragma unroll
        for ( int d = 0; d < 3; ++d )
        {
            const float a = pts[3 * (size_t)tris[3 * f] + d];
            const float b = pts[3 * (size_t)tris[3 * f + 1] + d];
            const float c = pts[3 * (size_t)tris[3 * f + 2] + d];
            float mn = FLT_MAX, mx = -FLT_MAX;
            if ( a < mn ) mn = a; if ( a > mx ) mx = a;
            if ( b < mn ) mn = b; if ( b > mx ) mx = b;
            if ( c < mn ) mn = c; if ( c > mx ) mx = c;
            nodes[i].bmin[d] = nextafterf( mn, -FLT_MAX );
            nodes[i].bmax[d] = nextafterf( mx, FLT_MAX );
            bb[d] = f2ord( mn ); bb[3 + d] = f2ord( mx );
        }
    }
#pragma unroll
    for ( int c = 0; c < 6; ++c )
    {
        const uint32_t r = c < 3 ? __reduce_min_sync( 0xFFFFFFFFu, bb[c] ) : __reduce_max_sync( 0xFFFFFFFFu, bb[c] );
        if ( ( threadIdx.x & 31 ) == 0 )
        {
            if ( c < 3 ) atomicMin( bbox6 + c, r ); else atomicMax( bbox6 + c, r );
        }
    }
}

// inner nodes of one depth: box = left box, include( right box ) (MRAABBTree.cpp:90-91); deeper levels are complete already
__global__ void kRefitInnerNodes( size_t nNodes, int level, const uint8_t* __restrict__ depth, RefNode* __restrict__ nodes )
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if ( i >= nNodes || depth[i] != level ) return;
    const int l = nodes[i].l, r = nodes[i].r;
    if ( r < 0 ) return;
#pragma unroll
    for ( int d = 0; d < 3; ++d )
    {
        float mn = nodes[l].bmin[d], mx = nodes[l].bmax[d];
        const float rmn = nodes[r].bmin[d], rmx = nodes[r].bmax[d];
        if ( rmn < mn ) mn = rmn; // Box::include( Box ), MRBox.h
        if ( rmx > mx ) mx = rmx;
        nodes[i].bmin[d] = mn; nodes[i].bmax[d] = mx;
    }
}

inline unsigned nblk( size_t n ) { return (unsigned)( ( n + TB - 1 ) / TB ); }

} // anonymous namespace

int initTreeAux( DeviceTree& t, int device )
{
    if ( !t.sideStream )
    {
        int lo = 0, hi = 0;
        B2_CUDA( cudaDeviceGetStreamPriorityRange( &lo, &hi ) ); // hi = greatest priority (numerically lowest)
        B2_CUDA( cudaStreamCreateWithPriority( &t.sideStream, cudaStreamNonBlocking, hi ) );
        B2_CUDA( cudaEventCreateWithFlags( &t.evFork, cudaEventDisableTiming ) );
        B2_CUDA( cudaEventCreateWithFlags( &t.evJoin, cudaEventDisableTiming ) );
    }
    if ( !t.pool )
    {
        cudaMemPoolProps props{};
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = device;
        B2_CUDA( cudaMemPoolCreate( &t.pool, &props ) );
        unsigned long long keep = ~0ull; // a host-output call synchronises per chunk: keep the scratch cached in between
        B2_CUDA( cudaMemPoolSetAttribute( t.pool, cudaMemPoolAttrReleaseThreshold, &keep ) );
    }
    return B2FWN_OK;
}

void freeTree( DeviceTree& t )
{
    cudaFree( t.points ); cudaFree( t.tris ); cudaFree( t.arena ); // refNodes, refDipoles, nodes, leafTris, leafFilt, nodeDepth
    cudaFree( t.faceNormals ); cudaFree( t.vertNormals ); cudaFree( t.faceAdj );
    if ( t.sideStream ) cudaStreamDestroy( t.sideStream );
    if ( t.evFork ) cudaEventDestroy( t.evFork );
    if ( t.evJoin ) cudaEventDestroy( t.evJoin );
    if ( t.pool ) cudaMemPoolDestroy( t.pool );
    t = DeviceTree{};
}

int buildTree( DeviceTree& t, cudaStream_t st, uint64_t* launches )
{
    uint64_t nl = 0;
    const size_t F = t.numFaces;
    t.numLeaves = t.numNodes = 0; t.depth = 0;
    if ( !t.sideStream || !t.pool )
    {
        int dev = 0;
        B2_CUDA( cudaGetDevice( &dev ) );
        if ( int e = initTreeAux( t, dev ) ) return e;
    }
    if ( F == 0 ) return B2FWN_OK;
    if ( F > (size_t)0x3FFFFFFF ) { setError( "too many faces" ); return B2FWN_ERR_INVALID; }

    // every temporary of the build is released on every way out (an out-of-memory failure half way through a 10 M-face build
    // would otherwise leak gigabytes: b2fwn_destroy only knows the arrays the tree keeps)
    struct Temps
    {
        std::vector<void*> p;
        ~Temps() { for ( void* q : p ) cudaFree( q ); }
        cudaError_t alloc( void** ptr, size_t bytes )
        {
            const cudaError_t e = cudaMalloc( ptr, bytes );
            if ( e == cudaSuccess ) p.push_back( *ptr );
            return e;
        }
        void release( void* q ) { for ( auto& x : p ) if ( x == q ) { cudaFree( q ); x = nullptr; } }
        void keep( void* q ) { for ( auto& x : p ) if ( x == q ) x = nullptr; }
    } temps;
#define T_ALLOC( ptr, bytes ) B2_CUDA( temps.alloc( (void**)&( ptr ), ( bytes ) ) )
    // 1. valid faces in increasing id order (MRAABBTree.cpp:35-46)
    uint8_t* dFlags = nullptr; int32_t* dValid = nullptr; int* dNum = nullptr;
    T_ALLOC( dFlags, F );
    T_ALLOC( dValid, F * sizeof( int32_t ) );
    T_ALLOC( dNum, 2 * sizeof( int ) );
    B2_CUDA( cudaMemsetAsync( dNum, 0, 2 * sizeof( int ), st ) );
    kValidFlags<<<nblk( F ), TB, 0, st>>>( t.tris, F, t.numVerts, dFlags, dNum + 1 ); ++nl;
    {
        size_t tmpBytes = 0;
        thrust::counting_iterator<int32_t> ids( 0 );
        cub::DeviceSelect::Flagged( nullptr, tmpBytes, ids, dFlags, dValid, dNum, (int)F, st );
        void* selTmp = nullptr;
        T_ALLOC( selTmp, tmpBytes );
        cub::DeviceSelect::Flagged( selTmp, tmpBytes, ids, dFlags, dValid, dNum, (int)F, st );
        int num[2] = { 0, 0 };
        B2_CUDA( cudaMemcpyAsync( num, dNum, sizeof( num ), cudaMemcpyDeviceToHost, st ) );
        B2_CUDA( cudaStreamSynchronize( st ) );
        temps.release( selTmp );
        if ( num[1] > 0 ) { setError( "triangulation refers to vertex ids beyond the vertex array" ); return B2FWN_ERR_INVALID; }
        t.numLeaves = (size_t)num[0];
    }
    temps.release( dFlags ); temps.release( dNum );
    const int L = (int)t.numLeaves;
    if ( L == 0 ) return B2FWN_OK;
    t.numNodes = 2 * (size_t)L - 1;
    const size_t N = t.numNodes;
    int levels = 0; // number of splitting levels = ceil(log2 L)
    while ( ( (size_t)1 << levels ) < (size_t)L ) ++levels;
    t.depth = levels;

    LeafBox* dLeaves = nullptr; LevelKey *dKeys = nullptr; Seg *dSegA = nullptr, *dSegB = nullptr;
    uint32_t* dSegBox = nullptr; int32_t *dSkip = nullptr, *dLeafNodeOfPos = nullptr, *dLeafSlotOfNode = nullptr;
    uint8_t* dDepth = nullptr;
    // the global passes stop at level0 (below it one block finishes each subtree): the segment tables only grow that far
    static const bool noSubtrees = devSwitch( "B2FWN_NO_SUBTREES" ) != nullptr; // dev switch: global passes down to the leaves
    const int level0 = noSubtrees ? levels + 1 : std::max( 0, levels - 10 ); // segments of level0 hold <= SUB_CAP = 2^10 leaves
    const size_t maxSegs = (size_t)1 << std::min( levels, level0 );
    size_t sortTmpBytes = 0; void* sortTmp = nullptr;
    cub::DeviceMergeSort::SortKeys( nullptr, sortTmpBytes, dKeys, L, LevelKeyLess(), st );
    // ONE allocation for all temporaries of the build and one for the arrays the tree keeps: cudaMalloc / cudaFree are slow and
    // device-synchronising, and their cost varies from box to box (r02, 10 M faces: 16 ms of the build on one box, ~70 ms on another,
    // with twenty separate allocations)
    uint32_t* dBBox = nullptr;
    {
        size_t off = 0;
        auto span = [&]( size_t bytes ) { const size_t at = off; off += ( bytes + 255 ) & ~(size_t)255; return at; };
        const size_t oLeaves = span( (size_t)L * sizeof( LeafBox ) ), oKeys = span( (size_t)L * sizeof( LevelKey ) );
        const size_t oSegA = span( 2 * maxSegs * sizeof( Seg ) ), oSegB = span( 2 * maxSegs * sizeof( Seg ) );
        const size_t oSegBox = span( maxSegs * 6 * sizeof( uint32_t ) ), oSkip = span( N * sizeof( int32_t ) );
        const size_t oPos = span( (size_t)L * sizeof( int32_t ) ), oSlot = span( N * sizeof( int32_t ) );
        const size_t oBBox = span( 6 * sizeof( uint32_t ) ), oSort = span( sortTmpBytes );
        uint8_t* arena = nullptr;
        T_ALLOC( arena, off );
        dLeaves = reinterpret_cast<LeafBox*>( arena + oLeaves ); dKeys = reinterpret_cast<LevelKey*>( arena + oKeys );
        dSegA = reinterpret_cast<Seg*>( arena + oSegA ); dSegB = reinterpret_cast<Seg*>( arena + oSegB );
        dSegBox = reinterpret_cast<uint32_t*>( arena + oSegBox ); dSkip = reinterpret_cast<int32_t*>( arena + oSkip );
        dLeafNodeOfPos = reinterpret_cast<int32_t*>( arena + oPos ); dLeafSlotOfNode = reinterpret_cast<int32_t*>( arena + oSlot );
        dBBox = reinterpret_cast<uint32_t*>( arena + oBBox ); sortTmp = arena + oSort;
    }
    {
        size_t off = 0;
        auto span = [&]( size_t bytes ) { const size_t at = off; off += ( bytes + 255 ) & ~(size_t)255; return at; };
        const size_t oRefN = span( N * sizeof( RefNode ) ), oRefD = span( N * sizeof( RefDipole ) ), oNodes = span( N * sizeof( TravNode ) );
        const size_t oTris = span( (size_t)L * sizeof( LeafTri ) ), oFilt = span( (size_t)L * sizeof( LeafFilt ) ), oDepth = span( N );
        B2_CUDA( cudaMalloc( &t.arena, off ) );
        uint8_t* a = static_cast<uint8_t*>( t.arena );
        t.refNodes = reinterpret_cast<RefNode*>( a + oRefN ); t.refDipoles = reinterpret_cast<RefDipole*>( a + oRefD );
        t.nodes = reinterpret_cast<TravNode*>( a + oNodes ); t.leafTris = reinterpret_cast<LeafTri*>( a + oTris );
        t.leafFilt = reinterpret_cast<LeafFilt*>( a + oFilt ); dDepth = a + oDepth;
    }

    kClearSegBoxes<<<1, 32, 0, st>>>( dBBox, 1 ); ++nl;
    kFaceBoxes<<<nblk( L ), TB, 0, st>>>( t.points, t.tris, dValid, L, dLeaves, dBBox ); ++nl;
    {
        uint32_t hb[6];
        B2_CUDA( cudaMemcpyAsync( hb, dBBox, sizeof( hb ), cudaMemcpyDeviceToHost, st ) );
        B2_CUDA( cudaStreamSynchronize( st ) );
        for ( int c = 0; c < 6; ++c )
        {
            const uint32_t u = ( hb[c] & 0x80000000u ) ? ( hb[c] & 0x7fffffffu ) : ~hb[c];
            memcpy( &t.bbox[c], &u, 4 );
        }
    }
    kInitPerm<<<nblk( L ), TB, 0, st>>>( L, dKeys, dLeaves ); ++nl;
    const Seg root = { 0, L, 0, 0 };
    B2_CUDA( cudaMemcpyAsync( dSegA, &root, sizeof( Seg ), cudaMemcpyHostToDevice, st ) );

    // 2. breadth-first build: one segmented sort per level while the segments are large, then one block per subtree
    for ( int lev = 0; lev <= levels; ++lev )
    {
        const size_t nSegs = (size_t)1 << lev;
        if ( lev == level0 )
        {
            B2_CUDA( cudaFuncSetAttribute( kSubtrees, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof( SubShared ) ) );
            kSubtrees<<<(unsigned)nSegs, SUB_TB, sizeof( SubShared ), st>>>( dSegA, level0, levels, dKeys, dLeaves, t.refNodes, dSkip, dDepth, dLeafNodeOfPos ); ++nl;
            break;
        }
        if ( lev < levels )
        {
            kClearSegBoxes<<<nblk( nSegs * 6 ), TB, 0, st>>>( dSegBox, nSegs ); ++nl;
            kSegBoxes<<<nblk( L ), TB, 0, st>>>( L, dKeys, dLeaves, dSegBox ); ++nl;
            kMakeKeys<<<nblk( L ), TB, 0, st>>>( L, dKeys, dLeaves, dSegBox, dSegA ); ++nl;
            B2_CUDA( cub::DeviceMergeSort::SortKeys( sortTmp, sortTmpBytes, dKeys, L, LevelKeyLess(), st ) ); nl += 2;
        }
        kEmitLevel<<<nblk( nSegs ), TB, 0, st>>>( (uint32_t)nSegs, lev, dSegA, dSegB, dSegBox, dKeys, dLeaves, t.refNodes,
                                                  dSkip, dDepth, dLeafNodeOfPos ); ++nl;
        if ( lev < levels )
        {
            kDescend<<<nblk( L ), TB, 0, st>>>( L, dKeys, dSegA ); ++nl;
        }
        Seg* tmp = dSegA; dSegA = dSegB; dSegB = tmp;
    }

    // 3. leaf triangles in pre-order leaf order, dipoles bottom-up, packed traversal nodes
    kLeafSlots<<<nblk( L ), TB, 0, st>>>( L, dLeafNodeOfPos, dLeafSlotOfNode, dKeys, t.points, t.tris, t.leafTris, t.leafFilt ); ++nl;
    kLeafDipoles<<<nblk( N ), TB, 0, st>>>( N, t.refNodes, t.points, t.tris, t.refDipoles ); ++nl;
    for ( int lev = levels - 1; lev >= 0; --lev )
    {
        kInnerDipoles<<<nblk( N ), TB, 0, st>>>( N, lev, dDepth, t.refNodes, t.refDipoles ); ++nl;
    }
    kFinishDipoles<<<nblk( N ), TB, 0, st>>>( N, t.refNodes, t.refDipoles, dSkip, dLeafSlotOfNode, t.nodes ); ++nl;
    B2_CUDA( cudaGetLastError() );
    B2_CUDA( cudaStreamSynchronize( st ) );

    t.nodeDepth = dDepth; // part of t.arena, kept for refitTree; the temporaries are released by temps
#undef T_ALLOC
    if ( launches ) *launches += nl;
    return B2FWN_OK;
}

// The tree keeps its topology (which face sits in which leaf, the order of the nodes, the links); boxes are recomputed bottom
// up from the new coordinates, then dipoles, traversal nodes and leaf triangles exactly as after a build.  The result is what
// AABBTree::refit with all vertices changed + calcDipoles give on the reference's side, bit for bit.
int refitTree( DeviceTree& t, cudaStream_t st, uint64_t* launches )
{
    // pseudonormals depend on the coordinates: rebuilt on next use
    cudaFree( t.faceNormals ); cudaFree( t.vertNormals ); cudaFree( t.faceAdj );
    t.faceNormals = t.vertNormals = nullptr; t.faceAdj = nullptr;
    const size_t N = t.numNodes;
    const int L = (int)t.numLeaves;
    if ( N == 0 ) return B2FWN_OK;
    if ( !t.nodeDepth ) { setError( "tree has no depth table" ); return B2FWN_ERR_INVALID; }
    uint64_t nl = 0;
    uint32_t* dBBox = nullptr;
    B2_CUDA( cudaMalloc( &dBBox, 6 * sizeof( uint32_t ) ) );
    kClearSegBoxes<<<1, 32, 0, st>>>( dBBox, 1 ); ++nl;
    kRefitLeafNodes<<<nblk( N ), TB, 0, st>>>( N, t.refNodes, t.points, t.tris, dBBox ); ++nl;
    for ( int lev = t.depth - 1; lev >= 0; --lev )
    {
        kRefitInnerNodes<<<nblk( N ), TB, 0, st>>>( N, lev, t.nodeDepth, t.refNodes ); ++nl;
    }
    kLeafSlots<<<nblk( L ), TB, 0, st>>>( L, nullptr, nullptr, nullptr, t.points, t.tris, t.leafTris, t.leafFilt ); ++nl;
    kLeafDipoles<<<nblk( N ), TB, 0, st>>>( N, t.refNodes, t.points, t.tris, t.refDipoles ); ++nl;
    for ( int lev = t.depth - 1; lev >= 0; --lev )
    {
        kInnerDipoles<<<nblk( N ), TB, 0, st>>>( N, lev, t.nodeDepth, t.refNodes, t.refDipoles ); ++nl;
    }
    kFinishDipoles<<<nblk( N ), TB, 0, st>>>( N, t.refNodes, t.refDipoles, nullptr, nullptr, t.nodes ); ++nl;
    uint32_t hb[6];
    B2_CUDA( cudaMemcpyAsync( hb, dBBox, sizeof( hb ), cudaMemcpyDeviceToHost, st ) );
    B2_CUDA( cudaStreamSynchronize( st ) );
    cudaFree( dBBox );
    for ( int c = 0; c < 6; ++c )
    {
        const uint32_t u = ( hb[c] & 0x80000000u ) ? ( hb[c] & 0x7fffffffu ) : ~hb[c];
        memcpy( &t.bbox[c], &u, 4 );
    }
    B2_CUDA( cudaGetLastError() );
    if ( launches ) *launches += nl;
    return B2FWN_OK;
}

} // namespace b2
