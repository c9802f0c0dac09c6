// b2fwn_mc.cu -- marching cubes on a device-resident float volume (hand-written for sm_100a).
//
// Replaces the reference's CPU-only VolumeMesher (source/MRVoxels/MRMarchingCubes.cpp):
//   pass 1 addPartBlock_ :548-662  (one vertex per sign-changing +X/+Y/+Z grid edge, per-block hash maps)
//   pass 2 finalize      :812-1102 (8-bit case -> <= 5 triangles via the published table, NaN-corner substitution)
//   ids    SeparationPointStorage::makeUniqueVids / getTriangulation (source/MRMesh/MRSeparationPoint.cpp:14-70)
// Output is IDENTICAL to the reference's TriMesh (vertex coordinates bit for bit, triangle triples, and the order of
// both): the reference numbers vertices by (voxel id, direction) and orders triangles by cell id regardless of its
// block / thread count, which is exactly what a prefix scan over voxel ids yields.
//
// HBM-bound streaming design (no hash maps):
//   kMcClassify  one thread per voxel, one warp per 32 consecutive voxel ids ("segment"): reads the 8 corner values,
//                writes a 1-byte code per voxel (bits 0-2: which of +X/+Y/+Z own a vertex, bits 3-5: triangles of the
//                cell) and the per-segment vertex / triangle totals (warp reduction)
//   cub::DeviceScan::ExclusiveSum over the per-segment totals (G/32 entries each)
//   kMcEmit      same mapping; warps whose 32 codes are all zero leave after one 32-B read; the others rank their
//                vertices / triangles with warp ballots (warp-level prefix scan) and write them compacted.
//                The vertex id of an edge owned by a neighbour voxel = segment prefix + popcount of the code bytes
//                before it in its segment (one 32-B read).
// Algorithmic bytes per voxel: 4 (value, read twice: 8) + 1 (code, written once, read once: 2) + 0.5 (segment totals)
// ~ 10.5 B/voxel + the compacted output.
#include "b2fwn_internal.h"
#include "b2fwn_math.cuh"
#include "b2fwn_mc_tables.h"
#include <cub/device/device_scan.cuh>

namespace b2
{

namespace
{

constexpr unsigned FULL = 0xFFFFFFFFu;
constexpr int MC_TB = 256;

__constant__ signed char cTriTable[256][B2_MC_ROW];
__constant__ unsigned char cTriCount[256];

// cEdgeIndicesMap (MRMarchingCubes.cpp:25-40): table edge -> (corner owning the vertex, direction 0=X 1=Y 2=Z);
// corner index bits: 1 = +x, 2 = +y, 4 = +z (cVoxelNeighbors :42-51); cMapNeighbors (:52-64): corner -> case bit
__constant__ unsigned char cEdgeCorner[12] = { 0, 1, 2, 0, 4, 5, 6, 4, 0, 1, 3, 2 };
__constant__ unsigned char cEdgeDir[12] = { 0, 1, 0, 1, 0, 1, 0, 1, 2, 2, 2, 2 };
__constant__ unsigned char cCornerBit[8] = { 0, 1, 3, 2, 4, 5, 7, 6 };
__constant__ unsigned char cNeiOrder[7] = { 1, 2, 4, 3, 5, 6, 7 }; // NaN substitution order (:951-959)

// The mesher works on a SLAB of the grid: the voxels of the layers [zo0, zo1) ("owners": each owns the vertices on its +X/+Y/+Z
// edges and the triangles of the cell whose lowest corner it is) get vertex / triangle ids baseV.. / baseT.. in voxel-id order.
// Their triangles also use vertices owned by voxels of layer zo1, so that layer is classified as well ("extra" layer: its codes
// give the ranks of its vertices, which continue the numbering: the next slab starts exactly there).  Values must be present for
// the layers [zv0, ..) up to min(dz, zo1 + 2).  The whole grid is the slab zo0 = 0, zo1 = dz.
// Thread / code index t: owners t in [0, nOwn) = voxel id minus zo0 * layer; the extra layer at [extraBase, extraBase + layer),
// extraBase = nOwn rounded up to the block size, so that a block never mixes the two.
struct McGrid
{
    int dx, dy, dz;
    size_t layer, total; // total = number of code slots (owners, padding, extra layer)
    float iso;
    float zero[3];  // origin + 0.5 * voxelSize (zeroPoint of a SimpleVolume, :566; MRVoxelsVolumeAccess.h:78)
    float vs[3];
    int lessInside;
    int zv0, zo0, zo1;
    size_t nOwn, extraBase;
    uint32_t baseV, baseT;
};

// code slot -> index into the slab's value array (layer zv0 first); slot must be an owner or an extra-layer voxel
__device__ __forceinline__ size_t valueIndex( const McGrid& g, size_t t )
{
    return t < g.nOwn ? t + (size_t)( g.zo0 - g.zv0 ) * g.layer : ( t - g.extraBase ) + (size_t)( g.zo1 - g.zv0 ) * g.layer;
}
// code slot of the voxel `off` voxel ids after owner slot t (off = +1, +dx, +layer and sums): layer zo1 lives at extraBase
__device__ __forceinline__ size_t slotAfter( const McGrid& g, size_t t, size_t off )
{
    const size_t r = t + off;
    return r < g.nOwn ? r : g.extraBase + ( r - g.nOwn );
}

__device__ __forceinline__ bool lowerThan( float v, float iso ) { return v < iso; }
__device__ __forceinline__ bool invalidVal( float v, float iso ) { return !( v < iso ) && !( v >= iso ); }
// a vertex sits on the edge v0 -> v1 iff one end is lower and the other same-or-higher (:613-640)
__device__ __forceinline__ bool hasSepPoint( float v0, float v1, float iso )
{
    if ( invalidVal( v0, iso ) ) return false;
    return v0 < iso ? ( v1 >= iso ) : ( v1 < iso );
}

struct Cell
{
    float v[8];    // corner values (corner index bits: 1=+x, 2=+y, 4=+z); only valid where inside the grid
    bool hasX, hasY, hasZ; // neighbour exists in +x / +y / +z
};

__device__ __forceinline__ void loadCell( const float* __restrict__ vol, const McGrid& g, size_t i, int x, int y, int z, Cell& c )
{
    c.hasX = x + 1 < g.dx; c.hasY = y + 1 < g.dy; c.hasZ = z + 1 < g.dz;
    const size_t sx = 1, sy = (size_t)g.dx, sz = g.layer;
    c.v[0] = __ldg( vol + i );
    c.v[1] = c.hasX ? __ldg( vol + i + sx ) : 0.f;
    c.v[2] = c.hasY ? __ldg( vol + i + sy ) : 0.f;
    c.v[4] = c.hasZ ? __ldg( vol + i + sz ) : 0.f;
    const bool full = c.hasX && c.hasY && c.hasZ;
    c.v[3] = full ? __ldg( vol + i + sx + sy ) : 0.f;
    c.v[5] = full ? __ldg( vol + i + sx + sz ) : 0.f;
    c.v[6] = full ? __ldg( vol + i + sy + sz ) : 0.f;
    c.v[7] = full ? __ldg( vol + i + sx + sy + sz ) : 0.f;
}

// which of the +X/+Y/+Z edges of the voxel own a vertex (pass 1)
__device__ __forceinline__ unsigned vertexMask( const Cell& c, float iso )
{
    unsigned m = 0;
    if ( c.hasX && hasSepPoint( c.v[0], c.v[1], iso ) ) m |= 1;
    if ( c.hasY && hasSepPoint( c.v[0], c.v[2], iso ) ) m |= 2;
    if ( c.hasZ && hasSepPoint( c.v[0], c.v[4], iso ) ) m |= 4;
    return m;
}

// case index of the cell after NaN substitution, or -1 when the cell yields no triangle (pass 2, :897-1045)
__device__ __forceinline__ int cellCase( const Cell& c, float iso )
{
    if ( !( c.hasX && c.hasY && c.hasZ ) ) return -1;
    unsigned low = 0, inv = 0;
#pragma unroll
    for ( int i = 0; i < 8; ++i )
    {
        if ( lowerThan( c.v[i], iso ) ) low |= 1u << i;
        if ( invalidVal( c.v[i], iso ) ) inv |= 1u << i;
    }
    if ( inv == 0xFFu ) return -1;
    unsigned cfg = 0;
    if ( inv == 0 )
    {
#pragma unroll
        for ( int i = 0; i < 8; ++i )
            if ( ( low >> i ) & 1 ) cfg |= 1u << cCornerBit[i];
        return ( cfg == 0 || cfg == 0xFF ) ? -1 : (int)cfg;
    }
    for ( int i = 0; i < 8; ++i )
    {
        unsigned l = ( low >> i ) & 1;
        if ( ( inv >> i ) & 1 )
            for ( int k = 0; k < 7; ++k )
            {
                const int j = i ^ cNeiOrder[k];
                if ( !( ( inv >> j ) & 1 ) ) { l = ( low >> j ) & 1; break; }
            }
        if ( l ) cfg |= 1u << cCornerBit[i];
    }
    if ( cfg == 0 || cfg == 0xFF ) return -1;
    // "ensure consistent nan voxel": every edge the plan uses must own a vertex
    const signed char* plan = cTriTable[cfg];
    for ( int k = 0; k < B2_MC_ROW && plan[k] >= 0; ++k )
    {
        const int c0 = cEdgeCorner[plan[k]], c1 = c0 | ( 1 << cEdgeDir[plan[k]] );
        if ( !hasSepPoint( c.v[c0], c.v[c1], iso ) ) return -1;
    }
    return (int)cfg;
}

// grid position of code slot t (owners: layers from zo0 on; extra layer: zo1)
__device__ __forceinline__ void voxelPos( const McGrid& g, size_t t, int& x, int& y, int& z )
{
    const bool extra = t >= g.extraBase;
    const size_t i = extra ? t - g.extraBase : t;
    const int lz = (int)( i / g.layer );
    const size_t r = i - (size_t)lz * g.layer;
    y = (int)( r / (size_t)g.dx );
    x = (int)( r - (size_t)y * g.dx );
    z = extra ? g.zo1 : g.zo0 + lz;
}

// (x, y, z) of the block's voxels without a 64-bit division per thread (ncu r02: kMcClassify was ALU-bound at 290 thread
// instructions per voxel, most of them the two divisions of voxelPos): thread 0 divides once for the block, the others step
__device__ __forceinline__ void blockVoxelPos( const McGrid& g, int* sPos, int& x, int& y, int& z )
{
    if ( threadIdx.x == 0 )
    {
        int bx, by, bz;
        voxelPos( g, blockIdx.x * (size_t)MC_TB, bx, by, bz );
        sPos[0] = bx; sPos[1] = by; sPos[2] = bz;
    }
    __syncthreads();
    x = sPos[0] + (int)threadIdx.x; y = sPos[1]; z = sPos[2];
    while ( x >= g.dx ) { x -= g.dx; ++y; }
    while ( y >= g.dy ) { y -= g.dy; ++z; }
}
// does code slot t hold a voxel (owner or extra layer)?
__device__ __forceinline__ bool slotInGrid( const McGrid& g, size_t t )
{
    return t < g.nOwn || ( t >= g.extraBase && t < g.total );
}

// The +x corners of a voxel are the own-x corners of the next voxel: inside a row the lane to the right has them already
__device__ __forceinline__ void loadCellShared( const float* __restrict__ vol, const McGrid& g, size_t i, int x, int y, int z, bool inGrid, Cell& c )
{
    const unsigned lane = threadIdx.x & 31;
    c.hasX = x + 1 < g.dx; c.hasY = y + 1 < g.dy; c.hasZ = z + 1 < g.dz;
    const size_t sy = (size_t)g.dx, sz = g.layer;
    const bool hy = inGrid && c.hasY, hz = inGrid && c.hasZ;
    c.v[0] = inGrid ? __ldg( vol + i ) : 0.f;
    c.v[2] = hy ? __ldg( vol + i + sy ) : 0.f;
    c.v[4] = hz ? __ldg( vol + i + sz ) : 0.f;
    c.v[6] = ( hy && hz ) ? __ldg( vol + i + sy + sz ) : 0.f;
    // lane + 1 holds voxel i + 1 = (x + 1, y, z) whenever x + 1 is still inside the row; its has-flags in y and z are ours
    const float n0 = __shfl_down_sync( FULL, c.v[0], 1 ), n2 = __shfl_down_sync( FULL, c.v[2], 1 );
    const float n4 = __shfl_down_sync( FULL, c.v[4], 1 ), n6 = __shfl_down_sync( FULL, c.v[6], 1 );
    const bool fromLane = lane < 31;
    const bool hx = inGrid && c.hasX;
    c.v[1] = !hx ? 0.f : fromLane ? n0 : __ldg( vol + i + 1 );
    c.v[3] = !( hx && c.hasY ) ? 0.f : fromLane ? n2 : __ldg( vol + i + 1 + sy );
    c.v[5] = !( hx && c.hasZ ) ? 0.f : fromLane ? n4 : __ldg( vol + i + 1 + sz );
    c.v[7] = !( hx && c.hasY && c.hasZ ) ? 0.f : fromLane ? n6 : __ldg( vol + i + 1 + sy + sz );
    // loadCell leaves v[3], v[5], v[6], v[7] at 0 unless the cell is complete; nothing reads them in that case
}

__global__ void __launch_bounds__( MC_TB ) kMcClassify( const float* __restrict__ vol, const McGrid g, uint8_t* __restrict__ code,
                                                        uint32_t* __restrict__ segV, uint32_t* __restrict__ segT )
{
    __shared__ int sPos[3];
    const size_t i = blockIdx.x * (size_t)MC_TB + threadIdx.x;
    unsigned vm = 0, nt = 0;
    int x, y, z;
    blockVoxelPos( g, sPos, x, y, z );
    const bool inGrid = slotInGrid( g, i );
    Cell c;
    loadCellShared( vol, g, inGrid ? valueIndex( g, i ) : 0, x, y, z, inGrid, c );
    if ( inGrid )
    {
        const bool full = c.hasX && c.hasY && c.hasZ;
        unsigned low = 0, inv = 0;
#pragma unroll
        for ( int k = 0; k < 8; ++k )
        {
            if ( lowerThan( c.v[k], g.iso ) ) low |= 1u << k;
            if ( invalidVal( c.v[k], g.iso ) ) inv |= 1u << k;
        }
        if ( full && inv == 0 )
        {
            // the common cell: eight valid corners.  Case bits = corner bits with 2 <-> 3 and 6 <-> 7 swapped (cCornerBit); an edge
            // owns a vertex iff its ends lie on different sides (hasSepPoint without NaN)
            const unsigned cfg = ( low & 0x33u ) | ( ( low & 0x44u ) << 1 ) | ( ( low & 0x88u ) >> 1 );
            nt = cTriCount[cfg]; // 0 for 0x00 and 0xFF
            const unsigned l0 = low & 1u;
            vm = ( ( ( low >> 1 ) & 1u ) ^ l0 ) | ( ( ( ( low >> 2 ) & 1u ) ^ l0 ) << 1 ) | ( ( ( ( low >> 4 ) & 1u ) ^ l0 ) << 2 );
        }
        else if ( !( full && inv == 0xFFu ) ) // a complete cell of NaNs yields nothing (cellCase :inv == 0xFF, vertexMask: v0 invalid)
        {
            vm = vertexMask( c, g.iso );
            const int cfg = cellCase( c, g.iso );
            nt = cfg < 0 ? 0 : cTriCount[cfg];
        }
    }
    if ( i < g.total ) code[i] = (uint8_t)( vm | ( nt << 3 ) ); // padding slots between the owners and the extra layer: 0
    const unsigned sv = __reduce_add_sync( FULL, __popc( vm ) );
    const unsigned stt = __reduce_add_sync( FULL, nt );
    if ( ( threadIdx.x & 31 ) == 0 && i < g.total )
    {
        segV[i >> 5] = sv;
        segT[i >> 5] = stt;
    }
}

// number of vertices owned by the voxels of `owner`'s segment that precede (owner, dir): popcount over the code bytes
__device__ __forceinline__ unsigned rankInSegment( const uint8_t* __restrict__ code, size_t owner, int dir, size_t total )
{
    const size_t seg0 = owner & ~(size_t)31;
    const unsigned pos = (unsigned)( owner & 31 );
    unsigned r = 0;
    if ( seg0 + 32 <= total )
    {
        const uint4* p = reinterpret_cast<const uint4*>( code + seg0 );
        const uint4 a = __ldg( p ), b = __ldg( p + 1 );
        const unsigned w[8] = { a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w };
#pragma unroll
        for ( int k = 0; k < 8; ++k )
        {
            unsigned m = w[k] & 0x07070707u;
            const unsigned first = 4u * k;
            if ( pos < first + 4 )
            {
                // keep the bytes before pos, and the directions before dir of byte pos
                const unsigned nb = pos > first ? pos - first : 0; // whole bytes kept (0..3)
                const unsigned keep = nb ? ( 0xFFFFFFFFu >> ( 32 - 8 * nb ) ) : 0u;
                const unsigned own = pos >= first ? ( ( ( 1u << dir ) - 1u ) << ( 8 * ( pos - first ) ) ) : 0u;
                m &= keep | own;
            }
            if ( first <= pos ) r += __popc( m );
        }
    }
    else
    {
        for ( size_t j = seg0; j < owner; ++j ) r += __popc( (unsigned)code[j] & 7u );
        r += __popc( (unsigned)code[owner] & ( ( 1u << dir ) - 1u ) );
    }
    return r;
}

__global__ void __launch_bounds__( MC_TB ) kMcEmit( const float* __restrict__ vol, const McGrid g, const uint8_t* __restrict__ code,
                                                    const uint32_t* __restrict__ segV, const uint32_t* __restrict__ segT,
                                                    float* __restrict__ outVerts, int32_t* __restrict__ outTris )
{
    __shared__ int sPos[3];
    const size_t i = blockIdx.x * (size_t)MC_TB + threadIdx.x;
    const unsigned lane = threadIdx.x & 31;
    const unsigned cd = i < g.nOwn ? code[i] : 0u; // owners only: the extra layer belongs to the next slab
    int x, y, z;
    blockVoxelPos( g, sPos, x, y, z );
    if ( __ballot_sync( FULL, cd != 0 ) == 0 ) return; // nothing on the surface in this segment
    const unsigned lt = ( 1u << lane ) - 1u;
    // warp-level exclusive prefix over 3-bit counts, bit plane by bit plane
    const unsigned vRank = __popc( __ballot_sync( FULL, cd & 1 ) & lt ) + __popc( __ballot_sync( FULL, cd & 2 ) & lt )
                           + __popc( __ballot_sync( FULL, cd & 4 ) & lt );
    const unsigned tRank = __popc( __ballot_sync( FULL, cd & 8 ) & lt ) + 2 * __popc( __ballot_sync( FULL, cd & 16 ) & lt )
                           + 4 * __popc( __ballot_sync( FULL, cd & 32 ) & lt );
    if ( cd == 0 ) return;
    Cell c;
    loadCell( vol, g, valueIndex( g, i ), x, y, z, c );
    const unsigned vm = cd & 7u;
    if ( vm )
    {
        size_t vid = (size_t)segV[i >> 5] + vRank; // slab-local: outVerts starts at the slab's first vertex
        float coords[3];
        coords[0] = FA( g.zero[0], FM( g.vs[0], (float)x ) );
        coords[1] = FA( g.zero[1], FM( g.vs[1], (float)y ) );
        coords[2] = FA( g.zero[2], FM( g.vs[2], (float)z ) );
#pragma unroll
        for ( int n = 0; n < 3; ++n )
        {
            if ( !( ( vm >> n ) & 1 ) ) continue;
            const float v0 = c.v[0], v1 = c.v[1 << n];
            const float ratio = __fdiv_rn( FS( g.iso, v0 ), FS( v1, v0 ) ); // positioner, :569-578
            const float om = FS( 1.0f, ratio );
#pragma unroll
            for ( int k = 0; k < 3; ++k )
            {
                const float nx = k == n ? FA( coords[k], g.vs[k] ) : coords[k];
                outVerts[3 * vid + k] = FA( FM( om, coords[k] ), FM( ratio, nx ) );
            }
            ++vid;
        }
    }
    const unsigned nt = cd >> 3;
    if ( nt )
    {
        const int cfg = cellCase( c, g.iso );
        const signed char* plan = cTriTable[cfg];
        size_t tid = (size_t)segT[i >> 5] + tRank;
        for ( unsigned t = 0; t < nt; ++t, ++tid )
        {
            int32_t vv[3];
#pragma unroll
            for ( int k = 0; k < 3; ++k )
            {
                const int e = plan[3 * t + k];
                const int cn = cEdgeCorner[e], dir = cEdgeDir[e];
                const size_t owner = slotAfter( g, i, ( cn & 1 ) + ( ( cn >> 1 ) & 1 ) * (size_t)g.dx + ( ( cn >> 2 ) & 1 ) * g.layer );
                vv[k] = (int32_t)( g.baseV + segV[owner >> 5] + rankInSegment( code, owner, dir, g.total ) );
            }
            outTris[3 * tid] = vv[0];
            outTris[3 * tid + 1] = g.lessInside ? vv[2] : vv[1]; // :1058-1069
            outTris[3 * tid + 2] = g.lessInside ? vv[1] : vv[2];
        }
    }
}

bool g_tablesReady[64] = {};

int uploadTables( int device )
{
    if ( device >= 0 && device < 64 && g_tablesReady[device] ) return B2FWN_OK;
    unsigned char cnt[256];
    for ( int c = 0; c < 256; ++c )
    {
        int n = 0;
        while ( n < B2_MC_ROW && B2_MC_TRI_TABLE[c][n] >= 0 ) ++n;
        cnt[c] = (unsigned char)( n / 3 );
    }
    B2_CUDA( cudaMemcpyToSymbol( cTriTable, B2_MC_TRI_TABLE, sizeof( B2_MC_TRI_TABLE ) ) );
    B2_CUDA( cudaMemcpyToSymbol( cTriCount, cnt, sizeof( cnt ) ) );
    if ( device >= 0 && device < 64 ) g_tablesReady[device] = true;
    return B2FWN_OK;
}

} // anonymous namespace

void freeMc( McResult& r )
{
    cudaFree( r.dVerts ); cudaFree( r.dTris );
    r = McResult{};
}

void freeMcWorkspace( McWorkspace& w )
{
    cudaFree( w.dCode ); cudaFree( w.dSegV ); cudaFree( w.dSegT ); cudaFree( w.dTmp );
    w = McWorkspace{};
}

namespace
{

// one slab between its two phases: count (classify + scans: how many vertices and triangles the owners yield) and emit
struct McPass
{
    McGrid g{};
    McWorkspace* w = nullptr;
    uint32_t nV = 0, nT = 0;
};

cudaError_t growBuffer( void** p, size_t& have, size_t need )
{
    if ( need <= have ) return cudaSuccess;
    cudaFree( *p );
    *p = nullptr; have = 0;
    const cudaError_t e = cudaMalloc( p, need );
    if ( e == cudaSuccess ) have = need;
    return e;
}

McGrid makeGrid( const int32_t dims[3], const b2fwn_mc_params& p, int zv0, int zo0, int zo1 )
{
    McGrid g{};
    g.dx = dims[0]; g.dy = dims[1]; g.dz = dims[2];
    g.layer = (size_t)dims[0] * dims[1];
    g.iso = p.iso;
    g.lessInside = p.less_inside;
    for ( int c = 0; c < 3; ++c )
    {
        g.vs[c] = p.voxel_size[c];
        volatile float half = 0.5f * p.voxel_size[c]; // separate rounding, as mult(shift, voxelSize) then +
        g.zero[c] = p.origin[c] + half;
    }
    g.zv0 = zv0; g.zo0 = zo0; g.zo1 = zo1;
    g.nOwn = (size_t)( zo1 - zo0 ) * g.layer;
    g.extraBase = ( g.nOwn + MC_TB - 1 ) / MC_TB * MC_TB;
    g.total = g.extraBase + ( zo1 < g.dz ? g.layer : 0 );
    return g;
}

// phase 1: dVol = values of the layers [zv0, min(dz, zo1 + 2)); fills ps.nV / ps.nT (synchronises the stream)
int mcCount( const float* dVol, McPass& ps, cudaStream_t st, uint64_t* launches )
{
    McGrid& g = ps.g;
    McWorkspace& w = *ps.w;
    const size_t S = ( g.total + 31 ) / 32;
    if ( S + 1 > 0x7FFFFFFFull ) { setError( "volume too large for one marching-cubes pass" ); return B2FWN_ERR_INVALID; }
#define MC_TRY( expr ) do { cudaError_t _e = ( expr ); if ( _e != cudaSuccess ) return cudaFail( _e, #expr ); } while ( 0 )
    MC_TRY( growBuffer( (void**)&w.dCode, w.codeBytes, ( g.total + 31 ) & ~(size_t)31 ) );
    if ( S + 1 > w.segEntries )
    {
        cudaFree( w.dSegV ); cudaFree( w.dSegT );
        w.dSegV = w.dSegT = nullptr; w.segEntries = 0;
        MC_TRY( cudaMalloc( &w.dSegV, ( S + 1 ) * sizeof( uint32_t ) ) );
        MC_TRY( cudaMalloc( &w.dSegT, ( S + 1 ) * sizeof( uint32_t ) ) );
        w.segEntries = S + 1;
    }
    MC_TRY( cudaMemsetAsync( w.dSegV + S, 0, sizeof( uint32_t ), st ) );
    MC_TRY( cudaMemsetAsync( w.dSegT + S, 0, sizeof( uint32_t ), st ) );
    const size_t blocks = ( g.total + MC_TB - 1 ) / MC_TB;
    if ( blocks > 0x7FFFFFFFull ) { setError( "volume too large for one launch" ); return B2FWN_ERR_INVALID; }
    kMcClassify<<<(unsigned)blocks, MC_TB, 0, st>>>( dVol, g, w.dCode, w.dSegV, w.dSegT );
    size_t tmpBytes = 0;
    MC_TRY( cub::DeviceScan::ExclusiveSum( nullptr, tmpBytes, w.dSegV, w.dSegV, (int)( S + 1 ), st ) );
    MC_TRY( growBuffer( &w.dTmp, w.tmpBytes, tmpBytes ) );
    MC_TRY( cub::DeviceScan::ExclusiveSum( w.dTmp, tmpBytes, w.dSegV, w.dSegV, (int)( S + 1 ), st ) );
    MC_TRY( cub::DeviceScan::ExclusiveSum( w.dTmp, tmpBytes, w.dSegT, w.dSegT, (int)( S + 1 ), st ) );
    // what the owners yield = the prefix at the first slot after them (extraBase is a multiple of 32)
    const size_t sOwn = g.extraBase / 32;
    MC_TRY( cudaMemcpyAsync( &ps.nV, w.dSegV + sOwn, sizeof( uint32_t ), cudaMemcpyDeviceToHost, st ) );
    MC_TRY( cudaMemcpyAsync( &ps.nT, w.dSegT + sOwn, sizeof( uint32_t ), cudaMemcpyDeviceToHost, st ) );
    MC_TRY( cudaStreamSynchronize( st ) );
    if ( launches ) *launches += 5; // classify + 2 x (cub scan = 2 kernels)
    return B2FWN_OK;
}

// phase 2: the owners' vertices and triangles to dVerts / dTris (this slab's first vertex / triangle first), ids from g.baseV on
int mcEmit( const float* dVol, McPass& ps, float* dVerts, int32_t* dTris, cudaStream_t st, uint64_t* launches )
{
    if ( ps.nV || ps.nT )
    {
        const size_t blocks = ( ps.g.nOwn + MC_TB - 1 ) / MC_TB;
        kMcEmit<<<(unsigned)blocks, MC_TB, 0, st>>>( dVol, ps.g, ps.w->dCode, ps.w->dSegV, ps.w->dSegT, dVerts, dTris );
        if ( launches ) ++*launches;
    }
    MC_TRY( cudaGetLastError() );
    return B2FWN_OK;
#undef MC_TRY
}

} // anonymous namespace

// growing output arrays of a mesh that arrives slab by slab
int mcReserve( McResult& out, size_t& capV, size_t& capT, size_t needV, size_t needT, cudaStream_t st )
{
    auto grow = [&]( void** p, size_t& cap, size_t used, size_t need, size_t elem ) -> cudaError_t
    {
        if ( need <= cap ) return cudaSuccess;
        const size_t nc = std::max( need, cap + cap / 2 );
        void* q = nullptr;
        cudaError_t e = cudaMalloc( &q, std::max<size_t>( nc, 1 ) * elem );
        if ( e != cudaSuccess ) return e;
        if ( used ) e = cudaMemcpyAsync( q, *p, used * elem, cudaMemcpyDeviceToDevice, st );
        if ( e == cudaSuccess ) e = cudaStreamSynchronize( st );
        cudaFree( *p );
        *p = q; cap = nc;
        return e;
    };
    B2_CUDA( grow( (void**)&out.dVerts, capV, out.nVerts, needV, 3 * sizeof( float ) ) );
    B2_CUDA( grow( (void**)&out.dTris, capT, out.nTris, needT, 3 * sizeof( int32_t ) ) );
    return B2FWN_OK;
}

int marchingCubesSlabs( const McVolumeSource& src, const int32_t dims[3], const b2fwn_mc_params& p, int layersPerSlab, int zBegin, int zEnd,
                        uint32_t baseV, uint32_t baseT, McResult& out, cudaStream_t st, uint64_t* launches, McWorkspace* work )
{
    McWorkspace local; // without a handle's workspace: scratch of this call only
    struct Guard { McWorkspace* w; ~Guard() { if ( w ) freeMcWorkspace( *w ); } } guard{ work ? nullptr : &local };
    if ( !work ) work = &local;
    freeMc( out );
    if ( dims[0] <= 0 || dims[1] <= 0 || dims[2] <= 0 ) return B2FWN_OK; // VolumeMesher::run :383-384
    if ( dims[2] <= 1 ) { setError( "a part must have at least two Z slices" ); return B2FWN_ERR_INVALID; } // :443-444
    int device = 0;
    B2_CUDA( cudaGetDevice( &device ) );
    if ( int e = uploadTables( device ) ) return e;
    if ( layersPerSlab <= 0 ) layersPerSlab = zEnd - zBegin;
    size_t capV = 0, capT = 0;
    for ( int z = zBegin; z < zEnd; z += layersPerSlab )
    {
        const int zo1 = std::min( zEnd, z + layersPerSlab );
        const int zv1 = std::min( dims[2], zo1 + 2 );
        const float* dVol = nullptr;
        if ( int e = src( z, zv1, &dVol ) ) return e; // values of the layers [z, zv1), pointer to layer z
        McPass ps;
        ps.w = work;
        ps.g = makeGrid( dims, p, z, z, zo1 );
        if ( int e = mcCount( dVol, ps, st, launches ) ) return e;
        if ( (uint64_t)baseV + out.nVerts + ps.nV > 0x7FFFFFFFull ) { setError( "more than 2^31 vertices" ); return B2FWN_ERR_INVALID; }
        ps.g.baseV = baseV + (uint32_t)out.nVerts;
        ps.g.baseT = baseT + (uint32_t)out.nTris;
        if ( int e = mcReserve( out, capV, capT, out.nVerts + ps.nV, out.nTris + ps.nT, st ) ) return e;
        const int rc = mcEmit( dVol, ps, out.dVerts + 3 * out.nVerts, out.dTris + 3 * out.nTris, st, launches );
        const cudaError_t se = cudaStreamSynchronize( st );
        if ( rc ) return rc;
        if ( se != cudaSuccess ) return cudaFail( se, "marching cubes slab" );
        out.nVerts += ps.nV; out.nTris += ps.nT;
    }
    return B2FWN_OK;
}

namespace
{
__global__ void __launch_bounds__( 256 ) kAddBase( int32_t* __restrict__ ids, size_t n, int32_t base )
{
    const size_t i = blockIdx.x * (size_t)256 + threadIdx.x;
    if ( i < n ) ids[i] += base;
}
} // namespace

int marchingCubesAddBase( McResult& r, int32_t base, cudaStream_t st, uint64_t* launches )
{
    const size_t n = 3 * (size_t)r.nTris;
    if ( n == 0 || base == 0 ) return B2FWN_OK;
    kAddBase<<<(unsigned)( ( n + 255 ) / 256 ), 256, 0, st>>>( r.dTris, n, base );
    if ( launches ) ++*launches;
    B2_CUDA( cudaGetLastError() );
    return B2FWN_OK;
}

int marchingCubesDevice( const float* dVolume, const int32_t dims[3], const b2fwn_mc_params& p, McResult& out,
                         cudaStream_t st, uint64_t* launches, McWorkspace* work, int layersPerSlab )
{
    const size_t layer = (size_t)std::max( dims[0], 0 ) * std::max( dims[1], 0 );
    return marchingCubesSlabs( [&]( int z0, int, const float** d ) { *d = dVolume + (size_t)z0 * layer; return B2FWN_OK; }, dims, p, layersPerSlab,
                               0, dims[2], 0, 0, out, st, launches, work );
}

} // namespace b2
