// b2fwn_topology.cu -- half-edge tables of a triangle mesh, built on the device (first slice).
//
// What it replaces: Mesh::fromTriMesh -> MeshBuilder::fromTriangles (source/MRMesh/MRMeshBuilder.cpp:371-540), the step that turns
// the marching-cubes TriMesh into a MeshTopology at the end of mcOffsetMesh (source/MRVoxels/MROffset.cpp:222-225).  The
// reference adds the triangles one by one to a growing half-edge structure (FaceAdder::add, MRMeshBuilder.cpp:125-214), in parts
// per hardware thread that are stitched afterwards: its half-edge NUMBERING depends on the thread count and is no contract.  The
// contract is MeshTopology's: records {next, prev, org, left} per half-edge (MRMeshTopology.h:581-594), e ^ 1 = sym(e), next /
// prev = the counter-clockwise / clockwise neighbour in the ring of half-edges with the same origin, the edges of a face f in
// order are e, prev(sym(e)), ...; plus one half-edge per vertex and per face.  Tables that satisfy MeshTopology::checkValidity
// (MRMeshTopology.cpp:2537-2640) and whose getTriangulation() returns the input triples ARE that mesh.
//
// Build (all in HBM, no incremental insertion):
//   1. one key per face corner (f, k): the undirected edge (min, max) of (v_k, v_k+1); radix sort of the 3 F keys
//   2. runs of equal keys = undirected edges, numbered in sorted order (prefix sum over run heads): half-edge 2u = (min -> max),
//      2u + 1 = (max -> min).  A run longer than two, or two corners running the same way, is a non-manifold / inconsistently
//      oriented edge: the build reports it and the caller keeps the reference's CPU builder for such a mesh
//   3. per corner: org, left; next[e_k+1] = sym(e_k), prev[sym(e_k)] = e_k+1 (the face (v0, v1, v2) makes v0->v2 the ccw
//      neighbour of v0->v1)
//   4. boundary half-edges (no left face) are chained around their origin to the sym of the boundary half-edge that ends there
//   5. one thread per vertex walks its ring: a ring shorter than the vertex's degree is a non-manifold vertex (two fans)
#include "b2fwn_ctx.h"
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

namespace b2
{

namespace
{

constexpr int TB = 256;
inline unsigned nblk( size_t n ) { return (unsigned)( ( n + TB - 1 ) / TB ); }

enum TopoError : unsigned
{
    ERR_DEGENERATE = 0,   // face with a repeated or out-of-range vertex
    ERR_EDGE = 1,         // undirected edge with more than two faces, or two faces on the same side
    ERR_VERTEX = 2,       // vertex with more than one fan
    ERR_COUNT = 4
};

__global__ void kCornerEdgeKeys( const int32_t* __restrict__ tris, size_t nCorners, int nVerts, uint64_t* __restrict__ keys,
                                 uint32_t* __restrict__ vals, unsigned* __restrict__ err )
{
    const size_t c = blockIdx.x * (size_t)TB + threadIdx.x;
    if ( c >= nCorners ) return;
    const size_t f = c / 3;
    const int k = (int)( c - 3 * f );
    const int a = tris[3 * f + k], b = tris[3 * f + ( k + 1 ) % 3];
    if ( a < 0 || b < 0 || a >= nVerts || b >= nVerts || a == b ) atomicAdd( err + ERR_DEGENERATE, 1u );
    const uint32_t lo = (uint32_t)min( a, b ), hi = (uint32_t)max( a, b );
    keys[c] = ( (uint64_t)lo << 32 ) | hi;
    vals[c] = (uint32_t)c;
}

__global__ void kRunHeads( const uint64_t* __restrict__ keys, size_t n, uint32_t* __restrict__ head )
{
    const size_t i = blockIdx.x * (size_t)TB + threadIdx.x;
    if ( i >= n ) return;
    head[i] = ( i == 0 || keys[i] != keys[i - 1] ) ? 1u : 0u;
}

// after the inclusive scan of the heads, rank[i] - 1 = undirected edge of sorted corner i
__global__ void kHalfEdgesOfCorners( const int32_t* __restrict__ tris, const uint64_t* __restrict__ keys, const uint32_t* __restrict__ vals,
                                     const uint32_t* __restrict__ rank, size_t n, int32_t* __restrict__ halfOfCorner,
                                     b2fwn_half_edge* __restrict__ edges, int32_t* __restrict__ degree, unsigned* __restrict__ err )
{
    const size_t i = blockIdx.x * (size_t)TB + threadIdx.x;
    if ( i >= n ) return;
    const uint64_t key = keys[i];
    const bool first = i == 0 || keys[i - 1] != key;
    const bool last = i + 1 == n || keys[i + 1] != key;
    const uint32_t c = vals[i];
    const size_t f = c / 3;
    const int k = (int)( c - 3 * f );
    const int a = tris[3 * f + k], b = tris[3 * f + ( k + 1 ) % 3];
    const int dir = a > b ? 1 : 0;
    const int32_t he = (int32_t)( 2 * ( rank[i] - 1 ) + dir );
    if ( !first && !last ) atomicAdd( err + ERR_EDGE, 1u ); // three or more corners on one edge
    if ( !first && last )
    {
        // second corner of the run: it must run the other way, and the run must have exactly two corners
        const uint32_t c0 = vals[i - 1];
        const size_t f0 = c0 / 3;
        const int k0 = (int)( c0 - 3 * f0 );
        const int a0 = tris[3 * f0 + k0], b0 = tris[3 * f0 + ( k0 + 1 ) % 3];
        if ( ( a0 > b0 ) == ( a > b ) ) atomicAdd( err + ERR_EDGE, 1u );
    }
    halfOfCorner[c] = he;
    edges[he].org = a;
    edges[he].left = (int32_t)f;
    atomicAdd( degree + a, 1 );
    if ( first && last )
    {
        // boundary edge: the half-edge without a face
        edges[he ^ 1].org = b;
        edges[he ^ 1].left = -1;
        edges[he ^ 1].next = -1; // set by kBoundaryLinks
        atomicAdd( degree + b, 1 );
    }
}

__global__ void kFaceLinks( const int32_t* __restrict__ halfOfCorner, size_t nFaces, b2fwn_half_edge* __restrict__ edges,
                            int32_t* __restrict__ edgePerFace, int32_t* __restrict__ edgePerVertex )
{
    const size_t f = blockIdx.x * (size_t)TB + threadIdx.x;
    if ( f >= nFaces ) return;
    int32_t e[3];
#pragma unroll
    for ( int k = 0; k < 3; ++k ) e[k] = halfOfCorner[3 * f + k];
#pragma unroll
    for ( int k = 0; k < 3; ++k )
    {
        const int32_t ek = e[k], en = e[( k + 1 ) % 3];
        edges[en].next = ek ^ 1;
        edges[ek ^ 1].prev = en;
        atomicMin( edgePerVertex + edges[ek].org, ek ); // deterministic choice: the smallest half-edge id leaving the vertex
    }
    edgePerFace[f] = e[0]; // the ring of the face starts at its first corner: getTriVerts returns the input rotation
}

// boundary half-edge h (no left face) ends at vertex dest(h) = org(sym(h)); a manifold vertex has at most one such
__global__ void kBoundaryIn( const b2fwn_half_edge* __restrict__ edges, size_t nHalf, int32_t* __restrict__ inAt, unsigned* __restrict__ err )
{
    const size_t h = blockIdx.x * (size_t)TB + threadIdx.x;
    if ( h >= nHalf || edges[h].left >= 0 ) return;
    const int w = edges[h ^ 1].org;
    if ( atomicCAS( inAt + w, -1, (int32_t)h ) != -1 ) atomicAdd( err + ERR_VERTEX, 1u );
}

__global__ void kBoundaryLinks( b2fwn_half_edge* __restrict__ edges, size_t nHalf, const int32_t* __restrict__ inAt,
                                int32_t* __restrict__ edgePerVertex, unsigned* __restrict__ err )
{
    const size_t h = blockIdx.x * (size_t)TB + threadIdx.x;
    if ( h >= nHalf || edges[h].left >= 0 ) return;
    const int w = edges[h].org;
    const int32_t hin = inAt[w];
    if ( hin < 0 ) { atomicAdd( err + ERR_VERTEX, 1u ); return; }
    edges[h].next = hin ^ 1;
    edges[hin ^ 1].prev = (int32_t)h;
    atomicMin( edgePerVertex + w, (int32_t)h );
}

// every half-edge leaving v must be reachable from edgePerVertex[v] by next (checkValidity: fromSameOriginRing)
__global__ void kVertexRings( const b2fwn_half_edge* __restrict__ edges, const int32_t* __restrict__ degree, size_t nVerts,
                              int32_t* __restrict__ edgePerVertex, unsigned* __restrict__ err )
{
    const size_t v = blockIdx.x * (size_t)TB + threadIdx.x;
    if ( v >= nVerts ) return;
    const int d = degree[v];
    if ( d == 0 ) { edgePerVertex[v] = -1; return; } // a vertex no triangle uses: invalid, as in the reference
    const int32_t e0 = edgePerVertex[v];
    int32_t e = e0;
    int n = 0;
    do { e = edges[e].next; ++n; } while ( e != e0 && e >= 0 && n <= d );
    if ( e != e0 || n != d ) atomicAdd( err + ERR_VERTEX, 1u );
}

} // anonymous namespace

void freeTopology( TopoResult& r )
{
    cudaFree( r.dEdges ); cudaFree( r.dEdgePerVertex ); cudaFree( r.dEdgePerFace );
    r = TopoResult{};
}

int buildTopologyDevice( const int32_t* dTris, size_t nFaces, size_t nVerts, TopoResult& out, cudaStream_t st, uint64_t* launches )
{
    freeTopology( out );
    out.nVerts = nVerts; out.nFaces = nFaces;
    if ( nFaces == 0 ) return B2FWN_OK;
    if ( 3 * nFaces > 0x7FFFFFF0ull || nVerts > 0x7FFFFFF0ull ) { setError( "mesh too large for 32-bit half-edge ids" ); return B2FWN_ERR_INVALID; }
    const size_t n = 3 * nFaces;
    uint64_t *dKeys = nullptr, *dKeys2 = nullptr; uint32_t *dVals = nullptr, *dVals2 = nullptr, *dRank = nullptr;
    int32_t *dHalfOfCorner = nullptr, *dDegree = nullptr, *dInAt = nullptr; unsigned* dErr = nullptr; void* dTmp = nullptr;
    auto cleanup = [&]() { cudaFree( dKeys ); cudaFree( dKeys2 ); cudaFree( dVals ); cudaFree( dVals2 ); cudaFree( dRank ); cudaFree( dHalfOfCorner );
                           cudaFree( dDegree ); cudaFree( dInAt ); cudaFree( dErr ); cudaFree( dTmp ); };
#define T_TRY( expr ) do { cudaError_t _e = ( expr ); if ( _e != cudaSuccess ) { cleanup(); freeTopology( out ); return cudaFail( _e, #expr ); } } while ( 0 )
    T_TRY( cudaMalloc( &dKeys, n * 8 ) ); T_TRY( cudaMalloc( &dKeys2, n * 8 ) );
    T_TRY( cudaMalloc( &dVals, n * 4 ) ); T_TRY( cudaMalloc( &dVals2, n * 4 ) ); T_TRY( cudaMalloc( &dRank, n * 4 ) );
    T_TRY( cudaMalloc( &dHalfOfCorner, n * 4 ) );
    T_TRY( cudaMalloc( &dDegree, std::max<size_t>( nVerts, 1 ) * 4 ) ); T_TRY( cudaMalloc( &dInAt, std::max<size_t>( nVerts, 1 ) * 4 ) );
    T_TRY( cudaMalloc( &dErr, ERR_COUNT * sizeof( unsigned ) ) );
    T_TRY( cudaMemsetAsync( dErr, 0, ERR_COUNT * sizeof( unsigned ), st ) );
    T_TRY( cudaMemsetAsync( dDegree, 0, std::max<size_t>( nVerts, 1 ) * 4, st ) );
    T_TRY( cudaMemsetAsync( dInAt, 0xFF, std::max<size_t>( nVerts, 1 ) * 4, st ) );
    uint64_t nl = 0;
    kCornerEdgeKeys<<<nblk( n ), TB, 0, st>>>( dTris, n, (int)nVerts, dKeys, dVals, dErr ); ++nl;
    int bits = 1;
    while ( ( (size_t)1 << bits ) < nVerts ) ++bits;
    size_t tmpSort = 0, tmpScan = 0;
    T_TRY( cub::DeviceRadixSort::SortPairs( nullptr, tmpSort, dKeys, dKeys2, dVals, dVals2, (int)n, 0, 32 + bits, st ) );
    T_TRY( cub::DeviceScan::InclusiveSum( nullptr, tmpScan, dRank, dRank, (int)n, st ) );
    T_TRY( cudaMalloc( &dTmp, std::max( tmpSort, tmpScan ) ) );
    T_TRY( cub::DeviceRadixSort::SortPairs( dTmp, tmpSort, dKeys, dKeys2, dVals, dVals2, (int)n, 0, 32 + bits, st ) ); nl += 8;
    kRunHeads<<<nblk( n ), TB, 0, st>>>( dKeys2, n, dRank ); ++nl;
    T_TRY( cub::DeviceScan::InclusiveSum( dTmp, tmpScan, dRank, dRank, (int)n, st ) ); nl += 2;
    uint32_t nUndirected = 0; unsigned hErr[ERR_COUNT] = {};
    T_TRY( cudaMemcpyAsync( &nUndirected, dRank + ( n - 1 ), 4, cudaMemcpyDeviceToHost, st ) );
    T_TRY( cudaMemcpyAsync( hErr, dErr, sizeof( hErr ), cudaMemcpyDeviceToHost, st ) );
    T_TRY( cudaStreamSynchronize( st ) );
    if ( hErr[ERR_DEGENERATE] )
    {
        cleanup(); freeTopology( out );
        setError( "triangulation has degenerate faces or vertex ids out of range" );
        return B2FWN_ERR_INVALID;
    }
    out.nHalfEdges = 2 * (size_t)nUndirected;
    T_TRY( cudaMalloc( &out.dEdges, out.nHalfEdges * sizeof( b2fwn_half_edge ) ) );
    T_TRY( cudaMalloc( &out.dEdgePerVertex, std::max<size_t>( nVerts, 1 ) * 4 ) );
    T_TRY( cudaMalloc( &out.dEdgePerFace, nFaces * 4 ) );
    T_TRY( cudaMemsetAsync( out.dEdgePerVertex, 0x7F, std::max<size_t>( nVerts, 1 ) * 4, st ) ); // 0x7F7F7F7F: above every half-edge id
    kHalfEdgesOfCorners<<<nblk( n ), TB, 0, st>>>( dTris, dKeys2, dVals2, dRank, n, dHalfOfCorner, out.dEdges, dDegree, dErr ); ++nl;
    T_TRY( cudaMemcpyAsync( hErr, dErr, sizeof( hErr ), cudaMemcpyDeviceToHost, st ) );
    T_TRY( cudaStreamSynchronize( st ) );
    if ( !hErr[ERR_EDGE] ) // with a broken edge the corner -> half-edge map is not injective: nothing further is meaningful
    {
        kFaceLinks<<<nblk( nFaces ), TB, 0, st>>>( dHalfOfCorner, nFaces, out.dEdges, out.dEdgePerFace, out.dEdgePerVertex ); ++nl;
        kBoundaryIn<<<nblk( out.nHalfEdges ), TB, 0, st>>>( out.dEdges, out.nHalfEdges, dInAt, dErr ); ++nl;
        kBoundaryLinks<<<nblk( out.nHalfEdges ), TB, 0, st>>>( out.dEdges, out.nHalfEdges, dInAt, out.dEdgePerVertex, dErr ); ++nl;
        T_TRY( cudaMemcpyAsync( hErr, dErr, sizeof( hErr ), cudaMemcpyDeviceToHost, st ) );
        T_TRY( cudaStreamSynchronize( st ) );
        if ( !hErr[ERR_VERTEX] )
        {
            kVertexRings<<<nblk( nVerts ), TB, 0, st>>>( out.dEdges, dDegree, nVerts, out.dEdgePerVertex, dErr ); ++nl;
            T_TRY( cudaMemcpyAsync( hErr, dErr, sizeof( hErr ), cudaMemcpyDeviceToHost, st ) );
            T_TRY( cudaStreamSynchronize( st ) );
        }
    }
    T_TRY( cudaGetLastError() );
#undef T_TRY
    cleanup();
    if ( launches ) *launches += nl;
    if ( hErr[ERR_EDGE] || hErr[ERR_VERTEX] )
    {
        freeTopology( out );
        setError( hErr[ERR_EDGE] ? "triangulation is not an oriented manifold: an edge has more than two faces or two faces on the same side"
                                 : "triangulation is not a manifold: a vertex has more than one fan of faces" );
        return B2FWN_ERR_INVALID;
    }
    return B2FWN_OK;
}

} // namespace b2
