// SignDetectionMode::ProjectionNormal on the device (second half): signedDistanceToMesh
// (source/MRMesh/MRMeshDistance.cpp:13-59) = closest point of the mesh (kFindProjections, the reference's visiting order, so
// face and barycentrics are the reference's) + the sign of dot( proj - pt, pseudonormal at the closest point ).
// The pseudonormals need what the reference reads from its half-edge topology -- the face across each edge and the ring of
// faces around each vertex.  Both are built here from the triangulation with two radix sorts (no half-edges on the device):
//   faceAdj[3 f + k]   = the face on the other side of the edge (v_k, v_k+1) of face f, -1 at a boundary
//   faceNormals[f]     = cross( b - a, c - a ).normalized()
//   vertNormals[v]     = ( sum of angle * n.normalized() over the faces at v ).normalized(), summed in increasing face id
//                        (the reference sums in ring order, which only the half-edge structure knows: the last bits of a
//                        vertex normal can differ; signs differ only where the dot product is within rounding of zero)
// Compiled with -fmad=false: the arithmetic is b2fwn_normals_math.h, operation by operation as the reference's.
#include "b2fwn_internal.h"
#include "b2fwn_math.cuh"
#include "b2fwn_normals_math.h"

#include <cub/cub.cuh>

#include <algorithm>
#include <cfloat>

namespace b2
{
namespace
{
using namespace b2n;

__device__ __forceinline__ V3 ldPoint( const float* __restrict__ pts, int v ) { return v3( pts[3 * v], pts[3 * v + 1], pts[3 * v + 2] ); }

// one entry per face corner i = 3 f + k: the undirected edge (v_k, v_k+1) and the vertex v_k as sort keys
__global__ void __launch_bounds__( 256 ) kCornerKeys( const int32_t* __restrict__ tris, const size_t nCorners, uint64_t* __restrict__ edgeKey,
                                                      uint32_t* __restrict__ vertKey, uint32_t* __restrict__ idx )
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if ( i >= nCorners ) return;
    const size_t f = i / 3;
    const int k = (int)( i - 3 * f );
    const int a = tris[3 * f + k], b = tris[3 * f + ( k + 1 ) % 3], c = tris[3 * f + ( k + 2 ) % 3];
    const bool valid = a >= 0 && b >= 0 && c >= 0;
    edgeKey[i] = valid ? ( ( (uint64_t)(uint32_t)min( a, b ) << 32 ) | (uint32_t)max( a, b ) ) : ~0ull;
    vertKey[i] = valid ? (uint32_t)a : 0xFFFFFFFFu;
    idx[i] = (uint32_t)i;
}

// equal neighbours in the sorted edge list are the two sides of one edge
__global__ void __launch_bounds__( 256 ) kLinkEdges( const uint64_t* __restrict__ key, const uint32_t* __restrict__ corner, const size_t n,
                                                     int32_t* __restrict__ faceAdj )
{
    const size_t j = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if ( j >= n ) return;
    const uint64_t k = key[j];
    int32_t other = -1;
    if ( k != ~0ull )
    {
        if ( j > 0 && key[j - 1] == k ) other = (int32_t)( corner[j - 1] / 3 );
        else if ( j + 1 < n && key[j + 1] == k ) other = (int32_t)( corner[j + 1] / 3 );
    }
    faceAdj[corner[j]] = other;
}

__global__ void __launch_bounds__( 256 ) kFaceNormals( const float* __restrict__ pts, const int32_t* __restrict__ tris, const size_t nFaces,
                                                       float* __restrict__ faceNormals )
{
    const size_t f = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if ( f >= nFaces ) return;
    const int a = tris[3 * f], b = tris[3 * f + 1], c = tris[3 * f + 2];
    V3 n = v3( 0.f, 0.f, 0.f );
    if ( a >= 0 && b >= 0 && c >= 0 ) n = faceNormal( ldPoint( pts, a ), ldPoint( pts, b ), ldPoint( pts, c ) );
    faceNormals[3 * f] = n.x; faceNormals[3 * f + 1] = n.y; faceNormals[3 * f + 2] = n.z;
}

// one thread per vertex: its corners are one run of the corner list sorted by vertex (stable sort: increasing face id)
__global__ void __launch_bounds__( 256 ) kVertexNormals( const float* __restrict__ pts, const int32_t* __restrict__ tris,
                                                         const uint32_t* __restrict__ vertKey, const uint32_t* __restrict__ corner,
                                                         const size_t nCorners, const size_t nVerts, float* __restrict__ vertNormals )
{
    const size_t v = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if ( v >= nVerts ) return;
    size_t lo = 0, hi = nCorners; // first entry with key >= v
    while ( lo < hi )
    {
        const size_t mid = ( lo + hi ) / 2;
        if ( vertKey[mid] < (uint32_t)v ) lo = mid + 1; else hi = mid;
    }
    V3 sum = v3( 0.f, 0.f, 0.f );
    for ( size_t j = lo; j < nCorners && vertKey[j] == (uint32_t)v; ++j )
    {
        const size_t i = corner[j], f = i / 3;
        const int k = (int)( i - 3 * f );
        const V3 a = ldPoint( pts, tris[3 * f + k] ), b = ldPoint( pts, tris[3 * f + ( k + 1 ) % 3] ), c = ldPoint( pts, tris[3 * f + ( k + 2 ) % 3] );
        sum = add( sum, cornerTerm( a, b, c ) );
    }
    const V3 n = normalized( sum );
    vertNormals[3 * v] = n.x; vertNormals[3 * v + 1] = n.y; vertNormals[3 * v + 2] = n.z;
}

// signedDistanceToMesh after findProjection (MRMeshDistance.cpp:26-38): NaN without a projection or outside [min, max) when
// asked for, else sqrt( distSq ) with the sign of isOutsideByProjNorm
__global__ void __launch_bounds__( 256 ) kSignByProjNormal( const b2fwn_projection* __restrict__ proj, const float* __restrict__ pts, const size_t n,
                                                            const int32_t* __restrict__ tris, const float* __restrict__ faceNormals,
                                                            const float* __restrict__ vertNormals, const int32_t* __restrict__ faceAdj,
                                                            const float minDistSq, const float maxDistSq, const int nullOutside,
                                                            float* __restrict__ out )
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if ( i >= n ) return;
    const float4 r0 = reinterpret_cast<const float4*>( proj )[2 * i], r1 = reinterpret_cast<const float4*>( proj )[2 * i + 1];
    const int f = __float_as_int( r0.x );
    const float distSq = r1.z;
    const float nan = __int_as_float( 0x7FC00000 );
    if ( f < 0 || ( nullOutside && ( distSq < minDistSq || distSq >= maxDistSq ) ) ) { out[i] = nan; return; }
    const float dist = sqrtf( distSq );
    const TriSite s = triSite( r1.x, r1.y );
    V3 nrm;
    if ( s.vertex >= 0 ) nrm = ldPoint( vertNormals, tris[3 * (size_t)f + s.vertex] );
    else
    {
        nrm = ldPoint( faceNormals, f );
        if ( s.edge >= 0 )
        {
            const int g = faceAdj[3 * (size_t)f + s.edge];
            nrm = edgeNormal( nrm, g >= 0 ? ldPoint( faceNormals, g ) : nrm, g >= 0 );
        }
    }
    const V3 pt = v3( pts[3 * i], pts[3 * i + 1], pts[3 * i + 2] );
    out[i] = outsideByNormal( v3( r0.y, r0.z, r0.w ), pt, nrm ) ? dist : -dist;
}

// The volume form on top of the brick walk (kGridDist<PROJ>): vol holds |d| (NaN = outside the limits), slot the leaf slot of
// the closest triangle.  One more closest-point evaluation on that triangle (closestPointInTriangle in double, as the walk did)
// gives the point and its barycentrics; then the same sign rule as kSignByProjNormal.  Where two triangles are equally close
// (closest point on their common edge or vertex) the walk may have kept another one than the reference's search: the
// pseudonormal of an edge or a vertex does not depend on which of its faces reported it.
__global__ void __launch_bounds__( 256 ) kSignBySlot( float* __restrict__ vol, const int32_t* __restrict__ slot, const LeafTri* __restrict__ leafTris,
                                                      const int32_t* __restrict__ tris, const float* __restrict__ faceNormals,
                                                      const float* __restrict__ vertNormals, const int32_t* __restrict__ faceAdj,
                                                      const float ox, const float oy, const float oz, const float vx, const float vy, const float vz,
                                                      const int dx, const int dy, const int z0, const size_t n )
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if ( i >= n ) return;
    const float d = vol[i];
    if ( d != d ) return;
    const int s = slot[i];
    if ( s < 0 ) { vol[i] = __int_as_float( 0x7FC00000 ); return; } // no projection found (MRMeshDistance.cpp:27-28)
    const size_t layer = (size_t)dx * dy;
    const int z = z0 + (int)( i / layer );
    const size_t r = i % layer;
    const int y = (int)( r / dx ), x = (int)( r % dx );
    const float3 p = make_float3( voxelCentre( ox, vx, x ), voxelCentre( oy, vy, y ), voxelCentre( oz, vz, z ) );
    const float4* tp = reinterpret_cast<const float4*>( leafTris + s );
    const float4 ta = __ldg( tp ), tb = __ldg( tp + 1 ), tc = __ldg( tp + 2 );
    const int f = __float_as_int( ta.w );
    double rx, ry, rz, va, vb;
    closestPointBary( p, make_float3( ta.x, ta.y, ta.z ), make_float3( tb.x, tb.y, tb.z ), make_float3( tc.x, tc.y, tc.z ), rx, ry, rz, va, vb );
    const TriSite site = triSite( __double2float_rn( va ), __double2float_rn( vb ) );
    V3 nrm;
    if ( site.vertex >= 0 ) nrm = ldPoint( vertNormals, tris[3 * (size_t)f + site.vertex] );
    else
    {
        nrm = ldPoint( faceNormals, f );
        if ( site.edge >= 0 )
        {
            const int g = faceAdj[3 * (size_t)f + site.edge];
            nrm = edgeNormal( nrm, g >= 0 ? ldPoint( faceNormals, g ) : nrm, g >= 0 );
        }
    }
    const V3 proj = v3( __double2float_rn( rx ), __double2float_rn( ry ), __double2float_rn( rz ) );
    vol[i] = outsideByNormal( proj, v3( p.x, p.y, p.z ), nrm ) ? d : -d;
}

__global__ void __launch_bounds__( 256 ) kVoxelCentres( const float ox, const float oy, const float oz, const float vx, const float vy, const float vz,
                                                        const int dx, const int dy, const int z0, const size_t n, float* __restrict__ out )
{
    const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if ( i >= n ) return;
    const size_t layer = (size_t)dx * dy;
    const int z = z0 + (int)( i / layer );
    const size_t r = i % layer;
    const int y = (int)( r / dx ), x = (int)( r % dx );
    out[3 * i] = voxelCentre( ox, vx, x ); out[3 * i + 1] = voxelCentre( oy, vy, y ); out[3 * i + 2] = voxelCentre( oz, vz, z );
}

inline unsigned nblk( size_t n ) { return (unsigned)( ( n + 255 ) / 256 ); }

} // anonymous namespace

int ensurePseudoNormals( DeviceTree& t, cudaStream_t st, uint64_t* launches )
{
    if ( t.faceAdj || t.numFaces == 0 ) return B2FWN_OK;
    const size_t F = t.numFaces, V = t.numVerts, n = 3 * F;
    if ( n > 0x7FFFFFFFull ) { setError( "too many faces for the pseudonormal tables" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaMalloc( &t.faceNormals, F * 3 * sizeof( float ) ) );
    B2_CUDA( cudaMalloc( &t.vertNormals, std::max<size_t>( V, 1 ) * 3 * sizeof( float ) ) );
    B2_CUDA( cudaMalloc( &t.faceAdj, n * sizeof( int32_t ) ) );
    size_t tmpE = 0, tmpV = 0;
    cub::DeviceRadixSort::SortPairs( nullptr, tmpE, (const uint64_t*)nullptr, (uint64_t*)nullptr, (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)n, 0, 64, st );
    cub::DeviceRadixSort::SortPairs( nullptr, tmpV, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const uint32_t*)nullptr, (uint32_t*)nullptr, (int)n, 0, 32, st );
    const size_t tmpBytes = ( std::max( tmpE, tmpV ) + 15 ) & ~(size_t)15;
    // scratch: edge keys in/out (8 B), vertex keys in/out (4 B), corner ids in/out/out (4 B), sort temp
    uint8_t* scratch = nullptr;
    const size_t nPad = ( n + 3 ) & ~(size_t)3;
    B2_CUDA( cudaMalloc( &scratch, nPad * ( 2 * 8 + 2 * 4 + 3 * 4 ) + tmpBytes ) );
    uint64_t* eIn = reinterpret_cast<uint64_t*>( scratch );
    uint64_t* eOut = eIn + nPad;
    uint32_t* vIn = reinterpret_cast<uint32_t*>( eOut + nPad );
    uint32_t* vOut = vIn + nPad;
    uint32_t* cIn = vOut + nPad;
    uint32_t* cOutE = cIn + nPad;
    uint32_t* cOutV = cOutE + nPad;
    void* tmp = cOutV + nPad;
    kCornerKeys<<<nblk( n ), 256, 0, st>>>( t.tris, n, eIn, vIn, cIn );
    size_t tb = tmpBytes;
    B2_CUDA( cub::DeviceRadixSort::SortPairs( tmp, tb, eIn, eOut, cIn, cOutE, (int)n, 0, 64, st ) );
    kLinkEdges<<<nblk( n ), 256, 0, st>>>( eOut, cOutE, n, t.faceAdj );
    tb = tmpBytes;
    B2_CUDA( cub::DeviceRadixSort::SortPairs( tmp, tb, vIn, vOut, cIn, cOutV, (int)n, 0, 32, st ) );
    kFaceNormals<<<nblk( F ), 256, 0, st>>>( t.points, t.tris, F, t.faceNormals );
    if ( V > 0 ) kVertexNormals<<<nblk( V ), 256, 0, st>>>( t.points, t.tris, vOut, cOutV, n, V, t.vertNormals );
    if ( launches ) *launches += 7;
    B2_CUDA( cudaGetLastError() );
    B2_CUDA( cudaStreamSynchronize( st ) );
    cudaFree( scratch );
    return B2FWN_OK;
}

int launchSignedDistProjNormal( DeviceTree& t, const float* dPts, size_t n, float minDistSq, float maxDistSq, int nullOutside,
                                b2fwn_projection* dProj, float* dOut, cudaStream_t st, uint64_t* launches )
{
    if ( n == 0 ) return B2FWN_OK;
    if ( int e = ensurePseudoNormals( t, st, launches ) ) return e;
    // the sign comes from the normal AT the closest point, so without the NaN rule the search may not stop early or far
    // (MRMeshDistance.cpp:19-24)
    const float lo = nullOutside ? minDistSq : 0.f, up = nullOutside ? maxDistSq : FLT_MAX;
    if ( int e = launchFindProjections( t, dPts, n, nullptr, nullptr, up, lo, dProj, st, launches ) ) return e;
    kSignByProjNormal<<<nblk( n ), 256, 0, st>>>( dProj, dPts, n, t.tris, t.faceNormals, t.vertNormals, t.faceAdj, minDistSq, maxDistSq,
                                                  nullOutside, dOut );
    if ( launches ) ++*launches;
    B2_CUDA( cudaGetLastError() );
    return B2FWN_OK;
}

int launchSignBySlot( DeviceTree& t, const int32_t dims[3], const float origin[3], const float voxelSize[3], int z0, int z1, float* dVol,
                      const int32_t* dSlot, cudaStream_t st, uint64_t* launches )
{
    const size_t n = (size_t)dims[0] * dims[1] * ( z1 - z0 );
    if ( n == 0 ) return B2FWN_OK;
    if ( int e = ensurePseudoNormals( t, st, launches ) ) return e; // no tables for a mesh without faces: every slot is -1 then
    kSignBySlot<<<nblk( n ), 256, 0, st>>>( dVol, dSlot, t.leafTris, t.tris, t.faceNormals, t.vertNormals, t.faceAdj, origin[0], origin[1], origin[2],
                                           voxelSize[0], voxelSize[1], voxelSize[2], dims[0], dims[1], z0, n );
    if ( launches ) ++*launches;
    B2_CUDA( cudaGetLastError() );
    return B2FWN_OK;
}

int launchVoxelCentres( const float origin[3], const float voxelSize[3], int dx, int dy, int z0, int z1, float* dPts, cudaStream_t st,
                        uint64_t* launches )
{
    const size_t n = (size_t)dx * dy * ( z1 - z0 );
    if ( n == 0 ) return B2FWN_OK;
    kVoxelCentres<<<nblk( n ), 256, 0, st>>>( origin[0], origin[1], origin[2], voxelSize[0], voxelSize[1], voxelSize[2], dx, dy, z0, n, dPts );
    if ( launches ) ++*launches;
    B2_CUDA( cudaGetLastError() );
    return B2FWN_OK;
}

} // namespace b2
