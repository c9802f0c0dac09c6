// oracle/ref_wrap.cpp -- TEST INFRASTRUCTURE ONLY.
//
// A thin extern "C" wrapper (our code) around the UNMODIFIED reference classes compiled from
// /root/reference/source/MRMesh (see oracle/Makefile, target `ref`).  It lets tests/ and bench.py's
// cpu_baseline / --impl reference leg drive the reference's own CPU (TBB) implementation of the path:
//   MR::FastWindingNumber::{calcFromVector,calcSelfIntersections,calcFromGrid,calcFromGridWithDistances}
//     (source/MRMesh/MRFastWindingNumber.cpp:28-94)
//   MR::AABBTree / MR::calcDipoles (source/MRMesh/MRAABBTree.cpp:31-60, MRDipole.cpp:23-71)
//   MR::marchingCubesAsTriMesh(SimpleVolume) (source/MRVoxels/MRMarchingCubes.cpp:1112-1115 -> VolumeMesher::run)
//   MR::makeTorus / MR::makeUVSphere (source/MRMesh/MRTorus.cpp:9-59, MRMakeSphereMesh.cpp:35-102)
// Nothing under meshlib_b200/ may link or load this file's output.
#include "MRMesh/MRMesh.h"
#include "MRMesh/MRAABBTree.h"
#include "MRMesh/MRDipole.h"
#include "MRMesh/MRFastWindingNumber.h"
#include "MRMesh/MRDistanceToMeshOptions.h"
#include "MRMesh/MRMeshProject.h"
#include "MRMesh/MRPointsToMeshProjector.h"
#include "MRMesh/MRTorus.h"
#include "MRMesh/MRMakeSphereMesh.h"
#include "MRMesh/MRBitSet.h"
#include "MRMesh/MRTimer.h"
#include "MRMesh/MRTriMesh.h"
#include "MRVoxels/MRMarchingCubes.h"
#include "MRVoxels/MRVoxelsVolume.h"
#include "MRMesh/MRMeshBuilder.h"
#include "MRMesh/MRMeshTopology.h"
#include "MRMesh/MRRegionBoundary.h"
#include "MRMesh/MRIOParsing.h"
#include <sstream>
#include <tbb/blocked_range.h>
#include <tbb/parallel_for.h>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <limits>
#include <thread>
#include <chrono>

// MRTimer.cpp needs the (absent) spdlog submodule; the path only uses the RAII start/finish pair.
namespace MR
{
void Timer::restart( std::string ) {}
void Timer::start( std::string ) {}
void Timer::start( const char* ) {}
void Timer::finish() {}
}

// MRIOParsing.cpp pulls in the text parsers (fast_float, boost); MeshTopology::read only needs this one function of it
// (source/MRMesh/MRIOParsing.cpp:103-110, restated)
namespace MR
{
std::streamoff getStreamSize( std::istream& in )
{
    const auto posStart = in.tellg();
    in.seekg( 0, std::ios::end );
    const auto posEnd = in.tellg();
    in.seekg( posStart );
    return posEnd - posStart;
}
}

using namespace MR;

struct RefMesh
{
    Mesh mesh;
    Triangulation tris; // reference vertex order per face (getTriangulation), -1 for invalid faces
};

static AffineXf3f toXf( const float* xf12 )
{
    AffineXf3f xf;
    xf.A.x = Vector3f( xf12[0], xf12[1], xf12[2] );
    xf.A.y = Vector3f( xf12[3], xf12[4], xf12[5] );
    xf.A.z = Vector3f( xf12[6], xf12[7], xf12[8] );
    xf.b = Vector3f( xf12[9], xf12[10], xf12[11] );
    return xf;
}

extern "C"
{

int mrref_hardware_threads() { return (int)std::thread::hardware_concurrency(); }

void* mrref_create( const float* pts, size_t nVerts, const int* tris, size_t nFaces )
{
    VertCoords points( nVerts );
    std::memcpy( (void*)points.data(), pts, nVerts * sizeof( Vector3f ) );
    Triangulation t( nFaces );
    for ( size_t i = 0; i < nFaces; ++i )
        t.vec_[i] = { VertId( tris[3 * i] ), VertId( tris[3 * i + 1] ), VertId( tris[3 * i + 2] ) };
    auto* r = new RefMesh;
    r->mesh = Mesh::fromTriangles( std::move( points ), t );
    r->tris = r->mesh.topology.getTriangulation();
    return r;
}

void* mrref_make_torus( float R, float r, int pRes, int sRes )
{
    auto* m = new RefMesh;
    m->mesh = makeTorus( R, r, pRes, sRes );
    m->tris = m->mesh.topology.getTriangulation();
    return m;
}

void* mrref_make_uv_sphere( float radius, int hRes, int vRes )
{
    auto* m = new RefMesh;
    m->mesh = makeUVSphere( radius, hRes, vRes );
    m->tris = m->mesh.topology.getTriangulation();
    return m;
}

void mrref_destroy( void* h ) { delete (RefMesh*)h; }

size_t mrref_num_verts( void* h ) { return ( (RefMesh*)h )->mesh.points.size(); }
size_t mrref_num_faces( void* h ) { return ( (RefMesh*)h )->tris.size(); }
int mrref_num_valid_faces( void* h ) { return ( (RefMesh*)h )->mesh.topology.numValidFaces(); }

void mrref_get_points( void* h, float* out )
{
    auto& m = ( (RefMesh*)h )->mesh;
    std::memcpy( out, m.points.data(), m.points.size() * sizeof( Vector3f ) );
}

void mrref_get_tris( void* h, int* out )
{
    auto& t = ( (RefMesh*)h )->tris;
    for ( size_t i = 0; i < t.size(); ++i )
        for ( int k = 0; k < 3; ++k )
            out[3 * i + k] = (int)t.vec_[i][k];
}

/// number of AABB tree nodes (builds the tree on first use, as Mesh::getAABBTree does)
size_t mrref_num_nodes( void* h ) { return ( (RefMesh*)h )->mesh.getAABBTree().nodes().size(); }

/// nodes: 8 x 4 bytes each = box.min xyz, box.max xyz (float), l, r (int32)
void mrref_get_nodes( void* h, void* out )
{
    const auto& nodes = ( (RefMesh*)h )->mesh.getAABBTree().nodes();
    static_assert( sizeof( nodes.vec_[0] ) == 32 );
    std::memcpy( out, nodes.vec_.data(), nodes.size() * 32 );
}

/// moved vertices, same topology: Mesh::updateCaches (source/MRMesh/MRMesh.cpp:565-578) = AABBTree::refit on the cached tree
/// (MRAABBTree.cpp:62-94) + dipoles recomputed on next use; every later call of this wrapper works on the refitted tree
int mrref_update_points( void* h, const float* pts, size_t nVerts )
{
    auto& m = ( (RefMesh*)h )->mesh;
    if ( nVerts != m.points.size() ) return 1;
    (void)m.getAABBTree(); // the tree of the old coordinates must exist to be refitted
    for ( size_t i = 0; i < nVerts; ++i ) m.points.vec_[i] = Vector3f( pts[3 * i], pts[3 * i + 1], pts[3 * i + 2] );
    const VertBitSet changed = m.topology.getValidVerts();
    m.updateCaches( changed );
    return 0;
}

/// dipoles: 8 floats each = pos xyz, area, dirArea xyz, rr
void mrref_get_dipoles( void* h, float* out )
{
    const auto& d = ( (RefMesh*)h )->mesh.getDipoles();
    std::memcpy( out, d.vec_.data(), d.size() * sizeof( Dipole ) );
}

int mrref_calc_from_vector( void* h, const float* pts, size_t n, float beta, int skipFace, float* out )
{
    FastWindingNumber fwn( ( (RefMesh*)h )->mesh );
    std::vector<Vector3f> points( n );
    std::memcpy( (void*)points.data(), pts, n * sizeof( Vector3f ) );
    std::vector<float> res;
    auto e = fwn.calcFromVector( res, points, beta, FaceId( skipFace ), {} );
    if ( !e )
        return 1;
    std::memcpy( out, res.data(), n * sizeof( float ) );
    return 0;
}

/// out_bits: ceil(faceSize/64) 64-bit words
int mrref_calc_self_intersections( void* h, float beta, uint64_t* out_bits )
{
    auto& m = ( (RefMesh*)h )->mesh;
    FastWindingNumber fwn( m );
    FaceBitSet res;
    auto e = fwn.calcSelfIntersections( res, beta, {} );
    if ( !e )
        return 1;
    const size_t nf = m.topology.faceSize();
    const size_t words = ( nf + 63 ) / 64;
    for ( size_t w = 0; w < words; ++w )
        out_bits[w] = 0;
    for ( auto f : res )
        out_bits[size_t( f ) / 64] |= uint64_t( 1 ) << ( size_t( f ) % 64 );
    return 0;
}

int mrref_calc_from_grid( void* h, const int* dims, const float* xf12, float beta, float* out )
{
    FastWindingNumber fwn( ( (RefMesh*)h )->mesh );
    std::vector<float> res;
    auto e = fwn.calcFromGrid( res, Vector3i( dims[0], dims[1], dims[2] ), toXf( xf12 ), beta, {} );
    if ( !e )
        return 1;
    std::memcpy( out, res.data(), res.size() * sizeof( float ) );
    return 0;
}

int mrref_calc_from_grid_with_distances( void* h, const int* dims, const float* xf12,
    float minDistSq, float maxDistSq, int nullOutsideMinMax, float threshold, float beta, float* out )
{
    FastWindingNumber fwn( ( (RefMesh*)h )->mesh );
    DistanceToMeshOptions o;
    o.minDistSq = minDistSq;
    o.maxDistSq = maxDistSq;
    o.nullOutsideMinMax = nullOutsideMinMax != 0;
    o.windingNumberThreshold = threshold;
    o.windingNumberBeta = beta;
    std::vector<float> res;
    auto e = fwn.calcFromGridWithDistances( res, Vector3i( dims[0], dims[1], dims[2] ), toXf( xf12 ), o, {} );
    if ( !e )
        return 1;
    std::memcpy( out, res.data(), res.size() * sizeof( float ) );
    return 0;
}

/// same per-voxel function on an arbitrary sub-range of z layers [z0, z1) -- used to time bounded samples
int mrref_calc_grid_with_distances_layers( void* h, const int* dims, const float* xf12,
    float minDistSq, float maxDistSq, int nullOutsideMinMax, float threshold, float beta, int z0, int z1, float* out )
{
    // a z-range of the full grid is itself a grid whose transform has b shifted by A*(0,0,z0)
    AffineXf3f xf = toXf( xf12 );
    xf.b = xf( Vector3f( 0, 0, float( z0 ) ) );
    float x2[12] = { xf.A.x.x, xf.A.x.y, xf.A.x.z, xf.A.y.x, xf.A.y.y, xf.A.y.z, xf.A.z.x, xf.A.z.y, xf.A.z.z, xf.b.x, xf.b.y, xf.b.z };
    int d2[3] = { dims[0], dims[1], z1 - z0 };
    return mrref_calc_from_grid_with_distances( h, d2, x2, minDistSq, maxDistSq, nullOutsideMinMax, threshold, beta, out );
}

/// unsigned squared distances (findProjection) for a vector of points
int mrref_find_projection_dist_sq( void* h, const float* pts, size_t n, float upDistLimitSq, float loDistLimitSq, float* out )
{
    auto& m = ( (RefMesh*)h )->mesh;
    (void)m.getAABBTree();
    for ( size_t i = 0; i < n; ++i )
        out[i] = findProjection( Vector3f( pts[3 * i], pts[3 * i + 1], pts[3 * i + 2] ), m, upDistLimitSq, nullptr, loDistLimitSq ).distSq;
    return 0;
}

/// marching cubes of a dense float volume by the reference's VolumeMesher (two calls: run -> counts, then fetch)
static TriMesh g_lastMc;
// IPointsToMeshProjector::findProjections of the reference's CPU class (MRPointsToMeshProjector.cpp:17-33):
// out = n records of 8 x 4 bytes { int face; float point[3]; float baryA, baryB; float distSq; int edge }.
// pointXf12 / treeXf12: row-major A then b, or null
int mrref_find_projections( void* h, const float* pts, size_t n, const float* pointXf12, const float* treeXf12,
                            float upDistLimitSq, float loDistLimitSq, void* out )
{
    auto& m = ( (RefMesh*)h )->mesh;
    (void)m.getAABBTree();
    auto toXf = []( const float* x )
    {
        AffineXf3f xf;
        xf.A = Matrix3f( Vector3f( x[0], x[1], x[2] ), Vector3f( x[3], x[4], x[5] ), Vector3f( x[6], x[7], x[8] ) );
        xf.b = Vector3f( x[9], x[10], x[11] );
        return xf;
    };
    AffineXf3f pxf, txf;
    if ( pointXf12 ) pxf = toXf( pointXf12 );
    if ( treeXf12 ) txf = toXf( treeXf12 );
    std::vector<Vector3f> points( n );
    for ( size_t i = 0; i < n; ++i ) points[i] = Vector3f( pts[3 * i], pts[3 * i + 1], pts[3 * i + 2] );
    std::vector<MeshProjectionResult> res;
    PointsToMeshProjector proj;
    proj.updateMeshData( &m );
    proj.findProjections( res, points, pointXf12 ? &pxf : nullptr, treeXf12 ? &txf : nullptr, upDistLimitSq, loDistLimitSq );
    struct Rec { int face; float p[3]; float a, b; float distSq; int edge; };
    Rec* o = (Rec*)out;
    for ( size_t i = 0; i < n; ++i )
    {
        const auto& r = res[i];
        o[i] = Rec{ (int)r.proj.face, { r.proj.point.x, r.proj.point.y, r.proj.point.z }, r.mtp.bary.a, r.mtp.bary.b, r.distSq, (int)r.mtp.e };
    }
    return 0;
}

// signedDistanceToMesh, SignDetectionMode::ProjectionNormal (source/MRMesh/MRMeshDistance.cpp:13-38), through the reference's
// own findProjection and Mesh::isOutsideByProjNorm; MRMeshDistance.cpp itself is not among the compiled TUs (its other sign
// modes need the ray-intersection code), so the ten lines around those two calls are restated here.
// outDist[i]: signed distance or NaN; outDot[i] (optional): dot( proj - pt, pseudonormal ), how far the decision is from a tie
int mrref_signed_distance_proj_normal( void* h, const float* pts, size_t n, float minDistSqIn, float maxDistSqIn, int nullOutside,
                                       float* outDist, float* outDot )
{
    auto& m = ( (RefMesh*)h )->mesh;
    (void)m.getAABBTree();
    const float nan = std::numeric_limits<float>::quiet_NaN();
    tbb::parallel_for( tbb::blocked_range<size_t>( 0, n ), [&]( const tbb::blocked_range<size_t>& range )
    {
        for ( size_t i = range.begin(); i < range.end(); ++i )
        {
            const Vector3f p( pts[3 * i], pts[3 * i + 1], pts[3 * i + 2] );
            float minDistSq = minDistSqIn, maxDistSq = maxDistSqIn;
            if ( !nullOutside ) { minDistSq = 0; maxDistSq = FLT_MAX; } // :19-24
            const auto proj = findProjection( p, m, maxDistSq, nullptr, minDistSq );
            if ( outDot ) outDot[i] = nan;
            if ( !proj ) { outDist[i] = nan; continue; } // :27-28
            if ( nullOutside && ( proj.distSq < minDistSq || proj.distSq >= maxDistSq ) ) { outDist[i] = nan; continue; } // :30-31
            float dist = std::sqrt( proj.distSq );
            if ( !m.isOutsideByProjNorm( p, proj, nullptr ) ) dist = -dist; // :36-38
            outDist[i] = dist;
            if ( outDot ) outDot[i] = dot( proj.proj.point - p, m.pseudonormal( proj.mtp, nullptr ) );
        }
    } );
    return 0;
}

// voxel centres of meshToDistanceFunctionVolume (source/MRVoxels/MRMeshToDistanceVolume.cpp:83-84), x fastest
int mrref_voxel_centres( const float* origin, const float* voxelSize, const int* dims, float* out )
{
    const Vector3f o( origin[0], origin[1], origin[2] ), vs( voxelSize[0], voxelSize[1], voxelSize[2] );
    size_t i = 0;
    for ( int z = 0; z < dims[2]; ++z )
        for ( int y = 0; y < dims[1]; ++y )
            for ( int x = 0; x < dims[0]; ++x, ++i )
            {
                const auto coord = Vector3f( Vector3i( x, y, z ) ) + Vector3f::diagonal( 0.5f );
                const auto c = o + mult( vs, coord );
                out[3 * i] = c.x; out[3 * i + 1] = c.y; out[3 * i + 2] = c.z;
            }
    return 0;
}

// pseudonormals of all faces (leftNormal of edgeWithLeft), vertices, and of the three edges of every face
// (source/MRMesh/MRMeshMath.cpp:510-547); edge k of face f = ( v_k, v_k+1 ) of getTriVerts( f )
int mrref_pseudonormals( void* h, float* faceN, float* vertN, float* edgeN )
{
    auto& m = ( (RefMesh*)h )->mesh;
    const auto& topo = m.topology;
    for ( FaceId f{ 0 }; f < (int)topo.faceSize(); ++f )
    {
        Vector3f n, e[3];
        if ( topo.hasFace( f ) )
        {
            const EdgeId e0 = topo.edgeWithLeft( f );
            n = m.leftNormal( e0 );
            // left ring of f starting at e0: e0 = (v0,v1), then (v1,v2), (v2,v0)
            EdgeId ek = e0;
            for ( int k = 0; k < 3; ++k )
            {
                e[k] = m.pseudonormal( ek.undirected(), nullptr );
                ek = topo.prev( ek.sym() );
            }
        }
        faceN[3 * (int)f] = n.x; faceN[3 * (int)f + 1] = n.y; faceN[3 * (int)f + 2] = n.z;
        for ( int k = 0; k < 3; ++k )
            for ( int c = 0; c < 3; ++c ) edgeN[9 * (int)f + 3 * k + c] = e[k][c];
    }
    for ( VertId v{ 0 }; v < (int)m.points.size(); ++v )
    {
        Vector3f n;
        if ( topo.hasVert( v ) ) n = m.pseudonormal( v, nullptr );
        vertN[3 * (int)v] = n.x; vertN[3 * (int)v + 1] = n.y; vertN[3 * (int)v + 2] = n.z;
    }
    return 0;
}

int mrref_marching_cubes( const float* vol, const int* dims, const float* voxelSize, const float* origin, float iso,
    int lessInside, size_t* nVerts, size_t* nTris )
{
    SimpleVolume v;
    v.dims = Vector3i( dims[0], dims[1], dims[2] );
    v.voxelSize = Vector3f( voxelSize[0], voxelSize[1], voxelSize[2] );
    v.data.resize( size_t( dims[0] ) * dims[1] * dims[2] );
    std::memcpy( (void*)v.data.data(), vol, v.data.size() * sizeof( float ) );
    MarchingCubesParams p;
    p.origin = Vector3f( origin[0], origin[1], origin[2] );
    p.iso = iso;
    p.lessInside = lessInside != 0;
    auto r = marchingCubesAsTriMesh( v, p );
    if ( !r )
        return 1;
    g_lastMc = std::move( *r );
    *nVerts = g_lastMc.points.size();
    *nTris = g_lastMc.tris.size();
    return 0;
}

void mrref_marching_cubes_fetch( float* outVerts, int* outTris )
{
    std::memcpy( outVerts, (const void*)g_lastMc.points.data(), g_lastMc.points.size() * sizeof( Vector3f ) );
    for ( size_t t = 0; t < g_lastMc.tris.size(); ++t )
        for ( int k = 0; k < 3; ++k )
            outTris[3 * t + k] = (int)g_lastMc.tris.vec_[t][k];
    g_lastMc = {};
}

// ---- half-edge tables ---------------------------------------------------------------------------------------
// Loads tables built elsewhere through the reference's own deserialiser: the stream layout of MeshTopology::write
// (MRMeshTopology.cpp:2461-2477) into MeshTopology::read (:2479-2534), which runs computeValidsFromEdges and checkValidity.
// Returns 0 and the triangulation the reference reads back from them (getTriangulation), the number of undirected edges and of
// boundary loops; 1 when the reference rejects the tables.
int mrref_topology_from_tables( const void* halfEdges16B, uint32_t nHalfEdges, const int* edgePerVertex, uint32_t nVerts,
    const int* edgePerFace, uint32_t nFaces, int* outTris, int* outUndirectedEdges, int* outBoundaryLoops, int* outValidVerts )
{
    std::stringstream ss( std::ios::in | std::ios::out | std::ios::binary );
    ss.write( (const char*)&nHalfEdges, 4 );
    ss.write( (const char*)halfEdges16B, size_t( nHalfEdges ) * 16 );
    ss.write( (const char*)&nVerts, 4 );
    ss.write( (const char*)edgePerVertex, size_t( nVerts ) * 4 );
    ss.write( (const char*)&nFaces, 4 );
    ss.write( (const char*)edgePerFace, size_t( nFaces ) * 4 );
    MeshTopology topo;
    if ( !topo.read( ss ) )
        return 1;
    const auto t = topo.getTriangulation();
    for ( size_t f = 0; f < nFaces && f < t.size(); ++f )
        for ( int k = 0; k < 3; ++k )
            outTris[3 * f + k] = (int)t.vec_[f][k];
    *outUndirectedEdges = (int)topo.computeNotLoneUndirectedEdges();
    *outBoundaryLoops = (int)findRightBoundary( topo ).size();
    *outValidVerts = topo.numValidVerts();
    return 0;
}

// the reference's own builder on the same triangles (Mesh::fromTriMesh's core, MRMeshBuilder.cpp:518-540): the counts the tables
// above must agree with, and the wall time of the CPU step they replace
int mrref_builder_stats( const int* tris, size_t nFaces, int* outValidFaces, int* outUndirectedEdges, int* outBoundaryLoops, int* outValidVerts,
    double* outSeconds )
{
    Triangulation t;
    t.vec_.resize( nFaces );
    for ( size_t f = 0; f < nFaces; ++f )
        t.vec_[f] = ThreeVertIds{ VertId( tris[3 * f] ), VertId( tris[3 * f + 1] ), VertId( tris[3 * f + 2] ) };
    const auto t0 = std::chrono::steady_clock::now();
    const MeshTopology topo = MeshBuilder::fromTriangles( t );
    *outSeconds = std::chrono::duration<double>( std::chrono::steady_clock::now() - t0 ).count();
    *outValidFaces = topo.numValidFaces();
    *outUndirectedEdges = (int)topo.computeNotLoneUndirectedEdges();
    *outBoundaryLoops = (int)findRightBoundary( topo ).size();
    *outValidVerts = topo.numValidVerts();
    return 0;
}

} // extern "C"
