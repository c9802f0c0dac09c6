// oracle/adapter_check.cpp -- TEST INFRASTRUCTURE ONLY.
//
// Drop-in check of meshlib_b200/host/b2fwn_meshlib_adapter.hpp against the UNMODIFIED reference classes
// (oracle/_ref/libmrref.so): the B200 object is used exactly the way MeshLib's callers use an IFastWindingNumber
//   * meshToDistanceVolume's call      (source/MRVoxels/MRMeshToDistanceVolume.cpp:52-56)
//   * mcOffsetMesh's by-parts branch   (source/MRVoxels/MROffset.cpp:194-226): calcFromGridWithDistancesByParts with
//     layerOverlap = 1 feeding the reference's own MarchingCubesByParts
// and every result is compared with MR::FastWindingNumber (the reference CPU/TBB implementation) on the same mesh.
// Built by `make -C oracle adapter` into oracle/_ref/adapter_check (travels to the GPU box); run by tests/ (-m gpu).
// Prints one "key value" line per check and exits non-zero on a parity failure.
#include "../meshlib_b200/host/b2fwn_meshlib_adapter.hpp"
#include "MRMesh/MRTorus.h"
#include "MRMesh/MRTriMesh.h"
#include "MRVoxels/MRMarchingCubes.h"
#include "MRVoxels/MRVoxelsVolume.h"
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <sstream>

using namespace MR;

static int g_fail = 0;
static void check( bool ok, const char* what )
{
    std::printf( "%s %s\n", what, ok ? "ok" : "FAIL" );
    if ( !ok ) ++g_fail;
}

// --bench pts.bin tris.bin nx ny nz ox oy oz voxel maxDistSq nullOutside devices(csv) [byparts]
// End to end through the C++ adapter the way a MeshLib caller sees it: a Mesh (built here by the reference's own
// Mesh::fromTriangles, not timed), then ONE calcFromGridWithDistances into a std::vector<float> (pageable), wall clock.  The
// first call includes prepare_ (upload + device tree build); the second reuses the handle.  `res` is pre-sized and touched
// before the timed call (the resize inside the adapter is then a no-op): value-initialising 34 GB on one thread is the
// caller's cost with the reference class as well and would hide everything else.
static double nowMs()
{
    return std::chrono::duration<double, std::milli>( std::chrono::steady_clock::now().time_since_epoch() ).count();
}
template <typename T>
static std::vector<T> readAll( const char* path )
{
    std::vector<T> v;
    FILE* f = std::fopen( path, "rb" );
    if ( !f ) return v;
    std::fseek( f, 0, SEEK_END );
    const long n = std::ftell( f );
    std::fseek( f, 0, SEEK_SET );
    v.resize( size_t( n ) / sizeof( T ) );
    if ( std::fread( v.data(), sizeof( T ), v.size(), f ) != v.size() ) v.clear();
    std::fclose( f );
    return v;
}
static int benchMain( int argc, char** argv )
{
    if ( argc < 14 ) { std::printf( "usage: --bench pts.bin tris.bin nx ny nz ox oy oz voxel maxDistSq nullOutside devices [byparts]\n" ); return 2; }
    auto pts = readAll<Vector3f>( argv[2] );
    auto tr = readAll<ThreeVertIds>( argv[3] );
    if ( pts.empty() || tr.empty() ) { std::printf( "cannot read the mesh files\n" ); return 2; }
    const Vector3i dims( std::atoi( argv[4] ), std::atoi( argv[5] ), std::atoi( argv[6] ) );
    const Vector3f origin( (float)std::atof( argv[7] ), (float)std::atof( argv[8] ), (float)std::atof( argv[9] ) );
    const float voxel = (float)std::atof( argv[10] );
    DistanceToMeshOptions opt;
    opt.maxDistSq = (float)std::atof( argv[11] );
    opt.nullOutsideMinMax = std::atoi( argv[12] ) != 0;
    std::vector<int> devices;
    { std::stringstream ss( argv[13] ); std::string t; while ( std::getline( ss, t, ',' ) ) devices.push_back( std::atoi( t.c_str() ) ); }
    const bool byParts = argc > 14 && !std::strcmp( argv[14], "byparts" );
    double t0 = nowMs();
    Triangulation tris;
    tris.vec_ = std::move( tr );
    VertCoords coords;
    coords.vec_ = std::move( pts );
    Mesh mesh = Mesh::fromTriangles( std::move( coords ), tris );
    std::printf( "bench.mesh_build_ms %.1f faces %zu\n", nowMs() - t0, (size_t)mesh.topology.numValidFaces() );
    const AffineXf3f basis{ Matrix3f::scale( voxel ), origin + Vector3f::diagonal( 0.5f * voxel ) };
    const size_t total = size_t( dims.x ) * dims.y * dims.z;
    std::vector<float> res( total, 1.f ); // sized and touched by the caller
    {
        // process warm-up on a tiny mesh: CUDA context, the library's staging ring (what a long-running MeshLib process has paid once)
        Mesh tiny = makeTorus( 2.f, 1.f, 16, 8 );
        B200::FastWindingNumber warm( tiny, devices );
        std::vector<float> w;
        t0 = nowMs();
        (void)warm.calcFromGridWithDistances( w, Vector3i( 16, 16, 16 ), AffineXf3f{ Matrix3f::scale( 0.4f ), Vector3f::diagonal( -3.f ) }, DistanceToMeshOptions{}, {} );
        std::printf( "bench.process_warmup_ms %.1f\n", nowMs() - t0 );
    }
    for ( int rep = 0; rep < 3; ++rep )
    {
        // rep 0: fresh object (prepare_ inside the call); rep 1: same object again; rep 2: fresh object, by-parts consumer copies into res
        static std::unique_ptr<B200::FastWindingNumber> fwn;
        if ( rep != 1 ) fwn = std::make_unique<B200::FastWindingNumber>( mesh, devices );
        t0 = nowMs();
        Expected<void> r;
        if ( rep == 2 )
        {
            if ( !byParts ) break;
            r = fwn->calcFromGridWithDistancesByParts( [&]( std::vector<float>&& data, const Vector3i& d, int z ) -> Expected<void>
            {
                std::memcpy( res.data() + size_t( z ) * d.x * d.y, data.data(), data.size() * sizeof( float ) );
                return {};
            }, dims, basis, opt, 0, {} );
        }
        else r = fwn->calcFromGridWithDistances( res, dims, basis, opt, {} );
        const double ms = nowMs() - t0;
        if ( !r ) { std::printf( "bench.call FAIL (%s)\n", r.error().c_str() ); return 1; }
        double sum = 0;
        for ( size_t i = 0; i < total; i += 4099 ) if ( res[i] == res[i] ) sum += std::abs( res[i] );
        std::printf( "bench.%s_ms %.1f voxels_per_s %.4g checksum %.6f\n", rep == 0 ? "first_call_with_prepare" : rep == 1 ? "second_call" : "by_parts_with_prepare",
                     ms, double( total ) / ( ms * 1e-3 ), sum );
    }
    return 0;
}

int main( int argc, char** argv )
{
    if ( argc > 1 && !std::strcmp( argv[1], "--bench" ) ) return benchMain( argc, argv );
    Mesh mesh = makeTorus( 2.f, 1.f, 64, 32 );
    // the reference-facing handle type used by every caller (params.fwn): std::shared_ptr<IFastWindingNumber>
    std::shared_ptr<IFastWindingNumber> ours = std::make_shared<B200::FastWindingNumber>( mesh );
    FastWindingNumber ref( mesh );

    // calcFromVector
    std::vector<Vector3f> pts;
    for ( int i = 0; i < 5000; ++i )
        pts.emplace_back( 3.5f * std::sin( 0.37f * i ), 3.5f * std::cos( 0.91f * i ), 1.5f * std::sin( 1.73f * i ) );
    std::vector<float> wa, wb;
    if ( auto e = ours->calcFromVector( wa, pts, 2.f ); !e )
    {
        std::printf( "calcFromVector.call FAIL (%s)\n", e.error().c_str() ); // e.g. no CUDA device: there is no CPU fallback
        return 2;
    }
    (void)ref.calcFromVector( wb, pts, 2.f, {}, {} );
    float maxDw = 0;
    for ( size_t i = 0; i < pts.size(); ++i ) maxDw = std::max( maxDw, std::abs( wa[i] - wb[i] ) );
    std::printf( "calcFromVector.max_abs_dw %g\n", maxDw );
    check( wa.size() == pts.size() && maxDw < 1e-5f, "calcFromVector.parity" );

    // calcSelfIntersections
    FaceBitSet sa, sb;
    check( ours->calcSelfIntersections( sa, 2.f ).has_value(), "calcSelfIntersections.call" );
    (void)ref.calcSelfIntersections( sb, 2.f, {} );
    check( sa.size() == sb.size() && sa == sb, "calcSelfIntersections.parity" );

    // calcFromGridWithDistances, as meshToDistanceVolume calls it (voxel centres)
    const Vector3i dims( 40, 40, 24 );
    const float voxel = 0.17f;
    const Vector3f origin( -3.4f, -3.4f, -2.04f );
    const AffineXf3f basis{ Matrix3f::scale( voxel ), origin + Vector3f::diagonal( 0.5f * voxel ) };
    DistanceToMeshOptions opt;
    std::vector<float> da, db, wgrid;
    check( ours->calcFromGridWithDistances( da, dims, basis, opt, {} ).has_value(), "calcFromGridWithDistances.call" );
    (void)ref.calcFromGridWithDistances( db, dims, basis, opt, {} );
    (void)ref.calcFromGrid( wgrid, dims, basis, opt.windingNumberBeta, {} );
    size_t signMism = 0, band = 0;
    float maxRel = 0;
    for ( size_t i = 0; i < db.size(); ++i )
    {
        if ( std::abs( wgrid[i] - opt.windingNumberThreshold ) < 1e-3f ) { ++band; continue; }
        if ( ( da[i] < 0 ) != ( db[i] < 0 ) ) ++signMism;
        maxRel = std::max( maxRel, std::abs( std::abs( da[i] ) - std::abs( db[i] ) ) / std::max( std::abs( db[i] ), 1e-30f ) );
    }
    std::printf( "calcFromGridWithDistances.sign_mismatch_outside_band %zu in_band %zu max_rel_dist_err %g\n", signMism, band, maxRel );
    check( da.size() == db.size() && signMism == 0 && maxRel <= 1e-5f, "calcFromGridWithDistances.parity" );

    // several GPUs of one process (b2fwn_config.n_devices): chunks dealt over the devices, every device copies into `res`
    // itself; the volume must be the single-device one bit for bit (same kernels, same per-voxel arithmetic)
    if ( b2fwn_device_count() >= 2 )
    {
        setenv( "B2FWN_MIN_CHUNK_VOXELS", "8000", 1 ); // test switch of the library: small chunks, so that both devices get several
        B200::FastWindingNumber multi( mesh, std::vector<int>{ 0, 1 } );
        std::vector<float> dm;
        const Vector3i big( 40, 40, 96 ); // enough layers for several chunks per device
        std::vector<float> d1;
        check( multi.calcFromGridWithDistances( dm, big, basis, opt, {} ).has_value(), "multiDevice.call" );
        check( ours->calcFromGridWithDistances( d1, big, basis, opt, {} ).has_value(), "multiDevice.single.call" );
        check( dm.size() == d1.size() && std::memcmp( dm.data(), d1.data(), dm.size() * sizeof( float ) ) == 0, "multiDevice.bit_identical" );
        unsetenv( "B2FWN_MIN_CHUNK_VOXELS" );
    }
    else std::printf( "multiDevice skipped (one device)\n" );

    // mcOffsetMesh's by-parts branch: our chunks -> the reference's MarchingCubesByParts
    const float offset = 0.1f;
    DistanceToMeshOptions band_;
    band_.minDistSq = ( offset - 1.001f * voxel ) > 0 ? ( offset - 1.001f * voxel ) * ( offset - 1.001f * voxel ) : 0.f;
    band_.maxDistSq = ( offset + 1.001f * voxel ) * ( offset + 1.001f * voxel );
    band_.nullOutsideMinMax = true;
    MarchingCubesParams mp;
    mp.origin = origin;
    mp.iso = offset;
    mp.lessInside = true;
    auto byParts = std::dynamic_pointer_cast<IFastWindingNumberByParts>( ours );
    check( bool( byParts ), "byParts.dynamic_pointer_cast" );
    MarchingCubesByParts mesher( dims, mp );
    int parts = 0, lastZ = -1;
    bool zOrder = true;
    auto addPart = [&]( std::vector<float>&& data, const Vector3i& d, int zOffset ) -> Expected<void>
    {
        zOrder = zOrder && zOffset == mesher.nextZ() && zOffset > lastZ;
        lastZ = zOffset;
        ++parts;
        SimpleVolume v;
        v.data = std::move( data );
        v.dims = d;
        v.voxelSize = Vector3f::diagonal( voxel );
        return mesher.addPart( v );
    };
    auto r = byParts->calcFromGridWithDistancesByParts( addPart, dims, basis, band_, 1, {} );
    check( r.has_value() && zOrder && parts >= 1, "byParts.contract" );
    auto triA = mesher.finalize();
    // same through the reference CPU FastWindingNumber + whole-volume marching cubes
    std::vector<float> full;
    (void)ref.calcFromGridWithDistances( full, dims, basis, band_, {} );
    SimpleVolume vol;
    vol.data = std::move( full );
    vol.dims = dims;
    vol.voxelSize = Vector3f::diagonal( voxel );
    auto triB = marchingCubesAsTriMesh( vol, mp );
    check( triA.has_value() && triB.has_value(), "marchingCubes.call" );
    if ( triA && triB )
    {
        std::printf( "offsetMesh.verts %zu %zu tris %zu %zu\n", triA->points.size(), triB->points.size(), triA->tris.size(), triB->tris.size() );
        check( triA->points.size() == triB->points.size() && triA->tris.size() == triB->tris.size() && triA->tris.size() > 0, "offsetMesh.counts" );
    }

    // the whole offset pipeline on the device, half-edge tables included (SURVEY 8a a19 + 8f N4): the Mesh must be valid for the
    // reference (MeshTopology::read ran checkValidity) and equal the TriMesh above in every count
    {
        auto b200 = std::dynamic_pointer_cast<B200::FastWindingNumber>( ours );
        auto m = b200->offsetMesh( offset, voxel, false );
        check( m.has_value(), "offsetMesh.device_pipeline.call" );
        if ( m )
        {
            // its own grid (calcOriginAndDimensions of the mesh box), so not the TriMesh above: an offset surface of a torus is a closed
            // orientable surface of genus 1 -- V - E + F = 0, no boundary -- and the reference must accept the tables
            const int V = m->topology.numValidVerts(), F = m->topology.numValidFaces();
            const int E = (int)m->topology.computeNotLoneUndirectedEdges();
            std::printf( "offsetMesh.device_pipeline verts %d faces %d edges %d\n", V, F, E );
            check( F > 1000 && (int)m->points.size() == V && m->topology.checkValidity(), "offsetMesh.device_pipeline.valid_mesh" );
            check( V - E + F == 0 && 2 * E == 3 * F, "offsetMesh.device_pipeline.closed_genus_1" );
            float maxDev = 0; // every vertex of the offset surface lies at distance |offset| from the torus (to a voxel's interpolation error)
            for ( const auto& p : m->points )
            {
                const float rho = std::sqrt( p.x * p.x + p.y * p.y ) - 2.f;
                maxDev = std::max( maxDev, std::abs( std::sqrt( rho * rho + p.z * p.z ) - ( 1.f + offset ) ) );
            }
            std::printf( "offsetMesh.device_pipeline.max_radial_deviation %g\n", maxDev );
            check( maxDev < 0.02f, "offsetMesh.device_pipeline.geometry" );
        }
    }

    // cancellation travels as the reference's error string
    auto c = ours->calcFromGridWithDistances( da, dims, basis, opt, []( float ) { return false; } );
    check( !c.has_value() && c.error() == "Operation was canceled", "cancel.error_string" );

    // makeSignedByWindingNumber: the reference's dense loop -- calcFromGrid, then negate every voxel with
    // w > threshold (MRVDBConversions.cpp:527-533, 552-565) -- against the one in-place call of our class
    {
        auto b200 = std::dynamic_pointer_cast<B200::FastWindingNumber>( ours );
        std::vector<float> ua( db.size() ), ub( db.size() );
        for ( size_t i = 0; i < db.size(); ++i ) ua[i] = ub[i] = std::abs( db[i] );
        const float thr = 0.4f, beta = 2.5f;
        std::vector<float> w;
        (void)ref.calcFromGrid( w, dims, basis, beta, {} );
        for ( size_t i = 0; i < ub.size(); ++i )
            if ( w[i] > thr ) ub[i] = -ub[i];
        check( b200 && b200->makeSignedByWindingNumber( ua.data(), dims, basis, thr, beta ).has_value(), "makeSignedByWindingNumber.call" );
        size_t mism = 0, inBand = 0, magDiff = 0, negatives = 0;
        for ( size_t i = 0; i < ub.size(); ++i )
        {
            if ( std::abs( ua[i] ) != std::abs( ub[i] ) ) ++magDiff;
            if ( ua[i] < 0 ) ++negatives;
            if ( std::abs( w[i] - thr ) < 1e-3f ) { ++inBand; continue; }
            if ( ( ua[i] < 0 ) != ( ub[i] < 0 ) ) ++mism;
        }
        std::printf( "makeSignedByWindingNumber.sign_mismatch_outside_band %zu in_band %zu magnitude_changed %zu negatives %zu\n", mism, inBand, magDiff, negatives );
        check( mism == 0 && magDiff == 0 && negatives > 0, "makeSignedByWindingNumber.parity" );
    }

    // IPointsToMeshProjector: our back-end against the reference's CPU class through the interface pointer,
    // the way findSignedDistances receives it (MRPointsToMeshProjector.cpp:35-62); rigid and non-rigid transform cases
    {
        std::unique_ptr<IPointsToMeshProjector> oursP = std::make_unique<B200::PointsToMeshProjector>();
        PointsToMeshProjector refP;
        oursP->updateMeshData( &mesh );
        refP.updateMeshData( &mesh );
        const AffineXf3f rigid( Matrix3f::rotation( Vector3f( 0.3f, 0.5f, 0.8f ).normalized(), 0.7f ), Vector3f( 0.2f, -0.1f, 0.3f ) );
        const AffineXf3f shear( Matrix3f( Vector3f( 1.2f, 0.1f, 0.f ), Vector3f( 0.f, 0.9f, 0.2f ), Vector3f( 0.05f, 0.f, 1.1f ) ), Vector3f( 0.3f, 0.1f, -0.2f ) );
        struct Case { const char* name; const AffineXf3f* xf; const AffineXf3f* refXf; float up, lo; };
        const Case cases[] = { { "plain", nullptr, nullptr, FLT_MAX, 0.f }, { "limits", nullptr, nullptr, 0.5f, 0.01f },
                               { "pointXf", &rigid, nullptr, FLT_MAX, 0.f }, { "rigidRefXf", &shear, &rigid, FLT_MAX, 0.f },
                               { "nonRigidRefXf", &rigid, &shear, 2.f, 0.f } };
        for ( const auto& c : cases )
        {
            std::vector<MeshProjectionResult> ra, rb;
            oursP->findProjections( ra, pts, c.xf, c.refXf, c.up, c.lo );
            refP.findProjections( rb, pts, c.xf, c.refXf, c.up, c.lo );
            bool same = ra.size() == rb.size() && ra.size() == pts.size();
            for ( size_t i = 0; same && i < ra.size(); ++i )
                same = ra[i].proj.face == rb[i].proj.face && ra[i].mtp.e == rb[i].mtp.e && ra[i].distSq == rb[i].distSq &&
                       ( !rb[i].proj.face.valid() || ( ra[i].proj.point == rb[i].proj.point && ra[i].mtp.bary.a == rb[i].mtp.bary.a && ra[i].mtp.bary.b == rb[i].mtp.bary.b ) );
            check( same, ( std::string( "findProjections." ) + c.name ).c_str() );
        }
        check( oursP->projectionsHeapBytes( 10 ) == 10 * ( 12 + 32 ), "projectionsHeapBytes" );
    }

    // moved vertices: Mesh::updateCaches refits the reference's cached tree, updatePoints() refits ours; the
    // two trees stay identical, so the winding numbers agree as tightly as after a build
    {
        auto b200 = std::dynamic_pointer_cast<B200::FastWindingNumber>( ours );
        for ( auto& p : mesh.points )
            p = Vector3f( 1.1f * p.x + 0.05f * std::sin( 3.f * p.y ), 0.9f * p.y, p.z + 0.1f * std::cos( 2.f * p.x ) );
        mesh.updateCaches( mesh.topology.getValidVerts() );
        check( b200 && b200->updatePoints().has_value(), "updatePoints.call" );
        FastWindingNumber ref2( mesh );
        std::vector<float> ua, ub;
        check( ours->calcFromVector( ua, pts, 2.f ).has_value(), "updatePoints.calcFromVector.call" );
        (void)ref2.calcFromVector( ub, pts, 2.f, {}, {} );
        float maxD = 0;
        for ( size_t i = 0; i < pts.size(); ++i ) maxD = std::max( maxD, std::abs( ua[i] - ub[i] ) );
        std::printf( "updatePoints.max_abs_dw %g\n", maxD );
        check( ua.size() == pts.size() && maxD < 1e-5f, "updatePoints.parity" );
    }
    return g_fail ? 1 : 0;
}
