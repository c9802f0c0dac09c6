// b2fwn_meshlib_adapter.hpp -- header-only C++ adapter that plugs libb2fwn.so (C ABI, include/b2fwn.h) in behind
// MeshLib's own operator interface.  Compile it INSIDE a MeshLib build (it includes MeshLib headers); it adds no
// dependency to libb2fwn itself.
//
//   MR::B200::PointsToMeshProjector : MR::IPointsToMeshProjector
//       (interface: source/MRMesh/MRPointsToMeshProjector.h:12-31; stands in for MR::Cuda::PointsToMeshProjector,
//        source/MRCuda/MRCudaPointsToMeshProjector.h)
//   MR::B200::FastWindingNumber : MR::IFastWindingNumber, MR::IFastWindingNumberByParts
//       (interface: source/MRMesh/MRFastWindingNumber.h:11-45, 74-107;
//        the class it stands in for: MR::Cuda::FastWindingNumber, source/MRCuda/MRCudaFastWindingNumber.h:17-38)
//
// Conventions preserved: outputs are caller-owned and RESIZED by the callee; by-parts chunks are moved
// into the consumer on the calling thread in increasing z; errors travel as Expected<void> strings ("NVIDIA GPU error:
// ...", "Operation was canceled"); device data is created lazily at the first call and cached, mesh edits afterwards are
// not detected (MRCudaFastWindingNumber.cpp:42-54).  Unlike the reference GPU class, calcFromGrid* DO resize `res`
// (Appendix A.11 documents the reference's omission as a defect).
#pragma once

#include "MRMesh/MRFastWindingNumber.h"
#include "MRMesh/MRMesh.h"
#include "MRMesh/MRBitSet.h"
#include "MRMesh/MRAffineXf3.h"
#include "MRMesh/MRVector3.h"
#include "MRMesh/MRDistanceToMeshOptions.h"
#include "MRMesh/MRPointsToMeshProjector.h"
#include "MRMesh/MRMeshProject.h"
#include "MRMesh/MRTriMesh.h"
#include "b2fwn.h"

#include <cfloat>
#include <cstring>
#include <deque>
#include <memory>
#include <new>
#include <sstream>
#include <string>
#include <vector>

namespace MR
{
namespace B200
{

class FastWindingNumber final : public IFastWindingNumber, public IFastWindingNumberByParts
{
public:
    /// stores a reference: the mesh must outlive this object (as MR::FastWindingNumber, MRFastWindingNumber.h:54-56)
    explicit FastWindingNumber( const Mesh& mesh, int device = 0 ) : mesh_( mesh ), devices_{ device } {}
    /// several GPUs of this process (b2fwn_config.n_devices): the grid calls below deal their z-chunks over all of them and
    /// every GPU copies its chunks straight into `res` / the parts; the reference class is tied to device 0
    /// (MRCudaFastWindingNumber.cpp:48)
    FastWindingNumber( const Mesh& mesh, std::vector<int> devices ) : mesh_( mesh ), devices_( std::move( devices ) )
    {
        if ( devices_.empty() ) devices_.push_back( 0 );
    }
    ~FastWindingNumber() override { b2fwn_destroy( h_ ); }
    FastWindingNumber( const FastWindingNumber& ) = delete;
    FastWindingNumber& operator=( const FastWindingNumber& ) = delete;

    Expected<void> calcFromVector( std::vector<float>& res, const std::vector<Vector3f>& points, float beta, FaceId skipFace = {},
                                   const ProgressCallback& cb = {} ) override
    {
        if ( auto e = prepare_(); !e ) return e;
        res.resize( points.size() ); // MRFastWindingNumber.cpp:31
        static_assert( sizeof( Vector3f ) == 3 * sizeof( float ) );
        return status_( b2fwn_calc_from_vector( h_, reinterpret_cast<const float*>( points.data() ), points.size(), beta,
                                                skipFace.valid() ? int( skipFace ) : -1, res.data(), &progress_, (void*)&cb ) );
    }

    Expected<void> calcSelfIntersections( FaceBitSet& res, float beta, const ProgressCallback& cb = {} ) override
    {
        if ( auto e = prepare_(); !e ) return e;
        const size_t n = mesh_.topology.faceSize();
        res.clear();
        res.resize( n ); // MRFastWindingNumber.cpp:43
        std::vector<uint64_t> words( ( n + 63 ) / 64 );
        if ( auto e = status_( b2fwn_calc_self_intersections( h_, beta, words.data(), &progress_, (void*)&cb ) ); !e ) return e;
        for ( size_t f = 0; f < n; ++f )
            if ( ( words[f >> 6] >> ( f & 63 ) ) & 1 )
                res.set( FaceId( int( f ) ) );
        return {};
    }

    Expected<void> calcFromGrid( std::vector<float>& res, const Vector3i& dims, const AffineXf3f& gridToMeshXf, float beta,
                                 const ProgressCallback& cb = {} ) override
    {
        if ( auto e = prepare_(); !e ) return e;
        res.resize( size_t( dims.x ) * dims.y * dims.z ); // MRFastWindingNumber.cpp:59
        const int32_t d[3] = { dims.x, dims.y, dims.z };
        float xf[12];
        packXf_( gridToMeshXf, xf );
        return status_( b2fwn_calc_from_grid( h_, d, xf, beta, res.data(), &progress_, (void*)&cb ) );
    }

    Expected<void> calcFromGridWithDistances( std::vector<float>& res, const Vector3i& dims, const AffineXf3f& gridToMeshXf,
                                              const DistanceToMeshOptions& options, const ProgressCallback& cb ) override
    {
        if ( auto e = prepare_(); !e ) return e;
        res.resize( size_t( dims.x ) * dims.y * dims.z ); // MRFastWindingNumber.cpp:83
        const int32_t d[3] = { dims.x, dims.y, dims.z };
        float xf[12];
        packXf_( gridToMeshXf, xf );
        const b2fwn_dist_options o = packOptions_( options );
        return status_( b2fwn_calc_from_grid_with_distances( h_, d, xf, &o, res.data(), &progress_, (void*)&cb ) );
    }

    Expected<void> calcFromGridByParts( GridByPartsFunc resFunc, const Vector3i& dims, const AffineXf3f& gridToMeshXf, float beta,
                                        int layerOverlap, const ProgressCallback& cb ) override
    {
        if ( auto e = prepare_(); !e ) return e;
        const int32_t d[3] = { dims.x, dims.y, dims.z };
        float xf[12];
        packXf_( gridToMeshXf, xf );
        PartCtx ctx{ &resFunc, {}, {} };
        const int rc = b2fwn_calc_from_grid_by_parts_into( h_, d, xf, beta, layerOverlap, 0, &partBuffer_, &part_, &ctx, &progress_, (void*)&cb );
        return ctx.error.empty() ? status_( rc ) : unexpected( ctx.error );
    }

    Expected<void> calcFromGridWithDistancesByParts( GridByPartsFunc resFunc, const Vector3i& dims, const AffineXf3f& gridToMeshXf,
                                                     const DistanceToMeshOptions& options, int layerOverlap, const ProgressCallback& cb ) override
    {
        if ( auto e = prepare_(); !e ) return e;
        const int32_t d[3] = { dims.x, dims.y, dims.z };
        float xf[12];
        packXf_( gridToMeshXf, xf );
        const b2fwn_dist_options o = packOptions_( options );
        PartCtx ctx{ &resFunc, {}, {} };
        const int rc = b2fwn_calc_from_grid_with_distances_by_parts_into( h_, d, xf, &o, layerOverlap, 0, &partBuffer_, &part_, &ctx, &progress_, (void*)&cb );
        return ctx.error.empty() ? status_( rc ) : unexpected( ctx.error );
    }

    /// Beyond the interface: the dense loop of makeSignedByWindingNumber (source/MRVoxels/MRVDBConversions.cpp:
    /// 517-568) in one call.  `volume` holds dims.x*dims.y*dims.z unsigned distances (x fastest, the VolumeIndexer order of
    /// the reference loop); every voxel whose winding number exceeds windingNumberThreshold is negated in place.  The
    /// winding numbers are never stored: no 4 B/voxel D2H + second pass over the grid as with calcFromGrid[ByParts].
    Expected<void> makeSignedByWindingNumber( float* volume, const Vector3i& dims, const AffineXf3f& gridToMeshXf,
                                              float windingNumberThreshold, float windingNumberBeta, const ProgressCallback& cb = {} )
    {
        if ( auto e = prepare_(); !e ) return e;
        const int32_t d[3] = { dims.x, dims.y, dims.z };
        float xf[12];
        packXf_( gridToMeshXf, xf );
        return status_( b2fwn_make_signed_by_winding_number( h_, d, xf, windingNumberBeta, windingNumberThreshold, volume,
                                                              &progress_, (void*)&cb ) );
    }

    /// Beyond the interface: signedDistanceToMesh with SignDetectionMode::ProjectionNormal for a vector of points
    /// (source/MRMesh/MRMeshDistance.cpp:13-59): res[i] = distance to the closest point of the mesh, negative where the
    /// pseudonormal at that point says inside, NaN outside [minDistSq, maxDistSq) when options.nullOutsideMinMax.
    /// options.signMode is not consulted; the winding-number fields are not used.
    Expected<void> signedDistancesByProjectionNormal( std::vector<float>& res, const std::vector<Vector3f>& points,
                                                      const DistanceToMeshOptions& options, const ProgressCallback& cb = {} )
    {
        if ( auto e = prepare_(); !e ) return e;
        res.resize( points.size() );
        const b2fwn_dist_options o = packOptions_( options );
        return status_( b2fwn_signed_distances_by_projection_normal( h_, reinterpret_cast<const float*>( points.data() ), points.size(), &o,
                                                                      res.data(), &progress_, (void*)&cb ) );
    }

    /// Beyond the interface: the mesh this object was made for had its vertices moved (same topology).
    /// Counterpart of Mesh::updateCaches( changedVerts ) (source/MRMesh/MRMesh.cpp:565-578): the device tree is refitted like
    /// AABBTree::refit (MRAABBTree.cpp:62-94), dipoles and leaf triangles are recomputed -- no re-upload of the triangulation,
    /// no rebuild.  The reference's Cuda::FastWindingNumber never notices such a change.
    Expected<void> updatePoints()
    {
        if ( !h_ ) return {}; // not uploaded yet: the first call will read the current coordinates
        return status_( b2fwn_update_points( h_, reinterpret_cast<const float*>( mesh_.points.data() ), mesh_.points.size() ) );
    }

    /// Beyond the interface (SURVEY 8a a19 + 8f N4): mcOffsetMesh / generalOffsetMesh( Mode::Standard ) with
    /// SignDetectionMode::HoleWindingRule (source/MRVoxels/MROffset.cpp:117-246) with every stage on the device: the distance band
    /// on the grid of calcOriginAndDimensions, marching cubes at iso = offset, and the half-edge tables of the result
    /// (Mesh::fromTriMesh, MROffset.cpp:222-225).  The host only receives the finished arrays: points, and the tables in the
    /// layout of MeshTopology::write, loaded (and validated) by MeshTopology::read.  A marching-cubes result that is not an
    /// oriented manifold goes through the reference's own Mesh::fromTriMesh instead.
    Expected<Mesh> offsetMesh( float offset, float voxelSize, bool closeHolesInHoleWindingNumber = true, float windingNumberThreshold = 0.5f,
                               float windingNumberBeta = 2.f, const ProgressCallback& cb = {} )
    {
        if ( auto e = prepare_(); !e ) return unexpected( e.error() );
        b2fwn_offset_params p{};
        p.voxel_size = voxelSize; p.offset = offset; p.close_holes = closeHolesInHoleWindingNumber ? 1 : 0;
        p.winding_number_threshold = windingNumberThreshold; p.winding_number_beta = windingNumberBeta;
        b2fwn_offset_result r{};
        if ( auto e = status_( b2fwn_mc_offset_mesh( h_, &p, &r, &progress_, (void*)&cb ) ); !e ) return unexpected( e.error() );
        Mesh res;
        res.points.resize( r.n_verts );
        uint64_t nHalf = 0, nV = 0, nF = 0;
        if ( b2fwn_build_topology( h_, nullptr, 0, 0, 0, &nHalf, &nV, &nF ) != B2FWN_OK )
        {
            TriMesh tm;
            tm.points.resize( r.n_verts );
            tm.tris.resize( r.n_tris );
            static_assert( sizeof( ThreeVertIds ) == 3 * sizeof( int32_t ) );
            if ( auto e = status_( b2fwn_marching_cubes_fetch( h_, reinterpret_cast<float*>( tm.points.data() ), reinterpret_cast<int32_t*>( tm.tris.data() ) ) ); !e )
                return unexpected( e.error() );
            return Mesh::fromTriMesh( std::move( tm ) );
        }
        if ( auto e = status_( b2fwn_marching_cubes_fetch( h_, reinterpret_cast<float*>( res.points.data() ), nullptr ) ); !e ) return unexpected( e.error() );
        // the stream layout of MeshTopology::write (MRMeshTopology.cpp:2461-2477)
        std::string buf( 12 + nHalf * sizeof( b2fwn_half_edge ) + ( nV + nF ) * sizeof( int32_t ), '\0' );
        char* q = buf.data();
        const uint32_t cnt[3] = { uint32_t( nHalf ), uint32_t( nV ), uint32_t( nF ) };
        std::memcpy( q, &cnt[0], 4 );
        auto* edges = reinterpret_cast<b2fwn_half_edge*>( q + 4 );
        char* qv = q + 4 + nHalf * sizeof( b2fwn_half_edge );
        std::memcpy( qv, &cnt[1], 4 );
        char* qf = qv + 4 + nV * sizeof( int32_t );
        std::memcpy( qf, &cnt[2], 4 );
        // q + 4 is 4-byte aligned only: fetch into aligned vectors, then copy
        {
            std::vector<b2fwn_half_edge> he( nHalf );
            std::vector<int32_t> epv( nV ), epf( nF );
            if ( auto e = status_( b2fwn_build_topology_fetch( h_, he.data(), epv.data(), epf.data() ) ); !e ) return unexpected( e.error() );
            std::memcpy( edges, he.data(), nHalf * sizeof( b2fwn_half_edge ) );
            std::memcpy( qv + 4, epv.data(), nV * sizeof( int32_t ) );
            std::memcpy( qf + 4, epf.data(), nF * sizeof( int32_t ) );
        }
        std::istringstream in( std::move( buf ), std::ios::binary );
        if ( auto e = res.topology.read( in ); !e ) return unexpected( e.error() );
        return res;
    }

private:
    struct PartCtx
    {
        GridByPartsFunc* func;
        std::string error;
        std::deque<std::pair<int, std::vector<float>>> pending; // parts being filled by the library, by zOffset
    };

    // lazy one-time upload + device tree/dipole build (replaces prepareData_, MRCudaFastWindingNumber.cpp:46-85:
    // no host AABB tree, no host dipoles are needed -- only points and the triangulation)
    Expected<void> prepare_()
    {
        if ( h_ ) return {};
        const auto tris = mesh_.topology.getTriangulation(); // invalid faces come out with invalid (-1) vertex ids
        static_assert( sizeof( ThreeVertIds ) == 3 * sizeof( int32_t ) );
        b2fwn_config cfg{};
        cfg.device = devices_[0];
        if ( devices_.size() > 1 )
        {
            if ( devices_.size() > B2FWN_MAX_DEVICES ) return unexpected( std::string( "too many devices" ) );
            cfg.n_devices = int32_t( devices_.size() );
            for ( size_t i = 0; i < devices_.size(); ++i ) cfg.devices[i] = devices_[i];
        }
        return status_( b2fwn_create( reinterpret_cast<const float*>( mesh_.points.data() ), mesh_.points.size(),
                                      reinterpret_cast<const int32_t*>( tris.data() ), tris.size(), &cfg, &h_ ) );
    }

    static Expected<void> status_( int rc )
    {
        if ( rc == B2FWN_OK ) return {};
        return unexpected( std::string( b2fwn_last_error() ) );
    }

    static int progress_( float v, void* user )
    {
        const auto& cb = *static_cast<const ProgressCallback*>( user );
        return ( !cb || cb( v ) ) ? 1 : 0;
    }

    // the consumer receives an owned vector it may keep or discard (MRCudaFastWindingNumber.cpp:340-354): the library asks for
    // the storage a few parts ahead and fills it (overlapped copies) while later parts are computed; no second copy here
    static float* partBuffer_( const int32_t dims[3], int32_t zOffset, void* user )
    {
        auto& ctx = *static_cast<PartCtx*>( user );
        try
        {
            ctx.pending.emplace_back( zOffset, std::vector<float>( size_t( dims[0] ) * dims[1] * dims[2] ) );
        }
        catch ( const std::bad_alloc& )
        {
            ctx.error = "not enough host memory for a part of the grid";
            return nullptr;
        }
        return ctx.pending.back().second.data();
    }

    static int part_( const float*, const int32_t dims[3], int32_t zOffset, void* user )
    {
        auto& ctx = *static_cast<PartCtx*>( user );
        if ( ctx.pending.empty() || ctx.pending.front().first != zOffset ) { ctx.error = "by-parts protocol error"; return 1; }
        std::vector<float> chunk = std::move( ctx.pending.front().second );
        ctx.pending.pop_front();
        auto r = ( *ctx.func )( std::move( chunk ), Vector3i( dims[0], dims[1], dims[2] ), zOffset );
        if ( r ) return 0;
        ctx.error = r.error();
        return 1;
    }

    static void packXf_( const AffineXf3f& xf, float out[12] )
    {
        const Vector3f* rows[3] = { &xf.A.x, &xf.A.y, &xf.A.z };
        for ( int r = 0; r < 3; ++r )
            for ( int c = 0; c < 3; ++c )
                out[3 * r + c] = ( *rows[r] )[c];
        for ( int c = 0; c < 3; ++c ) out[9 + c] = xf.b[c];
    }

    static b2fwn_dist_options packOptions_( const DistanceToMeshOptions& o )
    {
        b2fwn_dist_options r{};
        r.min_dist_sq = o.minDistSq;
        r.max_dist_sq = o.maxDistSq;
        r.null_outside_min_max = o.nullOutsideMinMax ? 1 : 0;
        r.winding_number_threshold = o.windingNumberThreshold;
        r.winding_number_beta = o.windingNumberBeta;
        return r;
    }

    const Mesh& mesh_;
    std::vector<int> devices_;
    b2fwn_handle h_ = nullptr;
};

/// GPU back-end of IPointsToMeshProjector: the closest point of the mesh to each given point, with face and barycentrics.
/// Same contract as MR::Cuda::PointsToMeshProjector (source/MRCuda/MRCudaPointsToMeshProjector.cpp:30-129): updateMeshData
/// uploads the mesh (here: builds tree and triangles on the device), findProjections resizes `result`, decomposes the two
/// transforms with the reference's own createProjectionTransforms, and fills MeshTriPoint::e from the mesh topology.
class PointsToMeshProjector final : public IPointsToMeshProjector
{
public:
    explicit PointsToMeshProjector( int device = 0 ) : device_( device ) {}
    ~PointsToMeshProjector() override { b2fwn_destroy( h_ ); }
    PointsToMeshProjector( const PointsToMeshProjector& ) = delete;
    PointsToMeshProjector& operator=( const PointsToMeshProjector& ) = delete;

    void updateMeshData( const Mesh* mesh ) override
    {
        b2fwn_destroy( h_ );
        h_ = nullptr;
        mesh_ = mesh;
        if ( !mesh_ ) return;
        const auto tris = mesh_->topology.getTriangulation();
        b2fwn_config cfg{};
        cfg.device = device_;
        if ( b2fwn_create( reinterpret_cast<const float*>( mesh_->points.data() ), mesh_->points.size(),
                           reinterpret_cast<const int32_t*>( tris.data() ), tris.size(), &cfg, &h_ ) != B2FWN_OK )
        {
            lastError_ = std::string( "NVIDIA GPU error: " ) + b2fwn_last_error();
            h_ = nullptr;
        }
    }

    void findProjections( std::vector<MeshProjectionResult>& result, const std::vector<Vector3f>& points,
                          const AffineXf3f* worldXf = nullptr, const AffineXf3f* worldRefXf = nullptr,
                          float upDistLimitSq = FLT_MAX, float loDistLimitSq = 0.0f ) override
    {
        if ( !mesh_ || !h_ ) return; // as the reference: nothing without mesh data (MRPointsToMeshProjector.cpp:19-20)
        result.resize( points.size() );
        AffineXf3f storage;
        const auto xfs = createProjectionTransforms( storage, worldXf, worldRefXf ); // MRMeshProject.cpp:104-121
        float pxf[12], txf[12];
        if ( xfs.rigidXfPoint ) packXf12_( *xfs.rigidXfPoint, pxf );
        if ( xfs.nonRigidXfTree ) packXf12_( *xfs.nonRigidXfTree, txf );
        std::vector<b2fwn_projection> raw( points.size() );
        if ( b2fwn_find_projections( h_, reinterpret_cast<const float*>( points.data() ), points.size(), xfs.rigidXfPoint ? pxf : nullptr,
                                     xfs.nonRigidXfTree ? txf : nullptr, upDistLimitSq, loDistLimitSq, raw.data(), nullptr, nullptr ) != B2FWN_OK )
        {
            lastError_ = b2fwn_last_error();
            return;
        }
        for ( size_t i = 0; i < raw.size(); ++i )
        {
            const auto& r = raw[i];
            MeshProjectionResult m;
            m.distSq = r.dist_sq;
            if ( r.face >= 0 )
            {
                m.proj.face = FaceId( r.face );
                m.proj.point = Vector3f( r.point[0], r.point[1], r.point[2] );
                m.mtp.e = mesh_->topology.edgeWithLeft( m.proj.face ); // MRCudaPointsToMeshProjector.cpp:117-123
                m.mtp.bary = TriPointf( r.bary_a, r.bary_b );
            }
            result[i] = m;
        }
    }

    size_t projectionsHeapBytes( size_t numProjections ) const override
    {
        return numProjections * ( 3 * sizeof( float ) + sizeof( b2fwn_projection ) ); // device staging per projection
    }

    const std::string& lastError() const { return lastError_; }

private:
    static void packXf12_( const AffineXf3f& xf, float out[12] )
    {
        const Vector3f* rows[3] = { &xf.A.x, &xf.A.y, &xf.A.z };
        for ( int r = 0; r < 3; ++r )
            for ( int c = 0; c < 3; ++c )
                out[3 * r + c] = ( *rows[r] )[c];
        for ( int c = 0; c < 3; ++c ) out[9 + c] = xf.b[c];
    }

    const Mesh* mesh_ = nullptr;
    int device_ = 0;
    b2fwn_handle h_ = nullptr;
    std::string lastError_;
};

} // namespace B200
} // namespace MR
