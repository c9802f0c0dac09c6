// b2fwn_api.cu -- the C ABI of include/b2fwn.h: handle, memory, streams, chunking / by-parts pipelines.
// Host-side orchestration only; all arithmetic lives in b2fwn_tree.cu / b2fwn_kernels.cu / b2fwn_mc.cu.
// There is deliberately no CPU implementation behind any entry point.
#include "b2fwn_ctx.h"
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

namespace b2
{

static thread_local std::string g_lastError;
static thread_local cudaError_t g_lastCudaError = cudaSuccess;

void setError( const std::string& msg ) { g_lastError = msg; g_lastCudaError = cudaSuccess; }
cudaError_t lastCudaError() { return g_lastCudaError; }

int cudaFail( cudaError_t e, const char* where )
{
    // reference wording: source/MRCuda/MRCudaBasic.cpp:109-112
    g_lastError = std::string( "NVIDIA GPU error: " ) + cudaGetErrorString( e ) + " [" + where + "]";
    g_lastCudaError = e;
    (void)cudaGetLastError();
    return B2FWN_ERR_CUDA;
}

} // namespace b2

using namespace b2;

namespace
{

const char* kCanceled = "Operation was canceled"; // source/MRMesh/MRExpected.h:42-51

bool report( b2fwn_progress cb, void* user, float v )
{
    return !cb || cb( v, user ) != 0;
}

int canceled()
{
    setError( kCanceled );
    return B2FWN_ERR_CANCELED;
}

GridXf toXf( const float xf[12] )
{
    GridXf g;
    for ( int i = 0; i < 9; ++i ) g.A[i] = xf[i];
    for ( int i = 0; i < 3; ++i ) g.b[i] = xf[9 + i];
    return g;
}

struct KernelTimer
{
    b2fwn_ctx* h;
    explicit KernelTimer( b2fwn_ctx* c ) : h( c ) { cudaEventRecord( h->evA, h->stream ); }
    void stop()
    {
        cudaEventRecord( h->evB, h->stream );
        if ( cudaEventSynchronize( h->evB ) == cudaSuccess )
        {
            float ms = 0;
            cudaEventElapsedTime( &ms, h->evA, h->evB );
            h->stats.last_kernel_ms = ms;
        }
    }
};

// the common shape of the host-output grid calls: whole range or a z-range into a contiguous host buffer
int runToHost( b2fwn_ctx* h, const int32_t dims[3], int zBegin, int zEnd, float* hostOut, b2fwn_progress cb, void* user, const ChunkLaunch& launch )
{
    ChunkedCall c;
    c.dims = dims; c.zBegin = zBegin; c.zEnd = zEnd; c.hostOut = hostOut; c.progress = cb; c.progressUser = user;
    return runChunked( h, c, launch );
}

int runByParts( b2fwn_ctx* h, const int32_t dims[3], int overlap, int maxLayers, b2fwn_part_buffer_cb buffer, b2fwn_part_cb part, void* partUser,
                b2fwn_progress cb, void* user, const ChunkLaunch& launch )
{
    ChunkedCall c;
    c.dims = dims; c.zBegin = 0; c.zEnd = dims[2]; c.overlap = overlap; c.layersPerPart = maxLayers;
    c.part = part; c.partBuffer = buffer; c.partUser = partUser; c.progress = cb; c.progressUser = user;
    return runChunked( h, c, launch );
}

} // anonymous namespace

namespace
{

void destroyCtx( b2fwn_ctx* h )
{
    if ( !h ) return;
    cudaSetDevice( h->device );
    cudaDeviceSynchronize();
    freeTree( h->tree );
    freeMc( h->mc );
    freeMcWorkspace( h->mcWork );
    freeTopology( h->topo );
    for ( int s = 0; s < 2; ++s )
    {
        if ( h->dScratch[s] ) cudaFree( h->dScratch[s] );
        if ( h->hPinned[s] ) cudaFreeHost( h->hPinned[s] );
        if ( h->partDone[s] ) cudaEventDestroy( h->partDone[s] );
        if ( h->computeDone[s] ) cudaEventDestroy( h->computeDone[s] );
    }
    if ( h->dMinMax ) cudaFree( h->dMinMax );
    if ( h->evA ) cudaEventDestroy( h->evA );
    if ( h->evB ) cudaEventDestroy( h->evB );
    if ( h->stream ) cudaStreamDestroy( h->stream );
    if ( h->stream2 ) cudaStreamDestroy( h->stream2 );
    if ( h->copyStream ) cudaStreamDestroy( h->copyStream );
    delete h;
}

// streams, events and the private stream-ordered pool of one device context; on failure the half-made context is destroyed
int newCtx( int dev, b2fwn_ctx** out )
{
    *out = nullptr;
    B2_CUDA( cudaSetDevice( dev ) );
    b2fwn_ctx* h = new b2fwn_ctx;
    h->device = dev;
#define B2_TRY( expr ) do { cudaError_t _e = ( expr ); if ( _e != cudaSuccess ) { const int rc = cudaFail( _e, #expr ); destroyCtx( h ); return rc; } } while ( 0 )
    B2_TRY( cudaStreamCreateWithFlags( &h->stream, cudaStreamNonBlocking ) );
    B2_TRY( cudaStreamCreateWithFlags( &h->stream2, cudaStreamNonBlocking ) );
    B2_TRY( cudaStreamCreateWithFlags( &h->copyStream, cudaStreamNonBlocking ) );
    B2_TRY( cudaEventCreate( &h->evA ) );
    B2_TRY( cudaEventCreate( &h->evB ) );
    for ( int s = 0; s < 2; ++s )
    {
        B2_TRY( cudaEventCreateWithFlags( &h->partDone[s], cudaEventDisableTiming ) );
        B2_TRY( cudaEventCreateWithFlags( &h->computeDone[s], cudaEventDisableTiming ) );
    }
#undef B2_TRY
    if ( int e = initTreeAux( h->tree, dev ) ) { destroyCtx( h ); return e; }
    *out = h;
    return B2FWN_OK;
}

// one device's share of b2fwn_create: context, mesh upload, tree + dipole build
int createOne( int dev, const float* points, size_t nVerts, const int32_t* tris, size_t nFaces, b2fwn_ctx** out )
{
    b2fwn_ctx* h = nullptr;
    if ( int e = newCtx( dev, &h ) ) return e;
#define B2_TRY( expr ) do { cudaError_t _e = ( expr ); if ( _e != cudaSuccess ) { const int rc = cudaFail( _e, #expr ); destroyCtx( h ); return rc; } } while ( 0 )
    h->tree.numVerts = nVerts; h->tree.numFaces = nFaces;
    if ( nVerts )
    {
        B2_TRY( cudaMalloc( &h->tree.points, nVerts * 3 * sizeof( float ) ) );
        B2_TRY( cudaMemcpyAsync( h->tree.points, points, nVerts * 3 * sizeof( float ), cudaMemcpyHostToDevice, h->stream ) );
    }
    if ( nFaces )
    {
        B2_TRY( cudaMalloc( &h->tree.tris, nFaces * 3 * sizeof( int32_t ) ) );
        B2_TRY( cudaMemcpyAsync( h->tree.tris, tris, nFaces * 3 * sizeof( int32_t ), cudaMemcpyHostToDevice, h->stream ) );
    }
#undef B2_TRY
    cudaEventRecord( h->evA, h->stream );
    if ( int e = buildTree( h->tree, h->stream, &h->stats.kernel_launches ) ) { destroyCtx( h ); return e; }
    cudaEventRecord( h->evB, h->stream );
    cudaEventSynchronize( h->evB );
    float ms = 0;
    cudaEventElapsedTime( &ms, h->evA, h->evB );
    h->stats.tree_build_ms = ms;
    h->stats.num_nodes = h->tree.numNodes;
    h->stats.num_leaves = h->tree.numLeaves;
    *out = h;
    return B2FWN_OK;
}

// runs fn( ctx ) for the primary on the calling thread and for every replica on a thread of its own; first error wins
template <typename Fn>
int onAllDevices( b2fwn_ctx* h, Fn fn )
{
    std::vector<int> rc( h->replicas.size(), B2FWN_OK );
    std::vector<std::string> err( h->replicas.size() );
    std::vector<std::thread> th;
    for ( size_t i = 0; i < h->replicas.size(); ++i )
        th.emplace_back( [&, i] { rc[i] = fn( h->replicas[i] ); if ( rc[i] ) err[i] = b2fwn_last_error(); } );
    const int rc0 = fn( h );
    for ( auto& t : th ) t.join();
    cudaSetDevice( h->device );
    if ( rc0 ) return rc0;
    for ( size_t i = 0; i < rc.size(); ++i )
        if ( rc[i] ) { setError( err[i] ); return rc[i]; }
    return B2FWN_OK;
}

} // namespace

#pragma GCC visibility push( default )
extern "C"
{

const char* b2fwn_last_error( void ) { return g_lastError.c_str(); }

int b2fwn_device_count( void )
{
    int n = 0;
    if ( cudaGetDeviceCount( &n ) != cudaSuccess ) { (void)cudaGetLastError(); return 0; }
    return n;
}

int b2fwn_create( const float* points, size_t nVerts, const int32_t* tris, size_t nFaces, const b2fwn_config* cfg,
                  b2fwn_handle* out )
{
    if ( !out || ( nVerts && !points ) || ( nFaces && !tris ) ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    *out = nullptr;
    std::vector<int> devices{ cfg ? cfg->device : 0 };
    if ( cfg && cfg->n_devices > 1 )
    {
        if ( cfg->n_devices > B2FWN_MAX_DEVICES ) { setError( "too many devices" ); return B2FWN_ERR_INVALID; }
        devices.assign( cfg->devices, cfg->devices + cfg->n_devices );
        for ( size_t i = 0; i < devices.size(); ++i )
            for ( size_t j = 0; j < i; ++j )
                if ( devices[i] == devices[j] ) { setError( "device listed twice" ); return B2FWN_ERR_INVALID; }
    }
    // Every device builds its own copy of the tree, all at once (one host thread per device): the build is deterministic, so the
    // copies are identical, and the wall time is that of one build.  (r02: building once and copying the arrays to the other
    // devices over NVLink needs peer access between the devices, and with peer access enabled every later cudaMalloc / cudaFree of
    // the process is mapped into the peers as well -- create + destroy of a 2-device handle cost 450 ms more than of one device.)
    std::vector<b2fwn_ctx*> made( devices.size(), nullptr );
    std::vector<int> rc( devices.size(), B2FWN_OK );
    std::vector<std::string> err( devices.size() );
    std::vector<std::thread> th;
    for ( size_t i = 1; i < devices.size(); ++i )
        th.emplace_back( [&, i] { rc[i] = createOne( devices[i], points, nVerts, tris, nFaces, &made[i] ); if ( rc[i] ) err[i] = b2fwn_last_error(); } );
    rc[0] = createOne( devices[0], points, nVerts, tris, nFaces, &made[0] );
    for ( auto& t : th ) t.join();
    for ( size_t i = 0; i < devices.size(); ++i )
        if ( rc[i] )
        {
            if ( i ) setError( err[i] );
            for ( b2fwn_ctx* c : made ) destroyCtx( c );
            return rc[i];
        }
    b2fwn_ctx* h = made[0];
    h->replicas.assign( made.begin() + 1, made.end() );
    cudaSetDevice( h->device );
    *out = h;
    return B2FWN_OK;
}

void b2fwn_destroy( b2fwn_handle h )
{
    if ( !h ) return;
    for ( b2fwn_ctx* r : h->replicas ) destroyCtx( r );
    destroyCtx( h );
}

int b2fwn_update_points( b2fwn_handle h, const float* points, size_t nVerts )
{
    if ( !h || ( nVerts && !points ) ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    if ( nVerts != h->tree.numVerts ) { setError( "the number of vertices must not change (same triangulation, new coordinates)" ); return B2FWN_ERR_INVALID; }
    return onAllDevices( h, [&]( b2fwn_ctx* d ) -> int
    {
        B2_CUDA( cudaSetDevice( d->device ) );
        B2_CUDA( cudaDeviceSynchronize() ); // nothing of an earlier call still reads the old coordinates
        if ( nVerts ) B2_CUDA( cudaMemcpyAsync( d->tree.points, points, nVerts * 3 * sizeof( float ), cudaMemcpyHostToDevice, d->stream ) );
        cudaEventRecord( d->evA, d->stream );
        if ( int e = refitTree( d->tree, d->stream, &d->stats.kernel_launches ) ) return e;
        cudaEventRecord( d->evB, d->stream );
        B2_CUDA( cudaEventSynchronize( d->evB ) );
        float ms = 0;
        if ( cudaEventElapsedTime( &ms, d->evA, d->evB ) == cudaSuccess ) d->stats.last_kernel_ms = ms;
        return B2FWN_OK;
    } );
}

size_t b2fwn_safe_memory_limit( int32_t device ) { return safeMemoryLimit( device ); }

int b2fwn_trim( b2fwn_handle h )
{
    if ( !h ) { setError( "null handle" ); return B2FWN_ERR_INVALID; }
    const int rc = onAllDevices( h, [&]( b2fwn_ctx* d ) -> int
    {
        B2_CUDA( cudaSetDevice( d->device ) );
        reclaimCaches( d, RECLAIM_ALL );
        return B2FWN_OK;
    } );
    cudaSetDevice( h->device );
    return rc;
}

int b2fwn_plan_parts( int32_t nLayers, int32_t layersPerPart, int32_t overlap, int32_t nDevices, int32_t capacity, int32_t* outBegin,
                      int32_t* outLayers, int32_t* outDevice, int32_t* nParts )
{
    if ( !nParts ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    if ( overlap < 0 ) { setError( "negative layer overlap" ); return B2FWN_ERR_INVALID; }
    if ( nLayers > 0 && layersPerPart <= overlap ) { setError( "part must be larger than the overlap" ); return B2FWN_ERR_INVALID; }
    // a grid with no more layers than the overlap: the calls refuse it (their one part could not be larger than the overlap), so does the plan
    if ( nLayers > 0 && nLayers <= overlap ) { setError( "part must be larger than the overlap" ); return B2FWN_ERR_INVALID; }
    const int count = partCount( nLayers, std::min( layersPerPart, std::max( nLayers, 1 ) ), overlap );
    *nParts = count;
    const int nd = std::max( 1, std::min( nDevices, count ) ), step = std::min( layersPerPart, std::max( nLayers, 1 ) ) - overlap;
    for ( int k = 0; k < count && k < capacity; ++k )
    {
        if ( outBegin ) outBegin[k] = k * step;
        if ( outLayers ) outLayers[k] = std::min( nLayers, k * step + std::min( layersPerPart, nLayers ) ) - k * step;
        if ( outDevice ) outDevice[k] = k % nd; // chunks are dealt round-robin: index into the handle's device list
    }
    return B2FWN_OK;
}

int b2fwn_plan_parts_for( b2fwn_handle h, const int32_t dims[3], int32_t zBegin, int32_t zEnd, int32_t overlap, int32_t maxLayersPerPart,
                          int32_t capacity, int32_t* outBegin, int32_t* outLayers, int32_t* outDevice, int32_t* nParts, int32_t* layersPerPart )
{
    if ( !h || !dims || !nParts ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    if ( dims[0] < 0 || dims[1] < 0 || zBegin < 0 || zEnd > dims[2] || zBegin > zEnd ) { setError( "invalid layer range" ); return B2FWN_ERR_INVALID; }
    if ( overlap < 0 ) { setError( "negative layer overlap" ); return B2FWN_ERR_INVALID; }
    *nParts = 0;
    if ( layersPerPart ) *layersPerPart = 0;
    const int nz = zEnd - zBegin;
    if ( nz == 0 || (size_t)dims[0] * dims[1] == 0 ) return B2FWN_OK; // the calls return at once as well
    const int lpp = layersPerPartOf( h, dims, nz, overlap, maxLayersPerPart );
    cudaSetDevice( h->device );
    if ( layersPerPart ) *layersPerPart = lpp;
    if ( int e = b2fwn_plan_parts( nz, lpp, overlap, 1 + (int)h->replicas.size(), capacity, outBegin, outLayers, outDevice, nParts ) ) return e;
    if ( outBegin )
        for ( int k = 0; k < *nParts && k < capacity; ++k ) outBegin[k] += zBegin;
    return B2FWN_OK;
}

int b2fwn_get_stats( b2fwn_handle h, b2fwn_stats* out )
{
    if ( !h || !out ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    *out = h->stats;
    for ( const b2fwn_ctx* r : h->replicas ) out->kernel_launches += r->stats.kernel_launches;
    return B2FWN_OK;
}

int b2fwn_get_tree( b2fwn_handle h, void* outNodes, void* outDipoles, size_t capacity, size_t* nNodes )
{
    if ( !h ) { setError( "null handle" ); return B2FWN_ERR_INVALID; }
    if ( nNodes ) *nNodes = h->tree.numNodes;
    if ( capacity < h->tree.numNodes || h->tree.numNodes == 0 ) return B2FWN_OK;
    B2_CUDA( cudaSetDevice( h->device ) );
    if ( outNodes ) B2_CUDA( cudaMemcpy( outNodes, h->tree.refNodes, h->tree.numNodes * sizeof( RefNode ), cudaMemcpyDeviceToHost ) );
    if ( outDipoles ) B2_CUDA( cudaMemcpy( outDipoles, h->tree.refDipoles, h->tree.numNodes * sizeof( RefDipole ), cudaMemcpyDeviceToHost ) );
    return B2FWN_OK;
}

// ---- winding numbers of a point vector ------------------------------------------------------------------------------

int b2fwn_calc_from_vector_device( b2fwn_handle h, const float* dPts, size_t n, float beta, int32_t skipFace, float* dOut )
{
    if ( !h || ( n && ( !dPts || !dOut ) ) ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    KernelTimer tm( h );
    if ( int e = launchFwnVector( h->tree, dPts, n, beta, skipFace, dOut, h->stream, &h->stats.kernel_launches ) ) return e;
    tm.stop();
    B2_CUDA( cudaStreamSynchronize( h->stream ) );
    return B2FWN_OK;
}

int b2fwn_calc_from_vector( b2fwn_handle h, const float* pts, size_t n, float beta, int32_t skipFace, float* out,
                            b2fwn_progress cb, void* user )
{
    if ( !h || ( n && ( !pts || !out ) ) ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    if ( !report( cb, user, 0.f ) ) return canceled();
    if ( n == 0 ) return B2FWN_OK;
    // chunks of 4 Mi points in two device slots: the D2H of chunk k runs on the copy stream while chunk k+1 is uploaded and
    // walked.  With pageable host memory (std::vector of the adapter) both copies are staged by the driver and block the
    // calling thread, so only pinned memory overlaps fully; the walk itself is 6.6 ms per 10 M points (C2) either way.
    const size_t chunk = (size_t)4 << 20;
    const size_t count = ( n + chunk - 1 ) / chunk;
    for ( int s = 0; s < 2; ++s )
        if ( int e = ensureScratch( h, s, std::min( n, chunk ) * 4 ) ) return e; // 3 floats in + 1 out per point
    B2_CUDA( cudaEventRecord( h->evA, h->stream ) );
    for ( size_t k = 0; k < count; ++k )
    {
        const int s = (int)( k & 1 );
        const size_t o = k * chunk, m = std::min( chunk, n - o );
        float* dIn = h->dScratch[s];
        float* dOut = h->dScratch[s] + 3 * std::min( n, chunk );
        if ( k >= 2 ) B2_CUDA( cudaStreamWaitEvent( h->stream, h->partDone[s], 0 ) ); // the slot's previous result has left
        B2_CUDA( cudaMemcpyAsync( dIn, pts + 3 * o, m * 3 * sizeof( float ), cudaMemcpyHostToDevice, h->stream ) );
        if ( int e = launchFwnVector( h->tree, dIn, m, beta, skipFace, dOut, h->stream, &h->stats.kernel_launches ) ) return e;
        B2_CUDA( cudaEventRecord( h->computeDone[s], h->stream ) );
        B2_CUDA( cudaStreamWaitEvent( h->copyStream, h->computeDone[s], 0 ) );
        B2_CUDA( cudaMemcpyAsync( out + o, dOut, m * sizeof( float ), cudaMemcpyDeviceToHost, h->copyStream ) );
        B2_CUDA( cudaEventRecord( h->partDone[s], h->copyStream ) );
        if ( cb )
        {
            B2_CUDA( cudaEventSynchronize( h->computeDone[s] ) );
            if ( !report( cb, user, float( k + 1 ) / float( count ) ) ) { cudaDeviceSynchronize(); return canceled(); }
        }
    }
    B2_CUDA( cudaEventRecord( h->evB, h->stream ) );
    B2_CUDA( cudaStreamSynchronize( h->stream ) );
    B2_CUDA( cudaStreamSynchronize( h->copyStream ) );
    {
        float ms = 0; // first upload to last kernel: includes the H2D copies between the kernels
        if ( cudaEventElapsedTime( &ms, h->evA, h->evB ) == cudaSuccess ) h->stats.last_kernel_ms = ms;
    }
    return B2FWN_OK;
}

// ---- self-intersections ----------------------------------------------------------------------------------------------

int b2fwn_calc_self_intersections_device( b2fwn_handle h, float beta, uint64_t* dBits, float* dW )
{
    if ( !h ) { setError( "null handle" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    KernelTimer tm( h );
    if ( int e = launchFwnFaces( h->tree, beta, dW, dBits, h->stream, &h->stats.kernel_launches ) ) return e;
    tm.stop();
    B2_CUDA( cudaStreamSynchronize( h->stream ) );
    return B2FWN_OK;
}

int b2fwn_calc_self_intersections( b2fwn_handle h, float beta, uint64_t* outBits, b2fwn_progress cb, void* user )
{
    if ( !h ) { setError( "null handle" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    if ( !report( cb, user, 0.f ) ) return canceled();
    const size_t words = ( h->tree.numFaces + 63 ) / 64;
    if ( words == 0 ) return B2FWN_OK;
    if ( !outBits ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    if ( int e = ensureScratch( h, 0, words * 2 ) ) return e;
    uint64_t* dBits = reinterpret_cast<uint64_t*>( h->dScratch[0] );
    if ( int e = b2fwn_calc_self_intersections_device( h, beta, dBits, nullptr ) ) return e;
    B2_CUDA( cudaMemcpy( outBits, dBits, words * sizeof( uint64_t ), cudaMemcpyDeviceToHost ) );
    if ( !report( cb, user, 1.f ) ) return canceled();
    return B2FWN_OK;
}

// ---- closest points with face and barycentrics (IPointsToMeshProjector) ---------------------------------------------------

int b2fwn_find_projections_device( b2fwn_handle h, const float* dPts, size_t n, const float pointXf[12], const float treeXf[12],
                                   float upDistLimitSq, float loDistLimitSq, b2fwn_projection* dOut )
{
    if ( !h || ( n && ( !dPts || !dOut ) ) ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    KernelTimer tm( h );
    if ( int e = launchFindProjections( h->tree, dPts, n, pointXf, treeXf, upDistLimitSq, loDistLimitSq, dOut, h->stream,
                                        &h->stats.kernel_launches ) ) return e;
    tm.stop();
    B2_CUDA( cudaStreamSynchronize( h->stream ) );
    return B2FWN_OK;
}

int b2fwn_find_projections( b2fwn_handle h, const float* pts, size_t n, const float pointXf[12], const float treeXf[12],
                            float upDistLimitSq, float loDistLimitSq, b2fwn_projection* out, b2fwn_progress cb, void* user )
{
    if ( !h || ( n && ( !pts || !out ) ) ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    if ( !report( cb, user, 0.f ) ) return canceled();
    if ( n == 0 ) return B2FWN_OK;
    // chunks of 4 Mi points (3 floats in, 8 floats out per point), the reference's splitByChunks loop (MRCudaPointsToMeshProjector.cpp:105-113)
    const size_t chunk = std::min<size_t>( n, (size_t)4 << 20 );
    const size_t inFloats = ( 3 * chunk + 3 ) / 4 * 4; // records start 16-B aligned (two float4 stores each)
    if ( int e = ensureScratch( h, 0, inFloats + 8 * chunk ) ) return e;
    float* dIn = h->dScratch[0];
    b2fwn_projection* dOut = reinterpret_cast<b2fwn_projection*>( h->dScratch[0] + inFloats );
    double kernelMs = 0;
    for ( size_t o = 0; o < n; o += chunk )
    {
        const size_t m = std::min( chunk, n - o );
        B2_CUDA( cudaMemcpyAsync( dIn, pts + 3 * o, m * 3 * sizeof( float ), cudaMemcpyHostToDevice, h->stream ) );
        cudaEventRecord( h->evA, h->stream );
        if ( int e = launchFindProjections( h->tree, dIn, m, pointXf, treeXf, upDistLimitSq, loDistLimitSq, dOut, h->stream,
                                            &h->stats.kernel_launches ) ) return e;
        cudaEventRecord( h->evB, h->stream );
        B2_CUDA( cudaMemcpyAsync( out + o, dOut, m * sizeof( b2fwn_projection ), cudaMemcpyDeviceToHost, h->stream ) );
        B2_CUDA( cudaStreamSynchronize( h->stream ) );
        float ms = 0;
        if ( cudaEventElapsedTime( &ms, h->evA, h->evB ) == cudaSuccess ) kernelMs += ms;
        if ( !report( cb, user, float( o + m ) / float( n ) ) ) return canceled();
    }
    h->stats.last_kernel_ms = kernelMs;
    return B2FWN_OK;
}

// ---- grids -------------------------------------------------------------------------------------------------------------

int b2fwn_calc_from_grid_device( b2fwn_handle h, const int32_t dims[3], const float xf[12], float beta, int32_t zBegin,
                                 int32_t zEnd, float* dOut )
{
    if ( !h || !dims || !xf ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    if ( !dOut && dims[0] > 0 && dims[1] > 0 && zEnd > zBegin ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    KernelTimer tm( h );
    if ( int e = launchFwnGrid( h->tree, dims, toXf( xf ), beta, zBegin, zEnd, dOut, h->stream, &h->stats.kernel_launches ) ) return e;
    tm.stop();
    B2_CUDA( cudaStreamSynchronize( h->stream ) );
    return B2FWN_OK;
}

int b2fwn_calc_from_grid_with_distances_device( b2fwn_handle h, const int32_t dims[3], const float xf[12],
                                                const b2fwn_dist_options* opt, int32_t zBegin, int32_t zEnd, float* dOut )
{
    if ( !h || !dims || !xf || !opt ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    if ( !dOut && dims[0] > 0 && dims[1] > 0 && zEnd > zBegin ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    KernelTimer tm( h );
    if ( int e = launchSdfGrid( h->tree, dims, toXf( xf ), *opt, zBegin, zEnd, dOut, h->stream, &h->stats.kernel_launches ) ) return e;
    tm.stop();
    B2_CUDA( cudaStreamSynchronize( h->stream ) );
    return B2FWN_OK;
}

int b2fwn_calc_from_grid_with_distances_device_chunks( b2fwn_handle h, const int32_t dims[3], const float xf[12],
                                                       const b2fwn_dist_options* opt, int32_t zBegin, int32_t zEnd,
                                                       int32_t zChunk, int32_t zStride, float* dVolumeAtZBegin )
{
    if ( !h || !dims || !xf || !opt ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    const ZChunks zc{ zChunk, zStride };
    KernelTimer tm( h );
    if ( int e = launchSdfGrid( h->tree, dims, toXf( xf ), *opt, zBegin, zEnd, dVolumeAtZBegin, h->stream, &h->stats.kernel_launches, &zc ) )
        return e;
    tm.stop();
    B2_CUDA( cudaStreamSynchronize( h->stream ) );
    return B2FWN_OK;
}

int b2fwn_calc_from_grid_with_distances_device_chunks_p2p( b2fwn_handle h, const int32_t dims[3], const float xf[12],
                                                           const b2fwn_dist_options* opt, int32_t zBegin, int32_t zEnd,
                                                           int32_t zChunk, int32_t zStride, float* dVolumeAtZBegin,
                                                           int32_t nPeers, float* const* dPeerVolumesAtZBegin )
{
    if ( !h || !dims || !xf || !opt || ( nPeers > 0 && !dPeerVolumesAtZBegin ) ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    if ( nPeers < 0 || nPeers > MAX_PEERS ) { setError( "too many peer volumes" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    const ZChunks zc{ zChunk, zStride };
    PeerOut po{};
    po.n = nPeers;
    for ( int k = 0; k < nPeers; ++k )
    {
        if ( !dPeerVolumesAtZBegin[k] ) { setError( "null peer volume" ); return B2FWN_ERR_INVALID; }
        po.p[k] = dPeerVolumesAtZBegin[k];
    }
    KernelTimer tm( h );
    if ( int e = launchSdfGrid( h->tree, dims, toXf( xf ), *opt, zBegin, zEnd, dVolumeAtZBegin, h->stream, &h->stats.kernel_launches, &zc, &po ) )
        return e;
    tm.stop();
    B2_CUDA( cudaStreamSynchronize( h->stream ) );
    return B2FWN_OK;
}

static_assert( sizeof( cudaIpcMemHandle_t ) == B2FWN_IPC_TOKEN_BYTES, "IPC token size" );

int b2fwn_shared_alloc( b2fwn_handle h, size_t bytes, void** dPtr, unsigned char tokenOut[B2FWN_IPC_TOKEN_BYTES] )
{
    if ( !h || !dPtr || !tokenOut ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    *dPtr = nullptr;
    B2_CUDA( cudaMalloc( dPtr, bytes ? bytes : 1 ) );
    cudaIpcMemHandle_t ipc;
    const cudaError_t e = cudaIpcGetMemHandle( &ipc, *dPtr );
    if ( e != cudaSuccess ) { cudaFree( *dPtr ); *dPtr = nullptr; return cudaFail( e, "cudaIpcGetMemHandle" ); }
    memcpy( tokenOut, &ipc, sizeof( ipc ) );
    return B2FWN_OK;
}

int b2fwn_shared_open( b2fwn_handle h, const unsigned char token[B2FWN_IPC_TOKEN_BYTES], void** dPeerPtr )
{
    if ( !h || !token || !dPeerPtr ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    cudaIpcMemHandle_t ipc;
    memcpy( &ipc, token, sizeof( ipc ) );
    B2_CUDA( cudaIpcOpenMemHandle( dPeerPtr, ipc, cudaIpcMemLazyEnablePeerAccess ) );
    return B2FWN_OK;
}

int b2fwn_shared_close( b2fwn_handle h, void* dPeerPtr )
{
    if ( !h ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    if ( dPeerPtr ) B2_CUDA( cudaIpcCloseMemHandle( dPeerPtr ) );
    return B2FWN_OK;
}

int b2fwn_shared_free( b2fwn_handle h, void* dPtr )
{
    if ( !h ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    if ( dPtr ) B2_CUDA( cudaFree( dPtr ) );
    return B2FWN_OK;
}

int b2fwn_calc_from_grid_by_parts_into( b2fwn_handle h, const int32_t dims[3], const float xf[12], float beta, int32_t overlap,
                                        int32_t maxLayers, b2fwn_part_buffer_cb buffer, b2fwn_part_cb part, void* partUser,
                                        b2fwn_progress cb, void* user )
{
    if ( !h || !dims || !xf || !part ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    if ( dims[0] < 0 || dims[1] < 0 || dims[2] < 0 ) { setError( "invalid grid dimensions" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    const GridXf g = toXf( xf );
    return runByParts( h, dims, overlap, maxLayers, buffer, part, partUser, cb, user, [&]( b2fwn_ctx* dev, int z0, int z1, float* d, cudaStream_t st )
        { return launchFwnGrid( dev->tree, dims, g, beta, z0, z1, d, st, &dev->stats.kernel_launches ); } );
}

int b2fwn_calc_from_grid_by_parts( b2fwn_handle h, const int32_t dims[3], const float xf[12], float beta, int32_t overlap,
                                   int32_t maxLayers, b2fwn_part_cb part, void* partUser, b2fwn_progress cb, void* user )
{
    return b2fwn_calc_from_grid_by_parts_into( h, dims, xf, beta, overlap, maxLayers, nullptr, part, partUser, cb, user );
}

int b2fwn_calc_from_grid_with_distances_by_parts_into( b2fwn_handle h, const int32_t dims[3], const float xf[12],
                                                       const b2fwn_dist_options* opt, int32_t overlap, int32_t maxLayers,
                                                       b2fwn_part_buffer_cb buffer, b2fwn_part_cb part, void* partUser,
                                                       b2fwn_progress cb, void* user )
{
    if ( !h || !dims || !xf || !opt || !part ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    if ( dims[0] < 0 || dims[1] < 0 || dims[2] < 0 ) { setError( "invalid grid dimensions" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    const GridXf g = toXf( xf );
    const b2fwn_dist_options o = *opt;
    return runByParts( h, dims, overlap, maxLayers, buffer, part, partUser, cb, user, [&]( b2fwn_ctx* dev, int z0, int z1, float* d, cudaStream_t st )
        { return launchSdfGrid( dev->tree, dims, g, o, z0, z1, d, st, &dev->stats.kernel_launches ); } );
}

int b2fwn_calc_from_grid_with_distances_by_parts( b2fwn_handle h, const int32_t dims[3], const float xf[12],
                                                  const b2fwn_dist_options* opt, int32_t overlap, int32_t maxLayers,
                                                  b2fwn_part_cb part, void* partUser, b2fwn_progress cb, void* user )
{
    return b2fwn_calc_from_grid_with_distances_by_parts_into( h, dims, xf, opt, overlap, maxLayers, nullptr, part, partUser, cb, user );
}

int b2fwn_calc_from_grid( b2fwn_handle h, const int32_t dims[3], const float xf[12], float beta, float* out,
                          b2fwn_progress cb, void* user )
{
    if ( !h || !dims || !xf ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    if ( dims[0] < 0 || dims[1] < 0 || dims[2] < 0 ) { setError( "invalid grid dimensions" ); return B2FWN_ERR_INVALID; }
    if ( (size_t)dims[0] * dims[1] * dims[2] != 0 && !out ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    const GridXf g = toXf( xf );
    return runToHost( h, dims, 0, dims[2], out, cb, user, [&]( b2fwn_ctx* dev, int z0, int z1, float* d, cudaStream_t st )
        { return launchFwnGrid( dev->tree, dims, g, beta, z0, z1, d, st, &dev->stats.kernel_launches ); } );
}

int b2fwn_calc_from_grid_with_distances_layers( b2fwn_handle h, const int32_t dims[3], const float xf[12],
                                                const b2fwn_dist_options* opt, int32_t zBegin, int32_t zEnd, float* out,
                                                b2fwn_progress cb, void* user )
{
    if ( !h || !dims || !xf || !opt ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    if ( dims[0] < 0 || dims[1] < 0 || dims[2] < 0 ) { setError( "invalid grid dimensions" ); return B2FWN_ERR_INVALID; }
    if ( zBegin < 0 || zEnd > dims[2] || zBegin > zEnd ) { setError( "invalid layer range" ); return B2FWN_ERR_INVALID; }
    if ( (size_t)dims[0] * dims[1] * ( zEnd - zBegin ) != 0 && !out ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    const GridXf g = toXf( xf );
    const b2fwn_dist_options o = *opt;
    return runToHost( h, dims, zBegin, zEnd, out, cb, user, [&]( b2fwn_ctx* dev, int z0, int z1, float* d, cudaStream_t st )
        { return launchSdfGrid( dev->tree, dims, g, o, z0, z1, d, st, &dev->stats.kernel_launches ); } );
}

int b2fwn_calc_from_grid_with_distances( b2fwn_handle h, const int32_t dims[3], const float xf[12],
                                         const b2fwn_dist_options* opt, float* out, b2fwn_progress cb, void* user )
{
    if ( !dims ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    return b2fwn_calc_from_grid_with_distances_layers( h, dims, xf, opt, 0, dims[2], out, cb, user );
}

// ---- makeSignedByWindingNumber: sign of a resident / streamed unsigned distance volume -----------------------------------

int b2fwn_make_signed_by_winding_number_device( b2fwn_handle h, const int32_t dims[3], const float xf[12], float beta,
                                                float threshold, int32_t zBegin, int32_t zEnd, float* dVolumeAtZBegin )
{
    if ( !h || !dims || !xf ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    if ( dims[0] < 0 || dims[1] < 0 || dims[2] < 0 || zBegin < 0 || zEnd > dims[2] || zBegin > zEnd )
    {
        setError( "invalid grid dimensions" );
        return B2FWN_ERR_INVALID;
    }
    if ( (size_t)dims[0] * dims[1] * ( zEnd - zBegin ) != 0 && !dVolumeAtZBegin ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    KernelTimer tm( h );
    if ( int e = launchSignGrid( h->tree, dims, toXf( xf ), beta, threshold, zBegin, zEnd, dVolumeAtZBegin, h->stream,
                                 &h->stats.kernel_launches ) )
        return e;
    tm.stop();
    B2_CUDA( cudaStreamSynchronize( h->stream ) );
    return B2FWN_OK;
}

int b2fwn_make_signed_by_winding_number( b2fwn_handle h, const int32_t dims[3], const float xf[12], float beta, float threshold,
                                         float* volume, b2fwn_progress cb, void* user )
{
    if ( !h || !dims || !xf ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    if ( dims[0] < 0 || dims[1] < 0 || dims[2] < 0 ) { setError( "invalid grid dimensions" ); return B2FWN_ERR_INVALID; }
    if ( (size_t)dims[0] * dims[1] * dims[2] != 0 && !volume ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    const GridXf g = toXf( xf );
    const size_t layer = (size_t)dims[0] * dims[1];
    // chunk k: H2D on its compute stream, walk, D2H (runChunked) back to where it came from; the slot is free again by then
    return runToHost( h, dims, 0, dims[2], volume, cb, user, [&]( b2fwn_ctx* dev, int z0, int z1, float* d, cudaStream_t st )
    {
        B2_CUDA( cudaMemcpyAsync( d, volume + (size_t)z0 * layer, (size_t)( z1 - z0 ) * layer * sizeof( float ), cudaMemcpyHostToDevice, st ) );
        // d holds the layers z0..z1 (its first float is voxel (0,0,z0)), like every slab the grid launchers write
        return launchSignGrid( dev->tree, dims, g, beta, threshold, z0, z1, d, st, &dev->stats.kernel_launches );
    } );
}

// ---- callers moved onto the device: meshToDistanceVolume, mcOffsetMesh --------------------------------------------------

namespace
{
__device__ __forceinline__ unsigned ordOf( float f )
{
    const unsigned u = __float_as_uint( f );
    return ( u & 0x80000000u ) ? ~u : ( u | 0x80000000u );
}
// parallelMinMax (source/MRMesh/MRParallelMinMax.h:14-42): NaN never passes `v < min` / `v > max`, so it is ignored
__global__ void __launch_bounds__( 256 ) kMinMax( const float* __restrict__ v, size_t n, unsigned* __restrict__ mm )
{
    unsigned lo = 0xFFFFFFFFu, hi = 0u;
    for ( size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x )
    {
        const float f = v[i];
        if ( f == f ) { const unsigned o = ordOf( f ); lo = min( lo, o ); hi = max( hi, o ); }
    }
    lo = __reduce_min_sync( 0xFFFFFFFFu, lo );
    hi = __reduce_max_sync( 0xFFFFFFFFu, hi );
    if ( ( threadIdx.x & 31 ) == 0 ) { atomicMin( mm, lo ); atomicMax( mm + 1, hi ); }
}
float fromOrd( unsigned o )
{
    const unsigned u = ( o & 0x80000000u ) ? ( o & 0x7fffffffu ) : ~o;
    float f;
    memcpy( &f, &u, 4 );
    return f;
}
} // namespace

int b2fwn_suggest_voxel_size( b2fwn_handle h, float approxNumVoxels, float* out )
{
    if ( !h || !out ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    if ( h->tree.numLeaves == 0 ) { setError( "mesh has no valid faces" ); return B2FWN_ERR_INVALID; }
    // Box::volume (MRBox.h:112-119): product of the float extents, left to right; then std::cbrt of the float quotient (MROffset.cpp:28-30)
    volatile float vol = 1.f;
    for ( int c = 0; c < 3; ++c )
    {
        volatile float ext = h->tree.bbox[3 + c] - h->tree.bbox[c];
        vol = vol * ext;
    }
    volatile float q = vol / approxNumVoxels;
    *out = cbrtf( q );
    return B2FWN_OK;
}

int b2fwn_get_bounding_box( b2fwn_handle h, float out[6] )
{
    if ( !h || !out ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    if ( h->tree.numLeaves == 0 ) { setError( "mesh has no valid faces" ); return B2FWN_ERR_INVALID; }
    for ( int c = 0; c < 6; ++c ) out[c] = h->tree.bbox[c];
    return B2FWN_OK;
}

int b2fwn_mesh_to_distance_volume( b2fwn_handle h, const b2fwn_volume_params* vol, const b2fwn_dist_options* opt, float* out,
                                   float* outMin, float* outMax, b2fwn_progress cb, void* user )
{
    if ( !h || !vol || !opt ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    // basis = ( scale( voxelSize ), origin + 0.5f * voxelSize )   (MRMeshToDistanceVolume.cpp:55)
    float xf[12] = { vol->voxel_size[0], 0, 0, 0, vol->voxel_size[1], 0, 0, 0, vol->voxel_size[2], 0, 0, 0 };
    for ( int c = 0; c < 3; ++c )
    {
        volatile float half = 0.5f * vol->voxel_size[c];
        xf[9 + c] = vol->origin[c] + half;
    }
    // running min / max per device (8 B each, part of the handle), combined on the host afterwards
    std::vector<b2fwn_ctx*> devs{ h };
    devs.insert( devs.end(), h->replicas.begin(), h->replicas.end() );
    const unsigned init[2] = { 0xFFFFFFFFu, 0u };
    for ( b2fwn_ctx* d : devs )
    {
        B2_CUDA( cudaSetDevice( d->device ) );
        if ( !d->dMinMax ) B2_CUDA( cudaMalloc( &d->dMinMax, sizeof( init ) ) );
        B2_CUDA( cudaMemcpy( d->dMinMax, init, sizeof( init ), cudaMemcpyHostToDevice ) );
    }
    B2_CUDA( cudaSetDevice( h->device ) );
    const GridXf g = toXf( xf );
    const b2fwn_dist_options o = *opt;
    const int32_t* dims = vol->dims;
    int rc = B2FWN_ERR_INVALID;
    if ( dims[0] < 0 || dims[1] < 0 || dims[2] < 0 ) setError( "invalid grid dimensions" );
    else if ( (size_t)dims[0] * dims[1] * dims[2] != 0 && !out ) setError( "null argument" );
    else
        rc = runToHost( h, dims, 0, dims[2], out, cb, user, [&]( b2fwn_ctx* dev, int z0, int z1, float* d, cudaStream_t st )
        {
            if ( int e = launchSdfGrid( dev->tree, dims, g, o, z0, z1, d, st, &dev->stats.kernel_launches ) ) return e;
            const size_t n = (size_t)dims[0] * dims[1] * ( z1 - z0 );
            kMinMax<<<(unsigned)std::min<size_t>( ( n + 255 ) / 256, 148 * 8 ), 256, 0, st>>>( d, n, dev->dMinMax );
            ++dev->stats.kernel_launches;
            return B2FWN_OK;
        } );
    if ( rc ) return rc;
    unsigned mm[2] = { 0xFFFFFFFFu, 0u };
    for ( b2fwn_ctx* d : devs )
    {
        unsigned one[2];
        B2_CUDA( cudaSetDevice( d->device ) );
        B2_CUDA( cudaMemcpy( one, d->dMinMax, sizeof( one ), cudaMemcpyDeviceToHost ) );
        mm[0] = std::min( mm[0], one[0] ); mm[1] = std::max( mm[1], one[1] );
    }
    B2_CUDA( cudaSetDevice( h->device ) );
    if ( outMin ) *outMin = mm[0] == 0xFFFFFFFFu ? FLT_MAX : fromOrd( mm[0] );
    if ( outMax ) *outMax = mm[1] == 0u ? -FLT_MAX : fromOrd( mm[1] );
    return B2FWN_OK;
}

// ---- SignDetectionMode::ProjectionNormal: signedDistanceToMesh per point / per voxel centre -------------------------------

namespace
{
// one batch of m points already on the device: scratch slot 0 = [ points 3m (padded) | records 8m | distances m ]
struct ProjNormalBatch
{
    float* dPts; b2fwn_projection* dProj; float* dDist;
};
int projNormalBatch( b2fwn_ctx* h, size_t m, ProjNormalBatch& b )
{
    const size_t inFloats = ( 3 * m + 3 ) / 4 * 4; // records start 16-B aligned
    if ( int e = ensureScratch( h, 0, inFloats + 8 * m + m ) ) return e;
    b.dPts = h->dScratch[0];
    b.dProj = reinterpret_cast<b2fwn_projection*>( h->dScratch[0] + inFloats );
    b.dDist = h->dScratch[0] + inFloats + 8 * m;
    return B2FWN_OK;
}
} // namespace

int b2fwn_signed_distances_by_projection_normal_device( b2fwn_handle h, const float* dPts, size_t n, const b2fwn_dist_options* opt,
                                                        float* dOut )
{
    if ( !h || !opt || ( n && ( !dPts || !dOut ) ) ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    if ( n == 0 ) return B2FWN_OK;
    b2fwn_projection* dProj = nullptr;
    B2_CUDA( cudaMallocFromPoolAsync( &dProj, n * sizeof( b2fwn_projection ), h->tree.pool, h->stream ) );
    KernelTimer tm( h );
    const int e = launchSignedDistProjNormal( h->tree, dPts, n, opt->min_dist_sq, opt->max_dist_sq, opt->null_outside_min_max, dProj, dOut,
                                              h->stream, &h->stats.kernel_launches );
    if ( !e ) tm.stop();
    cudaFreeAsync( dProj, h->stream );
    B2_CUDA( cudaStreamSynchronize( h->stream ) );
    return e;
}

int b2fwn_signed_distances_by_projection_normal( b2fwn_handle h, const float* pts, size_t n, const b2fwn_dist_options* opt, float* out,
                                                 b2fwn_progress cb, void* user )
{
    if ( !h || !opt || ( n && ( !pts || !out ) ) ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    if ( !report( cb, user, 0.f ) ) return canceled();
    if ( n == 0 ) return B2FWN_OK;
    const size_t chunk = std::min<size_t>( n, (size_t)4 << 20 );
    ProjNormalBatch b{};
    if ( int e = projNormalBatch( h, chunk, b ) ) return e;
    double kernelMs = 0;
    for ( size_t o = 0; o < n; o += chunk )
    {
        const size_t m = std::min( chunk, n - o );
        B2_CUDA( cudaMemcpyAsync( b.dPts, pts + 3 * o, m * 3 * sizeof( float ), cudaMemcpyHostToDevice, h->stream ) );
        cudaEventRecord( h->evA, h->stream );
        if ( int e = launchSignedDistProjNormal( h->tree, b.dPts, m, opt->min_dist_sq, opt->max_dist_sq, opt->null_outside_min_max, b.dProj,
                                                 b.dDist, h->stream, &h->stats.kernel_launches ) ) return e;
        cudaEventRecord( h->evB, h->stream );
        B2_CUDA( cudaMemcpyAsync( out + o, b.dDist, m * sizeof( float ), cudaMemcpyDeviceToHost, h->stream ) );
        B2_CUDA( cudaStreamSynchronize( h->stream ) );
        float ms = 0;
        if ( cudaEventElapsedTime( &ms, h->evA, h->evB ) == cudaSuccess ) kernelMs += ms;
        if ( !report( cb, user, float( o + m ) / float( n ) ) ) return canceled();
    }
    h->stats.last_kernel_ms = kernelMs;
    return B2FWN_OK;
}

int b2fwn_mesh_to_distance_volume_by_projection_normal( b2fwn_handle h, const b2fwn_volume_params* vol, const b2fwn_dist_options* opt,
                                                        float* out, float* outMin, float* outMax, b2fwn_progress cb, void* user )
{
    if ( !h || !vol || !opt ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    const int32_t* dims = vol->dims;
    if ( dims[0] < 0 || dims[1] < 0 || dims[2] < 0 ) { setError( "invalid grid dimensions" ); return B2FWN_ERR_INVALID; }
    const size_t layer = (size_t)dims[0] * dims[1], total = layer * dims[2];
    if ( total != 0 && !out ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    if ( !report( cb, user, 0.f ) ) return canceled();
    unsigned mm[2] = { 0xFFFFFFFFu, 0u };
    if ( total != 0 )
    {
        // whole layers, about 8 Mi voxels per batch.  Without a lower limit the distances come from the brick walk of the grid
        // kernels (the closest triangle's slot kept per voxel, 8 B of scratch per voxel); with minDistSq > 0 the reference's
        // early exit makes the answer depend on its visiting order, which only the per-point search reproduces (44 B per voxel)
        // The brick walk is a dev switch (B2FWN_PROJ_BRICKS=1), not the default: (1) where several triangles are exactly equally
        // close it may keep another one than the reference's search, and at a vertex the reference's mesh builder duplicated
        // (non-manifold: the pole of a UV sphere with faces removed) the duplicates have different pseudonormals -- 1 voxel of
        // 34,891 got the other sign (profiles/r01/diag_projbricks_n2.txt); (2) it measured 2 x faster over the full range of a
        // sphere but 3.5 x slower in a narrow band than the Morton-ordered per-point search (r01).
        const char* envBricks = devSwitch( "B2FWN_PROJ_BRICKS" );
        const bool bricks = envBricks && atoi( envBricks ) && ( !opt->null_outside_min_max || opt->min_dist_sq <= 0.f );
        const int layers = (int)std::min<size_t>( std::max<size_t>( ( (size_t)8 << 20 ) / layer, 1 ), (size_t)dims[2] );
        if ( layer * layers > 0xFFFFFFFFull ) { setError( "grid layer too large" ); return B2FWN_ERR_INVALID; }
        ProjNormalBatch b{};
        if ( int e = projNormalBatch( h, layer * layers, b ) ) return e;
        int32_t* dSlot = reinterpret_cast<int32_t*>( b.dPts ); // brick path: no point list, the slots take its place
        unsigned* dMM = nullptr;
        B2_CUDA( cudaMalloc( &dMM, sizeof( mm ) ) );
        cudaMemcpyAsync( dMM, mm, sizeof( mm ), cudaMemcpyHostToDevice, h->stream );
        double kernelMs = 0;
        int rc = B2FWN_OK;
        for ( int z0 = 0; z0 < dims[2] && rc == B2FWN_OK; z0 += layers )
        {
            const int z1 = std::min( dims[2], z0 + layers );
            const size_t m = layer * ( z1 - z0 );
            cudaEventRecord( h->evA, h->stream );
            if ( bricks )
            {
                rc = launchDistSlotGrid( h->tree, dims, vol->origin, vol->voxel_size, opt->null_outside_min_max ? opt->max_dist_sq : FLT_MAX,
                                         opt->null_outside_min_max, z0, z1, b.dDist, dSlot, h->stream, &h->stats.kernel_launches );
                if ( !rc ) rc = launchSignBySlot( h->tree, dims, vol->origin, vol->voxel_size, z0, z1, b.dDist, dSlot, h->stream,
                                                  &h->stats.kernel_launches );
            }
            else
            {
                rc = launchVoxelCentres( vol->origin, vol->voxel_size, dims[0], dims[1], z0, z1, b.dPts, h->stream, &h->stats.kernel_launches );
                if ( !rc ) rc = launchSignedDistProjNormal( h->tree, b.dPts, m, opt->min_dist_sq, opt->max_dist_sq, opt->null_outside_min_max, b.dProj,
                                                            b.dDist, h->stream, &h->stats.kernel_launches );
            }
            if ( rc ) break;
            kMinMax<<<(unsigned)std::min<size_t>( ( m + 255 ) / 256, 148 * 8 ), 256, 0, h->stream>>>( b.dDist, m, dMM );
            ++h->stats.kernel_launches;
            cudaEventRecord( h->evB, h->stream );
            cudaMemcpyAsync( out + (size_t)z0 * layer, b.dDist, m * sizeof( float ), cudaMemcpyDeviceToHost, h->stream );
            if ( cudaStreamSynchronize( h->stream ) != cudaSuccess ) { rc = cudaFail( cudaGetLastError(), "projection-normal volume batch" ); break; }
            float ms = 0;
            if ( cudaEventElapsedTime( &ms, h->evA, h->evB ) == cudaSuccess ) kernelMs += ms;
            if ( !report( cb, user, float( z1 ) / float( dims[2] ) ) ) rc = canceled();
        }
        if ( rc == B2FWN_OK ) cudaMemcpy( mm, dMM, sizeof( mm ), cudaMemcpyDeviceToHost );
        cudaFree( dMM );
        if ( rc ) return rc;
        h->stats.last_kernel_ms = kernelMs;
    }
    if ( outMin ) *outMin = mm[0] == 0xFFFFFFFFu ? FLT_MAX : fromOrd( mm[0] );
    if ( outMax ) *outMax = mm[1] == 0u ? -FLT_MAX : fromOrd( mm[1] );
    return B2FWN_OK;
}

int b2fwn_mc_offset_mesh( b2fwn_handle h, const b2fwn_offset_params* p, b2fwn_offset_result* res, b2fwn_progress cb, void* user )
{
    if ( !h || !p || !res ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    if ( !( p->voxel_size > 0 ) ) { setError( "invalid voxelSize value" ); return B2FWN_ERR_INVALID; } // MROffset.cpp:122-126
    if ( h->tree.numLeaves == 0 ) { setError( "mesh has no valid faces" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    memset( res, 0, sizeof( *res ) );
    if ( !report( cb, user, 0.f ) ) return canceled();
    const float v = p->voxel_size;
    const float absOffset = fabsf( p->offset );
    // box = bbox.expanded( |offset| ); calcOriginAndDimensions (MRCalcDims.h:18-27): origin = box.min - 2v,
    // dims = int( ( box.max + 2v - origin ) * ( 1 / v ) ) + 1  (Vector3f / float multiplies by the reciprocal, MRVector3.h:120-126)
    volatile float expansion = 2 * v, inv = 1 / v;
    int32_t dims[3];
    float origin[3];
    for ( int c = 0; c < 3; ++c )
    {
        volatile float bmin = h->tree.bbox[c] - absOffset, bmax = h->tree.bbox[3 + c] + absOffset;
        volatile float o = bmin - expansion;
        volatile float top = bmax + expansion;
        volatile float ext = top - o;
        volatile float q = inv * ext;
        origin[c] = o;
        dims[c] = (int32_t)q + 1;
        res->origin[c] = o; res->dims[c] = dims[c];
    }
    b2fwn_dist_options o{};
    {
        // MROffset.cpp:176-183 (1.001f guards against rounding, as there)
        volatile float lo = absOffset - 1.001f * v, hi = absOffset + 1.001f * v;
        const float lo0 = lo > 0.f ? (float)lo : 0.f;
        o.min_dist_sq = lo0 * lo0;
        o.max_dist_sq = hi * hi;
        o.null_outside_min_max = p->close_holes ? 0 : 1; // !isHoleWindingRule || !closeHolesInHoleWindingNumber
        o.winding_number_threshold = p->winding_number_threshold;
        o.winding_number_beta = p->winding_number_beta;
    }
    float xf[12] = { v, 0, 0, 0, v, 0, 0, 0, v, 0, 0, 0 };
    for ( int c = 0; c < 3; ++c )
    {
        volatile float half = 0.5f * v;
        xf[9 + c] = origin[c] + half; // basis = ( scale( voxelSize ), origin + 0.5f * voxelSize ), MROffset.cpp:202
    }
    b2fwn_mc_params mp{};
    for ( int c = 0; c < 3; ++c ) { mp.origin[c] = origin[c]; mp.voxel_size[c] = v; }
    mp.iso = p->offset;
    mp.less_inside = 1;
    const GridXf gxf = toXf( xf );

    // Every device of the handle takes a contiguous z-range of the grid and works through it in slabs: the distance band of the
    // slab's layers plus the two after them into a buffer, marching cubes on the slab (b2fwn_mc.cu), on to the next.  One slab
    // is the whole range when it fits the safe share of the free device memory (the usual case on a B200: 2048^3 needs 45 GB);
    // otherwise, or when the caller caps it (max_layers_per_slab), as many layers as fit -- the reference's MarchingCubesByParts
    // (MRMarchingCubes.h:96-122, MROffset.cpp:205-221) without the volume ever leaving HBM.  Ids are local to a device until
    // every device has counted; then the exclusive prefix of the counts is added to its triangles.
    std::vector<b2fwn_ctx*> devs{ h };
    devs.insert( devs.end(), h->replicas.begin(), h->replicas.end() );
    const int nd = (int)std::min<size_t>( devs.size(), (size_t)std::max( 1, dims[2] / 4 ) );
    devs.resize( nd );
    const size_t layer = (size_t)dims[0] * dims[1];
    struct DevRun { int z0 = 0, z1 = 0; double sdfMs = 0, totalMs = 0; int rc = B2FWN_OK; std::string err; };
    std::vector<DevRun> run( nd );
    for ( int d = 0; d < nd; ++d ) { run[d].z0 = (int)( (int64_t)dims[2] * d / nd ); run[d].z1 = (int)( (int64_t)dims[2] * ( d + 1 ) / nd ); }
    const int cap = p->max_layers_per_slab;
    auto work = [&]( int d )
    {
        b2fwn_ctx* c = devs[d];
        DevRun& R = run[d];
        auto fail = [&]( int rc ) { R.rc = rc; R.err = b2fwn_last_error(); };
        if ( cudaSetDevice( c->device ) != cudaSuccess ) { fail( cudaFail( cudaGetLastError(), "cudaSetDevice" ) ); return; }
        // values 4 B x (n + 2) layers, codes 1 B x (n + 1) layers, segment totals and scan scratch ~ 0.3 B: 5.5 B per voxel of a slab
        freeMc( c->mc ); // the previous result goes before the budget is taken
        // slot 0 and the mesher's workspace are reused; slot 1 and idle pool blocks are of no use here and go when the range does not fit
        auto fitLayers = [&]() -> size_t
        {
            const size_t budget = safeMemoryLimit( c->device ) + reclaimableBytes( c, RECLAIM_SCRATCH0 | RECLAIM_MC_WORK );
            return layer ? (size_t)( (double)budget / ( 5.5 * (double)layer ) ) : (size_t)( R.z1 - R.z0 ) + 2;
        };
        int n = R.z1 - R.z0;
        size_t fit = fitLayers();
        if ( fit < (size_t)n + 2 && reclaimableBytes( c, RECLAIM_SCRATCH1 | RECLAIM_POOL ) > 0 )
        {
            reclaimCaches( c, RECLAIM_SCRATCH1 | RECLAIM_POOL );
            fit = fitLayers();
        }
        if ( fit < (size_t)n + 2 ) n = fit > 3 ? (int)fit - 2 : 1;
        if ( cap > 0 ) n = std::min( n, cap );
        const size_t bufLayers = (size_t)std::min( dims[2] - R.z0, n + 2 );
        if ( int e = ensureScratch( c, 0, std::max<size_t>( bufLayers * layer, 1 ) ) ) { fail( e ); return; }
        float* dBuf = c->dScratch[0]; // the handle's cached slot: no 34-GB cudaMalloc / cudaFree per call
        cudaEvent_t e0 = nullptr, e1 = nullptr, s0 = nullptr, s1 = nullptr;
        cudaEventCreate( &e0 ); cudaEventCreate( &e1 ); cudaEventCreate( &s0 ); cudaEventCreate( &s1 );
        cudaEventRecord( e0, c->stream );
        const int rc = marchingCubesSlabs( [&]( int z0, int z1, const float** dl ) -> int
        {
            cudaEventRecord( s0, c->stream );
            if ( int e = launchSdfGrid( c->tree, dims, gxf, o, z0, z1, dBuf, c->stream, &c->stats.kernel_launches ) ) return e;
            cudaEventRecord( s1, c->stream );
            if ( cudaEventSynchronize( s1 ) != cudaSuccess ) return cudaFail( cudaGetLastError(), "offset: distance slab" );
            float ms = 0;
            cudaEventElapsedTime( &ms, s0, s1 );
            R.sdfMs += ms;
            *dl = dBuf;
            return B2FWN_OK;
        }, dims, mp, n, R.z0, R.z1, 0, 0, c->mc, c->stream, &c->stats.kernel_launches, &c->mcWork );
        cudaEventRecord( e1, c->stream );
        cudaEventSynchronize( e1 );
        float ms = 0;
        cudaEventElapsedTime( &ms, e0, e1 );
        R.totalMs = ms;
        cudaEventDestroy( e0 ); cudaEventDestroy( e1 ); cudaEventDestroy( s0 ); cudaEventDestroy( s1 );
        if ( rc ) fail( rc );
    };
    {
        std::vector<std::thread> th;
        for ( int d = 1; d < nd; ++d ) th.emplace_back( work, d );
        work( 0 );
        for ( auto& t : th ) t.join();
    }
    B2_CUDA( cudaSetDevice( h->device ) );
    for ( int d = 0; d < nd; ++d )
        if ( run[d].rc ) { setError( run[d].err ); return run[d].rc; }
    if ( !report( cb, user, 0.9f ) ) return canceled();
    // global ids: every device's triangles get the number of vertices of the devices before it
    uint64_t baseV = 0, totV = 0, totT = 0;
    for ( int d = 0; d < nd; ++d ) { totV += devs[d]->mc.nVerts; totT += devs[d]->mc.nTris; }
    if ( totV > 0x7FFFFFFFull ) { setError( "more than 2^31 vertices" ); return B2FWN_ERR_INVALID; }
    for ( int d = 0; d < nd; ++d )
    {
        b2fwn_ctx* c = devs[d];
        if ( d > 0 && c->mc.nTris )
        {
            B2_CUDA( cudaSetDevice( c->device ) );
            if ( int e = marchingCubesAddBase( c->mc, (int32_t)baseV, c->stream, &c->stats.kernel_launches ) ) return e;
            B2_CUDA( cudaStreamSynchronize( c->stream ) );
        }
        baseV += c->mc.nVerts;
        res->sdf_ms = std::max( res->sdf_ms, run[d].sdfMs );
        res->mc_ms = std::max( res->mc_ms, run[d].totalMs - run[d].sdfMs );
    }
    for ( size_t d = (size_t)nd; d < 1 + h->replicas.size(); ++d ) freeMc( ( d == 0 ? h : h->replicas[d - 1] )->mc ); // devices without a share
    B2_CUDA( cudaSetDevice( h->device ) );
    h->stats.last_kernel_ms = res->sdf_ms + res->mc_ms;
    res->n_verts = totV; res->n_tris = totT;
    if ( !report( cb, user, 1.f ) ) return canceled();
    return B2FWN_OK;
}

// ---- diagnostics: FP32 FMA-pipe peak of this GPU (the roofline denominator of the traversal kernels) -------------------

namespace
{
__global__ void __launch_bounds__( 256 ) kFmaPeak( float* out, int iters, float a, float b )
{
    float x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for ( int i = 0; i < iters; ++i )
    {
#pragma unroll
        for ( int u = 0; u < 16; ++u )
        {
            x0 = fmaf( x0, a, b ); x1 = fmaf( x1, a, b ); x2 = fmaf( x2, a, b ); x3 = fmaf( x3, a, b );
            x4 = fmaf( x4, a, b ); x5 = fmaf( x5, a, b ); x6 = fmaf( x6, a, b ); x7 = fmaf( x7, a, b );
        }
    }
    const float r = ( ( x0 + x1 ) + ( x2 + x3 ) ) + ( ( x4 + x5 ) + ( x6 + x7 ) );
    if ( r == 12345.678f ) out[0] = r; // never true; keeps the chain alive
}
} // namespace

int b2fwn_measure_fp32_peak( int32_t device, double* tflops )
{
    if ( !tflops ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( device ) );
    cudaDeviceProp prop;
    B2_CUDA( cudaGetDeviceProperties( &prop, device ) );
    float* d = nullptr;
    B2_CUDA( cudaMalloc( &d, 4 ) );
    cudaEvent_t a, b;
    cudaEventCreate( &a ); cudaEventCreate( &b );
    const int blocks = prop.multiProcessorCount * 8, iters = 4096;
    double best = 0;
    for ( int rep = 0; rep < 4; ++rep )
    {
        cudaEventRecord( a );
        kFmaPeak<<<blocks, 256>>>( d, iters, 1.0000001f, 1e-9f );
        cudaEventRecord( b );
        if ( cudaEventSynchronize( b ) != cudaSuccess ) break;
        float ms = 0;
        cudaEventElapsedTime( &ms, a, b );
        const double fl = 2.0 * 8 * 16 * (double)iters * 256.0 * blocks;
        if ( rep > 0 && ms > 0 ) best = std::max( best, fl / ( ms * 1e-3 ) / 1e12 );
    }
    cudaEventDestroy( a ); cudaEventDestroy( b ); cudaFree( d );
    B2_CUDA( cudaGetLastError() );
    *tflops = best;
    return B2FWN_OK;
}

// ---- marching cubes ------------------------------------------------------------------------------------------------------

int b2fwn_marching_cubes( b2fwn_handle h, const float* volume, int32_t onDevice, const int32_t dims[3], const b2fwn_mc_params* p,
                          uint64_t* nVerts, uint64_t* nTris )
{
    if ( !h || !dims || !p || !nVerts || !nTris ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    if ( dims[0] < 0 || dims[1] < 0 || dims[2] < 0 ) { setError( "invalid grid dimensions" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    const size_t total = (size_t)dims[0] * dims[1] * dims[2];
    *nVerts = *nTris = 0;
    freeMc( h->mc );
    for ( b2fwn_ctx* r : h->replicas ) { cudaSetDevice( r->device ); freeMc( r->mc ); }
    cudaSetDevice( h->device );
    if ( total == 0 ) return B2FWN_OK;
    if ( !volume ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    const float* dVol = volume;
    float* dTmp = nullptr;
    if ( !onDevice )
    {
        B2_CUDA( cudaMalloc( &dTmp, total * sizeof( float ) ) );
        cudaError_t e = cudaMemcpyAsync( dTmp, volume, total * sizeof( float ), cudaMemcpyHostToDevice, h->stream );
        if ( e != cudaSuccess ) { cudaFree( dTmp ); return cudaFail( e, "H2D volume" ); }
        dVol = dTmp;
    }
    KernelTimer tm( h );
    // the volume is the caller's or dTmp: both scratch slots can go when the mesher's own buffers do not fit
    const int rc = retryAfterReclaim( h, RECLAIM_SCRATCH0 | RECLAIM_SCRATCH1 | RECLAIM_POOL, [&]
    { return marchingCubesDevice( dVol, dims, *p, h->mc, h->stream, &h->stats.kernel_launches, &h->mcWork ); } );
    tm.stop();
    cudaStreamSynchronize( h->stream );
    if ( dTmp ) cudaFree( dTmp );
    if ( rc ) return rc;
    *nVerts = h->mc.nVerts; *nTris = h->mc.nTris;
    return B2FWN_OK;
}

int b2fwn_marching_cubes_fetch( b2fwn_handle h, float* outVerts, int32_t* outTris )
{
    if ( !h ) { setError( "null handle" ); return B2FWN_ERR_INVALID; }
    // the result of a multi-device offset lies on the devices in z order: the primary's share first
    std::vector<b2fwn_ctx*> devs{ h };
    devs.insert( devs.end(), h->replicas.begin(), h->replicas.end() );
    size_t v0 = 0, t0 = 0;
    for ( b2fwn_ctx* c : devs )
    {
        B2_CUDA( cudaSetDevice( c->device ) );
        if ( c->mc.nVerts && outVerts )
            B2_CUDA( cudaMemcpy( outVerts + 3 * v0, c->mc.dVerts, c->mc.nVerts * 3 * sizeof( float ), cudaMemcpyDeviceToHost ) );
        if ( c->mc.nTris && outTris )
            B2_CUDA( cudaMemcpy( outTris + 3 * t0, c->mc.dTris, c->mc.nTris * 3 * sizeof( int32_t ), cudaMemcpyDeviceToHost ) );
        v0 += c->mc.nVerts; t0 += c->mc.nTris;
    }
    B2_CUDA( cudaSetDevice( h->device ) );
    return B2FWN_OK;
}

// ---- half-edge tables of a triangle mesh ---------------------------------------------------------------------------------

int b2fwn_build_topology( b2fwn_handle h, const int32_t* tris, size_t nTris, int32_t onDevice, size_t nVerts, uint64_t* nHalfEdges,
                          uint64_t* nVertsOut, uint64_t* nFacesOut )
{
    if ( !h || !nHalfEdges ) { setError( "null argument" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    *nHalfEdges = 0;
    const int32_t* dTris = tris;
    int32_t* dTmp = nullptr;
    if ( !tris )
    {
        dTris = h->mc.dTris; nTris = h->mc.nTris; nVerts = h->mc.nVerts; // the marching-cubes result of this handle
        size_t moreT = 0, moreV = 0;
        for ( const b2fwn_ctx* r : h->replicas ) { moreT += r->mc.nTris; moreV += r->mc.nVerts; }
        if ( moreT || moreV )
        {
            // a multi-device offset: the other devices' triangles come over (peer copies; staged by the driver without peer access)
            B2_CUDA( cudaMalloc( &dTmp, ( nTris + moreT ) * 3 * sizeof( int32_t ) ) );
            cudaError_t e = nTris ? cudaMemcpyAsync( dTmp, h->mc.dTris, nTris * 3 * sizeof( int32_t ), cudaMemcpyDeviceToDevice, h->stream ) : cudaSuccess;
            size_t at = nTris;
            for ( const b2fwn_ctx* r : h->replicas )
            {
                if ( e == cudaSuccess && r->mc.nTris )
                    e = cudaMemcpyPeerAsync( dTmp + 3 * at, h->device, r->mc.dTris, r->device, r->mc.nTris * 3 * sizeof( int32_t ), h->stream );
                at += r->mc.nTris;
            }
            if ( e != cudaSuccess ) { cudaFree( dTmp ); return cudaFail( e, "gathering the triangles" ); }
            dTris = dTmp; nTris += moreT; nVerts += moreV;
        }
    }
    else if ( !onDevice && nTris )
    {
        B2_CUDA( cudaMalloc( &dTmp, nTris * 3 * sizeof( int32_t ) ) );
        const cudaError_t e = cudaMemcpyAsync( dTmp, tris, nTris * 3 * sizeof( int32_t ), cudaMemcpyHostToDevice, h->stream );
        if ( e != cudaSuccess ) { cudaFree( dTmp ); return cudaFail( e, "H2D triangles" ); }
        dTris = dTmp;
    }
    KernelTimer tm( h );
    // the triangles are the mesher's result or dTmp: every cache of the handle can go when the sort's buffers do not fit
    const int rc = retryAfterReclaim( h, RECLAIM_SCRATCH0 | RECLAIM_SCRATCH1 | RECLAIM_MC_WORK | RECLAIM_POOL, [&]
    { return buildTopologyDevice( dTris, nTris, nVerts, h->topo, h->stream, &h->stats.kernel_launches ); } );
    tm.stop();
    cudaStreamSynchronize( h->stream );
    if ( dTmp ) cudaFree( dTmp );
    if ( rc ) return rc;
    *nHalfEdges = h->topo.nHalfEdges;
    if ( nVertsOut ) *nVertsOut = h->topo.nVerts;
    if ( nFacesOut ) *nFacesOut = h->topo.nFaces;
    return B2FWN_OK;
}

int b2fwn_build_topology_fetch( b2fwn_handle h, b2fwn_half_edge* outEdges, int32_t* outEdgePerVertex, int32_t* outEdgePerFace )
{
    if ( !h ) { setError( "null handle" ); return B2FWN_ERR_INVALID; }
    B2_CUDA( cudaSetDevice( h->device ) );
    const TopoResult& t = h->topo;
    if ( t.nHalfEdges && outEdges ) B2_CUDA( cudaMemcpy( outEdges, t.dEdges, t.nHalfEdges * sizeof( b2fwn_half_edge ), cudaMemcpyDeviceToHost ) );
    if ( t.nVerts && t.dEdgePerVertex && outEdgePerVertex ) B2_CUDA( cudaMemcpy( outEdgePerVertex, t.dEdgePerVertex, t.nVerts * 4, cudaMemcpyDeviceToHost ) );
    if ( t.nFaces && t.dEdgePerFace && outEdgePerFace ) B2_CUDA( cudaMemcpy( outEdgePerFace, t.dEdgePerFace, t.nFaces * 4, cudaMemcpyDeviceToHost ) );
    return B2FWN_OK;
}

} // extern "C"
#pragma GCC visibility pop
