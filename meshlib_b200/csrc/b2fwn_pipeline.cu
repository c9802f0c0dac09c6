// b2fwn_pipeline.cu -- host side of the host-output and by-parts grid calls: z-chunks dealt over the devices of the handle,
// kernels of the next chunk enqueued before the copy of the current one is driven, results delivered into the caller's memory.
//
// What it replaces: the reference's blocking cudaMemcpy per chunk + TBB ping-pong (source/MRCuda/MRCudaPipeline.hpp:9-48,
// MRCudaFastWindingNumber.cpp:285-429), its single device (MRCudaFastWindingNumber.cpp:48) and its memory budget
// (MRCudaBasic.cpp:75-107).
//
// Destination memory decides the copy path:
//  * pinned (cudaMallocHost / cudaHostRegister'ed by the caller, or the library's own part buffers): one cudaMemcpyAsync per
//    chunk on the device's copy stream, straight into the destination;
//  * pageable (std::vector<float> of a MeshLib caller, a numpy array): a cudaMemcpyAsync into pageable memory only returns
//    when the copy is complete and runs at a fraction of the link rate, so the chunk is cut into 32-MiB pieces that travel
//    through a ring of pinned staging buffers (per device, allocated once per process) and are moved into the destination by
//    a few copier threads while the next pieces are in flight and the next chunk's kernels run.
// No arithmetic lives here.
#include "b2fwn_ctx.h"
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <map>
#include <mutex>
#include <thread>

namespace b2
{

namespace
{
size_t poolIdleBytes( cudaMemPool_t pool )
{
    if ( !pool ) return 0;
    uint64_t reserved = 0, used = 0;
    if ( cudaMemPoolGetAttribute( pool, cudaMemPoolAttrReservedMemCurrent, &reserved ) != cudaSuccess
         || cudaMemPoolGetAttribute( pool, cudaMemPoolAttrUsedMemCurrent, &used ) != cudaSuccess )
    {
        (void)cudaGetLastError();
        return 0;
    }
    return reserved > used ? (size_t)( reserved - used ) : 0;
}
}

size_t reclaimableBytes( const b2fwn_ctx* h, unsigned what )
{
    size_t bytes = 0;
    if ( what & RECLAIM_SCRATCH0 ) bytes += h->scratchFloats[0] * sizeof( float );
    if ( what & RECLAIM_SCRATCH1 ) bytes += h->scratchFloats[1] * sizeof( float );
    if ( what & RECLAIM_MC_WORK ) bytes += h->mcWork.codeBytes + 2 * h->mcWork.segEntries * sizeof( uint32_t ) + h->mcWork.tmpBytes;
    if ( what & RECLAIM_POOL ) bytes += poolIdleBytes( h->tree.pool );
    return bytes;
}

void reclaimCaches( b2fwn_ctx* h, unsigned what )
{
    cudaDeviceSynchronize(); // nothing in flight still reads what goes
    for ( int s = 0; s < 2; ++s )
    {
        if ( ( what & ( s ? RECLAIM_SCRATCH1 : RECLAIM_SCRATCH0 ) ) && h->dScratch[s] )
        {
            cudaFree( h->dScratch[s] );
            h->dScratch[s] = nullptr; h->scratchFloats[s] = 0;
        }
        if ( ( what & RECLAIM_PINNED ) && h->hPinned[s] )
        {
            cudaFreeHost( h->hPinned[s] );
            h->hPinned[s] = nullptr; h->pinnedFloats[s] = 0;
        }
    }
    if ( what & RECLAIM_MC_WORK ) freeMcWorkspace( h->mcWork );
    if ( ( what & RECLAIM_POOL ) && h->tree.pool ) cudaMemPoolTrimTo( h->tree.pool, 0 );
    (void)cudaGetLastError();
}

int ensureScratch( b2fwn_ctx* h, int slot, size_t floats )
{
    if ( h->scratchFloats[slot] >= floats ) return B2FWN_OK;
    if ( h->dScratch[slot] ) cudaFree( h->dScratch[slot] );
    h->dScratch[slot] = nullptr; h->scratchFloats[slot] = 0;
    // the other slot may hold the running call's data; the mesher never runs while a slot is being sized
    return retryAfterReclaim( h, RECLAIM_MC_WORK | RECLAIM_POOL, [&]() -> int
    {
        B2_CUDA( cudaMalloc( &h->dScratch[slot], floats * sizeof( float ) ) );
        h->scratchFloats[slot] = floats;
        return B2FWN_OK;
    } );
}

int ensurePinned( b2fwn_ctx* h, int slot, size_t floats )
{
    if ( h->pinnedFloats[slot] >= floats ) return B2FWN_OK;
    if ( h->hPinned[slot] ) cudaFreeHost( h->hPinned[slot] );
    h->hPinned[slot] = nullptr; h->pinnedFloats[slot] = 0;
    B2_CUDA( cudaMallocHost( &h->hPinned[slot], floats * sizeof( float ) ) );
    h->pinnedFloats[slot] = floats;
    return B2FWN_OK;
}

// splitByChunks (source/MRMesh/MRChunkIterator.cpp:25-49) over whole layers: part k covers [k*(part-overlap), +part), the last one shorter
int partCount( int nLayers, int layersPerPart, int overlap )
{
    if ( nLayers <= 0 || layersPerPart <= overlap ) return 0;
    const int step = layersPerPart - overlap;
    return nLayers <= overlap ? 1 : ( nLayers - overlap + step - 1 ) / step;
}

size_t safeMemoryLimit( int device )
{
    int cur = 0;
    cudaGetDevice( &cur );
    size_t memFree = 0, memTot = 0;
    if ( cudaSetDevice( device ) != cudaSuccess || cudaMemGetInfo( &memFree, &memTot ) != cudaSuccess )
    {
        (void)cudaGetLastError();
        cudaSetDevice( cur );
        return 0;
    }
    cudaSetDevice( cur );
    const size_t extra = (size_t)128 << 20; // "minus extra 128 MB", MRCudaBasic.cpp:82-83
    if ( memFree <= extra ) return 0;
    return size_t( (float)( memFree - extra ) * 0.80f ); // cMaxGpuMemoryUsage, MRCudaBasic.cpp:88-89
}

namespace
{

const char* kCanceled = "Operation was canceled"; // source/MRMesh/MRExpected.h:42-51

// ---- state of one call, shared by the calling thread, the device drivers and the copier threads --------------------------
struct CallState
{
    std::mutex m;
    std::condition_variable cv;
    std::vector<int> piecesLeft;   // staged copies: pieces of the chunk not yet in the destination
    std::vector<char> done;        // chunk is complete in its destination
    std::vector<float*> dst;       // destination of the chunk, nullptr = not known yet
    std::vector<char> consumed;    // by-parts: the consumer has returned (library part buffers become free again)
    bool stop = false;             // cancellation or failure: no further chunks
    int rc = B2FWN_OK;
    std::string error;

    explicit CallState( int count ) : piecesLeft( count, 0 ), done( count, 0 ), dst( count, nullptr ), consumed( count, 0 ) {}
    void fail( int code, const std::string& msg )
    {
        std::lock_guard<std::mutex> l( m );
        if ( rc == B2FWN_OK ) { rc = code; error = msg; }
        stop = true;
        cv.notify_all();
    }
    bool stopped()
    {
        std::lock_guard<std::mutex> l( m );
        return stop;
    }
};

// ---- pinned staging ring + copier threads of one device -------------------------------------------------------------------
struct CopyJob
{
    int slot;
    float* dst;
    size_t floats;
    CallState* call;
    int chunk;
    std::atomic<int>* outstanding; // jobs of the issuing driver thread still in the ring
};

struct HostRing
{
    static constexpr int R = 12;                          // staging slots
    size_t PIECE = (size_t)8 << 20;                       // floats per slot: 32 MiB (B2FWN_PIECE_FLOATS: test switch)
    int device = 0;
    float* slot[R] = {};
    cudaEvent_t ev[R] = {};
    bool busy[R] = {};
    std::mutex m;
    std::condition_variable cvSlot, cvJob;
    std::deque<CopyJob> jobs;
    bool ok = false;
    std::string error;

    explicit HostRing( int dev ) : device( dev )
    {
        if ( cudaSetDevice( dev ) != cudaSuccess ) { error = "cudaSetDevice"; return; }
        if ( const char* p = getenv( "B2FWN_PIECE_FLOATS" ) ) PIECE = std::max<size_t>( 1024, (size_t)atoll( p ) );
        for ( int s = 0; s < R; ++s )
        {
            cudaError_t e = cudaMallocHost( &slot[s], PIECE * sizeof( float ) );
            if ( e == cudaSuccess ) e = cudaEventCreateWithFlags( &ev[s], cudaEventDisableTiming );
            if ( e != cudaSuccess ) { error = cudaGetErrorString( e ); (void)cudaGetLastError(); release(); return; }
        }
        // measured on the 16-thread host of a B200 box (c5_band, 34 GB into pageable memory): 2 threads 1577 ms, 4 threads 1345 ms,
        // 8 threads 1133 ms against 1116 ms into pinned memory (profiles/r02)
        int threads = (int)std::max( 2u, std::min( 8u, std::thread::hardware_concurrency() / 2 ) );
        if ( const char* t = getenv( "B2FWN_COPY_THREADS" ) ) threads = std::max( 1, std::min( atoi( t ), R ) );
        for ( int t = 0; t < threads; ++t ) std::thread( [this] { threadMain(); } ).detach();
        ok = true;
    }

    // a ring whose set-up failed half way gives back what it got (no thread has been started yet)
    void release()
    {
        for ( int s = 0; s < R; ++s )
        {
            if ( slot[s] ) cudaFreeHost( slot[s] );
            if ( ev[s] ) cudaEventDestroy( ev[s] );
            slot[s] = nullptr; ev[s] = nullptr;
        }
        (void)cudaGetLastError();
    }

    int acquire()
    {
        std::unique_lock<std::mutex> l( m );
        for ( ;; )
        {
            for ( int s = 0; s < R; ++s )
                if ( !busy[s] ) { busy[s] = true; return s; }
            cvSlot.wait( l );
        }
    }

    void submit( const CopyJob& j )
    {
        { std::lock_guard<std::mutex> l( m ); jobs.push_back( j ); }
        cvJob.notify_one();
    }

    void threadMain()
    {
        cudaSetDevice( device );
        for ( ;; )
        {
            CopyJob j;
            {
                std::unique_lock<std::mutex> l( m );
                cvJob.wait( l, [this] { return !jobs.empty(); } );
                j = jobs.front();
                jobs.pop_front();
            }
            const cudaError_t e = cudaEventSynchronize( ev[j.slot] );
            if ( e == cudaSuccess ) memcpy( j.dst, slot[j.slot], j.floats * sizeof( float ) );
            else j.call->fail( B2FWN_ERR_CUDA, std::string( "NVIDIA GPU error: " ) + cudaGetErrorString( e ) + " [staged device-to-host copy]" );
            { std::lock_guard<std::mutex> l( m ); busy[j.slot] = false; }
            cvSlot.notify_one();
            {
                std::lock_guard<std::mutex> l( j.call->m );
                if ( --j.call->piecesLeft[j.chunk] == 0 ) { j.call->done[j.chunk] = 1; j.call->cv.notify_all(); }
            }
            j.outstanding->fetch_sub( 1, std::memory_order_release );
        }
    }
};

// one ring per device for the life of the process (384 MiB of pinned memory, allocated at the first pageable host-output call
// on that device; never freed once it stands: its threads outlive every handle)
HostRing* ringOf( int device, std::string& err )
{
    static std::mutex gm;
    static std::map<int, HostRing*> rings;
    std::lock_guard<std::mutex> l( gm );
    auto it = rings.find( device );
    if ( it == rings.end() ) it = rings.emplace( device, new HostRing( device ) ).first;
    if ( !it->second->ok )
    {
        // the failed ring is not kept: the call that needed it fails with the reason, the next one tries again (pinned memory may
        // have been short only for a moment)
        err = "NVIDIA GPU error: " + it->second->error + " [pinned staging ring]";
        delete it->second;
        rings.erase( it );
        return nullptr;
    }
    return it->second;
}

bool isPinned( const void* p )
{
    cudaPointerAttributes a{};
    if ( cudaPointerGetAttributes( &a, p ) != cudaSuccess ) { (void)cudaGetLastError(); return false; }
    return a.type == cudaMemoryTypeHost;
}

struct Geometry
{
    int nz = 0, count = 0, maxLayers = 0;
    size_t layer = 0;
    std::vector<int> begin, end; // layers [begin[k], end[k]) of the call's range
    int z0( int k ) const { return begin[k]; }
    int z1( int k ) const { return end[k]; }
    size_t floats( int k ) const { return (size_t)( end[k] - begin[k] ) * layer; }
    void add( int b, int e ) { begin.push_back( b ); end.push_back( e ); maxLayers = std::max( maxLayers, e - b ); count = (int)begin.size(); }
};

// layers per chunk when the caller leaves the choice to us: whole layers, one eighth of a device's share of the grid but
// between 32 Mi and 256 Mi voxels (128 MiB .. 1 GiB), so that the D2H of chunk k overlaps the kernels of chunk k+1, the copy
// left over after the last kernel is short, cancellation is polled many times per second -- and the per-launch costs
// (walk-length estimates, kernel tails, stream-ordered scratch) are paid tens of times per grid, not hundreds; and never more
// than fits twice (two slots) into the safe share of the free device memory (MRCudaBasic.cpp:86-107)
int chooseLayersPerPart( const std::vector<b2fwn_ctx*>& devs, const int32_t dims[3], int nz, int overlap, int wanted )
{
    const size_t layer = (size_t)dims[0] * dims[1];
    size_t budgetFloats = ~(size_t)0;
    for ( const b2fwn_ctx* d : devs )
    {
        // scratch slots this device already holds count as available (they are reused), and so does what ensureScratch takes back
        // from the handle's other caches when the allocation fails
        const size_t have = reclaimableBytes( d, RECLAIM_SCRATCH0 | RECLAIM_SCRATCH1 | RECLAIM_MC_WORK | RECLAIM_POOL );
        budgetFloats = std::min( budgetFloats, ( safeMemoryLimit( d->device ) + have ) / ( 2 * sizeof( float ) ) );
    }
    const size_t budgetLayers = std::max<size_t>( budgetFloats / layer, 1 );
    size_t layers;
    if ( wanted > 0 ) layers = (size_t)wanted;
    else
    {
        const size_t total = layer * (size_t)nz;
        size_t floor = (size_t)32 << 20;
        if ( const char* f = getenv( "B2FWN_MIN_CHUNK_VOXELS" ) ) floor = std::max<size_t>( 1, (size_t)atoll( f ) ); // test switch
        const size_t voxels = std::min<size_t>( std::max<size_t>( total / ( 8 * devs.size() ), floor ), (size_t)256 << 20 );
        layers = std::max<size_t>( voxels / layer, 2 );
    }
    layers = std::min( layers, budgetLayers );
    layers = std::max<size_t>( layers, (size_t)overlap + 1 );
    return (int)std::min<size_t>( layers, (size_t)nz );
}

// driver of one device: its chunks are k = d, d + nd, ...
void driveDevice( b2fwn_ctx* dev, const int d, const int nd, const ChunkedCall& c, const Geometry& g, const ChunkLaunch& launch, CallState& S,
                  const bool outPinned, double& kernelMs )
{
    std::atomic<int> outstanding{ 0 };
    auto cudaBad = [&]( cudaError_t e, const char* where )
    {
        if ( e == cudaSuccess ) return false;
        (void)cudaGetLastError();
        S.fail( B2FWN_ERR_CUDA, std::string( "NVIDIA GPU error: " ) + cudaGetErrorString( e ) + " [" + where + "]" );
        return true;
    };
    std::vector<int> mine;
    for ( int k = d; k < g.count; k += nd ) mine.push_back( k );
    const int n = (int)mine.size();
    if ( n == 0 ) return;
    if ( cudaBad( cudaSetDevice( dev->device ), "cudaSetDevice" ) ) return;
    const bool libraryParts = c.part && !c.partBuffer;
    const size_t chunkFloats = (size_t)g.maxLayers * g.layer;
    for ( int s = 0; s < 2 && s < n; ++s )
    {
        int e = ensureScratch( dev, s, chunkFloats );
        if ( !e && libraryParts ) e = ensurePinned( dev, s, chunkFloats );
        if ( e ) { S.fail( e, b2fwn_last_error() ); return; }
    }
    cudaStream_t cs[2] = { dev->stream, dev->stream2 };
    // chunks alternate between two compute streams: the kernels of the next chunk move in as the last (longest) walks of the
    // current one drain -- a chunk is a small launch, its tail would otherwise be paid in full every time
    cudaEventRecord( dev->evA, dev->stream );
    if ( cudaBad( cudaStreamWaitEvent( dev->stream2, dev->evA, 0 ), "cudaStreamWaitEvent" ) ) return;
    bool launched[2] = { false, false };
    auto launchIdx = [&]( int j ) -> bool
    {
        const int k = mine[j], s = j & 1;
        // the slot was read by the D2H of this device's chunk j-2
        if ( j >= 2 && cudaBad( cudaStreamWaitEvent( cs[s], dev->partDone[s], 0 ), "cudaStreamWaitEvent" ) ) return false;
        if ( int e = launch( dev, c.zBegin + g.z0( k ), c.zBegin + g.z1( k ), dev->dScratch[s], cs[s] ) )
        {
            S.fail( e, b2fwn_last_error() );
            return false;
        }
        launched[s] = true;
        return !cudaBad( cudaEventRecord( dev->computeDone[s], cs[s] ), "cudaEventRecord" );
    };
    HostRing* ring = nullptr;
    bool okSoFar = launchIdx( 0 );
    for ( int j = 0; j < n && okSoFar; ++j )
    {
        if ( S.stopped() ) break;
        if ( j + 1 < n && !launchIdx( j + 1 ) ) break;
        const int k = mine[j], s = j & 1;
        // destination of the chunk
        float* dst = nullptr;
        {
            std::unique_lock<std::mutex> l( S.m );
            if ( libraryParts )
            {
                if ( j >= 2 ) S.cv.wait( l, [&] { return S.stop || S.consumed[mine[j - 2]]; } );
                S.dst[k] = dev->hPinned[s];
            }
            else S.cv.wait( l, [&] { return S.stop || S.dst[k]; } );
            if ( S.stop ) break;
            dst = S.dst[k];
        }
        const size_t floats = g.floats( k );
        const float* src = dev->dScratch[s];
        if ( cudaBad( cudaStreamWaitEvent( dev->copyStream, dev->computeDone[s], 0 ), "cudaStreamWaitEvent" ) ) break;
        const bool direct = libraryParts || ( c.hostOut ? outPinned : isPinned( dst ) );
        if ( direct )
        {
            if ( cudaBad( cudaMemcpyAsync( dst, src, floats * sizeof( float ), cudaMemcpyDeviceToHost, dev->copyStream ), "device-to-host copy" ) ) break;
            if ( cudaBad( cudaEventRecord( dev->partDone[s], dev->copyStream ), "cudaEventRecord" ) ) break;
            if ( cudaBad( cudaEventSynchronize( dev->partDone[s] ), "device-to-host copy" ) ) break;
            std::lock_guard<std::mutex> l( S.m );
            S.done[k] = 1;
            S.cv.notify_all();
        }
        else
        {
            if ( !ring )
            {
                std::string err;
                ring = ringOf( dev->device, err );
                if ( !ring ) { S.fail( B2FWN_ERR_CUDA, err ); break; }
            }
            const size_t PIECE = ring->PIECE;
            const int pieces = (int)( ( floats + PIECE - 1 ) / PIECE );
            { std::lock_guard<std::mutex> l( S.m ); S.piecesLeft[k] = pieces; }
            bool bad = false;
            int issued = 0;
            for ( int p = 0; p < pieces && !bad; ++p )
            {
                const size_t off = (size_t)p * PIECE, len = std::min( PIECE, floats - off );
                const int slot = ring->acquire();
                bad = cudaBad( cudaMemcpyAsync( ring->slot[slot], src + off, len * sizeof( float ), cudaMemcpyDeviceToHost, dev->copyStream ), "device-to-host copy" )
                   || cudaBad( cudaEventRecord( ring->ev[slot], dev->copyStream ), "cudaEventRecord" );
                if ( bad )
                {
                    { std::lock_guard<std::mutex> l( ring->m ); ring->busy[slot] = false; }
                    ring->cvSlot.notify_one();
                    break;
                }
                outstanding.fetch_add( 1, std::memory_order_relaxed );
                ring->submit( CopyJob{ slot, dst + off, len, &S, k, &outstanding } );
                ++issued;
            }
            if ( bad ) break;
            if ( cudaBad( cudaEventRecord( dev->partDone[s], dev->copyStream ), "cudaEventRecord" ) ) break;
        }
    }
    // kernel time of the call on this device: from its first launch to the end of its last kernel (both streams)
    if ( launched[0] ) cudaStreamWaitEvent( dev->stream, dev->computeDone[0], 0 );
    if ( launched[1] ) cudaStreamWaitEvent( dev->stream, dev->computeDone[1], 0 );
    cudaEventRecord( dev->evB, dev->stream );
    cudaStreamSynchronize( dev->stream );
    cudaStreamSynchronize( dev->stream2 );
    cudaStreamSynchronize( dev->copyStream );
    while ( outstanding.load( std::memory_order_acquire ) != 0 ) std::this_thread::yield();
    float ms = 0;
    if ( cudaEventElapsedTime( &ms, dev->evA, dev->evB ) == cudaSuccess ) kernelMs = ms;
    (void)cudaGetLastError();
}

} // anonymous namespace

int layersPerPartOf( b2fwn_ctx* h, const int32_t dims[3], int nz, int overlap, int wanted )
{
    std::vector<b2fwn_ctx*> devs{ h };
    devs.insert( devs.end(), h->replicas.begin(), h->replicas.end() );
    return chooseLayersPerPart( devs, dims, nz, overlap, wanted );
}

int runChunked( b2fwn_ctx* h, const ChunkedCall& c, const ChunkLaunch& launch )
{
    const int32_t* dims = c.dims;
    if ( c.zBegin < 0 || c.zEnd > dims[2] || c.zBegin > c.zEnd ) { setError( "invalid layer range" ); return B2FWN_ERR_INVALID; }
    Geometry g{};
    g.nz = c.zEnd - c.zBegin;
    g.layer = (size_t)dims[0] * dims[1];
    if ( g.nz == 0 || g.layer == 0 ) return B2FWN_OK;
    if ( c.overlap < 0 ) { setError( "negative layer overlap" ); return B2FWN_ERR_INVALID; }
    std::vector<b2fwn_ctx*> devs{ h };
    devs.insert( devs.end(), h->replicas.begin(), h->replicas.end() );
    // (r02: chunk sizes that fall geometrically -- half of the rest per round, short tail -- were tried for the contiguous host
    // output and lost: 1778 ms against 1133 ms at c5_band into pageable memory.  The driver thread feeds the copies of a chunk
    // before it can launch the chunk after the next, so a 17-GB first chunk left the GPU idle behind its copy.)
    {
        const int layersPerPart = chooseLayersPerPart( devs, dims, g.nz, c.overlap, c.layersPerPart );
        if ( layersPerPart <= c.overlap ) { setError( "part must be larger than the overlap" ); return B2FWN_ERR_INVALID; }
        // splitByChunks (source/MRMesh/MRChunkIterator.cpp:25-49): offset_k = k*(chunk-overlap), last chunk shorter
        const int count = partCount( g.nz, layersPerPart, c.overlap ), step = layersPerPart - c.overlap;
        // (r02: cutting the LAST chunk of a contiguous host output into halves down to an eighth, so that less copy is left over
        // after the last kernel, changed nothing measurable: c5_band 1119 ms against 1100 - 1133 ms.)
        for ( int k = 0; k < count; ++k ) g.add( k * step, std::min( g.nz, k * step + layersPerPart ) );
    }
    auto report = [&]( float v ) { return !c.progress || c.progress( v, c.progressUser ) != 0; };
    if ( !report( 0.f ) ) { setError( kCanceled ); return B2FWN_ERR_CANCELED; }

    CallState S( g.count );
    bool outPinned = false, registered = false;
    if ( c.hostOut )
    {
        for ( int k = 0; k < g.count; ++k ) S.dst[k] = c.hostOut + (size_t)g.z0( k ) * g.layer;
        outPinned = isPinned( c.hostOut );
        // dev switch: pin the caller's buffer for the duration of the call instead of staging (page-locking tens of GB costs
        // seconds; measured against the staging ring in profiles/r02)
        static const bool wantRegister = devSwitch( "B2FWN_HOST_OUT" ) && !strcmp( devSwitch( "B2FWN_HOST_OUT" ), "register" );
        if ( !outPinned && wantRegister )
        {
            if ( cudaHostRegister( c.hostOut, (size_t)g.nz * g.layer * sizeof( float ), cudaHostRegisterPortable ) == cudaSuccess )
                outPinned = registered = true;
            else (void)cudaGetLastError();
        }
    }
    const int nd = (int)std::min<size_t>( devs.size(), (size_t)g.count );
    std::vector<double> kernelMs( nd, 0.0 );
    std::vector<std::thread> drivers;
    for ( int d = 0; d < nd; ++d )
        drivers.emplace_back( [&, d] { driveDevice( devs[d], d, nd, c, g, launch, S, outPinned, kernelMs[d] ); } );

    // the calling thread: destinations asked ahead, parts consumed in increasing z, progress reported, cancellation polled
    int nextProvide = 0;
    for ( int k = 0; k < g.count; ++k )
    {
        if ( c.part && c.partBuffer )
            for ( ; nextProvide < std::min( g.count, k + 2 * nd ); ++nextProvide )
            {
                const int32_t pd[3] = { dims[0], dims[1], g.z1( nextProvide ) - g.z0( nextProvide ) };
                float* p = c.partBuffer( pd, c.zBegin + g.z0( nextProvide ), c.partUser );
                if ( !p ) { S.fail( B2FWN_ERR_CANCELED, "by-parts consumer failed" ); break; }
                std::lock_guard<std::mutex> l( S.m );
                S.dst[nextProvide] = p;
                S.cv.notify_all();
            }
        {
            std::unique_lock<std::mutex> l( S.m );
            S.cv.wait( l, [&] { return S.stop || S.done[k]; } );
            if ( S.stop ) break;
        }
        if ( c.part )
        {
            const int32_t pd[3] = { dims[0], dims[1], g.z1( k ) - g.z0( k ) };
            if ( c.part( S.dst[k], pd, c.zBegin + g.z0( k ), c.partUser ) != 0 ) { S.fail( B2FWN_ERR_CANCELED, "by-parts consumer failed" ); break; }
            std::lock_guard<std::mutex> l( S.m );
            S.consumed[k] = 1;
            S.cv.notify_all();
        }
        if ( !report( float( k + 1 ) / float( g.count ) ) ) { S.fail( B2FWN_ERR_CANCELED, kCanceled ); break; }
    }
    for ( auto& t : drivers ) t.join();
    cudaSetDevice( h->device );
    if ( registered ) cudaHostUnregister( c.hostOut );
    if ( S.rc != B2FWN_OK ) { setError( S.error ); return S.rc; }
    h->stats.last_kernel_ms = *std::max_element( kernelMs.begin(), kernelMs.end() );
    return B2FWN_OK;
}

} // namespace b2
