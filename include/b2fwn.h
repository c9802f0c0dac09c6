/* b2fwn.h -- C ABI of libb2fwn.so: the B200-native (sm_100a) implementation of MeshLib's
 * mesh -> signed-distance-volume hot path (fast winding number + closest-point distance + marching cubes).
 *
 * The reference has NO C/FFI boundary for this path: the plug-in point is the C++ virtual interface
 * MR::IFastWindingNumber / MR::IFastWindingNumberByParts (source/MRMesh/MRFastWindingNumber.h:11-45, 74-107),
 * whose GPU implementation is MR::Cuda::FastWindingNumber (source/MRCuda/MRCudaFastWindingNumber.h:17-38).
 * Every entry point below states the reference method it replaces.  The C++ adapter that derives from the
 * reference interface and forwards to these calls is meshlib_b200/host/b2fwn_meshlib_adapter.hpp
 * (see INTEGRATION.md).
 *
 * Conventions
 *  - plain pointers and sizes only; every call returns B2FWN_OK (0) or an error code; the message of the
 *    last error of the calling thread is b2fwn_last_error().  CUDA failures read "NVIDIA GPU error: <text>"
 *    (reference: source/MRCuda/MRCudaBasic.cpp:109-112); cancellation reads "Operation was canceled"
 *    (reference: source/MRMesh/MRExpected.h:42-51).
 *  - "host" calls take host pointers and copy; "_device" calls take device pointers of the handle's GPU
 *    and only enqueue + synchronise the handle's stream.
 *  - host OUTPUT of the grid calls may be pageable (std::vector<float>, numpy) or pinned.  Pinned memory (cudaMallocHost /
 *    cudaHostRegister) receives each chunk by one asynchronous copy behind the next chunk's kernels.  Pageable memory is
 *    filled through a pinned staging ring (384 MiB per device, allocated at the first such call, kept for the life of the
 *    process) and up to 8 copier threads per device (B2FWN_COPY_THREADS), also behind the next chunk's kernels: both overlap.
 *  - grid linearisation id = x + dx*(y + dy*z) (source/MRMesh/MRVolumeIndexer.h:132-140); the query point of
 *    voxel (x,y,z) is A*(x,y,z)+b with xf = {A row-major (9), b (3)} (source/MRMesh/MRAffineXf.h:46).
 *  - there is no CPU fallback: without a CUDA device b2fwn_create fails.
 */
#ifndef B2FWN_H
#define B2FWN_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2FWN_OK 0
#define B2FWN_ERR_CUDA 1
#define B2FWN_ERR_CANCELED 2
#define B2FWN_ERR_INVALID 3

typedef struct b2fwn_ctx* b2fwn_handle;

/* progress callback: value in [0,1]; return 0 to cancel (MR::ProgressCallback, source/MRMesh/MRProgressCallback.h:14) */
typedef int ( *b2fwn_progress )( float progress, void* user );

/* by-parts consumer: data = dims[0]*dims[1]*dims[2] floats (host), z_offset = first layer of the part;
 * return 0 to continue, non-zero to abort (GridByPartsFunc, source/MRMesh/MRFastWindingNumber.h:88) */
typedef int ( *b2fwn_part_cb )( const float* data, const int32_t dims[3], int32_t z_offset, void* user );

/* Devices of a handle.  n_devices <= 1: the single device `device`; `devices[]` is not read (the reference always uses device 0,
 * MRCudaFastWindingNumber.cpp:48).  n_devices > 1: devices[0 .. n_devices) of ONE process, `device` is not read.  b2fwn_create
 * uploads the mesh to every listed device and each builds its own copy of the tree and the dipoles, all at the same time on one
 * host thread per device (the builds are deterministic, so the copies are identical; no peer access is needed).  What then
 * uses all of them:
 *  - the host-output grid calls (b2fwn_calc_from_grid, b2fwn_calc_from_grid_with_distances[_layers], the by-parts calls,
 *    b2fwn_make_signed_by_winding_number, b2fwn_mesh_to_distance_volume): z-chunks dealt round-robin, every device copying its
 *    chunks straight into the caller's buffer, no gather;
 *  - b2fwn_mc_offset_mesh: one z-range of the grid per device, vertex ids continued by the exclusive prefix of the devices'
 *    counts; b2fwn_marching_cubes_fetch reads every device's share, b2fwn_build_topology collects the triangles on devices[0];
 *  - b2fwn_update_points (every copy is refitted).
 * Everything else -- point vectors, face lists, projections and the ProjectionNormal volume, the *_device forms,
 * b2fwn_marching_cubes on a caller's volume -- runs on devices[0]; device pointers passed to a *_device call belong to devices[0]. */
#define B2FWN_MAX_DEVICES 16
typedef struct b2fwn_config
{
    int32_t device;
    int32_t n_devices;
    int32_t devices[B2FWN_MAX_DEVICES];
    int32_t reserved[6];
} b2fwn_config;

/* mirrors MR::DistanceToMeshOptions (source/MRMesh/MRDistanceToMeshOptions.h:10-33) */
typedef struct b2fwn_dist_options
{
    float min_dist_sq;            /* default 0 */
    float max_dist_sq;            /* default FLT_MAX */
    int32_t null_outside_min_max; /* default 1 */
    float winding_number_threshold; /* default 0.5 */
    float winding_number_beta;      /* default 2 */
} b2fwn_dist_options;

typedef struct b2fwn_stats
{
    double tree_build_ms;     /* device AABB tree + dipoles, last build */
    double last_kernel_ms;    /* device time of the last traversal launch(es), CUDA events */
    uint64_t num_nodes;
    uint64_t num_leaves;
    uint64_t kernel_launches; /* launches of our __global__ kernels since create */
    uint64_t reserved[3];
} b2fwn_stats;

const char* b2fwn_last_error( void );
int b2fwn_device_count( void );

/* replaces MR::Cuda::FastWindingNumber::FastWindingNumber(const Mesh&) + prepareData_
 * (source/MRCuda/MRCudaFastWindingNumber.cpp:42-85): uploads points (nVerts*3 floats) and the triangulation
 * (nFaces*3 int32, any negative id = deleted face, as MeshTopology::getTriangulation yields), then builds the
 * AABB tree (reference CPU: source/MRMesh/MRAABBTree.cpp:31-60, MRAABBTreeMaker.hpp:51-171) and the dipoles
 * (source/MRMesh/MRDipole.cpp:23-71) ON THE DEVICE.  cfg may be NULL. */
int b2fwn_create( const float* points_xyz, size_t n_verts, const int32_t* tris_abc, size_t n_faces,
                  const b2fwn_config* cfg, b2fwn_handle* out );
void b2fwn_destroy( b2fwn_handle h );

/* IFastWindingNumber::calcFromVector (MRFastWindingNumber.h:26; CPU MRFastWindingNumber.cpp:28-38;
 * GPU MRCudaFastWindingNumber.cpp:87-127).  skip_face < 0 = skip nothing. */
int b2fwn_calc_from_vector( b2fwn_handle h, const float* pts_xyz, size_t n, float beta, int32_t skip_face,
                            float* out_w, b2fwn_progress cb, void* user );
int b2fwn_calc_from_vector_device( b2fwn_handle h, const float* d_pts_xyz, size_t n, float beta, int32_t skip_face,
                                   float* d_out_w );

/* b2fwn_calc_from_vector_device for callers that must not trigger a device allocation.  It has no counterpart in the
 * reference (whose calcFromVector resizes its DynamicArrays on every call, MRCudaFastWindingNumber.cpp:87-127).
 * Who needs it: a host application that gave all but a sliver of the GPU's memory to its own arena at start-up (a later
 * cudaMalloc or memory-pool growth inside a plug-in then fails, or pushes the application's own data out), and loops with a
 * latency budget, for which an allocation is a synchronisation point of the whole device that stalls their other streams.
 * b2fwn_calc_from_vector_device borrows the index arrays of its Morton sort from the handle's memory pool; this call keeps them
 * in a fixed array inside the library image instead, and calls neither cudaMalloc / cudaFree nor any routine that allocates.
 *  - Same arguments, same results bit for bit (the order in which queries are walked does not enter the arithmetic).
 *  - n > B2FWN_NOALLOC_MAX_QUERIES is rejected with B2FWN_ERR_INVALID and a message that names the limit: split the vector.
 *  - Cost: 4 * 4 B * B2FWN_NOALLOC_MAX_QUERIES + 1 MiB = 17 MiB of device memory on EVERY GPU the process uses with this
 *    library loaded, from module load to process exit, whether or not the call is ever made.
 *  - The first call on a device looks up the array's address on that device (cudaGetSymbolAddress) and keeps it; make that first
 *    call before any stream capture begins.
 *  - All calls on one device share the one array, and it keeps whatever the previous call left there.  Nothing in it has to
 *    start at zero (each call writes every word before reading it), but two calls in flight on the same GPU at once would
 *    overwrite each other's sort: the library therefore lets one call per device through at a time, the others wait for it
 *    to finish.  Processes do not share the array. */
#define B2FWN_NOALLOC_MAX_QUERIES ( 1u << 20 )
int b2fwn_calc_from_vector_device_noalloc( b2fwn_handle h, const float* d_pts_xyz, size_t n, float beta, int32_t skip_face,
                                           float* d_out_w );

/* IFastWindingNumber::calcSelfIntersections (MRFastWindingNumber.h:31; CPU .cpp:40-52; GPU
 * MRCudaFastWindingNumber.cpp:129-173).  out_bits: ceil(n_faces/64) words, bit f set iff w<0 || w>1. */
int b2fwn_calc_self_intersections( b2fwn_handle h, float beta, uint64_t* out_bits, b2fwn_progress cb, void* user );
int b2fwn_calc_self_intersections_device( b2fwn_handle h, float beta, uint64_t* d_out_bits, float* d_out_w_or_null );

/* IPointsToMeshProjector::findProjections (source/MRMesh/MRPointsToMeshProjector.h:23-26; CPU .cpp:17-33; reference GPU
 * source/MRCuda/MRCudaPointsToMeshProjector.cpp:51-123): the closest point of the mesh to every query point, with its face
 * and barycentric coordinates (MeshProjectionResult, MRMeshProject.h:17-31).  point_xf / tree_xf (row-major A then b, or
 * NULL) are MeshProjectionTransforms::rigidXfPoint / nonRigidXfTree as createProjectionTransforms (MRMeshProject.cpp:104-121)
 * returns them: the first moves the query points, the second the tree's boxes and triangles.  No projection within
 * up_dist_limit_sq: face = -1, dist_sq = up_dist_limit_sq.  MeshTriPoint::e is topology (edgeWithLeft(face)) and is
 * filled by the caller, as the reference's GPU class does (MRCudaPointsToMeshProjector.cpp:117-123). */
typedef struct b2fwn_projection
{
    int32_t face;      /* PointOnFace::face, -1 = none */
    float point[3];    /* PointOnFace::point */
    float bary_a, bary_b; /* MeshTriPoint::bary */
    float dist_sq;     /* MeshProjectionResult::distSq */
    int32_t reserved;
} b2fwn_projection;
int b2fwn_find_projections( b2fwn_handle h, const float* points_xyz, size_t n, const float point_xf[12], const float tree_xf[12],
                            float up_dist_limit_sq, float lo_dist_limit_sq, b2fwn_projection* out, b2fwn_progress cb, void* user );
int b2fwn_find_projections_device( b2fwn_handle h, const float* d_points_xyz, size_t n, const float point_xf[12],
                                   const float tree_xf[12], float up_dist_limit_sq, float lo_dist_limit_sq,
                                   b2fwn_projection* d_out );

/* IFastWindingNumber::calcFromGrid (MRFastWindingNumber.h:38; CPU .cpp:54-67; GPU .cpp:175-228) */
int b2fwn_calc_from_grid( b2fwn_handle h, const int32_t dims[3], const float xf[12], float beta, float* out,
                          b2fwn_progress cb, void* user );
int b2fwn_calc_from_grid_device( b2fwn_handle h, const int32_t dims[3], const float xf[12], float beta,
                                 int32_t z_begin, int32_t z_end, float* d_out );

/* IFastWindingNumber::calcFromGridWithDistances (MRFastWindingNumber.h:44; CPU .cpp:69-94; GPU .cpp:230-283).
 * The _device form computes layers [z_begin, z_end) of the dims grid into d_out (first value = voxel
 * (0,0,z_begin)): the z-slab of one rank in the multi-GPU run. */
int b2fwn_calc_from_grid_with_distances( b2fwn_handle h, const int32_t dims[3], const float xf[12],
                                         const b2fwn_dist_options* opt, float* out, b2fwn_progress cb, void* user );
int b2fwn_calc_from_grid_with_distances_device( b2fwn_handle h, const int32_t dims[3], const float xf[12],
                                                const b2fwn_dist_options* opt, int32_t z_begin, int32_t z_end,
                                                float* d_out );

/* z-interleaved form for multi-GPU runs (no counterpart in the reference, which has one GPU): computes the layers
 * z_begin + c * z_stride + k (c = 0, 1, ...; 0 <= k < z_chunk) below z_end, and writes voxel (x, y, z) to
 * d_volume_at_z_begin[x + dims[0] * (y + dims[1] * (z - z_begin))] -- i.e. in place into a volume that holds all layers
 * from z_begin on.  With z_stride = n_ranks * z_chunk and z_begin = z0 + rank * z_chunk every rank gets an equal share of
 * every part of the grid (contiguous z-slabs differ in cost).  z_chunk must be a positive multiple of 4. */
int b2fwn_calc_from_grid_with_distances_device_chunks( b2fwn_handle h, const int32_t dims[3], const float xf[12],
                                                       const b2fwn_dist_options* opt, int32_t z_begin, int32_t z_end,
                                                       int32_t z_chunk, int32_t z_stride, float* d_volume_at_z_begin );

/* Multi-GPU gather fused into the kernels (one process per GPU of one node; replaces the all-gather that would follow
 * b2fwn_calc_from_grid_with_distances_device[_chunks]): every result is stored, besides d_volume_at_z_begin, at the same
 * offset of each of the n_peers peer volumes (device memory of other GPUs, mapped with b2fwn_shared_open).  When the call
 * has returned on every rank (barrier), every volume is complete.  n_peers <= 15. */
int b2fwn_calc_from_grid_with_distances_device_chunks_p2p( b2fwn_handle h, const int32_t dims[3], const float xf[12],
                                                           const b2fwn_dist_options* opt, int32_t z_begin, int32_t z_end,
                                                           int32_t z_chunk, int32_t z_stride, float* d_volume_at_z_begin,
                                                           int32_t n_peers, float* const* d_peer_volumes_at_z_begin );

/* Peer-visible device memory for such volumes: b2fwn_shared_alloc allocates `bytes` on the handle's device and returns a
 * 64-byte token another process of the node passes to b2fwn_shared_open to map the same memory (CUDA IPC; peer access over
 * NVLink is enabled by the mapping).  b2fwn_shared_close unmaps, b2fwn_shared_free releases the allocation. */
#define B2FWN_IPC_TOKEN_BYTES 64
int b2fwn_shared_alloc( b2fwn_handle h, size_t bytes, void** d_ptr, unsigned char token_out[B2FWN_IPC_TOKEN_BYTES] );
int b2fwn_shared_open( b2fwn_handle h, const unsigned char token[B2FWN_IPC_TOKEN_BYTES], void** d_peer_ptr );
int b2fwn_shared_close( b2fwn_handle h, void* d_peer_ptr );
int b2fwn_shared_free( b2fwn_handle h, void* d_ptr );

/* host-output form of one z-range [z_begin, z_end) of the grid (out[0] = voxel (0,0,z_begin)): what one rank of a
 * z-slab sharded run calls; b2fwn_calc_from_grid_with_distances is this with the full range */
int b2fwn_calc_from_grid_with_distances_layers( b2fwn_handle h, const int32_t dims[3], const float xf[12],
                                                const b2fwn_dist_options* opt, int32_t z_begin, int32_t z_end, float* out,
                                                b2fwn_progress cb, void* user );

/* makeSignedByWindingNumber (source/MRVoxels/MRVDBConversions.cpp:491-568), the consumer of calcFromGrid[ByParts] behind the
 * Smooth offset mode and doubleOffsetVdb: a dense volume of unsigned distances (dims voxels, x fastest; voxel (x,y,z) sits
 * at grid_to_mesh_xf * (x,y,z)) gets the sign of the generalized winding number: volume[i] = -volume[i] where
 * w(voxel i) > winding_number_threshold (MRVDBConversions.cpp:527-533).  The reference moves 4 B/voxel of winding numbers
 * to the host and walks the grid again on the CPU; here the volume is updated in place by the winding-number kernel, the
 * numbers never exist in memory.  NaN voxels stay as they are.  _device: d_volume is device memory of the handle's device
 * and the layers [z_begin, z_end) are updated (d_volume points at layer z_begin); the host form streams the volume through
 * the device in z-chunks (H2D, kernel, D2H overlapped). */
int b2fwn_make_signed_by_winding_number_device( b2fwn_handle h, const int32_t dims[3], const float grid_to_mesh_xf[12],
                                                float winding_number_beta, float winding_number_threshold,
                                                int32_t z_begin, int32_t z_end, float* d_volume_at_z_begin );
int b2fwn_make_signed_by_winding_number( b2fwn_handle h, const int32_t dims[3], const float grid_to_mesh_xf[12],
                                         float winding_number_beta, float winding_number_threshold, float* volume,
                                         b2fwn_progress cb, void* user );

/* SignDetectionMode::ProjectionNormal: signedDistanceToMesh (source/MRMesh/MRMeshDistance.cpp:13-59) for n points --
 * the closest point of the mesh by the reference's own search order, then the sign of
 * dot( closest - point, pseudonormal at the closest point ) (Mesh::isOutsideByProjNorm, source/MRMesh/MRMesh.cpp:165-168;
 * pseudonormal of a face, an edge or a vertex by TriPoint::onEdge / inVertex, source/MRMesh/MRMeshMath.cpp:510-557).
 * out[i] = signed distance, quiet NaN when null_outside_min_max and the squared distance is outside [min_dist_sq, max_dist_sq)
 * (or the mesh has no face).  Without null_outside_min_max the limits are ignored, as in the reference (:19-24).
 * winding_number_* of the options are not used.  Edge and vertex pseudonormals are built on the device from the
 * triangulation at the first call (no half-edge topology is needed). */
int b2fwn_signed_distances_by_projection_normal( b2fwn_handle h, const float* points_xyz, size_t n, const b2fwn_dist_options* opt,
                                                 float* out, b2fwn_progress cb, void* user );
int b2fwn_signed_distances_by_projection_normal_device( b2fwn_handle h, const float* d_points_xyz, size_t n,
                                                        const b2fwn_dist_options* opt, float* d_out );

/* IFastWindingNumberByParts::calcFromGridByParts / calcFromGridWithDistancesByParts
 * (MRFastWindingNumber.h:96-106; GPU MRCudaFastWindingNumber.cpp:285-429).  Parts are whole z-layers, delivered
 * strictly in increasing z on the calling thread, consecutive parts overlapping by layer_overlap layers
 * (chunk maths: source/MRMesh/MRChunkIterator.cpp:25-49).  max_layers_per_part <= 0: one eighth of a device's share of the
 * grid (32 Mi .. 256 Mi voxels); in either case never more than fits twice into the safe share of the free device memory,
 * 80 % of (free - 128 MB) as getCudaSafeMemoryLimit (source/MRCuda/MRCudaBasic.cpp:75-107).
 * The part callback receives library-owned pinned memory that is valid during the callback only. */
int b2fwn_calc_from_grid_by_parts( b2fwn_handle h, const int32_t dims[3], const float xf[12], float beta,
                                   int32_t layer_overlap, int32_t max_layers_per_part, b2fwn_part_cb part, void* part_user,
                                   b2fwn_progress cb, void* user );
int b2fwn_calc_from_grid_with_distances_by_parts( b2fwn_handle h, const int32_t dims[3], const float xf[12],
                                                  const b2fwn_dist_options* opt, int32_t layer_overlap,
                                                  int32_t max_layers_per_part, b2fwn_part_cb part, void* part_user,
                                                  b2fwn_progress cb, void* user );

/* The same with caller-owned part memory (what the C++ adapter uses: GridByPartsFunc takes a std::vector<float>&& it owns,
 * MRFastWindingNumber.h:88): `buffer` is asked on the calling thread, in increasing z and a few parts ahead of `part`, where the
 * part with the given dims / z_offset shall be delivered (dims[0]*dims[1]*dims[2] floats; NULL aborts the call); the part is
 * copied there while the next parts are computed, then `part` is called with that pointer.  buffer == NULL: as above. */
typedef float* ( *b2fwn_part_buffer_cb )( const int32_t dims[3], int32_t z_offset, void* user );
int b2fwn_calc_from_grid_by_parts_into( b2fwn_handle h, const int32_t dims[3], const float xf[12], float beta,
                                        int32_t layer_overlap, int32_t max_layers_per_part, b2fwn_part_buffer_cb buffer,
                                        b2fwn_part_cb part, void* part_user, b2fwn_progress cb, void* user );
int b2fwn_calc_from_grid_with_distances_by_parts_into( b2fwn_handle h, const int32_t dims[3], const float xf[12],
                                                       const b2fwn_dist_options* opt, int32_t layer_overlap,
                                                       int32_t max_layers_per_part, b2fwn_part_buffer_cb buffer,
                                                       b2fwn_part_cb part, void* part_user, b2fwn_progress cb, void* user );

/* The partition the by-parts / host-output calls use, without running anything (no device needed): splitByChunks
 * (source/MRMesh/MRChunkIterator.cpp:25-49) over whole layers -- part k begins at k * (layers_per_part - layer_overlap) and has
 * layers_per_part layers, the last one fewer -- and the device (index into the handle's device list) each part is computed on:
 * k mod min(n_devices, n_parts).  Writes at most `capacity` entries; *n_parts = the number of parts. */
int b2fwn_plan_parts( int32_t n_layers, int32_t layers_per_part, int32_t layer_overlap, int32_t n_devices, int32_t capacity,
                      int32_t* out_begin, int32_t* out_layers, int32_t* out_device, int32_t* n_parts );

/* b2fwn_plan_parts is arithmetic only: it is the partition of a call exactly when layers_per_part is the part size that call ends up
 * with.  The calls themselves may choose another one: max_layers_per_part <= 0 means "choose", and a wish that does not fit twice into
 * the memory the handle's devices can count on is cut (never below layer_overlap + 1).  b2fwn_plan_parts_for asks the handle: it runs
 * the same code as the host-output / by-parts calls do on entry, for the layers [z_begin, z_end) of a dims grid, and returns the
 * parts (out_begin = absolute first layer) and the part size it settled on (*layers_per_part, may be NULL).  The answer holds
 * for a call made in the same memory situation; a caller that sizes buffers from it should make the call next.  Both functions
 * refuse 0 < layers <= layer_overlap, as the calls do; no layers = no parts.  (No counterpart in the reference, which derives the
 * part size inside calcFromGridByParts, MRCudaFastWindingNumber.cpp:299-312, and does not tell.) */
int b2fwn_plan_parts_for( b2fwn_handle h, const int32_t dims[3], int32_t z_begin, int32_t z_end, int32_t layer_overlap,
                          int32_t max_layers_per_part, int32_t capacity, int32_t* out_begin, int32_t* out_layers, int32_t* out_device,
                          int32_t* n_parts, int32_t* layers_per_part );

/* getCudaSafeMemoryLimit (source/MRCuda/MRCudaBasic.cpp:75-90) of a device: 80 % of (free - 128 MB), in bytes; 0 on failure */
size_t b2fwn_safe_memory_limit( int32_t device );

/* Memory a handle keeps between calls, and how it is given back.  To spare the allocator (a cudaMalloc / cudaFree pair of tens of GB
 * costs more than the kernels), a handle caches and only ever grows: two device scratch slots of the host-output / by-parts / offset
 * calls (up to the whole offset band: 34 GB at 2048^3), the mesher's workspace (~1.3 B per voxel of the largest slab), two pinned
 * part buffers, and the idle blocks of its stream-ordered memory pool.  The reference frees its DynamicArrays with the object only
 * (source/MRCuda/MRCudaBasic.cuh:19-70); a long-lived host with several handles or other users of the device needs more:
 *  - b2fwn_trim releases all of these caches on every device of the handle (after synchronising them).  The tree, the dipoles
 *    and the last marching-cubes / topology results stay; the next call simply allocates again.  Call it after a large call
 *    whose like will not follow.
 *  - Without being asked, an allocation of the handle that fails with "out of memory" releases those caches the running call does
 *    not use and tries once more; the part-size budget of the host-output calls counts them as available for that reason.
 * A handle never touches another handle's caches: only b2fwn_trim / b2fwn_destroy of that handle return them. */
int b2fwn_trim( b2fwn_handle h );

/* debug / parity: download the device-built tree in the reference's layouts:
 * nodes = AABBTreeNode (32 B: box.min xyz, box.max xyz, l, r; source/MRMesh/MRAABBTreeNode.h:29-45),
 * dipoles = Dipole (32 B: pos xyz, area, dirArea xyz, rr; source/MRMesh/MRDipole.h:10-30). */
int b2fwn_get_tree( b2fwn_handle h, void* out_nodes_32B, void* out_dipoles_32B, size_t capacity_nodes, size_t* n_nodes );

/* Mesh edited in place (same triangulation, moved vertices): AABBTree::refit (source/MRMesh/MRAABBTree.cpp:62-94) with every
 * vertex marked as changed, followed by calcDipoles (MRDipole.cpp:23-71), on the device.  The tree keeps its topology; boxes,
 * dipoles, traversal nodes and leaf triangles are recomputed from points_xyz (n_verts must equal the handle's vertex
 * count).  b2fwn_get_tree afterwards returns what the reference's refit + calcDipoles produce, bit for bit; the reference's
 * CUDA class has no counterpart (it re-uploads and never invalidates).  b2fwn_stats.last_kernel_ms = refit time. */
int b2fwn_update_points( b2fwn_handle h, const float* points_xyz, size_t n_verts );

int b2fwn_get_stats( b2fwn_handle h, b2fwn_stats* out );

/* diagnostics: measured FP32 FMA throughput (TFLOP/s) of the device = the roofline denominator bench.py reports
 * the traversal kernels against (no reference counterpart) */
int b2fwn_measure_fp32_peak( int32_t device, double* tflops );

/* marching cubes over a dense float volume (replaces MR::marchingCubes(SimpleVolume) =
 * VolumeMesher::run, source/MRVoxels/MRMarchingCubes.cpp:382-398, 548-662, 812-1102).
 * Two-call protocol: b2fwn_marching_cubes(...) runs on the device and returns counts; then
 * b2fwn_marching_cubes_fetch copies vertices (n_verts*3 floats) and triangles (n_tris*3 int32). */
typedef struct b2fwn_mc_params
{
    float origin[3];
    float voxel_size[3];
    float iso;
    int32_t less_inside;   /* MarchingCubesParams::lessInside (MRMarchingCubes.h:33) */
    int32_t reserved[4];
} b2fwn_mc_params;
int b2fwn_marching_cubes( b2fwn_handle h, const float* volume, int32_t volume_on_device, const int32_t dims[3],
                          const b2fwn_mc_params* p, uint64_t* n_verts, uint64_t* n_tris );
int b2fwn_marching_cubes_fetch( b2fwn_handle h, float* out_verts_xyz, int32_t* out_tris_abc );

/* ---- callers of the interface, moved onto the device (a18, a19) -------------------------------------- */

/* exact bounding box of the vertices of all valid faces (min xyz, max xyz) = Mesh::computeBoundingBox() as mcOffsetMesh uses
 * it (source/MRMesh/MRMesh.cpp:213-226); B2FWN_ERR_INVALID for a mesh without valid faces */
int b2fwn_get_bounding_box( b2fwn_handle h, float out_min_max[6] );

/* MR::suggestVoxelSize (source/MRVoxels/MROffset.cpp:25-31): cbrt( volume of the mesh's bounding box / approx_num_voxels ), from the
 * box the device build already holds */
int b2fwn_suggest_voxel_size( b2fwn_handle h, float approx_num_voxels, float* out_voxel_size );

/* mirrors MR::DistanceVolumeParams (source/MRVoxels/MRDistanceVolumeParams.h:9-22) */
typedef struct b2fwn_volume_params
{
    float origin[3];      /* corner of voxel (0,0,0); values are taken at voxel centres origin + (i + 0.5) * voxel_size */
    float voxel_size[3];
    int32_t dims[3];
    int32_t reserved[3];
} b2fwn_volume_params;

/* MR::meshToDistanceVolume, SignDetectionMode::HoleWindingRule branch (source/MRVoxels/MRMeshToDistanceVolume.cpp:43-62):
 * out = dims[0]*dims[1]*dims[2] signed distances (host), out_min / out_max = parallelMinMax over the non-NaN values
 * (FLT_MAX / -FLT_MAX when there is none, as MinMaxArg's initial state) */
int b2fwn_mesh_to_distance_volume( b2fwn_handle h, const b2fwn_volume_params* vol, const b2fwn_dist_options* opt, float* out,
                                   float* out_min, float* out_max, b2fwn_progress cb, void* user );

/* MR::meshToDistanceVolume with SignDetectionMode::ProjectionNormal (MRMeshToDistanceVolume.cpp:64-91 + signedDistanceToMesh):
 * voxel centres origin + voxel_size * ( pos + 0.5 ) exactly as meshToDistanceFunctionVolume computes them (:83-84), generated
 * on the device; same outputs as b2fwn_mesh_to_distance_volume */
int b2fwn_mesh_to_distance_volume_by_projection_normal( b2fwn_handle h, const b2fwn_volume_params* vol,
                                                        const b2fwn_dist_options* opt, float* out, float* out_min, float* out_max,
                                                        b2fwn_progress cb, void* user );

/* MR::mcOffsetMesh / generalOffsetMesh( Mode::Standard ) with SignDetectionMode::HoleWindingRule
 * (source/MRVoxels/MROffset.cpp:117-246, grid by calcOriginAndDimensions, source/MRVoxels/MRCalcDims.h:18-27):
 * the signed-distance band is computed into HBM and meshed there (marching cubes at iso = offset, lessInside = true);
 * the volume never leaves the device.  A grid that does not fit the free device memory is worked through in z-slabs (the
 * reference's MarchingCubesByParts, MRMarchingCubes.h:96-122, fed in HBM); a multi-device handle gives every device a z-range of
 * the grid and numbers the vertices by the exclusive prefix of the devices' counts.  The mesh is identical in every case.
 * Fetch it with b2fwn_marching_cubes_fetch. */
typedef struct b2fwn_offset_params
{
    float voxel_size;               /* OffsetParameters::voxelSize, > 0 */
    float offset;
    int32_t close_holes;            /* OffsetParameters::closeHolesInHoleWindingNumber (default 1) */
    float winding_number_threshold; /* default 0.5 */
    float winding_number_beta;      /* default 2 */
    int32_t max_layers_per_slab;    /* 0: as many z-layers per slab as the free device memory allows (normally the whole grid) */
    int32_t reserved[2];
} b2fwn_offset_params;
typedef struct b2fwn_offset_result
{
    float origin[3];
    int32_t dims[3];
    uint64_t n_verts, n_tris;
    double sdf_ms, mc_ms;           /* device time of the two stages (CUDA events) */
} b2fwn_offset_result;
int b2fwn_mc_offset_mesh( b2fwn_handle h, const b2fwn_offset_params* p, b2fwn_offset_result* res, b2fwn_progress cb, void* user );

/* ---- after marching cubes: the half-edge tables of the result (first slice) ---------------------------- */

/* MeshTopology::HalfEdgeRecord (source/MRMesh/MRMeshTopology.h:581-594): sym(e) = e ^ 1; next / prev = counter-clockwise /
 * clockwise neighbour among the half-edges with the same origin; -1 = invalid id */
typedef struct b2fwn_half_edge
{
    int32_t next, prev, org, left;
} b2fwn_half_edge;

/* Replaces Mesh::fromTriMesh -> MeshBuilder::fromTriangles (source/MRMesh/MRMeshBuilder.cpp:371-540; the last step of mcOffsetMesh,
 * source/MRVoxels/MROffset.cpp:222-225) for an oriented manifold triangulation: half-edge records, one half-edge per vertex and
 * per face, built on the device by one radix sort over the face corners.  tris_abc = NULL: the triangles of the handle's last
 * marching-cubes result, still in HBM (n_tris / n_verts are then ignored); otherwise n_tris triples on the host
 * (tris_on_device = 0) or on the handle's device.  Half-edges are numbered by sorted (min vertex, max vertex): 2u = min -> max,
 * 2u + 1 = max -> min (the reference's numbering depends on its thread count and is no contract); faces and vertices keep their
 * ids, edge_per_face[f] starts at the first corner of f.  The three arrays in the layout of MeshTopology::write
 * (MRMeshTopology.cpp:2461-2477) load with MeshTopology::read, which also validates them.
 * Returns B2FWN_ERR_INVALID ("... not an oriented manifold ...") for meshes the direct construction does not cover (an edge with
 * three faces, two faces on the same side of an edge, a vertex with two fans): the caller keeps the CPU builder for those. */
int b2fwn_build_topology( b2fwn_handle h, const int32_t* tris_abc, size_t n_tris, int32_t tris_on_device, size_t n_verts,
                          uint64_t* n_half_edges, uint64_t* n_verts_out, uint64_t* n_faces_out );
int b2fwn_build_topology_fetch( b2fwn_handle h, b2fwn_half_edge* out_edges, int32_t* out_edge_per_vertex, int32_t* out_edge_per_face );

#ifdef __cplusplus
}
#endif
#endif /* B2FWN_H */
