// b2fwn_internal.h -- shared declarations between the translation units of libb2fwn.so.
// Not part of the public boundary (that is include/b2fwn.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <string>
#include <cstdlib>
#include <functional>
#include "../../include/b2fwn.h"

namespace b2
{

// Reference-layout node, 32 B (source/MRMesh/MRAABBTreeNode.h:29-45): leaf <=> r < 0, then l = FaceId.
struct RefNode
{
    float bmin[3];
    float bmax[3];
    int32_t l, r;
};

// Reference-layout dipole, 32 B (source/MRMesh/MRDipole.h:10-30).
struct RefDipole
{
    float pos[3];
    float area;
    float dirArea[3];
    float rr;
};

// Traversal node, 64 B = 4 x float4, one 128-B line holds two consecutive pre-order nodes.
//   q0 = { bmin.x, bmin.y, bmin.z, bmax.x }
//   q1 = { bmax.y, bmax.z, right, link }      <- distance half (one 32-B sector); right = index of the right child (int bits)
//   q2 = { pos.x, pos.y, pos.z, rr }
//   q3 = { dirArea.x, dirArea.y, dirArea.z, link }   <- winding-number half (one 32-B sector)
// link (int bits): internal node: index of the first node AFTER this node's subtree (pre-order "skip" pointer);
//                  leaf: -(leafSlot+1), where leafSlot indexes DeviceTree::tris; a leaf's skip is i+1.
struct __align__( 16 ) TravNode
{
    float4 q0, q1, q2, q3;
};

// Leaf triangle, 48 B: a.w carries the FaceId bits.
struct __align__( 16 ) LeafTri
{
    float4 a, b, c;
};

// Leaf filter record, 64 B: the triangle in its own orthonormal frame (computed in double, rounded once), used for the FP32
// distance estimate that decides whether a candidate needs the reference's exact (double) closest point at all.
//   f0 = { a.x, a.y, a.z, bx }        a = first vertex, bx = |b - a|
//   f1 = { u.x, u.y, u.z, cx }        u = (b - a) / bx,        (cx, cy) = c - a in the (u, v) frame, cy > 0
//   f2 = { n.x, n.y, n.z, cy }        n = unit normal,  v = n x u
//   f3 = { 1/|c-b|^2, 1/|c-a|^2, allowance, 0 }   allowance = absolute error budget of the estimate that depends on the
//        triangle alone; +inf for degenerate triangles (they always take the exact path)
struct __align__( 16 ) LeafFilt
{
    float4 f0, f1, f2, f3;
};

struct DeviceTree
{
    size_t numVerts = 0, numFaces = 0, numLeaves = 0, numNodes = 0;
    int depth = 0;
    float bbox[6] = { 0, 0, 0, 0, 0, 0 }; // exact box of the vertices of valid faces (min xyz, max xyz); valid iff numLeaves > 0
    void* arena = nullptr;       // one allocation behind refNodes, refDipoles, nodes, leafTris, leafFilt, nodeDepth
    float* points = nullptr;     // numVerts*3
    int32_t* tris = nullptr;     // numFaces*3
    RefNode* refNodes = nullptr; // numNodes
    RefDipole* refDipoles = nullptr;
    TravNode* nodes = nullptr;   // numNodes
    LeafTri* leafTris = nullptr; // numLeaves, in pre-order leaf order
    LeafFilt* leafFilt = nullptr; // numLeaves, same order
    // high-priority side stream of the grid launcher (walk-length estimates beside the distance kernels), forked / joined by events
    cudaStream_t sideStream = nullptr;
    cudaEvent_t evFork = nullptr, evJoin = nullptr;
    // private stream-ordered pool of the per-launch scratch (cached across synchronisations; the device's default pool and its
    // release threshold, which belong to the host application, are left alone)
    cudaMemPool_t pool = nullptr;
    // pseudonormal tables of SignDetectionMode::ProjectionNormal, built on first use (b2fwn_normals.cu)
    float* faceNormals = nullptr; // numFaces*3
    float* vertNormals = nullptr; // numVerts*3
    int32_t* faceAdj = nullptr;   // numFaces*3: face across the edge (v_k, v_k+1), -1 at a boundary
    uint8_t* nodeDepth = nullptr; // numNodes: depth of each node (bottom-up passes of build and refit)
};

struct GridXf
{
    float A[9];
    float b[3];
};

// errors -------------------------------------------------------------------------------------------------------
// Development switches (alternative kernels and orders kept for measurements: B2FWN_FWN5, _LINEAR, _NO_V8, _PROJ_BRICKS, _NO_SUBTREES,
// _QUERY_THREADS, _QUERY_WARPS, _NO_QUERY_SORT, _HOST_OUT, _STATS, _DEBUG) are read from the environment ONLY in a library built with
// `make DEV=1` (-DB2FWN_DEV_SWITCHES).  The product library ignores them: a plug-in must not let its host process's environment pick its
// kernels (B2FWN_PROJ_BRICKS=1 selects a walk that differs from the reference in a tie-dependent voxel).  What the product build still
// reads are three sizing knobs that cannot change a result: B2FWN_PIECE_FLOATS, B2FWN_COPY_THREADS (staging ring), B2FWN_MIN_CHUNK_VOXELS
// (smallest automatic part) -- the tests use them to drive small grids through the many-pieces / many-parts paths.
#ifdef B2FWN_DEV_SWITCHES
inline const char* devSwitch( const char* name ) { return getenv( name ); }
#else
inline const char* devSwitch( const char* ) { return nullptr; }
#endif

void setError( const std::string& msg );
int cudaFail( cudaError_t e, const char* where );
// the CUDA error code behind the calling thread's last cudaFail (cudaSuccess when its last failure was none of CUDA's)
cudaError_t lastCudaError();
#define B2_CUDA( expr )                                        \
    do {                                                       \
        cudaError_t _e = ( expr );                             \
        if ( _e != cudaSuccess )                               \
            return b2::cudaFail( _e, #expr );                  \
    } while ( 0 )

// tree build (b2fwn_tree.cu) -----------------------------------------------------------------------------------
// side stream, fork / join events and the private memory pool of a tree on `device` (idempotent)
int initTreeAux( DeviceTree& t, int device );
int buildTree( DeviceTree& t, cudaStream_t stream, uint64_t* launches );
// new vertex coordinates in t.points, same triangulation: boxes, dipoles, traversal nodes and leaf triangles recomputed on the
// unchanged tree topology (AABBTree::refit, MRAABBTree.cpp:62-94)
int refitTree( DeviceTree& t, cudaStream_t stream, uint64_t* launches );
void freeTree( DeviceTree& t );

// traversal launchers (b2fwn_kernels.cu) ------------------------------------------------------------------------
int launchFwnVector( const DeviceTree& t, const float* dPts, size_t n, float beta, int32_t skipFace, float* dOut,
                     cudaStream_t stream, uint64_t* launches );
int launchFwnFaces( const DeviceTree& t, float beta, float* dOutWOrNull, uint64_t* dBits, cudaStream_t stream,
                    uint64_t* launches );
// warp-cooperative form for large vectors in a spatially coherent order (b2fwn_grid.cu): dOrder = query index per slot (Morton order)
// or nullptr; faces: the queries are the centres of the leaf triangles, in leaf order, each skipping its own face
int launchFwnQueriesWarp( const DeviceTree& t, const float* dPts, const uint32_t* dOrder, size_t n, bool faces, float beta, int32_t skipFace,
                          float* dOutW, uint32_t* dBits32, cudaStream_t stream, uint64_t* launches );
// z-interleaved layer set of one launch (multi-GPU partition): chunks of `chunk` consecutive layers, `stride` layers apart
struct ZChunks
{
    int chunk, stride;
};
int launchFwnGrid( const DeviceTree& t, const int32_t dims[3], const GridXf& xf, float beta, int zBegin, int zEnd,
                   float* dOut, cudaStream_t stream, uint64_t* launches, const ZChunks* zc = nullptr );
// volumes of the other ranks (peer device memory) that receive every result besides the rank's own volume
constexpr int MAX_PEERS = 15;
struct PeerOut
{
    int n;
    float* p[MAX_PEERS];
};
int launchSdfGrid( const DeviceTree& t, const int32_t dims[3], const GridXf& xf, const b2fwn_dist_options& opt,
                   int zBegin, int zEnd, float* dOut, cudaStream_t stream, uint64_t* launches, const ZChunks* zc = nullptr,
                   const PeerOut* peers = nullptr );
// slab of (unsigned) distances -> negated in place where the voxel's winding number exceeds the threshold
int launchSignGrid( const DeviceTree& t, const int32_t dims[3], const GridXf& xf, float beta, float threshold, int zBegin, int zEnd,
                    float* dVolumeAtZBegin, cudaStream_t stream, uint64_t* launches );

// closest point with face and barycentrics (b2fwn_project.cu) -----------------------------------------------------------
int launchFindProjections( const DeviceTree& t, const float* dPts, size_t n, const float* pointXf12, const float* treeXf12,
                           float upDistLimitSq, float loDistLimitSq, b2fwn_projection* dOut, cudaStream_t stream, uint64_t* launches );

// signed distance by the pseudonormal at the closest point (b2fwn_normals.cu) ----------------------------------------------
// dProj: n records of scratch; dPts / dOut: n points in, n signed distances (NaN = none) out
int ensurePseudoNormals( DeviceTree& t, cudaStream_t stream, uint64_t* launches );
int launchSignedDistProjNormal( DeviceTree& t, const float* dPts, size_t n, float minDistSq, float maxDistSq, int nullOutside,
                                b2fwn_projection* dProj, float* dOut, cudaStream_t stream, uint64_t* launches );
// the same over the layers [z0, z1) of a volume by the brick walk: launchDistSlotGrid (b2fwn_grid.cu) leaves |d| / NaN in dVol and
// the leaf slot of the closest triangle in dSlot, launchSignBySlot (b2fwn_normals.cu) turns dVol into signed distances
int launchDistSlotGrid( const DeviceTree& t, const int32_t dims[3], const float origin[3], const float voxelSize[3], float maxDistSq,
                        int nullOutside, int zBegin, int zEnd, float* dOut, int32_t* dSlot, cudaStream_t stream, uint64_t* launches );
int launchSignBySlot( DeviceTree& t, const int32_t dims[3], const float origin[3], const float voxelSize[3], int z0, int z1, float* dVol,
                      const int32_t* dSlot, cudaStream_t stream, uint64_t* launches );
// voxel centres of the layers [z0, z1) of a volume, as meshToDistanceFunctionVolume computes them
int launchVoxelCentres( const float origin[3], const float voxelSize[3], int dx, int dy, int z0, int z1, float* dPts, cudaStream_t stream,
                        uint64_t* launches );

// marching cubes (b2fwn_mc.cu) -----------------------------------------------------------------------------------
struct McResult
{
    float* dVerts = nullptr;   // nVerts*3
    int32_t* dTris = nullptr;  // nTris*3
    uint64_t nVerts = 0, nTris = 0;
};
// classification codes, segment totals and scan scratch of the mesher: grow-only, kept by the handle between slabs and calls
// (r02, 2050^3: cudaMalloc / cudaFree of 10 GB of scratch per call cost more than the two kernels -- 170 .. 480 ms against ~110 ms)
struct McWorkspace
{
    uint8_t* dCode = nullptr; size_t codeBytes = 0;
    uint32_t *dSegV = nullptr, *dSegT = nullptr; size_t segEntries = 0;
    void* dTmp = nullptr; size_t tmpBytes = 0;
};
void freeMcWorkspace( McWorkspace& w );
// layersPerSlab > 0: the volume is meshed in slabs of that many owner layers (identical output; bounds the scratch memory)
int marchingCubesDevice( const float* dVolume, const int32_t dims[3], const b2fwn_mc_params& p, McResult& out,
                         cudaStream_t stream, uint64_t* launches, McWorkspace* work = nullptr, int layersPerSlab = 0 );
// The same for a volume that is produced slab by slab: src( z0, z1, &dPtr ) must make the values of the layers [z0, z1) of the grid
// available on the device (dPtr = layer z0) until its next call; it is asked for the owner layers of a slab plus the two layers
// after them.  Meshes the owner layers [zBegin, zEnd); vertex / triangle ids start at baseV / baseT (several devices each mesh
// a z-range of one grid: the bases are the exclusive prefix of their counts).
using McVolumeSource = std::function<int( int z0, int z1, const float** dLayers )>;
int marchingCubesSlabs( const McVolumeSource& src, const int32_t dims[3], const b2fwn_mc_params& p, int layersPerSlab, int zBegin, int zEnd,
                        uint32_t baseV, uint32_t baseT, McResult& out, cudaStream_t stream, uint64_t* launches, McWorkspace* work );
void freeMc( McResult& r );
// adds base to every vertex id of the triangles (a device's share of a multi-device mesh gets its global ids)
int marchingCubesAddBase( McResult& r, int32_t base, cudaStream_t stream, uint64_t* launches );

// half-edge tables of a triangle mesh (b2fwn_topology.cu) ----------------------------------------------------------------
struct TopoResult
{
    b2fwn_half_edge* dEdges = nullptr; // nHalfEdges records {next, prev, org, left}
    int32_t* dEdgePerVertex = nullptr; // nVerts, -1 = vertex without faces
    int32_t* dEdgePerFace = nullptr;   // nFaces
    size_t nHalfEdges = 0, nVerts = 0, nFaces = 0;
};
int buildTopologyDevice( const int32_t* dTris, size_t nFaces, size_t nVerts, TopoResult& out, cudaStream_t stream, uint64_t* launches );
void freeTopology( TopoResult& r );

} // namespace b2
