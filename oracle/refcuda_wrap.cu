// oracle/refcuda_wrap.cu -- TEST / BENCH INFRASTRUCTURE ONLY (never linked into the product).
//
// Our thin extern "C" driver around the reference's UNMODIFIED CUDA kernels
// (/root/reference/source/MRCuda/MRCudaFastWindingNumber.cu, compiled where it lies for sm_100 with the reference's own
// release flag -use_fast_math, cmake/Modules/CudaCompilerOptions.cmake:2).  It feeds them exactly what
// MR::Cuda::FastWindingNumber::prepareData_ uploads (source/MRCuda/MRCudaFastWindingNumber.cpp:46-85): the CPU-built
// AABB-tree nodes and dipoles, the points and the triangulation, and launches MR::Cuda::signedDistance /
// fastWindingNumberFromGrid through the reference's own launchers (.cu:245-274) over the whole grid in one chunk.
// Purpose: SURVEY.md 8(d) "second baseline" = the reference GPU path on the same B200, timed with CUDA events.
#include "MRCudaFastWindingNumber.cuh"
#include "MRMesh/MRDistanceToMeshOptions.h"
#include <cuda_runtime.h>
#include <cstdio>

using namespace MR;

namespace
{
struct Ctx
{
    Cuda::Dipole* dipoles = nullptr;
    Cuda::Node3* nodes = nullptr;
    float3* points = nullptr;
    Cuda::FaceToThreeVerts* faces = nullptr;
};
Cuda::Matrix4 toMatrix( const float* xf12 )
{
    Cuda::Matrix4 m;
    m.x = { xf12[0], xf12[1], xf12[2] };
    m.y = { xf12[3], xf12[4], xf12[5] };
    m.z = { xf12[6], xf12[7], xf12[8] };
    m.b = { xf12[9], xf12[10], xf12[11] };
    m.isIdentity = false;
    return m;
}
} // namespace

extern "C"
{

void* mrrefcuda_create( const void* nodes32, const void* dipoles32, size_t nNodes, const float* pts, size_t nVerts, const int* tris, size_t nFaces )
{
    static_assert( sizeof( Cuda::Node3 ) == 32 && sizeof( Cuda::Dipole ) == 32 && sizeof( Cuda::FaceToThreeVerts ) == 12 );
    Ctx* c = new Ctx;
    bool ok = cudaMalloc( &c->nodes, nNodes * 32 ) == cudaSuccess && cudaMalloc( &c->dipoles, nNodes * 32 ) == cudaSuccess
              && cudaMalloc( &c->points, nVerts * 12 ) == cudaSuccess && cudaMalloc( &c->faces, nFaces * 12 ) == cudaSuccess;
    ok = ok && cudaMemcpy( c->nodes, nodes32, nNodes * 32, cudaMemcpyHostToDevice ) == cudaSuccess
         && cudaMemcpy( c->dipoles, dipoles32, nNodes * 32, cudaMemcpyHostToDevice ) == cudaSuccess
         && cudaMemcpy( c->points, pts, nVerts * 12, cudaMemcpyHostToDevice ) == cudaSuccess
         && cudaMemcpy( c->faces, tris, nFaces * 12, cudaMemcpyHostToDevice ) == cudaSuccess;
    if ( !ok ) { std::fprintf( stderr, "mrrefcuda_create: %s\n", cudaGetErrorString( cudaGetLastError() ) ); return nullptr; }
    return c;
}

void mrrefcuda_destroy( void* h )
{
    Ctx* c = (Ctx*)h;
    if ( !c ) return;
    cudaFree( c->nodes ); cudaFree( c->dipoles ); cudaFree( c->points ); cudaFree( c->faces );
    delete c;
}

/// layers [z0, z1) of the grid through the reference's signedDistanceKernel; dOut = device buffer of (z1-z0)*dx*dy floats
/// (or null: allocated and freed here, results optionally copied to hostOut); returns the kernel time in ms, < 0 on error
double mrrefcuda_signed_distance( void* h, const int* dims, const float* xf12, float minDistSq, float maxDistSq, int nullOutside,
                                  float threshold, float beta, int z0, int z1, float* hostOut )
{
    Ctx* c = (Ctx*)h;
    const size_t layer = (size_t)dims[0] * dims[1], n = layer * ( z1 - z0 );
    float* d = nullptr;
    if ( cudaMalloc( &d, n * sizeof( float ) ) != cudaSuccess ) return -1;
    DistanceToMeshOptions o;
    o.minDistSq = minDistSq; o.maxDistSq = maxDistSq; o.nullOutsideMinMax = nullOutside != 0;
    o.windingNumberThreshold = threshold; o.windingNumberBeta = beta;
    cudaEvent_t a, b;
    cudaEventCreate( &a ); cudaEventCreate( &b );
    cudaEventRecord( a );
    Cuda::signedDistance( int3{ dims[0], dims[1], dims[2] }, toMatrix( xf12 ), { c->dipoles, c->nodes, c->points, c->faces }, d, n, layer * z0, o );
    cudaEventRecord( b );
    double ms = -1;
    if ( cudaEventSynchronize( b ) == cudaSuccess && cudaGetLastError() == cudaSuccess )
    {
        float f = 0;
        cudaEventElapsedTime( &f, a, b );
        ms = f;
        if ( hostOut ) cudaMemcpy( hostOut, d, n * sizeof( float ), cudaMemcpyDeviceToHost );
    }
    else
        std::fprintf( stderr, "mrrefcuda_signed_distance: %s\n", cudaGetErrorString( cudaGetLastError() ) );
    cudaEventDestroy( a ); cudaEventDestroy( b );
    cudaFree( d );
    return ms;
}

} // extern "C"
