// Arithmetic of SignDetectionMode::ProjectionNormal as plain host/device functions: the device kernels
// (b2fwn_normals.cu) are built from it, and the test suite compiles the same header for the host to hold it against the
// reference library where there is no GPU.  Everything is written
// operation by operation in the reference's order (no FMA contraction: the .cu is compiled with -fmad=false):
//   Vector3::normalized            source/MRMesh/MRVector3.h:74-80     ( 1 / len ) * v, zero vector for len <= 0
//   leftDirDblArea / leftNormal    source/MRMesh/MRMeshMath.cpp:140-149  cross( b - a, c - a ).normalized()
//   pseudonormal( VertId )         source/MRMesh/MRMeshMath.cpp:510-527  sum over the ring of angle * n.normalized()
//   pseudonormal( UndirectedEdge ) source/MRMesh/MRMeshMath.cpp:529-547  ( nl + nr ).normalized(), one normal at a boundary
//   pseudonormal( MeshTriPoint )   source/MRMesh/MRMeshMath.cpp:549-557  vertex / edge / face by TriPoint::inVertex / onEdge
//   TriPoint::inVertex / onEdge    source/MRMesh/MRTriPoint.h:91-115, eps = 10 * epsilon (MRSegmPoint.h:17)
//   Mesh::isOutsideByProjNorm      source/MRMesh/MRMesh.cpp:165-168    dot( proj - pt, pseudonormal ) <= 0
#pragma once
#include <cfloat>
#include <cmath>

#if defined( __CUDACC__ )
#define B2_HD __host__ __device__ __forceinline__
#else
#define B2_HD inline
#endif

namespace b2n
{

struct V3
{
    float x, y, z;
};

B2_HD V3 v3( float x, float y, float z ) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
B2_HD V3 sub( V3 a, V3 b ) { return v3( a.x - b.x, a.y - b.y, a.z - b.z ); }
B2_HD V3 add( V3 a, V3 b ) { return v3( a.x + b.x, a.y + b.y, a.z + b.z ); }
B2_HD V3 mul( float s, V3 a ) { return v3( s * a.x, s * a.y, s * a.z ); }
B2_HD float dot( V3 a, V3 b ) { return a.x * b.x + a.y * b.y + a.z * b.z; }
B2_HD V3 cross( V3 a, V3 b ) { return v3( a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x ); }
B2_HD float length( V3 a ) { return sqrtf( dot( a, a ) ); }
B2_HD V3 normalized( V3 a )
{
    const float len = length( a );
    if ( len <= 0 ) return v3( 0.f, 0.f, 0.f );
    return mul( 1 / len, a );
}

/// normal of the face (a, b, c), vertices in the order of the triangulation (= getLeftTriVerts of the face's edgeWithLeft)
B2_HD V3 faceNormal( V3 a, V3 b, V3 c ) { return normalized( cross( sub( b, a ), sub( c, a ) ) ); }

/// what one face adds to the pseudonormal of its corner a (b, c = the next two vertices of the face, counter-clockwise)
B2_HD V3 cornerTerm( V3 a, V3 b, V3 c )
{
    const V3 d0 = sub( b, a ), d1 = sub( c, a );
    const V3 n = cross( d0, d1 );
    const float angle = atan2f( length( n ), dot( d0, d1 ) ); // MR::angle, MRVector3.h:227-230
    return mul( angle, normalized( n ) );
}

/// where on its triangle a point with barycentric coordinates (a, b) lies: vertex 0..2, else edge k = (v_k, v_k+1), else inside
struct TriSite
{
    int vertex; // 0..2 or -1
    int edge;   // 0..2 or -1 (only if vertex < 0)
};
B2_HD TriSite triSite( float a, float b )
{
    const float eps = 10 * FLT_EPSILON;
    TriSite s; s.vertex = -1; s.edge = -1;
    const float rest = 1 - a - b;
    if ( a <= eps && b <= eps ) s.vertex = 0;
    else if ( rest <= eps )
    {
        if ( b <= eps ) s.vertex = 1;
        else if ( a <= eps ) s.vertex = 2;
    }
    if ( s.vertex >= 0 ) return s;
    // TriPoint::onEdge: 0 = edge [v1,v2], 1 = edge [v2,v0], 2 = edge [v0,v1]; here numbered by the edge's first vertex
    if ( rest <= eps ) s.edge = 1;
    else if ( a <= eps ) s.edge = 2;
    else if ( b <= eps ) s.edge = 0;
    return s;
}

/// pseudonormal of an edge from the normals of its face and of the face across it (hasOther = 0 at a boundary)
B2_HD V3 edgeNormal( V3 nFace, V3 nOther, int hasOther ) { return hasOther ? normalized( add( nFace, nOther ) ) : nFace; }

B2_HD bool outsideByNormal( V3 projPoint, V3 pt, V3 pseudoNormal ) { return dot( sub( projPoint, pt ), pseudoNormal ) <= 0; }

/// voxel centre of meshToDistanceFunctionVolume (source/MRVoxels/MRMeshToDistanceVolume.cpp:83-84):
/// origin + mult( voxelSize, Vector3f( pos ) + Vector3f::diagonal( 0.5f ) )
B2_HD float voxelCentre( float origin, float voxelSize, int pos ) { return origin + voxelSize * ( (float)pos + 0.5f ); }

} // namespace b2n
