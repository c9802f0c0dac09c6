// oracle/normals_host.cpp -- TEST INFRASTRUCTURE ONLY.
//
// The arithmetic of the ProjectionNormal kernels (meshlib_b200/csrc/b2fwn_normals_math.h, the header the device code is built
// from) run on the host, step for step like meshlib_b200/csrc/b2fwn_normals.cu (corner keys -> stable sorts -> adjacency,
// face normals, vertex sums in increasing face id, sign by the pseudonormal at the closest point), so that tests without a GPU
// can hold it against the reference library (tests/test_oracle.py).  Nothing under meshlib_b200/ links or loads this.
#include "../meshlib_b200/csrc/b2fwn_normals_math.h"

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <numeric>
#include <vector>

using namespace b2n;

static V3 ld( const float* p, size_t i ) { return v3( p[3 * i], p[3 * i + 1], p[3 * i + 2] ); }

extern "C" int b2n_host_tables( const float* pts, int nVerts, const int32_t* tris, int nFaces, float* faceN, float* vertN, int32_t* adj )
{
    const size_t n = 3 * (size_t)nFaces;
    std::vector<uint64_t> eKey( n );
    std::vector<uint32_t> vKey( n ), byEdge( n ), byVert( n );
    for ( size_t i = 0; i < n; ++i )
    {
        const size_t f = i / 3;
        const int k = (int)( i - 3 * f );
        const int a = tris[3 * f + k], b = tris[3 * f + ( k + 1 ) % 3], c = tris[3 * f + ( k + 2 ) % 3];
        const bool valid = a >= 0 && b >= 0 && c >= 0;
        eKey[i] = valid ? ( ( (uint64_t)(uint32_t)std::min( a, b ) << 32 ) | (uint32_t)std::max( a, b ) ) : ~0ull;
        vKey[i] = valid ? (uint32_t)a : 0xFFFFFFFFu;
    }
    std::iota( byEdge.begin(), byEdge.end(), 0u );
    std::iota( byVert.begin(), byVert.end(), 0u );
    std::stable_sort( byEdge.begin(), byEdge.end(), [&]( uint32_t x, uint32_t y ) { return eKey[x] < eKey[y]; } );
    std::stable_sort( byVert.begin(), byVert.end(), [&]( uint32_t x, uint32_t y ) { return vKey[x] < vKey[y]; } );
    for ( size_t j = 0; j < n; ++j )
    {
        const uint64_t k = eKey[byEdge[j]];
        int32_t other = -1;
        if ( k != ~0ull )
        {
            if ( j > 0 && eKey[byEdge[j - 1]] == k ) other = (int32_t)( byEdge[j - 1] / 3 );
            else if ( j + 1 < n && eKey[byEdge[j + 1]] == k ) other = (int32_t)( byEdge[j + 1] / 3 );
        }
        adj[byEdge[j]] = other;
    }
    for ( int f = 0; f < nFaces; ++f )
    {
        const int a = tris[3 * f], b = tris[3 * f + 1], c = tris[3 * f + 2];
        V3 nn = v3( 0.f, 0.f, 0.f );
        if ( a >= 0 && b >= 0 && c >= 0 ) nn = faceNormal( ld( pts, a ), ld( pts, b ), ld( pts, c ) );
        faceN[3 * f] = nn.x; faceN[3 * f + 1] = nn.y; faceN[3 * f + 2] = nn.z;
    }
    size_t j = 0;
    for ( int v = 0; v < nVerts; ++v )
    {
        while ( j < n && vKey[byVert[j]] < (uint32_t)v ) ++j;
        V3 sum = v3( 0.f, 0.f, 0.f );
        for ( ; j < n && vKey[byVert[j]] == (uint32_t)v; ++j )
        {
            const size_t i = byVert[j], f = i / 3;
            const int k = (int)( i - 3 * f );
            sum = add( sum, cornerTerm( ld( pts, tris[3 * f + k] ), ld( pts, tris[3 * f + ( k + 1 ) % 3] ), ld( pts, tris[3 * f + ( k + 2 ) % 3] ) ) );
        }
        const V3 nn = normalized( sum );
        vertN[3 * v] = nn.x; vertN[3 * v + 1] = nn.y; vertN[3 * v + 2] = nn.z;
    }
    return 0;
}

struct Rec
{
    int32_t face;
    float p[3];
    float a, b, distSq;
    int32_t reserved;
};

extern "C" int b2n_host_sign( const void* records, const float* pts, size_t n, const int32_t* tris, const float* faceN, const float* vertN,
                              const int32_t* adj, float minDistSq, float maxDistSq, int nullOutside, float* out )
{
    const Rec* rec = (const Rec*)records;
    uint32_t nanBits = 0x7FC00000u;
    float nan;
    std::memcpy( &nan, &nanBits, 4 );
    for ( size_t i = 0; i < n; ++i )
    {
        const Rec& r = rec[i];
        const int f = r.face;
        if ( f < 0 || ( nullOutside && ( r.distSq < minDistSq || r.distSq >= maxDistSq ) ) ) { out[i] = nan; continue; }
        const float dist = sqrtf( r.distSq );
        const TriSite s = triSite( r.a, r.b );
        V3 nrm;
        if ( s.vertex >= 0 ) nrm = ld( vertN, tris[3 * (size_t)f + s.vertex] );
        else
        {
            nrm = ld( faceN, f );
            if ( s.edge >= 0 )
            {
                const int g = adj[3 * (size_t)f + s.edge];
                nrm = edgeNormal( nrm, g >= 0 ? ld( faceN, g ) : nrm, g >= 0 );
            }
        }
        out[i] = outsideByNormal( v3( r.p[0], r.p[1], r.p[2] ), ld( pts, i ), nrm ) ? dist : -dist;
    }
    return 0;
}

extern "C" int b2n_host_voxel_centres( const float* origin, const float* voxelSize, const int32_t* dims, float* out )
{
    size_t i = 0;
    for ( int z = 0; z < dims[2]; ++z )
        for ( int y = 0; y < dims[1]; ++y )
            for ( int x = 0; x < dims[0]; ++x, ++i )
            {
                out[3 * i] = voxelCentre( origin[0], voxelSize[0], x );
                out[3 * i + 1] = voxelCentre( origin[1], voxelSize[1], y );
                out[3 * i + 2] = voxelCentre( origin[2], voxelSize[2], z );
            }
    return 0;
}
