/* oracle/port_mc.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Plain-C restatement of the reference's marching cubes over a dense float volume
 * (MR::marchingCubesAsTriMesh(SimpleVolume) = VolumeMesher::run, /root/reference/source/MRVoxels/MRMarchingCubes.cpp):
 *   pass 1  addPartBlock_   :548-662   one vertex per sign-changing +X/+Y/+Z grid edge, in voxel-id order
 *   pass 2  finalize        :812-1102  8-bit case per cell -> <= 5 triangles from the published table, NaN corners
 *                                      borrow the classification of the nearest valid corner (:941-975)
 *   ids     SeparationPointStorage::makeUniqueVids / getTriangulation (source/MRMesh/MRSeparationPoint.cpp:14-70):
 *           blocks are consecutive layer ranges filled sequentially, so global vertex ids are simply the rank of
 *           (voxel id, direction) and triangles are ordered by cell id -- independent of block / thread count.
 * Pinned against the compiled reference (oracle/_ref, tests/test_oracle.py) bit for bit: vertex coordinates and
 * triangle index triples, including order.  Compiled with -ffp-contract=off (the reference build has no FMA).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "../meshlib_b200/csrc/b2fwn_mc_tables.h"

/* cEdgeIndicesMap (MRMarchingCubes.cpp:25-40): table edge -> (corner that owns the vertex, direction 0=X 1=Y 2=Z);
 * corner index bits: 1 = +x, 2 = +y, 4 = +z (cVoxelNeighbors, :42-51) */
static const int EDGE_CORNER[12] = { 0, 1, 2, 0, 4, 5, 6, 4, 0, 1, 3, 2 };
static const int EDGE_DIR[12] = { 0, 1, 0, 1, 0, 1, 0, 1, 2, 2, 2, 2 };
/* cMapNeighbors (:52-64): corner index -> bit of the table's case index */
static const int CORNER_BIT[8] = { 0, 1, 3, 2, 4, 5, 7, 6 };
/* NaN substitution order (:951-959) */
static const int NEI_ORDER[7] = { 1, 2, 4, 3, 5, 6, 7 };

typedef struct
{
    float* verts;
    size_t nVerts;
    int32_t* tris;
    size_t nTris;
} McOut;

static McOut g_out;

static int isLower( float v, float iso ) { return v < iso; }
static int isNotLower( float v, float iso ) { return v >= iso; }
static int isInvalid( float v, float iso ) { return !isLower( v, iso ) && !isNotLower( v, iso ); }
/* vertex on the edge (v0 -> v1) iff exactly one end is lower and the other same-or-higher (:626-639) */
static int hasSepPoint( float v0, float v1, float iso )
{
    if ( isInvalid( v0, iso ) ) return 0;
    return isLower( v0, iso ) ? isNotLower( v1, iso ) : isLower( v1, iso );
}

int orc_marching_cubes( const float* vol, const int32_t* dims, const float* voxelSize, const float* origin, float iso,
                        int lessInside, size_t* nVertsOut, size_t* nTrisOut )
{
    const int dx = dims[0], dy = dims[1], dz = dims[2];
    free( g_out.verts ); free( g_out.tris );
    memset( &g_out, 0, sizeof( g_out ) );
    *nVertsOut = *nTrisOut = 0;
    if ( dx <= 0 || dy <= 0 || dz <= 0 ) return 0;
    if ( dz <= 1 ) return 1; /* "a part must have at least two Z slices" (:443-444) */
    const size_t layer = (size_t)dx * dy, total = layer * dz;
    const size_t step[3] = { 1, (size_t)dx, layer };
    /* first vertex id of every voxel (-1: none) and which directions have one */
    int32_t* firstVid = (int32_t*)malloc( total * sizeof( int32_t ) );
    uint8_t* dirMask = (uint8_t*)calloc( total, 1 );
    size_t capV = 1024, nV = 0;
    float* verts = (float*)malloc( capV * 3 * sizeof( float ) );
    /* zeroPoint = origin + mult( shift, voxelSize ), shift = 0.5 for SimpleVolume (:566; MRVoxelsVolumeAccess.h:78) */
    float zero[3];
    for ( int c = 0; c < 3; ++c ) zero[c] = origin[c] + ( 0.5f + 0.0f ) * voxelSize[c];

    size_t id = 0;
    for ( int z = 0; z < dz; ++z )
        for ( int y = 0; y < dy; ++y )
            for ( int x = 0; x < dx; ++x, ++id )
            {
                firstVid[id] = -1;
                const float v = vol[id];
                if ( isInvalid( v, iso ) ) continue;
                const int pos[3] = { x, y, z };
                const int lim[3] = { dx, dy, dz };
                float coords[3];
                for ( int c = 0; c < 3; ++c ) coords[c] = zero[c] + voxelSize[c] * (float)pos[c];
                for ( int n = 0; n < 3; ++n )
                {
                    if ( pos[n] + 1 >= lim[n] ) continue;
                    const float nv = vol[id + step[n]];
                    if ( !hasSepPoint( v, nv, iso ) ) continue;
                    float next[3] = { coords[0], coords[1], coords[2] };
                    next[n] += voxelSize[n];
                    const float ratio = ( iso - v ) / ( nv - v ); /* :569-578 */
                    if ( nV == capV ) { capV *= 2; verts = (float*)realloc( verts, capV * 3 * sizeof( float ) ); }
                    for ( int c = 0; c < 3; ++c ) verts[3 * nV + c] = ( 1.0f - ratio ) * coords[c] + ratio * next[c];
                    if ( firstVid[id] < 0 ) firstVid[id] = (int32_t)nV;
                    dirMask[id] |= (uint8_t)( 1 << n );
                    ++nV;
                }
            }

    size_t capT = 1024, nT = 0;
    int32_t* tris = (int32_t*)malloc( capT * 3 * sizeof( int32_t ) );
    for ( int z = 0; z + 1 < dz; ++z )
        for ( int y = 0; y + 1 < dy; ++y )
            for ( int x = 0; x + 1 < dx; ++x )
            {
                const size_t base = (size_t)x + (size_t)dx * ( (size_t)y + (size_t)dy * (size_t)z );
                size_t cid[8];
                int low[8], inv[8], anyInv = 0, allInv = 1;
                for ( int i = 0; i < 8; ++i )
                {
                    cid[i] = base + ( i & 1 ) + ( ( i >> 1 ) & 1 ) * step[1] + ( ( i >> 2 ) & 1 ) * step[2];
                    const float v = vol[cid[i]];
                    low[i] = isLower( v, iso );
                    inv[i] = isInvalid( v, iso );
                    anyInv |= inv[i];
                    allInv &= inv[i];
                }
                if ( allInv ) continue;
                int vx[8], atLeastOneNan = 0, cfg = 0;
                for ( int i = 0; i < 8; ++i )
                {
                    int l = low[i];
                    if ( inv[i] )
                    {
                        /* nearest valid corner in the fixed order; one exists because not all 8 are invalid */
                        for ( int k = 0; k < 7; ++k )
                        {
                            const int j = i ^ NEI_ORDER[k];
                            if ( !inv[j] ) { l = low[j]; break; }
                        }
                        atLeastOneNan = 1;
                    }
                    vx[i] = l;
                    if ( l ) cfg |= 1 << CORNER_BIT[i];
                }
                (void)anyInv;
                if ( cfg == 0 || cfg == 0xff ) continue;
                const signed char* plan = B2_MC_TRI_TABLE[cfg];
                int valid = 1;
                if ( atLeastOneNan ) /* "ensure consistent nan voxel" (:1031-1045): every used edge must own a vertex */
                    for ( int i = 0; plan[i] >= 0 && valid; ++i )
                        valid = ( dirMask[cid[EDGE_CORNER[plan[i]]]] >> EDGE_DIR[plan[i]] ) & 1;
                if ( !valid ) continue;
                for ( int i = 0; plan[i] >= 0; i += 3 )
                {
                    int32_t t[3];
                    for ( int k = 0; k < 3; ++k )
                    {
                        const int e = plan[i + k];
                        const size_t owner = cid[EDGE_CORNER[e]];
                        const int d = EDGE_DIR[e];
                        int rank = 0;
                        for ( int b = 0; b < d; ++b ) rank += ( dirMask[owner] >> b ) & 1;
                        t[k] = firstVid[owner] + rank;
                    }
                    if ( nT == capT ) { capT *= 2; tris = (int32_t*)realloc( tris, capT * 3 * sizeof( int32_t ) ); }
                    tris[3 * nT] = t[0];
                    tris[3 * nT + 1] = lessInside ? t[2] : t[1]; /* :1058-1069 */
                    tris[3 * nT + 2] = lessInside ? t[1] : t[2];
                    ++nT;
                }
                (void)vx;
            }
    free( firstVid ); free( dirMask );
    g_out.verts = verts; g_out.nVerts = nV; g_out.tris = tris; g_out.nTris = nT;
    *nVertsOut = nV; *nTrisOut = nT;
    return 0;
}

void orc_marching_cubes_fetch( float* outVerts, int32_t* outTris )
{
    if ( g_out.nVerts ) memcpy( outVerts, g_out.verts, g_out.nVerts * 3 * sizeof( float ) );
    if ( g_out.nTris ) memcpy( outTris, g_out.tris, g_out.nTris * 3 * sizeof( int32_t ) );
}

/* the table itself, for the row-by-row check against the reference's cTriangleTable */
const signed char* orc_mc_table( void ) { return &B2_MC_TRI_TABLE[0][0]; }
