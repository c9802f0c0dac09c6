/* oracle/port.c -- TEST INFRASTRUCTURE ONLY (never linked into or loaded by the product, meshlib_b200/).
 *
 * Plain-C CPU restatement of the reference's mesh -> signed-distance hot path.  Every function cites the
 * reference lines it follows (paths relative to /root/reference/source).  Compile with -O2 -ffp-contract=off and
 * no -march/-ffast-math so that each float operation rounds like the reference's release build
 * (cmake/Modules/CompilerOptions.cmake: no fast-math, baseline x86-64 => no FMA contraction).
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this port bit-for-bit (tree nodes, dipoles, winding numbers,
 * squared distances, signed-distance grids) against oracle/_ref/libmrref.so = the unmodified reference sources
 * compiled in the build container (oracle/Makefile, target ref), against the reference's known-answer tests
 * (MRTest/MRDipoleTests.cpp:25-51, MRAABBTreeTests.cpp:10-27, MRMesh/MRChunkIterator.cpp:52-80), and against the
 * committed fixtures under tests/golden/ generated from the reference (tests/golden/make_golden.py).
 *
 * Besides results it returns OPERATION COUNTS of the reference algorithm (nodes visited, dipoles accepted, exact
 * solid angles, box tests, closest-point evaluations): bench.py turns them into the algorithmic flops/voxel of the
 * roofline.
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { float mn[3], mx[3]; int32_t l, r; } Node;       /* MRMesh/MRAABBTreeNode.h:29-45; leaf <=> r < 0 */
typedef struct { float pos[3], area, dirArea[3], rr; } Dipole;  /* MRMesh/MRDipole.h:10-30 */
typedef struct { float mn[3], mx[3]; int32_t face; } BoxedLeaf; /* MRMesh/MRAABBTreeMaker.h BoxedLeaf */

typedef struct
{
    size_t nv, nf;
    float* pts;
    int32_t* tris;
    int32_t nNodes;
    Node* nodes;
    Dipole* dip;
} OMesh;

/* counters: 0 fwn nodes visited, 1 dipoles accepted, 2 exact solid angles, 3 box tests, 4 closest-point calls */
typedef struct { uint64_t c[5]; } Counters;

/* ---- small vector helpers, written out so that the operation order is explicit ---------------------------------- */
static inline float dot3f( const float* a, const float* b ) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
static inline double dot3d( const double* a, const double* b ) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

static void triPoints( const OMesh* m, int32_t f, float* a, float* b, float* c )
{
    const int32_t* t = m->tris + 3 * (size_t)f;
    memcpy( a, m->pts + 3 * (size_t)t[0], 12 );
    memcpy( b, m->pts + 3 * (size_t)t[1], 12 );
    memcpy( c, m->pts + 3 * (size_t)t[2], 12 );
}

/* ---- AABB tree: MRMesh/MRAABBTree.cpp:15-60, MRAABBTreeMaker.hpp:51-171 --------------------------------------------- */

/* comparator of partitionLeaves_ (MRAABBTreeMaker.hpp:56-69): strict total order */
typedef struct { int dims[3]; } SortDims;
static inline int leafLess( const SortDims* sd, const BoxedLeaf* a, const BoxedLeaf* b )
{
    for ( int i = 2; i >= 0; --i )
    {
        const int d = sd->dims[i];
        const float ad = a->mn[d] + a->mx[d];
        const float bd = b->mn[d] + b->mx[d];
        if ( ad != bd )
            return ad < bd;
    }
    return a->face < b->face;
}

/* findSortedBoxDims (MRMesh/MRBox.h:376-396): bubble sort of pairs (extent, axis), ascending */
static SortDims sortedBoxDims( const float* mn, const float* mx )
{
    float e[3]; int ax[3] = { 0, 1, 2 };
    for ( int i = 0; i < 3; ++i ) e[i] = mx[i] - mn[i];
    for ( int i = 0; i + 1 < 3; ++i )
        for ( int j = i + 1; j < 3; ++j )
            if ( e[j] < e[i] || ( !( e[i] < e[j] ) && ax[j] < ax[i] ) )
            {
                float te = e[i]; e[i] = e[j]; e[j] = te;
                int ta = ax[i]; ax[i] = ax[j]; ax[j] = ta;
            }
    SortDims sd = { { ax[0], ax[1], ax[2] } };
    return sd;
}

/* std::nth_element replacement: any correct selection yields the same left/right SETS because the order is total */
static void nthElement( BoxedLeaf* v, int n, int k, const SortDims* sd )
{
    int lo = 0, hi = n - 1;
    while ( lo < hi )
    {
        /* median of three pivot */
        const int mid = lo + ( hi - lo ) / 2;
        BoxedLeaf t;
        if ( leafLess( sd, &v[mid], &v[lo] ) ) { t = v[mid]; v[mid] = v[lo]; v[lo] = t; }
        if ( leafLess( sd, &v[hi], &v[lo] ) ) { t = v[hi]; v[hi] = v[lo]; v[lo] = t; }
        if ( leafLess( sd, &v[hi], &v[mid] ) ) { t = v[hi]; v[hi] = v[mid]; v[mid] = t; }
        const BoxedLeaf pivot = v[mid];
        int i = lo, j = hi;
        while ( i <= j )
        {
            while ( leafLess( sd, &v[i], &pivot ) ) ++i;
            while ( leafLess( sd, &pivot, &v[j] ) ) --j;
            if ( i <= j ) { t = v[i]; v[i] = v[j]; v[j] = t; ++i; --j; }
        }
        if ( k <= j ) hi = j;
        else if ( k >= i ) lo = i;
        else return;
    }
}

/* makeNode_/makeSubtree (MRAABBTreeMaker.hpp:73-137): node box = union of leaf boxes, split at n/2,
 * l = root+1, r = root+1+getNumNodes(left) = root + 2*left */
static void makeSubtree( Node* nodes, BoxedLeaf* leaves, int root, int first, int n )
{
    while ( 1 )
    {
        Node* nd = &nodes[root];
        if ( n == 1 )
        {
            memcpy( nd->mn, leaves[first].mn, 12 );
            memcpy( nd->mx, leaves[first].mx, 12 );
            nd->l = leaves[first].face; /* setLeafId, MRAABBTreeNode.h:41 */
            nd->r = -1;
            return;
        }
        for ( int d = 0; d < 3; ++d ) { nd->mn[d] = FLT_MAX; nd->mx[d] = -FLT_MAX; }
        for ( int i = 0; i < n; ++i )
            for ( int d = 0; d < 3; ++d )
            {
                if ( leaves[first + i].mn[d] < nd->mn[d] ) nd->mn[d] = leaves[first + i].mn[d];
                if ( leaves[first + i].mx[d] > nd->mx[d] ) nd->mx[d] = leaves[first + i].mx[d];
            }
        const SortDims sd = sortedBoxDims( nd->mn, nd->mx );
        const int nl = n / 2;
        nthElement( leaves + first, n, nl, &sd );
        nd->l = root + 1;
        nd->r = root + 2 * nl;
        makeSubtree( nodes, leaves, nd->l, first, nl );
        /* tail call for the right child */
        root = nd->r; first += nl; n -= nl;
    }
}

static void buildTree( OMesh* m )
{
    int nLeaves = 0;
    for ( size_t f = 0; f < m->nf; ++f )
        if ( m->tris[3 * f] >= 0 && m->tris[3 * f + 1] >= 0 && m->tris[3 * f + 2] >= 0 ) ++nLeaves;
    m->nNodes = 0; m->nodes = NULL; m->dip = NULL;
    if ( nLeaves <= 0 ) return;
    BoxedLeaf* leaves = (BoxedLeaf*)malloc( sizeof( BoxedLeaf ) * (size_t)nLeaves );
    int k = 0;
    for ( size_t f = 0; f < m->nf; ++f )
    {
        if ( !( m->tris[3 * f] >= 0 && m->tris[3 * f + 1] >= 0 && m->tris[3 * f + 2] >= 0 ) ) continue;
        /* computeFaceBox (MRAABBTree.cpp:15-29) + insignificantlyExpanded (MRBox.h:260-270) */
        float p[3][3];
        triPoints( m, (int32_t)f, p[0], p[1], p[2] );
        BoxedLeaf* L = &leaves[k++];
        L->face = (int32_t)f;
        for ( int d = 0; d < 3; ++d )
        {
            float mn = FLT_MAX, mx = -FLT_MAX;
            for ( int v = 0; v < 3; ++v )
            {
                if ( p[v][d] < mn ) mn = p[v][d];
                if ( p[v][d] > mx ) mx = p[v][d];
            }
            L->mn[d] = nextafterf( mn, -FLT_MAX );
            L->mx[d] = nextafterf( mx, FLT_MAX );
        }
    }
    m->nNodes = 2 * nLeaves - 1;
    m->nodes = (Node*)malloc( sizeof( Node ) * (size_t)m->nNodes );
    makeSubtree( m->nodes, leaves, 0, 0, nLeaves );
    free( leaves );
}

/* ---- dipoles: MRMesh/MRDipole.cpp:11-71 --------------------------------------------------------------------------------- */

static void calcDipoles( OMesh* m )
{
    const int N = m->nNodes;
    if ( N <= 0 ) return;
    m->dip = (Dipole*)calloc( (size_t)N, sizeof( Dipole ) );
    for ( int i = 0; i < N; ++i )
    {
        const Node* nd = &m->nodes[i];
        if ( nd->r >= 0 ) continue;
        float a[3], b[3], c[3];
        triPoints( m, nd->l, a, b, c );
        /* dirDblArea = cross(b-a, c-a) (MRMeshMath.cpp:140-149); da = 0.5f * that */
        const float u[3] = { b[0] - a[0], b[1] - a[1], b[2] - a[2] };
        const float v[3] = { c[0] - a[0], c[1] - a[1], c[2] - a[2] };
        const float cr[3] = { u[1] * v[2] - u[2] * v[1], u[2] * v[0] - u[0] * v[2], u[0] * v[1] - u[1] * v[0] };
        const float da[3] = { 0.5f * cr[0], 0.5f * cr[1], 0.5f * cr[2] };
        const float area = sqrtf( dot3f( da, da ) );
        /* triCenter = (1/3.0f) * (v0 + v1 + v2) (MRMeshMath.cpp:119-124) */
        const float third = 1 / 3.0f;
        const float ctr[3] = { third * ( ( a[0] + b[0] ) + c[0] ), third * ( ( a[1] + b[1] ) + c[1] ), third * ( ( a[2] + b[2] ) + c[2] ) };
        Dipole* d = &m->dip[i];
        for ( int k = 0; k < 3; ++k ) { d->pos[k] = area * ctr[k]; d->dirArea[k] = da[k]; }
        d->area = area;
    }
    /* serial reverse sweep (MRDipole.cpp:47-61): children have larger indices than their parent */
    for ( int i = N - 1; i >= 0; --i )
    {
        const Node* nd = &m->nodes[i];
        if ( nd->r < 0 ) continue;
        const Dipole *dl = &m->dip[nd->l], *dr = &m->dip[nd->r];
        Dipole* d = &m->dip[i];
        for ( int k = 0; k < 3; ++k ) { d->pos[k] = dl->pos[k] + dr->pos[k]; d->dirArea[k] = dl->dirArea[k] + dr->dirArea[k]; }
        d->area = dl->area + dr->area;
    }
    for ( int i = 0; i < N; ++i )
    {
        const Node* nd = &m->nodes[i];
        Dipole* d = &m->dip[i];
        if ( d->area > 0 )
        {
            const float inv = 1 / d->area; /* Vector3 /= is *= (1/b), MRVector3.h:131-137 */
            d->pos[0] *= inv; d->pos[1] *= inv; d->pos[2] *= inv;
        }
        float rr = 0; /* distToFarthestCornerSq, MRDipole.cpp:11-21 */
        for ( int k = 0; k < 3; ++k )
        {
            const float tmn = d->pos[k] - nd->mn[k], tmx = d->pos[k] - nd->mx[k];
            const float a2 = tmn * tmn, b2 = tmx * tmx;
            rr += a2 < b2 ? b2 : a2; /* std::max(toMinSq, toMaxSq) */
        }
        d->rr = rr;
    }
}

/* ---- winding number: MRMesh/MRDipole.h:18-27, MRDipole.cpp:81-128 ------------------------------------------------------ */

static float triangleSolidAngle( const float* p, const float* ta, const float* tb, const float* tc )
{
    const float a[3] = { ta[0] - p[0], ta[1] - p[1], ta[2] - p[2] };
    const float b[3] = { tb[0] - p[0], tb[1] - p[1], tb[2] - p[2] };
    const float c[3] = { tc[0] - p[0], tc[1] - p[1], tc[2] - p[2] };
    const float x = sqrtf( dot3f( a, a ) ), y = sqrtf( dot3f( b, b ) ), z = sqrtf( dot3f( c, c ) );
    const float den = x * y * z + dot3f( a, b ) * z + dot3f( b, c ) * x + dot3f( c, a ) * y;
    /* Matrix3::det (MRMatrix3.h:227-233) */
    const float det = a[0] * ( b[1] * c[2] - b[2] * c[1] ) - a[1] * ( b[0] * c[2] - b[2] * c[0] ) + a[2] * ( b[0] * c[1] - b[1] * c[0] );
    return 2 * atan2f( det, den );
}

static float fastWindingNumber( const OMesh* m, const float* q, float beta, int32_t skipFace, Counters* cnt )
{
    if ( m->nNodes <= 0 ) return 0;
    const float betaSq = beta * beta;
    int32_t stack[64]; int sp = 0;
    stack[sp++] = 0;
    float res = 0;
    uint64_t visits = 0, accepts = 0, leavesN = 0;
    while ( sp > 0 )
    {
        const int32_t i = stack[--sp];
        const Node* nd = &m->nodes[i];
        const Dipole* d = &m->dip[i];
        ++visits;
        /* addIfGoodApprox */
        const float dp[3] = { d->pos[0] - q[0], d->pos[1] - q[1], d->pos[2] - q[2] };
        const float dd = dot3f( dp, dp );
        if ( !( dd <= betaSq * d->rr ) )
        {
            const float dist = sqrtf( dd );
            if ( dist > 0 )
                res += dot3f( dp, d->dirArea ) / ( dist * dd );
            ++accepts;
            continue;
        }
        if ( nd->r >= 0 )
        {
            stack[sp++] = nd->r;
            stack[sp++] = nd->l;
            continue;
        }
        if ( nd->l != skipFace )
        {
            float a[3], b[3], c[3];
            triPoints( m, nd->l, a, b, c );
            res += triangleSolidAngle( q, a, b, c );
            ++leavesN;
        }
    }
    if ( cnt ) { cnt->c[0] += visits; cnt->c[1] += accepts; cnt->c[2] += leavesN; }
    const float INV_4PI = 1.0f / ( 4 * 3.14159265358979323846f );
    return INV_4PI * res;
}

/* ---- closest point: MRMesh/MRClosestPointInTriangle.h:27-87 (T = double), MRBox.h:207-220, MRMeshProject.cpp:13-103 ----- */

static void closestPointInTriangleD( const double* p, const double* a, const double* b, const double* c, double* out )
{
    const double ab[3] = { b[0] - a[0], b[1] - a[1], b[2] - a[2] };
    const double ac[3] = { c[0] - a[0], c[1] - a[1], c[2] - a[2] };
    const double ap[3] = { p[0] - a[0], p[1] - a[1], p[2] - a[2] };
    const double d1 = dot3d( ab, ap ), d2 = dot3d( ac, ap );
    if ( d1 <= 0 && d2 <= 0 ) { memcpy( out, a, 24 ); return; }
    const double bp[3] = { p[0] - b[0], p[1] - b[1], p[2] - b[2] };
    const double d3 = dot3d( ab, bp ), d4 = dot3d( ac, bp );
    if ( d3 >= 0 && d4 <= d3 ) { memcpy( out, b, 24 ); return; }
    const double cp[3] = { p[0] - c[0], p[1] - c[1], p[2] - c[2] };
    const double d5 = dot3d( ab, cp ), d6 = dot3d( ac, cp );
    if ( d6 >= 0 && d5 <= d6 ) { memcpy( out, c, 24 ); return; }
    const double vc = d1 * d4 - d3 * d2;
    if ( vc <= 0 && d1 >= 0 && d3 <= 0 )
    {
        const double v = d1 / ( d1 - d3 );
        for ( int k = 0; k < 3; ++k ) out[k] = a[k] + v * ab[k];
        return;
    }
    const double vb = d5 * d2 - d1 * d6;
    if ( vb <= 0 && d6 <= 0 && d2 >= 0 )
    {
        const double v = d2 / ( d2 - d6 );
        for ( int k = 0; k < 3; ++k ) out[k] = a[k] + v * ac[k];
        return;
    }
    const double va = d3 * d6 - d5 * d4;
    if ( va <= 0 )
    {
        if ( d4 < d3 ) { memcpy( out, b, 24 ); return; }
        if ( d5 < d6 ) { memcpy( out, c, 24 ); return; }
        const double v = ( d4 - d3 ) / ( ( d4 - d3 ) + ( d5 - d6 ) );
        for ( int k = 0; k < 3; ++k ) out[k] = b[k] + v * ( c[k] - b[k] );
        return;
    }
    const double denom = 1 / ( va + vb + vc );
    const double v = vb * denom, w = vc * denom;
    for ( int k = 0; k < 3; ++k ) out[k] = a[k] + v * ab[k] + w * ac[k];
}

static inline float boxDistSq( const Node* nd, const float* pt )
{
    float res = 0;
    for ( int i = 0; i < 3; ++i )
    {
        if ( pt[i] < nd->mn[i] ) { const float d = pt[i] - nd->mn[i]; res += d * d; }
        else if ( pt[i] > nd->mx[i] ) { const float d = pt[i] - nd->mx[i]; res += d * d; }
    }
    return res;
}

static float triDistSq( const OMesh* m, int32_t face, const float* pt )
{
    float a[3], b[3], c[3];
    triPoints( m, face, a, b, c );
    const double pd[3] = { pt[0], pt[1], pt[2] }, ad[3] = { a[0], a[1], a[2] }, bd[3] = { b[0], b[1], b[2] }, cd[3] = { c[0], c[1], c[2] };
    double proj[3];
    closestPointInTriangleD( pd, ad, bd, cd, proj );
    const float d[3] = { (float)proj[0] - pt[0], (float)proj[1] - pt[1], (float)proj[2] - pt[2] };
    return dot3f( d, d );
}

/* findProjectionSubtree with region = xf = validFaces = validProjections = null (the path's call,
 * MRFastWindingNumber.cpp:71, MRMeshProject.cpp:123-127); returns distSq (initially upDistLimitSq) */
static float findProjectionDistSq( const OMesh* m, const float* pt, float upDistLimitSq, float loDistLimitSq, Counters* cnt )
{
    float best = upDistLimitSq;
    if ( m->nNodes <= 0 ) return best;
    struct { int32_t n; float d; } stack[64]; int sp = 0;
    uint64_t boxTests = 0, triTests = 0;
    {
        const float d = boxDistSq( &m->nodes[0], pt ); ++boxTests;
        if ( d < best ) { stack[sp].n = 0; stack[sp].d = d; ++sp; }
    }
    while ( sp > 0 )
    {
        --sp;
        const int32_t n = stack[sp].n; const float sd = stack[sp].d;
        const Node* nd = &m->nodes[n];
        if ( sd >= best ) continue;
        if ( nd->r < 0 )
        {
            const float cand = triDistSq( m, nd->l, pt ); ++triTests;
            if ( cand < best )
            {
                best = cand;
                if ( best <= loDistLimitSq ) break;
            }
            continue;
        }
        const float d1 = boxDistSq( &m->nodes[nd->l], pt ), d2 = boxDistSq( &m->nodes[nd->r], pt );
        boxTests += 2;
        if ( d1 < d2 )
        {
            if ( d2 < best ) { stack[sp].n = nd->r; stack[sp].d = d2; ++sp; }
            if ( d1 < best ) { stack[sp].n = nd->l; stack[sp].d = d1; ++sp; }
        }
        else
        {
            if ( d1 < best ) { stack[sp].n = nd->l; stack[sp].d = d1; ++sp; }
            if ( d2 < best ) { stack[sp].n = nd->r; stack[sp].d = d2; ++sp; }
        }
    }
    if ( cnt ) { cnt->c[3] += boxTests; cnt->c[4] += triTests; }
    return best;
}

/* ---- full projection result: IPointsToMeshProjector::findProjections ----------------------------------
 * MRPointsToMeshProjector.cpp:17-33 -> findProjection -> findProjectionSubtree (MRMeshProject.cpp:13-103) with the optional
 * transforms of createProjectionTransforms (MRMeshProject.cpp:104-121): pointXf moves the query point, treeXf (non-rigid)
 * moves node boxes (transformed(): hull of the 8 corners, MRBox.h:336-345, corner order MRBox.h:289-303) and triangles. */
typedef struct { int32_t face; float p[3]; float a, b; float distSq; int32_t edge; } OProj;

static inline void xfApply( const float* x, const float* v, float* o ) /* AffineXf3f::operator(): A * v + b */
{
    for ( int r = 0; r < 3; ++r )
        o[r] = ( x[3 * r] * v[0] + x[3 * r + 1] * v[1] + x[3 * r + 2] * v[2] ) + x[9 + r];
}

static inline float boxDistSqXf( const Node* nd, const float* treeXf, const float* pt )
{
    if ( !treeXf ) return boxDistSq( nd, pt );
    Node t;
    for ( int k = 0; k < 3; ++k ) { t.mn[k] = FLT_MAX; t.mx[k] = -FLT_MAX; } /* Box default: min = max(), max = lowest() */
    for ( int c = 0; c < 8; ++c )
    {
        const float v[3] = { ( c & 1 ) ? nd->mx[0] : nd->mn[0], ( c & 2 ) ? nd->mx[1] : nd->mn[1], ( c & 4 ) ? nd->mx[2] : nd->mn[2] };
        float q[3];
        xfApply( treeXf, v, q );
        for ( int k = 0; k < 3; ++k ) { if ( q[k] < t.mn[k] ) t.mn[k] = q[k]; if ( q[k] > t.mx[k] ) t.mx[k] = q[k]; }
    }
    return boxDistSq( &t, pt );
}

/* closestPointInTriangle<double> with its barycentric pair (MRClosestPointInTriangle.h:27-87) */
static void closestPointBaryD( const double* p, const double* a, const double* b, const double* c, double* out, double* bary )
{
    const double ab[3] = { b[0] - a[0], b[1] - a[1], b[2] - a[2] };
    const double ac[3] = { c[0] - a[0], c[1] - a[1], c[2] - a[2] };
    const double ap[3] = { p[0] - a[0], p[1] - a[1], p[2] - a[2] };
    const double d1 = dot3d( ab, ap ), d2 = dot3d( ac, ap );
    if ( d1 <= 0 && d2 <= 0 ) { memcpy( out, a, 24 ); bary[0] = 0; bary[1] = 0; return; }
    const double bp[3] = { p[0] - b[0], p[1] - b[1], p[2] - b[2] };
    const double d3 = dot3d( ab, bp ), d4 = dot3d( ac, bp );
    if ( d3 >= 0 && d4 <= d3 ) { memcpy( out, b, 24 ); bary[0] = 1; bary[1] = 0; return; }
    const double cp[3] = { p[0] - c[0], p[1] - c[1], p[2] - c[2] };
    const double d5 = dot3d( ab, cp ), d6 = dot3d( ac, cp );
    if ( d6 >= 0 && d5 <= d6 ) { memcpy( out, c, 24 ); bary[0] = 0; bary[1] = 1; return; }
    const double vc = d1 * d4 - d3 * d2;
    if ( vc <= 0 && d1 >= 0 && d3 <= 0 )
    {
        const double v = d1 / ( d1 - d3 );
        for ( int k = 0; k < 3; ++k ) out[k] = a[k] + v * ab[k];
        bary[0] = v; bary[1] = 0; return;
    }
    const double vb = d5 * d2 - d1 * d6;
    if ( vb <= 0 && d6 <= 0 && d2 >= 0 )
    {
        const double v = d2 / ( d2 - d6 );
        for ( int k = 0; k < 3; ++k ) out[k] = a[k] + v * ac[k];
        bary[0] = 0; bary[1] = v; return;
    }
    const double va = d3 * d6 - d5 * d4;
    if ( va <= 0 )
    {
        if ( d4 < d3 ) { memcpy( out, b, 24 ); bary[0] = 1; bary[1] = 0; return; }
        if ( d5 < d6 ) { memcpy( out, c, 24 ); bary[0] = 0; bary[1] = 1; return; }
        const double v = ( d4 - d3 ) / ( ( d4 - d3 ) + ( d5 - d6 ) );
        for ( int k = 0; k < 3; ++k ) out[k] = b[k] + v * ( c[k] - b[k] );
        bary[0] = 1 - v; bary[1] = v; return;
    }
    const double denom = 1 / ( va + vb + vc );
    const double v = vb * denom, w = vc * denom;
    for ( int k = 0; k < 3; ++k ) out[k] = a[k] + v * ab[k] + w * ac[k];
    bary[0] = v; bary[1] = w;
}

static OProj findProjectionFull( const OMesh* m, const float* pt, const float* treeXf, float upDistLimitSq, float loDistLimitSq )
{
    OProj res; memset( &res, 0, sizeof( res ) );
    res.face = -1; res.edge = -1; res.distSq = upDistLimitSq;
    if ( m->nNodes <= 0 ) return res;
    struct { int32_t n; float d; } stack[64]; int sp = 0;
    {
        const float d = boxDistSqXf( &m->nodes[0], treeXf, pt );
        if ( d < res.distSq ) { stack[sp].n = 0; stack[sp].d = d; ++sp; }
    }
    while ( sp > 0 )
    {
        --sp;
        const int32_t n = stack[sp].n; const float sd = stack[sp].d;
        const Node* nd = &m->nodes[n];
        if ( sd >= res.distSq ) continue;
        if ( nd->r < 0 )
        {
            float a[3], b[3], c[3];
            triPoints( m, nd->l, a, b, c );
            if ( treeXf )
            {
                float t[3];
                xfApply( treeXf, a, t ); memcpy( a, t, 12 );
                xfApply( treeXf, b, t ); memcpy( b, t, 12 );
                xfApply( treeXf, c, t ); memcpy( c, t, 12 );
            }
            const double pd[3] = { pt[0], pt[1], pt[2] }, ad[3] = { a[0], a[1], a[2] }, bd[3] = { b[0], b[1], b[2] }, cd[3] = { c[0], c[1], c[2] };
            double proj[3], bary[2];
            closestPointBaryD( pd, ad, bd, cd, proj, bary );
            const float pf[3] = { (float)proj[0], (float)proj[1], (float)proj[2] };
            const float d[3] = { pf[0] - pt[0], pf[1] - pt[1], pf[2] - pt[2] };
            const float cand = dot3f( d, d );
            if ( cand < res.distSq )
            {
                res.face = nd->l; memcpy( res.p, pf, 12 ); res.a = (float)bary[0]; res.b = (float)bary[1]; res.distSq = cand;
                if ( res.distSq <= loDistLimitSq ) break;
            }
            continue;
        }
        const float d1 = boxDistSqXf( &m->nodes[nd->l], treeXf, pt ), d2 = boxDistSqXf( &m->nodes[nd->r], treeXf, pt );
        if ( d1 < d2 )
        {
            if ( d2 < res.distSq ) { stack[sp].n = nd->r; stack[sp].d = d2; ++sp; }
            if ( d1 < res.distSq ) { stack[sp].n = nd->l; stack[sp].d = d1; ++sp; }
        }
        else
        {
            if ( d1 < res.distSq ) { stack[sp].n = nd->l; stack[sp].d = d1; ++sp; }
            if ( d2 < res.distSq ) { stack[sp].n = nd->r; stack[sp].d = d2; ++sp; }
        }
    }
    return res;
}

/* ---- exported C API (loaded by oracle/portlib.py through ctypes) -------------------------------------------------------- */

/* pointXf12 / treeXf12 = row-major A then b, or NULL; out = n OProj records (edge is always -1: no topology here) */
void orc_find_projections( void* h, const float* pts, size_t n, const float* pointXf12, const float* treeXf12, float up, float lo, void* out );

void* orc_create( const float* pts, size_t nv, const int32_t* tris, size_t nf )
{
    OMesh* m = (OMesh*)calloc( 1, sizeof( OMesh ) );
    m->nv = nv; m->nf = nf;
    m->pts = (float*)malloc( nv * 12 + 4 ); memcpy( m->pts, pts, nv * 12 );
    m->tris = (int32_t*)malloc( nf * 12 + 4 ); memcpy( m->tris, tris, nf * 12 );
    buildTree( m );
    calcDipoles( m );
    return m;
}

/* Mesh::updateCaches with every vertex changed (MRMesh/MRMesh.cpp:565-578): AABBTree::refit (MRAABBTree.cpp:62-94) -- leaf boxes
 * by computeFaceBox at the new coordinates, then inner boxes in decreasing node order, box = left box, include( right box ) --
 * and the dipoles recomputed on the refitted tree (dipolesOwner_.reset(), calcDipoles on next use).  The topology stays. */
int orc_update_points( void* h, const float* pts, size_t nv )
{
    OMesh* m = (OMesh*)h;
    if ( nv != m->nv ) return 1;
    memcpy( m->pts, pts, nv * 12 );
    const int N = m->nNodes;
    for ( int i = 0; i < N; ++i )
    {
        Node* nd = &m->nodes[i];
        if ( nd->r >= 0 ) continue;
        float p[3][3];
        triPoints( m, nd->l, p[0], p[1], p[2] );
        for ( int d = 0; d < 3; ++d )
        {
            float mn = FLT_MAX, mx = -FLT_MAX;
            for ( int v = 0; v < 3; ++v )
            {
                if ( p[v][d] < mn ) mn = p[v][d];
                if ( p[v][d] > mx ) mx = p[v][d];
            }
            nd->mn[d] = nextafterf( mn, -FLT_MAX );
            nd->mx[d] = nextafterf( mx, FLT_MAX );
        }
    }
    for ( int i = N - 1; i >= 0; --i )
    {
        Node* nd = &m->nodes[i];
        if ( nd->r < 0 ) continue;
        const Node *l = &m->nodes[nd->l], *r = &m->nodes[nd->r];
        for ( int d = 0; d < 3; ++d )
        {
            nd->mn[d] = l->mn[d]; nd->mx[d] = l->mx[d];
            if ( r->mn[d] < nd->mn[d] ) nd->mn[d] = r->mn[d]; /* Box::include( Box ), MRBox.h */
            if ( r->mx[d] > nd->mx[d] ) nd->mx[d] = r->mx[d];
        }
    }
    free( m->dip ); m->dip = NULL;
    calcDipoles( m );
    return 0;
}

void orc_destroy( void* h )
{
    OMesh* m = (OMesh*)h;
    if ( !m ) return;
    free( m->pts ); free( m->tris ); free( m->nodes ); free( m->dip ); free( m );
}

size_t orc_num_nodes( void* h ) { return (size_t)( (OMesh*)h )->nNodes; }
void orc_get_nodes( void* h, void* out ) { OMesh* m = (OMesh*)h; if ( m->nNodes > 0 ) memcpy( out, m->nodes, sizeof( Node ) * (size_t)m->nNodes ); }
void orc_get_dipoles( void* h, void* out ) { OMesh* m = (OMesh*)h; if ( m->nNodes > 0 ) memcpy( out, m->dip, sizeof( Dipole ) * (size_t)m->nNodes ); }

static void addCounters( uint64_t* dst, const Counters* c )
{
    if ( !dst ) return;
    for ( int k = 0; k < 5; ++k )
    {
#pragma omp atomic
        dst[k] += c->c[k];
    }
}

/* FastWindingNumber::calcFromVector (MRMesh/MRFastWindingNumber.cpp:28-38) */
void orc_calc_from_vector( void* h, const float* pts, size_t n, float beta, int32_t skipFace, float* out, uint64_t* counters )
{
    const OMesh* m = (const OMesh*)h;
#pragma omp parallel
    {
        Counters c; memset( &c, 0, sizeof c );
#pragma omp for schedule( dynamic, 256 )
        for ( long long i = 0; i < (long long)n; ++i )
            out[i] = fastWindingNumber( m, pts + 3 * i, beta, skipFace, &c );
        addCounters( counters, &c );
    }
}

/* FastWindingNumber::calcSelfIntersections (MRFastWindingNumber.cpp:40-52): query triCenter(f), skipFace = f,
 * bit set iff w < 0 || w > 1; out_w optional */
void orc_calc_self_intersections( void* h, float beta, uint64_t* outBits, float* outW )
{
    const OMesh* m = (const OMesh*)h;
    const size_t words = ( m->nf + 63 ) / 64;
    memset( outBits, 0, words * 8 );
#pragma omp parallel for schedule( dynamic, 4 )
    for ( long long wd = 0; wd < (long long)words; ++wd )
    {
        uint64_t bits = 0;
        for ( int b = 0; b < 64; ++b )
        {
            const size_t f = (size_t)wd * 64 + b;
            if ( f >= m->nf ) break;
            const int32_t* t = m->tris + 3 * f;
            float w = 0;
            if ( t[0] >= 0 && t[1] >= 0 && t[2] >= 0 )
            {
                float a[3], bb[3], c[3];
                triPoints( m, (int32_t)f, a, bb, c );
                const float third = 1 / 3.0f;
                const float q[3] = { third * ( ( a[0] + bb[0] ) + c[0] ), third * ( ( a[1] + bb[1] ) + c[1] ), third * ( ( a[2] + bb[2] ) + c[2] ) };
                w = fastWindingNumber( m, q, beta, (int32_t)f, NULL );
                if ( w < 0 || w > 1 ) bits |= (uint64_t)1 << b;
            }
            if ( outW ) outW[f] = w;
        }
        outBits[wd] = bits;
    }
}

/* AffineXf3f::operator() = A*x + b with Matrix3*Vector3 = row dots (MRMesh/MRAffineXf.h:46, MRMatrix3.h:127-130) */
static inline void applyXf( const float* xf, float x, float y, float z, float* out )
{
    out[0] = ( xf[0] * x + xf[1] * y + xf[2] * z ) + xf[9];
    out[1] = ( xf[3] * x + xf[4] * y + xf[5] * z ) + xf[10];
    out[2] = ( xf[6] * x + xf[7] * y + xf[8] * z ) + xf[11];
}

/* FastWindingNumber::calcFromGrid (MRFastWindingNumber.cpp:54-67) for layers [z0,z1); out[0] = voxel (0,0,z0) */
void orc_calc_from_grid( void* h, const int32_t* dims, const float* xf, float beta, int z0, int z1, float* out, uint64_t* counters )
{
    const OMesh* m = (const OMesh*)h;
    const long long dx = dims[0], dy = dims[1];
    const long long total = dx * dy * ( z1 - z0 );
#pragma omp parallel
    {
        Counters c; memset( &c, 0, sizeof c );
#pragma omp for schedule( dynamic, 512 )
        for ( long long i = 0; i < total; ++i )
        {
            const int x = (int)( i % dx ), y = (int)( ( i / dx ) % dy ), z = z0 + (int)( i / ( dx * dy ) );
            float p[3];
            applyXf( xf, (float)x, (float)y, (float)z, p );
            out[i] = fastWindingNumber( m, p, beta, -1, &c );
        }
        addCounters( counters, &c );
    }
}

/* FastWindingNumber::calcWithDistances / calcFromGridWithDistances (MRFastWindingNumber.cpp:69-94), stride: evaluate only
 * voxels whose linear index is a multiple of stride (others left untouched) -- used for bounded samples */
void orc_calc_from_grid_with_distances( void* h, const int32_t* dims, const float* xf, float minDistSq, float maxDistSq,
    int nullOutsideMinMax, float threshold, float beta, int z0, int z1, long long stride, float* out, uint64_t* counters )
{
    const OMesh* m = (const OMesh*)h;
    const long long dx = dims[0], dy = dims[1];
    const long long total = dx * dy * ( z1 - z0 );
    if ( stride < 1 ) stride = 1;
    const long long nSamples = ( total + stride - 1 ) / stride;
#pragma omp parallel
    {
        Counters c; memset( &c, 0, sizeof c );
#pragma omp for schedule( dynamic, 512 )
        for ( long long s = 0; s < nSamples; ++s )
        {
            const long long i = s * stride;
            const int x = (int)( i % dx ), y = (int)( ( i / dx ) % dy ), z = z0 + (int)( i / ( dx * dy ) );
            float p[3];
            applyXf( xf, (float)x, (float)y, (float)z, p );
            const float resSq = findProjectionDistSq( m, p, maxDistSq, minDistSq, &c );
            if ( nullOutsideMinMax && ( resSq < minDistSq || resSq >= maxDistSq ) )
            {
                union { uint32_t u; float f; } nan = { 0x7fc00000u }; /* cQuietNan */
                out[i] = nan.f;
                continue;
            }
            const float sign = fastWindingNumber( m, p, beta, -1, &c ) > threshold ? -1.f : +1.f;
            out[i] = sign * sqrtf( resSq );
        }
        addCounters( counters, &c );
    }
}

/* findProjection(...).distSq for a vector of points (MRMeshProject.cpp:123-127) */
void orc_find_projection_dist_sq( void* h, const float* pts, size_t n, float up, float lo, float* out )
{
    const OMesh* m = (const OMesh*)h;
#pragma omp parallel for schedule( dynamic, 256 )
    for ( long long i = 0; i < (long long)n; ++i )
        out[i] = findProjectionDistSq( m, pts + 3 * i, up, lo, NULL );
}

float orc_triangle_solid_angle( const float* p, const float* tri9 ) { return triangleSolidAngle( p, tri9, tri9 + 3, tri9 + 6 ); }

/* chunkCount / ChunkIterator::operator* (MRMesh/MRChunkIterator.cpp:25-49) */
size_t orc_chunk_count( size_t totalSize, size_t chunkSize, size_t overlap )
{
    if ( totalSize == 0 || chunkSize == 0 || chunkSize <= overlap ) return 0;
    const size_t size = totalSize - overlap, step = chunkSize - overlap;
    return ( size + step - 1 ) / step;
}
void orc_chunk( size_t totalSize, size_t chunkSize, size_t overlap, size_t index, size_t* offset, size_t* size )
{
    *offset = index * ( chunkSize - overlap );
    *size = chunkSize < totalSize - *offset ? chunkSize : totalSize - *offset;
}


void orc_find_projections( void* h, const float* pts, size_t n, const float* pointXf12, const float* treeXf12, float up, float lo, void* out )
{
    const OMesh* m = (const OMesh*)h;
    OProj* o = (OProj*)out;
    for ( size_t i = 0; i < n; ++i )
    {
        float q[3] = { pts[3 * i], pts[3 * i + 1], pts[3 * i + 2] };
        if ( pointXf12 ) { float t[3]; xfApply( pointXf12, q, t ); memcpy( q, t, 12 ); }
        o[i] = findProjectionFull( m, q, treeXf12, up, lo );
    }
}
