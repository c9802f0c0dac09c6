// tests/brick_walk_host.cpp -- TEST INFRASTRUCTURE.  Host replay of the v6 winding-number walk of meshlib_b200/csrc/b2fwn_grid.cu
// (kGridFwn6): same brick layout (8 x 8 x 4 voxels per warp), same brick-level tests (the product header b2fwn_brick_math.h,
// compiled here for the host), same per-voxel decisions and term arithmetic, operation by operation in float
// (-ffp-contract=off, explicit fmaf).  Only rsqrt.approx is replaced by 1 / sqrtf.  Used by tests/test_brick_math.py to check,
// without a GPU, that (a) the uniform answers ALL_NEAR / ALL_FAR never contradict the exactly rounded per-voxel decision of
// Dipole::addIfGoodApprox (source/MRMesh/MRDipole.h:18-27) and (b) the two-mode walk gives every voxel the reference's winding
// number.  Node / dipole layouts are the reference's (AABBTreeNode 32 B, Dipole 32 B) as the oracle hands them out.
#include "../meshlib_b200/csrc/b2fwn_brick_math.h"
#include <climits>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace
{
struct Node { float mn[3], mx[3]; int32_t l, r; };
struct Dipole { float pos[3], area, dirArea[3], rr; };

inline float dot3( float ax, float ay, float az, float bx, float by, float bz ) { return ( ax * bx + ay * by ) + az * bz; }
// AffineXf3f::operator() with the operation order of b2fwn_math.cuh applyXf
inline void applyXf( const float* xf, float x, float y, float z, float* o )
{
    o[0] = dot3( xf[0], xf[1], xf[2], x, y, z ) + xf[9];
    o[1] = dot3( xf[3], xf[4], xf[5], x, y, z ) + xf[10];
    o[2] = dot3( xf[6], xf[7], xf[8], x, y, z ) + xf[11];
}
float triangleSolidAngle( const float* p, const float* ta, const float* tb, const float* tc )
{
    const float a[3] = { ta[0] - p[0], ta[1] - p[1], ta[2] - p[2] };
    const float b[3] = { tb[0] - p[0], tb[1] - p[1], tb[2] - p[2] };
    const float c[3] = { tc[0] - p[0], tc[1] - p[1], tc[2] - p[2] };
    const float x = sqrtf( dot3( a[0], a[1], a[2], a[0], a[1], a[2] ) ), y = sqrtf( dot3( b[0], b[1], b[2], b[0], b[1], b[2] ) ),
                z = sqrtf( dot3( c[0], c[1], c[2], c[0], c[1], c[2] ) );
    float den = ( x * y ) * z;
    den = den + dot3( a[0], a[1], a[2], b[0], b[1], b[2] ) * z;
    den = den + dot3( b[0], b[1], b[2], c[0], c[1], c[2] ) * x;
    den = den + dot3( c[0], c[1], c[2], a[0], a[1], a[2] ) * y;
    const float t0 = a[0] * ( b[1] * c[2] - b[2] * c[1] ), t1 = a[1] * ( b[0] * c[2] - b[2] * c[0] ), t2 = a[2] * ( b[0] * c[1] - b[1] * c[0] );
    return 2.f * atan2f( ( t0 - t1 ) + t2, den );
}
} // namespace

extern "C"
{

// counters: [0] uniform ALL_NEAR visits [1] uniform ALL_FAR visits [2] diverged sections [3] visits inside sections
// [4] ALL_NEAR answers contradicted by a voxel [5] ALL_FAR answers contradicted by a voxel (both must stay 0)
// valid: optional dims[0]*dims[1]*dims[2] bytes, 0 = the voxel takes no part (NaN distance in mode 1)
void bw_calc_from_grid( const Node* nodes, const Dipole* dip, int N, const float* points, const int32_t* tris, const int32_t* dims,
                        const float* xf12, float beta, const uint8_t* valid, float* outW, uint64_t* counters )
{
    using namespace b2;
    const float betaSq = beta * beta;
    std::vector<int> skip( N > 0 ? N : 1 );
    for ( int i = N - 1; i >= 0; --i ) skip[i] = nodes[i].r < 0 ? i + 1 : skip[nodes[i].r];
    const int FXv = 8, BY = 8, BZ = 4;
    const int nbx = ( dims[0] + 7 ) / 8, nby = ( dims[1] + BY - 1 ) / BY, nbz = ( dims[2] + BZ - 1 ) / BZ;
    for ( int bz = 0; bz < nbz; ++bz ) for ( int by = 0; by < nby; ++by ) for ( int bx = 0; bx < nbx; ++bx )
    {
        // lanes: y = by*8 + (lane & 7), z = bz*4 + (lane >> 3); voxels x0 .. x0+7
        float px[32][8], py[32], pz[32], sum[32][8];
        bool ok[32][8];
        int fR[32][8];
        const int x0 = bx * 8;
        float pc[3];
        applyXf( xf12, (float)x0 + 3.5f, (float)( by * BY ) + 3.5f, (float)( bz * BZ ) + 1.5f, pc );
        float radSq = 0.f;
        bool any = false;
        for ( int lane = 0; lane < 32; ++lane )
        {
            const int y = by * BY + ( lane & 7 ), z = bz * BZ + ( lane >> 3 );
            float p0[3];
            applyXf( xf12, (float)x0, (float)y, (float)z, p0 );
            py[lane] = p0[1]; pz[lane] = p0[2];
            const float ey = py[lane] - pc[1], ez = pz[lane] - pc[2];
            const float eyz = fmaf( ey, ey, ez * ez );
            for ( int v = 0; v < FXv; ++v )
            {
                float p[3];
                applyXf( xf12, (float)( x0 + v ), (float)y, (float)z, p );
                px[lane][v] = p[0];
                sum[lane][v] = 0.f;
                bool in = y < dims[1] && z < dims[2] && x0 + v < dims[0];
                if ( in && valid ) in = valid[(size_t)( x0 + v ) + (size_t)dims[0] * ( y + (size_t)dims[1] * z )] != 0;
                ok[lane][v] = in;
                if ( in )
                {
                    any = true;
                    const float ex = px[lane][v] - pc[0];
                    radSq = fmaxf( radSq, fmaf( ex, ex, eyz ) );
                }
            }
        }
        if ( !any ) continue;
        const BrickBall ball = makeBrickBall( pc[0], pc[1], pc[2], radSq );

        // the term of one voxel, as termPair / fwn6AddAll compute it
        auto term = [&]( int lane, int v, const Dipole& d, float& dot, float& r3, float& ddc, float& ddExact )
        {
            const float dy = d.pos[1] - py[lane], dz = d.pos[2] - pz[lane];
            const float dy2 = dy * dy, dz2 = dz * dz;
            const float syz = fmaf( dy, dy, dz2 );
            const float yz = fmaf( dy, d.dirArea[1], dz * d.dirArea[2] );
            const float dx = d.pos[0] - px[lane][v];
            ddExact = ( dx * dx + dy2 ) + dz2;
            ddc = fmaf( dx, dx, syz );
            const float r = 1.f / sqrtf( ddc );
            r3 = ( r * r ) * r;
            dot = fmaf( dx, d.dirArea[0], yz );
        };

        int i = 0;
        while ( i < N )
        {
            const Dipole& d = dip[i];
            const bool leaf = nodes[i].r < 0;
            const float thr = betaSq * d.rr;
            const float s = brickDistSq( ball, d.pos[0], d.pos[1], d.pos[2] ), rhs = s * ball.r4;
            const bool near = brickAllNear( ball, s, rhs, thr ), far = brickAllFar( ball, s, rhs, thr );
            if ( near || far )
            {
                // brute-force audit of the uniform answer against the exactly rounded decision of every participating voxel
                for ( int lane = 0; lane < 32; ++lane ) for ( int v = 0; v < FXv; ++v )
                {
                    if ( !ok[lane][v] ) continue;
                    float dot, r3, ddc, dd;
                    term( lane, v, d, dot, r3, ddc, dd );
                    if ( near && !( dd <= thr ) ) ++counters[4];
                    if ( far && near == false && dd <= thr ) ++counters[5];
                }
            }
            if ( !leaf && near ) { ++counters[0]; ++i; continue; }
            const int next = skip[i];
            if ( far )
            {
                ++counters[1];
                for ( int lane = 0; lane < 32; ++lane ) for ( int v = 0; v < FXv; ++v )
                {
                    float dot, r3, ddc, dd;
                    term( lane, v, d, dot, r3, ddc, dd );
                    sum[lane][v] = fmaf( dot, r3, sum[lane][v] );
                }
                i = next;
                continue;
            }
            ++counters[2];
            for ( int lane = 0; lane < 32; ++lane ) for ( int v = 0; v < FXv; ++v ) fR[lane][v] = ok[lane][v] ? i : INT_MAX;
            while ( i < next )
            {
                ++counters[3];
                const Dipole& e = dip[i];
                const bool leaf2 = nodes[i].r < 0;
                const int link = skip[i];
                const float thr2 = betaSq * e.rr;
                int m = INT_MAX;
                for ( int lane = 0; lane < 32; ++lane ) for ( int v = 0; v < FXv; ++v )
                {
                    float dot, r3, ddc, dd;
                    term( lane, v, e, dot, r3, ddc, dd );
                    const bool a = i >= fR[lane][v], farV = !( dd <= thr2 );
                    if ( !leaf2 )
                    {
                        if ( a && farV )
                        {
                            if ( thr2 >= 2.f * FLT_MIN || ( dd >= FLT_MIN && ddc >= FLT_MIN ) ) sum[lane][v] = fmaf( dot, r3, sum[lane][v] );
                            fR[lane][v] = link;
                        }
                    }
                    else
                    {
                        if ( a && farV && dd >= FLT_MIN && ddc >= FLT_MIN ) sum[lane][v] = fmaf( dot, r3, sum[lane][v] );
                        if ( a && !farV )
                        {
                            const int32_t f = nodes[i].l;
                            const float p[3] = { px[lane][v], py[lane], pz[lane] };
                            sum[lane][v] = sum[lane][v] + triangleSolidAngle( p, points + 3 * tris[3 * f], points + 3 * tris[3 * f + 1], points + 3 * tris[3 * f + 2] );
                        }
                        if ( a ) fR[lane][v] = i + 1;
                    }
                    if ( fR[lane][v] < m ) m = fR[lane][v];
                }
                i = m > i + 1 ? m : i + 1;
            }
            i = next;
        }
        const float INV_4PI = 1.0f / ( 4 * 3.14159265358979323846f );
        for ( int lane = 0; lane < 32; ++lane ) for ( int v = 0; v < FXv; ++v )
        {
            const int y = by * BY + ( lane & 7 ), z = bz * BZ + ( lane >> 3 );
            if ( !( y < dims[1] && z < dims[2] && x0 + v < dims[0] ) ) continue;
            const size_t id = (size_t)( x0 + v ) + (size_t)dims[0] * ( y + (size_t)dims[1] * z );
            if ( ok[lane][v] ) outW[id] = INV_4PI * sum[lane][v];
        }
    }
}

// random audit of the brick tests alone: n balls / nodes drawn around the given scales; returns the number of wrong uniform answers
uint64_t bw_audit_brick_tests( uint64_t seed, uint64_t n, float coordScale, float voxel, uint64_t* nearCount, uint64_t* farCount )
{
    using namespace b2;
    uint64_t bad = 0, s64 = seed;
    auto rnd = [&]() { s64 = s64 * 6364136223846793005ull + 1442695040888963407ull; return (float)( ( s64 >> 40 ) & 0xFFFFFF ) / 16777216.f; };
    for ( uint64_t it = 0; it < n; ++it )
    {
        // brick of 8 x 8 x 4 voxel centres at a random place
        const float ox = ( rnd() - 0.5f ) * coordScale, oy = ( rnd() - 0.5f ) * coordScale, oz = ( rnd() - 0.5f ) * coordScale;
        float P[256][3];
        int k = 0;
        for ( int z = 0; z < 4; ++z ) for ( int y = 0; y < 8; ++y ) for ( int x = 0; x < 8; ++x, ++k )
        {
            P[k][0] = voxel * (float)x + ox; P[k][1] = voxel * (float)y + oy; P[k][2] = voxel * (float)z + oz;
        }
        const float pc[3] = { voxel * 3.5f + ox, voxel * 3.5f + oy, voxel * 1.5f + oz };
        float radSq = 0.f;
        for ( k = 0; k < 256; ++k )
        {
            const float ex = P[k][0] - pc[0], ey = P[k][1] - pc[1], ez = P[k][2] - pc[2];
            radSq = fmaxf( radSq, fmaf( ex, ex, fmaf( ey, ey, ez * ez ) ) );
        }
        const BrickBall ball = makeBrickBall( pc[0], pc[1], pc[2], radSq );
        const float rad = sqrtf( radSq );
        // a node whose beta-sphere passes close to the brick: distance and threshold chosen so that sqrt(thr) ~ dist +- 1.5 rad
        const float dist = rad * ( 0.2f + 40.f * rnd() * rnd() );
        float dir[3] = { rnd() - 0.5f, rnd() - 0.5f, rnd() - 0.5f };
        const float dl = sqrtf( dir[0] * dir[0] + dir[1] * dir[1] + dir[2] * dir[2] ) + 1e-9f;
        const float q[3] = { pc[0] + dist * dir[0] / dl, pc[1] + dist * dir[1] / dl, pc[2] + dist * dir[2] / dl };
        const float st = fmaxf( dist + rad * ( 3.f * rnd() - 1.5f ), 0.f );
        // snap the threshold onto the edge of the uniform regions every other time (where a missing margin would show)
        float thr = st * st;
        if ( it & 1 )
        {
            const float e = ( it & 2 ) ? dist + rad : fmaxf( dist - rad, 0.f );
            thr = e * e * ( 1.f + ( rnd() - 0.5f ) * 1e-4f );
        }
        const int cls = brickClassify( ball, q[0], q[1], q[2], thr );
        if ( cls == BRICK_MIXED ) continue;
        if ( cls == BRICK_ALL_NEAR ) ++*nearCount; else ++*farCount;
        for ( k = 0; k < 256; ++k )
        {
            const float dx = q[0] - P[k][0], dy = q[1] - P[k][1], dz = q[2] - P[k][2];
            const float dd = ( dx * dx + dy * dy ) + dz * dz;
            const bool nearV = dd <= thr;
            if ( ( cls == BRICK_ALL_NEAR ) != nearV ) { ++bad; break; }
        }
    }
    return bad;
}

} // extern "C"
