// b2fwn_math.cuh -- per-query arithmetic of the path, written so that every float/double operation rounds as in
// the reference's scalar CPU code (explicit _rn intrinsics = no FMA contraction).  With an identical tree and an
// identical summation order this makes the winding number differ from the CPU reference only through atan2f, and
// the distance not at all.
//   Dipole::addIfGoodApprox        source/MRMesh/MRDipole.h:18-27
//   triangleSolidAngle             source/MRMesh/MRDipole.cpp:81-92   (Matrix3::det: MRMatrix3.h:227-233)
//   Box::getDistanceSq(pt)         source/MRMesh/MRBox.h:207-220
//   closestPointInTriangle<double> source/MRMesh/MRClosestPointInTriangle.h:27-87, called at MRMeshProject.cpp:68-74
#pragma once
#include <cuda_runtime.h>

namespace b2
{

#define FM( a, b ) __fmul_rn( ( a ), ( b ) )
#define FA( a, b ) __fadd_rn( ( a ), ( b ) )
#define FS( a, b ) __fsub_rn( ( a ), ( b ) )
#define DM( a, b ) __dmul_rn( ( a ), ( b ) )
#define DA( a, b ) __dadd_rn( ( a ), ( b ) )
#define DS( a, b ) __dsub_rn( ( a ), ( b ) )

__device__ __forceinline__ float dot3( float ax, float ay, float az, float bx, float by, float bz )
{
    return FA( FA( FM( ax, bx ), FM( ay, by ) ), FM( az, bz ) );
}
__device__ __forceinline__ double ddot3( double ax, double ay, double az, double bx, double by, double bz )
{
    return DA( DA( DM( ax, bx ), DM( ay, by ) ), DM( az, bz ) );
}

// AffineXf3f::operator() : A * x + b  (MRAffineXf.h:46, MRMatrix3.h:127-130)
__device__ __forceinline__ float3 applyXf( const float* __restrict__ A, const float* __restrict__ b, float x, float y, float z )
{
    float3 r;
    r.x = FA( dot3( A[0], A[1], A[2], x, y, z ), b[0] );
    r.y = FA( dot3( A[3], A[4], A[5], x, y, z ), b[1] );
    r.z = FA( dot3( A[6], A[7], A[8], x, y, z ), b[2] );
    return r;
}

// squared distance from p to the box {q0.xyz, (q0.w, q1.x, q1.y)}
__device__ __forceinline__ float boxDistSq( const float3 p, const float4 q0, const float4 q1 )
{
    float res = 0.f;
    if ( p.x < q0.x ) { const float d = FS( p.x, q0.x ); res = FA( res, FM( d, d ) ); }
    else if ( p.x > q0.w ) { const float d = FS( p.x, q0.w ); res = FA( res, FM( d, d ) ); }
    if ( p.y < q0.y ) { const float d = FS( p.y, q0.y ); res = FA( res, FM( d, d ) ); }
    else if ( p.y > q1.x ) { const float d = FS( p.y, q1.x ); res = FA( res, FM( d, d ) ); }
    if ( p.z < q0.z ) { const float d = FS( p.z, q0.z ); res = FA( res, FM( d, d ) ); }
    else if ( p.z > q1.y ) { const float d = FS( p.z, q1.y ); res = FA( res, FM( d, d ) ); }
    return res;
}

// returns true when the dipole {q2 = pos.xyz, rr; q3 = dirArea.xyz} is a good approximation at q and adds its term
__device__ __forceinline__ bool dipoleAddIfGoodApprox( const float3 q, const float betaSq, const float4 q2, const float4 q3, float& addTo )
{
    const float dx = FS( q2.x, q.x ), dy = FS( q2.y, q.y ), dz = FS( q2.z, q.z );
    const float dd = dot3( dx, dy, dz, dx, dy, dz );
    if ( dd <= FM( betaSq, q2.w ) )
        return false;
    const float d = __fsqrt_rn( dd );
    if ( d > 0 )
        addTo = FA( addTo, __fdiv_rn( dot3( dx, dy, dz, q3.x, q3.y, q3.z ), FM( d, dd ) ) );
    return true;
}

__device__ __forceinline__ float triangleSolidAngle( const float3 p, const float4 ta, const float4 tb, const float4 tc )
{
    const float ax = FS( ta.x, p.x ), ay = FS( ta.y, p.y ), az = FS( ta.z, p.z );
    const float bx = FS( tb.x, p.x ), by = FS( tb.y, p.y ), bz = FS( tb.z, p.z );
    const float cx = FS( tc.x, p.x ), cy = FS( tc.y, p.y ), cz = FS( tc.z, p.z );
    const float x = __fsqrt_rn( dot3( ax, ay, az, ax, ay, az ) );
    const float y = __fsqrt_rn( dot3( bx, by, bz, bx, by, bz ) );
    const float z = __fsqrt_rn( dot3( cx, cy, cz, cx, cy, cz ) );
    // den = x*y*z + dot(a,b)*z + dot(b,c)*x + dot(c,a)*y
    float den = FM( FM( x, y ), z );
    den = FA( den, FM( dot3( ax, ay, az, bx, by, bz ), z ) );
    den = FA( den, FM( dot3( bx, by, bz, cx, cy, cz ), x ) );
    den = FA( den, FM( dot3( cx, cy, cz, ax, ay, az ), y ) );
    // det of rows a, b, c
    const float t0 = FM( ax, FS( FM( by, cz ), FM( bz, cy ) ) );
    const float t1 = FM( ay, FS( FM( bx, cz ), FM( bz, cx ) ) );
    const float t2 = FM( az, FS( FM( bx, cy ), FM( by, cx ) ) );
    const float det = FA( FS( t0, t1 ), t2 );
    return FM( 2.f, atan2f( det, den ) );
}

// closestPointInTriangle<double> (MRClosestPointInTriangle.h:27-87) with its barycentric pair; same operations, same order
__device__ __forceinline__ void closestPointBary( const float3 pf, const float3 fa, const float3 fb, const float3 fc,
                                                  double& rx, double& ry, double& rz, double& ba, double& bb )
{
    const double px = pf.x, py = pf.y, pz = pf.z;
    const double ax = fa.x, ay = fa.y, az = fa.z, bx = fb.x, by = fb.y, bz = fb.z, cx = fc.x, cy = fc.y, cz = fc.z;
    const double abx = DS( bx, ax ), aby = DS( by, ay ), abz = DS( bz, az );
    const double acx = DS( cx, ax ), acy = DS( cy, ay ), acz = DS( cz, az );
    const double apx = DS( px, ax ), apy = DS( py, ay ), apz = DS( pz, az );
    const double d1 = ddot3( abx, aby, abz, apx, apy, apz );
    const double d2 = ddot3( acx, acy, acz, apx, apy, apz );
    if ( d1 <= 0 && d2 <= 0 ) { rx = ax; ry = ay; rz = az; ba = 0; bb = 0; return; } // #1
    const double bpx = DS( px, bx ), bpy = DS( py, by ), bpz = DS( pz, bz );
    const double d3 = ddot3( abx, aby, abz, bpx, bpy, bpz );
    const double d4 = ddot3( acx, acy, acz, bpx, bpy, bpz );
    if ( d3 >= 0 && d4 <= d3 ) { rx = bx; ry = by; rz = bz; ba = 1; bb = 0; return; } // #2
    const double cpx = DS( px, cx ), cpy = DS( py, cy ), cpz = DS( pz, cz );
    const double d5 = ddot3( abx, aby, abz, cpx, cpy, cpz );
    const double d6 = ddot3( acx, acy, acz, cpx, cpy, cpz );
    if ( d6 >= 0 && d5 <= d6 ) { rx = cx; ry = cy; rz = cz; ba = 0; bb = 1; return; } // #3
    const double vc = DS( DM( d1, d4 ), DM( d3, d2 ) );
    if ( vc <= 0 && d1 >= 0 && d3 <= 0 )
    {
        const double v = d1 / DS( d1, d3 );
        rx = DA( ax, DM( v, abx ) ); ry = DA( ay, DM( v, aby ) ); rz = DA( az, DM( v, abz ) ); ba = v; bb = 0; return; // #4
    }
    const double vb = DS( DM( d5, d2 ), DM( d1, d6 ) );
    if ( vb <= 0 && d6 <= 0 && d2 >= 0 )
    {
        const double v = d2 / DS( d2, d6 );
        rx = DA( ax, DM( v, acx ) ); ry = DA( ay, DM( v, acy ) ); rz = DA( az, DM( v, acz ) ); ba = 0; bb = v; return; // #5
    }
    const double va = DS( DM( d3, d6 ), DM( d5, d4 ) );
    if ( va <= 0 )
    {
        if ( d4 < d3 ) { rx = bx; ry = by; rz = bz; ba = 1; bb = 0; return; }
        if ( d5 < d6 ) { rx = cx; ry = cy; rz = cz; ba = 0; bb = 1; return; }
        const double e = DS( d4, d3 ), g = DS( d5, d6 );
        const double v = e / DA( e, g );
        rx = DA( bx, DM( v, DS( cx, bx ) ) ); ry = DA( by, DM( v, DS( cy, by ) ) ); rz = DA( bz, DM( v, DS( cz, bz ) ) );
        ba = DS( 1.0, v ); bb = v; return; // #6
    }
    const double denom = 1.0 / DA( DA( va, vb ), vc );
    const double v = DM( vb, denom ), w = DM( vc, denom );
    rx = DA( DA( ax, DM( v, abx ) ), DM( w, acx ) );
    ry = DA( DA( ay, DM( v, aby ) ), DM( w, acy ) );
    rz = DA( DA( az, DM( v, abz ) ), DM( w, acz ) ); // #0
    ba = v; bb = w;
}

// squared distance from p to triangle (a,b,c) exactly as the reference computes it: closest point in DOUBLE,
// rounded to float, difference and squared length in float.
__device__ __forceinline__ float triDistSqExact( const float3 pf, const float4 ta, const float4 tb, const float4 tc )
{
    const double px = pf.x, py = pf.y, pz = pf.z;
    const double ax = ta.x, ay = ta.y, az = ta.z;
    const double bx = tb.x, by = tb.y, bz = tb.z;
    const double cx = tc.x, cy = tc.y, cz = tc.z;
    const double abx = DS( bx, ax ), aby = DS( by, ay ), abz = DS( bz, az );
    const double acx = DS( cx, ax ), acy = DS( cy, ay ), acz = DS( cz, az );
    const double apx = DS( px, ax ), apy = DS( py, ay ), apz = DS( pz, az );
    double rx, ry, rz;
    const double d1 = ddot3( abx, aby, abz, apx, apy, apz );
    const double d2 = ddot3( acx, acy, acz, apx, apy, apz );
    bool done = false;
    if ( d1 <= 0 && d2 <= 0 ) { rx = ax; ry = ay; rz = az; done = true; } // #1
    double d3 = 0, d4 = 0, d5 = 0, d6 = 0;
    if ( !done )
    {
        const double bpx = DS( px, bx ), bpy = DS( py, by ), bpz = DS( pz, bz );
        d3 = ddot3( abx, aby, abz, bpx, bpy, bpz );
        d4 = ddot3( acx, acy, acz, bpx, bpy, bpz );
        if ( d3 >= 0 && d4 <= d3 ) { rx = bx; ry = by; rz = bz; done = true; } // #2
    }
    if ( !done )
    {
        const double cpx = DS( px, cx ), cpy = DS( py, cy ), cpz = DS( pz, cz );
        d5 = ddot3( abx, aby, abz, cpx, cpy, cpz );
        d6 = ddot3( acx, acy, acz, cpx, cpy, cpz );
        if ( d6 >= 0 && d5 <= d6 ) { rx = cx; ry = cy; rz = cz; done = true; } // #3
    }
    if ( !done )
    {
        const double vc = DS( DM( d1, d4 ), DM( d3, d2 ) );
        if ( vc <= 0 && d1 >= 0 && d3 <= 0 )
        {
            const double v = d1 / DS( d1, d3 );
            rx = DA( ax, DM( v, abx ) ); ry = DA( ay, DM( v, aby ) ); rz = DA( az, DM( v, abz ) ); // #4
            done = true;
        }
        if ( !done )
        {
            const double vb = DS( DM( d5, d2 ), DM( d1, d6 ) );
            if ( vb <= 0 && d6 <= 0 && d2 >= 0 )
            {
                const double v = d2 / DS( d2, d6 );
                rx = DA( ax, DM( v, acx ) ); ry = DA( ay, DM( v, acy ) ); rz = DA( az, DM( v, acz ) ); // #5
                done = true;
            }
            if ( !done )
            {
                const double va = DS( DM( d3, d6 ), DM( d5, d4 ) );
                if ( va <= 0 )
                {
                    if ( d4 < d3 ) { rx = bx; ry = by; rz = bz; }
                    else if ( d5 < d6 ) { rx = cx; ry = cy; rz = cz; }
                    else
                    {
                        const double e = DS( d4, d3 ), g = DS( d5, d6 );
                        const double v = e / DA( e, g );
                        rx = DA( bx, DM( v, DS( cx, bx ) ) ); ry = DA( by, DM( v, DS( cy, by ) ) ); rz = DA( bz, DM( v, DS( cz, bz ) ) ); // #6
                    }
                }
                else
                {
                    const double denom = 1.0 / DA( DA( va, vb ), vc );
                    const double v = DM( vb, denom ), w = DM( vc, denom );
                    // a + v * ab + w * ac   (left to right)
                    rx = DA( DA( ax, DM( v, abx ) ), DM( w, acx ) );
                    ry = DA( DA( ay, DM( v, aby ) ), DM( w, acy ) );
                    rz = DA( DA( az, DM( v, abz ) ), DM( w, acz ) ); // #0
                }
            }
        }
    }
    const float fx = FS( __double2float_rn( rx ), pf.x ), fy = FS( __double2float_rn( ry ), pf.y ), fz = FS( __double2float_rn( rz ), pf.z );
    return dot3( fx, fy, fz, fx, fy, fz );
}


// ---- branch-free / throughput forms used by the brick kernel ---------------------------------------------------------

// Box::getDistanceSq, branch-free and bit-identical to boxDistSq above: per axis d = max(lo - p, p - hi, 0) has the
// same magnitude as the reference's (p - lo) or (p - hi), and adding the exact zeros of the untouched axes changes nothing
__device__ __forceinline__ float boxDistSqBF( const float3 p, const float4 q0, const float4 q1 )
{
    const float dx = fmaxf( fmaxf( FS( q0.x, p.x ), FS( p.x, q0.w ) ), 0.f );
    const float dy = fmaxf( fmaxf( FS( q0.y, p.y ), FS( p.y, q1.x ) ), 0.f );
    const float dz = fmaxf( fmaxf( FS( q0.z, p.z ), FS( p.z, q1.y ) ), 0.f );
    return FA( FA( FM( dx, dx ), FM( dy, dy ) ), FM( dz, dz ) );
}

// far-field test of Dipole::addIfGoodApprox with the reference's exact rounding for the DECISION (dd <= beta^2 rr),
// and a throughput form for the accepted TERM: dot(dp, dirArea) * rsqrt(dd)^3 on the MUFU unit
// (relative error <= ~1e-6 per term against the reference's dot / (sqrt(dd) * dd); the sign contract is |w - thr| >= 1e-3)
__device__ __forceinline__ bool dipoleFar( const float3 q, const float thr, const float4 q2, float& dx, float& dy, float& dz, float& dd )
{
    dx = FS( q2.x, q.x ); dy = FS( q2.y, q.y ); dz = FS( q2.z, q.z );
    dd = dot3( dx, dy, dz, dx, dy, dz );
    return !( dd <= thr );
}
__device__ __forceinline__ float rsqrtApprox( const float x )
{
    float r;
    asm( "rsqrt.approx.ftz.f32 %0, %1;" : "=f"( r ) : "f"( x ) ); // one MUFU.RSQ, <= 2^-22 relative error
    return r;
}
// only meaningful for dd >= FLT_MIN (the caller masks the rest out, as the reference's `d > 0` guard does)
__device__ __forceinline__ float dipoleTermFast( const float dx, const float dy, const float dz, const float dd, const float4 q3 )
{
    const float r = rsqrtApprox( dd );
    return fmaf( dx, q3.x, fmaf( dy, q3.y, dz * q3.z ) ) * ( r * r * r );
}

// FP32 estimate of the distance from p to a triangle given in its own frame (LeafFilt): decides whether a candidate can
// possibly beat the running best before the reference's double-precision closest point is spent on it.
//   w = p - a;  (x, y, h) = (w.u, w.v, w.n);  d^2 = h^2 + (2-D distance from (x, y) to the triangle (0,0) (bx,0) (cx,cy))^2
// Every quantity is a first-order function of w with error <= ~4e-7 |w| (frame vectors are double-computed and rounded once;
// no barycentric solve, so nothing is ill-conditioned for far points or thin triangles); the 2-D point-segment distances add
// a few ulp of (|w| + edge length).  Allowance used by the caller, in DISTANCE units:
//   m = 3e-6 |w|_1 + f3.z + pmag,  f3.z = 3e-6 (|ab| + |ac|) + 1e-6 |tri|_1,  pmag = 1e-6 |p|_1
// (>= 3x the bound above, plus the reference's own float rounding of the projection, <= 2^-24 * coordinate magnitude per
// axis).  A candidate is rejected only if estimate - m >= sqrt(best); rejected candidates cannot be the minimum, so the
// result stays the reference's value bit for bit.
__device__ __forceinline__ bool estimateRejects( const float3 p, const float pmag, const float sqrtBest,
                                                 const float4 f0, const float4 f1, const float4 f2, const float4 f3, const float3 vv )
{
    const float wx = p.x - f0.x, wy = p.y - f0.y, wz = p.z - f0.z;
    const float x = fmaf( wx, f1.x, fmaf( wy, f1.y, wz * f1.z ) );
    const float y = fmaf( wx, vv.x, fmaf( wy, vv.y, wz * vv.z ) );
    const float h = fmaf( wx, f2.x, fmaf( wy, f2.y, wz * f2.z ) );
    const float bx = f0.w, cx = f1.w, cy = f2.w;
    // edge a-b (the x axis)
    const float tab = fminf( fmaxf( x, 0.f ), bx );
    const float dab = fmaf( x - tab, x - tab, y * y );
    // edge b-c
    const float ex = cx - bx, xb = x - bx;
    const float sbc = __saturatef( fmaf( xb, ex, y * cy ) * f3.x );
    const float rbx = fmaf( -sbc, ex, xb ), rby = fmaf( -sbc, cy, y );
    const float dbc = fmaf( rbx, rbx, rby * rby );
    // edge a-c
    const float sca = __saturatef( fmaf( x, cx, y * cy ) * f3.y );
    const float rcx = fmaf( -sca, cx, x ), rcy = fmaf( -sca, cy, y );
    const float dca = fmaf( rcx, rcx, rcy * rcy );
    // inside the (counter-clockwise) triangle?
    const bool inside = y >= 0.f && fmaf( ex, y, -cy * xb ) >= 0.f && fmaf( cy, x, -cx * y ) >= 0.f;
    const float ip = inside ? 0.f : fminf( dab, fminf( dbc, dca ) );
    const float d2 = fmaf( h, h, ip );
    const float m = fmaf( 3e-6f, fabsf( wx ) + fabsf( wy ) + fabsf( wz ), f3.z + pmag );
    const float lim = sqrtBest + m;
    return d2 >= lim * lim;
}

} // namespace b2
