// b2fwn_brick_math.h -- brick-level far-field tests of the winding-number walk (host/device).
// A brick is a block of voxels walked by one warp; its voxel centres lie inside the ball (centre c, radius rad).  For a tree
// node with dipole position q and threshold thr = beta^2 * rr (Dipole::addIfGoodApprox, source/MRMesh/MRDipole.h:18-27: the
// dipole is used when dd > thr, dd = |q - p|^2 in float) the walk asks once per node, for the whole brick:
//   ALL_NEAR: dd <= thr for EVERY voxel  -> nobody accepts the dipole, everybody descends
//   ALL_FAR : dd >  thr for EVERY voxel  -> everybody accepts the dipole, nobody descends
//   MIXED   : anything else              -> the voxels decide one by one with the reference's exactly rounded dd
// Both uniform answers must be CONSERVATIVE against the exactly rounded per-voxel decision: they may say MIXED too often,
// never ALL_NEAR / ALL_FAR wrongly -- then every voxel accepts exactly the reference's set of dipoles.
//
// No square root: with s = |q - c|^2
//   (sqrt(s) + rad)^2 <= thr  <=>  u >= 0 and u^2 >= 4 rad^2 s,  u = thr - s - rad^2
//   (sqrt(s) - rad)^2 >  thr  <=>  w >  0 and w^2 >  4 rad^2 s,  w = s + rad^2 - thr      (and s > rad^2: the ball does not contain q)
// Every term of u and w carries a relative margin of BRICK_EPS = 1e-5 towards MIXED: the float error of the few operations
// here (<= 4 * 2^-24 of the largest term) and of the reference's own dd (<= 3 * 2^-24 dd) is a hundred times smaller, and
// rad itself is padded by 1.5e-5 against the rounding of the voxel centres.  tests/test_brick_math.py checks the claim by
// brute force against the exactly rounded decision of every voxel.
#pragma once
#include <math.h>
#include <float.h>
#ifdef __CUDACC__
#define B2_HD __host__ __device__ __forceinline__
#else
#define B2_HD inline
#endif

namespace b2
{

constexpr float BRICK_EPS = 1e-5f;
enum BrickClass { BRICK_MIXED = 0, BRICK_ALL_NEAR = 1, BRICK_ALL_FAR = 2 };

struct BrickBall
{
    float cx, cy, cz;   // centre
    float r2p, r2m, r4; // rad^2 (1 + eps), rad^2 (1 - eps), 4 rad^2 (1 + 3 eps)  -- rad already padded
};

// radSqMax = max over the voxels of |p - c|^2 as computed in float
B2_HD BrickBall makeBrickBall( const float cx, const float cy, const float cz, const float radSqMax )
{
    BrickBall B;
    B.cx = cx; B.cy = cy; B.cz = cz;
    const float r2 = radSqMax * 1.00003f; // radius padded by 1.5e-5
    B.r2p = r2 * ( 1.f + BRICK_EPS );
    B.r2m = r2 * ( 1.f - BRICK_EPS );
    B.r4 = 4.f * r2 * ( 1.f + 3.f * BRICK_EPS );
    return B;
}

// the two uniform answers as predicates over s = |q - c|^2 and rhs = s * B.r4 (the walk kernel branches on them one after the other)
B2_HD float brickDistSq( const BrickBall& B, const float qx, const float qy, const float qz )
{
    const float ex = qx - B.cx, ey = qy - B.cy, ez = qz - B.cz;
    return fmaf( ex, ex, fmaf( ey, ey, ez * ez ) );
}
B2_HD bool brickAllNear( const BrickBall& B, const float s, const float rhs, const float thr )
{
    const float u = fmaf( -s, 1.f + BRICK_EPS, fmaf( thr, 1.f - BRICK_EPS, -B.r2p ) );
    return u >= 0.f && u * u >= rhs;
}
// thr < FLT_MIN (degenerate node, rr ~ 0) never answers ALL_FAR: such nodes keep the per-voxel guard of the reference's `d > 0`
B2_HD bool brickAllFar( const BrickBall& B, const float s, const float rhs, const float thr )
{
    const float w = fmaf( -thr, 1.f + BRICK_EPS, fmaf( s, 1.f - BRICK_EPS, B.r2m ) );
    return w > 0.f && w * w > rhs && s > B.r2p && thr >= FLT_MIN;
}
B2_HD int brickClassify( const BrickBall& B, const float qx, const float qy, const float qz, const float thr )
{
    const float s = brickDistSq( B, qx, qy, qz );
    const float rhs = s * B.r4;
    if ( brickAllNear( B, s, rhs, thr ) ) return BRICK_ALL_NEAR;
    if ( brickAllFar( B, s, rhs, thr ) ) return BRICK_ALL_FAR;
    return BRICK_MIXED;
}

} // namespace b2
