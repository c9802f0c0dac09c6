// oracle/shadow/MRPch/MROpenVDB.h -- TEST INFRASTRUCTURE ONLY (our file, not reference code).
//
// Stand-in for the reference's precompiled-header wrapper of OpenVDB (source/MRPch/MROpenVDB.h), put first on the
// include path when oracle/Makefile compiles source/MRVoxels/MRMarchingCubes.cpp verbatim.  OpenVDB itself cannot be
// built here (needs generated version.h + Boost + Blosc).  Only the SimpleVolume entry points of the
// reference's marching cubes are called by the oracle; the VdbVolume overloads merely have to COMPILE, which needs
// the handful of names below.  None of these bodies ever runs.
#pragma once
#include <MRPch/MRTBB.h>
#include <cstddef>
#include <memory>

namespace openvdb
{
struct Coord
{
    int v[3];
    Coord( int x = 0, int y = 0, int z = 0 ) : v{ x, y, z } {}
    int x() const { return v[0]; }
    int y() const { return v[1]; }
    int z() const { return v[2]; }
};
struct CoordBBox
{
    Coord a, b;
    CoordBBox() = default;
    CoordBBox( Coord lo, Coord hi ) : a( lo ), b( hi ) {}
    const Coord& min() const { return a; }
    const Coord& max() const { return b; }
};
struct FloatGrid
{
    using Ptr = std::shared_ptr<FloatGrid>;
    struct ConstAccessor
    {
        bool probeValue( const Coord&, float& ) const { return false; }
    };
    ConstAccessor getConstAccessor() const { return {}; }
    CoordBBox evalActiveVoxelBoundingBox() const { return {}; }
    size_t memUsage() const { return 0; }
};
} // namespace openvdb
