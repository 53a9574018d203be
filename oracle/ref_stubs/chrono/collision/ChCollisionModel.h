// Stub of chrono/collision/ChCollisionModel.h for the oracle build: only the shape-type enumeration
// the multicore collision code switches on (order as in src/chrono/collision/ChCollisionShape.h:31-53).
#pragma once
namespace chrono {
struct ChCollisionShape {
    enum Type {
        SPHERE, ELLIPSOID, BOX, CYLINDER, CYLSHELL, CONVEXHULL, TRIANGLEMESH, BARREL, POINT, SEGMENT,
        TRIANGLE, CONNECTEDTRIANGLE, CAPSULE, CONE, ROUNDEDBOX, ROUNDEDCYL, TETRAHEDRON, PATH2D,
        SEGMENT2D, ARC2D, UNKNOWN_SHAPE
    };
};
}  // namespace chrono
