// Stub of chrono/multicore_math/utility.h for the oracle build: the two frame-transform helpers
// (real ones at src/chrono/multicore_math/utility.h:48-57); the rest needs Eigen-backed ChVector3.
#pragma once
#include "chrono/multicore_math/matrix.h"
#include "chrono/multicore_math/types.h"
namespace chrono {
static inline real3 TransformLocalToParent(const real3& p, const quaternion& q, const real3& rl) {
    return p + Rotate(rl, q);
}
static inline real3 TransformParentToLocal(const real3& p, const quaternion& q, const real3& rp) {
    return RotateT(rp - p, q);
}
}  // namespace chrono
