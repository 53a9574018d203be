// Stub of chrono/collision/ChCollisionInfo.h for the oracle build
// (default value from src/chrono/collision/ChCollisionInfo.cpp:19).
#pragma once
namespace chrono {
struct ChCollisionInfo {
    static double GetDefaultEffectiveCurvatureRadius() { return 0.1; }
};
}  // namespace chrono
