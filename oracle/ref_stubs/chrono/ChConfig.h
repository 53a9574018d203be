// Stub of the CMake-generated chrono/ChConfig.h, used only to compile the reference's own collision
// translation units for the oracle (see oracle/Makefile). Test infrastructure, not product code.
// Leaves CHRONO_SIMD_ENABLED / CHRONO_HAS_AVX undefined so the reference takes its scalar math path.
#pragma once
#define CHRONO_OPENMP_ENABLED
#define CHRONO_HAS_THRUST
#define USE_COLLISION_DOUBLE
#ifndef THRUST_DEVICE_SYSTEM
#define THRUST_DEVICE_SYSTEM THRUST_DEVICE_SYSTEM_OMP
#endif
#ifndef THRUST_HOST_SYSTEM
#define THRUST_HOST_SYSTEM THRUST_HOST_SYSTEM_OMP
#endif
