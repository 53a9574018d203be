// Stub of chrono/physics/ChContactContainer.h for the oracle build: the collision data header only
// needs the math constants and the shape enumeration from it.
#pragma once
#include "chrono/utils/ChConstants.h"
#include "chrono/collision/ChCollisionModel.h"
