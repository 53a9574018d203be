// Stub of chrono/core/ChApiCE.h for the oracle build: no DLL decoration, plus the standard headers
// the real file pulls in transitively.
#pragma once
#include <cmath>
#include <iostream>
#include <memory>
#include <string>
#include <vector>
#define ChApi
