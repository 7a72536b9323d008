// ORACLE BUILD SHIM: sv4/Material.h:3 includes "Maths.h"; on the authors' case-insensitive file system
// that is the variant's own maths.h.  Forward to it (and keep OptixUtils/ off the include path).
#pragma once
#include "maths.h"
