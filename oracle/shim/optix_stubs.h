// ORACLE BUILD SHIM: empty stand-in for the OptiX SDK's optix_stubs.h (sv4/CUDABuffer.h:21).
#pragma once
