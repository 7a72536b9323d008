// ORACLE BUILD SHIM: replaces sutil/Exception.h (which drags in glad/OpenGL) for host-compiling
// sv4/CUDABuffer.h + sv4/Probe.h.  No CUDA call is ever made by the oracle.
#pragma once
#include <stdexcept>
#define CUDA_CHECK(call) (void)(call)
#define CUDA_SYNC_CHECK() ((void)0)
#define OPTIX_CHECK(call) (void)(call)
