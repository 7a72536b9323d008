// ORACLE BUILD SHIM (test infrastructure).  The reference includes <optix.h> only for the opaque
// traversable handle type (sv4/LaunchParams.h:19,80) and, through sv4/CUDABuffer.h, for CUdeviceptr;
// the OptiX SDK is not available here.
#pragma once
typedef unsigned long long OptixTraversableHandle;
typedef unsigned long long CUdeviceptr;
