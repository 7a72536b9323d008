// ORACLE — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
// "reference" shading backend: host-compiles the reference's OWN headers, read in place from
// /root/reference (never copied into this repo), so that the oracle's shading arithmetic is the
// reference's code rather than a restatement:
//   cuda/random.h (tea/lcg/rnd), sutil/vec_math.h, sv4/maths.h (Random, sampling), sv4/sample.h (Sample2D),
//   sv4/Material.h, sv4/Probe.cuh + sv4/Probe.h (ProbeSample/ProbeEval/BuildCDF), sv4/Disney.cuh,
//   cuda/helpers.h (toSRGB/make_color).
// Built only by `make -C oracle ref` into oracle/_ref/ (git-ignored) when /root/reference exists.
// Shims (oracle/shim/): stub optix.h / optix_stubs.h, Maths.h -> maths.h forwarder, sutil/Exception.h.
// <stdlib.h>/<math.h>/<cmath> MUST come first so unqualified abs(float) is not C's int abs(int)
// (SURVEY.md §8c trap 4).
#pragma once
#include <stdlib.h>
#include <math.h>
#include <cmath>
#include <cfloat>
#include <cstdint>
#include <cuda_runtime.h>
using std::isfinite;
#include "random.h"            // /root/reference/cuda/random.h
#include <sutil/vec_math.h>    // /root/reference/sutil/vec_math.h
#include "Disney.cuh"          // sv4: pulls maths.h, sample.h, LaunchParams.h, Probe.cuh, Material.h
#include "Probe.h"             // sv4: ProbeData::BuildCDF
#include <cuda/helpers.h>      // make_color (compiled with -D__device__=)
#define ORACLE_BACKEND_NAME "reference"
