// ref_prelude.h — force-included (-include) in front of every reference translation unit that is
// compiled for the host under oracle/_ref.  TEST INFRASTRUCTURE (see oracle/ref/README.md).
//
// The reference's __both__ headers are written for nvcc/MSVC.  This prelude supplies what those
// two give implicitly, WITHOUT touching the reference's text:
//   * `struct random` (random.hpp:34) shares its name with POSIX random(3); under g++ the function
//     declaration of <stdlib.h> would hide the type.  All system headers the reference pulls in
//     are included here once with the POSIX name moved out of the way.
//   * CUDA qualifiers and handle types (device_global.hpp:4 includes <cuda_runtime.h>; a stand-in
//     of that name lives beside this file).
//   * ::max / ::min on floats (math.hpp:71 calls `max(0.0f, ...)` unqualified: a CUDA global).
#pragma once
#define random posix_random_hidden
#define srandom posix_srandom_hidden
#include <stdlib.h>
#include <cstdlib>
#undef random
#undef srandom
#include <math.h>
#include <cmath>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <algorithm>
#include <map>
#include <string>
#include <vector>

#ifndef __CUDACC__
#define __constant__
#define __device__
#define __host__
#define __global__
#define __forceinline__ inline
#endif

typedef unsigned long long cudaTextureObject_t;
typedef unsigned long long OptixTraversableHandle;

using std::max;
using std::min;

// macros.hpp:8,11,14 abort a thread with the PTX instruction `asm("trap;")`.  The host assembler is
// given a macro of that name (x86 `ud2` = SIGILL: a failed bounds check still kills the run).
#if defined(__x86_64__) && !defined(__CUDACC__)
asm(".macro trap\n\tud2\n.endm");
#endif
