// devmath.cuh -- qualifiers and intrinsic shims that let the per-particle 3x3 math headers
// (svd3.cuh, constitutive.cuh) compile both as sm_100a device code and, for the host-side numerics
// tests only (tests/host_math.cpp), as plain C++.  The product path always compiles them with nvcc.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#include <cuda_runtime.h>
#define MPMK_DEV __device__ __forceinline__
#define MPMK_FAST_RCP(x) __frcp_rn(x)
#else
#define MPMK_DEV static inline
static inline float rsqrtf(float x) { return 1.0f / sqrtf(x); }
#define MPMK_FAST_RCP(x) (1.0f / (x))
#endif
