// Host stand-ins for the few CUDA built-ins the line-FFT routines of
// pyscf_isdf_b200/csrc/fft_core.cuh use, so that the same source compiles with g++ and runs a
// "CTA" as real threads meeting at a barrier (TEST INFRASTRUCTURE: checks index logic and missing
// barriers on the CPU; it says nothing about speed).
#pragma once
#include <barrier>
#include <cmath>
#include <cstddef>

struct double2 { double x, y; };
static inline double2 make_double2(double x, double y) { return double2{x, y}; }

#define __device__
#define __host__
#define __forceinline__ inline
#define __restrict__

struct EmulDim3 { int x = 1, y = 1, z = 1; };
static thread_local EmulDim3 threadIdx;
static EmulDim3 blockDim;
static std::barrier<>* emul_barrier = nullptr;
static inline void __syncthreads() { emul_barrier->arrive_and_wait(); }

static inline int __float2int_rz(float f) { return (int)f; }        // C++ float -> int conversion truncates
static inline float __int2float_rn(int a) { return (float)a; }      // round to nearest (default mode)
