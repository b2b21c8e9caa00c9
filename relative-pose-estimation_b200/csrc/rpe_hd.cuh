// rpe_hd.cuh — portability shims so the per-element arithmetic of every kernel is written
// once as __host__ __device__ code.  The product compiles these for sm_100a only; the
// tests additionally compile the same headers with g++ (tests/hostcheck) so that the exact
// integer / f32 / f64 arithmetic can be diffed against the oracle without a GPU.  The host
// instantiation is TEST-ONLY: librpe_b200.so never runs it (there is no CPU fallback).
#pragma once
#include <stdint.h>
#include <math.h>

#if defined(__CUDACC__)
#define RPE_HD __host__ __device__ __forceinline__
#define RPE_HDN __host__ __device__ __noinline__
#else
#define RPE_HD inline
#define RPE_HDN inline
#endif

// read-only tables visible from __host__ __device__ functions
#if defined(__CUDACC__)
#define RPE_DEFINE_TABLE(type, name, n, ...)          \
  static const type name##_host[n] = __VA_ARGS__;     \
  static __device__ __constant__ type name##_dev[n] = __VA_ARGS__;
#else
#define RPE_DEFINE_TABLE(type, name, n, ...) static const type name##_host[n] = __VA_ARGS__;
#endif
#if defined(__CUDA_ARCH__)
#define RPE_TBL(name) name##_dev
#else
#define RPE_TBL(name) name##_host
#endif

// individually rounded f32 operations (the reference arithmetic is non-fused unless stated;
// nvcc would otherwise contract a*b+c into an FMA).  Host builds use -ffp-contract=off.
RPE_HD float f32_mul(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fmul_rn(a, b);
#else
  return a * b;
#endif
}
RPE_HD float f32_add(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fadd_rn(a, b);
#else
  return a + b;
#endif
}
RPE_HD float f32_sub(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fsub_rn(a, b);
#else
  return a - b;
#endif
}
RPE_HD float f32_div(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fdiv_rn(a, b);
#else
  return a / b;
#endif
}
RPE_HD float f32_fma(float a, float b, float c) { return fmaf(a, b, c); }
// individually rounded f64 operations, for the places where a discrete decision of the reference hangs on the last
// bit (cv2 is built without FMA contraction; nvcc contracts a*b+c by default)
RPE_HD double f64_mul(double a, double b) {
#if defined(__CUDA_ARCH__)
  return __dmul_rn(a, b);
#else
  return a * b;
#endif
}
RPE_HD double f64_add(double a, double b) {
#if defined(__CUDA_ARCH__)
  return __dadd_rn(a, b);
#else
  return a + b;
#endif
}
// round half to even (cvRound)
RPE_HD int f32_rint(float a) {
#if defined(__CUDA_ARCH__)
  return __float2int_rn(a);
#else
  return (int)lrintf(a);
#endif
}
RPE_HD int f64_rint(double a) {
#if defined(__CUDA_ARCH__)
  return __double2int_rn(a);
#else
  return (int)lrint(a);
#endif
}
RPE_HD int imin(int a, int b) { return a < b ? a : b; }
RPE_HD int imax(int a, int b) { return a > b ? a : b; }
