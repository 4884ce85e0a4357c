// compat.cuh -- one source, two builds.
//
//   nvcc (product):      the kernels below are real sm_100a CUDA kernels.
//   g++ -DMPDE_EMU=1:    the same kernel SOURCE runs on a host-side SIMT
//                        emulator (tests/emu/emu_runtime.h: one fiber per CUDA
//                        thread, a scheduler pass per __syncthreads interval).
//                        The emulator build exists ONLY so that tests/ can
//                        check kernel indexing / synchronisation logic in the
//                        CPU-only container.  It is never loaded by the
//                        meta_pde_b200 package (see _lib.py) and is not a
//                        fallback: the product library is the nvcc build.
#pragma once
#include <stdint.h>
#include <stddef.h>

#ifdef MPDE_EMU
// ------------------------------------------------------------------ emulator
#include <cmath>
#include <cstring>
#include <cstdlib>
#include "emu_runtime.h"

#define MPDE_HD inline
#define MPDE_DEV inline
#define MPDE_GLOBAL static
#define MPDE_LAUNCH_BOUNDS(t, b)
#define MPDE_RESTRICT __restrict__
#define MPDE_UNROLL
#define MPDE_NOUNROLL

struct float4 { float x, y, z, w; };
struct float2 { float x, y; };
static inline float4 make_float4(float a, float b, float c, float d) { return float4{a, b, c, d}; }
static inline float2 make_float2(float a, float b) { return float2{a, b}; }

#define threadIdx (emu::tls().tid)
#define blockIdx (emu::tls().bid)
#define blockDim (emu::tls().bdim)
#define gridDim (emu::tls().gdim)
#define __syncthreads() emu::syncthreads()

static inline unsigned char* mpde_smem() { return emu::tls().smem; }
#ifdef MPDE_LIB_SINCOS
static inline void mpde_sincosf(float x, float* s, float* c) { *s = sinf(x); *c = cosf(x); }
#endif
static inline void cp_async16(void* dst, const void* src) { memcpy(dst, src, 16); }
static inline void cp_async_commit() {}
template <int N> static inline void cp_async_wait() {}
static inline float mpde_ldg(const float* p) { return *p; }

typedef void* cudaStream_t;
typedef int cudaError_t;
#define cudaSuccess 0

#else
// ---------------------------------------------------------------------- CUDA
#include <cuda_runtime.h>

#define MPDE_HD __host__ __device__ __forceinline__
#define MPDE_DEV __device__ __forceinline__
#define MPDE_GLOBAL __global__
#define MPDE_LAUNCH_BOUNDS(t, b) __launch_bounds__(t, b)
#define MPDE_RESTRICT __restrict__
#define MPDE_UNROLL _Pragma("unroll")
#define MPDE_NOUNROLL _Pragma("unroll 1")

extern __shared__ __align__(16) unsigned char mpde_smem_raw[];
__device__ __forceinline__ unsigned char* mpde_smem() { return mpde_smem_raw; }
#ifdef MPDE_LIB_SINCOS
// the CUDA library's sincosf (accurate everywhere, but its slow-path branch sits inside every activation)
__host__ __device__ __forceinline__ void mpde_sincosf(float x, float* s, float* c) { sincosf(x, s, c); }
#endif
__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
  unsigned d = (unsigned)__cvta_generic_to_shared(dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}
__device__ __forceinline__ float mpde_ldg(const float* p) { return __ldg(p); }
#endif

#ifndef MPDE_LIB_SINCOS
// sin and cos of the SIREN pre-activation omega * z, branch-free (shared by the CUDA and the emulator build).
// The library sincosf carries a slow-path branch (Payne-Hanek, |x| > 105615) into every activation: the convergence
// barrier keeps the compiler from interleaving the independent sincos chains of a thread's register tile, and the
// epilogues are latency-bound (3x64 at a saturating grid: LOSS launch 4.96 -> 2.9 ms, GRAD 12.1 -> 10.3 ms).
// Here: k = rint(x * 2/pi); Cody-Waite reduction with FMAs (the first step is exact, the rounding of the second is kept
// in e and fed back to first order: sin(r + e) = sin r + e cos r, cos(r + e) = cos r - e r); least-squares polynomials
// fitted on |r| <= 0.9 (a k that is off by one near a half-integer only pushes r past pi/4); cos as 1 - t/2 + t^2 Q(t)
// with the exact remainders of 1 - t/2 and of t = r * r carried in the tail; quadrant fix-up by selects.
// Measured against float64 sin/cos on the CPU (4e6 random arguments per decade): max abs error 4.6e-8 for |x| <= 1e6
// (0.75 ulp at 1; glibc: 3.3e-8; the CUDA library documents 2 ulp) -- never __sinf/__cosf.  Beyond 1e6 a float32
// argument is itself spaced >= 0.06 rad; there the result is only bounded (|r| clamped).  NaN / Inf give NaN (x - x).
MPDE_HD void mpde_sincosf(float x, float* sp, float* cp) {
  const float k = rintf(x * 0.636619772367581343f);
  const float r1 = fmaf(k, -1.57079601287841796875f, x);
  float r = fmaf(k, -3.13916473527819e-07f, r1);
  float e = fmaf(k, -3.13916473527819e-07f, r1 - r);
  e = fmaf(k, -5.39030252995776476554e-15f, e);
  r = fminf(fmaxf(r, -0.9f), 0.9f) + (x - x);
  const int q = (int)k;
  const float t = r * r, tl = fmaf(r, r, -t);
  float s = fmaf(t, 2.715350092e-06f, -1.983922266e-04f);
  s = fmaf(s, t, 8.333330043e-03f);
  s = fmaf(s, t, -1.666666716e-01f);
  float c = fmaf(t, 2.435894385e-05f, -1.388665405e-03f);
  c = fmaf(c, t, 4.166663066e-02f);
  const float hz = 0.5f * t, w = 1.0f - hz;
  const float tail = fmaf(t * t, c, fmaf(-r, e, fmaf(-0.5f, tl, (1.0f - w) - hz)));
  const float cr = w + tail;
  const float sr = r + fmaf(s * t, r, e * w);
  const bool sw = (q & 1) != 0;
  float ss = sw ? cr : sr, cc = sw ? sr : cr;
  if (q & 2) ss = -ss;
  if ((q + 1) & 2) cc = -cc;
  *sp = ss; *cp = cc;
}
#endif
