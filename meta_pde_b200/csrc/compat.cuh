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
static inline void mpde_sincosf(float x, float* s, float* c) { *s = sinf(x); *c = cosf(x); }
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
// accurate (range-reduced) sincosf: arguments reach |omega*z| ~ 1e2 and the
// parity bar is rel 1e-4 on Hessian-vector products -- never __sinf/__cosf.
__host__ __device__ __forceinline__ void mpde_sincosf(float x, float* s, float* c) { sincosf(x, s, c); }
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
