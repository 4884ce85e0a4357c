// tc_probe.cu -- bring-up probe for the tcgen05 path (run on a B200 through gpurun):
//   1. split-precision GEMM correctness, K-major operands:  kind::f16 (fp16 hi/lo) and kind::tf32, 1 and 3 terms
//   2. the weight-gradient contraction with MN-major operands out of the same slabs
//   3. tcgen05.mma issue throughput (all SMs), 4. L2 -> shared bulk-copy streaming bandwidth
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -o tools/tc_probe tools/tc_probe.cu
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../meta_pde_b200/csrc/tc_common.cuh"

using namespace mpde::tc;

#define CK(x)                                                                          \
  do {                                                                                 \
    cudaError_t e_ = (x);                                                              \
    if (e_ != cudaSuccess) {                                                           \
      printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
      exit(2);                                                                         \
    }                                                                                  \
  } while (0)

// ---------------------------------------------------------------------------------------------------------
// 1. D[128][N] = A[128][K] * B[N][K]^T.   KIND 0: fp16 hi/lo (kind::f16), KIND 1: tf32 hi/lo (kind::tf32)
// ---------------------------------------------------------------------------------------------------------
template <int KIND, int TERMS, int N, int K>
__global__ void __launch_bounds__(128, 1) gemm_kmajor_kernel(const float* A, const float* B, float* D) {
  constexpr int EB = KIND == 0 ? 2 : 4;          // operand element bytes
  constexpr int EPS = 128 / EB;                  // elements per slab row
  constexpr int NSLAB = K / EPS;
  constexpr int KSTEP = 32 / EB;                 // K per MMA (32 bytes)
  constexpr uint32_t A_SLAB = 128 * 128, B_SLAB = N * 128;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_hi = smem;
  uint8_t* a_lo = a_hi + NSLAB * A_SLAB;
  uint8_t* b_hi = a_lo + NSLAB * A_SLAB;
  uint8_t* b_lo = b_hi + NSLAB * B_SLAB;
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
  if (warp == 0) tmem_alloc(&tmem_slot, 256);
  // operands: split and scatter into the swizzled slabs
  for (int mat = 0; mat < 2; ++mat) {
    const float* src = mat == 0 ? A : B;
    const int rows = mat == 0 ? 128 : N;
    uint8_t* hi = mat == 0 ? a_hi : b_hi;
    uint8_t* lo = mat == 0 ? a_lo : b_lo;
    const uint32_t slab = mat == 0 ? A_SLAB : B_SLAB;
    for (int i = tid; i < rows * K; i += 128) {
      const int r = i / K, k = i % K;
      const float x = src[i];
      const uint32_t off = (k / EPS) * slab + sw128_off(r, (k % EPS) * EB);
      if (KIND == 0) {
        __half h, l;
        split_f16(x, h, l);
        *reinterpret_cast<__half*>(hi + off) = h;
        *reinterpret_cast<__half*>(lo + off) = l;
      } else {
        const float h = __uint_as_float(__float_as_uint(x) & 0xffffe000u);
        *reinterpret_cast<float*>(hi + off) = h;
        *reinterpret_cast<float*>(lo + off) = x - h;
      }
    }
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = tmem_slot;

  if (tid == 0) {
    constexpr uint32_t idesc = make_idesc(KIND == 0 ? FMT_F16 : FMT_TF32, 128, N, 0, 0);
    uint32_t acc = 0;
    for (int term = 0; term < TERMS; ++term) {
      const uint8_t* pa = (term == 1) ? a_lo : a_hi;
      const uint8_t* pb = (term == 2) ? b_lo : b_hi;
      for (int s = 0; s < NSLAB; ++s)
        for (int kk = 0; kk < EPS / KSTEP; ++kk) {
          const uint64_t ad = desc_kmajor(smem_u32(pa + s * A_SLAB) + kk * 32);
          const uint64_t bd = desc_kmajor(smem_u32(pb + s * B_SLAB) + kk * 32);
          if (KIND == 0) mma_f16_ss(tbase, ad, bd, idesc, acc); else mma_tf32_ss(tbase, ad, bd, idesc, acc);
          acc = 1;
        }
    }
    tc_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  for (int c0 = 0; c0 < N; c0 += 32) {
    uint32_t v[32];
    tmem_ld32(tmem_addr(tbase, 32 * warp, c0), v);
    tmem_ld_wait();
    for (int j = 0; j < 32; ++j) D[(size_t)(32 * warp + lane) * N + c0 + j] = __uint_as_float(v[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 256);
}

// ---------------------------------------------------------------------------------------------------------
// 2. Weight-gradient contraction D[m][n] = sum_r H[r][m] * Z[r][n], M = 128, N = 256, both operands MN-major
//    (fp16 hi/lo, 3 terms), read from slabs laid out exactly as the K-major activation buffers [row][k].
// ---------------------------------------------------------------------------------------------------------
template <int R, int KIND>
__global__ void __launch_bounds__(128, 1) gemm_mnmajor_kernel(const float* H, const float* Z, float* D) {
  constexpr int M = 128, N = 256;
  constexpr int EB = KIND == 0 ? 2 : 4, CPS = 128 / EB;   // columns per slab
  constexpr int KSTEP = 32 / EB;                          // rows of r per MMA
  constexpr uint32_t SLAB = R * 128;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* h_hi = smem;
  uint8_t* h_lo = h_hi + (M / CPS) * SLAB;
  uint8_t* z_hi = h_lo + (M / CPS) * SLAB;
  uint8_t* z_lo = z_hi + (N / CPS) * SLAB;
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
  if (warp == 0) tmem_alloc(&tmem_slot, 256);
  for (int mat = 0; mat < 2; ++mat) {
    const float* src = mat == 0 ? H : Z;
    const int cols = mat == 0 ? M : N;
    uint8_t* hi = mat == 0 ? h_hi : z_hi;
    uint8_t* lo = mat == 0 ? h_lo : z_lo;
    for (int i = tid; i < R * cols; i += 128) {
      const int r = i / cols, c = i % cols;
      const uint32_t off = (c / CPS) * SLAB + sw128_off(r, (c % CPS) * EB);
      if (KIND == 0) {
        __half h, l;
        split_f16(src[i], h, l);
        *reinterpret_cast<__half*>(hi + off) = h;
        *reinterpret_cast<__half*>(lo + off) = l;
      } else {
        const float x = src[i];
        const float h = __uint_as_float(__float_as_uint(x) & 0xffffe000u);
        *reinterpret_cast<float*>(hi + off) = h;
        *reinterpret_cast<float*>(lo + off) = x - h;
      }
    }
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = tmem_slot;
  if (tid == 0) {
    constexpr uint32_t idesc = make_idesc(KIND == 0 ? FMT_F16 : FMT_TF32, M, N, 1, 1);
    uint32_t acc = 0;
    for (int term = 0; term < 3; ++term) {
      const uint8_t* pa = (term == 1) ? h_lo : h_hi;
      const uint8_t* pb = (term == 2) ? z_lo : z_hi;
      for (int j = 0; j < R / KSTEP; ++j) {
        const uint64_t ad = desc_mnmajor(smem_u32(pa) + j * (KSTEP / 8) * 1024, SLAB);
        const uint64_t bd = desc_mnmajor(smem_u32(pb) + j * (KSTEP / 8) * 1024, SLAB);
        if (KIND == 0) mma_f16_ss(tbase, ad, bd, idesc, acc); else mma_tf32_ss(tbase, ad, bd, idesc, acc);
        acc = 1;
      }
    }
    tc_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  for (int c0 = 0; c0 < N; c0 += 32) {
    uint32_t v[32];
    tmem_ld32(tmem_addr(tbase, 32 * warp, c0), v);
    tmem_ld_wait();
    for (int j = 0; j < 32; ++j) D[(size_t)(32 * warp + lane) * N + c0 + j] = __uint_as_float(v[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 256);
}

// ---------------------------------------------------------------------------------------------------------
// 3. MMA issue throughput: every CTA re-issues M=128 x N x 32-byte-K MMAs on resident operands.
// ---------------------------------------------------------------------------------------------------------
template <int KIND, int N>
__global__ void __launch_bounds__(128, 1) mma_rate_kernel(int batches, float* sink) {
  constexpr uint32_t A_SLAB = 128 * 128, B_SLAB = N * 128;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a = smem;
  uint8_t* b = a + A_SLAB;
  __shared__ uint64_t bar[2];
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  if (tid == 0) { mbar_init(&bar[0], 1); mbar_init(&bar[1], 1); mbar_fence_init(); }
  if (warp == 0) tmem_alloc(&tmem_slot, 512);
  for (uint32_t i = tid; i < (A_SLAB + B_SLAB) / 4; i += 128) {
    uint32_t h = i * 2654435761u + 12345u;
    if (KIND == 0) {
      const __half2 v = __floats2half2_rn(((h & 0xffff) / 65536.f) - 0.5f, ((h >> 16) / 65536.f) - 0.5f);
      reinterpret_cast<__half2*>(smem)[i] = v;
    } else {
      reinterpret_cast<float*>(smem)[i] = ((h & 0xffffff) / 16777216.f) - 0.5f;
    }
  }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = tmem_slot;
  if (tid == 0) {
    constexpr uint32_t idesc = make_idesc(KIND == 0 ? FMT_F16 : FMT_TF32, 128, N, 0, 0);
    for (int bt = 0; bt < batches; ++bt) {
      if (bt >= 2) mbar_wait(&bar[bt & 1], ((bt >> 1) - 1) & 1);   // keep at most two batches in flight
      const uint32_t d = tbase + (bt & 1) * 256;
      for (int i = 0; i < 32; ++i) {
        const uint64_t ad = desc_kmajor(smem_u32(a) + (i & 3) * 32);
        const uint64_t bd = desc_kmajor(smem_u32(b) + (i & 3) * 32);
        if (KIND == 0) mma_f16_ss(d, ad, bd, idesc, i > 0); else mma_tf32_ss(d, ad, bd, idesc, i > 0);
      }
      tc_commit(&bar[bt & 1]);
    }
    for (int q = 0; q < 2; ++q) {
      const int bt = batches - 2 + q;
      if (bt >= 0) mbar_wait(&bar[bt & 1], (bt >> 1) & 1);
    }
  }
  __syncthreads();
  tc_fence_after();
  if (tid < 32) {
    uint32_t v[16];
    tmem_ld16(tmem_addr(tbase, 0, 0), v);
    tmem_ld_wait();
    if (sink != nullptr && v[0] == 0x12345678u) sink[blockIdx.x] = __uint_as_float(v[1]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tbase, 512);
}

// ---------------------------------------------------------------------------------------------------------
// 4. L2 -> shared streaming: each CTA pulls `chunks` bulk copies of CH bytes through an S-stage ring.
// ---------------------------------------------------------------------------------------------------------
template <int CH, int S>
__global__ void __launch_bounds__(128, 1) bulk_stream_kernel(const uint8_t* src, size_t task_bytes, int ntask, int chunks,
                                                             unsigned* sink) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t full[S];
  const int tid = threadIdx.x;
  if (tid == 0) { for (int s = 0; s < S; ++s) mbar_init(&full[s], 1); mbar_fence_init(); }
  __syncthreads();
  const uint8_t* base = src + (size_t)(blockIdx.x % ntask) * task_bytes;
  const int per_task = (int)(task_bytes / CH);
  if (tid == 0) {
    for (int c = 0; c < chunks + S; ++c) {
      if (c >= S) mbar_wait(&full[(c - S) % S], ((c - S) / S) & 1);   // chunk c-S landed: its slot is free again
      if (c < chunks) {
        mbar_arrive_expect_tx(&full[c % S], CH);
        bulk_g2s(smem + (size_t)(c % S) * CH, base + (size_t)(c % per_task) * CH, CH, &full[c % S]);
      }
    }
  }
  __syncthreads();
  if (tid == 0 && sink != nullptr) sink[blockIdx.x] = smem[17];
}

// ---------------------------------------------------------------------------------------------------------
static float frand(unsigned& s) { s = s * 1664525u + 1013904223u; return ((s >> 8) / 8388608.0f) - 1.0f; }

static double rel_l2(const std::vector<float>& got, const std::vector<double>& ref) {
  double num = 0, den = 0;
  for (size_t i = 0; i < ref.size(); ++i) { const double d = got[i] - ref[i]; num += d * d; den += ref[i] * ref[i]; }
  return std::sqrt(num / (den > 0 ? den : 1));
}

template <int KIND, int TERMS, int N, int K>
static void run_kmajor(const char* name) {
  unsigned seed = 1234u + KIND * 17 + TERMS;
  std::vector<float> A(128 * K), B(N * K), D(128 * N);
  for (auto& x : A) x = frand(seed);
  for (auto& x : B) x = frand(seed);
  std::vector<double> ref(128 * N);
  for (int m = 0; m < 128; ++m)
    for (int n = 0; n < N; ++n) {
      double s = 0;
      for (int k = 0; k < K; ++k) s += (double)A[m * K + k] * (double)B[n * K + k];
      ref[m * N + n] = s;
    }
  float *dA, *dB, *dD;
  CK(cudaMalloc(&dA, A.size() * 4)); CK(cudaMalloc(&dB, B.size() * 4)); CK(cudaMalloc(&dD, D.size() * 4));
  CK(cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dB, B.data(), B.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemset(dD, 0, D.size() * 4));
  constexpr int EB = KIND == 0 ? 2 : 4;
  const int smem = 2 * (128 + N) * K * EB + 1024;
  auto kern = gemm_kmajor_kernel<KIND, TERMS, N, K>;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  kern<<<1, 128, smem>>>(dA, dB, dD);
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
  printf("{\"probe\": \"gemm_kmajor\", \"kind\": \"%s\", \"terms\": %d, \"N\": %d, \"K\": %d, \"rel_l2\": %.3e}\n", name,
         TERMS, N, K, rel_l2(D, ref));
  fflush(stdout);
  cudaFree(dA); cudaFree(dB); cudaFree(dD);
}

template <int R, int KIND>
static void run_mnmajor() {
  unsigned seed = 99u + R;
  std::vector<float> H(R * 128), Z(R * 256), D(128 * 256);
  for (auto& x : H) x = frand(seed);
  for (auto& x : Z) x = frand(seed);
  std::vector<double> ref(128 * 256);
  for (int m = 0; m < 128; ++m)
    for (int n = 0; n < 256; ++n) {
      double s = 0;
      for (int r = 0; r < R; ++r) s += (double)H[r * 128 + m] * (double)Z[r * 256 + n];
      ref[m * 256 + n] = s;
    }
  float *dH, *dZ, *dD;
  CK(cudaMalloc(&dH, H.size() * 4)); CK(cudaMalloc(&dZ, Z.size() * 4)); CK(cudaMalloc(&dD, D.size() * 4));
  CK(cudaMemcpy(dH, H.data(), H.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(dZ, Z.data(), Z.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemset(dD, 0, D.size() * 4));
  const int smem = 2 * (128 + 256) * R * (KIND == 0 ? 2 : 4) + 1024;
  auto kern = gemm_mnmajor_kernel<R, KIND>;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  kern<<<1, 128, smem>>>(dH, dZ, dD);
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());
  CK(cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost));
  printf("{\"probe\": \"gemm_mnmajor_wgrad\", \"kind\": \"%s\", \"R\": %d, \"rel_l2\": %.3e}\n", KIND == 0 ? "f16x3" : "tf32x3", R, rel_l2(D, ref));
  fflush(stdout);
  cudaFree(dH); cudaFree(dZ); cudaFree(dD);
}

template <int KIND, int N>
static void run_rate(const char* name, int ctas) {
  const int batches = 2000;
  const int smem = (128 + N) * 128 + 1024;
  auto kern = mma_rate_kernel<KIND, N>;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  kern<<<ctas, 128, smem>>>(50, nullptr);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  kern<<<ctas, 128, smem>>>(batches, nullptr);
  CK(cudaEventRecord(e1));
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());
  float ms;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  const double kper = KIND == 0 ? 16 : 8;
  const double flops = 2.0 * 128 * N * kper * 32.0 * batches * ctas;
  printf("{\"probe\": \"mma_rate\", \"kind\": \"%s\", \"M\": 128, \"N\": %d, \"ctas\": %d, \"ms\": %.3f, \"tflops\": %.1f}\n",
         name, N, ctas, ms, flops / ms * 1e-9);
  fflush(stdout);
}

template <int CH, int S>
static void run_stream(int ctas, int ntask, size_t task_bytes) {
  uint8_t* src;
  CK(cudaMalloc(&src, task_bytes * ntask));
  CK(cudaMemset(src, 1, task_bytes * ntask));
  const int chunks = 4096;
  const int smem = CH * S + 1024;
  auto kern = bulk_stream_kernel<CH, S>;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  kern<<<ctas, 128, smem>>>(src, task_bytes, ntask, 256, nullptr);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(e0));
  kern<<<ctas, 128, smem>>>(src, task_bytes, ntask, chunks, nullptr);
  CK(cudaEventRecord(e1));
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());
  float ms;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  const double bytes = (double)CH * chunks * ctas;
  printf("{\"probe\": \"bulk_stream_l2\", \"chunk\": %d, \"stages\": %d, \"ctas\": %d, \"ntask\": %d, \"task_kb\": %zu, "
         "\"ms\": %.3f, \"gbs\": %.1f, \"bytes_per_clk_per_sm_at_1965\": %.1f}\n",
         CH, S, ctas, ntask, task_bytes / 1024, ms, bytes / ms * 1e-6, bytes / ms * 1e-6 / ctas / 1.965);
  fflush(stdout);
  cudaFree(src);
}

int main(int argc, char** argv) {
  int what = argc > 1 ? atoi(argv[1]) : 15;
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  printf("{\"device\": \"%s\", \"sms\": %d}\n", prop.name, prop.multiProcessorCount);
  const int sms = prop.multiProcessorCount;
  if (what & 1) {
    run_kmajor<0, 1, 128, 128>("f16");
    run_kmajor<0, 3, 128, 128>("f16");
    run_kmajor<1, 1, 128, 64>("tf32");
    run_kmajor<1, 3, 128, 64>("tf32");
    run_kmajor<0, 3, 64, 128>("f16");
    run_kmajor<0, 3, 256, 64>("f16");
  }
  if (what & 2) { run_mnmajor<64, 0>(); run_mnmajor<128, 0>(); run_mnmajor<32, 1>(); run_mnmajor<64, 1>(); }
  if (what & 4) {
    run_rate<0, 256>("f16", sms); run_rate<0, 128>("f16", sms); run_rate<0, 64>("f16", sms);
    run_rate<1, 256>("tf32", sms); run_rate<1, 128>("tf32", sms); run_rate<1, 64>("tf32", sms);
  }
  if (what & 8) {
    run_stream<32768, 4>(sms, 64, 1 << 20);
    run_stream<32768, 6>(sms, 64, 1 << 20);
    run_stream<16384, 8>(sms, 64, 1 << 20);
    run_stream<32768, 4>(sms, 1, 1 << 20);
    run_stream<32768, 4>(sms, 2048, 1 << 20);   // 2 GB: HBM-resident
  }
  return 0;
}
