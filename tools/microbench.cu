// microbench.cu -- measured pipe peaks on the box (the FP32-FMA roofline denominator is not in
// MEASURED_PEAKS.json): FP32 FFMA issue rate and legacy mma.sync tf32 / bf16 rates.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/microbench tools/microbench.cu
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_bf16.h>

__global__ void ffma_kernel(float* out, int iters) {
  float a[16];
  for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 1e-3f + i;
  const float b = 1.000001f, c = 1e-7f;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
      for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], b, c);
  }
  float s = 0;
  for (int i = 0; i < 16; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void mma_tf32_kernel(float* out, int iters) {
  float d[8][4];
  for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) d[i][j] = 0.f;
  unsigned a0 = threadIdx.x, a1 = 1, a2 = 2, a3 = 3, b0 = 4, b1 = 5;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                   : "+f"(d[i][0]), "+f"(d[i][1]), "+f"(d[i][2]), "+f"(d[i][3])
                   : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  float s = 0;
  for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) s += d[i][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void mma_bf16_kernel(float* out, int iters) {
  float d[8][4];
  for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) d[i][j] = 0.f;
  unsigned a0 = threadIdx.x, a1 = 1, a2 = 2, a3 = 3, b0 = 4, b1 = 5;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
                   : "+f"(d[i][0]), "+f"(d[i][1]), "+f"(d[i][2]), "+f"(d[i][3])
                   : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  float s = 0;
  for (int i = 0; i < 8; ++i) for (int j = 0; j < 4; ++j) s += d[i][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F> float time_ms(F f) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  f(); cudaDeviceSynchronize();
  float best = 1e30f;
  for (int r = 0; r < 5; ++r) {
    cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
  }
  return best;
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  const int nsm = p.multiProcessorCount;
  float* out; cudaMalloc(&out, sizeof(float) * nsm * 8 * 1024);
  const int iters = 4096;
  for (int warps = 8; warps <= 32; warps *= 2) {
    const int nt = warps * 32, blocks = nsm * 2;
    float ms = time_ms([&] { ffma_kernel<<<blocks, nt>>>(out, iters); });
    double flop = 2.0 * blocks * nt * (double)iters * 8 * 16;
    printf("{\"bench\": \"ffma\", \"warps_per_cta\": %d, \"ctas\": %d, \"ms\": %.3f, \"tflops\": %.2f}\n", warps, blocks, ms, flop / ms * 1e-9);
  }
  for (int warps = 4; warps <= 16; warps *= 2) {
    const int nt = warps * 32, blocks = nsm * 2;
    float ms = time_ms([&] { mma_tf32_kernel<<<blocks, nt>>>(out, iters); });
    double flop = 2.0 * 16 * 8 * 8 * (double)blocks * warps * iters * 8;
    printf("{\"bench\": \"mma_sync_tf32_m16n8k8\", \"warps_per_cta\": %d, \"ms\": %.3f, \"tflops\": %.2f}\n", warps, ms, flop / ms * 1e-9);
    ms = time_ms([&] { mma_bf16_kernel<<<blocks, nt>>>(out, iters); });
    flop = 2.0 * 16 * 8 * 16 * (double)blocks * warps * iters * 8;
    printf("{\"bench\": \"mma_sync_bf16_m16n8k16\", \"warps_per_cta\": %d, \"ms\": %.3f, \"tflops\": %.2f}\n", warps, ms, flop / ms * 1e-9);
  }
  printf("{\"sm_count\": %d, \"clock_khz\": %d}\n", nsm, p.clockRate);
  return 0;
}
