// tc_inst.cu -- instantiation and launch of the tensor-core (tcgen05) jet kernels for widths 128 and 256.
#include "host_common.h"
#include "tc_kernels.cuh"
#include "tc_wgrad.cuh"
#include "tc_hvp.cuh"
#include <mutex>

namespace mpde {

using namespace tc;

template <int W, int PDE, int MODE>
static int tc_run_inst(int op, const TcParams* p, int gx, int gy, void* stream, KernelInfo* info) {
  typedef TcGeo<W> G;
  auto kfn = tc_pde_kernel<W, PDE, MODE>;
  static PerDeviceOnce once;
  if (once.run([&]() { return cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G::SMEM); }) != cudaSuccess) return 4;
  if (op == 0) {
    info->nt = G::NT; info->smem = G::SMEM; info->occ = 1; info->np = TR / 4;
    return 0;
  }
  g_launches++;
  kfn<<<dim3(gx, gy, 1), dim3(G::NT, 1, 1), G::SMEM, (cudaStream_t)stream>>>(*p);
  return cudaGetLastError() != cudaSuccess ? 4 : 0;
}

template <int W>
static int tc_run_width(int pde, int mode, int op, const TcParams* p, int gx, int gy, void* stream, KernelInfo* info) {
  if (mode == MODE_GRAD) {
    switch (pde) {
      case PDE_POISSON: return tc_run_inst<W, PDE_POISSON, MODE_GRAD>(op, p, gx, gy, stream, info);
      case PDE_BURGERS: return tc_run_inst<W, PDE_BURGERS, MODE_GRAD>(op, p, gx, gy, stream, info);
      case PDE_HYPER:   return tc_run_inst<W, PDE_HYPER, MODE_GRAD>(op, p, gx, gy, stream, info);
    }
    return 1;
  }
  if (mode != MODE_LOSS) return 1;
  switch (pde) {
    case PDE_POISSON: return tc_run_inst<W, PDE_POISSON, MODE_LOSS>(op, p, gx, gy, stream, info);
    case PDE_BURGERS: return tc_run_inst<W, PDE_BURGERS, MODE_LOSS>(op, p, gx, gy, stream, info);
    case PDE_HYPER:   return tc_run_inst<W, PDE_HYPER, MODE_LOSS>(op, p, gx, gy, stream, info);
    case PDE_EVAL1:   return tc_run_inst<W, PDE_EVAL1, MODE_LOSS>(op, p, gx, gy, stream, info);
    case PDE_EVAL2:   return tc_run_inst<W, PDE_EVAL2, MODE_LOSS>(op, p, gx, gy, stream, info);
  }
  return 1;
}

template <int W, int PDE>
static int tc_run_hvp_inst(int op, const HvParams* p, int gx, int gy, void* stream, KernelInfo* info) {
  typedef HvGeo<W> G;
  auto kfn = tc_hvp_kernel<W, PDE>;
  static PerDeviceOnce once;
  if (once.run([&]() { return cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G::SMEM); }) != cudaSuccess) return 4;
  if (op == 0) {
    info->nt = G::NT; info->smem = G::SMEM; info->occ = 1; info->np = HR / 4;
    return 0;
  }
  g_launches++;
  kfn<<<dim3(gx, gy, 1), dim3(G::NT, 1, 1), G::SMEM, (cudaStream_t)stream>>>(*p);
  return cudaGetLastError() != cudaSuccess ? 4 : 0;
}
template <int W>
static int tc_run_hvp_width(int pde, int op, const HvParams* p, int gx, int gy, void* stream, KernelInfo* info) {
  switch (pde) {
    case PDE_POISSON: return tc_run_hvp_inst<W, PDE_POISSON>(op, p, gx, gy, stream, info);
    case PDE_BURGERS: return tc_run_hvp_inst<W, PDE_BURGERS>(op, p, gx, gy, stream, info);
    case PDE_HYPER:   return tc_run_hvp_inst<W, PDE_HYPER>(op, p, gx, gy, stream, info);
  }
  return 1;
}
int dispatch_tc_hvp(int W, int pde, int op, const HvParams* p, int gx, int gy, void* stream, KernelInfo* info) {
  if (W == 128) return tc_run_hvp_width<128>(pde, op, p, gx, gy, stream, info);
  if (W == 256) return tc_run_hvp_width<256>(pde, op, p, gx, gy, stream, info);
  return 1;
}

// modes the tensor-core path implements for this width / depth (the FP32-FMA kernels serve everything else)
bool tc_supported(int W, int L, int mode) {
  return (W == 128 || W == 256) && L >= 2 && L <= TC_MAX_L && (mode == MODE_LOSS || mode == MODE_GRAD || mode == MODE_HVP);
}

int dispatch_tc(int W, int pde, int mode, int op, const TcParams* p, int gx, int gy, void* stream, KernelInfo* info) {
  if (W == 128) return tc_run_width<128>(pde, mode, op, p, gx, gy, stream, info);
  if (W == 256) return tc_run_width<256>(pde, mode, op, p, gx, gy, stream, info);
  return 1;
}

size_t tc_image_bytes(int W, int L, int with_bwd) {
  return W == 128 ? TcGeo<128>::image_bytes(L, with_bwd != 0) : TcGeo<256>::image_bytes(L, with_bwd != 0);
}

int tc_launch_prep(int W, int L, int OUT, int T, int with_bwd, const float* theta, long long theta_stride, const float* th,
                   long long th_stride, uint8_t* wimg, long long wimg_stride, int* wshift, long long wshift_stride,
                   void* stream) {
  TcPrepParams pp;
  pp.W = W; pp.L = L; pp.OUT = OUT; pp.T = T; pp.with_bwd = with_bwd;
  pp.theta = theta; pp.theta_stride = theta_stride; pp.th = th; pp.th_stride = th_stride;
  pp.wimg = wimg; pp.wimg_stride = wimg_stride; pp.wshift = wshift; pp.wshift_stride = wshift_stride;
  const int stages = (int)(tc_image_bytes(W, L, with_bwd) / STAGE_BYTES);
  g_launches++;
  tc_prep_kernel<<<dim3(stages, T, 1), dim3(256, 1, 1), 0, (cudaStream_t)stream>>>(pp);
  return cudaGetLastError() != cudaSuccess ? 4 : 0;
}


// ---- hidden-layer weight gradients from the spills (tc_wgrad.cuh) ----
__global__ void __launch_bounds__(256) tc_wgrad_reduce_kernel(int W, int L, int rs, const float* part, long long part_stride,
                                                              float* gout, long long g_stride) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x, l = 1 + blockIdx.y, t = blockIdx.z;
  if (e >= W * W) return;
  float* o = gout + (size_t)t * g_stride + nat_off_layer(W, l) + e;
  float acc = *o;
  for (int z = 1; z < rs; ++z) acc += part[(size_t)(z - 1) * part_stride + ((size_t)t * (L - 1) + (l - 1)) * W * W + e];
  *o = acc;
}

template <int W, bool HV> static cudaError_t wgrad_attr() {
  return cudaFuncSetAttribute(tc_wgrad_kernel<W, HV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WgGeo<W, HV>::SMEM);
}
template <int W, bool HV> static void wgrad_launch(const TcWgradParams* p, dim3 grid, void* stream) {
  tc_wgrad_kernel<W, HV><<<grid, dim3(WgGeo<W, HV>::NT, 1, 1), WgGeo<W, HV>::SMEM, (cudaStream_t)stream>>>(*p);
}
int tc_launch_wgrad(const TcWgradParams* p, void* stream) {
  static PerDeviceOnce once;
  const cudaError_t attr_err = once.run([&]() {
    cudaError_t e = wgrad_attr<128, false>();
    if (e == cudaSuccess) e = wgrad_attr<128, true>();
    if (e == cudaSuccess) e = wgrad_attr<256, false>();
    if (e == cudaSuccess) e = wgrad_attr<256, true>();
    return e;
  });
  if (attr_err != cudaSuccess) return 4;
  g_launches++;
  const int rs = p->row_splits > 1 ? p->row_splits : 1;
  const dim3 grid((p->W / 128) * (p->L - 1), p->T, rs);
  if (p->W == 128) { if (p->want_hv) wgrad_launch<128, true>(p, grid, stream); else wgrad_launch<128, false>(p, grid, stream); }
  else { if (p->want_hv) wgrad_launch<256, true>(p, grid, stream); else wgrad_launch<256, false>(p, grid, stream); }
  if (cudaGetLastError() != cudaSuccess) return 4;
  if (rs > 1) {   // gout(hidden kernels) += parts of splits 1..rs-1, fixed order
    g_launches++;
    const dim3 rgrid((p->W * p->W + 255) / 256, p->L - 1, p->T);
    tc_wgrad_reduce_kernel<<<rgrid, 256, 0, (cudaStream_t)stream>>>(p->W, p->L, rs, p->part, p->part_stride, p->gout, p->g_stride);
    if (cudaGetLastError() != cudaSuccess) return 4;
  }
  return 0;
}

// ---- fixed-order sum of the per-CTA small-parameter partials into the natural gradient layout ----
struct ReduceSmallParams { int W, L, OUT, PS, G; const float* ps; float* gout; long long g_stride; };
__global__ void __launch_bounds__(256) tc_reduce_small_kernel(ReduceSmallParams p) {
  const int t = blockIdx.y;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= p.PS) return;
  const float* src = p.ps + (size_t)t * p.G * p.PS + i;
  float s = 0.f;
  for (int c = 0; c < p.G; ++c) s += src[(size_t)c * p.PS];
  const int W = p.W, L = p.L, OUT = p.OUT;
  int dst;
  if (i < 3 * W) dst = i;                                                     // w0 [2][W], b0 [W]
  else if (i < (L + 2) * W) { const int l = (i - 2 * W) / W, n = i % W; dst = nat_off_layer(W, l) + W * W + n; }   // b_l
  else dst = nat_off_out(W, L) + (i - (L + 2) * W);                           // wL [W][OUT], bL [OUT]
  p.gout[(size_t)t * p.g_stride + dst] = s;
}
int tc_launch_reduce_small(int W, int L, int OUT, int PS, int G, int T, const float* ps, float* gout, long long g_stride,
                           void* stream) {
  ReduceSmallParams rp{W, L, OUT, PS, G, ps, gout, g_stride};
  g_launches++;
  tc_reduce_small_kernel<<<dim3((PS + 255) / 256, T, 1), dim3(256, 1, 1), 0, (cudaStream_t)stream>>>(rp);
  return cudaGetLastError() != cudaSuccess ? 4 : 0;
}

}  // namespace mpde
