// tc_inst.cu -- instantiation and launch of the tensor-core (tcgen05) jet kernels for widths 128 and 256.
#include "host_common.h"
#include "tc_kernels.cuh"
#include <mutex>

namespace mpde {

using namespace tc;

template <int W, int PDE, int MODE>
static int tc_run_inst(int op, const TcParams* p, int gx, int gy, void* stream, KernelInfo* info) {
  typedef TcGeo<W> G;
  auto kfn = tc_pde_kernel<W, PDE, MODE>;
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [&]() { attr_err = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G::SMEM); });
  if (attr_err != cudaSuccess) return 4;
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

// modes the tensor-core path implements for this width / depth (the FP32-FMA kernels serve everything else)
bool tc_supported(int W, int L, int mode) {
  return (W == 128 || W == 256) && L >= 2 && L <= TC_MAX_L && mode == MODE_LOSS;
}

int dispatch_tc(int W, int pde, int mode, int op, const TcParams* p, int gx, int gy, void* stream, KernelInfo* info) {
  if (W == 128) return tc_run_width<128>(pde, mode, op, p, gx, gy, stream, info);
  if (W == 256) return tc_run_width<256>(pde, mode, op, p, gx, gy, stream, info);
  return 1;
}

size_t tc_image_bytes(int W, int L) { return W == 128 ? TcGeo<128>::image_bytes(L) : TcGeo<256>::image_bytes(L); }

int tc_launch_prep(int W, int L, int OUT, int T, const float* theta, long long theta_stride, const float* th,
                   long long th_stride, uint8_t* wimg, long long wimg_stride, int* wshift, long long wshift_stride,
                   void* stream) {
  TcPrepParams pp;
  pp.W = W; pp.L = L; pp.OUT = OUT; pp.T = T;
  pp.theta = theta; pp.theta_stride = theta_stride; pp.th = th; pp.th_stride = th_stride;
  pp.wimg = wimg; pp.wimg_stride = wimg_stride; pp.wshift = wshift; pp.wshift_stride = wshift_stride;
  const int stages = (int)(tc_image_bytes(W, L) / STAGE_BYTES);
  g_launches++;
  tc_prep_kernel<<<dim3(stages, T, 1), dim3(256, 1, 1), 0, (cudaStream_t)stream>>>(pp);
  return cudaGetLastError() != cudaSuccess ? 4 : 0;
}

}  // namespace mpde
