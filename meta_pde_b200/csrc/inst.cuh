// inst.cuh -- instantiates pde_kernel for one width and dispatches on (pde, mode).
#pragma once
#include "host_common.h"
#include <mutex>

namespace mpde {

template <int W, int NTY, int PDE, int MODE, int ACT>
int run_inst(int op, const KParams* p, int gx, int gy, void* stream, KernelInfo* info) {
  typedef Geo<W, NTY> G;
  auto kfn = KernelSel<W, NTY, PDE, MODE, ACT>::fn();
  const size_t smem = G::smem_bytes(MODE);
  static int occ = 0;
#ifndef MPDE_EMU
  static PerDeviceOnce once;
  const cudaError_t attr_err = once.run([&]() {
    cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int o = 0;
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, kfn, G::NT, smem);
    if (e == cudaSuccess) occ = o;      // identical on every device of the node (same GPU model)
    return e;
  });
  if (attr_err != cudaSuccess || occ < 1) return 4;
#else
  occ = 2;
#endif
  if (op == 0) {
    info->nt = G::NT; info->smem = smem; info->occ = occ; info->np = G::NP;
    return 0;
  }
  KParams kp = *p;
  MPDE_LAUNCH(kfn, dim3(gx, gy, 1), dim3(G::NT, 1, 1), smem, stream, kp);
  return MPDE_CHECK_LAUNCH() ? 4 : 0;
}

template <int W, int NTY, int PDE, int ACT>
int run_mode(int mode, int op, const KParams* p, int gx, int gy, void* stream, KernelInfo* info) {
  switch (mode) {
    case MODE_LOSS: return run_inst<W, NTY, PDE, MODE_LOSS, ACT>(op, p, gx, gy, stream, info);
    case MODE_GRAD: return run_inst<W, NTY, PDE, MODE_GRAD, ACT>(op, p, gx, gy, stream, info);
    case MODE_HVP:  return run_inst<W, NTY, PDE, MODE_HVP, ACT>(op, p, gx, gy, stream, info);
  }
  return 1;
}

template <int W, int NTY, int ACT = ACT_SIN>
int run_width(int pde, int mode, int op, const KParams* p, int gx, int gy, void* stream, KernelInfo* info) {
  switch (pde) {
    case PDE_POISSON: return run_mode<W, NTY, PDE_POISSON, ACT>(mode, op, p, gx, gy, stream, info);
    case PDE_BURGERS: return run_mode<W, NTY, PDE_BURGERS, ACT>(mode, op, p, gx, gy, stream, info);
    case PDE_HYPER:   return run_mode<W, NTY, PDE_HYPER, ACT>(mode, op, p, gx, gy, stream, info);
    case PDE_EVAL1:   return run_inst<W, NTY, PDE_EVAL1, MODE_LOSS, ACT>(op, p, gx, gy, stream, info);
    case PDE_EVAL2:   return run_inst<W, NTY, PDE_EVAL2, MODE_LOSS, ACT>(op, p, gx, gy, stream, info);
  }
  return 1;
}

}  // namespace mpde
