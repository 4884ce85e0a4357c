// host_common.h -- host-side glue shared by the translation units of libmetapde_b200.
#pragma once
#include "compat.cuh"
#include "pde_kernels.cuh"
#include <atomic>
#include <mutex>

namespace mpde {

struct KernelInfo {
  int nt;            // threads per CTA
  size_t smem;       // dynamic shared memory bytes
  int occ;           // resident CTAs per SM
  int np;            // points per channel-jet tile (value-only tiles hold 4*np)
};

extern std::atomic<long long> g_launches;
int device_sm_count();

#ifndef MPDE_EMU
// One-time per-DEVICE kernel attributes (cudaFuncSetAttribute is per device: a process that drives several GPUs must
// set the dynamic shared memory limit on each of them).  run(f) calls f once for the calling thread's current device.
struct PerDeviceOnce {
  static constexpr int MAXDEV = 64;
  std::once_flag flags[MAXDEV];
  cudaError_t err[MAXDEV];
  template <class F> cudaError_t run(F&& f) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAXDEV) return cudaErrorInvalidDevice;
    std::call_once(flags[dev], [&]() { err[dev] = f(); });
    return err[dev];
  }
};
#endif

// op 0: fill *info; op 1: launch.  One definition per width (inst_w*.cu) so the widths compile in parallel.
int dispatch_w32(int pde, int mode, int op, const KParams* p, int gx, int gy, void* stream, KernelInfo* info);
int dispatch_w64(int pde, int mode, int op, const KParams* p, int gx, int gy, void* stream, KernelInfo* info);
int dispatch_w128(int pde, int mode, int op, const KParams* p, int gx, int gy, void* stream, KernelInfo* info);
int dispatch_w256(int pde, int mode, int op, const KParams* p, int gx, int gy, void* stream, KernelInfo* info);
// the same kernels with swish hidden layers (`--nosiren`, inst_sw_w*.cu)
int dispatch_sw_w32(int pde, int mode, int op, const KParams* p, int gx, int gy, void* stream, KernelInfo* info);
int dispatch_sw_w64(int pde, int mode, int op, const KParams* p, int gx, int gy, void* stream, KernelInfo* info);
int dispatch_sw_w128(int pde, int mode, int op, const KParams* p, int gx, int gy, void* stream, KernelInfo* info);
int dispatch_sw_w256(int pde, int mode, int op, const KParams* p, int gx, int gy, void* stream, KernelInfo* info);

// tensor-core (tcgen05) path, widths 128 / 256 (tc_inst.cu; not part of the emulator build)
#ifndef MPDE_EMU
namespace tc { struct TcParams; struct TcWgradParams; struct HvParams; }
int dispatch_tc_hvp(int W, int pde, int op, const tc::HvParams* p, int gx, int gy, void* stream, KernelInfo* info);
bool tc_supported(int W, int L, int mode);
int dispatch_tc(int W, int pde, int mode, int op, const tc::TcParams* p, int gx, int gy, void* stream, KernelInfo* info);
size_t tc_image_bytes(int W, int L, int with_bwd);
int tc_launch_wgrad(const tc::TcWgradParams* p, void* stream);
int tc_launch_reduce_small(int W, int L, int OUT, int PS, int G, int T, const float* ps, float* gout, long long g_stride,
                           void* stream);
int tc_launch_prep(int W, int L, int OUT, int T, int with_bwd, const float* theta, long long theta_stride, const float* th,
                   long long th_stride, uint8_t* wimg, long long wimg_stride, int* wshift, long long wshift_stride,
                   void* stream);
#endif

#ifdef MPDE_EMU
#define MPDE_LAUNCH(kfn, grid, block, smem, stream, ...)                                   \
  do { ::mpde::g_launches++; emu::launch(grid, block, smem, [=]() { kfn(__VA_ARGS__); }); } while (0)
#define MPDE_CHECK_LAUNCH() 0
#else
#define MPDE_LAUNCH(kfn, grid, block, smem, stream, ...)                                   \
  do { ::mpde::g_launches++; kfn<<<grid, block, smem, (cudaStream_t)(stream)>>>(__VA_ARGS__); } while (0)
#define MPDE_CHECK_LAUNCH() (cudaGetLastError() != cudaSuccess)
#endif

}  // namespace mpde
