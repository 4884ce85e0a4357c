// width-128 instantiation of the fused jet kernels for the `--nosiren` (swish) network
#include "inst.cuh"
namespace mpde {
int dispatch_sw_w128(int pde, int mode, int op, const KParams* p, int gx, int gy, void* stream, KernelInfo* info) {
  return run_width<128, 8, ACT_SWISH>(pde, mode, op, p, gx, gy, stream, info);
}
}  // namespace mpde
