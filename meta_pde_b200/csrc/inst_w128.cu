// width-128 instantiation of the fused jet kernels (NTY = 8 row groups: 32 points per tile)
#include "inst.cuh"
namespace mpde {
int dispatch_w128(int pde, int mode, int op, const KParams* p, int gx, int gy, void* stream, KernelInfo* info) {
  return run_width<128, 8>(pde, mode, op, p, gx, gy, stream, info);
}
}  // namespace mpde
