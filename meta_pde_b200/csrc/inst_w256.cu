// width-256 instantiation of the fused jet kernels (NTY = 4 row groups: 16 points per tile)
#include "inst.cuh"
namespace mpde {
int dispatch_w256(int pde, int mode, int op, const KParams* p, int gx, int gy, void* stream, KernelInfo* info) {
  return run_width<256, 4>(pde, mode, op, p, gx, gy, stream, info);
}
}  // namespace mpde
