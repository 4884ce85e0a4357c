// width-32 instantiation of the fused jet kernels (NTY = 8 row groups: 32 points per tile)
#include "inst.cuh"
namespace mpde {
int dispatch_w32(int pde, int mode, int op, const KParams* p, int gx, int gy, void* stream, KernelInfo* info) {
  return run_width<32, 8>(pde, mode, op, p, gx, gy, stream, info);
}
}  // namespace mpde
