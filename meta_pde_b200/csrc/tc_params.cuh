// tc_params.cuh -- launch parameters of the tensor-core jet kernels (shared by api.cu and tc_inst.cu).
#pragma once
#include "pde_kernels.cuh"

namespace mpde {
namespace tc {

constexpr int TC_MAX_L = 16;     // deepest network the tensor-core path takes (per-layer shift tables live in shared memory)

struct TcParams {
  KParams k;                       // segments, points, pde params, theta-hat (first layer, biases, output layer)
  const uint8_t* wimg;             // [T or 1][image_bytes]   forward weight image (tc_prep_kernel)
  long long wimg_stride;           // bytes between tasks (0: shared)
  const int* wshift;               // [T or 1][TC_MAX_L]     weight shift per step (hidden layers 1..L-1, output layer at L)
  long long wshift_stride;
};

}  // namespace tc
}  // namespace mpde
