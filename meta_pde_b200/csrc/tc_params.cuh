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
  const int* wshift;               // [T or 1][2*TC_MAX_L]   weight shift per step (hidden layers 1..L-1, output layer at L)
  long long wshift_stride;
  // GRAD / HVP (spills consumed by tc_wgrad_kernel; all indexed per task)
  float* rec;                      // activation records: per (seg tile, layer): [NF][W/128][PTS][128]
  long long rec_task_stride;       // floats
  long long rec_seg_base[MAX_SEG]; // floats, first tile of the segment inside the task's block
  float* zsp;                      // pre-activation adjoints zbar_l, l = 1..L-1: per (task tile, l-1): [TR rows][W]
  long long zsp_task_stride;       // floats
  float* psmall;                   // [T][G][PS] per-CTA partials of the small parameters (first/last layer, biases)
  int PS;
  int rec_prefetch;                // the record-prefetch warp asks L2 for the next reverse layer's records (MPDE_REC_PREFETCH=1; off by default)
};

// per-CTA partial layout of the small parameters: w0 [2][W], b0 [W], b_l [W] (l = 1..L-1), wL [W][OUT], bL [OUT]
MPDE_HD int ps_off_bias(int W, int l) { return 2 * W + l * W; }                  // l = 0..L-1
MPDE_HD int ps_off_wl(int W, int L) { return (L + 2) * W; }
MPDE_HD int ps_count(int W, int L, int OUT) { return (L + 2) * W + W * OUT + OUT; }
// Record block of one (tile, layer), round-2 layout: [k tile][step of 16 jet rows][field][point of the step][128 neurons]
// (PPS = points per 16-row step: 4 jet points or 16 value points), so that everything tc_wgrad_kernel needs for one
// pipeline step of one k tile -- a range of fields of the step's points -- is ONE contiguous piece (it was one 2 KB piece
// per field).  Offset (floats) of (field f, point pl) of neuron n; NFX = fields per point, PTS = points per tile.
MPDE_HD size_t rec_off(int NFX, int PTS, int PPS, int f, int pl, int n) {
  return (size_t)(n >> 7) * (size_t)(NFX * PTS * 128) + ((size_t)(pl / PPS) * NFX + f) * (size_t)(PPS * 128) +
         (size_t)(pl % PPS) * 128 + (size_t)(n & 127);
}
// floats of activation records per (tile, layer): NF * PTS * W
// (GRAD tiles: 64 rows = 16 jet points or 64 value points)
MPDE_HD int rec_fields_points(int jk) { return jk == JET_V ? 2 * 64 : (jk == JET_G ? 4 * 16 : 5 * 16); }
// HVP tiles (64 rows): saved fields + the tangent pre-activations zdot
MPDE_HD int rec_fields_points_hvp(int jk) { return jk == JET_V ? 3 * 64 : (jk == JET_G ? 7 * 16 : 9 * 16); }


struct TcWgradParams {
  int W, L, T, n_seg;
  float omega;
  int tile_rows;                            // jet rows per spilled tile (64: GRAD and HVP spills)
  int hvp_records;                          // records carry the tangent pre-activations zdot after the NF saved fields
  int want_hv;                              // 0: grad W_l = h^T zbar;  1: (H v)_W_l = hdot^T zbar + h^T zbardot
  int seg_tile_begin[MAX_SEG + 1];
  int seg_jk[MAX_SEG];
  long long rec_seg_base[MAX_SEG];
  const float* rec; long long rec_task_stride;
  const float* zsp; long long zsp_task_stride;     // zbar_l
  const float* zspd;                               // zbardot_l (HVP spills), same layout and stride
  float* gout; long long g_stride;          // [T][P] natural layout; hidden kernels are written (not accumulated)
  int row_splits;                           // grid.z: split z contracts tiles [n_tiles*z/splits, n_tiles*(z+1)/splits)
  float* part; long long part_stride;       // results of splits 1.. : [split-1][T][L-1][W][W], summed into gout in fixed order
};

struct HvParams {
  TcParams t;                      // primal: segments, theta-hat, W image + shifts, records, zbar spill, small partials (grad)
  const uint8_t* wimg_d;           // direction image (same layout as t.wimg, built from v / theta-hat-dot)
  long long wimg_d_stride;
  const int* wshift_d; long long wshift_d_stride;
  float* zspd;                     // zbardot spill (layout of t.zsp)
  float* psmall_h;                 // [T][G][PS] per-CTA partials of (H v) on the small parameters
  int want_grad;                   // also accumulate the small-parameter gradient into t.psmall
};

}  // namespace tc
}  // namespace mpde
