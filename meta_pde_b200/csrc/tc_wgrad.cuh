// tc_wgrad.cuh -- hidden-layer weight gradients (and Hessian-vector products) of the tensor-core path.
//
//   grad W_l [k][n]   = sum over the task's jet rows r of  h_{l-1}[r][k] * zbar_l[r][n]                 
//   (H v)_W_l [k][n]  = sum_r  hdot_{l-1}[r][k] * zbar_l[r][n] + h_{l-1}[r][k] * zbardot_l[r][n]
//
// tc_pde_kernel / tc_hvp_kernel spill the activation records of every layer and zbar_l (zbardot_l) in fp32; this
// kernel is the contraction over rows.  One CTA per (task, layer, 128-wide k tile) walks the task's rows in steps of
// 32: 256 converter threads recompute h (hdot) from the records, split both operands into tf32 hi/lo
// (round-to-nearest, x = hi + lo) and write them K-major into a two-slot shared-memory pipeline (slab row = neuron,
// the 128 bytes run over the 32 rows; kind::tf32 with MN-major swizzled operands returned zeros on B200,
// tools/tc_probe.cu); one thread issues the three kind::tf32 MMAs hi*hi + lo*hi + hi*lo per 8 rows.  tf32 keeps the
// fp32 exponent, so rows of very different magnitude need no scaling.
//
// The tensor core's accumulator is trusted for 4 steps only (48 accumulations): its additions are not
// round-to-nearest and a sum over all rows of a task drifted by 3e-4 at 16k points (measured).  Every 4 steps the
// converter threads add the TMEM accumulator into an fp32 master copy in registers (two TMEM buffers alternate).
// Fixed summation order: deterministic.
#pragma once
#include "tc_kernels.cuh"

namespace mpde {
namespace tc {

TC_DEV float tf32_rn(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
TC_DEV void split_tf32(float x, float& hi, float& lo) { hi = tf32_rn(x); lo = tf32_rn(x - hi); }

constexpr int WG_QR = 32;                    // rows per pipeline step
constexpr int WG_GROUP = 4;                  // steps per accumulator generation

template <int W> struct WgGeo {
  static constexpr int NMT = W / 128;
  static constexpr int HBYTES = 128 * 128, ZBYTES = W * 128;  // one K-major slab each: [128 k][32 r], [W n][32 r] tf32
  static constexpr int SLOT = 2 * (HBYTES + ZBYTES);          // hi + lo of both operands
  static constexpr int NGRP = 2, NCONV = 256 * NGRP, NT = NCONV + 32;   // converter groups alternate pipeline steps
  static constexpr int SMEM = 2 * SLOT + 96;
};

// 16-byte chunk `chunk` (4 consecutive rows) of slab row `row`, swizzled
TC_DEV uint32_t wg_chunk_off(int row, int chunk) {
  return (uint32_t)(row >> 3) * 1024u + (uint32_t)(row & 7) * 128u + (((uint32_t)chunk ^ (uint32_t)(row & 7)) << 4);
}
TC_DEV void wg_store4(uint8_t* hi_base, uint8_t* lo_base, uint32_t off, const float (&x)[4]) {
  float hi[4], lo[4];
  MPDE_UNROLL
  for (int i = 0; i < 4; ++i) split_tf32(x[i], hi[i], lo[i]);
  *reinterpret_cast<float4*>(hi_base + off) = make_float4(hi[0], hi[1], hi[2], hi[3]);
  *reinterpret_cast<float4*>(lo_base + off) = make_float4(lo[0], lo[1], lo[2], lo[3]);
}

// B operand: 32 rows x W of zbar (or zbardot), fp32 [row][n] contiguous -> slab row n, chunk = group of 4 rows.
// Iterations [J0, J0 + NJ) of the group's 256 threads; loads and stores are separate calls so that the caller can
// keep several batches of loads in flight.
template <int W, int J0, int NJ>
TC_DEV void wg_load_z(const float* zq, int tid, float (&x)[NJ][4]) {
  MPDE_UNROLL
  for (int j = 0; j < NJ; ++j) {
    const int idx = tid + 256 * (J0 + j), nn = idx % W, g = idx / W;
    MPDE_UNROLL
    for (int i = 0; i < 4; ++i) x[j][i] = zq[(size_t)(4 * g + i) * W + nn];
  }
}
template <int W, int J0, int NJ>
TC_DEV void wg_store_z(uint8_t* z_hi, uint8_t* z_lo, int tid, const float (&x)[NJ][4]) {
  MPDE_UNROLL
  for (int j = 0; j < NJ; ++j) {
    const int idx = tid + 256 * (J0 + j), nn = idx % W, g = idx / W;
    wg_store4(z_hi, z_lo, wg_chunk_off(nn, g), x[j]);
  }
}

// A operand: h (TAN = false) or hdot (TAN = true) of 32 rows x 128 neurons, recomputed from the records
// [field][point][W]; this thread: neuron k (slab row), every other chunk starting at `half`.
template <int JK, bool TAN, bool HVPREC>
TC_DEV void wg_convert_h(const float* rp, int W, int PTS, int p0, float om, uint8_t* h_hi, uint8_t* h_lo, int k, int half) {
  typedef JetTraits<JK> JT;
  constexpr int NF = JT::NF, C = JT::C;
  if (JK == JET_V) {
    // 32 points = 32 rows; chunk g = points 4g..4g+3.  h = S (field 0); hdot = om * Cc * zdot (fields 1, NF)
    float s0[4][4], s1[4][4];
    MPDE_UNROLL
    for (int j = 0; j < 4; ++j)
      MPDE_UNROLL
      for (int i = 0; i < 4; ++i) {
        const int pl = p0 + 4 * (half + 2 * j) + i;
        s0[j][i] = rp[(size_t)((TAN ? 1 : 0) * PTS + pl) * W];
        if (TAN) s1[j][i] = rp[(size_t)(NF * PTS + pl) * W];
      }
    MPDE_UNROLL
    for (int j = 0; j < 4; ++j) {
      float x[4];
      MPDE_UNROLL
      for (int i = 0; i < 4; ++i) x[i] = TAN ? om * s0[j][i] * s1[j][i] : s0[j][i];
      wg_store4(h_hi, h_lo, wg_chunk_off(k, half + 2 * j), x);
    }
  } else {
    // 8 points; chunk pq = the 4 channel rows of point pq
    constexpr int NL = NF + ((TAN && HVPREC) ? C : 0);
    float f[4][NL];
    MPDE_UNROLL
    for (int j = 0; j < 4; ++j) {
      const int pl = p0 + half + 2 * j;
      MPDE_UNROLL
      for (int q = 0; q < NL; ++q) f[j][q] = rp[(size_t)(q * PTS + pl) * W];
    }
    MPDE_UNROLL
    for (int j = 0; j < 4; ++j) {
      float a[4] = {0.f, 0.f, 0.f, 0.f};
      if (TAN) jet_act_tan<JK>(om, f[j], f[j] + NF, a);
      else jet_act_from_saved<JK>(om, f[j], a);
      wg_store4(h_hi, h_lo, wg_chunk_off(k, half + 2 * j), a);
    }
  }
}

template <int W>
__global__ void __launch_bounds__(WgGeo<W>::NT, 1) tc_wgrad_kernel(const __grid_constant__ TcWgradParams p) {
  typedef WgGeo<W> G;
  uint8_t* smem = tc_smem_raw;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 2 * G::SLOT);   // full[2], empty[2], dfull[2], dfree[2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int task = blockIdx.y;
  const int l = 1 + blockIdx.x / G::NMT, kt = blockIdx.x % G::NMT;
  const int L = p.L;
  const int n_tiles = p.seg_tile_begin[p.n_seg];
  const int qpt = p.tile_rows / WG_QR;                 // 32-row quarters per tile
  const int npair = p.want_hv ? 2 : 1;
  const int nsteps = n_tiles * qpt * npair;
  if (tid == 0) {
    if (smem_u32(smem) & 1023u) { printf("tc_wgrad_kernel: dynamic shared memory is not 1024-byte aligned\n"); __trap(); }
    mbar_init(&bars[0], 256); mbar_init(&bars[1], 256);
    mbar_init(&bars[2], 1); mbar_init(&bars[3], 1);
    mbar_init(&bars[4], 1); mbar_init(&bars[5], 1);
    mbar_init(&bars[6], G::NCONV); mbar_init(&bars[7], G::NCONV);
    mbar_fence_init();
  }
  if (warp == 0) tmem_alloc(tmem_slot, 2 * W);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = *tmem_slot;

  if (warp < G::NCONV / 32) {
    // ===== converters: group g (8 warps) fills pipeline slot g on steps st = g, g + 2, ... =====
    const float* rec_t = p.rec + (size_t)task * p.rec_task_stride;
    const int grp = warp >> 3, gtid = tid & 255;
    const int k = gtid & 127, half = gtid >> 7;
    // fp32 master copy of this thread's part of the result: TMEM lane 32 * (warp & 3) + lane, W/4 columns
    constexpr int MC = W / 4;
    float master[MC];
    MPDE_UNROLL
    for (int j = 0; j < MC; ++j) master[j] = 0.f;
    auto promote = [&](int gen) {
      const int b = gen & 1;
      mbar_wait(&bars[4 + b], (gen >> 1) & 1);
      tc_fence_after();
      MPDE_UNROLL
      for (int c = 0; c < MC; c += 32) {
        uint32_t v[32];
        tmem_ld32(tmem_addr(tbase, 32u * (uint32_t)(warp & 3), (uint32_t)(b * W + (warp >> 2) * MC + c)), v);
        tmem_ld_wait();
        MPDE_UNROLL
        for (int j = 0; j < 32; ++j) master[c + j] += __uint_as_float(v[j]);
      }
      tc_fence_before();
      mbar_arrive(&bars[6 + b]);
    };
    int done_gen = 0;                                   // generations 0 .. done_gen-1 are in the master copy
    for (int st = grp; st < nsteps; st += G::NGRP) {
      const int slot = st & 1;
      uint8_t* h_hi = smem + (size_t)slot * G::SLOT;
      uint8_t* h_lo = h_hi + G::HBYTES;
      uint8_t* z_hi = h_lo + G::HBYTES;
      uint8_t* z_lo = z_hi + G::ZBYTES;
      const int pair = st % npair, q = st / npair;
      const int t = q / qpt, qq = q % qpt;
      int s = 0;
      while (s + 1 < p.n_seg && t >= p.seg_tile_begin[s + 1]) ++s;
      const int jk = p.seg_jk[s];
      const int tile = t - p.seg_tile_begin[s];
      // pair 0 of an H v product: (hdot, zbar); pair 1: (h, zbardot); gradient: (h, zbar)
      const bool tan = p.want_hv && pair == 0;
      const float* zsrc = (p.want_hv && pair == 1) ? p.zspd : p.zsp;
      const float* zq = zsrc + (size_t)task * p.zsp_task_stride + ((size_t)t * (L - 1) + (l - 1)) * p.tile_rows * W +
                        (size_t)qq * WG_QR * W;
      // all z loads are in flight while the h records are loaded and converted
      constexpr int NIT = W * (WG_QR / 4) / 256, NJA = NIT / 2, NJB = NIT - NJA;
      float zA[NJA][4], zB[NJB][4];
      wg_load_z<W, 0, NJA>(zq, gtid, zA);
      wg_load_z<W, NJA, NJB>(zq, gtid, zB);
      // Promotion lags the MMAs by two steps (no pipeline drain at a generation boundary); the slot is only needed
      // for the stores, so the z loads above are already in flight while this thread waits.
      while (st >= 2 && done_gen < (st - 2) / WG_GROUP) promote(done_gen++);
      if (st >= 2) mbar_wait(&bars[2 + slot], ((st >> 1) - 1) & 1);
      {
        const int CP = jk == JET_V ? 1 : 4, PTS = p.tile_rows / CP, p0 = qq * (WG_QR / CP);
        const int NF = jk == JET_V ? 2 : (jk == JET_G ? 4 : 5);
        const int NFT = NF + (p.hvp_records ? (jk == JET_V ? 1 : (jk == JET_G ? 3 : 4)) : 0);
        const float* rp = rec_t + p.rec_seg_base[s] + ((size_t)tile * L + (l - 1)) * (size_t)(NFT * PTS) * W + kt * 128 + k;
        const float om = p.omega;
        if (!tan) {
          if (jk == JET_V) wg_convert_h<JET_V, false, false>(rp, W, PTS, p0, om, h_hi, h_lo, k, half);
          else if (jk == JET_G) wg_convert_h<JET_G, false, false>(rp, W, PTS, p0, om, h_hi, h_lo, k, half);
          else if (jk == JET_LAP) wg_convert_h<JET_LAP, false, false>(rp, W, PTS, p0, om, h_hi, h_lo, k, half);
          else wg_convert_h<JET_XX, false, false>(rp, W, PTS, p0, om, h_hi, h_lo, k, half);
        } else {
          if (jk == JET_V) wg_convert_h<JET_V, true, true>(rp, W, PTS, p0, om, h_hi, h_lo, k, half);
          else if (jk == JET_G) wg_convert_h<JET_G, true, true>(rp, W, PTS, p0, om, h_hi, h_lo, k, half);
          else if (jk == JET_LAP) wg_convert_h<JET_LAP, true, true>(rp, W, PTS, p0, om, h_hi, h_lo, k, half);
          else wg_convert_h<JET_XX, true, true>(rp, W, PTS, p0, om, h_hi, h_lo, k, half);
        }
      }
      wg_store_z<W, 0, NJA>(z_hi, z_lo, gtid, zA);
      wg_store_z<W, NJA, NJB>(z_hi, z_lo, gtid, zB);
      fence_proxy_async_smem();
      mbar_arrive(&bars[slot]);
    }
    // ===== remaining generations, then the result: [128 k][W n] -> grad W_l =====
    while (done_gen <= (nsteps - 1) / WG_GROUP) promote(done_gen++);
    float* out = p.gout + (size_t)task * p.g_stride + nat_off_layer(W, l) + (size_t)(kt * 128 + 32 * (warp & 3) + lane) * W +
                 (warp >> 2) * MC;
    MPDE_UNROLL
    for (int j = 0; j < MC; ++j) out[j] = master[j];   // gout rows are only 4-byte aligned (P is odd-ish)
  } else {
    // ===== MMA issuer =====
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(FMT_TF32, 128, W, 0, 0);
      for (int st = 0; st < nsteps; ++st) {
        const int slot = st & 1;
        const int gen = st / WG_GROUP, b = gen & 1;     // accumulator generation and its TMEM buffer
        const uint32_t d = tbase + (uint32_t)(b * W);
        uint32_t acc = (st % WG_GROUP) ? 1u : 0u;       // a fresh accumulator per generation
        if ((st % WG_GROUP) == 0 && gen >= 2) { mbar_wait(&bars[6 + b], ((gen >> 1) - 1) & 1); tc_fence_after(); }
        mbar_wait(&bars[slot], (st >> 1) & 1);
        tc_fence_after();
        const uint32_t h_hi = smem_u32(smem + (size_t)slot * G::SLOT), h_lo = h_hi + G::HBYTES;
        const uint32_t z_hi = h_lo + G::HBYTES, z_lo = z_hi + G::ZBYTES;
        MPDE_UNROLL
        for (int j = 0; j < WG_QR / 8; ++j) {           // 8 rows of r (32 bytes of K) per MMA
          const uint64_t ah = desc_kmajor(h_hi + j * 32), al = desc_kmajor(h_lo + j * 32);
          const uint64_t bh = desc_kmajor(z_hi + j * 32), bl = desc_kmajor(z_lo + j * 32);
          mma_tf32_ss(d, ah, bh, idesc, acc); acc = 1;
          mma_tf32_ss(d, al, bh, idesc, 1);
          mma_tf32_ss(d, ah, bl, idesc, 1);
        }
        tc_commit(&bars[2 + slot]);
        if ((st % WG_GROUP) == WG_GROUP - 1 || st == nsteps - 1) tc_commit(&bars[4 + b]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 0) tmem_dealloc(tbase, 2 * W);
}

}  // namespace tc
}  // namespace mpde
