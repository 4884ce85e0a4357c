// tc_wgrad.cuh -- hidden-layer weight gradients of the tensor-core path.
//
//   grad W_l [k][n] = sum over the task's jet rows r of  h_{l-1}[r][k] * zbar_l[r][n]        
//
// tc_pde_kernel (GRAD) spills the activation records of every layer and zbar_l in fp32; this kernel is the
// contraction over rows: one CTA per (task, layer, 128-wide k tile) keeps the [128 x W] fp32 result in tensor memory
// for the whole task (fixed summation order: deterministic), recomputes h_{l-1} from the records, splits both operands
// into tf32 hi/lo on the fly (x = hi + lo, hi = x with the low 13 mantissa bits cleared) and issues the three
// kind::tf32 MMAs hi*hi + lo*hi + hi*lo.  The converters write both operands K-major (slab row = neuron, the 128 bytes
// run over 32 jet rows): kind::tf32 with MN-major 128B-swizzled operands returned zeros on B200 (tools/tc_probe.cu).
// No scaling is needed (tf32 has the fp32 exponent), so tiles of very different magnitude accumulate exactly.
#pragma once
#include "tc_kernels.cuh"

namespace mpde {
namespace tc {

// x = hi + lo, both exactly representable in tf32 (round-to-nearest split: the dropped lo*lo term and the residual
// of lo are unbiased; a truncating split leaves a systematic error of the sign of the product)
TC_DEV float tf32_rn(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
TC_DEV void split_tf32(float x, float& hi, float& lo) { hi = tf32_rn(x); lo = tf32_rn(x - hi); }

constexpr int WG_QR = 32;                    // rows per pipeline step (a quarter tile)

template <int W> struct WgGeo {
  static constexpr int NMT = W / 128;
  static constexpr int HBYTES = 128 * 128, ZBYTES = W * 128;  // one K-major slab each: [128 k][32 r], [W n][32 r] tf32
  static constexpr int SLOT = 2 * (HBYTES + ZBYTES);          // hi + lo of both operands
  static constexpr int NCONV = 256, NT = NCONV + 32;
  static constexpr int SMEM = 2 * SLOT + 96;
};

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
  if (tid == 0) {
    if (smem_u32(smem) & 1023u) { printf("tc_wgrad_kernel: dynamic shared memory is not 1024-byte aligned\n"); __trap(); }
    mbar_init(&bars[0], G::NCONV); mbar_init(&bars[1], G::NCONV);
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
  const int nq = n_tiles * (TR / WG_QR);

  if (warp < G::NCONV / 32) {
    // ===== converters: global fp32 -> tf32 hi/lo slabs =====
    const float* rec_t = p.rec + (size_t)task * p.rec_task_stride;
    const float* zsp_t = p.zsp + (size_t)task * p.zsp_task_stride;
    const int k = tid & 127, half = tid >> 7;
    // fp32 master copy of this thread's part of the result: TMEM lane 32 * (warp & 3) + lane, W/2 columns.  The tensor
    // core's own accumulator is only trusted for one tile (48 accumulation steps): its additions are not
    // round-to-nearest, and a sum over all rows of a task drifted by 3e-4 at 16k points (measured).
    float master[W / 2];
    MPDE_UNROLL
    for (int j = 0; j < W / 2; ++j) master[j] = 0.f;
    auto promote = [&](int tt) {
      const int b = tt & 1;
      mbar_wait(&bars[4 + b], (tt >> 1) & 1);
      tc_fence_after();
      MPDE_UNROLL
      for (int c = 0; c < W / 2; c += 32) {
        uint32_t v[32];
        tmem_ld32(tmem_addr(tbase, 32u * (uint32_t)(warp & 3), (uint32_t)(b * W + (warp >> 2) * (W / 2) + c)), v);
        tmem_ld_wait();
        MPDE_UNROLL
        for (int j = 0; j < 32; ++j) master[c + j] += __uint_as_float(v[j]);
      }
      tc_fence_before();
      mbar_arrive(&bars[6 + b]);
    };
    for (int q = 0; q < nq; ++q) {
      const int slot = q & 1;
      if ((q & 3) == 0 && q >= 4) promote(q / 4 - 1);
      if (q >= 2) mbar_wait(&bars[2 + slot], ((q >> 1) - 1) & 1);
      uint8_t* h_hi = smem + (size_t)slot * G::SLOT;
      uint8_t* h_lo = h_hi + G::HBYTES;
      uint8_t* z_hi = h_lo + G::HBYTES;
      uint8_t* z_lo = z_hi + G::ZBYTES;
      const int t = q / (TR / WG_QR), qq = q % (TR / WG_QR);
      int s = 0;
      while (s + 1 < p.n_seg && t >= p.seg_tile_begin[s + 1]) ++s;
      const int jk = p.seg_jk[s];
      const int tile = t - p.seg_tile_begin[s];
      // zbar quarter [32 rows][W] fp32 -> B operand, K-major: slab row = output neuron n, the 128 bytes run over the
      // 32 rows r.  A thread takes one n and four consecutive r (four coalesced loads, one 16-byte store).
      {
        const float* zq = zsp_t + ((size_t)t * (L - 1) + (l - 1)) * TR * W + (size_t)qq * WG_QR * W;
        MPDE_UNROLL
        for (int j = 0; j < W * (WG_QR / 4) / G::NCONV; ++j) {
          const int idx = tid + G::NCONV * j;
          const int nn = idx % W, g = idx / W;                 // g: group of 4 rows
          float x[4], hi[4], lo[4];
          MPDE_UNROLL
          for (int i = 0; i < 4; ++i) x[i] = zq[(size_t)(4 * g + i) * W + nn];
          MPDE_UNROLL
          for (int i = 0; i < 4; ++i) split_tf32(x[i], hi[i], lo[i]);
          const uint32_t off = (uint32_t)(nn >> 3) * 1024u + (uint32_t)(nn & 7) * 128u + (((uint32_t)g ^ (uint32_t)(nn & 7)) << 4);
          *reinterpret_cast<float4*>(z_hi + off) = make_float4(hi[0], hi[1], hi[2], hi[3]);
          *reinterpret_cast<float4*>(z_lo + off) = make_float4(lo[0], lo[1], lo[2], lo[3]);
        }
      }
      // h_{l-1} quarter, neurons kt*128 .. +128, recomputed from the records of layer l-1 -> A operand, K-major:
      // slab row = input neuron k, the 128 bytes run over the 32 rows r (4 channels of a point = one 16-byte chunk)
      {
        const int NF = jk == JET_V ? 2 : (jk == JET_G ? 4 : 5);
        const int CP = jk == JET_V ? 1 : 4, PTS = TR / CP, PPQ = WG_QR / CP;
        const float* rp = rec_t + p.rec_seg_base[s] + ((size_t)tile * L + (l - 1)) * (size_t)(NF * PTS) * W + kt * 128 + k;
        const uint32_t rowoff = (uint32_t)(k >> 3) * 1024u + (uint32_t)(k & 7) * 128u;
        if (jk == JET_V) {
          // 32 points = 32 rows: this thread's half takes the chunks g = half, half + 2, ... (4 points each)
          for (int g = half; g < WG_QR / 4; g += 2) {
            float x[4], hi[4], lo[4];
            MPDE_UNROLL
            for (int i = 0; i < 4; ++i) x[i] = rp[(size_t)(qq * PPQ + 4 * g + i) * W];
            MPDE_UNROLL
            for (int i = 0; i < 4; ++i) split_tf32(x[i], hi[i], lo[i]);
            const uint32_t off = rowoff + (((uint32_t)g ^ (uint32_t)(k & 7)) << 4);
            *reinterpret_cast<float4*>(h_hi + off) = make_float4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<float4*>(h_lo + off) = make_float4(lo[0], lo[1], lo[2], lo[3]);
          }
        } else {
          for (int pq = half; pq < PPQ; pq += 2) {           // one point = rows 4 pq .. 4 pq + 3 = chunk pq
            const int pl = qq * PPQ + pq;
            float a[4] = {0.f, 0.f, 0.f, 0.f}, hi[4], lo[4];
            float sv[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
            for (int f = 0; f < NF; ++f) sv[f] = rp[(size_t)(f * PTS + pl) * W];
            if (jk == JET_G) jet_act_from_saved<JET_G>(p.omega, sv, a);
            else if (jk == JET_LAP) jet_act_from_saved<JET_LAP>(p.omega, sv, a);
            else jet_act_from_saved<JET_XX>(p.omega, sv, a);
            MPDE_UNROLL
            for (int i = 0; i < 4; ++i) split_tf32(a[i], hi[i], lo[i]);
            const uint32_t off = rowoff + (((uint32_t)pq ^ (uint32_t)(k & 7)) << 4);
            *reinterpret_cast<float4*>(h_hi + off) = make_float4(hi[0], hi[1], hi[2], hi[3]);
            *reinterpret_cast<float4*>(h_lo + off) = make_float4(lo[0], lo[1], lo[2], lo[3]);
          }
        }
      }
      fence_proxy_async_smem();
      mbar_arrive(&bars[slot]);
    }
    // ===== last tile, then the result: [128 k][W n] -> grad W_l =====
    promote(n_tiles - 1);
    float* out = p.gout + (size_t)task * p.g_stride + nat_off_layer(W, l) + (size_t)(kt * 128 + 32 * (warp & 3) + lane) * W +
                 (warp >> 2) * (W / 2);
    MPDE_UNROLL
    for (int j = 0; j < W / 2; ++j) out[j] = master[j];   // gout rows are only 4-byte aligned (P is odd-ish)
  } else {
    // ===== MMA issuer =====
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(FMT_TF32, 128, W, 0, 0);
      for (int q = 0; q < nq; ++q) {
        const int slot = q & 1;
        const int tt = q >> 2, b = tt & 1;              // tile and its accumulator buffer
        const uint32_t d = tbase + (uint32_t)(b * W);
        uint32_t acc = (q & 3) ? 1u : 0u;               // a fresh accumulator per tile
        if ((q & 3) == 0 && tt >= 2) { mbar_wait(&bars[6 + b], ((tt >> 1) - 1) & 1); tc_fence_after(); }
        mbar_wait(&bars[slot], (q >> 1) & 1);
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
        if ((q & 3) == 3) tc_commit(&bars[4 + b]);
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
