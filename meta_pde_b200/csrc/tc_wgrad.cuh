// tc_wgrad.cuh -- hidden-layer weight gradients (and Hessian-vector products) of the tensor-core path.
//
//   grad W_l [k][n]   = sum over the task's jet rows r of  h_{l-1}[r][k] * zbar_l[r][n]                 
//   (H v)_W_l [k][n]  = sum_r  hdot_{l-1}[r][k] * zbar_l[r][n] + h_{l-1}[r][k] * zbardot_l[r][n]
//
// tc_pde_kernel / tc_hvp_kernel spill the activation records of every layer and zbar_l (zbardot_l) in fp32; this
// kernel is the contraction over rows, and it is bound by reading those spills once from HBM.  One CTA per
// (task, layer, 128-wide k tile) walks the task's rows in steps of 16:
//   * a producer warp streams the raw fp32 spills of a step (16 rows of zbar: one contiguous bulk copy; the record
//     fields of the step's points: ONE contiguous bulk copy, records are laid out
//     [layer][k tile][step][field][point][128 neurons], rec_off in tc_params.cuh) into a 3/5-stage shared-memory ring (cp.async.bulk + mbarrier): the
//     memory latency is hidden by the ring, not by registers;
//   * two groups of 8 converter warps alternate steps: read the raw stage (conflict-free LDS), recompute h (hdot)
//     from the records, split both operands into tf32 hi/lo (round-to-nearest, x = hi + lo) and write them K-major
//     into the group's operand slot (slab row = neuron, 128 bytes = 16 rows of hi then 16 rows of lo; kind::tf32 with
//     MN-major swizzled operands returned zeros on B200, tools/tc_probe.cu);
//   * one thread issues the three kind::tf32 MMAs hi*hi + lo*hi + hi*lo per 8 rows.  tf32 keeps the fp32 exponent,
//     so rows of very different magnitude need no scaling.
//
// The tensor core's accumulator is trusted for 128 rows only (48 accumulations): its additions are not
// round-to-nearest and a sum over all rows of a task drifted by 3e-4 at 16k points (measured).  Every 8 steps the
// converter threads add the TMEM accumulator into an fp32 master copy in registers (two TMEM buffers alternate;
// the promotion lags the MMAs by two steps so the pipeline never drains).  Fixed summation order: deterministic.
#pragma once
#include "tc_kernels.cuh"

namespace mpde {
namespace tc {

// Round-to-nearest (ties away, like cvt.rna.tf32.f32) to the 10-bit tf32 mantissa with integer arithmetic: ptxas
// expands cvt.rna.tf32.f32 into 4 instructions (add, inf/nan test, select, mask: ncu r01e); the operands here are
// finite, so the add and the mask suffice.
TC_DEV float tf32_rn(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u); }
// lo = x - hi is exact; the tensor core reads only its upper 19 bits (truncation: 2^-21 |x|, far below the 2^-11 of
// a single tf32 term), so it is stored unrounded
TC_DEV void split_tf32(float x, float& hi, float& lo) { hi = tf32_rn(x); lo = x - hi; }

constexpr int WG_QR = 16;                    // rows per pipeline step
constexpr int WG_GROUP = 8;                  // steps per accumulator generation (128 rows)

// HV = true: H v products (tangent pairs need up to 9 record fields per point); false: gradients (5 fields).  The raw
// ring takes what is left of the 227 KB after the two operand slots: 3 stages of 34 KB, or 5 stages of 26 KB (the
// kernel is latency-bound on the spill reads: ncu r01f, 63 % of the raw_full waits miss with 3 stages).
template <int W, bool HV> struct WgGeo {
  static constexpr int NMT = W / 128;
  static constexpr int HBYTES = 128 * 128, ZBYTES = W * 128;  // K-major slabs [128 k] / [W n] x (16 r hi | 16 r lo) tf32
  static constexpr int SLOT = HBYTES + ZBYTES;
  static constexpr int RAWZ = WG_QR * W * 4;                  // 16 rows of zbar, fp32
  static constexpr int RAWH = (HV ? 9 : 5) * 4 * 512;         // fields x 4 points (or <= 2 fields x 16 points) x 128 neurons
  static constexpr int RAW = RAWZ + RAWH;
  static constexpr int NS = HV ? 3 : 5;                       // raw staging stages
  static constexpr int NGRP = 2, NCONV = 256 * NGRP, NT = NCONV + 64;   // + MMA warp + producer warp
  static constexpr int NBAR = 2 * NS + 8;                     // raw full/empty, operand full/empty, dfull, dfree
  static constexpr int SMEM = 2 * SLOT + NS * RAW + NBAR * 8 + 16;
  static_assert(SMEM <= 232448, "tc_wgrad_kernel: shared memory");
};

// 16-byte chunk `chunk` (4 consecutive rows; chunks 0..3 = hi, 4..7 = lo) of slab row `row`, swizzled
TC_DEV uint32_t wg_chunk_off(int row, int chunk) {
  return (uint32_t)(row >> 3) * 1024u + (uint32_t)(row & 7) * 128u + (((uint32_t)chunk ^ (uint32_t)(row & 7)) << 4);
}
TC_DEV void wg_store4(uint8_t* slab, int row, int chunk, const float (&x)[4]) {
  float hi[4], lo[4];
  MPDE_UNROLL
  for (int i = 0; i < 4; ++i) split_tf32(x[i], hi[i], lo[i]);
  *reinterpret_cast<float4*>(slab + wg_chunk_off(row, chunk)) = make_float4(hi[0], hi[1], hi[2], hi[3]);
  *reinterpret_cast<float4*>(slab + wg_chunk_off(row, chunk + 4)) = make_float4(lo[0], lo[1], lo[2], lo[3]);
}

// B operand: 16 rows x W of zbar (or zbardot) from the raw stage [row][n] -> slab row n, chunk = group of 4 rows.
template <int W>
TC_DEV void wg_convert_z(const float* zraw, uint8_t* zslab, int gtid) {
  constexpr int NIT = W * (WG_QR / 4) / 256;
  MPDE_UNROLL
  for (int j = 0; j < NIT; ++j) {
    const int idx = gtid + 256 * j, nn = idx % W, g = idx / W;
    float x[4];
    MPDE_UNROLL
    for (int i = 0; i < 4; ++i) x[i] = zraw[(4 * g + i) * W + nn];
    wg_store4(zslab, nn, g, x);
  }
}

// Record fields of one step in the raw stage: [slot][point][128 neurons]; which record fields go to which slot:
//   value sets: h needs S (field 0) -> slot 0; hdot needs Cc and zdot (fields 1, NF = 2) -> slots 0, 1
//   jet sets:   h needs fields 0..NF-1; hdot needs all NF + C fields
TC_DEV int wg_raw_fields(int jk, bool tan) {
  if (jk == JET_V) return tan ? 2 : 1;
  const int NF = jk == JET_G ? 4 : 5, C = jk == JET_G ? 3 : 4;
  return tan ? NF + C : NF;
}
TC_DEV int wg_raw_field_src(int jk, bool tan, int slot) {
  if (jk == JET_V) return tan ? (slot == 0 ? 1 : 2) : 0;
  return slot;
}

// A operand: h (TAN = false) or hdot (TAN = true) of 16 rows x 128 neurons, recomputed from the raw record fields;
// this thread: neuron k (slab row), chunks `half` and `half + 2`.
template <int JK, bool TAN>
TC_DEV void wg_convert_h(const float* hraw, float om, uint8_t* hslab, int k, int half) {
  typedef JetTraits<JK> JT;
  constexpr int NF = JT::NF, C = JT::C;
  if (JK == JET_V) {
    // 16 points = 16 rows; chunk g = points 4g..4g+3.  h = S; hdot = om * Cc * zdot
    MPDE_UNROLL
    for (int j = 0; j < 2; ++j) {
      const int g = half + 2 * j;
      float x[4];
      MPDE_UNROLL
      for (int i = 0; i < 4; ++i) {
        const float s0 = hraw[(4 * g + i) * 128 + k];
        x[i] = TAN ? om * s0 * hraw[(WG_QR + 4 * g + i) * 128 + k] : s0;
      }
      wg_store4(hslab, k, g, x);
    }
  } else {
    // 4 points; chunk pq = the 4 channel rows of point pq
    constexpr int NL = TAN ? NF + C : NF;
    MPDE_UNROLL
    for (int j = 0; j < 2; ++j) {
      const int pq = half + 2 * j;
      float f[NL];
      MPDE_UNROLL
      for (int q = 0; q < NL; ++q) f[q] = hraw[(q * 4 + pq) * 128 + k];
      float a[4] = {0.f, 0.f, 0.f, 0.f};
      if (TAN) jet_act_tan<JK>(om, f, f + NF, a);
      else jet_act_from_saved<JK>(om, f, a);
      wg_store4(hslab, k, pq, a);
    }
  }
}

// step st -> what it contracts
struct WgStep { int s, jk, tile, t, qq; bool tan, zd; };
// qpt (steps per tile: 4 or 8) and npair (1 or 2) are powers of two: shifts, no integer division in the step loop
TC_DEV WgStep wg_decode_step(const TcWgradParams& p, int st, int qpt_log2, int npair_log2) {
  WgStep w;
  const int pair = st & ((1 << npair_log2) - 1), q = st >> npair_log2;
  w.t = q >> qpt_log2; w.qq = q & ((1 << qpt_log2) - 1);
  int s = 0;
  while (s + 1 < p.n_seg && w.t >= p.seg_tile_begin[s + 1]) ++s;
  w.s = s; w.jk = p.seg_jk[s]; w.tile = w.t - p.seg_tile_begin[s];
  // pair 0 of an H v product: (hdot, zbar); pair 1: (h, zbardot); gradient: (h, zbar)
  w.tan = p.want_hv && pair == 0;
  w.zd = p.want_hv && pair == 1;
  return w;
}

template <int W, bool HV>
__global__ void __launch_bounds__(WgGeo<W, HV>::NT, 1) tc_wgrad_kernel(const __grid_constant__ TcWgradParams p) {
  typedef WgGeo<W, HV> G;
  constexpr int WG_NS = G::NS;
  uint8_t* smem = tc_smem_raw;
  uint8_t* raw = smem + 2 * G::SLOT;
  uint64_t* bars = reinterpret_cast<uint64_t*>(raw + WG_NS * G::RAW);
  uint64_t* raw_full = bars;                 // [WG_NS] producer -> converters (transaction bytes)
  uint64_t* raw_empty = bars + WG_NS;        // [WG_NS] converters (256) -> producer
  uint64_t* op_full = bars + 2 * WG_NS;      // [2] converters (256) -> MMA
  uint64_t* op_empty = op_full + 2;          // [2] MMA commit -> converters
  uint64_t* dfull = op_full + 4;             // [2] MMA commit -> promotion
  uint64_t* dfree = op_full + 6;             // [2] promotion (512) -> MMA
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + G::NBAR);
  const int tid = threadIdx.x, warp = warp_idx_uniform(), lane = tid & 31;
  const int task = blockIdx.y;
  const int l = 1 + blockIdx.x / G::NMT, kt = blockIdx.x % G::NMT;
  const int L = p.L;
  const int n_tiles = p.seg_tile_begin[p.n_seg];
  const int qpt = p.tile_rows / WG_QR;                 // 16-row steps per tile
  const int npair = p.want_hv ? 2 : 1;
  const int qpt_log2 = 31 - __clz(qpt), npair_log2 = npair - 1;
  // row split blockIdx.z: tiles [t_lo, t_hi) of the task; steps are counted from 0 inside the split
  const int t_lo = (int)((long long)n_tiles * blockIdx.z / gridDim.z), t_hi = (int)((long long)n_tiles * (blockIdx.z + 1) / gridDim.z);
  const int st0 = t_lo * qpt * npair;
  const int nsteps = (t_hi - t_lo) * qpt * npair;
  if (tid == 0) {
    if (smem_u32(smem) & 1023u) { printf("tc_wgrad_kernel: dynamic shared memory is not 1024-byte aligned\n"); __trap(); }
    for (int i = 0; i < WG_NS; ++i) { mbar_init(&raw_full[i], 1); mbar_init(&raw_empty[i], 256); }
    mbar_init(&op_full[0], 256); mbar_init(&op_full[1], 256);
    mbar_init(&op_empty[0], 1); mbar_init(&op_empty[1], 1);
    mbar_init(&dfull[0], 1); mbar_init(&dfull[1], 1);
    mbar_init(&dfree[0], G::NCONV); mbar_init(&dfree[1], G::NCONV);
    mbar_fence_init();
  }
  if (warp == 0) tmem_alloc(tmem_slot, 2 * W);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = *tmem_slot;

  if (warp < G::NCONV / 32) {
    // ===== converters: group g (8 warps) converts steps st = g, g + 2, ... into operand slot g =====
    const int grp = warp >> 3, gtid = tid & 255;
    const int k = gtid & 127, half = gtid >> 7;
    uint8_t* hslab = smem + (size_t)grp * G::SLOT;
    uint8_t* zslab = hslab + G::HBYTES;
    // fp32 master copy of this thread's part of the result: TMEM lane 32 * (warp & 3) + lane, W/4 columns
    constexpr int MC = W / 4;
    float master[MC];
    MPDE_UNROLL
    for (int j = 0; j < MC; ++j) master[j] = 0.f;
    auto promote = [&](int gen) {
      const int b = gen & 1;
      mbar_wait(&dfull[b], (gen >> 1) & 1);
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
      mbar_arrive(&dfree[b]);
    };
    int done_gen = 0;                                   // generations 0 .. done_gen-1 are in the master copy
    const float om = p.omega;
    int cseg = 0;
    for (int st = grp; st < nsteps; st += G::NGRP) {
      // segment look-up is monotonic in st: no full decode per step
      const int t_st = (st0 + st) >> (qpt_log2 + npair_log2);
      while (cseg + 1 < p.n_seg && t_st >= p.seg_tile_begin[cseg + 1]) ++cseg;
      WgStep w;
      w.jk = p.seg_jk[cseg];
      w.tan = HV && (((st0 + st) & 1) == 0);               // HV: steps alternate (hdot, zbar) / (h, zbardot)
      // promotion lags the MMAs by two steps (no pipeline drain at a generation boundary)
      while (st >= 2 && done_gen < (st - 2) / WG_GROUP) promote(done_gen++);
      const int stage = (int)((unsigned)st % (unsigned)WG_NS);
      const float* zraw = reinterpret_cast<const float*>(raw + (size_t)stage * G::RAW);
      const float* hraw = reinterpret_cast<const float*>(raw + (size_t)stage * G::RAW + G::RAWZ);
      mbar_wait(&raw_full[stage], (st / WG_NS) & 1);                      // the step's raw spills have landed
      if (st >= 2) mbar_wait(&op_empty[grp], ((st >> 1) - 1) & 1);        // the MMAs of step st - 2 have read the slot
      wg_convert_z<W>(zraw, zslab, gtid);
      if (!HV || !w.tan) {
        if (w.jk == JET_V) wg_convert_h<JET_V, false>(hraw, om, hslab, k, half);
        else if (w.jk == JET_G) wg_convert_h<JET_G, false>(hraw, om, hslab, k, half);
        else if (w.jk == JET_LAP) wg_convert_h<JET_LAP, false>(hraw, om, hslab, k, half);
        else wg_convert_h<JET_XX, false>(hraw, om, hslab, k, half);
      } else {
        if (w.jk == JET_V) wg_convert_h<JET_V, true>(hraw, om, hslab, k, half);
        else if (w.jk == JET_G) wg_convert_h<JET_G, true>(hraw, om, hslab, k, half);
        else if (w.jk == JET_LAP) wg_convert_h<JET_LAP, true>(hraw, om, hslab, k, half);
        else wg_convert_h<JET_XX, true>(hraw, om, hslab, k, half);
      }
      mbar_arrive(&raw_empty[stage]);                                     // raw stage consumed (values are in registers / stored)
      fence_proxy_async_smem();
      mbar_arrive(&op_full[grp]);
    }
    // ===== remaining generations, then the result: [128 k][W n] -> grad W_l =====
    while (nsteps > 0 && done_gen <= (nsteps - 1) / WG_GROUP) promote(done_gen++);
    float* out = (blockIdx.z == 0)
                     ? p.gout + (size_t)task * p.g_stride + nat_off_layer(W, l)
                     : p.part + (size_t)(blockIdx.z - 1) * p.part_stride + ((size_t)task * (L - 1) + (l - 1)) * W * W;
    out += (size_t)(kt * 128 + 32 * (warp & 3) + lane) * W + (warp >> 2) * MC;
    MPDE_UNROLL
    for (int j = 0; j < MC; ++j) out[j] = master[j];   // gout rows are only 4-byte aligned (P is odd-ish)
  } else if (warp == G::NCONV / 32) {
    // ===== MMA issuer: warp-uniform loop, one elected lane issues (see elect_one in tc_common.cuh) =====
    constexpr uint32_t idesc = make_idesc(FMT_TF32, 128, W, 0, 0);
    const uint32_t smem0 = smem_u32(smem);
    for (int st = 0; st < nsteps; ++st) {
      const int slot = st & 1;
      const int gen = st / WG_GROUP, b = gen & 1;     // accumulator generation and its TMEM buffer
      const uint32_t d = tbase + (uint32_t)(b * W);
      const uint32_t acc0 = (st % WG_GROUP) ? 1u : 0u;  // a fresh accumulator per generation
      if ((st % WG_GROUP) == 0 && gen >= 2) { mbar_wait(&dfree[b], ((gen >> 1) - 1) & 1); tc_fence_after(); }
      mbar_wait(&op_full[slot], (st >> 1) & 1);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t hs = desc_lo_kmajor(smem0 + (uint32_t)slot * (uint32_t)G::SLOT), zs = hs + ((uint32_t)G::HBYTES >> 4);
        MPDE_UNROLL
        for (int j = 0; j < WG_QR / 8; ++j) {           // 8 rows of r (32 bytes of K) per MMA; lo parts 64 bytes further
          const uint64_t ah = desc_join(hs + j * 2), al = desc_join(hs + 4 + j * 2);
          const uint64_t bh = desc_join(zs + j * 2), bl = desc_join(zs + 4 + j * 2);
          mma_tf32_ss(d, ah, bh, idesc, j ? 1u : acc0);
          mma_tf32_ss(d, al, bh, idesc, 1);
          mma_tf32_ss(d, ah, bl, idesc, 1);
        }
        tc_commit(&op_empty[slot]);
        if ((st % WG_GROUP) == WG_GROUP - 1 || st == nsteps - 1) tc_commit(&dfull[b]);
      }
      __syncwarp();
    }
  } else {
    // ===== producer: raw spills of step st -> stage st % WG_NS (lane 0: zbar rows, lane 1: the step's record fields).
    // Everything that only changes per tile is computed per tile and the per-step work is a handful of adds: round 2's
    // profile showed this ONE warp busy 100 % of the time (a full step decode, 64-bit multiplies and a modulo per step)
    // while the 16 converter warps waited for data 19 % of theirs -- the producer's instruction stream paced the kernel. =====
    const float* rec_t = p.rec + (size_t)task * p.rec_task_stride;
    const float* zp_t = p.zsp + (size_t)task * p.zsp_task_stride;
    const float* zd_t = p.want_hv ? p.zspd + (size_t)task * p.zsp_task_stride : zp_t;
    int seg = 0;
    uint32_t stage = 0, ephase = 1;                      // ring stage of the next step; parity to wait on raw_empty with
    int st = 0;
    for (int t = t_lo; t < t_hi; ++t) {
      while (seg + 1 < p.n_seg && t >= p.seg_tile_begin[seg + 1]) ++seg;
      const int jk = p.seg_jk[seg], tile = t - p.seg_tile_begin[seg];
      const int CP = jk == JET_V ? 1 : 4, PTS = p.tile_rows / CP, npts = WG_QR / CP;
      const int NF = jk == JET_V ? 2 : (jk == JET_G ? 4 : 5);
      const int NFT = NF + (p.hvp_records ? (jk == JET_V ? 1 : (jk == JET_G ? 3 : 4)) : 0);
      const uint32_t fbytes = (uint32_t)npts * 512u;
      // this tile's zbar rows of layer l and record block of layer l - 1, k tile kt
      const size_t zoff = ((size_t)t * (L - 1) + (l - 1)) * (size_t)p.tile_rows * W;
      const float* rec_blk = rec_t + p.rec_seg_base[seg] + ((size_t)tile * L + (l - 1)) * (size_t)(NFT * PTS) * W +
                             (size_t)kt * (size_t)(NFT * PTS * 128);
      const int step_floats = NFT * npts * 128;          // records of one 16-row step
      // per pair (0: (hdot, zbar) / gradient (h, zbar); 1: (h, zbardot)): fields and their first field
      const int nfl0 = wg_raw_fields(jk, p.want_hv != 0), f00 = wg_raw_field_src(jk, p.want_hv != 0, 0);
      const int nfl1 = wg_raw_fields(jk, false), f01 = wg_raw_field_src(jk, false, 0);
      for (int qq = 0; qq < qpt; ++qq) {
        for (int pair = 0; pair < npair; ++pair, ++st) {
          const int nfl = pair ? nfl1 : nfl0, f0 = pair ? f01 : f00;
          if (st >= WG_NS) mbar_wait(&raw_empty[stage], ephase);
          uint8_t* dst = raw + (size_t)stage * G::RAW;
          const uint32_t rbytes = (uint32_t)nfl * fbytes;
          if (lane == 0) {
            mbar_arrive_expect_tx(&raw_full[stage], (uint32_t)G::RAWZ + rbytes);
            bulk_g2s(dst, (pair ? zd_t : zp_t) + zoff + (size_t)qq * (WG_QR * W), (uint32_t)G::RAWZ, &raw_full[stage]);
            // the step's record fields f0 .. f0 + nfl - 1 of this k tile are one contiguous piece (rec_off in tc_params.cuh)
            bulk_g2s(dst + G::RAWZ, rec_blk + (size_t)qq * step_floats + (size_t)f0 * (npts * 128), rbytes, &raw_full[stage]);
          }
          if (++stage == (uint32_t)WG_NS) { stage = 0; if (st >= WG_NS - 1) ephase ^= 1u; }
        }
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
