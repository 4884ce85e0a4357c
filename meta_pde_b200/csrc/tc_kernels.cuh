// tc_kernels.cuh -- tensor-core (tcgen05 / TMEM) path of the fused SIREN jet kernels for layer widths 128 and 256.
//
// Same math as pde_kernels.cuh (reference: src/nets/field.py:47-72,146-199 and the three loss_fn, see there), but the
// per-layer [jet rows x W] x [W x W] contractions run on the 5th-generation tensor cores in split precision:
// every fp32 operand x is carried as fp16 (hi, lo) with hi + lo = x * 2^shift to 22 bits, and a product is the three
// kind::f16 MMAs hi*hi + lo*hi + hi*lo accumulated in fp32 in tensor memory (measured 7e-7 rel-L2 on B200,
// profiles/r01_tc_probe.jsonl).  Block scaling: one power-of-two shift per (tile, layer, jet channel) for the
// activations (from an exact-or-upper bound of the channel maximum, so nothing overflows fp16) and one per layer
// for the weights.
//
// Orientation ("transposed"): D[neuron][row] = sum_k Wt[neuron][k] * X[row][k].  The weights are the A operand
// (M = 128 neurons per MMA, streamed from L2 through a bulk-copy ring, pre-swizzled by tc_prep_kernel), the tile's
// jet rows are the MN-major B operand (N = 64 rows per tile, two tiles resident in shared memory), and a TMEM lane is a neuron: the epilogue
// thread of neuron n reads its accumulator row with tcgen05.ld and has all jet channels of every point of the tile
// in registers, so the sin-jet algebra is thread-local exactly as in the FP32-FMA kernel.
//
// Warp roles per CTA: W/32 epilogue warps (thread = neuron), one bulk-copy producer warp, one MMA-issuer warp.
#pragma once
#include "compat.cuh"
#include "jet_math.cuh"
#include "pde_kernels.cuh"
#include "tc_common.cuh"
#include "tc_params.cuh"

namespace mpde {
namespace tc {

constexpr int TR = 64;                  // jet rows per tile (MMA N of one tile)
constexpr int NSLOT = 2;                // epilogue warp groups per CTA; group g owns operand buffer X[g]
constexpr int NPAIR = 2;                // pairs of tiles in flight: one pair in the tensor core, the other in the epilogue
constexpr int STAGE_BYTES = 32768;      // one ring stage: [128 neurons][64 k] fp16 hi (16 KB) then lo (16 KB)
constexpr int NSTAGE = 3;

template <int W> struct TcGeo {
  static constexpr int NMT = W / 128;                 // M tiles (128 neurons each)
  static constexpr int KSL = W / 64;                  // 64-wide k slabs
  static constexpr int NEPI = W;                      // epilogue threads PER GROUP (thread = neuron); NSLOT groups of them
  // + one control warpgroup: producer, MMA issuer, record prefetcher, one idle warp.  W = 256: 16 epilogue warps + 4 control
  // warps launch at 96 registers; the control warpgroup then drops to 32 and the epilogue warps rise to 112 (setmaxnreg)
  static constexpr bool REGSHIFT = (W == 256);
  // the pool of setmaxnreg is what the CTA was launched with: 128 control threads x (96 - 32) = 512 epilogue threads x (112 - 96)
  static constexpr int EPI_REGS = 112, CTL_REGS = 32;
  static constexpr int NT = NSLOT * W + 128;
  static constexpr int SLAB = TR * 128;               // bytes of one [64 rows][64 k] fp16 slab
  static constexpr int XBYTES = 2 * KSL * SLAB;       // one group's operand buffer: hi slabs then lo slabs
  static constexpr int RING = NSTAGE * STAGE_BYTES;
  // per group (bytes): ybuf[TR][2] f32, red[TR] f32;  per (pair, group): cmax[2][4] u32
  static constexpr int GRP_MISC = 2 * TR * 4 + TR * 4;
  // shared: terms[NSLOT][4] f32, barriers (2*NSTAGE + 2*NPAIR) u64, tmem slot u32
  static constexpr int MISC = NSLOT * GRP_MISC + NPAIR * NSLOT * 32 + NSLOT * 16 + (2 * NSTAGE + 2 * NPAIR) * 8 + 16;
  static constexpr int SMEM = NSLOT * XBYTES + RING + MISC;
  // TMEM: pair P, M tile mt, group g -> columns P * PAIR_COLS + mt * (NSLOT * TR) + g * TR: one N = 128 MMA per (pair, M tile)
  static constexpr int PAIR_COLS = NMT * NSLOT * TR;
  static constexpr int TMEM_COLS = NPAIR * PAIR_COLS; // 256 or 512 (power of two)
  // global weight image per task: hidden layers 1..L-1 (forward copy: NMT*KSL stages each), then the output layer
  static constexpr size_t layer_bytes() { return (size_t)NMT * KSL * STAGE_BYTES; }
  static constexpr size_t out_bytes() { return (size_t)KSL * STAGE_BYTES; }
  static size_t image_bytes(int L, bool with_bwd) { return (size_t)(L - 1) * layer_bytes() * (with_bwd ? 2 : 1) + out_bytes(); }
};

// ---------------------------------------------------------------------------------------------------------
// tc_prep_kernel: natural theta (hidden kernels [n_in][n_out]) + theta-hat (output layer) -> swizzled fp16 hi/lo
// stages in streaming order.  grid (stages_per_task, T), 256 threads.  Each block recomputes its layer's max.
// ---------------------------------------------------------------------------------------------------------
struct TcPrepParams {
  int W, L, OUT, T;
  int with_bwd;                                 // also write the adjoint copies (layers L-1..1, A operand = W_l [k][n])
  const float* theta; long long theta_stride;   // natural parameters (or the HVP direction)
  const float* th;    long long th_stride;      // theta-hat (effective output layer wLe [W][OUT])
  uint8_t* wimg; long long wimg_stride;
  int* wshift; long long wshift_stride;
};

// shift s with max * 2^s in [2^13, 2^14); 0 for an all-zero block
TC_DEV int shift_for_max(float mx, int target = 14) {
  if (!(mx > 0.f)) return 0;
  const int e = (int)((__float_as_uint(mx) >> 23) & 0xffu) - 126;   // mx < 2^e
  int s = target - e;
  s = s < -100 ? -100 : (s > 100 ? 100 : s);
  return s;
}
TC_DEV float pow2i(int s) { return __uint_as_float((uint32_t)(s + 127) << 23); }

__global__ void __launch_bounds__(256) tc_prep_kernel(TcPrepParams p) {
  const int W = p.W, L = p.L, OUT = p.OUT;
  const int NMT = W / 128, KSL = W / 64;
  const int t = blockIdx.y;
  const int st = blockIdx.x;                          // stage index inside the task image
  const int per_layer = NMT * KSL;
  const int nst_fwd = (L - 1) * per_layer + KSL;
  const bool is_bwd = st >= nst_fwd;
  const bool is_out = !is_bwd && st >= (L - 1) * per_layer;
  // step index: hidden layer l, or L = output layer; adjoint copies run l = L-1 .. 1
  const int l = is_bwd ? (L - 1) - (st - nst_fwd) / per_layer : (is_out ? L : 1 + st / per_layer);
  const int rem = is_bwd ? (st - nst_fwd) % per_layer : (is_out ? st - (L - 1) * per_layer : st % per_layer);
  const int mt = is_out ? 0 : rem / KSL, s = is_out ? rem : rem % KSL;
  const float* src = is_out ? p.th + (size_t)t * p.th_stride + ph_off_out(W, L)
                            : p.theta + (size_t)t * p.theta_stride + nat_off_layer(W, l);
  const int count = is_out ? W * OUT : W * W;
  __shared__ float red[256];
  float mx = 0.f;
  for (int i = threadIdx.x; i < count; i += 256) mx = fmaxf(mx, fabsf(src[i]));
  red[threadIdx.x] = mx;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) red[threadIdx.x] = fmaxf(red[threadIdx.x], red[threadIdx.x + o]);
    __syncthreads();
  }
  const int sh = shift_for_max(red[0]);
  const float sc = pow2i(sh);
  if (rem == 0 && !is_bwd && threadIdx.x == 0) p.wshift[(size_t)t * p.wshift_stride + l] = sh;
  uint8_t* dst = p.wimg + (size_t)t * p.wimg_stride + (size_t)st * STAGE_BYTES;
  for (int i = threadIdx.x; i < 128 * 64; i += 256) {
    const int m = i >> 6, kk = i & 63;
    const int n = mt * 128 + m, k = s * 64 + kk;
    float w;
    if (is_out) w = (m < OUT) ? src[k * OUT + m] : 0.f;
    else if (is_bwd) w = src[(size_t)n * W + k];     // adjoint copy: row m = input neuron, K runs over output neurons
    else w = src[(size_t)k * W + n];
    __half h, lo;
    split_f16(w * sc, h, lo);
    const uint32_t off = sw128_off(m, kk * 2);
    *reinterpret_cast<__half*>(dst + off) = h;
    *reinterpret_cast<__half*>(dst + 16384 + off) = lo;
  }
}

// ---------------------------------------------------------------------------------------------------------
// shared-memory carve-up and per-CTA context
// ---------------------------------------------------------------------------------------------------------
extern __shared__ __align__(1024) unsigned char tc_smem_raw[];

template <int W> struct TcCtx {
  typedef TcGeo<W> G;
  uint8_t* xbuf0; uint8_t* ring; uint8_t* misc0;
  uint32_t* cmax0; float* terms;
  uint64_t* full; uint64_t* empty; uint64_t* xready; uint64_t* dfull;     // xready[NPAIR], dfull[NPAIR]
  uint32_t* tmem_slot;
  TC_DEV void carve() {
    uint8_t* s = tc_smem_raw;
    xbuf0 = s; ring = s + NSLOT * G::XBYTES;
    misc0 = ring + G::RING;
    cmax0 = reinterpret_cast<uint32_t*>(misc0 + NSLOT * G::GRP_MISC);
    terms = reinterpret_cast<float*>(cmax0 + NPAIR * NSLOT * 8);
    full = reinterpret_cast<uint64_t*>(terms + 4 * NSLOT);
    empty = full + NSTAGE; xready = empty + NSTAGE; dfull = xready + NPAIR;
    tmem_slot = reinterpret_cast<uint32_t*>(dfull + NPAIR);
  }
  // per-group / per-tile pieces (runtime indices: plain address arithmetic, no indexed register arrays)
  TC_DEV uint8_t* xbuf(int grp) const { return xbuf0 + (size_t)grp * G::XBYTES; }
  TC_DEV float* ybuf(int grp) const { return reinterpret_cast<float*>(misc0 + (size_t)grp * G::GRP_MISC); }
  TC_DEV float* red(int grp) const { return ybuf(grp) + 2 * TR; }
  TC_DEV uint32_t* cmax(int pair, int grp) const { return cmax0 + (pair * NSLOT + grp) * 8; }
};

// barrier among the epilogue warps of one group (named barrier 1 + group)
template <int NEPI> TC_DEV void epi_sync(int grp = 0) { asm volatile("bar.sync %0, %1;" ::"r"(1 + grp), "n"(NEPI) : "memory"); }

// warp max of non-negative floats (uint order == float order), folded into shared cmax[c]
TC_DEV void fold_max(uint32_t* slot, float v, int lane) {
  const uint32_t m = __reduce_max_sync(0xffffffffu, __float_as_uint(v));
  if (lane == 0 && m != 0u) atomicMax(slot, m);
}

// ---------------------------------------------------------------------------------------------------------
// Per-tile epilogue work of one epilogue thread (= neuron n).  HEAD selects jet kind / channel count.
// Row r of the tile: channel sets r = 4 * point + channel (32 points, 3-channel jets padded to 4 rows);
// value sets r = point (128 points).
// ---------------------------------------------------------------------------------------------------------
template <int W, int HEAD, int MODE>
struct TcTile {
  typedef TcGeo<W> G;
  typedef HeadTraits<HEAD> HT;
  static constexpr int JK = HT::JK, OUT = HT::OUT;
  typedef JetTraits<JK> JT;
  static constexpr int C = JT::C;
  static constexpr bool VAL = (JK == JET_V);
  static constexpr int CP = VAL ? 1 : 4;           // rows per point
  static constexpr int PTS = TR / CP;              // points per tile

  // The tile's jet rows are the MN-major B operand: blocks of 64 rows, each [k = neuron][128 B = 64 rows] with 8 neurons
  // per 1024-byte swizzle atom, so 8 consecutive rows of one neuron are one 16-byte chunk: the epilogue thread of neuron
  // n stores its 4 channels of a point (or 8 value rows) with one vector store for the hi and one for the lo parts
  // (the K-major layout needed two 2-byte stores per value: ncu r01h, 3.5 % of the warp instructions were STS).
  static constexpr uint32_t NBLK_BYTES = (uint32_t)W * 128u;             // one block of 64 rows
  static constexpr uint32_t LO_OFF = (uint32_t)(TR / 64) * NBLK_BYTES;   // lo parts behind the hi parts
  static TC_DEV uint32_t xoff(int r, int n) {
    return (uint32_t)(r >> 6) * NBLK_BYTES + (uint32_t)(n >> 3) * 1024u + (uint32_t)(n & 7) * 128u +
           (((((uint32_t)r & 63u) >> 3) ^ (uint32_t)(n & 7)) << 4) + (uint32_t)(r & 7) * 2u;
  }
  static TC_DEV void store_x(uint8_t* xbuf, int r, int n, float y) {
    __half h, l;
    split_f16(y, h, l);
    const uint32_t o = xoff(r, n);
    *reinterpret_cast<__half*>(xbuf + o) = h;
    *reinterpret_cast<__half*>(xbuf + LO_OFF + o) = l;
  }
  static TC_DEV void split2(float a, float b, uint32_t& hi, uint32_t& lo) {
    const __half2 h = __floats2half2_rn(a, b);
    const float2 f = __half22float2(h);
    const __half2 l = __floats2half2_rn(a - f.x, b - f.y);
    hi = *reinterpret_cast<const uint32_t*>(&h);
    lo = *reinterpret_cast<const uint32_t*>(&l);
  }
  // rows r0 .. r0+3 (r0 % 4 == 0) of neuron n
  static TC_DEV void store_x4(uint8_t* xbuf, int r0, int n, const float (&y)[4]) {
    uint2 hv, lv;
    split2(y[0], y[1], hv.x, lv.x);
    split2(y[2], y[3], hv.y, lv.y);
    const uint32_t o = xoff(r0, n);
    *reinterpret_cast<uint2*>(xbuf + o) = hv;
    *reinterpret_cast<uint2*>(xbuf + LO_OFF + o) = lv;
  }
  // rows r0 .. r0+7 (r0 % 8 == 0) of neuron n
  static TC_DEV void store_x8(uint8_t* xbuf, int r0, int n, const float (&y)[8]) {
    uint4 hv, lv;
    split2(y[0], y[1], hv.x, lv.x);
    split2(y[2], y[3], hv.y, lv.y);
    split2(y[4], y[5], hv.z, lv.z);
    split2(y[6], y[7], hv.w, lv.w);
    const uint32_t o = xoff(r0, n);
    *reinterpret_cast<uint4*>(xbuf + o) = hv;
    *reinterpret_cast<uint4*>(xbuf + LO_OFF + o) = lv;
  }

  // ---- staging in tensor memory (round 2) --------------------------------------------------------------------
  // The operand buffer X[g] of a group is shared by its tiles of the two pairs in flight: while the tensor core reads it for
  // pair P, the group already computes the next operand of pair P' -- and parks it, in fp32, in the accumulator columns of
  // P' it has just consumed (thread n owns lane n % 128 of them).  When pair P's contraction has completed, every thread
  // reads its parked values back, scales them by the agreed block shifts and writes the fp16 hi/lo operand into X[g].
  static TC_DEV uint32_t chunk_off(int r0, int n) {      // chunk of rows r0..r0+7 (r0 % 8 == 0) of neuron n, hi part
    return (uint32_t)(r0 >> 6) * NBLK_BYTES + (uint32_t)(n >> 3) * 1024u + (uint32_t)(n & 7) * 128u +
           (((((uint32_t)r0 & 63u) >> 3) ^ (uint32_t)(n & 7)) << 4);
  }
  // all TR parked rows of this thread's neuron -> fp16 hi / lo operand rows, scaled by ys[channel]
  static TC_DEV void convert_from_tmem(uint32_t tlane, uint8_t* xb, int n, const float* ys) {
    MPDE_NOUNROLL
    for (int ch = 0; ch < TR / 16; ++ch) {
      uint32_t v[16];
      tmem_ld16(tlane + ch * 16, v);
      tmem_ld_wait();
      MPDE_UNROLL
      for (int h = 0; h < 2; ++h) {
        float f[8];
        MPDE_UNROLL
        for (int i = 0; i < 8; ++i) f[i] = __uint_as_float(v[8 * h + i]) * ys[VAL ? 0 : (i & 3)];
        uint4 hv, lv;
        split2(f[0], f[1], hv.x, lv.x); split2(f[2], f[3], hv.y, lv.y);
        split2(f[4], f[5], hv.z, lv.z); split2(f[6], f[7], hv.w, lv.w);
        const uint32_t o = chunk_off(ch * 16 + 8 * h, n);
        *reinterpret_cast<uint4*>(xb + o) = hv;
        *reinterpret_cast<uint4*>(xb + LO_OFF + o) = lv;
      }
    }
  }
  // fold the channel maxima, barrier among the group's warps, read the shifts.  FWD: the value channel is sin, |a_0| <= 1.
  // `cmx` = this tile's cmax[2][4]; the buffer of agreement seq - 1 stays intact until the barrier of agreement seq, so the
  // next phase of the tile re-derives the shifts from it (prev_shifts) instead of carrying them in registers.
  template <bool FWD>
  static TC_DEV void agree_shifts(uint32_t* cmx, int grp, int& seq, const float (&mx)[4], float (&ys)[4], int tid, int lane) {
    uint32_t* cm = cmx + (seq & 1) * 4;
    MPDE_UNROLL
    for (int c = 0; c < 4; ++c) {
      const bool used = VAL ? (!FWD && c == 0) : (c < C && !(FWD && c == 0));
      if (used) fold_max(cm + c, mx[c], lane);
    }
    epi_sync<G::NEPI>(grp);
    MPDE_UNROLL
    for (int c = 0; c < 4; ++c) ys[c] = pow2i((FWD && c == 0) ? 14 : shift_for_max(__uint_as_float(cm[c])));
    if (tid == 0) {
      uint32_t* other = cmx + ((seq & 1) ^ 1) * 4;
      other[0] = other[1] = other[2] = other[3] = 0u;
    }
    ++seq;
  }
  template <bool FWD>
  static TC_DEV void prev_shifts(const uint32_t* cmx, int seq, int (&sh)[4]) {
    const uint32_t* cm = cmx + ((seq - 1) & 1) * 4;
    MPDE_UNROLL
    for (int c = 0; c < 4; ++c) sh[c] = (FWD && c == 0) ? 14 : shift_for_max(__uint_as_float(cm[c]));
  }
  // one point's outputs into the chunk being parked (`qs`: index of the point inside its 16-row chunk, compile-time)
  static TC_DEV void park_point(uint32_t (&o)[16], int qs, const float* a, float (&mx)[4]) {
    if (VAL) {
      o[qs] = __float_as_uint(a[0]);
      mx[0] = fmaxf(mx[0], fabsf(a[0]));
    } else {
      MPDE_UNROLL
      for (int c = 0; c < 4; ++c) o[4 * qs + c] = (c < C) ? __float_as_uint(a[c]) : 0u;
      MPDE_UNROLL
      for (int c = 0; c < C; ++c) mx[c] = fmaxf(mx[c], fabsf(a[c]));
    }
  }
  // the point's coordinates (L2-resident; every lane of the warp reads the same address)
  static TC_DEV float2 point_xy(const float* pts, int gp, int npts) {
    const bool ok = gp < npts;
    const float2 v = __ldg(reinterpret_cast<const float2*>(pts + 2 * (size_t)(ok ? gp : npts - 1)));   // read-only path: L1-cached
    return ok ? v : make_float2(0.f, 0.f);
  }
  // E2: wait until the other pair's contraction has released X[g], convert the parked rows, publish the operand
  static TC_DEV void publish(TcCtx<W>& cx, int grp, int pair, uint32_t tlane, int n, const float* ys, int wait_other,
                             uint32_t par_other) {
    tmem_st_wait();                                      // this thread's parked rows are in tensor memory
    if (wait_other) { mbar_wait(cx.dfull + (pair ^ 1), par_other); tc_fence_after(); }
    convert_from_tmem(tlane, cx.xbuf(grp), n, ys);
    fence_proxy_async_smem();
    tc_fence_before();
    mbar_arrive(cx.xready + pair);
  }

  // ---- one PHASE of one tile (the W epilogue threads of group `grp`) -----------------------------------------
  // Phases of a tile (L layers); phase ph waits for contraction ph of its pair (ph >= 1) and, unless it is the last, ends by
  // publishing the operand of contraction ph + 1:
  //   0          points, layer 0 (direct)
  //   l = 1..L-1 accumulators of hidden layer l -> activation jets
  //   L          output layer -> head (loss terms, seeds); GRAD: reverse layer L-1 (seeds, no contraction)
  //   L + j      GRAD: adjoint accumulators -> reverse layer L-1-j (j = 1..L-1); layer 0 publishes nothing
  // par_self: parity of the pair's dfull barrier for contraction ph; wait_other / par_other: see publish().
  static TC_DEV void phase(const TcParams& p, TcCtx<W>& cx, const SegDesc& sg, int seg_id, int task, int tile, int grp,
                           int pair, int ph, uint32_t tbase, int& seq, uint32_t par_self, int wait_other, uint32_t par_other,
                           const int* wshift) {
    const KParams& kp = p.k;
    const int tid = (int)threadIdx.x - grp * G::NEPI, lane = tid & 31;   // indices inside the group
    const int warp = warp_idx_uniform() - grp * (G::NEPI / 32);
    const int n = tid;                          // this thread's neuron
    const int mt = n >> 7;
    const float om = kp.omega;
    const int L = kp.L;
    const float* th = kp.th + (size_t)task * kp.th_stride;
    const float* prm = kp.prm + (size_t)task * kp.nprm;
    const int pt0 = tile * PTS;
    const float* pts = sg.pts + (size_t)task * sg.task_stride;
    // a warp may only touch the TMEM lane quarter of its hardware warp id % 4; groups start at multiples of 4 warps
    const uint32_t tlane = tmem_addr(tbase, 32u * (uint32_t)(warp & 3),
                                     (uint32_t)pair * G::PAIR_COLS + (uint32_t)mt * (NSLOT * TR) + (uint32_t)grp * TR);
    float* ybuf = cx.ybuf(grp);
    float* red = cx.red(grp);
    uint32_t* cmx = cx.cmax(pair, grp);
    uint64_t* dfull = cx.dfull + pair;
    constexpr int NF = JT::NF;
    constexpr int TCH = 16;                             // accumulator columns (rows) per chunk
    constexpr int PPC = TCH / CP;                       // points per chunk
    // activation records of this tile: [layer][field][point][neuron]
    float* recp = nullptr;
    if (MODE != MODE_LOSS)
      recp = p.rec + (size_t)task * p.rec_task_stride + p.rec_seg_base[seg_id] + (size_t)tile * L * (NF * PTS) * W;   // see rec_off

    if (ph == 0) {
      // ---- layer 0 (in_dim = 2): direct ----
      float ys[4];
      const float w00 = mpde_ldg(th + n), w01 = mpde_ldg(th + W + n), b0 = mpde_ldg(th + 2 * W + n);
      float mx[4] = {0.f, 0.f, 0.f, 0.f};
      MPDE_NOUNROLL
      for (int ch = 0; ch < TR / TCH; ++ch) {
        uint32_t o[16];
        MPDE_UNROLL
        for (int q = 0; q < PPC; ++q) {
          const int pl = ch * PPC + q;
          const float2 xy = point_xy(pts, pt0 + pl, sg.n);
          float z[4], a[4], sv[5];
          z[0] = fmaf(w00, xy.x, fmaf(w01, xy.y, b0));
          z[1] = w00; z[2] = w01; z[3] = 0.f;
          jet_act_fwd<JK>(om, z, a, sv);
          if (MODE != MODE_LOSS) {
            MPDE_UNROLL
            for (int f = 0; f < NF; ++f) recp[rec_off(NF, PTS, 16 / CP, f, pl, n)] = sv[f];
          }
          park_point(o, q, a, mx);
        }
        tmem_st16(tlane + ch * TCH, o);
      }
      agree_shifts<true>(cmx, grp, seq, mx, ys, tid, lane);
      publish(cx, grp, pair, tlane, n, ys, wait_other, par_other);
      return;
    }

    if (ph < L) {
      // ---- hidden layer l = ph: accumulators from TMEM -> activation, parked in the same columns ----
      const int l = ph;
      int sh[4];
      prev_shifts<true>(cmx, seq, sh);
      const float bias = mpde_ldg(th + ph_off_layer(W, l) + 2 * W * W + n);
      float zs[4], ys[4];
      { const int sw = wshift[l]; for (int c = 0; c < 4; ++c) zs[c] = pow2i(-sh[c] - sw); }
      mbar_wait(dfull, par_self);                       // every MMA of this contraction has completed
      tc_fence_after();
      float mx[4] = {0.f, 0.f, 0.f, 0.f};
      MPDE_NOUNROLL
      for (int ch = 0; ch < TR / TCH; ++ch) {
        uint32_t v[TCH], o[16];
        tmem_ld16(tlane + ch * TCH, v);
        tmem_ld_wait();
        MPDE_UNROLL
        for (int q = 0; q < PPC; ++q) {
          const int pl = ch * PPC + q;
          float z[4] = {0.f, 0.f, 0.f, 0.f}, a[4], sv[5];
          MPDE_UNROLL
          for (int c = 0; c < C; ++c) z[c] = __uint_as_float(v[CP * q + c]) * zs[c];
          z[0] += bias;
          jet_act_fwd<JK>(om, z, a, sv);
          if (MODE != MODE_LOSS) {
            MPDE_UNROLL
            for (int f = 0; f < NF; ++f) recp[(size_t)l * NF * (PTS * W) + rec_off(NF, PTS, 16 / CP, f, pl, n)] = sv[f];
          }
          park_point(o, q, a, mx);
        }
        tmem_st16(tlane + ch * TCH, o);
      }
      agree_shifts<true>(cmx, grp, seq, mx, ys, tid, lane);
      publish(cx, grp, pair, tlane, n, ys, wait_other, par_other);
      return;
    }

    if (ph == L) {
      // ---- output layer: D0 lanes 0..OUT-1 hold y[o][row] ----
      int sh[4];
      prev_shifts<true>(cmx, seq, sh);
      mbar_wait(dfull, par_self);
      tc_fence_after();
      if (warp == 0) {
        float zs[4];
        { const int sw = wshift[L]; for (int c = 0; c < 4; ++c) zs[c] = pow2i(-sh[c] - sw); }
        MPDE_NOUNROLL
        for (int ch = 0; ch < TR / TCH; ++ch) {
          uint32_t v[TCH];
          tmem_ld16(tlane + ch * TCH, v);
          tmem_ld_wait();
          if (lane < OUT) {
            MPDE_UNROLL
            for (int q = 0; q < TCH; ++q) ybuf[(ch * TCH + q) * 2 + lane] = __uint_as_float(v[q]) * zs[VAL ? 0 : (q & 3)];
          }
        }
      }
      tc_fence_before();
      epi_sync<G::NEPI>(grp);

      // ---- head: one thread per point ----
      if (tid < PTS) {
        const int pl = tid;
        const float* bL = th + ph_off_out(W, L) + W * OUT;
        float u[OUT * C], ub[OUT * C];
        MPDE_UNROLL
        for (int o = 0; o < OUT; ++o)
          MPDE_UNROLL
          for (int c = 0; c < C; ++c) {
            float y = ybuf[(CP * pl + c) * 2 + o];
            if (c == 0) y += mpde_ldg(bL + o);
            u[o * C + c] = y;
          }
        const int gp = pt0 + pl;
        const bool valid = gp < sg.n;
        const float2 xy = point_xy(pts, gp, sg.n);
        float res = 0.f;
        float val = head_eval<HEAD>(u, xy.x, xy.y, prm, sg.coef, ub, &res);
        if (!valid) val = 0.f;
        red[pl] = val;
        if (MODE != MODE_LOSS) {
          // ybuf now carries the seeds ybar[row][o] of the reverse pass (zero for padding points / channels)
          MPDE_UNROLL
          for (int o = 0; o < OUT; ++o)
            MPDE_UNROLL
            for (int c = 0; c < CP; ++c) ybuf[(CP * pl + c) * 2 + o] = (valid && c < C) ? ub[o * C + c] : 0.f;
        }
        if (MODE == MODE_LOSS && valid && sg.res_out != nullptr) {
          float* ro = sg.res_out + (size_t)task * sg.res_stride;
          if (HEAD == HEAD_EVAL || HEAD == HEAD_EVAL2) {
            MPDE_UNROLL
            for (int o = 0; o < OUT; ++o) ro[(size_t)gp * OUT + o] = u[o * C];
          } else {
            ro[gp] = res;
          }
        }
      }
      epi_sync<G::NEPI>(grp);
      if (warp == 0) {
        float s = 0.f;
        MPDE_UNROLL
        for (int j = 0; j < (PTS + 31) / 32; ++j) s += (lane + 32 * j < PTS) ? red[lane + 32 * j] : 0.f;
        MPDE_UNROLL
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) cx.terms[grp * 4 + seg_id] += s;
      }
      if (MODE == MODE_LOSS) { epi_sync<G::NEPI>(grp); return; }   // ybuf / red are rewritten by the group's next phase
    }

    // =====================================================================================================
    // reverse pass: thread n owns neuron n of every layer.  abar_l comes from the seeds (l = L-1) or
    // from the adjoint contraction D[k][row] = sum_n W_{l+1}[k][n] zbar_{l+1}[row][n] in TMEM; zbar_l goes (a) in fp32
    // to the spill read by tc_wgrad_kernel and (b) parked, then as fp16 hi/lo, into X: the B operand of the next
    // adjoint contraction.  One layer per phase: l = L-1 in phase L (right after the head), l = 2L-1-ph after it.
    // =====================================================================================================
    if (MODE == MODE_LOSS) return;
    const int l = 2 * L - 1 - ph;
    // the layer kind is a template argument: no runtime branch inside the per-point loops
    if (ph == L) reverse_layer<RK_LAST>(p, cx, sg, seg_id, task, tile, grp, pair, l, tlane, recp, seq, par_self, wait_other, par_other, wshift);
    else if (l == 0) reverse_layer<RK_FIRST>(p, cx, sg, seg_id, task, tile, grp, pair, l, tlane, recp, seq, par_self, wait_other, par_other, wshift);
    else reverse_layer<RK_MID>(p, cx, sg, seg_id, task, tile, grp, pair, l, tlane, recp, seq, par_self, wait_other, par_other, wshift);
  }

  enum { RK_LAST = 0, RK_MID = 1, RK_FIRST = 2 };   // layer L-1 (seeds, no accumulators) / 0 < l < L-1 / layer 0 (no X to build)
  template <int RK>
  static TC_DEV void reverse_layer(const TcParams& p, TcCtx<W>& cx, const SegDesc& sg, int seg_id, int task, int tile, int grp,
                                   int pair, int l, uint32_t tlane, const float* recp, int& seq, uint32_t par_self,
                                   int wait_other, uint32_t par_other, const int* wshift) {
    constexpr bool LAST = (RK == RK_LAST), FIRST = (RK == RK_FIRST);
    const KParams& kp = p.k;
    const int tid = (int)threadIdx.x - grp * G::NEPI, lane = tid & 31;
    const int n = tid;
    const float om = kp.omega;
    const int L = kp.L;
    const float* th = kp.th + (size_t)task * kp.th_stride;
    const float* pts = sg.pts + (size_t)task * sg.task_stride;
    const int pt0 = tile * PTS;
    float* ybuf = cx.ybuf(grp);
    uint32_t* cmx = cx.cmax(pair, grp);
    constexpr int NF = JT::NF;
    constexpr int TCH = 16, PPC = TCH / CP, SB = PPC;
    float* ps = p.psmall + (((size_t)task * gridDim.x + blockIdx.x) * NSLOT + grp) * p.PS;   // per (CTA, group) partial
    float wl[OUT];
    MPDE_UNROLL
    for (int o = 0; o < OUT; ++o) wl[o] = 0.f;
    if (LAST) {
      MPDE_UNROLL
      for (int o = 0; o < OUT; ++o) wl[o] = mpde_ldg(th + ph_off_out(W, L) + n * OUT + o);
      if (tid < OUT) {   // output bias: sum of the value-channel seeds
        float g = 0.f;
        for (int pl = 0; pl < PTS; ++pl) g += ybuf[(CP * pl) * 2 + tid];
        ps[ps_off_wl(W, L) + W * OUT + tid] += g;
      }
    }
    const float* rl = recp + (size_t)l * NF * (PTS * W);       // this layer's record block (layout: rec_off)
    float zs[4] = {0.f, 0.f, 0.f, 0.f};
    // the first batch of record loads is issued before the accumulator wait (it does not depend on it)
    float svb[SB][NF];
    MPDE_UNROLL
    for (int qq = 0; qq < SB; ++qq)
      MPDE_UNROLL
      for (int f = 0; f < NF; ++f) svb[qq][f] = rl[rec_off(NF, PTS, 16 / CP, f, qq, n)];
    if (!LAST) {
      int shz[4];
      prev_shifts<false>(cmx, seq, shz);
      const int sw = wshift[l + 1];
      for (int c = 0; c < 4; ++c) zs[c] = pow2i(-shz[c] - sw);
      mbar_wait(cx.dfull + pair, par_self);           // all adjoint MMAs of this contraction have completed
      tc_fence_after();
    }
    float mx[4] = {0.f, 0.f, 0.f, 0.f};
    float bsum = 0.f, g0 = 0.f, g1 = 0.f, gw[OUT];
    MPDE_UNROLL
    for (int o = 0; o < OUT; ++o) gw[o] = 0.f;
    // zbar_l spill for tc_wgrad_kernel (layers >= 1)
    float* zl = p.zsp + (size_t)task * p.zsp_task_stride +
                ((size_t)(kp.seg_tile_begin[seg_id] + tile) * (L - 1) + (FIRST ? 0 : l - 1)) * TR * W + n;
    MPDE_NOUNROLL
    for (int ch = 0; ch < TR / TCH; ++ch) {
      uint32_t v[TCH], o[16];
      if (!LAST) { tmem_ld16(tlane + ch * TCH, v); tmem_ld_wait(); }
      // next chunk's records (clamped on the last chunk: a repeated, valid load instead of a branch)
      const int chn = (ch + 1 < TR / TCH) ? ch + 1 : ch;
      MPDE_UNROLL
      for (int q = 0; q < PPC; ++q) {
        const int pl = ch * PPC + q;
        float ab[4] = {0.f, 0.f, 0.f, 0.f}, sv[5], zb[4] = {0.f, 0.f, 0.f, 0.f};
        MPDE_UNROLL
        for (int c = 0; c < C; ++c) {
          if (!LAST) {
            ab[c] = __uint_as_float(v[CP * q + c]) * zs[c];
          } else {
            float a_ = 0.f;
            MPDE_UNROLL
            for (int oo = 0; oo < OUT; ++oo) a_ = fmaf(wl[oo], ybuf[(CP * pl + c) * 2 + oo], a_);
            ab[c] = a_;
          }
        }
        MPDE_UNROLL
        for (int f = 0; f < NF; ++f) sv[f] = svb[q][f];
        // this slot's record of the NEXT chunk: in flight while the chunk is processed
        MPDE_UNROLL
        for (int f = 0; f < NF; ++f) svb[q][f] = rl[rec_off(NF, PTS, 16 / CP, f, chn * PPC + q, n)];
        jet_act_bwd<JK>(om, sv, ab, zb);
        bsum += zb[0];
        if (LAST) {
          float a[4];
          jet_act_from_saved<JK>(om, sv, a);
          MPDE_UNROLL
          for (int oo = 0; oo < OUT; ++oo)
            MPDE_UNROLL
            for (int c = 0; c < C; ++c) gw[oo] = fmaf(a[c], ybuf[(CP * pl + c) * 2 + oo], gw[oo]);
        }
        if (FIRST) {
          const float2 xy = point_xy(pts, pt0 + pl, sg.n);
          if (VAL) { g0 = fmaf(xy.x, zb[0], g0); g1 = fmaf(xy.y, zb[0], g1); }
          else { g0 += fmaf(xy.x, zb[0], zb[1]); g1 += fmaf(xy.y, zb[0], zb[2]); }
        } else {
          MPDE_UNROLL
          for (int c = 0; c < CP; ++c) zl[(size_t)(CP * pl + c) * W] = c < C ? zb[c] : 0.f;   // fp32 spill for tc_wgrad_kernel
          park_point(o, q, zb, mx);
        }
      }
      if (!FIRST) tmem_st16(tlane + ch * TCH, o);
    }
    // small-parameter gradients: owner-computes read-modify-write into this (CTA, group) partial
    ps[ps_off_bias(W, l) + n] += bsum;
    if (LAST) {
      MPDE_UNROLL
      for (int oo = 0; oo < OUT; ++oo) ps[ps_off_wl(W, L) + n * OUT + oo] += gw[oo];
    }
    if (FIRST) { ps[n] += g0; ps[W + n] += g1; epi_sync<G::NEPI>(grp); return; }   // ybuf is rewritten by the group's next phase
    float ysz[4];
    agree_shifts<false>(cmx, grp, seq, mx, ysz, tid, lane);
    publish(cx, grp, pair, tlane, n, ysz, wait_other, par_other);
  }

  // a tile that does not exist (tail of the CTA's tile list): keep the pair's barriers in step -- one arrival per
  // contraction, each after the previous contraction of the pair has completed
  static TC_DEV void dummy_phase(TcCtx<W>& cx, int pair, int ph, int nph, uint32_t par_self) {
    if (ph >= 1) mbar_wait(cx.dfull + pair, par_self);
    if (ph < nph - 1) mbar_arrive(cx.xready + pair);
  }
};

// ---------------------------------------------------------------------------------------------------------
// kernel: grid = (ctas_per_task, T); each CTA walks the task's tiles t_i = blockIdx.x + i * gridDim.x, FOUR at a time:
// pair A = (t_{4q}, t_{4q+1}), pair B = (t_{4q+2}, t_{4q+3}); group g of W epilogue threads (thread = neuron) owns tile g
// of both pairs and the operand buffer X[g].  A pair is one N = 128 MMA operand (X[0] | X[1]), so every weight stage is
// read once per 128 rows.  Round 1 ran one 128-row tile strictly serially (contraction, then epilogue, ...: tensor pipe
// 19-30 % active).  Now the tensor core runs pair A's contraction while both groups are in the epilogue of pair B: the
// epilogue parks its fp32 outputs in the accumulator columns it has just read (tensor memory, 4 accumulator sets = 512
// columns at W = 256) and converts them into X[g] as soon as the other pair's contraction has released the buffer.
// Tiles that do not exist at the tail are "dummy" (barrier arrivals only).
// ---------------------------------------------------------------------------------------------------------
template <int W> TC_DEV int tc_step_first_stage(int step, int L) {
  constexpr int PER = TcGeo<W>::NMT * TcGeo<W>::KSL;
  // hidden layer `step` (1..L-1), output layer (step L), adjoint copies of layers L-1..1 (steps L+1..2L-1)
  return step < L ? (step - 1) * PER : (step == L ? (L - 1) * PER : (L - 1) * PER + TcGeo<W>::KSL + (step - L - 1) * PER);
}

template <int W, int PDE, int MODE>
TC_DEV void tc_dispatch_phase(const TcParams& p, TcCtx<W>& cx, int t, int task, int grp, int pair, int ph, uint32_t tbase,
                              int& seq, uint32_t par_self, int wait_other, uint32_t par_other, const int* wshift) {
  const KParams& kp = p.k;
  int s = 0;
  while (s + 1 < kp.n_seg && t >= kp.seg_tile_begin[s + 1]) ++s;
  const int tile = t - kp.seg_tile_begin[s];
  const SegDesc& sg = kp.seg[s];
#define TC_PHASE(H, M) TcTile<W, H, M>::phase(p, cx, sg, s, task, tile, grp, pair, ph, tbase, seq, par_self, wait_other, par_other, wshift)
  if (PDE == PDE_POISSON) {
    if (sg.head == HEAD_POISSON_DOMAIN) TC_PHASE(HEAD_POISSON_DOMAIN, MODE); else TC_PHASE(HEAD_POISSON_BOUNDARY, MODE);
  } else if (PDE == PDE_BURGERS) {
    if (sg.head == HEAD_BURGERS_DOMAIN) TC_PHASE(HEAD_BURGERS_DOMAIN, MODE); else TC_PHASE(HEAD_BURGERS_IC, MODE);
  } else if (PDE == PDE_HYPER) {
    if (sg.head == HEAD_HYPER_DOMAIN) TC_PHASE(HEAD_HYPER_DOMAIN, MODE);
    else if (sg.head == HEAD_HYPER_TOP) TC_PHASE(HEAD_HYPER_TOP, MODE);
    else TC_PHASE(HEAD_HYPER_BOTTOM, MODE);
  } else if (PDE == PDE_EVAL1) {
    TC_PHASE(HEAD_EVAL, MODE_LOSS);
  } else {
    TC_PHASE(HEAD_EVAL2, MODE_LOSS);
  }
#undef TC_PHASE
}

template <int W, int PDE, int MODE>
__global__ void __launch_bounds__(TcGeo<W>::NT, 1) tc_pde_kernel(const __grid_constant__ TcParams p) {
  typedef TcGeo<W> G;
  TcCtx<W> cx;
  cx.carve();
  const KParams& kp = p.k;
  const int tid = threadIdx.x, warp = warp_idx_uniform(), lane = tid & 31;
  const int task = blockIdx.y;
  const int cta = blockIdx.y * gridDim.x + blockIdx.x;
  const int L = kp.L;
  const int n_tiles = kp.seg_tile_begin[kp.n_seg];
  const int gx = (int)gridDim.x, bx = (int)blockIdx.x;

  if (tid == 0) {
    if (smem_u32(tc_smem_raw) & 1023u) { printf("tc_pde_kernel: dynamic shared memory is not 1024-byte aligned\n"); __trap(); }
    for (int i = 0; i < NSTAGE; ++i) { mbar_init(&cx.full[i], 1); mbar_init(&cx.empty[i], 1); }
    for (int i = 0; i < NPAIR; ++i) { mbar_init(&cx.xready[i], NSLOT * G::NEPI); mbar_init(&cx.dfull[i], 1); }
    mbar_fence_init();
    for (int i = 0; i < NPAIR * NSLOT * 8; ++i) cx.cmax0[i] = 0u;
    for (int i = 0; i < NSLOT * 4; ++i) cx.terms[i] = 0.f;
  }
  if (warp == 0) tmem_alloc(cx.tmem_slot, G::TMEM_COLS);
  if (MODE != MODE_LOSS) {
    float* ps = p.psmall + (size_t)cta * NSLOT * p.PS;
    for (int i = tid; i < NSLOT * p.PS; i += G::NT) ps[i] = 0.f;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = *cx.tmem_slot;
  const int nsteps = (MODE != MODE_LOSS) ? 2 * L - 1 : L;    // contractions per tile
  const int nph = nsteps + 1;                                // epilogue phases per tile
  // tile-list positions of a quad: i0 .. i0 + 3; pair B exists if position i0 + 2 does
  const int n_mine = (n_tiles > bx) ? (n_tiles - bx + gx - 1) / gx : 0;   // tiles of this CTA

  constexpr int NEW = NSLOT * G::NEPI / 32;                  // epilogue warps (both groups)
  if (G::REGSHIFT && MODE != MODE_LOSS) {
    if (warp < NEW) setmaxnreg_inc<G::EPI_REGS>(); else setmaxnreg_dec<G::CTL_REGS>();
  }
  if (warp < NEW) {
    // ===== epilogue warps: group g = tile g of pair A, then tile g of pair B, phase by phase =====
    const int grp = warp / (G::NEPI / 32);
    int seqA = 0, seqB = 0;
    uint32_t cA = 0, cB = 0;                                  // contractions of pair A / B completed before this quad
    const int* wshift = p.wshift + (size_t)task * p.wshift_stride;
    for (int i0 = 0; i0 < n_mine; i0 += NPAIR * NSLOT) {
      const bool pairB = i0 + NSLOT < n_mine;
      const bool hasA = i0 + grp < n_mine, hasB = i0 + NSLOT + grp < n_mine;
      const int tA = bx + (i0 + grp) * gx, tB = bx + (i0 + NSLOT + grp) * gx;
      MPDE_NOUNROLL
      for (int ph = 0; ph < nph; ++ph) {
        // ONE call site for both pairs (pair is a runtime value): a second inlined copy of the phase code thrashed the
        // instruction cache (21.8 k instructions, 8 % no-instruction stalls)
        MPDE_NOUNROLL
        for (int pr = 0; pr < (pairB ? NPAIR : 1); ++pr) {
          // pair A: its contraction ph (1-based) is the (cA + ph - 1)-th completion of dfull[0]; before publishing the operand
          // of contraction ph + 1 it waits for pair B's contraction ph (the last reader of X[g]).  Pair B: before publishing
          // it waits for pair A's contraction ph + 1.
          const uint32_t cs = pr ? cB : cA, co = pr ? cA : cB;
          const uint32_t par_self = (cs + (uint32_t)ph - 1u) & 1u;
          const uint32_t par_other = pr ? ((co + (uint32_t)ph) & 1u) : ((co + (uint32_t)ph - 1u) & 1u);
          const int wait_other = pr ? 1 : ((ph >= 1 && pairB) ? 1 : 0);
          int seq = pr ? seqB : seqA;
          if (pr ? hasB : hasA)
            tc_dispatch_phase<W, PDE, MODE>(p, cx, pr ? tB : tA, task, grp, pr, ph, tbase, seq, par_self, wait_other, par_other, wshift);
          else
            TcTile<W, HEAD_EVAL, MODE_LOSS>::dummy_phase(cx, pr, ph, nph, par_self);
          if (pr) seqB = seq; else seqA = seq;
        }
      }
      cA += (uint32_t)nsteps;
      if (pairB) cB += (uint32_t)nsteps;
    }
  } else if (warp == NEW) {
    // ===== producer: per quad and contraction step, the step's stages once per pair =====
    if (lane == 0) {
      const uint8_t* img = p.wimg + (size_t)task * p.wimg_stride;
      uint32_t stage = 0, phase = 0;
      for (int i0 = 0; i0 < n_mine; i0 += NPAIR * NSLOT) {
        const int npair = (i0 + NSLOT < n_mine) ? 2 : 1;
        for (int step = 1; step <= nsteps; ++step) {
          const int first = tc_step_first_stage<W>(step, L);
          const int cnt = (step != L) ? G::NMT * G::KSL : G::KSL;
          for (int pr = 0; pr < npair; ++pr) {
            for (int i = 0; i < cnt; ++i) {
              mbar_wait(&cx.empty[stage], phase ^ 1u);
              mbar_arrive_expect_tx(&cx.full[stage], STAGE_BYTES);
              bulk_g2s(cx.ring + (size_t)stage * STAGE_BYTES, img + (size_t)(first + i) * STAGE_BYTES, STAGE_BYTES, &cx.full[stage]);
              if (++stage == NSTAGE) { stage = 0; phase ^= 1u; }
            }
          }
        }
      }
    }
  } else if (warp == NEW + 1) {
    // ===== MMA issuer: the whole warp runs the loop (warp-uniform), one elected lane issues (see elect_one) =====
    // A (weights) K-major, B (jet rows) MN-major: X[0] | X[1] are the two 64-row blocks of ONE N = 128 operand (block
    // stride XBYTES); a dummy tile's block holds stale rows, its accumulator columns are never read
    constexpr uint32_t idesc = make_idesc(FMT_F16, 128, NSLOT * TR, 0, 1);
    constexpr uint32_t NBLK_BYTES = (uint32_t)W * 128u, LO_OFF = (uint32_t)(TR / 64) * NBLK_BYTES;
    const uint32_t xb = smem_u32(cx.xbuf0), ring0 = smem_u32(cx.ring);
    const uint32_t bh_0 = desc_lo_mnmajor(xb, (uint32_t)G::XBYTES), bl_0 = desc_lo_mnmajor(xb + LO_OFF, (uint32_t)G::XBYTES);
    uint32_t stage = 0, phase = 0, xpar0 = 0, xpar1 = 0;
    for (int i0 = 0; i0 < n_mine; i0 += NPAIR * NSLOT) {
      const int npair = (i0 + NSLOT < n_mine) ? 2 : 1;
      for (int step = 1; step <= nsteps; ++step) {
        const int nmt = (step != L) ? G::NMT : 1;   // step L is the output layer (one M tile)
        for (int pr = 0; pr < npair; ++pr) {
          if (pr) { mbar_wait(&cx.xready[1], xpar1); xpar1 ^= 1u; } else { mbar_wait(&cx.xready[0], xpar0); xpar0 ^= 1u; }
          tc_fence_after();
          for (int mt = 0; mt < nmt; ++mt) {
            const uint32_t d = tbase + (uint32_t)pr * (uint32_t)G::PAIR_COLS + (uint32_t)mt * (NSLOT * TR);
            for (int s = 0; s < G::KSL; ++s) {
              mbar_wait(&cx.full[stage], phase);
              tc_fence_after();
              if (elect_one()) {
                const uint32_t ah_0 = desc_lo_kmajor(ring0 + stage * (uint32_t)STAGE_BYTES), al_0 = ah_0 + (16384u >> 4);
                const uint32_t koff = (uint32_t)s * (8192u >> 4);              // k0 = 64 s: 8 atoms of 8 neurons
                MPDE_UNROLL
                for (int kk = 0; kk < 4; ++kk) {
                  const uint64_t ah = desc_join(ah_0 + kk * 2), al = desc_join(al_0 + kk * 2);
                  const uint64_t bh = desc_join(bh_0 + koff + kk * 128), bl = desc_join(bl_0 + koff + kk * 128);
                  mma_f16_ss(d, ah, bh, idesc, (s | kk) ? 1u : 0u);
                  mma_f16_ss(d, al, bh, idesc, 1);
                  mma_f16_ss(d, ah, bl, idesc, 1);
                }
                tc_commit(&cx.empty[stage]);
              }
              __syncwarp();
              if (++stage == NSTAGE) { stage = 0; phase ^= 1u; }
            }
          }
          if (elect_one()) tc_commit(&cx.dfull[pr]);         // one completion per contraction and pair: all M tiles
          __syncwarp();
        }
      }
    }
  }
  else if (MODE != MODE_LOSS && warp == NEW + 2) {
    // ===== record prefetcher: asks L2 for the records of layer l one contraction before the reverse pass of the pair's
    // tiles reads them =====
    if (lane == 0) {
      uint32_t xpar0 = 0, xpar1 = 0;
      for (int i0 = 0; i0 < n_mine; i0 += NPAIR * NSLOT) {
        const int npair = (i0 + NSLOT < n_mine) ? 2 : 1;
        for (int step = 1; step <= nsteps; ++step) {
          for (int pr = 0; pr < npair; ++pr) {
            if (pr) { mbar_wait(&cx.xready[1], xpar1); xpar1 ^= 1u; } else { mbar_wait(&cx.xready[0], xpar0); xpar0 ^= 1u; }
            const int l = 2 * L - 2 - step;                // step L + j (j = 0 .. L-2) -> layer L-2-j
            if (p.rec_prefetch && step >= L && l >= 0) {
              for (int g = 0; g < NSLOT; ++g) {
                if (i0 + pr * NSLOT + g >= n_mine) break;
                const int t = bx + (i0 + pr * NSLOT + g) * gx;
                int sgi = 0;
                while (sgi + 1 < kp.n_seg && t >= kp.seg_tile_begin[sgi + 1]) ++sgi;
                const int tile = t - kp.seg_tile_begin[sgi];
                const int head = kp.seg[sgi].head;
                const int jk = (head == HEAD_POISSON_DOMAIN) ? JET_LAP : (head == HEAD_BURGERS_DOMAIN) ? JET_XX : (head == HEAD_HYPER_DOMAIN) ? JET_G : JET_V;
                const size_t per_layer = (size_t)rec_fields_points(jk) * W;           // floats of one (tile, layer) record block
                const float* rec_tile = p.rec + (size_t)task * p.rec_task_stride + p.rec_seg_base[sgi] + (size_t)tile * L * per_layer;
                bulk_prefetch_l2(rec_tile + (size_t)l * per_layer, (uint32_t)(per_layer * sizeof(float)));
              }
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 0) tmem_dealloc(tbase, G::TMEM_COLS);
  if (tid < MAX_SEG) {
    float a = 0.f;
    for (int sl = 0; sl < NSLOT; ++sl) a += cx.terms[sl * 4 + tid];
    kp.partial_terms[(size_t)cta * MAX_SEG + tid] = a;
  }
}

}  // namespace tc
}  // namespace mpde
