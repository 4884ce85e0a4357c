// tc_hvp.cuh -- tensor-core path of the Hessian-vector product (MODE_HVP): tangent-of-(forward + adjoint) of the
// fused SIREN jet kernel (reference: the second-order pass of src/nets/maml.py:195-202).
//
// Same machinery as tc_pde_kernel (tc_kernels.cuh) with a tangent stream next to the primal one:
//   forward   z    = W^T h,      zdot    = W^T hdot    + Wdot^T h
//   reverse   abar = W zbar,     abardot = W zbardot   + Wdot zbar
// A tile is 64 jet rows (16 points x 4 channels, or 64 value points): the primal rows X (h / zbar) and the tangent rows
// Xd (hdot / zbardot) are two 64-row blocks of one MN-major B operand (64 KB each in shared memory), so W * [X | Xd] is
// a single N = 128 MMA that reads the weight stage once, and Wdot * X an N = 64 one that ACCUMULATES into the tangent
// columns (shared exponent, see agree_shifts).  Two accumulator sets per tile (primal, tangent) and TWO TILES IN FLIGHT
// (2 x NMT x 128 = 512 TMEM columns at W = 256): the tensor core runs tile A's contraction while the epilogue works on
// tile B; the epilogue parks its fp32 outputs in the accumulator columns it has just read and converts them into X / Xd
// once the other tile's contraction has released the buffers (same scheme as tc_pde_kernel).
// Spills for tc_wgrad_kernel: records (saved fields + zdot) of every layer, zbar_l and zbardot_l.
#pragma once
#include "tc_kernels.cuh"

namespace mpde {
namespace tc {

constexpr int HR = 64;                   // jet rows per HVP tile (MMA N of one row block)
constexpr int HNT = 2;                   // tiles in flight

template <int W> struct HvGeo {
  static constexpr int NMT = W / 128, KSL = W / 64, NEPI = W;
  // Two epilogue GROUPS of W threads (thread = neuron): group g owns rows [32 g, 32 g + 32) of every tile.  + one control
  // warpgroup (producer, MMA issuer, record prefetcher, one idle warp).  W = 256: 20 warps launch at 96 registers; the
  // control warpgroup drops to 32 and the 16 epilogue warps rise to 112 (setmaxnreg: 128 x 64 = 512 x 16).
  static constexpr int NG = 2, GR = HR / NG;            // groups, jet rows per group
  static constexpr bool REGSHIFT = (W == 256);
  static constexpr int EPI_REGS = 112, CTL_REGS = 32;
  static constexpr int NT = NG * W + 128;
  static constexpr int SLAB = HR * 128;                 // [64 rows][64 k] fp16
  static constexpr int XBYTES = 2 * KSL * SLAB;         // hi slabs then lo slabs (one operand)
  static constexpr int RING = NSTAGE * STAGE_BYTES;
  // per tile in flight: ybuf[HR][2], ydbuf[HR][2], red[HR] f32; per (tile, group): cmax[2][8] u32
  static constexpr int TILE_MISC = 2 * HR * 4 + 2 * HR * 4 + HR * 4;
  // shared: terms[4] f32, barriers (2*NSTAGE + 2*HNT) u64, tmem slot
  static constexpr int MISC = HNT * TILE_MISC + HNT * NG * 64 + 16 + (2 * NSTAGE + 2 * HNT) * 8 + 16;
  static constexpr int SMEM = 2 * XBYTES + RING + MISC;
  static_assert(SMEM <= 232448, "tc_hvp_kernel: shared memory");
  // TMEM columns: tile slot P, M tile mt: [W*primal (64) | W*tangent + Wdot*primal (64)] at P * TILE_COLS + mt * 128
  static constexpr int TILE_COLS = NMT * 2 * HR;
  static constexpr int TMEM_COLS = HNT * TILE_COLS;     // 256 or 512
};

template <int W> struct HvCtx {
  typedef HvGeo<W> G;
  uint8_t* xbuf; uint8_t* xdbuf; uint8_t* ring; uint8_t* misc0;
  uint32_t* cmax0; float* terms;
  uint64_t* full; uint64_t* empty; uint64_t* xready; uint64_t* dfull;     // xready[HNT], dfull[HNT]
  uint32_t* tmem_slot;
  TC_DEV void carve() {
    uint8_t* s = tc_smem_raw;
    xbuf = s; xdbuf = s + G::XBYTES; ring = s + 2 * G::XBYTES;
    misc0 = ring + G::RING;
    cmax0 = reinterpret_cast<uint32_t*>(misc0 + HNT * G::TILE_MISC);
    terms = reinterpret_cast<float*>(cmax0 + HNT * G::NG * 16);
    full = reinterpret_cast<uint64_t*>(terms + 4);
    empty = full + NSTAGE; xready = empty + NSTAGE; dfull = xready + HNT;
    tmem_slot = reinterpret_cast<uint32_t*>(dfull + HNT);
  }
  TC_DEV float* ybuf(int ts) const { return reinterpret_cast<float*>(misc0 + (size_t)ts * G::TILE_MISC); }
  TC_DEV float* ydbuf(int ts) const { return ybuf(ts) + 2 * HR; }
  TC_DEV float* red(int ts) const { return ybuf(ts) + 4 * HR; }
  TC_DEV uint32_t* cmax(int ts, int grp) const { return cmax0 + (ts * G::NG + grp) * 16; }
};

// ---------------------------------------------------------------------------------------------------------
template <int W, int HEAD>
struct HvTile {
  typedef HvGeo<W> G;
  typedef HeadTraits<HEAD> HT;
  static constexpr int JK = HT::JK, OUT = HT::OUT;
  typedef JetTraits<JK> JT;
  static constexpr int C = JT::C, NF = JT::NF, NFT = JT::NF + JT::C;
  static constexpr bool VAL = (JK == JET_V);
  static constexpr int CP = VAL ? 1 : 4;
  static constexpr int PTS = HR / CP;
  static constexpr int TCH = 8;                         // accumulator columns per chunk (2 jet points: 112 registers)
  static constexpr int PPC = TCH / CP;                  // points per chunk
  static constexpr int PG = G::GR / CP;                 // points of one group

  // MN-major B operands (see TcTile in tc_kernels.cuh): one block of 64 rows, [k = neuron][128 B = 64 rows]; 8 consecutive
  // rows of a neuron are one 16-byte chunk
  static constexpr uint32_t LO_OFF = (uint32_t)W * 128u;                 // lo parts behind the hi parts
  static TC_DEV void split2(float a, float b, uint32_t& hi, uint32_t& lo) {
    const __half2 h = __floats2half2_rn(a, b);
    const float2 f = __half22float2(h);
    const __half2 l = __floats2half2_rn(a - f.x, b - f.y);
    hi = *reinterpret_cast<const uint32_t*>(&h);
    lo = *reinterpret_cast<const uint32_t*>(&l);
  }
  static TC_DEV uint32_t chunk_off(int j, int n) {      // rows 8j .. 8j+7 of neuron n, hi part
    return (uint32_t)(n >> 3) * 1024u + (uint32_t)(n & 7) * 128u + (((uint32_t)j ^ (uint32_t)(n & 7)) << 4);
  }
  // the group's GR parked rows of this thread's neuron (TMEM columns tcol .. tcol + GR) -> fp16 hi / lo operand rows
  static TC_DEV void convert_from_tmem(uint32_t tcol, uint8_t* xb, int n, int grp, const float* ys) {
    MPDE_UNROLL
    for (int jj = 0; jj < G::GR / 8; ++jj) {
      uint32_t v[8];
      tmem_ld8(tcol + jj * 8, v);
      tmem_ld_wait();
      float f[8];
      MPDE_UNROLL
      for (int i = 0; i < 8; ++i) f[i] = __uint_as_float(v[i]) * ys[VAL ? 0 : (i & 3)];
      uint4 hv, lv;
      split2(f[0], f[1], hv.x, lv.x); split2(f[2], f[3], hv.y, lv.y);
      split2(f[4], f[5], hv.z, lv.z); split2(f[6], f[7], hv.w, lv.w);
      const uint32_t o = chunk_off(grp * (G::GR / 8) + jj, n);
      *reinterpret_cast<uint4*>(xb + o) = hv;
      *reinterpret_cast<uint4*>(xb + LO_OFF + o) = lv;
    }
  }
  // Shared exponent (round 2): W * tangent and Wdot * primal are accumulated in ONE set of accumulator columns, so their
  // scales must agree: sh_tangent + sw = sh_primal + swd, i.e. sh_tangent = sh_primal + delta with delta = swd - sw of the
  // contraction that consumes the operand.  Per channel e = min(max-safe tangent shift, max-safe primal shift + delta);
  // tangent shift = e, primal shift = e - delta: never above a max-safe shift (no fp16 overflow); the operand scaled below
  // its own optimum belongs to the term that is the smaller one at the common scale.  cm = channel maxima [8] (uint bits).
  template <bool FWD>
  static TC_DEV void equalize(const uint32_t* cm, int delta, int (&sh)[8]) {
    MPDE_UNROLL
    for (int c = 0; c < 4; ++c) {
      const uint32_t mp = cm[c], md = cm[4 + c];
      // an all-zero block puts no constraint on the common exponent; FWD: the primal value channel is sin, |a_0| <= 1
      const int sp = (FWD && c == 0) ? 14 : (mp ? shift_for_max(__uint_as_float(mp)) : 100);
      const int sd = md ? shift_for_max(__uint_as_float(md)) : 100;
      int e = sp + delta < sd ? sp + delta : sd;
      e = e < -90 ? -90 : (e > 90 ? 90 : e);
      int ep = e - delta;
      ep = ep < -90 ? -90 : (ep > 90 ? 90 : ep);
      sh[4 + c] = ep + delta;
      sh[c] = ep;
    }
  }
  // fold the channel maxima (mx[0..3] primal, mx[4..7] tangent), barrier among the group's warps, read the shifts.
  // `cmx` = cmax[2][8] of (tile slot, group); the buffer of agreement seq - 1 stays intact until the barrier of agreement
  // seq, so the next phase re-derives the shifts from it (prev_shifts) instead of carrying them in registers.
  template <bool FWD>
  static TC_DEV void agree_shifts(uint32_t* cmx, int grp, int& seq, int delta, const float (&mx)[8], float (&ys)[8], int tid, int lane) {
    uint32_t* cm = cmx + (seq & 1) * 8;
    MPDE_UNROLL
    for (int c = 0; c < 8; ++c) {
      const bool used = VAL ? (c == 4 || (!FWD && c == 0)) : ((c & 3) < C && !(FWD && c == 0));
      if (used) fold_max(cm + c, mx[c], lane);
    }
    epi_sync<G::NEPI>(grp);                             // the group's own barrier: its rows carry their own shifts
    int sh[8];
    equalize<FWD>(cm, delta, sh);
    MPDE_UNROLL
    for (int c = 0; c < 8; ++c) ys[c] = pow2i(sh[c]);
    if (tid == 0) {
      uint32_t* other = cmx + ((seq & 1) ^ 1) * 8;
      MPDE_UNROLL
      for (int c = 0; c < 8; ++c) other[c] = 0u;
    }
    ++seq;
  }
  template <bool FWD>
  static TC_DEV void prev_shifts(const uint32_t* cmx, int seq, int delta, int (&sh)[8]) {
    equalize<FWD>(cmx + ((seq - 1) & 1) * 8, delta, sh);
  }
  // one point's outputs into the chunks being parked (`qs`: index of the point inside its chunk, compile-time)
  static TC_DEV void park_point(uint32_t (&o)[TCH], uint32_t (&od)[TCH], int qs, const float* a, const float* ad, float (&mx)[8]) {
    if (VAL) {
      o[qs] = __float_as_uint(a[0]); od[qs] = __float_as_uint(ad[0]);
      mx[0] = fmaxf(mx[0], fabsf(a[0])); mx[4] = fmaxf(mx[4], fabsf(ad[0]));
    } else {
      MPDE_UNROLL
      for (int c = 0; c < 4; ++c) {
        o[4 * qs + c] = (c < C) ? __float_as_uint(a[c]) : 0u;
        od[4 * qs + c] = (c < C) ? __float_as_uint(ad[c]) : 0u;
      }
      MPDE_UNROLL
      for (int c = 0; c < C; ++c) { mx[c] = fmaxf(mx[c], fabsf(a[c])); mx[4 + c] = fmaxf(mx[4 + c], fabsf(ad[c])); }
    }
  }
  static TC_DEV void tmem_st8(uint32_t taddr, const uint32_t (&v)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                 ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
  }
  // the point's coordinates (L1 / L2-resident; every lane of the warp reads the same address)
  static TC_DEV float2 point_xy(const float* pts, int gp, int npts) {
    const bool ok = gp < npts;
    const float2 v = __ldg(reinterpret_cast<const float2*>(pts + 2 * (size_t)(ok ? gp : npts - 1)));
    return ok ? v : make_float2(0.f, 0.f);
  }
  // barrier among ALL epilogue threads (both groups): named barrier 3
  static TC_DEV void all_sync() { asm volatile("bar.sync 3, %0;" ::"n"(G::NG * G::NEPI) : "memory"); }
  // wait until the other tile's contraction has released X / Xd, convert the parked rows, publish the operand
  static TC_DEV void publish(HvCtx<W>& cx, int grp, int ts, uint32_t tp, uint32_t tt, int n, const float* ys, int wait_other,
                             uint32_t par_other) {
    tmem_st_wait();                                      // this thread's parked rows are in tensor memory
    if (wait_other) { mbar_wait(cx.dfull + (ts ^ 1), par_other); tc_fence_after(); }
    convert_from_tmem(tp, cx.xbuf, n, grp, ys);
    convert_from_tmem(tt, cx.xdbuf, n, grp, ys + 4);
    fence_proxy_async_smem();
    tc_fence_before();
    mbar_arrive(cx.xready + ts);
  }

  // ---- one PHASE of one tile (both groups of W epilogue threads run it, each on its rows) -------------------
  // Phases of a tile (L layers); phase ph waits for contraction ph of its tile (ph >= 1) and, unless it is the last, ends
  // by publishing the operands of contraction ph + 1:
  //   0          layer 0 (direct)            l = 1..L-1  hidden layer l
  //   L          output layer -> head (value, seeds and their tangents), reverse layer L-1 (seeds, no contraction)
  //   L + j      adjoint accumulators -> reverse layer L-1-j (j = 1..L-1); layer 0 publishes nothing
  static TC_DEV void phase(const HvParams& hp, HvCtx<W>& cx, const SegDesc& sg, int seg_id, int task, int tile, int ts, int ph,
                           uint32_t tbase, int& seq, uint32_t par_self, int wait_other, uint32_t par_other, const int* wshift,
                           const int* wshift_d) {
    const TcParams& p = hp.t;
    const KParams& kp = p.k;
    const int tid = threadIdx.x, lane = tid & 31;
    const int hwarp = warp_idx_uniform();
    const int grp = hwarp / (G::NEPI / 32);             // epilogue group: rows [GR * grp, GR * grp + GR) of the tile
    const int n = tid - grp * G::NEPI, mt = n >> 7;     // this thread's neuron
    const int warp = hwarp - grp * (G::NEPI / 32);      // warp inside the group (groups start at multiples of 4 warps)
    const float om = kp.omega;
    const int L = kp.L;
    const float* th = kp.th + (size_t)task * kp.th_stride;
    const float* thd = kp.thd + (size_t)task * kp.thd_stride;
    const float* prm = kp.prm + (size_t)task * kp.nprm;
    const int pt0 = tile * PTS;
    const float* pts = sg.pts + (size_t)task * sg.task_stride;
    // this thread's accumulator columns: primal tp, tangent tt (its group's rows of its M tile of tile slot ts)
    const uint32_t tp = tmem_addr(tbase, 32u * (uint32_t)(warp & 3),
                                  (uint32_t)ts * G::TILE_COLS + (uint32_t)mt * 2u * HR + (uint32_t)(grp * G::GR));
    const uint32_t tt = tp + HR;
    float* ybuf = cx.ybuf(ts);
    float* ydbuf = cx.ydbuf(ts);
    float* red = cx.red(ts);
    uint32_t* cmx = cx.cmax(ts, grp);
    uint64_t* dfull = cx.dfull + ts;
    float* recp = p.rec + (size_t)task * p.rec_task_stride + p.rec_seg_base[seg_id] + (size_t)tile * L * (NFT * PTS) * W;   // see rec_off
    const int pg0 = grp * PG;                           // first point of this group

    if (ph == 0) {
      // ---- layer 0 (in_dim = 2): direct ----
      float ys[8];
      const float w00 = mpde_ldg(th + n), w01 = mpde_ldg(th + W + n), b0 = mpde_ldg(th + 2 * W + n);
      const float d00 = mpde_ldg(thd + n), d01 = mpde_ldg(thd + W + n), db0 = mpde_ldg(thd + 2 * W + n);
      float mx[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      MPDE_NOUNROLL
      for (int ch = 0; ch < G::GR / TCH; ++ch) {
        uint32_t o[TCH], od[TCH];
        MPDE_UNROLL
        for (int q = 0; q < PPC; ++q) {
          const int pl = pg0 + ch * PPC + q;
          const float2 xy = point_xy(pts, pt0 + pl, sg.n);
          float z[4], zd[4], a[4], ad[4], sv[5];
          z[0] = fmaf(w00, xy.x, fmaf(w01, xy.y, b0)); z[1] = w00; z[2] = w01; z[3] = 0.f;
          zd[0] = fmaf(d00, xy.x, fmaf(d01, xy.y, db0)); zd[1] = d00; zd[2] = d01; zd[3] = 0.f;
          jet_act_fwd<JK>(om, z, a, sv);
          jet_act_tan<JK>(om, sv, zd, ad);
          MPDE_UNROLL
          for (int f = 0; f < NF; ++f) recp[rec_off(NFT, PTS, 16 / CP, f, pl, n)] = sv[f];
          MPDE_UNROLL
          for (int c = 0; c < C; ++c) recp[rec_off(NFT, PTS, 16 / CP, NF + c, pl, n)] = zd[c];
          park_point(o, od, q, a, ad, mx);
        }
        tmem_st8(tp + ch * TCH, o); tmem_st8(tt + ch * TCH, od);
      }
      agree_shifts<true>(cmx, grp, seq, wshift_d[1] - wshift[1], mx, ys, n, lane);
      publish(cx, grp, ts, tp, tt, n, ys, wait_other, par_other);
      return;
    }

    if (ph < L) {
      // ---- hidden layer l = ph ----
      const int l = ph;
      int sh[8];
      prev_shifts<true>(cmx, seq, wshift_d[l] - wshift[l], sh);
      const float bias = mpde_ldg(th + ph_off_layer(W, l) + 2 * W * W + n);
      const float biasd = mpde_ldg(thd + ph_off_layer(W, l) + 2 * W * W + n);
      float zs[4], zs1[4], ys[8];
      {
        const int sw = wshift[l];
        for (int c = 0; c < 4; ++c) { zs[c] = pow2i(-sh[c] - sw); zs1[c] = pow2i(-sh[4 + c] - sw); }
      }
      mbar_wait(dfull, par_self);                       // every MMA of this contraction has completed
      tc_fence_after();
      float mx[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      MPDE_NOUNROLL
      for (int ch = 0; ch < G::GR / TCH; ++ch) {
        uint32_t v0[TCH], v1[TCH], o[TCH], od[TCH];
        tmem_ld8(tp + ch * TCH, v0); tmem_ld8(tt + ch * TCH, v1);
        tmem_ld_wait();
        MPDE_UNROLL
        for (int q = 0; q < PPC; ++q) {
          const int pl = pg0 + ch * PPC + q;
          float z[4] = {0.f, 0.f, 0.f, 0.f}, zd[4] = {0.f, 0.f, 0.f, 0.f}, a[4], ad[4], sv[5];
          MPDE_UNROLL
          for (int c = 0; c < C; ++c) {
            z[c] = __uint_as_float(v0[CP * q + c]) * zs[c];
            zd[c] = __uint_as_float(v1[CP * q + c]) * zs1[c];
          }
          z[0] += bias; zd[0] += biasd;
          jet_act_fwd<JK>(om, z, a, sv);
          jet_act_tan<JK>(om, sv, zd, ad);
          MPDE_UNROLL
          for (int f = 0; f < NF; ++f) recp[(size_t)l * NFT * (PTS * W) + rec_off(NFT, PTS, 16 / CP, f, pl, n)] = sv[f];
          MPDE_UNROLL
          for (int c = 0; c < C; ++c) recp[(size_t)l * NFT * (PTS * W) + rec_off(NFT, PTS, 16 / CP, NF + c, pl, n)] = zd[c];
          park_point(o, od, q, a, ad, mx);
        }
        tmem_st8(tp + ch * TCH, o); tmem_st8(tt + ch * TCH, od);
      }
      agree_shifts<true>(cmx, grp, seq, wshift_d[l + 1] - wshift[l + 1], mx, ys, n, lane);
      publish(cx, grp, ts, tp, tt, n, ys, wait_other, par_other);
      return;
    }

    if (ph == L) {
      // ---- output layer: lanes 0..OUT-1 of the first M tile hold y, ydot ----
      int sh[8];
      prev_shifts<true>(cmx, seq, wshift_d[L] - wshift[L], sh);
      mbar_wait(dfull, par_self);
      tc_fence_after();
      if (warp == 0) {
        // each group's first warp: the output rows of ITS rows (descaled with the group's own shifts)
        float zs[4], zs1[4];
        {
          const int sw = wshift[L];
          for (int c = 0; c < 4; ++c) { zs[c] = pow2i(-sh[c] - sw); zs1[c] = pow2i(-sh[4 + c] - sw); }
        }
        MPDE_NOUNROLL
        for (int ch = 0; ch < G::GR / TCH; ++ch) {
          uint32_t v0[TCH], v1[TCH];
          tmem_ld8(tp + ch * TCH, v0); tmem_ld8(tt + ch * TCH, v1);
          tmem_ld_wait();
          if (lane < OUT) {
            MPDE_UNROLL
            for (int q = 0; q < TCH; ++q) {
              const int c = VAL ? 0 : (q & 3);
              const int r = grp * G::GR + ch * TCH + q;
              ybuf[r * 2 + lane] = __uint_as_float(v0[q]) * zs[c];
              ydbuf[r * 2 + lane] = __uint_as_float(v1[q]) * zs1[c];
            }
          }
        }
      }
      tc_fence_before();
      all_sync();

      // ---- head: value, seeds ybar and their tangents ybardot ----
      if (tid < PTS) {
        const int pl = tid;
        const float* bL = th + ph_off_out(W, L) + W * OUT;
        const float* bLd = thd + ph_off_out(W, L) + W * OUT;
        float u[OUT * C], ud[OUT * C], ub[OUT * C], ubd[OUT * C];
        MPDE_UNROLL
        for (int o = 0; o < OUT; ++o)
          MPDE_UNROLL
          for (int c = 0; c < C; ++c) {
            float y = ybuf[(CP * pl + c) * 2 + o], yd = ydbuf[(CP * pl + c) * 2 + o];
            if (c == 0) { y += mpde_ldg(bL + o); yd += mpde_ldg(bLd + o); }
            u[o * C + c] = y; ud[o * C + c] = yd;
          }
        const bool valid = (pt0 + pl) < sg.n;
        const float2 xy = point_xy(pts, pt0 + pl, sg.n);
        float res = 0.f;
        float val = head_eval<HEAD>(u, xy.x, xy.y, prm, sg.coef, ub, &res);
        head_tangent<HEAD>(u, ud, xy.x, xy.y, prm, sg.coef, ubd);
        if (!valid) val = 0.f;
        red[pl] = val;
        MPDE_UNROLL
        for (int o = 0; o < OUT; ++o)
          MPDE_UNROLL
          for (int c = 0; c < CP; ++c) {
            ybuf[(CP * pl + c) * 2 + o] = (valid && c < C) ? ub[o * C + c] : 0.f;
            ydbuf[(CP * pl + c) * 2 + o] = (valid && c < C) ? ubd[o * C + c] : 0.f;
          }
      }
      all_sync();
      if (tid < 32) {
        float s = 0.f;
        MPDE_UNROLL
        for (int j = 0; j < (PTS + 31) / 32; ++j) s += (lane + 32 * j < PTS) ? red[lane + 32 * j] : 0.f;
        MPDE_UNROLL
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) cx.terms[seg_id] += s;
      }
    }

    // ==================== reverse pass and its tangent ====================
    // one instantiation per layer kind (l = L-1: seeds, no accumulators / 0 < l < L-1 / l = 0: nothing to publish): no
    // runtime branch inside the per-point loops
    const int l = 2 * L - 1 - ph;
    if (ph == L) reverse_layer<RK_LAST>(hp, cx, sg, seg_id, task, tile, ts, l, tp, tt, recp, seq, par_self, wait_other, par_other, wshift, wshift_d);
    else if (l == 0) reverse_layer<RK_FIRST>(hp, cx, sg, seg_id, task, tile, ts, l, tp, tt, recp, seq, par_self, wait_other, par_other, wshift, wshift_d);
    else reverse_layer<RK_MID>(hp, cx, sg, seg_id, task, tile, ts, l, tp, tt, recp, seq, par_self, wait_other, par_other, wshift, wshift_d);
  }

  enum { RK_LAST = 0, RK_MID = 1, RK_FIRST = 2 };
  template <int RK>
  static TC_DEV void reverse_layer(const HvParams& hp, HvCtx<W>& cx, const SegDesc& sg, int seg_id, int task, int tile, int ts,
                                   int l, uint32_t tp, uint32_t tt, const float* recp, int& seq, uint32_t par_self,
                                   int wait_other, uint32_t par_other, const int* wshift, const int* wshift_d) {
    constexpr bool LAST = (RK == RK_LAST), FIRST = (RK == RK_FIRST);
    const TcParams& p = hp.t;
    const KParams& kp = p.k;
    const int tid = threadIdx.x, lane = tid & 31;
    const int grp = warp_idx_uniform() / (G::NEPI / 32);
    const int n = tid - grp * G::NEPI;
    const float om = kp.omega;
    const int L = kp.L;
    const float* th = kp.th + (size_t)task * kp.th_stride;
    const float* thd = kp.thd + (size_t)task * kp.thd_stride;
    const float* pts = sg.pts + (size_t)task * sg.task_stride;
    const int pt0 = tile * PTS;
    const float* ybuf = cx.ybuf(ts);
    const float* ydbuf = cx.ydbuf(ts);
    uint32_t* cmx = cx.cmax(ts, grp);
    const int pg0 = grp * PG;
    const bool wg = hp.want_grad != 0;
    // per (CTA, group) partials of the small parameters: both groups own a thread for neuron n
    const size_t pso = (((size_t)task * gridDim.x + blockIdx.x) * G::NG + grp) * p.PS;
    float* psg = p.psmall + pso;
    float* psh = hp.psmall_h + pso;
    float wl[OUT], wld[OUT];
    MPDE_UNROLL
    for (int o = 0; o < OUT; ++o) { wl[o] = 0.f; wld[o] = 0.f; }
    if (LAST) {
      MPDE_UNROLL
      for (int o = 0; o < OUT; ++o) {
        wl[o] = mpde_ldg(th + ph_off_out(W, L) + n * OUT + o);
        wld[o] = mpde_ldg(thd + ph_off_out(W, L) + n * OUT + o);
      }
      if (tid < OUT) {                                  // group 0 only: the whole tile's seeds
        float g = 0.f, h = 0.f;
        for (int pl = 0; pl < PTS; ++pl) { g += ybuf[(CP * pl) * 2 + tid]; h += ydbuf[(CP * pl) * 2 + tid]; }
        psh[ps_off_wl(W, L) + W * OUT + tid] += h;
        if (wg) psg[ps_off_wl(W, L) + W * OUT + tid] += g;
      }
    }
    const float* rl = recp + (size_t)l * NFT * (PTS * W);   // this layer's record block (layout: rec_off)
    float zs[4] = {0.f, 0.f, 0.f, 0.f}, zs1[4] = {0.f, 0.f, 0.f, 0.f};
    // the first chunk's record loads are issued before the accumulator wait (they do not depend on it)
    float svb[PPC][NF + C];
    MPDE_UNROLL
    for (int q = 0; q < PPC; ++q)
      MPDE_UNROLL
      for (int f = 0; f < NF + C; ++f) svb[q][f] = rl[rec_off(NFT, PTS, 16 / CP, f, pg0 + q, n)];
    if (!LAST) {
      int shz[8];
      prev_shifts<false>(cmx, seq, wshift_d[l + 1] - wshift[l + 1], shz);
      const int sw = wshift[l + 1];
      for (int c = 0; c < 4; ++c) { zs[c] = pow2i(-shz[c] - sw); zs1[c] = pow2i(-shz[4 + c] - sw); }
      mbar_wait(cx.dfull + ts, par_self);              // all adjoint MMAs of this contraction have completed
      tc_fence_after();
    }
    float mx[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float bsum = 0.f, bsumd = 0.f, g0 = 0.f, g1 = 0.f, h0 = 0.f, h1 = 0.f, gw[OUT], hw[OUT];
    MPDE_UNROLL
    for (int o = 0; o < OUT; ++o) { gw[o] = 0.f; hw[o] = 0.f; }
    const size_t zoff = (size_t)task * p.zsp_task_stride +
                        ((size_t)(kp.seg_tile_begin[seg_id] + tile) * (L - 1) + (FIRST ? 0 : l - 1)) * HR * W + n;
    float* zl = p.zsp + zoff;                              // only dereferenced for l >= 1
    float* zld = hp.zspd + zoff;
    MPDE_NOUNROLL
    for (int ch = 0; ch < G::GR / TCH; ++ch) {
      uint32_t v0[TCH], v1[TCH], o[TCH], od[TCH];
      if (!LAST) {
        tmem_ld8(tp + ch * TCH, v0); tmem_ld8(tt + ch * TCH, v1);
        tmem_ld_wait();
      }
      const int chn = (ch + 1 < G::GR / TCH) ? ch + 1 : ch;   // next chunk's records (clamped: a repeated valid load, no branch)
      MPDE_UNROLL
      for (int q = 0; q < PPC; ++q) {
        const int pl = pg0 + ch * PPC + q;
        float ab[4] = {0.f, 0.f, 0.f, 0.f}, abd[4] = {0.f, 0.f, 0.f, 0.f}, sv[5], zd[4] = {0.f, 0.f, 0.f, 0.f};
        float zb[4] = {0.f, 0.f, 0.f, 0.f}, zbd[4] = {0.f, 0.f, 0.f, 0.f};
        MPDE_UNROLL
        for (int c = 0; c < C; ++c) {
          if (!LAST) {
            ab[c] = __uint_as_float(v0[CP * q + c]) * zs[c];
            abd[c] = __uint_as_float(v1[CP * q + c]) * zs1[c];
          } else {
            float a_ = 0.f, ad_ = 0.f;
            MPDE_UNROLL
            for (int oo = 0; oo < OUT; ++oo) {
              const float yb = ybuf[(CP * pl + c) * 2 + oo], ybd = ydbuf[(CP * pl + c) * 2 + oo];
              a_ = fmaf(wl[oo], yb, a_);
              ad_ = fmaf(wl[oo], ybd, fmaf(wld[oo], yb, ad_));
            }
            ab[c] = a_; abd[c] = ad_;
          }
        }
        MPDE_UNROLL
        for (int f = 0; f < NF; ++f) sv[f] = svb[q][f];
        MPDE_UNROLL
        for (int c = 0; c < C; ++c) zd[c] = svb[q][NF + c];
        // this slot's record of the NEXT chunk: in flight while the chunk is processed
        MPDE_UNROLL
        for (int f = 0; f < NF + C; ++f) svb[q][f] = rl[rec_off(NFT, PTS, 16 / CP, f, pg0 + chn * PPC + q, n)];
        jet_act_bwd<JK>(om, sv, ab, zb);
        jet_act_bwd_tan<JK>(om, sv, zd, ab, abd, zbd);
        bsum += zb[0]; bsumd += zbd[0];
        if (LAST) {
          float a[4], ad[4];
          jet_act_from_saved<JK>(om, sv, a);
          jet_act_tan<JK>(om, sv, zd, ad);
          MPDE_UNROLL
          for (int oo = 0; oo < OUT; ++oo)
            MPDE_UNROLL
            for (int c = 0; c < C; ++c) {
              const float yb = ybuf[(CP * pl + c) * 2 + oo], ybd = ydbuf[(CP * pl + c) * 2 + oo];
              gw[oo] = fmaf(a[c], yb, gw[oo]);
              hw[oo] = fmaf(ad[c], yb, fmaf(a[c], ybd, hw[oo]));
            }
        }
        if (FIRST) {
          const float2 xy = point_xy(pts, pt0 + pl, sg.n);
          if (VAL) {
            g0 = fmaf(xy.x, zb[0], g0); g1 = fmaf(xy.y, zb[0], g1);
            h0 = fmaf(xy.x, zbd[0], h0); h1 = fmaf(xy.y, zbd[0], h1);
          } else {
            g0 += fmaf(xy.x, zb[0], zb[1]); g1 += fmaf(xy.y, zb[0], zb[2]);
            h0 += fmaf(xy.x, zbd[0], zbd[1]); h1 += fmaf(xy.y, zbd[0], zbd[2]);
          }
        } else {
          // fp32 spill for tc_wgrad_kernel (written once, never read back here) + the parked copy for the next contraction
          MPDE_UNROLL
          for (int c = 0; c < CP; ++c) {
            zl[(size_t)(CP * pl + c) * W] = c < C ? zb[c] : 0.f;
            zld[(size_t)(CP * pl + c) * W] = c < C ? zbd[c] : 0.f;
          }
          park_point(o, od, q, zb, zbd, mx);
        }
      }
      if (!FIRST) { tmem_st8(tp + ch * TCH, o); tmem_st8(tt + ch * TCH, od); }
    }
    psh[ps_off_bias(W, l) + n] += bsumd;
    if (wg) psg[ps_off_bias(W, l) + n] += bsum;
    if (LAST) {
      MPDE_UNROLL
      for (int oo = 0; oo < OUT; ++oo) {
        psh[ps_off_wl(W, L) + n * OUT + oo] += hw[oo];
        if (wg) psg[ps_off_wl(W, L) + n * OUT + oo] += gw[oo];
      }
    }
    if (FIRST) {
      psh[n] += h0; psh[W + n] += h1;
      if (wg) { psg[n] += g0; psg[W + n] += g1; }
      return;
    }
    float ysz[8];
    agree_shifts<false>(cmx, grp, seq, wshift_d[l] - wshift[l], mx, ysz, n, lane);
    publish(cx, grp, ts, tp, tt, n, ysz, wait_other, par_other);
  }
};

// ---------------------------------------------------------------------------------------------------------
// kernel: grid = (ctas_per_task, T); each CTA walks the task's tiles t_i = blockIdx.x + i * gridDim.x two at a time
// (tile slots A = t_{2q}, B = t_{2q+1}); see the header for the overlap scheme
// ---------------------------------------------------------------------------------------------------------
template <int W, int PDE>
TC_DEV void hv_dispatch_phase(const HvParams& hp, HvCtx<W>& cx, int t, int task, int ts, int ph, uint32_t tbase, int& seq,
                              uint32_t par_self, int wait_other, uint32_t par_other, const int* wshift, const int* wshift_d) {
  const KParams& kp = hp.t.k;
  int s = 0;
  while (s + 1 < kp.n_seg && t >= kp.seg_tile_begin[s + 1]) ++s;
  const int tile = t - kp.seg_tile_begin[s];
  const SegDesc& sg = kp.seg[s];
#define HV_PHASE(H) HvTile<W, H>::phase(hp, cx, sg, s, task, tile, ts, ph, tbase, seq, par_self, wait_other, par_other, wshift, wshift_d)
  if (PDE == PDE_POISSON) {
    if (sg.head == HEAD_POISSON_DOMAIN) HV_PHASE(HEAD_POISSON_DOMAIN); else HV_PHASE(HEAD_POISSON_BOUNDARY);
  } else if (PDE == PDE_BURGERS) {
    if (sg.head == HEAD_BURGERS_DOMAIN) HV_PHASE(HEAD_BURGERS_DOMAIN); else HV_PHASE(HEAD_BURGERS_IC);
  } else {
    if (sg.head == HEAD_HYPER_DOMAIN) HV_PHASE(HEAD_HYPER_DOMAIN);
    else if (sg.head == HEAD_HYPER_TOP) HV_PHASE(HEAD_HYPER_TOP);
    else HV_PHASE(HEAD_HYPER_BOTTOM);
  }
#undef HV_PHASE
}

template <int W, int PDE>
__global__ void __launch_bounds__(HvGeo<W>::NT, 1) tc_hvp_kernel(const __grid_constant__ HvParams hp) {
  typedef HvGeo<W> G;
  HvCtx<W> cx;
  cx.carve();
  const TcParams& p = hp.t;
  const KParams& kp = p.k;
  const int tid = threadIdx.x, warp = warp_idx_uniform(), lane = tid & 31;
  const int task = blockIdx.y;
  const int cta = blockIdx.y * gridDim.x + blockIdx.x;
  const int L = kp.L;
  const int n_tiles = kp.seg_tile_begin[kp.n_seg];
  const int gx = (int)gridDim.x, bx = (int)blockIdx.x;

  if (tid == 0) {
    if (smem_u32(tc_smem_raw) & 1023u) { printf("tc_hvp_kernel: dynamic shared memory is not 1024-byte aligned\n"); __trap(); }
    for (int i = 0; i < NSTAGE; ++i) { mbar_init(&cx.full[i], 1); mbar_init(&cx.empty[i], 1); }
    for (int i = 0; i < HNT; ++i) { mbar_init(&cx.xready[i], G::NG * G::NEPI); mbar_init(&cx.dfull[i], 1); }
    mbar_fence_init();
    for (int i = 0; i < HNT * G::NG * 16; ++i) cx.cmax0[i] = 0u;
    for (int i = 0; i < MAX_SEG; ++i) cx.terms[i] = 0.f;
  }
  if (warp == 0) tmem_alloc(cx.tmem_slot, G::TMEM_COLS);
  {
    float* psg = p.psmall + (size_t)cta * G::NG * p.PS;   // one partial per (CTA, epilogue group)
    float* psh = hp.psmall_h + (size_t)cta * G::NG * p.PS;
    for (int i = tid; i < G::NG * p.PS; i += G::NT) { psh[i] = 0.f; if (hp.want_grad) psg[i] = 0.f; }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tbase = *cx.tmem_slot;
  constexpr int PER = G::NMT * G::KSL;                       // stages per hidden layer (per image)
  const int nst_fwd = (L - 1) * PER + G::KSL;
  const int nsteps = 2 * L - 1;                              // L-1 hidden, output, L-1 adjoint contractions
  const int nph = nsteps + 1;
  const int n_mine = (n_tiles > bx) ? (n_tiles - bx + gx - 1) / gx : 0;   // tiles of this CTA

  constexpr int NEW = G::NG * G::NEPI / 32;                  // epilogue warps (both groups)
  if (G::REGSHIFT) {
    if (warp < NEW) setmaxnreg_inc<G::EPI_REGS>(); else setmaxnreg_dec<G::CTL_REGS>();
  }
  if (warp < NEW) {
    // ===== epilogue warps (both groups): tile A phase ph, tile B phase ph, ... =====
    int seqA = 0, seqB = 0;
    uint32_t cA = 0, cB = 0;                                  // contractions of slot A / B completed before this pair of tiles
    const int* wshift = p.wshift + (size_t)task * p.wshift_stride;
    const int* wshift_d = hp.wshift_d + (size_t)task * hp.wshift_d_stride;
    for (int i0 = 0; i0 < n_mine; i0 += HNT) {
      const bool hasB = i0 + 1 < n_mine;
      const int tA = bx + i0 * gx, tB = bx + (i0 + 1) * gx;
      MPDE_NOUNROLL
      for (int ph = 0; ph < nph; ++ph) {
        MPDE_NOUNROLL
        for (int ts = 0; ts < (hasB ? HNT : 1); ++ts) {      // ONE call site: the tile slot is a runtime value
          // slot A: contraction ph (1-based) is completion cA + ph - 1 of dfull[0]; before publishing the operands of
          // contraction ph + 1 it waits for slot B's contraction ph (the last reader of X / Xd).  Slot B: before
          // publishing it waits for slot A's contraction ph + 1.
          const uint32_t cs = ts ? cB : cA, co = ts ? cA : cB;
          const uint32_t par_self = (cs + (uint32_t)ph - 1u) & 1u;
          const uint32_t par_other = ts ? ((co + (uint32_t)ph) & 1u) : ((co + (uint32_t)ph - 1u) & 1u);
          const int wait_other = ts ? 1 : ((ph >= 1 && hasB) ? 1 : 0);
          int seq = ts ? seqB : seqA;
          hv_dispatch_phase<W, PDE>(hp, cx, ts ? tB : tA, task, ts, ph, tbase, seq, par_self, wait_other, par_other, wshift, wshift_d);
          if (ts) seqB = seq; else seqA = seq;
        }
      }
      cA += (uint32_t)nsteps;
      if (hasB) cB += (uint32_t)nsteps;
    }
  } else if (warp == NEW) {
    // ===== producer: per contraction, tile slot and M tile, the stages of W then the stages of Wdot =====
    if (lane == 0) {
      const uint8_t* img = p.wimg + (size_t)task * p.wimg_stride;
      const uint8_t* imgd = hp.wimg_d + (size_t)task * hp.wimg_d_stride;
      uint32_t stage = 0, phase = 0;
      for (int i0 = 0; i0 < n_mine; i0 += HNT) {
        const int nts = (i0 + 1 < n_mine) ? HNT : 1;
        for (int step = 1; step <= nsteps; ++step) {
          const int nmt = (step != L) ? G::NMT : 1;
          // first stage of this contraction in the image: hidden layer `step`, output layer, adjoint copies (L-1 .. 1)
          const int base = (step < L) ? (step - 1) * PER : (step == L ? (L - 1) * PER : nst_fwd + (step - L - 1) * PER);
          for (int ts = 0; ts < nts; ++ts)
            for (int mt = 0; mt < nmt; ++mt)
              for (int which = 0; which < 2; ++which)
                for (int s = 0; s < G::KSL; ++s) {
                  const uint8_t* src = (which ? imgd : img) + (size_t)(base + mt * G::KSL + s) * STAGE_BYTES;
                  mbar_wait(&cx.empty[stage], phase ^ 1u);
                  mbar_arrive_expect_tx(&cx.full[stage], STAGE_BYTES);
                  bulk_g2s(cx.ring + (size_t)stage * STAGE_BYTES, src, STAGE_BYTES, &cx.full[stage]);
                  if (++stage == NSTAGE) { stage = 0; phase ^= 1u; }
                }
        }
      }
    }
  } else if (warp == NEW + 1) {
    // ===== MMA issuer: the whole warp runs the loop (warp-uniform), one elected lane issues =====
    constexpr uint32_t idesc = make_idesc(FMT_F16, 128, HR, 0, 1);       // A (weights) K-major, B (jet rows) MN-major
    constexpr uint32_t idesc2 = make_idesc(FMT_F16, 128, 2 * HR, 0, 1);  // N = 128: rows of X then rows of Xd
    constexpr uint32_t LO_OFF = (uint32_t)W * 128u;
    const uint32_t xb = smem_u32(cx.xbuf);                // Xd sits XBYTES behind X
    const uint32_t ring0 = smem_u32(cx.ring);
    // lower descriptor words of the B operand at k = 0: W * [X | Xd] (second 64-row block XBYTES further) and Wdot * X
    const uint32_t bh2_0 = desc_lo_mnmajor(xb, (uint32_t)G::XBYTES), bl2_0 = desc_lo_mnmajor(xb + LO_OFF, (uint32_t)G::XBYTES);
    const uint32_t bh_0 = desc_lo_mnmajor(xb, LO_OFF), bl_0 = desc_lo_mnmajor(xb + LO_OFF, LO_OFF);
    uint32_t stage = 0, phase = 0, xpar0 = 0, xpar1 = 0;
    for (int i0 = 0; i0 < n_mine; i0 += HNT) {
      const int nts = (i0 + 1 < n_mine) ? HNT : 1;
      for (int step = 1; step <= nsteps; ++step) {
        const int nmt = (step != L) ? G::NMT : 1;
        for (int ts = 0; ts < nts; ++ts) {
          if (ts) { mbar_wait(&cx.xready[1], xpar1); xpar1 ^= 1u; } else { mbar_wait(&cx.xready[0], xpar0); xpar0 ^= 1u; }
          tc_fence_after();
          for (int mt = 0; mt < nmt; ++mt) {
            // Wdot * X accumulates into the tangent columns
            const uint32_t d01 = tbase + (uint32_t)ts * (uint32_t)G::TILE_COLS + (uint32_t)mt * 2u * HR, d2 = d01 + HR;
            for (int which = 0; which < 2; ++which) {
              for (int s = 0; s < G::KSL; ++s) {
                mbar_wait(&cx.full[stage], phase);
                tc_fence_after();
                if (elect_one()) {
                  const uint32_t ah_0 = desc_lo_kmajor(ring0 + stage * (uint32_t)STAGE_BYTES), al_0 = ah_0 + (16384u >> 4);
                  const uint32_t koff = (uint32_t)s * (8192u >> 4);            // k0 = 64 s: 8 atoms of 8 neurons
                  MPDE_UNROLL
                  for (int kk = 0; kk < 4; ++kk) {
                    const uint64_t ah = desc_join(ah_0 + kk * 2), al = desc_join(al_0 + kk * 2);      // 32 bytes of K per MMA
                    if (which == 0) {
                      const uint32_t acc = (s | kk) ? 1u : 0u;
                      const uint64_t bh2 = desc_join(bh2_0 + koff + kk * 128), bl2 = desc_join(bl2_0 + koff + kk * 128);
                      mma_f16_ss(d01, ah, bh2, idesc2, acc); mma_f16_ss(d01, al, bh2, idesc2, 1); mma_f16_ss(d01, ah, bl2, idesc2, 1);
                    } else {
                      const uint64_t bh = desc_join(bh_0 + koff + kk * 128), bl = desc_join(bl_0 + koff + kk * 128);
                      mma_f16_ss(d2, ah, bh, idesc, 1); mma_f16_ss(d2, al, bh, idesc, 1); mma_f16_ss(d2, ah, bl, idesc, 1);   // += Wdot * primal
                    }
                  }
                  tc_commit(&cx.empty[stage]);
                }
                __syncwarp();
                if (++stage == NSTAGE) { stage = 0; phase ^= 1u; }
              }
            }
          }
          if (elect_one()) tc_commit(&cx.dfull[ts]);          // one completion per contraction and tile slot
          __syncwarp();
        }
      }
    }
  }
  else if (warp == NEW + 2) {
    // ===== record prefetcher: asks L2 for the activation records of layer l one contraction before the reverse pass of
    // the tile reads them =====
    if (lane == 0) {
      uint32_t xpar0 = 0, xpar1 = 0;
      for (int i0 = 0; i0 < n_mine; i0 += HNT) {
        const int nts = (i0 + 1 < n_mine) ? HNT : 1;
        for (int step = 1; step <= nsteps; ++step) {
          for (int ts = 0; ts < nts; ++ts) {
            if (ts) { mbar_wait(&cx.xready[1], xpar1); xpar1 ^= 1u; } else { mbar_wait(&cx.xready[0], xpar0); xpar0 ^= 1u; }
            const int l = 2 * L - 2 - step;                // step L + j (j = 0 .. L-2) -> layer L-2-j
            if (p.rec_prefetch && step >= L && l >= 0) {
              const int t = bx + (i0 + ts) * gx;
              int sgi = 0;
              while (sgi + 1 < kp.n_seg && t >= kp.seg_tile_begin[sgi + 1]) ++sgi;
              const int tile = t - kp.seg_tile_begin[sgi];
              const int head = kp.seg[sgi].head;
              const int jk = (head == HEAD_POISSON_DOMAIN) ? JET_LAP : (head == HEAD_BURGERS_DOMAIN) ? JET_XX : (head == HEAD_HYPER_DOMAIN) ? JET_G : JET_V;
              const size_t per_layer = (size_t)rec_fields_points_hvp(jk) * W;       // floats of one (tile, layer) record block
              const float* rec_tile = p.rec + (size_t)task * p.rec_task_stride + p.rec_seg_base[sgi] + (size_t)tile * L * per_layer;
              bulk_prefetch_l2(rec_tile + (size_t)l * per_layer, (uint32_t)(per_layer * sizeof(float)));
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
  if (tid < MAX_SEG) kp.partial_terms[(size_t)cta * MAX_SEG + tid] = cx.terms[tid];
}

}  // namespace tc
}  // namespace mpde
