// tc_hvp.cuh -- tensor-core path of the Hessian-vector product (MODE_HVP): tangent-of-(forward + adjoint) of the
// fused SIREN jet kernel (reference: the second-order pass of src/nets/maml.py:195-202).
//
// Same machinery as tc_pde_kernel (tc_kernels.cuh) with a tangent stream next to the primal one:
//   forward   z    = W^T h,      zdot    = W^T hdot    + Wdot^T h
//   reverse   abar = W zbar,     abardot = W zbardot   + Wdot zbar
// A tile is 64 jet rows (16 points x 4 channels, or 64 value points): the primal rows X (h / zbar) and the tangent rows
// Xd (hdot / zbardot) are two 64-row blocks of one MN-major B operand (64 KB each in shared memory), so W * [X | Xd] is
// a single N = 128 MMA that reads the weight stage once, and Wdot * X an N = 64 one.  Three accumulator sets live in
// TMEM (W * primal, W * tangent, Wdot * primal: the two tangent products carry different block shifts and are added in
// the epilogue), and the weight ring streams, per M tile, the stages of W followed by the stages of Wdot.
// Spills for tc_wgrad_kernel: records (saved fields + zdot) of every layer, zbar_l and zbardot_l.
#pragma once
#include "tc_kernels.cuh"

namespace mpde {
namespace tc {

constexpr int HR = 64;                   // jet rows per HVP tile (MMA N)

template <int W> struct HvGeo {
  static constexpr int NMT = W / 128, KSL = W / 64, NEPI = W;
  // Two epilogue GROUPS of W threads (thread = neuron): group g owns rows [32 g, 32 g + 32) of the tile -- the same MMAs,
  // operand buffers and accumulators as before, twice the warps on the sin-jet algebra (round 2: an epilogue warp issued
  // one instruction every ~5 cycles, a dependent chain; 2 warps per scheduler could not hide it).  + one control warpgroup
  // (producer, MMA issuer, record prefetcher, one idle warp).  W = 256: 20 warps launch at 96 registers; the control
  // warpgroup drops to 32 and the 16 epilogue warps rise to 112 (setmaxnreg: 128 x 64 = 512 x 16).
  static constexpr int NG = 2, GR = HR / NG;            // groups, jet rows per group
  static constexpr bool REGSHIFT = (W == 256);
  static constexpr int EPI_REGS = 112, CTL_REGS = 32;
  static constexpr int NT = NG * W + 128;
  static constexpr int SLAB = HR * 128;                 // [64 rows][64 k] fp16
  static constexpr int XBYTES = 2 * KSL * SLAB;         // hi slabs then lo slabs (one operand)
  static constexpr int RING = NSTAGE * STAGE_BYTES;
  // misc: xs[2][HR], ybuf[HR][2], ydbuf[HR][2], red[HR] f32; cmax[NG][2][8] u32; terms[4] f32; barriers; tmem slot
  static constexpr int MISC = 2 * HR * 4 + 2 * HR * 4 + 2 * HR * 4 + HR * 4 + NG * 64 + 16 + (2 * NSTAGE + 3) * 8 + 16;
  static constexpr int SMEM = 2 * XBYTES + RING + MISC;
  // TMEM columns: per M tile mt the pair [W*primal (64) | W*tangent + Wdot*primal (64)] at mt * 128: one N = 128 MMA over
  // the two row blocks X | Xd, which share the A operand, then the N = 64 MMAs of Wdot * X accumulate into the tangent columns
  static constexpr int TMEM_COLS = 2 * NMT * HR;        // 128 or 256
};

template <int W> struct HvCtx {
  typedef HvGeo<W> G;
  uint8_t* xbuf; uint8_t* xdbuf; uint8_t* ring;
  float* xs; float* ybuf; float* ydbuf; float* red; uint32_t* cmax; float* terms;
  uint64_t* full; uint64_t* empty; uint64_t* xready; uint64_t* dfull;
  uint32_t* tmem_slot;
  TC_DEV void carve() {
    uint8_t* s = tc_smem_raw;
    xbuf = s; xdbuf = s + G::XBYTES; ring = s + 2 * G::XBYTES;
    xs = reinterpret_cast<float*>(ring + G::RING);
    ybuf = xs + 2 * HR; ydbuf = ybuf + 2 * HR; red = ydbuf + 2 * HR;
    cmax = reinterpret_cast<uint32_t*>(red + HR);
    terms = reinterpret_cast<float*>(cmax + 16 * G::NG);
    full = reinterpret_cast<uint64_t*>(terms + 4);
    empty = full + NSTAGE; xready = empty + NSTAGE; dfull = xready + 1;
    tmem_slot = reinterpret_cast<uint32_t*>(dfull + 2);
  }
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

  // MN-major B operands (see TcTile in tc_kernels.cuh): one block of 64 rows, [k = neuron][128 B = 64 rows]; 8 consecutive
  // rows of a neuron are one 16-byte chunk, so a point's 4 channels (or 8 rows) are stored with one vector store
  static constexpr uint32_t LO_OFF = (uint32_t)W * 128u;                 // lo parts behind the hi parts
  static TC_DEV uint32_t xoff(int r, int n) {
    return (uint32_t)(n >> 3) * 1024u + (uint32_t)(n & 7) * 128u + ((((uint32_t)r >> 3) ^ (uint32_t)(n & 7)) << 4) +
           (uint32_t)(r & 7) * 2u;
  }
  static TC_DEV void store_x(uint8_t* xb, int r, int n, float y) {
    __half h, l;
    split_f16(y, h, l);
    const uint32_t o = xoff(r, n);
    *reinterpret_cast<__half*>(xb + o) = h;
    *reinterpret_cast<__half*>(xb + LO_OFF + o) = l;
  }
  static TC_DEV void split2(float a, float b, uint32_t& hi, uint32_t& lo) {
    const __half2 h = __floats2half2_rn(a, b);
    const float2 f = __half22float2(h);
    const __half2 l = __floats2half2_rn(a - f.x, b - f.y);
    hi = *reinterpret_cast<const uint32_t*>(&h);
    lo = *reinterpret_cast<const uint32_t*>(&l);
  }
  static TC_DEV void store_x4(uint8_t* xb, int r0, int n, const float (&y)[4]) {
    uint2 hv, lv;
    split2(y[0], y[1], hv.x, lv.x);
    split2(y[2], y[3], hv.y, lv.y);
    const uint32_t o = xoff(r0, n);
    *reinterpret_cast<uint2*>(xb + o) = hv;
    *reinterpret_cast<uint2*>(xb + LO_OFF + o) = lv;
  }
  static TC_DEV void store_x8(uint8_t* xb, int r0, int n, const float (&y)[8]) {
    uint4 hv, lv;
    split2(y[0], y[1], hv.x, lv.x);
    split2(y[2], y[3], hv.y, lv.y);
    split2(y[4], y[5], hv.z, lv.z);
    split2(y[6], y[7], hv.w, lv.w);
    const uint32_t o = xoff(r0, n);
    *reinterpret_cast<uint4*>(xb + o) = hv;
    *reinterpret_cast<uint4*>(xb + LO_OFF + o) = lv;
  }
  // ---- in-place staging (round 2) ------------------------------------------------------------------------
  // A neuron's part of an operand is 128 B of hi parts + 128 B of lo parts = 64 fp32: exactly its 64 rows in fp32.  The
  // epilogue therefore parks the fp32 values of a tile IN the operand rows it owns (rows 8j..8j+3 in the 16-byte chunk j
  // of the hi part, rows 8j+4..8j+7 in chunk j of the lo part), all threads agree on the EXACT channel maxima, and
  // each thread converts its own chunks to fp16 hi/lo in place.  One pass over tensor memory (the bound pass of round 1
  // is gone), one sincos per element, and no re-read of zbar from global memory in the reverse pass.
  static TC_DEV uint32_t chunk_off(int j, int n) {
    return (uint32_t)(n >> 3) * 1024u + (uint32_t)(n & 7) * 128u + (((uint32_t)j ^ (uint32_t)(n & 7)) << 4);
  }
  // 4 consecutive rows r0..r0+3 (r0 % 4 == 0) of neuron n, fp32
  static TC_DEV void stage4(uint8_t* xb, int r0, int n, float a, float b, float c, float d) {
    *reinterpret_cast<float4*>(xb + ((r0 & 4) ? LO_OFF : 0u) + chunk_off(r0 >> 3, n)) = make_float4(a, b, c, d);
  }
  // all 64 staged rows of neuron n -> fp16 hi / lo operand, scaled by ys[channel]
  static TC_DEV void convert_rows(uint8_t* xb, int n, int grp, const float* ys) {
    MPDE_UNROLL
    for (int jj = 0; jj < G::GR / 8; ++jj) {
      const int j = grp * (G::GR / 8) + jj;             // this group's rows: chunks 4 g .. 4 g + 3
      const uint32_t o = chunk_off(j, n);
      const float4 a = *reinterpret_cast<const float4*>(xb + o);
      const float4 b = *reinterpret_cast<const float4*>(xb + LO_OFF + o);
      uint4 hv, lv;
      if (VAL) {
        split2(a.x * ys[0], a.y * ys[0], hv.x, lv.x); split2(a.z * ys[0], a.w * ys[0], hv.y, lv.y);
        split2(b.x * ys[0], b.y * ys[0], hv.z, lv.z); split2(b.z * ys[0], b.w * ys[0], hv.w, lv.w);
      } else {
        split2(a.x * ys[0], a.y * ys[1], hv.x, lv.x); split2(a.z * ys[2], a.w * ys[3], hv.y, lv.y);
        split2(b.x * ys[0], b.y * ys[1], hv.z, lv.z); split2(b.z * ys[2], b.w * ys[3], hv.w, lv.w);
      }
      *reinterpret_cast<uint4*>(xb + o) = hv;
      *reinterpret_cast<uint4*>(xb + LO_OFF + o) = lv;
    }
  }
  // fold the channel maxima (mx[0..3] primal, mx[4..7] tangent), barrier among the group's warps, read the shifts.
  // FWD: the primal value channel is sin: |a_0| <= 1, shift at most 14.
  // Shared exponent (round 2): W * tangent and Wdot * primal are accumulated in ONE set of accumulator columns, so their
  // scales must agree: sh_tangent + sw = sh_primal + swd, i.e. sh_tangent = sh_primal + delta with delta = swd - sw of the
  // contraction that consumes this operand.  Per channel e = min(max-safe tangent shift, max-safe primal shift + delta);
  // tangent shift = e, primal shift = e - delta: never above a max-safe shift (no fp16 overflow); the operand that is
  // scaled below its own optimum belongs to the term that is the smaller one at the common scale.
  template <bool FWD>
  static TC_DEV void agree_shifts(HvCtx<W>& cx, int grp, int& seq, int delta, const float (&mx)[8], float (&ys)[8], int (&sh)[8],
                                  int tid, int lane) {
    uint32_t* cm = cx.cmax + grp * 16 + (seq & 1) * 8;
    MPDE_UNROLL
    for (int c = 0; c < 8; ++c) {
      const bool used = VAL ? (c == 4 || (!FWD && c == 0)) : ((c & 3) < C && !(FWD && c == 0));
      if (used) fold_max(cm + c, mx[c], lane);
    }
    epi_sync<G::NEPI>(grp);                             // the group's own barrier: its rows carry their own shifts
    MPDE_UNROLL
    for (int c = 0; c < 4; ++c) {
      const uint32_t mp = cm[c], md = cm[4 + c];
      // an all-zero block puts no constraint on the common exponent
      const int sp = (FWD && c == 0) ? 14 : (mp ? shift_for_max(__uint_as_float(mp)) : 100);
      const int sd = md ? shift_for_max(__uint_as_float(md)) : 100;
      int e = sp + delta < sd ? sp + delta : sd;
      e = e < -90 ? -90 : (e > 90 ? 90 : e);
      int ep = e - delta;
      ep = ep < -90 ? -90 : (ep > 90 ? 90 : ep);
      sh[4 + c] = ep + delta;                           // (clamped consistently: tangent = primal + delta)
      sh[c] = ep;
    }
    MPDE_UNROLL
    for (int c = 0; c < 8; ++c) ys[c] = pow2i(sh[c]);
    if (tid == 0) {
      uint32_t* other = cx.cmax + grp * 16 + ((seq & 1) ^ 1) * 8;
      MPDE_UNROLL
      for (int c = 0; c < 8; ++c) other[c] = 0u;
    }
    ++seq;
  }
  // `qs`: the point's index inside its unrolled group (compile-time at every call site; pl % 4 == qs % 4)
  static TC_DEV void stage_point(uint8_t* xb, uint8_t* xdb, int pl, int qs, int n, const float* a, const float* ad, float (&mx)[8],
                                 float (&va)[4], float (&vd)[4]) {
    if (VAL) {
      va[qs & 3] = a[0]; vd[qs & 3] = ad[0];
      mx[0] = fmaxf(mx[0], fabsf(a[0])); mx[4] = fmaxf(mx[4], fabsf(ad[0]));
      if ((qs & 3) == 3) {
        stage4(xb, pl - 3, n, va[0], va[1], va[2], va[3]);
        stage4(xdb, pl - 3, n, vd[0], vd[1], vd[2], vd[3]);
      }
    } else {
      MPDE_UNROLL
      for (int c = 0; c < C; ++c) { mx[c] = fmaxf(mx[c], fabsf(a[c])); mx[4 + c] = fmaxf(mx[4 + c], fabsf(ad[c])); }
      stage4(xb, 4 * pl, n, a[0], a[1], a[2], C > 3 ? a[3] : 0.f);
      stage4(xdb, 4 * pl, n, ad[0], ad[1], ad[2], C > 3 ? ad[3] : 0.f);
    }
  }

  // barrier among ALL epilogue threads (both groups): named barrier 3
  static TC_DEV void all_sync() { asm volatile("bar.sync 3, %0;" ::"n"(G::NG * G::NEPI) : "memory"); }

  // ---- the epilogue side of one tile (both groups of W epilogue threads) --------------------------------
  static TC_DEV void run(const HvParams& hp, HvCtx<W>& cx, const SegDesc& sg, int seg_id, int task, int tile,
                         uint32_t tbase, int& seq, uint32_t& dpar, const int* wshift, const int* wshift_d) {
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
    const uint32_t tl = tmem_addr(tbase, 32u * (uint32_t)(warp & 3), (uint32_t)(grp * G::GR));
    const uint32_t c0 = (uint32_t)mt * 2u * HR, c1 = c0 + HR;   // accumulator columns: primal, tangent
    float* xs = cx.xs;
    float* recp = p.rec + (size_t)task * p.rec_task_stride + p.rec_seg_base[seg_id] + (size_t)tile * L * (NFT * PTS) * W +
             (size_t)(n >> 7) * (PTS * 128) + (n & 127);   // record layout [layer][field][k tile][point][128 neurons]
    constexpr int TCH = 8;                              // accumulator columns per chunk (2 jet points: 112 registers)
    constexpr int PPC = TCH / CP;                       // points per chunk
    constexpr int PG = G::GR / CP;                      // points of one group
    const int pg0 = grp * PG;                           // first point of this group

    for (int i = tid; i < PTS; i += G::NG * G::NEPI) {
      const int gp = pt0 + i;
      float x0 = 0.f, x1 = 0.f;
      if (gp < sg.n) { x0 = pts[2 * (size_t)gp]; x1 = pts[2 * (size_t)gp + 1]; }
      xs[i] = x0; xs[HR + i] = x1;
    }
    // ---- layer 0 (in_dim = 2): direct ----
    int sh[8];
    float ys[8];
    {
      const float w00 = mpde_ldg(th + n), w01 = mpde_ldg(th + W + n), b0 = mpde_ldg(th + 2 * W + n);
      const float d00 = mpde_ldg(thd + n), d01 = mpde_ldg(thd + W + n), db0 = mpde_ldg(thd + 2 * W + n);
      all_sync();                                       // xs visible (written by threads of both groups)
      float mx[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, va[4], vd[4];
      MPDE_NOUNROLL
      for (int pl0 = pg0; pl0 < pg0 + PG; pl0 += 4) {
        MPDE_UNROLL
        for (int q = 0; q < 4; ++q) {
          const int pl = pl0 + q;
          float z[4], zd[4], a[4], ad[4], sv[5];
          z[0] = fmaf(w00, xs[pl], fmaf(w01, xs[HR + pl], b0)); z[1] = w00; z[2] = w01; z[3] = 0.f;
          zd[0] = fmaf(d00, xs[pl], fmaf(d01, xs[HR + pl], db0)); zd[1] = d00; zd[2] = d01; zd[3] = 0.f;
          jet_act_fwd<JK>(om, z, a, sv);
          jet_act_tan<JK>(om, sv, zd, ad);
          MPDE_UNROLL
          for (int f = 0; f < NF; ++f) recp[(size_t)(f) * (PTS * W) + (pl) * 128] = sv[f];
          MPDE_UNROLL
          for (int c = 0; c < C; ++c) recp[(size_t)((NF + c)) * (PTS * W) + (pl) * 128] = zd[c];
          stage_point(cx.xbuf, cx.xdbuf, pl, q, n, a, ad, mx, va, vd);
        }
      }
      agree_shifts<true>(cx, grp, seq, wshift_d[1] - wshift[1], mx, ys, sh, n, lane);
      convert_rows(cx.xbuf, n, grp, ys);
      convert_rows(cx.xdbuf, n, grp, ys + 4);
    }
    fence_proxy_async_smem();
    tc_fence_before();
    mbar_arrive(cx.xready);

    // ---- hidden layers ----
    for (int l = 1; l < L; ++l) {
      const float bias = mpde_ldg(th + ph_off_layer(W, l) + 2 * W * W + n);
      const float biasd = mpde_ldg(thd + ph_off_layer(W, l) + 2 * W * W + n);
      float zs[4], zs1[4];
      {
        const int sw = wshift[l];
        for (int c = 0; c < 4; ++c) { zs[c] = pow2i(-sh[c] - sw); zs1[c] = pow2i(-sh[4 + c] - sw); }
      }
      mbar_wait(cx.dfull, dpar); dpar ^= 1u;             // every MMA of this layer has completed: X / Xd may be rewritten
      tc_fence_after();
      float mx[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, va[4], vd[4];
      MPDE_NOUNROLL
      for (int ch = 0; ch < G::GR / TCH; ++ch) {
        uint32_t v0[TCH], v1[TCH];
        tmem_ld8(tl + c0 + ch * TCH, v0); tmem_ld8(tl + c1 + ch * TCH, v1);
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
          for (int f = 0; f < NF; ++f) recp[(size_t)((l * NFT + f)) * (PTS * W) + (pl) * 128] = sv[f];
          MPDE_UNROLL
          for (int c = 0; c < C; ++c) recp[(size_t)((l * NFT + NF + c)) * (PTS * W) + (pl) * 128] = zd[c];
          stage_point(cx.xbuf, cx.xdbuf, pl, q, n, a, ad, mx, va, vd);
        }
      }
      agree_shifts<true>(cx, grp, seq, wshift_d[l + 1] - wshift[l + 1], mx, ys, sh, n, lane);
      convert_rows(cx.xbuf, n, grp, ys);
      convert_rows(cx.xdbuf, n, grp, ys + 4);
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(cx.xready);
    }

    // ---- output layer: lanes 0..OUT-1 of the first M tile hold y, ydot ----
    mbar_wait(cx.dfull, dpar); dpar ^= 1u;
    tc_fence_after();
    if (warp == 0) {
      float zs[4], zs1[4];
      {
        const int sw = wshift[L];
        for (int c = 0; c < 4; ++c) { zs[c] = pow2i(-sh[c] - sw); zs1[c] = pow2i(-sh[4 + c] - sw); }
      }
      // each group's first warp: the output rows of ITS rows (descaled with the group's own shifts)
      MPDE_NOUNROLL
      for (int ch = 0; ch < G::GR / TCH; ++ch) {
        uint32_t v0[TCH], v1[TCH];
        tmem_ld8(tl + c0 + ch * TCH, v0); tmem_ld8(tl + c1 + ch * TCH, v1);
        tmem_ld_wait();
        if (lane < OUT) {
          MPDE_UNROLL
          for (int q = 0; q < TCH; ++q) {
            const int c = VAL ? 0 : (q & 3);
            const int r = grp * G::GR + ch * TCH + q;
            cx.ybuf[r * 2 + lane] = __uint_as_float(v0[q]) * zs[c];
            cx.ydbuf[r * 2 + lane] = __uint_as_float(v1[q]) * zs1[c];
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
          float y = cx.ybuf[(CP * pl + c) * 2 + o], yd = cx.ydbuf[(CP * pl + c) * 2 + o];
          if (c == 0) { y += mpde_ldg(bL + o); yd += mpde_ldg(bLd + o); }
          u[o * C + c] = y; ud[o * C + c] = yd;
        }
      const bool valid = (pt0 + pl) < sg.n;
      float res = 0.f;
      float val = head_eval<HEAD>(u, xs[pl], xs[HR + pl], prm, sg.coef, ub, &res);
      head_tangent<HEAD>(u, ud, xs[pl], xs[HR + pl], prm, sg.coef, ubd);
      if (!valid) val = 0.f;
      cx.red[pl] = val;
      MPDE_UNROLL
      for (int o = 0; o < OUT; ++o)
        MPDE_UNROLL
        for (int c = 0; c < CP; ++c) {
          cx.ybuf[(CP * pl + c) * 2 + o] = (valid && c < C) ? ub[o * C + c] : 0.f;
          cx.ydbuf[(CP * pl + c) * 2 + o] = (valid && c < C) ? ubd[o * C + c] : 0.f;
        }
    }
    all_sync();
    if (tid < 32) {
      float s = 0.f;
      MPDE_UNROLL
      for (int j = 0; j < (PTS + 31) / 32; ++j) s += (lane + 32 * j < PTS) ? cx.red[lane + 32 * j] : 0.f;
      MPDE_UNROLL
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (lane == 0) cx.terms[seg_id] += s;
    }

    // ==================== reverse pass and its tangent ====================
    // one instantiation per layer kind (l = L-1: seeds, no accumulators / 0 < l < L-1 / l = 0: no X to build): no runtime
    // branch inside the per-point loops (round 2: the in-loop `l == ...` tests cost 11 % of the GRAD kernel)
    int shz[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    reverse_layer<RK_LAST>(hp, cx, sg, seg_id, task, tile, tbase, L - 1, recp, shz, seq, dpar, wshift, wshift_d);
    for (int l = L - 2; l >= 1; --l)
      reverse_layer<RK_MID>(hp, cx, sg, seg_id, task, tile, tbase, l, recp, shz, seq, dpar, wshift, wshift_d);
    reverse_layer<RK_FIRST>(hp, cx, sg, seg_id, task, tile, tbase, 0, recp, shz, seq, dpar, wshift, wshift_d);
  }

  enum { RK_LAST = 0, RK_MID = 1, RK_FIRST = 2 };
  template <int RK>
  static TC_DEV void reverse_layer(const HvParams& hp, HvCtx<W>& cx, const SegDesc& sg, int seg_id, int task, int tile,
                                   uint32_t tbase, int l, const float* recp, int (&shz)[8], int& seq, uint32_t& dpar,
                                   const int* wshift, const int* wshift_d) {
    constexpr bool LAST = (RK == RK_LAST), FIRST = (RK == RK_FIRST);
    const TcParams& p = hp.t;
    const KParams& kp = p.k;
    const int tid = threadIdx.x, lane = tid & 31;
    const int hwarp = warp_idx_uniform();
    const int grp = hwarp / (G::NEPI / 32);
    const int n = tid - grp * G::NEPI, mt = n >> 7;
    const int warp = hwarp - grp * (G::NEPI / 32);
    const float om = kp.omega;
    const int L = kp.L;
    const float* th = kp.th + (size_t)task * kp.th_stride;
    const float* thd = kp.thd + (size_t)task * kp.thd_stride;
    const uint32_t tl = tmem_addr(tbase, 32u * (uint32_t)(warp & 3), (uint32_t)(grp * G::GR));
    const uint32_t c0 = (uint32_t)mt * 2u * HR, c1 = c0 + HR;   // accumulator columns: primal, tangent
    float* xs = cx.xs;
    constexpr int TCH = 8, PPC = TCH / CP, PG = G::GR / CP;
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
        for (int pl = 0; pl < PTS; ++pl) { g += cx.ybuf[(CP * pl) * 2 + tid]; h += cx.ydbuf[(CP * pl) * 2 + tid]; }
        psh[ps_off_wl(W, L) + W * OUT + tid] += h;
        if (wg) psg[ps_off_wl(W, L) + W * OUT + tid] += g;
      }
    }
    const float* rl = recp + (size_t)l * NFT * (PTS * W);   // this layer's records: [field][k tile][point][128]
    float zs[4] = {0.f, 0.f, 0.f, 0.f}, zs1[4] = {0.f, 0.f, 0.f, 0.f};
    // the first chunk's record loads are issued before the accumulator wait (they do not depend on it)
    float svb[PPC][NF + C];
    MPDE_UNROLL
    for (int q = 0; q < PPC; ++q)
      MPDE_UNROLL
      for (int f = 0; f < NF + C; ++f) svb[q][f] = rl[f * (PTS * W) + (pg0 + q) * 128];
    if (!LAST) {
      const int sw = wshift[l + 1];
      for (int c = 0; c < 4; ++c) { zs[c] = pow2i(-shz[c] - sw); zs1[c] = pow2i(-shz[4 + c] - sw); }
      mbar_wait(cx.dfull, dpar); dpar ^= 1u;           // all adjoint MMAs that read X / Xd have completed
      tc_fence_after();
    }
    float mx[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, va[4], vd[4];
    float bsum = 0.f, bsumd = 0.f, g0 = 0.f, g1 = 0.f, h0 = 0.f, h1 = 0.f, gw[OUT], hw[OUT];
    MPDE_UNROLL
    for (int o = 0; o < OUT; ++o) { gw[o] = 0.f; hw[o] = 0.f; }
    const size_t zoff = (size_t)task * p.zsp_task_stride +
                        ((size_t)(kp.seg_tile_begin[seg_id] + tile) * (L - 1) + (FIRST ? 0 : l - 1)) * HR * W + n;
    float* zl = p.zsp + zoff;                              // only dereferenced for l >= 1
    float* zld = hp.zspd + zoff;
    MPDE_NOUNROLL
    for (int ch = 0; ch < G::GR / TCH; ++ch) {
      uint32_t v0[TCH], v1[TCH];
      if (!LAST) {
        tmem_ld8(tl + c0 + ch * TCH, v0); tmem_ld8(tl + c1 + ch * TCH, v1);
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
            for (int o = 0; o < OUT; ++o) {
              const float yb = cx.ybuf[(CP * pl + c) * 2 + o], ybd = cx.ydbuf[(CP * pl + c) * 2 + o];
              a_ = fmaf(wl[o], yb, a_);
              ad_ = fmaf(wl[o], ybd, fmaf(wld[o], yb, ad_));
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
        for (int f = 0; f < NF + C; ++f) svb[q][f] = rl[f * (PTS * W) + (pg0 + chn * PPC + q) * 128];
        jet_act_bwd<JK>(om, sv, ab, zb);
        jet_act_bwd_tan<JK>(om, sv, zd, ab, abd, zbd);
        bsum += zb[0]; bsumd += zbd[0];
        if (LAST) {
          float a[4], ad[4];
          jet_act_from_saved<JK>(om, sv, a);
          jet_act_tan<JK>(om, sv, zd, ad);
          MPDE_UNROLL
          for (int o = 0; o < OUT; ++o)
            MPDE_UNROLL
            for (int c = 0; c < C; ++c) {
              const float yb = cx.ybuf[(CP * pl + c) * 2 + o], ybd = cx.ydbuf[(CP * pl + c) * 2 + o];
              gw[o] = fmaf(a[c], yb, gw[o]);
              hw[o] = fmaf(ad[c], yb, fmaf(a[c], ybd, hw[o]));
            }
        }
        if (FIRST) {
          if (VAL) {
            g0 = fmaf(xs[pl], zb[0], g0); g1 = fmaf(xs[HR + pl], zb[0], g1);
            h0 = fmaf(xs[pl], zbd[0], h0); h1 = fmaf(xs[HR + pl], zbd[0], h1);
          } else {
            g0 += fmaf(xs[pl], zb[0], zb[1]); g1 += fmaf(xs[HR + pl], zb[0], zb[2]);
            h0 += fmaf(xs[pl], zbd[0], zbd[1]); h1 += fmaf(xs[HR + pl], zbd[0], zbd[2]);
          }
        } else {
          // fp32 spill for tc_wgrad_kernel (written once, never read back here) + the staged copy for the next contraction
          MPDE_UNROLL
          for (int c = 0; c < CP; ++c) {
            zl[(size_t)(CP * pl + c) * W] = c < C ? zb[c] : 0.f;
            zld[(size_t)(CP * pl + c) * W] = c < C ? zbd[c] : 0.f;
          }
          stage_point(cx.xbuf, cx.xdbuf, pl, q, n, zb, zbd, mx, va, vd);
        }
      }
    }
    psh[ps_off_bias(W, l) + n] += bsumd;
    if (wg) psg[ps_off_bias(W, l) + n] += bsum;
    if (LAST) {
      MPDE_UNROLL
      for (int o = 0; o < OUT; ++o) {
        psh[ps_off_wl(W, L) + n * OUT + o] += hw[o];
        if (wg) psg[ps_off_wl(W, L) + n * OUT + o] += gw[o];
      }
    }
    if (FIRST) {
      psh[n] += h0; psh[W + n] += h1;
      if (wg) { psg[n] += g0; psg[W + n] += g1; }
      all_sync();                                       // xs / ybuf are rewritten by the next tile
      return;
    }
    float ysz[8];
    agree_shifts<false>(cx, grp, seq, wshift_d[l] - wshift[l], mx, ysz, shz, n, lane);
    convert_rows(cx.xbuf, n, grp, ysz);
    convert_rows(cx.xdbuf, n, grp, ysz + 4);
    fence_proxy_async_smem();
    tc_fence_before();
    mbar_arrive(cx.xready);
  }
};

// ---------------------------------------------------------------------------------------------------------
// kernel: grid = (ctas_per_task, T)
// ---------------------------------------------------------------------------------------------------------
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

  if (tid == 0) {
    if (smem_u32(tc_smem_raw) & 1023u) { printf("tc_hvp_kernel: dynamic shared memory is not 1024-byte aligned\n"); __trap(); }
    for (int i = 0; i < NSTAGE; ++i) { mbar_init(&cx.full[i], 1); mbar_init(&cx.empty[i], 1); }
    mbar_init(cx.xready, G::NG * G::NEPI);
    mbar_init(cx.dfull, 1);
    mbar_fence_init();
    for (int i = 0; i < 16 * G::NG; ++i) cx.cmax[i] = 0u;
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

  constexpr int NEW = G::NG * G::NEPI / 32;                  // epilogue warps (both groups)
  if (G::REGSHIFT) {
    if (warp < NEW) setmaxnreg_inc<G::EPI_REGS>(); else setmaxnreg_dec<G::CTL_REGS>();
  }
  if (warp < NEW) {
    int seq = 0; uint32_t dpar = 0;
    const int* wshift = p.wshift + (size_t)task * p.wshift_stride;
    const int* wshift_d = hp.wshift_d + (size_t)task * hp.wshift_d_stride;
    for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
      int s = 0;
      while (s + 1 < kp.n_seg && t >= kp.seg_tile_begin[s + 1]) ++s;
      const int tile = t - kp.seg_tile_begin[s];
      const SegDesc& sg = kp.seg[s];
      if (PDE == PDE_POISSON) {
        if (sg.head == HEAD_POISSON_DOMAIN) HvTile<W, HEAD_POISSON_DOMAIN>::run(hp, cx, sg, s, task, tile, tbase, seq, dpar, wshift, wshift_d);
        else HvTile<W, HEAD_POISSON_BOUNDARY>::run(hp, cx, sg, s, task, tile, tbase, seq, dpar, wshift, wshift_d);
      } else if (PDE == PDE_BURGERS) {
        if (sg.head == HEAD_BURGERS_DOMAIN) HvTile<W, HEAD_BURGERS_DOMAIN>::run(hp, cx, sg, s, task, tile, tbase, seq, dpar, wshift, wshift_d);
        else HvTile<W, HEAD_BURGERS_IC>::run(hp, cx, sg, s, task, tile, tbase, seq, dpar, wshift, wshift_d);
      } else {
        if (sg.head == HEAD_HYPER_DOMAIN) HvTile<W, HEAD_HYPER_DOMAIN>::run(hp, cx, sg, s, task, tile, tbase, seq, dpar, wshift, wshift_d);
        else if (sg.head == HEAD_HYPER_TOP) HvTile<W, HEAD_HYPER_TOP>::run(hp, cx, sg, s, task, tile, tbase, seq, dpar, wshift, wshift_d);
        else HvTile<W, HEAD_HYPER_BOTTOM>::run(hp, cx, sg, s, task, tile, tbase, seq, dpar, wshift, wshift_d);
      }
    }
  } else if (warp == NEW) {
    // ===== producer: per contraction and M tile, the stages of W then the stages of Wdot =====
    if (lane == 0) {
      const uint8_t* img = p.wimg + (size_t)task * p.wimg_stride;
      const uint8_t* imgd = hp.wimg_d + (size_t)task * hp.wimg_d_stride;
      uint32_t stage = 0, phase = 0;
      for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        for (int step = 1; step <= nsteps; ++step) {
          const int nmt = (step != L) ? G::NMT : 1;
          // first stage of this contraction in the image: hidden layer `step`, output layer, adjoint copies (L-1 .. 1)
          const int base = (step < L) ? (step - 1) * PER : (step == L ? (L - 1) * PER : nst_fwd + (step - L - 1) * PER);
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
    uint32_t stage = 0, phase = 0, xpar = 0;
    for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
      for (int step = 1; step <= nsteps; ++step) {
        mbar_wait(cx.xready, xpar); xpar ^= 1u;
        tc_fence_after();
        const int nmt = (step != L) ? G::NMT : 1;
        for (int mt = 0; mt < nmt; ++mt) {
          const uint32_t d01 = tbase + (uint32_t)mt * 2u * HR, d2 = d01 + HR;   // Wdot * X accumulates into the tangent columns
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
                  const uint32_t acc = (s | kk) ? 1u : 0u;
                  if (which == 0) {
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
        if (elect_one()) tc_commit(cx.dfull);            // one completion per contraction step: all M tiles
        __syncwarp();
      }
    }
  }
  else if (warp == NEW + 2) {
    // ===== record prefetcher: the reverse pass reads the activation records of layer l one full contraction step after
    // this thread asked L2 for them (they were written ~100 us earlier and the per-SM share of L2 does not hold a
    // whole tile's spills: without the prefetch every batch of record loads pays an HBM round trip) =====
    if (lane == 0) {
      uint32_t xpar = 0;
      for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        int sgi = 0;
        while (sgi + 1 < kp.n_seg && t >= kp.seg_tile_begin[sgi + 1]) ++sgi;
        const int tile = t - kp.seg_tile_begin[sgi];
        const int head = kp.seg[sgi].head;
        const int jk = (head == HEAD_POISSON_DOMAIN) ? JET_LAP : (head == HEAD_BURGERS_DOMAIN) ? JET_XX : (head == HEAD_HYPER_DOMAIN) ? JET_G : JET_V;
        const size_t per_layer = (size_t)rec_fields_points_hvp(jk) * W;       // floats of one (tile, layer) record block
        const float* rec_tile = p.rec + (size_t)task * p.rec_task_stride + p.rec_seg_base[sgi] + (size_t)tile * L * per_layer;
        for (int step = 1; step <= nsteps; ++step) {
          mbar_wait(cx.xready, xpar); xpar ^= 1u;
          const int l = 2 * L - 2 - step;                // step L + j (j = 0 .. L-2) -> layer L-2-j
          if (step >= L && l >= 0) bulk_prefetch_l2(rec_tile + (size_t)l * per_layer, (uint32_t)(per_layer * sizeof(float)));
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
