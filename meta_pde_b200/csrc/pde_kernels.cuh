// pde_kernels.cuh -- fused SIREN forward + Taylor-jet + PDE loss, its adjoint
// (parameter gradient) and the tangent of that adjoint (Hessian-vector product).
//
// One CTA = one tile of collocation points of ONE task.  Per hidden layer the
// tile does a [rows x W] x [W x W] contraction on the FP32 FMA pipe with the
// jet rows of the tile resident in shared memory and the (per-task) weights
// streamed in k-slabs through a cp.async double buffer; the sin-jet algebra,
// the PDE head and every adjoint/tangent rule run in registers between the
// contractions.  Nothing but the per-layer (sin, cos, p, q) records leaves the
// SM between forward and backward, and those go to a per-CTA scratch that
// stays L2 resident.
//
// Replaces (reference, /root/reference): nf_apply under vmap(jacfwd(jacrev)) /
// jacfwd(jacfwd) / jacfwd (src/nets/field.py:47-72,146-199), the three
// loss_fn (src/poisson/poisson_common.py:26-39, src/burgers/td_burgers_common.py:47-91,
// src/elasticity/hyper_elasticity_common.py:23-86), jax.value_and_grad of them
// (src/nets/maml.py:66-70) and the second-order pass (src/nets/maml.py:195-202).
//
// Thread geometry (W = layer width, NTY row groups):
//   NTX = W/4 column groups, NT = NTX*NTY threads, NP = 4*NTY points per tile.
//   A thread owns 4 points x (all CR row blocks) x 4 neurons n_j = tx + NTX*j.
//   Row blocks are the jet channels (CR = C) or, for value-only sets, 4
//   independent groups of NP points (CR = 4).
//   Shared activation buffers are [k][RS], RS = 4*NP + 4, row r = cr*NP + pt.
#pragma once
#include "compat.cuh"
#include "jet_math.cuh"

namespace mpde {

constexpr int MAX_SEG = 4;
enum Mode { MODE_LOSS = 0, MODE_GRAD = 1, MODE_HVP = 2 };
enum PdeKind { PDE_POISSON = 0, PDE_BURGERS = 1, PDE_HYPER = 2, PDE_EVAL1 = 3, PDE_EVAL2 = 4 };

struct SegDesc {
  const float* pts;        // [T][n][2]
  long long task_stride;   // floats between tasks (0: shared)
  int n;                   // points per task
  int head;                // HeadKind
  float coef;              // loss_weight * term_scale / n_elems
  float* res_out;          // optional per-point output (MODE_LOSS): residual / psi / u
  long long res_stride;    // floats between tasks
};

struct KParams {
  int L;                   // hidden layers
  int P;                   // natural parameter count (incl. log scales)
  int PH;                  // prepped (theta-hat) floats per task
  float omega;
  int T, n_seg;
  SegDesc seg[MAX_SEG];
  int seg_tile_begin[MAX_SEG + 1];
  const float* prm; int nprm;
  const float* th;  long long th_stride;   // prepped theta-hat
  const float* thd; long long thd_stride;  // prepped direction-hat (HVP)
  float* scratch;   long long scratch_per_cta;
  float* partial_g;                        // [T][gridDim.x][P]   (effective-parameter gradient)
  float* partial_h;                        // [T][gridDim.x][P]   (effective H v)
  float* partial_terms;                    // [T][gridDim.x][MAX_SEG]
  int want_grad;                           // HVP: also accumulate hidden-layer gradients
};

template <int W_, int NTY_> struct Geo {
  static constexpr int W = W_, NTY = NTY_, NTX = W / 4, NT = NTX * NTY, NP = 4 * NTY;
  static constexpr int RS = 4 * NP + 4;
  static constexpr int KS = (W >= 256 ? 8 : 16);
  static constexpr int BUF = W * RS;          // floats per activation buffer
  static constexpr int SLAB = 2 * KS * W;     // floats, double buffer
  static constexpr int ROWS = 4 * NP;
  static constexpr int NPART = (NT / ROWS) > 0 ? (NT / ROWS) : 1;
  // misc shared: xs[2][ROWS], ybuf[ROWS][2], ydbuf[ROWS][2], wl[4*W] (wLe, wLd), terms[MAX_SEG].
  // ypart[2][NPART][ROWS][2] and red[ROWS] alias the weight slab (no contraction runs during the head).
  static constexpr int MISC = 2 * ROWS + 2 * ROWS + 2 * ROWS + 4 * W + 8;
  static_assert(2 * NPART * ROWS * 2 + ROWS <= SLAB, "head scratch must fit in the slab");
  static constexpr int nbuf(int mode) { return mode == MODE_HVP ? 3 : (mode == MODE_GRAD ? 2 : 1); }
  static constexpr size_t smem_bytes(int mode) { return sizeof(float) * (size_t)(nbuf(mode) * BUF + SLAB + MISC); }
  // prepped layout
  static constexpr int PH_LAYER = 2 * W * W + W;
  static constexpr int NAT_LAYER = W * W + W;
};

// natural (C ABI) parameter offsets
MPDE_HD int nat_off_layer(int W, int l) { return 3 * W + (l - 1) * (W * W + W); }          // l = 1..L-1: kernel then bias
MPDE_HD int nat_off_out(int W, int L) { return 3 * W + (L - 1) * (W * W + W); }           // wL [W][OUT], bL [OUT]
MPDE_HD int nat_count(int W, int L, int OUT, int log_scale) {
  return nat_off_out(W, L) + W * OUT + OUT + (log_scale ? 2 + OUT : 0);
}
// prepped offsets: w0e [2][W], b0 [W]; per hidden layer: Wp [W*W], WTp [W*W], b [W]; wLe [W][OUT], bLe [OUT]
MPDE_HD int ph_off_layer(int W, int l) { return 3 * W + (l - 1) * (2 * W * W + W); }
MPDE_HD int ph_off_out(int W, int L) { return 3 * W + (L - 1) * (2 * W * W + W); }
MPDE_HD int ph_count(int W, int L, int OUT) { return (ph_off_out(W, L) + W * OUT + OUT + 3) & ~3; }

// ---------------------------------------------------------------------------
// acc[cr][i][j] += sum_k A[k][cr*NP + ty*4 + i] * B[k][col(tx, j)]
// A: shared [W][RS].  B: global, row-major [W][W] with columns pre-permuted so
// that the thread's 4 columns are contiguous at tx*4 (prep kernel).
// ---------------------------------------------------------------------------
template <int W, int NTY, int CR>
MPDE_DEV void gemm_F(float (&acc)[CR][4][4], const float* bufA, const float* MPDE_RESTRICT gB, float* slab,
                     int tid, int tx, int ty) {
  typedef Geo<W, NTY> G;
  constexpr int KS = G::KS, NSLAB = W / KS, V4 = KS * W / 4;
  static_assert(V4 % G::NT == 0, "slab copy must divide evenly");
  {
    const float4* src = reinterpret_cast<const float4*>(gB);
    float4* dst = reinterpret_cast<float4*>(slab);
    MPDE_UNROLL
    for (int i = tid; i < V4; i += G::NT) cp_async16(dst + i, src + i);
    cp_async_commit();
  }
  MPDE_NOUNROLL
  for (int s = 0; s < NSLAB; ++s) {
    if (s + 1 < NSLAB) {
      const float4* src = reinterpret_cast<const float4*>(gB + (size_t)(s + 1) * KS * W);
      float4* dst = reinterpret_cast<float4*>(slab + ((s + 1) & 1) * KS * W);
      MPDE_UNROLL
      for (int i = tid; i < V4; i += G::NT) cp_async16(dst + i, src + i);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    const float* sb = slab + (s & 1) * KS * W + tx * 4;
    const float* sa = bufA + (size_t)(s * KS) * G::RS + ty * 4;
    MPDE_UNROLL
    for (int kk = 0; kk < KS; ++kk) {
      const float4 b = *reinterpret_cast<const float4*>(sb + kk * W);
      MPDE_UNROLL
      for (int cr = 0; cr < CR; ++cr) {
        const float4 a = *reinterpret_cast<const float4*>(sa + kk * G::RS + cr * G::NP);
        acc[cr][0][0] = fmaf(a.x, b.x, acc[cr][0][0]); acc[cr][0][1] = fmaf(a.x, b.y, acc[cr][0][1]);
        acc[cr][0][2] = fmaf(a.x, b.z, acc[cr][0][2]); acc[cr][0][3] = fmaf(a.x, b.w, acc[cr][0][3]);
        acc[cr][1][0] = fmaf(a.y, b.x, acc[cr][1][0]); acc[cr][1][1] = fmaf(a.y, b.y, acc[cr][1][1]);
        acc[cr][1][2] = fmaf(a.y, b.z, acc[cr][1][2]); acc[cr][1][3] = fmaf(a.y, b.w, acc[cr][1][3]);
        acc[cr][2][0] = fmaf(a.z, b.x, acc[cr][2][0]); acc[cr][2][1] = fmaf(a.z, b.y, acc[cr][2][1]);
        acc[cr][2][2] = fmaf(a.z, b.z, acc[cr][2][2]); acc[cr][2][3] = fmaf(a.z, b.w, acc[cr][2][3]);
        acc[cr][3][0] = fmaf(a.w, b.x, acc[cr][3][0]); acc[cr][3][1] = fmaf(a.w, b.y, acc[cr][3][1]);
        acc[cr][3][2] = fmaf(a.w, b.z, acc[cr][3][2]); acc[cr][3][3] = fmaf(a.w, b.w, acc[cr][3][3]);
      }
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------
// gout[k][n] += sum_{r < rows} H[k][r] * Z[n][r]     (weight-gradient contraction over the tile's rows)
// Owner-computes read-modify-write into this CTA's private partial buffer: deterministic, no atomics.
// ---------------------------------------------------------------------------
template <int W, int NTY>
MPDE_DEV void gemm_G(const float* bufH, const float* bufZ, float* MPDE_RESTRICT gout, int rows, int tid) {
  typedef Geo<W, NTY> G;
  constexpr int GN = (W >= 64 ? 8 : W / 8), NKY = G::NT / 8, GK = W / NKY, NPASS = W / (8 * GN);
  static_assert(NKY * GK == W && NPASS * 8 * GN == W, "gemm_G tiling");
  const int nx = tid & 7, ky = tid >> 3;
  MPDE_NOUNROLL
  for (int pass = 0; pass < NPASS; ++pass) {
    float acc[GK][GN];
    MPDE_UNROLL
    for (int i = 0; i < GK; ++i)
      MPDE_UNROLL
      for (int j = 0; j < GN; ++j) acc[i][j] = 0.f;
    const float* hb = bufH + (size_t)ky * G::RS;
    const float* zb = bufZ + (size_t)(pass * 8 * GN + nx) * G::RS;
    MPDE_NOUNROLL
    for (int r = 0; r < rows; r += 4) {
      float4 h[GK], z[GN];
      MPDE_UNROLL
      for (int i = 0; i < GK; ++i) h[i] = *reinterpret_cast<const float4*>(hb + (size_t)(NKY * i) * G::RS + r);
      MPDE_UNROLL
      for (int j = 0; j < GN; ++j) z[j] = *reinterpret_cast<const float4*>(zb + (size_t)(8 * j) * G::RS + r);
      MPDE_UNROLL
      for (int i = 0; i < GK; ++i)
        MPDE_UNROLL
        for (int j = 0; j < GN; ++j) {
          float a = acc[i][j];
          a = fmaf(h[i].x, z[j].x, a); a = fmaf(h[i].y, z[j].y, a);
          a = fmaf(h[i].z, z[j].z, a); a = fmaf(h[i].w, z[j].w, a);
          acc[i][j] = a;
        }
    }
    MPDE_UNROLL
    for (int i = 0; i < GK; ++i)
      MPDE_UNROLL
      for (int j = 0; j < GN; ++j) {
        float* o = gout + (size_t)(ky + NKY * i) * W + pass * 8 * GN + nx + 8 * j;
        *o += acc[i][j];
      }
  }
}

// ---------------------------------------------------------------------------
// per-thread register tile <-> shared buffer / scratch helpers
// ---------------------------------------------------------------------------
template <int W, int NTY, int CR>
MPDE_DEV void store_tile(float* buf, const float (&a)[CR][4][4], int tx, int ty) {
  typedef Geo<W, NTY> G;
  MPDE_UNROLL
  for (int j = 0; j < 4; ++j) {
    float* row = buf + (size_t)(tx + G::NTX * j) * G::RS + ty * 4;
    MPDE_UNROLL
    for (int cr = 0; cr < CR; ++cr)
      *reinterpret_cast<float4*>(row + cr * G::NP) = make_float4(a[cr][0][j], a[cr][1][j], a[cr][2][j], a[cr][3][j]);
  }
}

// The scratch record of one layer: fields [NFT][W][PTS] with PTS = NP (channel jets) or 4*NP (value sets).
template <int JK, int MODE> struct Rec {
  typedef JetTraits<JK> T;
  static constexpr int NF = T::NF;
  static constexpr int NFT = (MODE == MODE_HVP) ? (T::NF + T::C) : T::NF;
};

// Per-tile processing ---------------------------------------------------------
template <int W, int NTY, int HEAD, int MODE, int ACT = ACT_SIN>
struct Tile {
  typedef Geo<W, NTY> G;
  typedef HeadTraits<HEAD> HT;
  static constexpr int JK = HT::JK, OUT = HT::OUT;
  typedef JetTraits<JK> JT;
  static constexpr int C = JT::C, NF = JT::NF;
  static constexpr bool VAL = (JK == JET_V);
  static constexpr int CR = VAL ? 4 : C;               // register row blocks
  static constexpr int PTS = VAL ? 4 * G::NP : G::NP;  // points per tile
  static constexpr int ROWS = CR * G::NP;              // live rows in the shared buffers
  static constexpr int NFT = Rec<JK, MODE>::NFT;
  static constexpr int NE = VAL ? 4 : 1;               // independent elements per (i, j): row blocks of a value set

  // scratch index of field f, neuron n, first point of this thread (block cr for value sets)
  static MPDE_DEV size_t sidx(int l, int f, int n, int ty, int e) {
    return ((size_t)(l * NFT + f) * W + n) * PTS + (VAL ? e * G::NP : 0) + ty * 4;
  }

  // activation forward on the register tile: acc (pre-activations) -> acc (activations); records -> scratch
  static MPDE_DEV void act_forward(float (&acc)[CR][4][4], float om, float* scr, int l, int tx, int ty) {
    MPDE_UNROLL
    for (int j = 0; j < 4; ++j) {
      const int n = tx + G::NTX * j;
      MPDE_UNROLL
      for (int e = 0; e < NE; ++e) {
        float svv[NF][4];
        MPDE_UNROLL
        for (int i = 0; i < 4; ++i) {
          float z[C], a[C], sv[NF];
          MPDE_UNROLL
          for (int c = 0; c < C; ++c) z[c] = acc[VAL ? e : c][i][j];
          jet_act_fwd<JK, ACT>(om, z, a, sv);
          MPDE_UNROLL
          for (int c = 0; c < C; ++c) acc[VAL ? e : c][i][j] = a[c];
          MPDE_UNROLL
          for (int f = 0; f < NF; ++f) svv[f][i] = sv[f];
        }
        if (MODE != MODE_LOSS) {
          MPDE_UNROLL
          for (int f = 0; f < NF; ++f)
            *reinterpret_cast<float4*>(scr + sidx(l, f, n, ty, e)) = make_float4(svv[f][0], svv[f][1], svv[f][2], svv[f][3]);
        }
      }
    }
  }

  // tangent of the activation: accd (zdot) -> accd (adot); zdot recorded
  static MPDE_DEV void act_tangent(float (&accd)[CR][4][4], float om, float* scr, int l, int tx, int ty, bool record) {
    MPDE_UNROLL
    for (int j = 0; j < 4; ++j) {
      const int n = tx + G::NTX * j;
      MPDE_UNROLL
      for (int e = 0; e < NE; ++e) {
        float4 svv[NF];
        MPDE_UNROLL
        for (int f = 0; f < NF; ++f) svv[f] = *reinterpret_cast<const float4*>(scr + sidx(l, f, n, ty, e));
        float zdv[C][4];
        MPDE_UNROLL
        for (int i = 0; i < 4; ++i) {
          float sv[NF], zd[C], ad[C];
          MPDE_UNROLL
          for (int f = 0; f < NF; ++f) sv[f] = (i == 0 ? svv[f].x : i == 1 ? svv[f].y : i == 2 ? svv[f].z : svv[f].w);
          MPDE_UNROLL
          for (int c = 0; c < C; ++c) { zd[c] = accd[VAL ? e : c][i][j]; zdv[c][i] = zd[c]; }
          jet_act_tan<JK, ACT>(om, sv, zd, ad);
          MPDE_UNROLL
          for (int c = 0; c < C; ++c) accd[VAL ? e : c][i][j] = ad[c];
        }
        if (record) {
          MPDE_UNROLL
          for (int c = 0; c < C; ++c)
            *reinterpret_cast<float4*>(scr + sidx(l, NF + c, n, ty, e)) = make_float4(zdv[c][0], zdv[c][1], zdv[c][2], zdv[c][3]);
        }
      }
    }
  }

  // recompute h_l (or hdot_l) from the record of layer l into a register tile
  static MPDE_DEV void act_recompute(float (&out)[CR][4][4], float om, const float* scr, int l, int tx, int ty, bool tangent) {
    MPDE_UNROLL
    for (int j = 0; j < 4; ++j) {
      const int n = tx + G::NTX * j;
      MPDE_UNROLL
      for (int e = 0; e < NE; ++e) {
        float4 svv[NF], zdv[C];
        MPDE_UNROLL
        for (int f = 0; f < NF; ++f) svv[f] = *reinterpret_cast<const float4*>(scr + sidx(l, f, n, ty, e));
        if (MODE == MODE_HVP && tangent) {
          MPDE_UNROLL
          for (int c = 0; c < C; ++c) zdv[c] = *reinterpret_cast<const float4*>(scr + sidx(l, NF + c, n, ty, e));
        }
        MPDE_UNROLL
        for (int i = 0; i < 4; ++i) {
          float sv[NF], a[C];
          MPDE_UNROLL
          for (int f = 0; f < NF; ++f) sv[f] = (i == 0 ? svv[f].x : i == 1 ? svv[f].y : i == 2 ? svv[f].z : svv[f].w);
          if (MODE == MODE_HVP && tangent) {
            float zd[C];
            MPDE_UNROLL
            for (int c = 0; c < C; ++c) zd[c] = (i == 0 ? zdv[c].x : i == 1 ? zdv[c].y : i == 2 ? zdv[c].z : zdv[c].w);
            jet_act_tan<JK, ACT>(om, sv, zd, a);
          } else {
            jet_act_from_saved<JK, ACT>(om, sv, a);
          }
          MPDE_UNROLL
          for (int c = 0; c < C; ++c) out[VAL ? e : c][i][j] = a[c];
        }
      }
    }
  }

  // adjoint through the activation: ab (abar) -> ab (zbar); with tangent: abd (abardot) -> abd (zbardot)
  static MPDE_DEV void act_backward(float (&ab)[CR][4][4], float (*abd)[4][4], float om, const float* scr, int l, int tx, int ty) {
    MPDE_UNROLL
    for (int j = 0; j < 4; ++j) {
      const int n = tx + G::NTX * j;
      MPDE_UNROLL
      for (int e = 0; e < NE; ++e) {
        float4 svv[NF], zdv[C];
        MPDE_UNROLL
        for (int f = 0; f < NF; ++f) svv[f] = *reinterpret_cast<const float4*>(scr + sidx(l, f, n, ty, e));
        if (MODE == MODE_HVP) {
          MPDE_UNROLL
          for (int c = 0; c < C; ++c) zdv[c] = *reinterpret_cast<const float4*>(scr + sidx(l, NF + c, n, ty, e));
        }
        MPDE_UNROLL
        for (int i = 0; i < 4; ++i) {
          float sv[NF], a_[C], zb[C];
          MPDE_UNROLL
          for (int f = 0; f < NF; ++f) sv[f] = (i == 0 ? svv[f].x : i == 1 ? svv[f].y : i == 2 ? svv[f].z : svv[f].w);
          MPDE_UNROLL
          for (int c = 0; c < C; ++c) a_[c] = ab[VAL ? e : c][i][j];
          jet_act_bwd<JK, ACT>(om, sv, a_, zb);
          if (MODE == MODE_HVP) {
            float zd[C], ad_[C], zbd[C];
            MPDE_UNROLL
            for (int c = 0; c < C; ++c) {
              zd[c] = (i == 0 ? zdv[c].x : i == 1 ? zdv[c].y : i == 2 ? zdv[c].z : zdv[c].w);
              ad_[c] = abd[VAL ? e : c][i][j];
            }
            jet_act_bwd_tan<JK, ACT>(om, sv, zd, a_, ad_, zbd);
            MPDE_UNROLL
            for (int c = 0; c < C; ++c) abd[VAL ? e : c][i][j] = zbd[c];
          }
          MPDE_UNROLL
          for (int c = 0; c < C; ++c) ab[VAL ? e : c][i][j] = zb[c];
        }
      }
    }
  }

  // ------------------------------------------------------------------------
  static MPDE_DEV void run(const KParams& p, const SegDesc& sg, int seg_id, int task, int tile, float* smem,
                           float* scr, float* pg, float* ph) {
    const int tid = threadIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    constexpr int WX = G::NTX / 8;
    const int tx = (warp % WX) * 8 + (lane & 7);
    const int ty = (warp / WX) * 4 + (lane >> 3);
    const float om = p.omega;
    const int L = p.L;

    float* bufA = smem;
    float* bufB = smem + G::BUF;                                   // GRAD/HVP
    float* bufC = smem + 2 * G::BUF;                               // HVP
    float* slab = smem + G::nbuf(MODE) * G::BUF;
    float* misc = slab + G::SLAB;
    float* xs = misc;                           // [2][ROWS]
    float* ybuf = xs + 2 * G::ROWS;             // [ROWS][2]
    float* ydbuf = ybuf + 2 * G::ROWS;          // [ROWS][2]
    float* ypart = slab;                        // [2][NPART][ROWS][2]  (aliases the slab)
    float* red = slab + 2 * G::NPART * G::ROWS * 2;   // [ROWS]
    float* wl = ydbuf + 2 * G::ROWS;            // wLe [W][OUT] at 0, wLd [W][OUT] at 2W
    float* terms = wl + 4 * W;                  // [MAX_SEG]

    const float* th = p.th + (size_t)task * p.th_stride;
    const float* thd = (MODE == MODE_HVP) ? p.thd + (size_t)task * p.thd_stride : nullptr;
    const float* prm = p.prm + (size_t)task * p.nprm;
    const int pt0 = tile * PTS;                                    // first point of the tile within the set
    const float* pts = sg.pts + (size_t)task * sg.task_stride;

    // ---- points ----------------------------------------------------------
    for (int i = tid; i < PTS; i += G::NT) {
      const int gp = pt0 + i;
      float x0 = 0.f, x1 = 0.f;
      if (gp < sg.n) { x0 = pts[2 * (size_t)gp]; x1 = pts[2 * (size_t)gp + 1]; }
      xs[i] = x0; xs[G::ROWS + i] = x1;
    }
    __syncthreads();

    // ---- layer 0: direct (in_dim = 2) -------------------------------------
    float acc[CR][4][4];
    float accd[CR][4][4];   // tangent tile (dead outside MODE_HVP)
    {
      float w0[2][4], b0[4], w0d[2][4], b0d[4];
      MPDE_UNROLL
      for (int j = 0; j < 4; ++j) {
        const int n = tx + G::NTX * j;
        w0[0][j] = mpde_ldg(th + n); w0[1][j] = mpde_ldg(th + W + n); b0[j] = mpde_ldg(th + 2 * W + n);
        if (MODE == MODE_HVP) { w0d[0][j] = mpde_ldg(thd + n); w0d[1][j] = mpde_ldg(thd + W + n); b0d[j] = mpde_ldg(thd + 2 * W + n); }
      }
      MPDE_UNROLL
      for (int e = 0; e < NE; ++e)
        MPDE_UNROLL
        for (int i = 0; i < 4; ++i) {
          const int pl = (VAL ? e * G::NP : 0) + ty * 4 + i;
          const float x0 = xs[pl], x1 = xs[G::ROWS + pl];
          MPDE_UNROLL
          for (int j = 0; j < 4; ++j) {
            acc[VAL ? e : 0][i][j] = fmaf(w0[0][j], x0, fmaf(w0[1][j], x1, b0[j]));
            if (!VAL) {
              acc[1][i][j] = w0[0][j]; acc[2][i][j] = w0[1][j];
              if constexpr (C == 4) acc[3][i][j] = 0.f;
            }
            if (MODE == MODE_HVP) {
              accd[VAL ? e : 0][i][j] = fmaf(w0d[0][j], x0, fmaf(w0d[1][j], x1, b0d[j]));
              if (!VAL) {
                accd[1][i][j] = w0d[0][j]; accd[2][i][j] = w0d[1][j];
                if constexpr (C == 4) accd[3][i][j] = 0.f;
              }
            }
          }
        }
    }
    act_forward(acc, om, scr, 0, tx, ty);
    store_tile<W, NTY, CR>(bufA, acc, tx, ty);
    if (MODE == MODE_HVP) {
      act_tangent(accd, om, scr, 0, tx, ty, true);
      store_tile<W, NTY, CR>(bufB, accd, tx, ty);
    }
    __syncthreads();

    // ---- hidden layers 1..L-1 ----------------------------------------------
    // HVP buffers: h in bufH, hdot in bufB, spare bufS (h and spare swap every layer)
    float* bufH = bufA;
    float* bufS = bufC;
    for (int l = 1; l < L; ++l) {
      const float* lw = th + ph_off_layer(W, l);
      MPDE_UNROLL
      for (int cr = 0; cr < CR; ++cr)
        MPDE_UNROLL
        for (int i = 0; i < 4; ++i)
          MPDE_UNROLL
          for (int j = 0; j < 4; ++j) acc[cr][i][j] = 0.f;
      gemm_F<W, NTY, CR>(acc, bufH, lw, slab, tid, tx, ty);
      {
        const float* bl = lw + 2 * W * W;
        MPDE_UNROLL
        for (int j = 0; j < 4; ++j) {
          const float b = mpde_ldg(bl + tx + G::NTX * j);
          MPDE_UNROLL
          for (int e = 0; e < NE; ++e)
            MPDE_UNROLL
            for (int i = 0; i < 4; ++i) acc[VAL ? e : 0][i][j] += b;
        }
      }
      act_forward(acc, om, scr, l, tx, ty);
      if (MODE == MODE_HVP) {
        // h_l -> spare buffer (nobody reads it during the tangent contractions)
        store_tile<W, NTY, CR>(bufS, acc, tx, ty);
        const float* lwd = thd + ph_off_layer(W, l);
        MPDE_UNROLL
        for (int cr = 0; cr < CR; ++cr)
          MPDE_UNROLL
          for (int i = 0; i < 4; ++i)
            MPDE_UNROLL
            for (int j = 0; j < 4; ++j) accd[cr][i][j] = 0.f;
        gemm_F<W, NTY, CR>(accd, bufB, lw, slab, tid, tx, ty);    // W^T hdot
        gemm_F<W, NTY, CR>(accd, bufH, lwd, slab, tid, tx, ty);   // Wdot^T h
        {
          const float* bld = lwd + 2 * W * W;
          MPDE_UNROLL
          for (int j = 0; j < 4; ++j) {
            const float b = mpde_ldg(bld + tx + G::NTX * j);
            MPDE_UNROLL
            for (int e = 0; e < NE; ++e)
              MPDE_UNROLL
              for (int i = 0; i < 4; ++i) accd[VAL ? e : 0][i][j] += b;
          }
        }
        act_tangent(accd, om, scr, l, tx, ty, true);
        // gemm_F ended with a barrier: every thread is done reading bufB / bufH
        store_tile<W, NTY, CR>(bufB, accd, tx, ty);
        float* t_ = bufH; bufH = bufS; bufS = t_;
        __syncthreads();
      } else {
        // gemm_F ended with a barrier: safe to overwrite bufA in place
        store_tile<W, NTY, CR>(bufH, acc, tx, ty);
        __syncthreads();
      }
    }

    // ---- output layer: y[r][o] = sum_n h[n][r] * wLe[n][o] (+ bLe on value rows) ------------
    {
      constexpr int NPART = G::NPART;
      constexpr int NCH = W / NPART;
      for (int idx = tid; idx < ROWS * NPART; idx += G::NT) {
        const int r = idx % ROWS, part = idx / ROWS;
        float y[2] = {0.f, 0.f}, yd[2] = {0.f, 0.f};
        for (int n = part * NCH; n < (part + 1) * NCH; ++n) {
          const float h = bufH[(size_t)n * G::RS + r];
          MPDE_UNROLL
          for (int o = 0; o < OUT; ++o) y[o] = fmaf(h, wl[n * OUT + o], y[o]);
          if (MODE == MODE_HVP) {
            const float hd = bufB[(size_t)n * G::RS + r];
            MPDE_UNROLL
            for (int o = 0; o < OUT; ++o) yd[o] = fmaf(hd, wl[n * OUT + o], fmaf(h, wl[2 * W + n * OUT + o], yd[o]));
          }
        }
        MPDE_UNROLL
        for (int o = 0; o < OUT; ++o) {
          ypart[(part * G::ROWS + r) * 2 + o] = y[o];
          if (MODE == MODE_HVP) ypart[((NPART + part) * G::ROWS + r) * 2 + o] = yd[o];
        }
      }
      __syncthreads();
    }

    // ---- head: one thread per point -----------------------------------------
    {
      const float* bL = th + ph_off_out(W, L) + W * OUT;
      const float* bLd = (MODE == MODE_HVP) ? thd + ph_off_out(W, L) + W * OUT : nullptr;
      for (int pl = tid; pl < PTS; pl += G::NT) {
        float u[OUT * C], ud[OUT * C], ub[OUT * C], ubd[OUT * C];
        MPDE_UNROLL
        for (int o = 0; o < OUT; ++o)
          MPDE_UNROLL
          for (int c = 0; c < C; ++c) {
            const int r = VAL ? pl : c * G::NP + pl;
            float y = 0.f, yd = 0.f;
            MPDE_UNROLL
            for (int part = 0; part < G::NPART; ++part) {
              y += ypart[(part * G::ROWS + r) * 2 + o];
              if (MODE == MODE_HVP) yd += ypart[((G::NPART + part) * G::ROWS + r) * 2 + o];
            }
            if (c == 0) { y += mpde_ldg(bL + o); if (MODE == MODE_HVP) yd += mpde_ldg(bLd + o); }
            u[o * C + c] = y; ud[o * C + c] = yd;
          }
        const int gp = pt0 + pl;
        const bool valid = gp < sg.n;
        const float x0 = xs[pl], x1 = xs[G::ROWS + pl];
        float res = 0.f;
        float val = head_eval<HEAD>(u, x0, x1, prm, sg.coef, ub, &res);
        if (MODE == MODE_HVP) head_tangent<HEAD>(u, ud, x0, x1, prm, sg.coef, ubd);
        if (!valid) {
          val = 0.f;
          MPDE_UNROLL
          for (int q = 0; q < OUT * C; ++q) { ub[q] = 0.f; ubd[q] = 0.f; }
        }
        red[pl] = val;
        if (MODE == MODE_LOSS && valid && sg.res_out != nullptr) {
          float* ro = sg.res_out + (size_t)task * sg.res_stride;
          if (HEAD == HEAD_EVAL || HEAD == HEAD_EVAL2) {
            MPDE_UNROLL
            for (int o = 0; o < OUT; ++o) ro[(size_t)gp * OUT + o] = u[o * C];
          } else {
            ro[gp] = res;
          }
        }
        if (MODE != MODE_LOSS) {
          MPDE_UNROLL
          for (int o = 0; o < OUT; ++o)
            MPDE_UNROLL
            for (int c = 0; c < C; ++c) {
              const int r = VAL ? pl : c * G::NP + pl;
              ybuf[r * 2 + o] = ub[o * C + c];
              if (MODE == MODE_HVP) ydbuf[r * 2 + o] = ubd[o * C + c];
            }
        }
      }
      __syncthreads();
      if (tid == 0) {
        float s = 0.f;
        for (int i = 0; i < PTS; ++i) s += red[i];
        terms[seg_id] += s;
      }
    }
    if (MODE == MODE_LOSS) { __syncthreads(); return; }

    // ---- output layer adjoint ------------------------------------------------
    {
      const int oL = nat_off_out(W, L);
      for (int n = tid; n < W; n += G::NT) {
        float g[2] = {0.f, 0.f}, hacc[2] = {0.f, 0.f};
        for (int r = 0; r < ROWS; ++r) {
          const float h = bufH[(size_t)n * G::RS + r];
          MPDE_UNROLL
          for (int o = 0; o < OUT; ++o) g[o] = fmaf(h, ybuf[r * 2 + o], g[o]);
          if (MODE == MODE_HVP) {
            const float hd = bufB[(size_t)n * G::RS + r];
            MPDE_UNROLL
            for (int o = 0; o < OUT; ++o) hacc[o] = fmaf(hd, ybuf[r * 2 + o], fmaf(h, ydbuf[r * 2 + o], hacc[o]));
          }
        }
        MPDE_UNROLL
        for (int o = 0; o < OUT; ++o) {
          pg[oL + n * OUT + o] += g[o];
          if (MODE == MODE_HVP) ph[oL + n * OUT + o] += hacc[o];
        }
      }
      if (tid < OUT) {
        float g = 0.f, hh = 0.f;
        const int nv = VAL ? ROWS : G::NP;   // value rows
        for (int r = 0; r < nv; ++r) { g += ybuf[r * 2 + tid]; if (MODE == MODE_HVP) hh += ydbuf[r * 2 + tid]; }
        pg[oL + W * OUT + tid] += g;
        if (MODE == MODE_HVP) ph[oL + W * OUT + tid] += hh;
      }
      // abar[r][n] = sum_o wLe[n][o] ybar[r][o]  (and its tangent)
      MPDE_UNROLL
      for (int j = 0; j < 4; ++j) {
        const int n = tx + G::NTX * j;
        MPDE_UNROLL
        for (int cr = 0; cr < CR; ++cr)
          MPDE_UNROLL
          for (int i = 0; i < 4; ++i) {
            const int r = cr * G::NP + ty * 4 + i;
            float a = 0.f, ad = 0.f;
            MPDE_UNROLL
            for (int o = 0; o < OUT; ++o) {
              a = fmaf(wl[n * OUT + o], ybuf[r * 2 + o], a);
              if (MODE == MODE_HVP) ad = fmaf(wl[n * OUT + o], ydbuf[r * 2 + o], fmaf(wl[2 * W + n * OUT + o], ybuf[r * 2 + o], ad));
            }
            acc[cr][i][j] = a;
            if (MODE == MODE_HVP) accd[cr][i][j] = ad;
          }
      }
    }
    __syncthreads();   // everyone done with bufH / bufB (h_{L-1}, hdot_{L-1}) and ybuf

    // ---- hidden layers, backward ---------------------------------------------
    // GRAD: Z = bufB, H = bufA.   HVP: Z = bufA', Zd = bufB, H/Hd = spare.
    float* bufZ = (MODE == MODE_HVP) ? bufH : bufB;
    float* bufZd = bufB;
    float* bufR = (MODE == MODE_HVP) ? bufS : bufA;  // recomputed h_{l-1} / hdot_{l-1}
    for (int l = L - 1; l >= 1; --l) {
      act_backward(acc, MODE == MODE_HVP ? accd : nullptr, om, scr, l, tx, ty);
      store_tile<W, NTY, CR>(bufZ, acc, tx, ty);
      if (MODE == MODE_HVP) {
        store_tile<W, NTY, CR>(bufZd, accd, tx, ty);
      }
      {
        float hreg[CR][4][4];
        act_recompute(hreg, om, scr, l - 1, tx, ty, false);
        store_tile<W, NTY, CR>(bufR, hreg, tx, ty);
      }
      __syncthreads();
      const int on = nat_off_layer(W, l);
      // bias gradients: sum over the value rows of Z
      for (int n = tid; n < W; n += G::NT) {
        const int nv = VAL ? ROWS : G::NP;
        float g = 0.f, hh = 0.f;
        for (int r = 0; r < nv; ++r) { g += bufZ[(size_t)n * G::RS + r]; if (MODE == MODE_HVP) hh += bufZd[(size_t)n * G::RS + r]; }
        pg[on + W * W + n] += g;
        if (MODE == MODE_HVP) ph[on + W * W + n] += hh;
      }
      if (MODE == MODE_GRAD || p.want_grad) gemm_G<W, NTY>(bufR, bufZ, pg + on, ROWS, tid);
      if (MODE == MODE_HVP) {
        gemm_G<W, NTY>(bufR, bufZd, ph + on, ROWS, tid);     // h zbardot^T
        __syncthreads();
        {
          float hreg[CR][4][4];
          act_recompute(hreg, om, scr, l - 1, tx, ty, true);
          store_tile<W, NTY, CR>(bufR, hreg, tx, ty);
        }
        __syncthreads();
        gemm_G<W, NTY>(bufR, bufZ, ph + on, ROWS, tid);      // hdot zbar^T
      }
      // abar_{l-1} = W zbar ; abardot_{l-1} = W zbardot + Wdot zbar
      const float* lwt = th + ph_off_layer(W, l) + W * W;
      if (MODE == MODE_HVP) {
        const float* lwtd = thd + ph_off_layer(W, l) + W * W;
        MPDE_UNROLL
        for (int cr = 0; cr < CR; ++cr)
          MPDE_UNROLL
          for (int i = 0; i < 4; ++i)
            MPDE_UNROLL
            for (int j = 0; j < 4; ++j) accd[cr][i][j] = 0.f;
        gemm_F<W, NTY, CR>(accd, bufZd, lwt, slab, tid, tx, ty);
        gemm_F<W, NTY, CR>(accd, bufZ, lwtd, slab, tid, tx, ty);
      }
      MPDE_UNROLL
      for (int cr = 0; cr < CR; ++cr)
        MPDE_UNROLL
        for (int i = 0; i < 4; ++i)
          MPDE_UNROLL
          for (int j = 0; j < 4; ++j) acc[cr][i][j] = 0.f;
      gemm_F<W, NTY, CR>(acc, bufZ, lwt, slab, tid, tx, ty);
      // gemm_F ends with a barrier: Z / Zd / R may be overwritten by the next iteration
    }

    // ---- layer 0, backward ------------------------------------------------------
    act_backward(acc, MODE == MODE_HVP ? accd : nullptr, om, scr, 0, tx, ty);
    store_tile<W, NTY, CR>(bufZ, acc, tx, ty);
    if (MODE == MODE_HVP) store_tile<W, NTY, CR>(bufZd, accd, tx, ty);
    __syncthreads();
    for (int n = tid; n < W; n += G::NT) {
      float g0 = 0.f, g1 = 0.f, gb = 0.f, h0 = 0.f, h1 = 0.f, hb = 0.f;
      const float* zr = bufZ + (size_t)n * G::RS;
      const float* zdr = bufZd + (size_t)n * G::RS;
      if (VAL) {
        for (int r = 0; r < ROWS; ++r) {
          const float x0 = xs[r], x1 = xs[G::ROWS + r];
          const float z = zr[r];
          g0 = fmaf(x0, z, g0); g1 = fmaf(x1, z, g1); gb += z;
          if (MODE == MODE_HVP) { const float zd = zdr[r]; h0 = fmaf(x0, zd, h0); h1 = fmaf(x1, zd, h1); hb += zd; }
        }
      } else {
        for (int r = 0; r < G::NP; ++r) {
          const float x0 = xs[r], x1 = xs[G::ROWS + r];
          const float z = zr[r];
          g0 += fmaf(x0, z, zr[G::NP + r]); g1 += fmaf(x1, z, zr[2 * G::NP + r]); gb += z;
          if (MODE == MODE_HVP) {
            const float zd = zdr[r];
            h0 += fmaf(x0, zd, zdr[G::NP + r]); h1 += fmaf(x1, zd, zdr[2 * G::NP + r]); hb += zd;
          }
        }
      }
      pg[n] += g0; pg[W + n] += g1; pg[2 * W + n] += gb;
      if (MODE == MODE_HVP) { ph[n] += h0; ph[W + n] += h1; ph[2 * W + n] += hb; }
    }
    __syncthreads();
  }
};

// ---------------------------------------------------------------------------
// kernel: grid = (ctas_per_task, T); each CTA walks the task's tiles with stride gridDim.x
// ---------------------------------------------------------------------------
template <int W, int NTY, int PDE, int MODE, int ACT>
MPDE_DEV void pde_kernel_body(const KParams& p) {
  typedef Geo<W, NTY> G;
  float* smem = reinterpret_cast<float*>(mpde_smem());
  const int tid = threadIdx.x;
  const int task = blockIdx.y;
  const int cta = blockIdx.y * gridDim.x + blockIdx.x;
  constexpr int OUT = (PDE == PDE_HYPER || PDE == PDE_EVAL2) ? 2 : 1;

  float* slab = smem + G::nbuf(MODE) * G::BUF;
  float* misc = slab + G::SLAB;
  float* wl = misc + 2 * G::ROWS + 2 * G::ROWS + 2 * G::ROWS;
  float* terms = wl + 4 * W;

  float* scr = (MODE != MODE_LOSS) ? p.scratch + (size_t)cta * p.scratch_per_cta : nullptr;
  float* pg = (MODE != MODE_LOSS) ? p.partial_g + (size_t)cta * p.P : nullptr;
  float* ph = (MODE == MODE_HVP) ? p.partial_h + (size_t)cta * p.P : nullptr;

  // stage the output-layer weights; zero this CTA's partials
  {
    const float* th = p.th + (size_t)task * p.th_stride;
    const int oo = ph_off_out(W, p.L);
    for (int i = tid; i < W * OUT; i += G::NT) {
      wl[i] = th[oo + i];
      if (MODE == MODE_HVP) wl[2 * W + i] = (p.thd + (size_t)task * p.thd_stride)[oo + i];
    }
    if (tid < MAX_SEG) terms[tid] = 0.f;
    if (MODE != MODE_LOSS) {
      for (int i = tid; i < p.P; i += G::NT) { pg[i] = 0.f; if (MODE == MODE_HVP) ph[i] = 0.f; }
    }
  }
  __syncthreads();

  const int n_tiles = p.seg_tile_begin[p.n_seg];
  for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
    int s = 0;
    while (s + 1 < p.n_seg && t >= p.seg_tile_begin[s + 1]) ++s;
    const int tile = t - p.seg_tile_begin[s];
    const SegDesc& sg = p.seg[s];
    if (PDE == PDE_POISSON) {
      if (sg.head == HEAD_POISSON_DOMAIN) Tile<W, NTY, HEAD_POISSON_DOMAIN, MODE, ACT>::run(p, sg, s, task, tile, smem, scr, pg, ph);
      else Tile<W, NTY, HEAD_POISSON_BOUNDARY, MODE, ACT>::run(p, sg, s, task, tile, smem, scr, pg, ph);
    } else if (PDE == PDE_BURGERS) {
      if (sg.head == HEAD_BURGERS_DOMAIN) Tile<W, NTY, HEAD_BURGERS_DOMAIN, MODE, ACT>::run(p, sg, s, task, tile, smem, scr, pg, ph);
      else Tile<W, NTY, HEAD_BURGERS_IC, MODE, ACT>::run(p, sg, s, task, tile, smem, scr, pg, ph);
    } else if (PDE == PDE_HYPER) {
      if (sg.head == HEAD_HYPER_DOMAIN) Tile<W, NTY, HEAD_HYPER_DOMAIN, MODE, ACT>::run(p, sg, s, task, tile, smem, scr, pg, ph);
      else if (sg.head == HEAD_HYPER_TOP) Tile<W, NTY, HEAD_HYPER_TOP, MODE, ACT>::run(p, sg, s, task, tile, smem, scr, pg, ph);
      else Tile<W, NTY, HEAD_HYPER_BOTTOM, MODE, ACT>::run(p, sg, s, task, tile, smem, scr, pg, ph);
    } else if (PDE == PDE_EVAL1) {
      Tile<W, NTY, HEAD_EVAL, MODE_LOSS, ACT>::run(p, sg, s, task, tile, smem, scr, pg, ph);
    } else {
      Tile<W, NTY, HEAD_EVAL2, MODE_LOSS, ACT>::run(p, sg, s, task, tile, smem, scr, pg, ph);
    }
  }
  __syncthreads();
  if (tid < MAX_SEG) p.partial_terms[(size_t)cta * MAX_SEG + tid] = terms[tid];
}

// SIREN network (the published configurations)
template <int W, int NTY, int PDE, int MODE>
MPDE_GLOBAL void MPDE_LAUNCH_BOUNDS((Geo<W, NTY>::NT), 1) pde_kernel(KParams p) {
  pde_kernel_body<W, NTY, PDE, MODE, ACT_SIN>(p);
}
// `--nosiren` network: swish hidden layers, no omega (maml_pde.py:115, field.py:187-188)
template <int W, int NTY, int PDE, int MODE>
MPDE_GLOBAL void MPDE_LAUNCH_BOUNDS((Geo<W, NTY>::NT), 1) pde_kernel_swish(KParams p) {
  pde_kernel_body<W, NTY, PDE, MODE, ACT_SWISH>(p);
}
template <int W, int NTY, int PDE, int MODE, int ACT> struct KernelSel;
template <int W, int NTY, int PDE, int MODE> struct KernelSel<W, NTY, PDE, MODE, ACT_SIN> {
  static constexpr auto fn() { return pde_kernel<W, NTY, PDE, MODE>; }
};
template <int W, int NTY, int PDE, int MODE> struct KernelSel<W, NTY, PDE, MODE, ACT_SWISH> {
  static constexpr auto fn() { return pde_kernel_swish<W, NTY, PDE, MODE>; }
};

}  // namespace mpde
