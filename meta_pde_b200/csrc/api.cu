// api.cu -- the C ABI of libmetapde_b200 (include/metapde_b200.h): argument validation,
// workspace carving, launch planning, and the stream-ordered MAML / LEAP rollouts.
// No allocation, no synchronisation, no mutable global state besides one-time
// kernel attributes and a launch counter.
#include "../../include/metapde_b200.h"
#include "host_common.h"
#include "aux_kernels.cuh"
#include "sampler_kernels.cuh"
#ifndef MPDE_EMU
#include "tc_params.cuh"
#endif
#include <cstdlib>
#include <mutex>
#include <vector>
#ifndef MPDE_EMU
#include <nvtx3/nvToolsExt.h>
#endif

namespace mpde {

// NVTX ranges around the entry points and every loss / grad / H.v evaluation inside the rollouts (SURVEY.md section 5:
// the reference's own tracing is absl logging + TensorBoard scalars).  Off unless MPDE_NVTX=1 (read once per process):
// with it, `ncu --nvtx --nvtx-include "mpde:maml_outer/"` profiles one rollout, and a timeline tool shows the
// K forward evaluations and K reverse H.v evaluations of an outer step by name.  No NVTX call is made when it is off.
#ifndef MPDE_EMU
struct NvtxRange {
  bool on;
  explicit NvtxRange(const char* name) : on(enabled()) { if (on) nvtxRangePushA(name); }
  ~NvtxRange() { if (on) nvtxRangePop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
  static bool enabled() {
    static const bool e = []() { const char* s = getenv("MPDE_NVTX"); return s && s[0] == '1'; }();
    return e;
  }
};
#else
struct NvtxRange { explicit NvtxRange(const char*) {} };
#endif

std::atomic<long long> g_launches{0};

int device_sm_count() {
#ifdef MPDE_EMU
  return 4;
#else
  static std::atomic<int> cache[PerDeviceOnce::MAXDEV];   // zero-initialised; per device
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= PerDeviceOnce::MAXDEV) return 148;
  int n = cache[dev].load(std::memory_order_relaxed);
  if (n < 1) {
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n < 1) n = 148;
    cache[dev].store(n, std::memory_order_relaxed);
  }
  return n;
#endif
}

// ---- optional event timing of the jet kernel (mpde_prof_enable / mpde_prof_read) ----
#ifndef MPDE_EMU
struct ProfState {
  std::mutex mu;
  std::atomic<bool> on{false};
  std::vector<cudaEvent_t> ev[4];   // pairs: [2i] start, [2i+1] stop; slot 3 = tc_wgrad_kernel launches
  size_t used[4] = {0, 0, 0, 0};
};
static ProfState g_prof;
static bool prof_begin(int mode, void* stream, cudaEvent_t* stop) {
  if (!g_prof.on.load(std::memory_order_relaxed)) return false;
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;   // events cannot be recorded into a capturing stream
  if (cudaStreamIsCapturing((cudaStream_t)stream, &cap) != cudaSuccess || cap != cudaStreamCaptureStatusNone) return false;
  std::lock_guard<std::mutex> lk(g_prof.mu);
  auto& v = g_prof.ev[mode];
  if (g_prof.used[mode] + 2 > v.size()) {
    cudaEvent_t a, b;
    if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) return false;
    v.push_back(a); v.push_back(b);
  }
  cudaEvent_t start = v[g_prof.used[mode]];
  *stop = v[g_prof.used[mode] + 1];
  g_prof.used[mode] += 2;
  cudaEventRecord(start, (cudaStream_t)stream);
  return true;
}
#endif

static int dispatch(int W, int act, int pde, int mode, int op, const KParams* p, int gx, int gy, void* stream, KernelInfo* info) {
  if (act == MPDE_ACT_SWISH) {       // `--nosiren` network: FP32-FMA kernels at every width
    switch (W) {
      case 32: return dispatch_sw_w32(pde, mode, op, p, gx, gy, stream, info);
      case 64: return dispatch_sw_w64(pde, mode, op, p, gx, gy, stream, info);
      case 128: return dispatch_sw_w128(pde, mode, op, p, gx, gy, stream, info);
      case 256: return dispatch_sw_w256(pde, mode, op, p, gx, gy, stream, info);
    }
    return MPDE_UNSUPPORTED_NET;
  }
  switch (W) {
    case 32: return dispatch_w32(pde, mode, op, p, gx, gy, stream, info);
    case 64: return dispatch_w64(pde, mode, op, p, gx, gy, stream, info);
    case 128: return dispatch_w128(pde, mode, op, p, gx, gy, stream, info);
    case 256: return dispatch_w256(pde, mode, op, p, gx, gy, stream, info);
  }
  return MPDE_UNSUPPORTED_NET;
}

// The tensor-core path serves widths 128/256 where it is implemented; MPDE_DISABLE_TC=1 forces the FP32-FMA kernels.
static bool use_tc_path(int W, int L, int mode) {
#ifdef MPDE_EMU
  (void)W; (void)L; (void)mode;
  return false;
#else
  static const bool disabled = []() { const char* e = getenv("MPDE_DISABLE_TC"); return e && e[0] == '1'; }();
  return !disabled && tc_supported(W, L, mode);
#endif
}

static inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

struct Carver {
  char* base; size_t off, cap;
  Carver(void* b, size_t c) : base((char*)b), off(0), cap(c) {}
  float* take(size_t nfloats) {
    size_t bytes = align_up(nfloats * sizeof(float));
    float* p = (float*)(base ? base + off : nullptr);
    off += bytes;
    return p;
  }
  bool ok() const { return off <= cap; }
};

static int pde_out_dim(int kind) { return kind == MPDE_PDE_HYPER_ELASTICITY ? 2 : 1; }

static int check_net(const mpde_net_desc* net, int out_dim_required) {
  if (!net) return MPDE_INVALID_ARG;
  if (net->num_layers < 1 || net->num_layers > 32) return MPDE_UNSUPPORTED_NET;
  if (net->in_dim != 2) return MPDE_UNSUPPORTED_NET;
  if (net->layer_size != 32 && net->layer_size != 64 && net->layer_size != 128 && net->layer_size != 256)
    return MPDE_UNSUPPORTED_NET;
  if (net->out_dim != 1 && net->out_dim != 2) return MPDE_UNSUPPORTED_NET;
  if (out_dim_required && net->out_dim != out_dim_required) return MPDE_UNSUPPORTED_NET;
  if (net->nonlinearity != MPDE_ACT_SIN && net->nonlinearity != MPDE_ACT_SWISH) return MPDE_UNSUPPORTED_NET;
  return MPDE_OK;
}

// -------------------------------------------------------------------------
// evaluation plan: segments, tiles, CTAs per task, workspace
// -------------------------------------------------------------------------
struct EvalPlan {
  int W, L, OUT, P, PH, pde, mode, T;
  int n_seg;
  int head[MAX_SEG], npts[MAX_SEG], term_of_seg[MAX_SEG];
  float term_scale[MAX_SEG], weight[MAX_SEG], inv_elems[MAX_SEG];
  int tile_begin[MAX_SEG + 1];
  int n_terms, domain_seg;
  int G;                    // CTAs per task
  KernelInfo ki;
  size_t scratch_per_cta;   // floats
  // workspace sizes (floats)
  size_t n_th, n_thd, n_scratch, n_pg, n_ph, n_pt, n_gtmp;
  bool tc;                  // tensor-core path
  int act;                  // mpde_nonlinearity
  size_t n_wimg, n_wshift;  // floats: tc weight image [T][image], shifts [T][2*TC_MAX_L]
  size_t n_rec, n_zsp, n_psmall;          // tc GRAD spills (floats) and per-CTA small-parameter partials
  size_t n_wimg_d, n_wshift_d, n_zspd, n_psmall_h;   // tc HVP: direction image / shifts, zbardot spill, H v partials
  size_t n_wpart; int wg_splits;                     // tc_wgrad row splits: partial results of splits 1..wg_splits-1
  long long rec_seg_base[MAX_SEG]; long long rec_per_task, zsp_per_task; int PS;
  size_t bytes;
};

static int kernel_pde(int kind) { return kind == MPDE_PDE_POISSON ? PDE_POISSON : kind == MPDE_PDE_TD_BURGERS ? PDE_BURGERS : PDE_HYPER; }

static int make_plan(const mpde_net_desc* net, const mpde_pde_desc* pde, int T, int mode, bool need_gtmp, EvalPlan* pl) {
  if (!pde || T < 1) return MPDE_INVALID_ARG;
  if (pde->kind < 0 || pde->kind > 2) return MPDE_INVALID_ARG;
  if (mode < 0 || mode > 2) return MPDE_INVALID_ARG;
  int rc = check_net(net, pde_out_dim(pde->kind));
  if (rc) return rc;
  pl->W = net->layer_size; pl->L = net->num_layers; pl->OUT = net->out_dim;
  pl->P = mpde_param_count(net);
  pl->PH = ph_count(pl->W, pl->L, pl->OUT);
  pl->pde = kernel_pde(pde->kind); pl->mode = mode; pl->T = T;
  const float bw = pde->bc_weight;
  if (pde->kind == MPDE_PDE_POISSON) {
    pl->n_seg = 2; pl->n_terms = 2; pl->domain_seg = 1;
    const int heads[2] = {HEAD_POISSON_BOUNDARY, HEAD_POISSON_DOMAIN};
    for (int s = 0; s < 2; ++s) { pl->head[s] = heads[s]; pl->term_of_seg[s] = s; pl->term_scale[s] = 1.f; }
    pl->weight[0] = bw; pl->weight[1] = 1.f;
  } else if (pde->kind == MPDE_PDE_TD_BURGERS) {
    pl->n_seg = 4; pl->n_terms = 4; pl->domain_seg = 3;
    const int heads[4] = {HEAD_BURGERS_IC, HEAD_BURGERS_IC, HEAD_BURGERS_IC, HEAD_BURGERS_DOMAIN};
    const int terms[4] = {1, 2, 0, 3};   // sets: left, right, initial, domain -> terms: initial, left, right, domain
    for (int s = 0; s < 4; ++s) { pl->head[s] = heads[s]; pl->term_of_seg[s] = terms[s]; pl->term_scale[s] = 1.f; pl->weight[s] = (s < 3) ? bw : 1.f; }
  } else {
    pl->n_seg = 3; pl->n_terms = 3; pl->domain_seg = 2;
    const int heads[3] = {HEAD_HYPER_TOP, HEAD_HYPER_BOTTOM, HEAD_HYPER_DOMAIN};
    const int terms[3] = {1, 0, 2};      // sets: top, bottom, domain -> terms: bottom, top, domain
    for (int s = 0; s < 3; ++s) { pl->head[s] = heads[s]; pl->term_of_seg[s] = terms[s]; }
    pl->term_scale[0] = 1000.f; pl->term_scale[1] = 1000.f; pl->term_scale[2] = 1.f;
    pl->weight[0] = bw; pl->weight[1] = bw; pl->weight[2] = 1.f;
  }
  pl->act = net->nonlinearity;
  pl->tc = pl->act == MPDE_ACT_SIN && use_tc_path(pl->W, pl->L, mode);   // the tcgen05 epilogues are sin-jet only
#ifndef MPDE_EMU
  if (pl->tc && mode == MODE_HVP) rc = dispatch_tc_hvp(pl->W, pl->pde, 0, nullptr, 0, 0, nullptr, &pl->ki);
  else if (pl->tc) rc = dispatch_tc(pl->W, pl->pde, mode, 0, nullptr, 0, 0, nullptr, &pl->ki);
  else
#endif
  rc = dispatch(pl->W, pl->act, pl->pde, mode, 0, nullptr, 0, 0, nullptr, &pl->ki);
  if (rc) return rc;
  const int NP = pl->ki.np;
  pl->tile_begin[0] = 0;
  for (int s = 0; s < pl->n_seg; ++s) {
    const int n = pde->n_points[s];
    if (n < 1) return MPDE_INVALID_ARG;
    pl->npts[s] = n;
    const bool domain = (s == pl->domain_seg);
    const int per_tile = domain ? NP : 4 * NP;
    pl->tile_begin[s + 1] = pl->tile_begin[s] + (n + per_tile - 1) / per_tile;
    const float elems = domain ? (float)n : (float)n * (float)pl->OUT;   // np.mean over [N] or [N, out_dim]
    pl->inv_elems[s] = 1.f / elems;
  }
  // CTAs per task: minimise waves * tiles-per-CTA; prefer fewer CTAs on ties (fewer partial buffers)
  const int tiles = pl->tile_begin[pl->n_seg];
  const long long slots = (long long)device_sm_count() * pl->ki.occ;
  int best_g = 1; double best_cost = 1e300;
  const int gmax = tiles < 512 ? tiles : 512;
  for (int g = 1; g <= gmax; ++g) {
    const long long ctas = (long long)T * g;
    const long long waves = (ctas + slots - 1) / slots;
    const int per = (tiles + g - 1) / g;
    const double cost = (double)waves * ((double)per + 0.15);   // 0.15 tile-equivalents of per-CTA overhead
    if (cost < best_cost * (1.0 - 1e-9)) { best_cost = cost; best_g = g; }
  }
  pl->G = best_g;
  const size_t ctas = (size_t)T * pl->G;
  const int nft_val = (mode == MODE_HVP) ? 3 : 2, nft_ch = (mode == MODE_HVP) ? 9 : 5;
  const size_t per_layer = (size_t)pl->W * NP * (size_t)((4 * nft_val > nft_ch) ? 4 * nft_val : nft_ch);
  pl->scratch_per_cta = (mode == MODE_LOSS) ? 0 : per_layer * pl->L;
  pl->n_th = (size_t)T * pl->PH;
  pl->n_thd = (mode == MODE_HVP) ? (size_t)T * pl->PH : 0;
  pl->n_scratch = ctas * pl->scratch_per_cta;
  pl->n_pg = (mode == MODE_LOSS) ? 0 : ctas * pl->P;
  pl->n_ph = (mode == MODE_HVP) ? ctas * pl->P : 0;
  pl->n_pt = ctas * MAX_SEG;
  pl->n_gtmp = need_gtmp ? (size_t)T * pl->P : 0;
  pl->n_wimg = 0; pl->n_wshift = 0;
  pl->n_rec = pl->n_zsp = pl->n_psmall = 0; pl->PS = 0; pl->rec_per_task = pl->zsp_per_task = 0;
  pl->n_wimg_d = pl->n_wshift_d = pl->n_zspd = pl->n_psmall_h = 0;
  pl->n_wpart = 0; pl->wg_splits = 1;
#ifndef MPDE_EMU
  if (pl->tc) {
    pl->n_wimg = (size_t)T * (tc_image_bytes(pl->W, pl->L, mode != MODE_LOSS) / 4);
    pl->n_wshift = (size_t)T * tc::TC_MAX_L * 2;
    if (mode != MODE_LOSS) {
      pl->n_scratch = 0; pl->n_pg = 0; pl->n_ph = 0;     // FP32-FMA kernel buffers are not used
      long long off = 0;
      for (int s = 0; s < pl->n_seg; ++s) {
        pl->rec_seg_base[s] = off;
        const int jk = (pl->head[s] == HEAD_POISSON_DOMAIN) ? JET_LAP : (pl->head[s] == HEAD_BURGERS_DOMAIN) ? JET_XX
                     : (pl->head[s] == HEAD_HYPER_DOMAIN) ? JET_G : JET_V;
        const int rfp = (mode == MODE_HVP) ? tc::rec_fields_points_hvp(jk) : tc::rec_fields_points(jk);
        off += (long long)(pl->tile_begin[s + 1] - pl->tile_begin[s]) * pl->L * rfp * pl->W;
      }
      pl->rec_per_task = off;
      pl->zsp_per_task = (long long)tiles * (pl->L - 1) * 64 * pl->W;   // tiles of 64 jet rows in both modes
      pl->PS = tc::ps_count(pl->W, pl->L, pl->OUT);
      pl->n_rec = (size_t)T * off; pl->n_zsp = (size_t)T * pl->zsp_per_task; pl->n_psmall = ctas * pl->PS * 2;   // one partial per (CTA, epilogue group)
      {
        // tc_wgrad_kernel: one CTA per (task, layer, k tile) and row split; splits chosen to fill whole waves of the
        // SMs (512 items at T = 64: 3.46 waves -> 2 splits = 6.92; 64 items at T = 8: 0.43 -> 2 splits)
        const long long items = (long long)T * (pl->L - 1) * (pl->W / 128), sms = device_sm_count();
        int best = 1; double best_cost = 1e300;
        for (int rs = 1; rs <= 8 && rs <= tiles; ++rs) {
          const double cost = (double)((items * rs + sms - 1) / sms) / rs + 0.02 * rs;
          if (cost < best_cost * (1.0 - 1e-9)) { best_cost = cost; best = rs; }
        }
        pl->wg_splits = best;
        pl->n_wpart = (size_t)(best - 1) * T * (pl->L - 1) * pl->W * pl->W;
      }
      if (mode == MODE_HVP) {
        pl->n_wimg_d = pl->n_wimg; pl->n_wshift_d = pl->n_wshift; pl->n_zspd = pl->n_zsp; pl->n_psmall_h = pl->n_psmall;
      }
    }
  }
#endif
  Carver c(nullptr, 0);
  c.take(pl->n_th); c.take(pl->n_thd); c.take(pl->n_scratch); c.take(pl->n_pg); c.take(pl->n_ph); c.take(pl->n_pt); c.take(pl->n_gtmp);
  c.take(pl->n_wimg); c.take(pl->n_wshift); c.take(pl->n_rec); c.take(pl->n_zsp); c.take(pl->n_psmall);
  c.take(pl->n_wimg_d); c.take(pl->n_wshift_d); c.take(pl->n_zspd); c.take(pl->n_psmall_h); c.take(pl->n_wpart);
  pl->bytes = c.off + 256;
  return MPDE_OK;
}

template <typename K, typename... Args>
static int launch_simple(K kfn, int gx, int gy, int nt, size_t smem, void* stream, Args... args) {
  MPDE_LAUNCH(kfn, dim3(gx, gy, 1), dim3(nt, 1, 1), smem, stream, args...);
  return MPDE_CHECK_LAUNCH() ? MPDE_CUDA_ERROR : MPDE_OK;
}

static int grid_for(long long n, int nt = 256) {
  long long g = (n + nt - 1) / nt;
  const long long cap = (long long)device_sm_count() * 8;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

// One evaluation of loss / loss+grad / loss+grad+Hv for T tasks.
static int run_eval(const mpde_net_desc* net, const mpde_pde_desc* pde, int T, int mode, const float* theta,
                    long long theta_stride, const float* v, long long v_stride, const mpde_points* points,
                    const float* pde_params, float* loss, long long loss_stride, float* terms, long long terms_stride,
                    float* domain_residual, float* grad, float* hv, int want_grad, void* workspace, size_t ws_bytes,
                    void* stream) {
  if (!net || !pde || !theta || !points || !pde_params) return MPDE_INVALID_ARG;
  NvtxRange nvtx(mode == MODE_LOSS ? "mpde:loss" : mode == MODE_GRAD ? "mpde:loss_grad" : "mpde:loss_hvp");
  if (mode == MODE_GRAD && !grad) return MPDE_INVALID_ARG;
  if (mode == MODE_HVP && (!hv || !v)) return MPDE_INVALID_ARG;
  EvalPlan pl;
  const bool need_gtmp = (mode == MODE_HVP) && !grad;
  int rc = make_plan(net, pde, T, mode, need_gtmp, &pl);
  if (rc) return rc;
  if (!workspace || ws_bytes < pl.bytes) return MPDE_WORKSPACE_TOO_SMALL;
  for (int s = 0; s < pl.n_seg; ++s) if (!points->set[s]) return MPDE_INVALID_ARG;
  char* wsb = (char*)workspace;
  wsb += (256 - ((uintptr_t)wsb & 255)) & 255;
  Carver c(wsb, ws_bytes);
  float* th = c.take(pl.n_th); float* thd = c.take(pl.n_thd); float* scratch = c.take(pl.n_scratch);
  float* pg = c.take(pl.n_pg); float* ph = c.take(pl.n_ph); float* pt = c.take(pl.n_pt); float* gtmp = c.take(pl.n_gtmp);
  float* wimg = c.take(pl.n_wimg); float* wshift = c.take(pl.n_wshift);
  float* rec = c.take(pl.n_rec); float* zsp = c.take(pl.n_zsp); float* psmall = c.take(pl.n_psmall);
  float* wimg_d = c.take(pl.n_wimg_d); float* wshift_d = c.take(pl.n_wshift_d); float* zspd = c.take(pl.n_zspd);
  float* psmall_h = c.take(pl.n_psmall_h); float* wpart = c.take(pl.n_wpart);
  (void)wimg; (void)wshift; (void)rec; (void)zsp; (void)psmall; (void)wimg_d; (void)wshift_d; (void)zspd; (void)psmall_h; (void)wpart;
  if (mode != MODE_HVP) thd = nullptr;

  // 1. prep: theta -> theta-hat (shared theta: prepared once)
  const int T_prep = (theta_stride == 0 && (mode != MODE_HVP || v_stride == 0)) ? 1 : T;
  {
    PrepParams pp;
    pp.W = pl.W; pp.L = pl.L; pp.OUT = pl.OUT; pp.log_scale = net->log_scale; pp.P = pl.P; pp.PH = pl.PH; pp.T = T_prep;
    pp.theta = theta; pp.theta_stride = theta_stride;
    pp.v = (mode == MODE_HVP) ? v : nullptr; pp.v_stride = v_stride;
    pp.th = th; pp.thd = thd;
    int gx = grid_for(pl.PH);
    if (gx > 64) gx = 64;
    rc = launch_simple(prep_kernel, gx, T_prep, 256, 0, stream, pp);
    if (rc) return rc;
  }
  // 2. the fused jet kernel
  KParams kp;
  kp.L = pl.L; kp.P = pl.P; kp.PH = pl.PH; kp.omega = net->omega; kp.T = T; kp.n_seg = pl.n_seg;
  for (int s = 0; s < pl.n_seg; ++s) {
    kp.seg[s].pts = points->set[s];
    kp.seg[s].task_stride = points->task_stride[s];
    kp.seg[s].n = pl.npts[s];
    kp.seg[s].head = pl.head[s];
    kp.seg[s].coef = pl.weight[s] * pl.term_scale[s] * pl.inv_elems[s];
    kp.seg[s].res_out = (s == pl.domain_seg && mode == MODE_LOSS) ? domain_residual : nullptr;
    kp.seg[s].res_stride = pl.npts[s];
  }
  for (int s = 0; s <= pl.n_seg; ++s) kp.seg_tile_begin[s] = pl.tile_begin[s];
  kp.prm = pde_params; kp.nprm = mpde_num_pde_params(pde->kind);
  kp.th = th; kp.th_stride = (T_prep == 1) ? 0 : pl.PH;
  kp.thd = thd; kp.thd_stride = (T_prep == 1) ? 0 : pl.PH;
  kp.scratch = scratch; kp.scratch_per_cta = (long long)pl.scratch_per_cta;
  kp.partial_g = pg; kp.partial_h = ph; kp.partial_terms = pt;
  kp.want_grad = want_grad;
#ifndef MPDE_EMU
  if (pl.tc) {
    // weight image for the tensor-core kernels (hidden kernels from natural theta, output layer from theta-hat)
    rc = tc_launch_prep(pl.W, pl.L, pl.OUT, T_prep, mode != MODE_LOSS, theta, theta_stride, th, kp.th_stride, (uint8_t*)wimg,
                        (long long)tc_image_bytes(pl.W, pl.L, mode != MODE_LOSS), (int*)wshift, 2 * tc::TC_MAX_L, stream);
    if (rc) return rc;
    if (mode == MODE_HVP) {   // the direction's image: hidden kernels of v, output layer of theta-hat-dot
      rc = tc_launch_prep(pl.W, pl.L, pl.OUT, T_prep, 1, v, v_stride, thd, kp.thd_stride, (uint8_t*)wimg_d,
                          (long long)tc_image_bytes(pl.W, pl.L, 1), (int*)wshift_d, 2 * tc::TC_MAX_L, stream);
      if (rc) return rc;
    }
  }
  cudaEvent_t prof_stop;
  const bool prof = prof_begin(mode, stream, &prof_stop);
#endif
#ifndef MPDE_EMU
  if (pl.tc) {
    const long long img_bytes = (long long)tc_image_bytes(pl.W, pl.L, mode != MODE_LOSS);
    tc::TcParams tp;
    tp.k = kp;
    tp.wimg = (const uint8_t*)wimg; tp.wimg_stride = (T_prep == 1) ? 0 : img_bytes;
    tp.wshift = (const int*)wshift; tp.wshift_stride = (T_prep == 1) ? 0 : 2 * tc::TC_MAX_L;
    tp.rec = rec; tp.rec_task_stride = pl.rec_per_task;
    for (int s = 0; s < MAX_SEG; ++s) tp.rec_seg_base[s] = (s < pl.n_seg) ? pl.rec_seg_base[s] : 0;
    tp.zsp = zsp; tp.zsp_task_stride = pl.zsp_per_task;
    tp.psmall = psmall; tp.PS = pl.PS;
    {
      // off by default since round 2: with two units in flight per SM the prefetched lines were evicted before the reverse
      // pass read them (DRAM reads of the HVP launch 77 -> 124 GB, +2.5..4 % time); MPDE_REC_PREFETCH=1 re-enables it
      static const int pf = []() { const char* e = getenv("MPDE_REC_PREFETCH"); return (e && e[0] == '1') ? 1 : 0; }();
      tp.rec_prefetch = pf;
    }
    if (mode == MODE_HVP) {
      tc::HvParams hp;
      hp.t = tp;
      hp.wimg_d = (const uint8_t*)wimg_d; hp.wimg_d_stride = tp.wimg_stride;
      hp.wshift_d = (const int*)wshift_d; hp.wshift_d_stride = tp.wshift_stride;
      hp.zspd = zspd; hp.psmall_h = psmall_h;
      hp.want_grad = 1;   // the small-parameter gradient is always needed (chain rule of H v through the io scales)
      rc = dispatch_tc_hvp(pl.W, pl.pde, 1, &hp, pl.G, T, stream, nullptr);
    } else
    rc = dispatch_tc(pl.W, pl.pde, mode, 1, &tp, pl.G, T, stream, nullptr);
  } else
#endif
  rc = dispatch(pl.W, pl.act, pl.pde, mode, 1, &kp, pl.G, T, stream, nullptr);
#ifndef MPDE_EMU
  if (prof) cudaEventRecord(prof_stop, (cudaStream_t)stream);
#endif
  if (rc) return rc;
  // 3. loss terms
  {
    LossFinParams lp;
    lp.G = pl.G; lp.n_seg = pl.n_seg; lp.n_terms = pl.n_terms;
    for (int s = 0; s < pl.n_seg; ++s) {
      lp.term_of_seg[s] = pl.term_of_seg[s]; lp.term_scale[s] = pl.term_scale[s];
      lp.inv_elems[s] = pl.inv_elems[s]; lp.weight[s] = pl.weight[s];
    }
    lp.partial_terms = pt; lp.loss = loss; lp.terms = terms; lp.loss_stride = loss_stride; lp.terms_stride = terms_stride;
    if (loss || terms) {
      rc = launch_simple(loss_finalize_kernel, (T + 127) / 128, 1, 128, 0, stream, lp, T);
      if (rc) return rc;
    }
  }
  if (mode == MODE_LOSS) return MPDE_OK;
  // 4. reduce the per-CTA partials, then the chain rule back to theta
  float* gout = grad ? grad : gtmp;
#ifndef MPDE_EMU
  if (pl.tc) {
    // hidden kernels: contraction over the task's rows from the spills; small parameters: fixed-order sum over CTAs
    tc::TcWgradParams wp;
    wp.W = pl.W; wp.L = pl.L; wp.T = T; wp.n_seg = pl.n_seg; wp.omega = net->omega;
    wp.tile_rows = 64; wp.hvp_records = (mode == MODE_HVP); wp.want_hv = 0; wp.zspd = zspd;
    for (int s = 0; s <= pl.n_seg; ++s) wp.seg_tile_begin[s] = pl.tile_begin[s];
    for (int s = 0; s < pl.n_seg; ++s) {
      wp.seg_jk[s] = (pl.head[s] == HEAD_POISSON_DOMAIN) ? JET_LAP : (pl.head[s] == HEAD_BURGERS_DOMAIN) ? JET_XX
                   : (pl.head[s] == HEAD_HYPER_DOMAIN) ? JET_G : JET_V;
      wp.rec_seg_base[s] = pl.rec_seg_base[s];
    }
    wp.rec = rec; wp.rec_task_stride = pl.rec_per_task; wp.zsp = zsp; wp.zsp_task_stride = pl.zsp_per_task;
    wp.gout = gout; wp.g_stride = pl.P;
    wp.row_splits = pl.wg_splits; wp.part = wpart; wp.part_stride = (long long)T * (pl.L - 1) * pl.W * pl.W;
    if (mode == MODE_GRAD || want_grad) {
      cudaEvent_t ws; const bool wprof = prof_begin(3, stream, &ws);
      rc = tc_launch_wgrad(&wp, stream);
      if (wprof) cudaEventRecord(ws, (cudaStream_t)stream);
      if (rc) return rc;
    }
    rc = tc_launch_reduce_small(pl.W, pl.L, pl.OUT, pl.PS, pl.G * 2, T, psmall, gout, pl.P, stream);
    if (rc) return rc;
    if (mode == MODE_HVP) {
      wp.want_hv = 1; wp.gout = hv;
      cudaEvent_t ws; const bool wprof = prof_begin(3, stream, &ws);
      rc = tc_launch_wgrad(&wp, stream);
      if (wprof) cudaEventRecord(ws, (cudaStream_t)stream);
      if (rc) return rc;
      rc = tc_launch_reduce_small(pl.W, pl.L, pl.OUT, pl.PS, pl.G * 2, T, psmall_h, hv, pl.P, stream);
      if (rc) return rc;
    }
  } else
#endif
  {
    ReduceParams rp;
    rp.P = pl.P; rp.G = pl.G;
    rp.part_a = pg; rp.out_a = gout; rp.out_a_stride = pl.P;
    rp.part_b = (mode == MODE_HVP) ? ph : nullptr; rp.out_b = hv; rp.out_b_stride = pl.P;
    rc = launch_simple(reduce_partials_kernel, (pl.P + 255) / 256, T, 256, 0, stream, rp);
    if (rc) return rc;
  }
  {
    ChainParams cp;
    cp.W = pl.W; cp.L = pl.L; cp.OUT = pl.OUT; cp.log_scale = net->log_scale; cp.P = pl.P; cp.PH = pl.PH;
    cp.theta = theta; cp.theta_stride = theta_stride; cp.v = v; cp.v_stride = v_stride;
    cp.th = th; cp.th_stride = kp.th_stride; cp.thd = thd; cp.thd_stride = kp.thd_stride;
    cp.g = gout; cp.g_stride = pl.P; cp.h = (mode == MODE_HVP) ? hv : nullptr; cp.h_stride = pl.P;
    if (net->log_scale) {
      rc = launch_simple(chain_kernel, T, 1, 256, 256 * sizeof(float), stream, cp);
      if (rc) return rc;
    }
  }
  return MPDE_OK;
}

}  // namespace mpde

using namespace mpde;

// =========================================================================== C ABI
extern "C" {

int mpde_param_count(const mpde_net_desc* net) {
  if (!net) return -1;
  return nat_count(net->layer_size, net->num_layers, net->out_dim, net->log_scale);
}
int mpde_num_point_sets(int k) { return k == MPDE_PDE_POISSON ? 2 : k == MPDE_PDE_TD_BURGERS ? 4 : k == MPDE_PDE_HYPER_ELASTICITY ? 3 : -1; }
int mpde_num_terms(int k) { return mpde_num_point_sets(k); }
int mpde_num_pde_params(int k) { return k == MPDE_PDE_POISSON ? NPRM_POISSON : k == MPDE_PDE_TD_BURGERS ? NPRM_BURGERS : k == MPDE_PDE_HYPER_ELASTICITY ? NPRM_HYPER : -1; }

const char* mpde_status_string(int s) {
  switch (s) {
    case MPDE_OK: return "ok";
    case MPDE_INVALID_ARG: return "invalid argument";
    case MPDE_UNSUPPORTED_NET: return "unsupported network (sin or swish hidden layers, in_dim 2, layer_size in {32,64,128,256}, out_dim matching the PDE)";
    case MPDE_WORKSPACE_TOO_SMALL: return "workspace too small";
    case MPDE_CUDA_ERROR: return "CUDA error";
    case MPDE_STATIC_SCRATCH_EXCEEDED: return "problem needs more scratch than the library's static device array holds (mpde_static_scratch_bytes): use the workspace entry point";
  }
  return "unknown status";
}

int64_t mpde_launch_count(void) { return (int64_t)g_launches.load(); }
const char* mpde_backend(void) {
#ifdef MPDE_EMU
  return "emu";
#else
  return "cuda";
#endif
}

int mpde_prof_enable(int on) {
#ifndef MPDE_EMU
  std::lock_guard<std::mutex> lk(g_prof.mu);
  g_prof.on = on != 0;
  if (on) for (int m = 0; m < 4; ++m) g_prof.used[m] = 0;
#else
  (void)on;
#endif
  return MPDE_OK;
}
int mpde_prof_read(int mode, double* total_ms, int64_t* launches) {
  if (mode < 0 || mode > 3 || !total_ms || !launches) return MPDE_INVALID_ARG;
  *total_ms = 0.0; *launches = 0;
#ifndef MPDE_EMU
  std::lock_guard<std::mutex> lk(g_prof.mu);
  auto& v = g_prof.ev[mode];
  for (size_t i = 0; i + 1 < g_prof.used[mode]; i += 2) {
    if (cudaEventSynchronize(v[i + 1]) != cudaSuccess) return MPDE_CUDA_ERROR;
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, v[i], v[i + 1]) != cudaSuccess) return MPDE_CUDA_ERROR;
    *total_ms += ms; *launches += 1;
  }
  g_prof.used[mode] = 0;
#endif
  return MPDE_OK;
}

size_t mpde_workspace_bytes(const mpde_net_desc* net, const mpde_pde_desc* pde, int T, int mode) {
  EvalPlan pl;
  if (make_plan(net, pde, T, mode, mode == MODE_HVP, &pl)) return 0;
  return pl.bytes;
}

int mpde_loss(const mpde_net_desc* net, const mpde_pde_desc* pde, int T, const float* theta, int64_t theta_task_stride,
              const mpde_points* points, const float* pde_params, float* loss, float* terms, float* domain_residual,
              void* workspace, size_t workspace_bytes, void* stream) {
  if (!loss && !terms && !domain_residual) return MPDE_INVALID_ARG;
  const int nt = pde ? mpde_num_terms(pde->kind) : 0;
  return run_eval(net, pde, T, MODE_LOSS, theta, theta_task_stride, nullptr, 0, points, pde_params, loss, 1, terms, nt,
                  domain_residual, nullptr, nullptr, 0, workspace, workspace_bytes, stream);
}

int mpde_loss_grad(const mpde_net_desc* net, const mpde_pde_desc* pde, int T, const float* theta,
                   int64_t theta_task_stride, const mpde_points* points, const float* pde_params, float* loss,
                   float* terms, float* grad, void* workspace, size_t workspace_bytes, void* stream) {
  const int nt = pde ? mpde_num_terms(pde->kind) : 0;
  return run_eval(net, pde, T, MODE_GRAD, theta, theta_task_stride, nullptr, 0, points, pde_params, loss, 1, terms, nt,
                  nullptr, grad, nullptr, 1, workspace, workspace_bytes, stream);
}

int mpde_loss_hvp(const mpde_net_desc* net, const mpde_pde_desc* pde, int T, const float* theta,
                  int64_t theta_task_stride, const float* v, int64_t v_task_stride, const mpde_points* points,
                  const float* pde_params, float* loss, float* terms, float* grad, float* hv, void* workspace,
                  size_t workspace_bytes, void* stream) {
  const int nt = pde ? mpde_num_terms(pde->kind) : 0;
  return run_eval(net, pde, T, MODE_HVP, theta, theta_task_stride, v, v_task_stride, points, pde_params, loss, 1, terms,
                  nt, nullptr, grad, hv, grad ? 1 : 0, workspace, workspace_bytes, stream);
}

// ------------------------------------------------------------------ field eval
size_t mpde_field_eval_workspace_bytes(const mpde_net_desc* net, int T) {
  if (check_net(net, 0) || T < 1) return 0;
  const size_t PH = ph_count(net->layer_size, net->num_layers, net->out_dim);
  return align_up((size_t)T * PH * sizeof(float)) + align_up((size_t)T * 4096 * MAX_SEG * sizeof(float)) + 512;
}

int mpde_field_eval(const mpde_net_desc* net, int T, const float* theta, int64_t theta_task_stride, const float* coords,
                    int64_t coords_task_stride, int n, float* out, void* workspace, size_t workspace_bytes, void* stream) {
  NvtxRange nvtx("mpde:field_eval");
  int rc = check_net(net, 0);
  if (rc) return rc;
  if (!theta || !coords || !out || T < 1 || n < 1) return MPDE_INVALID_ARG;
  if (!workspace || workspace_bytes < mpde_field_eval_workspace_bytes(net, T)) return MPDE_WORKSPACE_TOO_SMALL;
  const int W = net->layer_size, L = net->num_layers, OUT = net->out_dim;
  const int PH = ph_count(W, L, OUT);
  char* wsb = (char*)workspace;
  wsb += (256 - ((uintptr_t)wsb & 255)) & 255;
  Carver c(wsb, workspace_bytes);
  const int T_prep = theta_task_stride == 0 ? 1 : T;
  float* th = c.take((size_t)T * PH);
  float* pt = c.take((size_t)T * 4096 * MAX_SEG);
  PrepParams pp;
  pp.W = W; pp.L = L; pp.OUT = OUT; pp.log_scale = net->log_scale; pp.P = mpde_param_count(net); pp.PH = PH; pp.T = T_prep;
  pp.theta = theta; pp.theta_stride = theta_task_stride; pp.v = nullptr; pp.v_stride = 0; pp.th = th; pp.thd = nullptr;
  int gx = grid_for(PH); if (gx > 64) gx = 64;
  rc = launch_simple(prep_kernel, gx, T_prep, 256, 0, stream, pp);
  if (rc) return rc;
  const int kpde = (OUT == 2) ? PDE_EVAL2 : PDE_EVAL1;
  KernelInfo ki;
  rc = dispatch(W, net->nonlinearity, kpde, MODE_LOSS, 0, nullptr, 0, 0, nullptr, &ki);
  if (rc) return rc;
  KParams kp;
  kp.L = L; kp.P = pp.P; kp.PH = PH; kp.omega = net->omega; kp.T = T; kp.n_seg = 1;
  kp.seg[0].pts = coords; kp.seg[0].task_stride = coords_task_stride; kp.seg[0].n = n;
  kp.seg[0].head = (OUT == 2) ? HEAD_EVAL2 : HEAD_EVAL; kp.seg[0].coef = 0.f;
  kp.seg[0].res_out = out; kp.seg[0].res_stride = (long long)n * OUT;
  const int per_tile = 4 * ki.np;
  const int tiles = (n + per_tile - 1) / per_tile;
  kp.seg_tile_begin[0] = 0; kp.seg_tile_begin[1] = tiles;
  static const float dummy_prm[16] = {0};
  (void)dummy_prm;
  kp.prm = theta; kp.nprm = 0;   // heads of the eval kernels never read pde params
  kp.th = th; kp.th_stride = (T_prep == 1) ? 0 : PH; kp.thd = nullptr; kp.thd_stride = 0;
  kp.scratch = nullptr; kp.scratch_per_cta = 0; kp.partial_g = nullptr; kp.partial_h = nullptr; kp.partial_terms = pt;
  kp.want_grad = 0;
  int G = tiles < 4096 ? tiles : 4096;
  return dispatch(W, net->nonlinearity, kpde, MODE_LOSS, 1, &kp, G, T, stream, nullptr);
}

// ------------------------------------------------------------------ samplers: jax.random on the GPU
int mpde_random_bits(int S, const uint32_t* keys, int64_t n, uint32_t* out, void* stream) {
  if (S < 1 || n < 1 || !keys || !out) return MPDE_INVALID_ARG;
  int gx = grid_for(n); if (gx > 1024) gx = 1024;
  return launch_simple(random_bits_kernel, gx, S, 256, 0, stream, keys, (long long)n, out);
}
int mpde_random_uniform(int S, const uint32_t* keys, int64_t n, float minval, float maxval, float* out, void* stream) {
  if (S < 1 || n < 1 || !keys || !out) return MPDE_INVALID_ARG;
  int gx = grid_for(n); if (gx > 1024) gx = 1024;
  return launch_simple(random_uniform_kernel, gx, S, 256, 0, stream, keys, (long long)n, minval, maxval, out);
}
int mpde_random_split(int S, const uint32_t* keys, int num, uint32_t* out, void* stream) {
  if (S < 1 || num < 1 || !keys || !out) return MPDE_INVALID_ARG;
  return launch_simple(random_split_kernel, S, 1, 32, 0, stream, keys, num, out);
}
int mpde_sample_poisson_candidates(int S, const uint32_t* keys, int n, const float* geo, float xmin, float xmax, float ymin,
                                   float ymax, float* boundary, float* cand_xy, float* cand_g, void* stream) {
  if (S < 1 || n < 1 || !keys || !geo || !boundary || !cand_xy || !cand_g) return MPDE_INVALID_ARG;
  PoissonSampleParams p;
  p.S = S; p.n = n; p.keys = keys; p.geo = geo; p.xmin = xmin; p.xmax = xmax; p.ymin = ymin; p.ymax = ymax;
  p.boundary = boundary; p.cand_xy = cand_xy; p.cand_g = cand_g;
  int gx = grid_for(4LL * n); if (gx > 256) gx = 256;
  return launch_simple(poisson_sample_kernel, gx, S, 256, 0, stream, p);
}

// ------------------------------------------------------------------ K1f without any allocation
// mpde_field_eval with its scratch (theta-hat + per-CTA partial terms) in a file-scope device array instead of a caller
// workspace: for callers that cannot allocate (CUDA-graph capture, serving loops).  No counterpart in the reference.
// Every GPU that loads the library holds the array until the process exits.
#ifndef MPDE_EMU
static __device__ __align__(256) unsigned char g_static_scratch[MPDE_STATIC_SCRATCH_BYTES];
static void* g_static_scratch_ptr[PerDeviceOnce::MAXDEV];   // DEVICE address of the array, per device (the symbol's name is
                                                            // only a host-side stand-in: never pass it to a kernel)
static int static_scratch_address(void** out) {
  static PerDeviceOnce once;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= PerDeviceOnce::MAXDEV) return MPDE_CUDA_ERROR;
  const cudaError_t e = once.run([&]() {          // once per device, with that device current
    void* ptr = nullptr;
    const cudaError_t ee = cudaGetSymbolAddress(&ptr, g_static_scratch);
    if (ee == cudaSuccess) g_static_scratch_ptr[dev] = ptr;
    return ee;
  });
  if (e != cudaSuccess || g_static_scratch_ptr[dev] == nullptr) return MPDE_CUDA_ERROR;
  *out = g_static_scratch_ptr[dev];
  return MPDE_OK;
}
#endif

size_t mpde_static_scratch_bytes(void) { return (size_t)MPDE_STATIC_SCRATCH_BYTES; }

int mpde_field_eval_static(const mpde_net_desc* net, int T, const float* theta, int64_t theta_task_stride,
                           const float* coords, int64_t coords_task_stride, int n, float* out, void* stream) {
#ifdef MPDE_EMU
  (void)net; (void)T; (void)theta; (void)theta_task_stride; (void)coords; (void)coords_task_stride; (void)n; (void)out; (void)stream;
  return MPDE_INVALID_ARG;      // the emulator build has no device memory
#else
  int rc = check_net(net, 0);
  if (rc) return rc;
  if (!theta || !coords || !out || T < 1 || n < 1) return MPDE_INVALID_ARG;
  const size_t need = mpde_field_eval_workspace_bytes(net, T);
  if (need == 0) return MPDE_UNSUPPORTED_NET;
  if (need > (size_t)MPDE_STATIC_SCRATCH_BYTES) return MPDE_STATIC_SCRATCH_EXCEEDED;
  void* scratch = nullptr;
  rc = static_scratch_address(&scratch);
  if (rc) return rc;
  // nothing in the scratch has to start at zero: prep_kernel writes every theta-hat entry the jet kernel reads and the
  // kernel writes (never accumulates) its per-CTA partial terms, so the previous call's contents are simply overwritten
  return mpde_field_eval(net, T, theta, theta_task_stride, coords, coords_task_stride, n, out, scratch,
                         (size_t)MPDE_STATIC_SCRATCH_BYTES, stream);
#endif
}

// ------------------------------------------------------------------ K4
int mpde_maml_inner_update(int T, int P, const float* theta, int64_t theta_task_stride, const float* grad,
                           const float* alpha, float alpha_scalar, int softplus_lrs, float inner_lr, float clip,
                           float* theta_next, float* gnorm, void* stream) {
  if (T < 1 || P < 1 || !theta || !grad || !theta_next) return MPDE_INVALID_ARG;
  InnerUpdParams p;
  p.P = P; p.theta = theta; p.theta_stride = theta_task_stride; p.grad = grad; p.alpha = alpha; p.alpha_scalar = alpha_scalar;
  p.softplus = softplus_lrs; p.inner_lr = inner_lr; p.clip = clip; p.theta_next = theta_next; p.gnorm = gnorm; p.gnorm_stride = 1;
  // this entry point has no workspace argument: one chunk per task, the partial is parked in gnorm[t]
  if (!gnorm) return MPDE_INVALID_ARG;
  p.part = gnorm;
  int rc = launch_simple(maml_inner_norm_kernel, 1, T, 1024, 1024 * sizeof(float), stream, p);
  if (rc) return rc;
  return launch_simple(maml_inner_update_kernel, 1, T, 1024, 0, stream, p);
}

// chunks per task of the two-stage inner-step kernels: ~2 CTAs of 256 threads per SM over all tasks, >= 2048 elements each
static int inner_splits(int T, int P) {
  long long s = (2LL * device_sm_count() + T - 1) / T;
  const long long smax = (P + 2047) / 2048;
  if (s > smax) s = smax;
  if (s > INNER_MAX_SPLITS) s = INNER_MAX_SPLITS;
  if (s < 1) s = 1;
  return (int)s;
}

size_t mpde_sumsq_workspace_bytes(int64_t n) { (void)n; return 1024 * sizeof(float) + 256; }
int mpde_sumsq(const float* x, int64_t n, float* out, void* workspace, size_t workspace_bytes, void* stream) {
  if (!x || !out || n < 1) return MPDE_INVALID_ARG;
  if (!workspace || workspace_bytes < mpde_sumsq_workspace_bytes(n)) return MPDE_WORKSPACE_TOO_SMALL;
  int nb = grid_for(n, 256); if (nb > 1024) nb = 1024;
  int rc = launch_simple(sumsq_stage1_kernel, nb, 1, 256, 256 * sizeof(float), stream, x, (long long)n, (float*)workspace);
  if (rc) return rc;
  return launch_simple(sumsq_stage2_kernel, 1, 1, 32, 0, stream, (const float*)workspace, nb, out);
}
int mpde_adam_step(float* x, const float* g, float* m, float* v, int64_t n, int t, float lr, float b1, float b2,
                   float eps, const float* sumsq, float clip, void* stream) {
  if (!x || !g || !m || !v || n < 1 || t < 1) return MPDE_INVALID_ARG;
  const float bc1 = 1.f - powf(b1, (float)t), bc2 = 1.f - powf(b2, (float)t);
  return launch_simple(adam_kernel, grid_for(n), 1, 256, 0, stream, x, g, m, v, (long long)n, lr, b1, b2, eps, bc1, bc2, sumsq, clip);
}
int mpde_sgd_step(float* x, const float* g, int64_t n, float lr, const float* sumsq, float clip, void* stream) {
  if (!x || !g || n < 1) return MPDE_INVALID_ARG;
  return launch_simple(sgd_kernel, grid_for(n), 1, 256, 0, stream, x, g, (long long)n, lr, sumsq, clip);
}

int mpde_task_sum(float* out, const float* x, int T, int64_t P, float scale, void* stream) {
  if (!out || !x || T < 1 || P < 1) return MPDE_INVALID_ARG;
  return launch_simple(task_sum_kernel, grid_for(P), 1, 256, 0, stream, out, x, T, (long long)P, scale);
}

int mpde_outer_step(float* theta, int64_t n_theta, float* lrs, int64_t n_lrs, const float* g_packed, float* m_theta,
                    float* v_theta, float* m_lrs, float* v_lrs, int optimizer, int t_theta, float lr_theta, float b1,
                    float b2, float eps, int t_lrs, float lr_lrs, float b1_lrs, float b2_lrs, float eps_lrs, float clip,
                    float* sumsq_out, void* stream) {
  NvtxRange nvtx("mpde:outer_step");
  if (!theta || !g_packed || n_theta < 1 || n_lrs < 0 || t_theta < 1) return MPDE_INVALID_ARG;
  if (optimizer != 0 && optimizer != 1) return MPDE_INVALID_ARG;
  if (optimizer == 1 && (!m_theta || !v_theta)) return MPDE_INVALID_ARG;
  if (n_lrs > 0 && (!lrs || !m_lrs || !v_lrs || t_lrs < 1)) return MPDE_INVALID_ARG;
  OuterStepParams p;
  p.n_theta = n_theta; p.n_lrs = n_lrs; p.theta = theta; p.lrs = lrs; p.g = g_packed;
  p.m_theta = m_theta; p.v_theta = v_theta; p.m_lrs = m_lrs; p.v_lrs = v_lrs; p.opt = optimizer;
  p.lr_theta = lr_theta; p.b1 = b1; p.b2 = b2; p.eps = eps;
  p.bc1 = 1.f - powf(b1, (float)t_theta); p.bc2 = 1.f - powf(b2, (float)t_theta);
  p.lr_lrs = lr_lrs; p.b1l = b1_lrs; p.b2l = b2_lrs; p.epsl = eps_lrs;
  p.bc1l = 1.f - powf(b1_lrs, (float)(t_lrs > 0 ? t_lrs : 1)); p.bc2l = 1.f - powf(b2_lrs, (float)(t_lrs > 0 ? t_lrs : 1));
  p.clip = clip; p.sumsq_out = sumsq_out;
  // every CTA re-reads g_theta for the norm (L2-resident): keep the grid at two CTAs per SM for large trees
  int gx = grid_for(n_theta + n_lrs);
  const int cap = 2 * device_sm_count();
  if (gx > cap) gx = cap;
  return launch_simple(outer_step_kernel, gx, 1, 256, 256 * sizeof(float), stream, p);
}

// ------------------------------------------------------------------ MAML rollout
struct MamlWs {
  size_t eval_bytes;
  size_t n_thetas, n_grads, n_lam, n_gbar, n_tmp, n_galpha, n_gnorm, n_M, n_loss1, n_part;
  size_t bytes;
};
static int maml_ws(const mpde_net_desc* net, const mpde_pde_desc* pi, const mpde_pde_desc* po, const mpde_maml_cfg* cfg,
                   int T, bool outer, MamlWs* w) {
  if (!net || !pi || !cfg || T < 1 || cfg->inner_steps < 1) return MPDE_INVALID_ARG;
  const int K = cfg->inner_steps;
  const size_t P = (size_t)mpde_param_count(net);
  size_t e = mpde_workspace_bytes(net, pi, T, outer ? MODE_HVP : MODE_GRAD);
  if (!e) return MPDE_UNSUPPORTED_NET;
  if (outer) {
    if (!po) return MPDE_INVALID_ARG;
    size_t e2 = mpde_workspace_bytes(net, po, T, MODE_GRAD);
    if (!e2) return MPDE_UNSUPPORTED_NET;
    if (e2 > e) e = e2;
    size_t e3 = mpde_workspace_bytes(net, pi, T, MODE_GRAD);
    if (e3 > e) e = e3;
  }
  w->eval_bytes = align_up(e);
  w->n_thetas = (size_t)K * T * P;                 // theta_1..theta_K
  w->n_grads = outer ? (size_t)K * T * P : (size_t)T * P;   // g_k kept for the reverse sweep
  w->n_lam = outer ? (size_t)T * P : 0;
  w->n_gbar = outer ? (size_t)T * P : 0;
  w->n_tmp = outer ? (size_t)T * P : 0;
  w->n_galpha = outer ? (size_t)T * P : 0;
  w->n_gnorm = (size_t)K * T;
  w->n_M = (size_t)K * T;
  w->n_loss1 = (size_t)T;
  w->n_part = (size_t)T * INNER_MAX_SPLITS;
  Carver c(nullptr, 0);
  c.take(w->n_thetas); c.take(w->n_grads); c.take(w->n_lam); c.take(w->n_gbar); c.take(w->n_tmp); c.take(w->n_galpha);
  c.take(w->n_gnorm); c.take(w->n_M); c.take(w->n_loss1); c.take(w->n_part);
  w->bytes = w->eval_bytes + c.off + 512;
  return MPDE_OK;
}

size_t mpde_maml_workspace_bytes(const mpde_net_desc* net, const mpde_pde_desc* pde_inner, const mpde_pde_desc* pde_outer,
                                 const mpde_maml_cfg* cfg, int T) {
  MamlWs w;
  if (maml_ws(net, pde_inner, pde_outer, cfg, T, pde_outer != nullptr, &w)) return 0;
  return w.bytes;
}

static int maml_impl(const mpde_net_desc* net, const mpde_pde_desc* pi, const mpde_pde_desc* po, const mpde_maml_cfg* cfg,
                     int T, const float* theta0, const float* inner_lrs, const mpde_points* inner_points,
                     const mpde_points* outer_points, const mpde_points* final_points, const mpde_points* aux_points,
                     const float* pde_params, float grad_scale, float* meta_grad_theta, float* meta_grad_lrs,
                     float* losses, float* meta_losses, float* aux_terms, float* theta_final, bool outer,
                     void* workspace, size_t workspace_bytes, void* stream) {
  MamlWs w;
  NvtxRange nvtx(outer ? "mpde:maml_outer" : "mpde:maml_adapt");
  int rc = maml_ws(net, pi, po, cfg, T, outer, &w);
  if (rc) return rc;
  if (!theta0 || !inner_points || !final_points || !pde_params || !losses) return MPDE_INVALID_ARG;
  if (outer && (!outer_points || !aux_points || !meta_grad_theta || !meta_losses)) return MPDE_INVALID_ARG;
  if (!workspace || workspace_bytes < w.bytes) return MPDE_WORKSPACE_TOO_SMALL;
  const int K = cfg->inner_steps;
  const int P = mpde_param_count(net);
  const long long TP = (long long)T * P;
  char* wsb = (char*)workspace;
  wsb += (256 - ((uintptr_t)wsb & 255)) & 255;
  void* eval_ws = wsb;
  Carver c(wsb + w.eval_bytes, workspace_bytes - w.eval_bytes);
  float* thetas = c.take(w.n_thetas); float* grads = c.take(w.n_grads); float* lam = c.take(w.n_lam);
  float* gbar = c.take(w.n_gbar); float* tmp = c.take(w.n_tmp); float* galpha = c.take(w.n_galpha);
  float* gnorm = c.take(w.n_gnorm); float* Mk = c.take(w.n_M); float* loss1 = c.take(w.n_loss1);
  float* part = c.take(w.n_part);
  (void)loss1;
  const int S = inner_splits(T, P);
  const int nterms = mpde_num_terms(pi->kind);

  // ---- forward sweep: K inner SGD steps (maml.py:132-155) ----
  for (int k = 0; k < K; ++k) {
    const float* th_k = (k == 0) ? theta0 : thetas + (size_t)(k - 1) * TP;
    const long long st_k = (k == 0) ? 0 : P;
    float* g_k = outer ? grads + (size_t)k * TP : grads;
    rc = run_eval(net, pi, T, MODE_GRAD, th_k, st_k, nullptr, 0, &inner_points[k], pde_params, losses + k, K + 1,
                  nullptr, 0, nullptr, g_k, nullptr, 1, eval_ws, w.eval_bytes, stream);
    if (rc) return rc;
    InnerUpdParams up;
    up.P = P; up.theta = th_k; up.theta_stride = st_k; up.grad = g_k;
    up.alpha = inner_lrs ? inner_lrs + (size_t)k * P : nullptr; up.alpha_scalar = 1.0f; up.softplus = cfg->softplus_lrs;
    up.inner_lr = cfg->inner_lr; up.clip = cfg->inner_grad_clip;
    up.theta_next = thetas + (size_t)k * TP; up.gnorm = gnorm + (size_t)k * T; up.gnorm_stride = 1; up.part = part;
    rc = launch_simple(maml_inner_norm_kernel, S, T, 256, 256 * sizeof(float), stream, up);
    if (rc) return rc;
    rc = launch_simple(maml_inner_update_kernel, S, T, 256, 0, stream, up);
    if (rc) return rc;
  }
  const float* th_K = thetas + (size_t)(K - 1) * TP;
  // final inner loss (maml.py:158) -> losses[:, K]
  rc = run_eval(net, pi, T, MODE_LOSS, th_K, P, nullptr, 0, final_points, pde_params, losses + K, K + 1, nullptr, 0,
                nullptr, nullptr, nullptr, 0, eval_ws, w.eval_bytes, stream);
  if (rc) return rc;
  if (theta_final) {
    rc = launch_simple(axpby_kernel, grid_for(TP), 1, 256, 0, stream, theta_final, th_K, 1.0f, 0.0f, TP);
    if (rc) return rc;
  }
  if (!outer) return MPDE_OK;

  // aux-only outer loss at theta_K (maml.py:192)
  if (aux_terms) {
    rc = run_eval(net, po, T, MODE_LOSS, th_K, P, nullptr, 0, aux_points, pde_params, nullptr, 0, aux_terms, nterms,
                  nullptr, nullptr, nullptr, 0, eval_ws, w.eval_bytes, stream);
    if (rc) return rc;
  }

  // ---- reverse sweep ----
  rc = launch_simple(fill_kernel, grid_for(TP), 1, 256, 0, stream, lam, 0.0f, TP);
  if (rc) return rc;
  float coef = 1.0f;   // decay^(K-k-1)
  for (int k = K - 1; k >= 0; --k) {
    const float* th_k1 = thetas + (size_t)k * TP;
    // (M_{k+1}, grad M_{k+1}) at theta_{k+1} on the outer points of step k
    rc = run_eval(net, po, T, MODE_GRAD, th_k1, P, nullptr, 0, &outer_points[k], pde_params, Mk + k, K, nullptr, 0,
                  nullptr, tmp, nullptr, 1, eval_ws, w.eval_bytes, stream);
    if (rc) return rc;
    InnerVjpParams vp;
    vp.P = P; vp.lam = lam; vp.add = tmp; vp.c_add = coef; vp.grad = grads + (size_t)k * TP;
    vp.alpha = inner_lrs ? inner_lrs + (size_t)k * P : nullptr; vp.alpha_scalar = 1.0f; vp.softplus = cfg->softplus_lrs;
    vp.gnorm = gnorm + (size_t)k * T; vp.gnorm_stride = 1; vp.inner_lr = cfg->inner_lr; vp.clip = cfg->inner_grad_clip;
    vp.gbar = gbar; vp.galpha = (inner_lrs && meta_grad_lrs) ? galpha : nullptr; vp.part = part;
    rc = launch_simple(maml_inner_vjp_dot_kernel, S, T, 256, 256 * sizeof(float), stream, vp);
    if (rc) return rc;
    rc = launch_simple(maml_inner_vjp_kernel, S, T, 256, 0, stream, vp);
    if (rc) return rc;
    if (vp.galpha) {
      rc = launch_simple(task_sum_kernel, grid_for(P), 1, 256, 0, stream, meta_grad_lrs + (size_t)k * P, (const float*)galpha, T, (long long)P, grad_scale);
      if (rc) return rc;
    }
    const float* th_k = (k == 0) ? theta0 : thetas + (size_t)(k - 1) * TP;
    const long long st_k = (k == 0) ? 0 : P;
    // H(theta_k) gbar on the inner points of step k
    rc = run_eval(net, pi, T, MODE_HVP, th_k, st_k, gbar, P, &inner_points[k], pde_params, nullptr, 0, nullptr, 0,
                  nullptr, nullptr, tmp, 0, eval_ws, w.eval_bytes, stream);
    if (rc) return rc;
    rc = launch_simple(axpby_kernel, grid_for(TP), 1, 256, 0, stream, lam, (const float*)tmp, 1.0f, 1.0f, TP);
    if (rc) return rc;
    coef *= cfg->outer_loss_decay;
  }
  rc = launch_simple(task_sum_kernel, grid_for(P), 1, 256, 0, stream, meta_grad_theta, (const float*)lam, T, (long long)P, grad_scale);
  if (rc) return rc;
  // Mk is laid out [T][K] via (ptr + k, stride K)
  return launch_simple(meta_loss_kernel, (T + 127) / 128, 1, 128, 0, stream, meta_losses, (const float*)Mk, T, K, cfg->outer_loss_decay);
}

int mpde_maml_outer(const mpde_net_desc* net, const mpde_pde_desc* pde_inner, const mpde_pde_desc* pde_outer,
                    const mpde_maml_cfg* cfg, int T, const float* theta0, const float* inner_lrs,
                    const mpde_points* inner_points, const mpde_points* outer_points, const mpde_points* final_points,
                    const mpde_points* aux_points, const float* pde_params, float grad_scale, float* meta_grad_theta,
                    float* meta_grad_lrs, float* losses, float* meta_losses, float* aux_terms, float* theta_final,
                    void* workspace, size_t workspace_bytes, void* stream) {
  if (!pde_outer) return MPDE_INVALID_ARG;
  return maml_impl(net, pde_inner, pde_outer, cfg, T, theta0, inner_lrs, inner_points, outer_points, final_points,
                   aux_points, pde_params, grad_scale, meta_grad_theta, meta_grad_lrs, losses, meta_losses, aux_terms,
                   theta_final, true, workspace, workspace_bytes, stream);
}

int mpde_maml_adapt(const mpde_net_desc* net, const mpde_pde_desc* pde_inner, const mpde_maml_cfg* cfg, int T,
                    const float* theta0, const float* inner_lrs, const mpde_points* inner_points,
                    const mpde_points* final_points, const float* pde_params, float* theta_final, float* losses,
                    void* workspace, size_t workspace_bytes, void* stream) {
  return maml_impl(net, pde_inner, nullptr, cfg, T, theta0, inner_lrs, inner_points, nullptr, final_points, nullptr,
                   pde_params, 0.f, nullptr, nullptr, losses, nullptr, nullptr, theta_final, false, workspace,
                   workspace_bytes, stream);
}

// ------------------------------------------------------------------ LEAP rollout
struct LeapWs { size_t eval_bytes, n_tp, n_T, bytes; };
static int leap_ws(const mpde_net_desc* net, const mpde_pde_desc* pde, const mpde_leap_cfg* cfg, int T, LeapWs* w) {
  if (!net || !pde || !cfg || T < 1 || cfg->inner_steps < 1) return MPDE_INVALID_ARG;
  size_t e = mpde_workspace_bytes(net, pde, T, MODE_GRAD);
  if (!e) return MPDE_UNSUPPORTED_NET;
  w->eval_bytes = align_up(e);
  w->n_tp = (size_t)T * mpde_param_count(net);
  w->n_T = (size_t)T;
  Carver c(nullptr, 0);
  for (int i = 0; i < 6; ++i) c.take(w->n_tp);   // theta_a, theta_b, grad, accum, m, v
  c.take(w->n_T); c.take(w->n_T);
  w->bytes = w->eval_bytes + c.off + 512;
  return MPDE_OK;
}
size_t mpde_leap_workspace_bytes(const mpde_net_desc* net, const mpde_pde_desc* pde, const mpde_leap_cfg* cfg, int T) {
  LeapWs w;
  if (leap_ws(net, pde, cfg, T, &w)) return 0;
  return w.bytes;
}

int mpde_leap_outer(const mpde_net_desc* net, const mpde_pde_desc* pde, const mpde_leap_cfg* cfg, int T,
                    const float* theta0, const mpde_points* loss0_points, const mpde_points* step_points,
                    const mpde_points* new_points, const float* pde_params, float grad_scale, float* meta_grad,
                    float* losses, float* theta_final, void* workspace, size_t workspace_bytes, void* stream) {
  LeapWs w;
  NvtxRange nvtx("mpde:leap_outer");
  int rc = leap_ws(net, pde, cfg, T, &w);
  if (rc) return rc;
  if (!theta0 || !loss0_points || !step_points || !new_points || !pde_params || !meta_grad || !losses) return MPDE_INVALID_ARG;
  if (!workspace || workspace_bytes < w.bytes) return MPDE_WORKSPACE_TOO_SMALL;
  const int K = cfg->inner_steps, P = mpde_param_count(net);
  const long long TP = (long long)T * P;
  char* wsb = (char*)workspace;
  wsb += (256 - ((uintptr_t)wsb & 255)) & 255;
  void* eval_ws = wsb;
  Carver c(wsb + w.eval_bytes, workspace_bytes - w.eval_bytes);
  float* th_a = c.take(w.n_tp); float* th_b = c.take(w.n_tp); float* grad = c.take(w.n_tp); float* accum = c.take(w.n_tp);
  float* m = c.take(w.n_tp); float* v = c.take(w.n_tp); float* lossk = c.take(w.n_T); float* dummy = c.take(w.n_T);
  (void)dummy;
  // loss0 (leap.py:126)
  rc = run_eval(net, pde, T, MODE_LOSS, theta0, 0, nullptr, 0, loss0_points, pde_params, losses, K + 1, nullptr, 0, nullptr,
                nullptr, nullptr, 0, eval_ws, w.eval_bytes, stream);
  if (rc) return rc;
  rc = launch_simple(fill_kernel, grid_for(TP), 1, 256, 0, stream, accum, 0.0f, TP); if (rc) return rc;
  if (cfg->inner_opt == 1) {
    rc = launch_simple(fill_kernel, grid_for(TP), 1, 256, 0, stream, m, 0.0f, TP); if (rc) return rc;
    rc = launch_simple(fill_kernel, grid_for(TP), 1, 256, 0, stream, v, 0.0f, TP); if (rc) return rc;
  }
  const float* th_k = theta0; long long st_k = 0;
  float* th_next = th_a;
  for (int k = 0; k < K; ++k) {
    rc = run_eval(net, pde, T, MODE_GRAD, th_k, st_k, nullptr, 0, &step_points[k], pde_params, lossk, 1, nullptr, 0, nullptr,
                  grad, nullptr, 1, eval_ws, w.eval_bytes, stream);
    if (rc) return rc;
    LeapStepParams sp;
    sp.P = P; sp.theta = th_k; sp.theta_stride = st_k; sp.grad = grad; sp.theta_next = th_next; sp.m = m; sp.v = v;
    sp.inner_opt = cfg->inner_opt; sp.t_step = k + 1; sp.lr = cfg->inner_lr; sp.clip = cfg->inner_grad_clip;
    sp.b1 = cfg->beta1; sp.b2 = cfg->beta2; sp.eps = cfg->eps;
    rc = launch_simple(leap_inner_update_kernel, T, 1, 1024, 1024 * sizeof(float), stream, sp);
    if (rc) return rc;
    // new_loss on fresh points (leap.py:93) -> losses[:, k+1]
    rc = run_eval(net, pde, T, MODE_LOSS, th_next, P, nullptr, 0, &new_points[k], pde_params, losses + (k + 1), K + 1, nullptr,
                  0, nullptr, nullptr, nullptr, 0, eval_ws, w.eval_bytes, stream);
    if (rc) return rc;
    LeapIncParams ip;
    ip.P = P; ip.theta = th_k; ip.theta_stride = st_k; ip.theta_next = th_next; ip.grad = grad;
    ip.loss = lossk; ip.loss_stride = 1; ip.new_loss = losses + (k + 1); ip.new_loss_stride = K + 1; ip.accum = accum;
    ip.norm = cfg->norm; ip.loss_in_distance = cfg->loss_in_distance; ip.stabilize = cfg->stabilize;
    rc = launch_simple(leap_increment_kernel, T, 1, 1024, 1024 * sizeof(float), stream, ip);
    if (rc) return rc;
    th_k = th_next; st_k = P;
    th_next = (th_next == th_a) ? th_b : th_a;
  }
  if (theta_final) {
    rc = launch_simple(axpby_kernel, grid_for(TP), 1, 256, 0, stream, theta_final, th_k, 1.0f, 0.0f, TP);
    if (rc) return rc;
  }
  return launch_simple(task_sum_kernel, grid_for(P), 1, 256, 0, stream, meta_grad, (const float*)accum, T, (long long)P, grad_scale);
}

}  // extern "C"
