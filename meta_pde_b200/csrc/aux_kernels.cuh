// aux_kernels.cuh -- parameter-sized kernels around the fused jet kernels:
//   prep      natural theta -> "theta-hat": io scales folded into the first/last layer,
//             hidden kernels stored column-permuted (+ transposed copy) for the tile contractions
//   reduce    deterministic fixed-order sum of the per-CTA partial gradients
//   chain     chain rule from theta-hat back to theta (log_input_scale / log_output_scale,
//             src/nets/field.py:160-164,193-197)
//   finalize  per-task loss terms and the driver loss (src/maml_pde.py:76-85)
//   K4        maml_inner_step tail (src/nets/maml.py:72-96) and its VJP,
//             LEAP increment (src/nets/leap.py:190-228), Adam/SGD steps, norms.
// All memory-bound, tiny next to the jet kernels.
#pragma once
#include "compat.cuh"
#include "pde_kernels.cuh"

namespace mpde {

struct PrepParams {
  int W, L, OUT, log_scale, P, PH, T;
  const float* theta; long long theta_stride;
  const float* v;     long long v_stride;        // optional direction
  float* th; float* thd;                          // [T][PH]
};

// grid (blocks, T), 256 threads
MPDE_GLOBAL void prep_kernel(PrepParams p) {
  const int t = blockIdx.y;
  const float* th_n = p.theta + (size_t)t * p.theta_stride;
  const float* v_n = p.v ? p.v + (size_t)t * p.v_stride : nullptr;
  float* out = p.th + (size_t)t * p.PH;
  float* outd = p.thd ? p.thd + (size_t)t * p.PH : nullptr;
  const int W = p.W, L = p.L, OUT = p.OUT, NTX = W / 4;
  const int lin = nat_off_out(W, L) + W * OUT + OUT;   // log_input_scale
  const int lout = lin + 2;                             // log_output_scale
  const int oo_ph = ph_off_out(W, L), oo_nat = nat_off_out(W, L);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < p.PH; i += gridDim.x * blockDim.x) {
    float val = 0.f, vald = 0.f;
    if (i < 2 * W) {
      const int r = i / W;
      const float s = p.log_scale ? expf(th_n[lin + r]) : 1.f;
      val = s * th_n[i];
      if (v_n) vald = s * (v_n[i] + (p.log_scale ? th_n[i] * v_n[lin + r] : 0.f));
    } else if (i < 3 * W) {
      val = th_n[i];
      if (v_n) vald = v_n[i];
    } else if (i < oo_ph) {
      const int per = 2 * W * W + W;
      const int l = 1 + (i - 3 * W) / per;
      const int j = (i - 3 * W) % per;
      const int nb = nat_off_layer(W, l);
      int src;
      if (j < W * W) {                 // column-permuted kernel: Wp[k][tx*4+jj] = W[k][tx + NTX*jj]
        const int k = j / W, c = j % W;
        src = nb + k * W + (c / 4) + NTX * (c % 4);
      } else if (j < 2 * W * W) {      // column-permuted transpose: WTp[n][tx*4+jj] = W[tx + NTX*jj][n]
        const int n = (j - W * W) / W, c = (j - W * W) % W;
        src = nb + ((c / 4) + NTX * (c % 4)) * W + n;
      } else {
        src = nb + W * W + (j - 2 * W * W);
      }
      val = th_n[src];
      if (v_n) vald = v_n[src];
    } else if (i < oo_ph + W * OUT + OUT) {
      const int q = i - oo_ph;
      const int o = (q < W * OUT) ? (q % OUT) : (q - W * OUT);
      const float e = p.log_scale ? expf(th_n[lout + o]) : 1.f;
      val = e * th_n[oo_nat + q];
      if (v_n) vald = e * (v_n[oo_nat + q] + (p.log_scale ? th_n[oo_nat + q] * v_n[lout + o] : 0.f));
    }
    out[i] = val;
    if (outd) outd[i] = vald;
  }
}

struct ReduceParams {
  int P, G;
  const float* part_a; float* out_a; long long out_a_stride;   // [T][G][P] -> [T][P]
  const float* part_b; float* out_b; long long out_b_stride;   // optional second array
};
// grid (ceil(P/256), T)
MPDE_GLOBAL void reduce_partials_kernel(ReduceParams p) {
  const int t = blockIdx.y;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= p.P) return;
  {
    const float* src = p.part_a + (size_t)t * p.G * p.P + i;
    float s = 0.f;
    for (int c = 0; c < p.G; ++c) s += src[(size_t)c * p.P];
    p.out_a[(size_t)t * p.out_a_stride + i] = s;
  }
  if (p.part_b) {
    const float* src = p.part_b + (size_t)t * p.G * p.P + i;
    float s = 0.f;
    for (int c = 0; c < p.G; ++c) s += src[(size_t)c * p.P];
    p.out_b[(size_t)t * p.out_b_stride + i] = s;
  }
}

// deterministic block sum (blockDim.x a power of two <= 1024); result valid in every thread
MPDE_DEV float block_sum(float v, float* sh) {
  const int tid = threadIdx.x;
  __syncthreads();
  sh[tid] = v;
  __syncthreads();
  for (int s = blockDim.x >> 1; s > 0; s >>= 1) {
    if (tid < s) sh[tid] += sh[tid + s];
    __syncthreads();
  }
  const float r = sh[0];
  __syncthreads();
  return r;
}

struct ChainParams {
  int W, L, OUT, log_scale, P, PH;
  const float* theta; long long theta_stride;
  const float* v; long long v_stride;             // HVP only
  const float* th; long long th_stride;           // prepped theta-hat
  const float* thd; long long thd_stride;         // prepped direction-hat (HVP)
  float* g; long long g_stride;                   // in: effective gradient, out: gradient
  float* h; long long h_stride;                   // in: effective H v,      out: H v   (optional)
};
// grid T, 256 threads, smem 256 floats
MPDE_GLOBAL void chain_kernel(ChainParams p) {
  float* sh = reinterpret_cast<float*>(mpde_smem());
  const int t = blockIdx.x, tid = threadIdx.x, nth = blockDim.x;
  const int W = p.W, L = p.L, OUT = p.OUT;
  float* g = p.g + (size_t)t * p.g_stride;
  float* h = p.h ? p.h + (size_t)t * p.h_stride : nullptr;
  const int oo = nat_off_out(W, L);
  const int lin = oo + W * OUT + OUT, lout = lin + 2;
  if (!p.log_scale) return;   // theta-hat == theta apart from the layout: nothing to do
  const float* theta = p.theta + (size_t)t * p.theta_stride;
  const float* v = h ? p.v + (size_t)t * p.v_stride : nullptr;
  const float* th = p.th + (size_t)t * p.th_stride;
  const float* thd = h ? p.thd + (size_t)t * p.thd_stride : nullptr;
  const int oo_ph = ph_off_out(W, L);
  float glin[2], hlin[2], glo[2], hlo[2];
  for (int r = 0; r < 2; ++r) {
    float a = 0.f, b = 0.f;
    for (int n = tid; n < W; n += nth) {
      a += th[r * W + n] * g[r * W + n];
      if (h) b += th[r * W + n] * h[r * W + n] + thd[r * W + n] * g[r * W + n];
    }
    glin[r] = block_sum(a, sh);
    hlin[r] = h ? block_sum(b, sh) : 0.f;
  }
  for (int o = 0; o < OUT; ++o) {
    float a = 0.f, b = 0.f;
    for (int n = tid; n <= W; n += nth) {   // n == W: the bias entry
      const int q = (n < W) ? n * OUT + o : W * OUT + o;
      a += th[oo_ph + q] * g[oo + q];
      if (h) b += th[oo_ph + q] * h[oo + q] + thd[oo_ph + q] * g[oo + q];
    }
    glo[o] = block_sum(a, sh);
    hlo[o] = h ? block_sum(b, sh) : 0.f;
  }
  // rescale the first / last layer blocks (H v first: it needs the effective gradient)
  for (int i = tid; i < 2 * W; i += nth) {
    const int r = i / W;
    const float s = expf(theta[lin + r]);
    const float ge = g[i];
    if (h) h[i] = s * (h[i] + v[lin + r] * ge);
    g[i] = s * ge;
  }
  for (int q = tid; q < W * OUT + OUT; q += nth) {
    const int o = (q < W * OUT) ? (q % OUT) : (q - W * OUT);
    const float e = expf(theta[lout + o]);
    const float ge = g[oo + q];
    if (h) h[oo + q] = e * (h[oo + q] + v[lout + o] * ge);
    g[oo + q] = e * ge;
  }
  if (tid < 2) { g[lin + tid] = glin[tid]; if (h) h[lin + tid] = hlin[tid]; }
  if (tid < OUT) { g[lout + tid] = glo[tid]; if (h) h[lout + tid] = hlo[tid]; }
}

struct LossFinParams {
  int G, n_seg, n_terms;
  int term_of_seg[MAX_SEG];
  float term_scale[MAX_SEG];   // term = term_scale * sum / n_elems
  float inv_elems[MAX_SEG];
  float weight[MAX_SEG];       // loss += weight * term
  const float* partial_terms;  // [T][G][MAX_SEG]
  float* loss; float* terms;   // [T], [T][n_terms]  (either may be null)
  long long loss_stride, terms_stride;   // element strides between tasks
};
// grid ceil(T/128), 128 threads: one thread per task
MPDE_GLOBAL void loss_finalize_kernel(LossFinParams p, int T) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  float loss = 0.f;
  for (int s = 0; s < p.n_seg; ++s) {
    const float* src = p.partial_terms + (size_t)t * p.G * MAX_SEG + s;
    float a = 0.f;
    for (int c = 0; c < p.G; ++c) a += src[(size_t)c * MAX_SEG];
    const float term = p.term_scale[s] * (a * p.inv_elems[s]);
    if (p.terms) p.terms[(size_t)t * p.terms_stride + p.term_of_seg[s]] = term;
    loss += p.weight[s] * term;
  }
  if (p.loss) p.loss[(size_t)t * p.loss_stride] = loss;
}

// ------------------------------------------------------------------ K4
MPDE_HD float softplus_f(float x) { return (x > 20.f) ? x : log1pf(expf(x)); }
MPDE_HD float sigmoid_f(float x) { return 1.f / (1.f + expf(-x)); }

constexpr int INNER_MAX_SPLITS = 64;
struct InnerUpdParams {
  int P;
  const float* theta; long long theta_stride;
  const float* grad;                 // [T][P]
  const float* alpha; float alpha_scalar; int softplus;
  float inner_lr, clip;
  float* theta_next;                 // [T][P]
  float* gnorm; long long gnorm_stride;
  float* part;                       // [T][gridDim.x] partial sums of squares (stage 1 -> stage 2)
};
// Two stages, grid (S, T) x 256 threads each (round 1 ran one 1024-thread CTA per task: 8 CTAs on 148 SMs at 8 tasks per
// GPU, the non-scaling tail of the 8-GPU step).  Chunk s of task t = parameters [s * ceil(P / S), ...); the partials are
// re-added in chunk order by every CTA of stage 2: deterministic, no atomics.   maml.py:72-96
MPDE_DEV void inner_chunk(int P, int& lo, int& hi) {
  const int S = gridDim.x, chunk = (P + S - 1) / S;
  lo = (int)blockIdx.x * chunk;
  hi = lo + chunk < P ? lo + chunk : P;
}
MPDE_DEV float inner_part_total(const float* part, int t) {
  float tot = 0.f;
  for (int s = 0; s < (int)gridDim.x; ++s) tot += part[(size_t)t * gridDim.x + s];
  return tot;
}
MPDE_GLOBAL void maml_inner_norm_kernel(InnerUpdParams p) {
  float* sh = reinterpret_cast<float*>(mpde_smem());
  const int t = blockIdx.y, tid = threadIdx.x, nth = blockDim.x;
  const float* g = p.grad + (size_t)t * p.P;
  const float s_scalar = p.softplus ? softplus_f(p.alpha_scalar) : p.alpha_scalar;
  int lo, hi;
  inner_chunk(p.P, lo, hi);
  float ss = 0.f;
  for (int i = lo + tid; i < hi; i += nth) {
    const float s = p.alpha ? (p.softplus ? softplus_f(p.alpha[i]) : p.alpha[i]) : s_scalar;
    const float gs = g[i] * s;
    ss = fmaf(gs, gs, ss);
  }
  ss = block_sum(ss, sh);
  if (tid == 0) p.part[(size_t)t * gridDim.x + blockIdx.x] = ss;
}
MPDE_GLOBAL void maml_inner_update_kernel(InnerUpdParams p) {
  const int t = blockIdx.y, tid = threadIdx.x, nth = blockDim.x;
  const float* th = p.theta + (size_t)t * p.theta_stride;
  const float* g = p.grad + (size_t)t * p.P;
  float* out = p.theta_next + (size_t)t * p.P;
  const float s_scalar = p.softplus ? softplus_f(p.alpha_scalar) : p.alpha_scalar;
  const float n = sqrtf(inner_part_total(p.part, t));
  int lo, hi;
  inner_chunk(p.P, lo, hi);
  for (int i = lo + tid; i < hi; i += nth) {
    const float s = p.alpha ? (p.softplus ? softplus_f(p.alpha[i]) : p.alpha[i]) : s_scalar;
    float gs = g[i] * s;
    if (n > p.clip) gs = p.clip * gs / n;
    out[i] = th[i] - p.inner_lr * gs;
  }
  __syncthreads();   // part may alias gnorm (mpde_maml_inner_update: one chunk, partial parked in gnorm[t])
  if (tid == 0 && blockIdx.x == 0 && p.gnorm) p.gnorm[(size_t)t * p.gnorm_stride] = n;
}

struct InnerVjpParams {
  int P;
  float* lam;                        // [T][P] in/out: lam += c_add * add  (then used)
  const float* add; float c_add;     // optional
  const float* grad;                 // g_k [T][P]
  const float* alpha; float alpha_scalar; int softplus;
  const float* gnorm; long long gnorm_stride;   // n_k [T]
  float inner_lr, clip;
  float* gbar;                       // [T][P]: direction for the HVP
  float* galpha;                     // optional [T][P]: per-task lr-tree gradient of this step
  float* part;                       // [T][gridDim.x] partial dot products (stage 1 -> stage 2)
};
// VJP of theta_{k+1} = theta_k - eta * clip(softplus(alpha) * g_k)  (SURVEY.md 9.4), two stages, grid (S, T) x 256 threads:
// stage 1 folds `add` into lam and, for clipped tasks, leaves the chunk's part of <g * s, -eta * lam>; stage 2 finishes.
MPDE_GLOBAL void maml_inner_vjp_dot_kernel(InnerVjpParams p) {
  float* sh = reinterpret_cast<float*>(mpde_smem());
  const int t = blockIdx.y, tid = threadIdx.x, nth = blockDim.x;
  float* lam = p.lam + (size_t)t * p.P;
  const float* add = p.add ? p.add + (size_t)t * p.P : nullptr;
  const float* g = p.grad + (size_t)t * p.P;
  const float n = p.gnorm[(size_t)t * p.gnorm_stride];
  const bool clipped = n > p.clip;
  const float s_scalar = p.softplus ? softplus_f(p.alpha_scalar) : p.alpha_scalar;
  int lo, hi;
  inner_chunk(p.P, lo, hi);
  float dot = 0.f;
  for (int i = lo + tid; i < hi; i += nth) {
    float l = lam[i];
    if (add) { l = fmaf(p.c_add, add[i], l); lam[i] = l; }
    if (clipped) {
      const float s = p.alpha ? (p.softplus ? softplus_f(p.alpha[i]) : p.alpha[i]) : s_scalar;
      dot = fmaf(g[i] * s, -p.inner_lr * l, dot);
    }
  }
  dot = block_sum(dot, sh);
  if (tid == 0) p.part[(size_t)t * gridDim.x + blockIdx.x] = dot;
}
MPDE_GLOBAL void maml_inner_vjp_kernel(InnerVjpParams p) {
  const int t = blockIdx.y, tid = threadIdx.x, nth = blockDim.x;
  const float* lam = p.lam + (size_t)t * p.P;
  const float* g = p.grad + (size_t)t * p.P;
  float* gbar = p.gbar + (size_t)t * p.P;
  float* ga = p.galpha ? p.galpha + (size_t)t * p.P : nullptr;
  const float n = p.gnorm[(size_t)t * p.gnorm_stride];
  const bool clipped = n > p.clip;
  const float dot = clipped ? inner_part_total(p.part, t) : 0.f;
  const float cn = clipped ? p.clip / n : 1.f;
  const float dn = clipped ? dot / (n * n) : 0.f;
  int lo, hi;
  inner_chunk(p.P, lo, hi);
  for (int i = lo + tid; i < hi; i += nth) {
    const float a = p.alpha ? p.alpha[i] : p.alpha_scalar;
    const float s = p.softplus ? softplus_f(a) : a;
    const float ds = p.softplus ? sigmoid_f(a) : 1.f;
    float u = -p.inner_lr * lam[i];
    if (clipped) u = cn * (u - g[i] * s * dn);
    if (ga) ga[i] = ds * g[i] * u;
    gbar[i] = s * u;
  }
}

// y[i] = a * x[i] + b * y[i]
MPDE_GLOBAL void axpby_kernel(float* y, const float* x, float a, float b, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    y[i] = (b == 0.f) ? a * x[i] : fmaf(a, x[i], b * y[i]);   // b == 0: plain scaled copy (y may be uninitialised)
}

// out[i] = scale * sum_t x[t][i]   (fixed task order)
MPDE_GLOBAL void task_sum_kernel(float* out, const float* x, int T, long long P, float scale) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < P; i += (long long)gridDim.x * blockDim.x) {
    float s = 0.f;
    for (int t = 0; t < T; ++t) s += x[(size_t)t * P + i];
    out[i] = scale * s;
  }
}

// meta_loss[t] = fold_k ( M[t][k] + decay * meta )   (maml.py:136-139)
MPDE_GLOBAL void meta_loss_kernel(float* meta, const float* M, int T, int K, float decay) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  float m = 0.f;
  for (int k = 0; k < K; ++k) m = M[(size_t)t * K + k] + m * decay;
  meta[t] = m;
}

// sum of squares, two deterministic stages
MPDE_GLOBAL void sumsq_stage1_kernel(const float* x, long long n, float* part) {
  float* sh = reinterpret_cast<float*>(mpde_smem());
  float s = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    s = fmaf(x[i], x[i], s);
  s = block_sum(s, sh);
  if (threadIdx.x == 0) part[blockIdx.x] = s;
}
MPDE_GLOBAL void sumsq_stage2_kernel(const float* part, int nb, float* out) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    float s = 0.f;
    for (int i = 0; i < nb; ++i) s += part[i];
    out[0] = s;
  }
}

MPDE_GLOBAL void adam_kernel(float* x, const float* g, float* m, float* v, long long n, float lr, float b1, float b2,
                             float eps, float bc1, float bc2, const float* sumsq, float clip) {
  float scale = 1.f, nrm = 0.f;
  bool clipped = false;
  if (sumsq) { nrm = sqrtf(sumsq[0]); clipped = nrm > clip; }
  (void)scale;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float gi = g[i];
    if (clipped) gi = clip * gi / nrm;
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi; v[i] = vi;
    x[i] = x[i] - lr * (mi / bc1) / (sqrtf(vi / bc2) + eps);
  }
}

MPDE_GLOBAL void sgd_kernel(float* x, const float* g, long long n, float lr, const float* sumsq, float clip) {
  float nrm = 0.f;
  bool clipped = false;
  if (sumsq) { nrm = sqrtf(sumsq[0]); clipped = nrm > clip; }
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float gi = g[i];
    if (clipped) gi = clip * gi / nrm;
    x[i] = x[i] - lr * gi;
  }
}

// Fused outer step: ||g_theta||_2, clip of the WHOLE packed tree by that norm, optimizer step on
// theta and Adam step on the lr tree in ONE launch (src/maml_pde.py:210-225; src/leap_pde.py:180-194 and
// src/nn_pde_maml.py:150-169 with n_lrs = 0).  Every CTA recomputes the norm of g_theta in the same fixed order
// (thread-strided partial sums, then the block tree): identical bits in every CTA and on every rank, no grid-wide
// barrier, no float atomics.  g is packed [g_theta (n_theta) | g_lrs (n_lrs)], the layout of the NCCL all-reduce.
struct OuterStepParams {
  long long n_theta, n_lrs;
  float* theta; float* lrs;
  const float* g;
  float* m_theta; float* v_theta; float* m_lrs; float* v_lrs;
  int opt;                                   // 0: SGD (Momentum beta 0), 1: Adam on theta
  float lr_theta, b1, b2, eps, bc1, bc2;     // theta: flax Adam, bias corrections 1 - b^t
  float lr_lrs, b1l, b2l, epsl, bc1l, bc2l;  // lr tree: jax.experimental.optimizers.adam at its own step index
  float clip;
  float* sumsq_out;                          // [1]: sum(g_theta^2) (meta_grad_norm^2), written by CTA 0
};
MPDE_GLOBAL void outer_step_kernel(OuterStepParams p) {
  float* sh = reinterpret_cast<float*>(mpde_smem());
  float ss = 0.f;
  for (long long i = threadIdx.x; i < p.n_theta; i += blockDim.x) ss = fmaf(p.g[i], p.g[i], ss);
  ss = block_sum(ss, sh);
  if (blockIdx.x == 0 && threadIdx.x == 0 && p.sumsq_out) p.sumsq_out[0] = ss;
  const float nrm = sqrtf(ss);
  const bool clipped = nrm > p.clip;
  const long long n = p.n_theta + p.n_lrs;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float gi = p.g[i];
    if (clipped) gi = p.clip * gi / nrm;
    if (i < p.n_theta) {
      if (p.opt == 1) {
        const float mi = p.b1 * p.m_theta[i] + (1.f - p.b1) * gi;
        const float vi = p.b2 * p.v_theta[i] + (1.f - p.b2) * gi * gi;
        p.m_theta[i] = mi; p.v_theta[i] = vi;
        p.theta[i] = p.theta[i] - p.lr_theta * (mi / p.bc1) / (sqrtf(vi / p.bc2) + p.eps);
      } else {
        p.theta[i] = p.theta[i] - p.lr_theta * gi;
      }
    } else {
      const long long j = i - p.n_theta;
      const float mi = p.b1l * p.m_lrs[j] + (1.f - p.b1l) * gi;
      const float vi = p.b2l * p.v_lrs[j] + (1.f - p.b2l) * gi * gi;
      p.m_lrs[j] = mi; p.v_lrs[j] = vi;
      p.lrs[j] = p.lrs[j] - p.lr_lrs * (mi / p.bc1l) / (sqrtf(vi / p.bc2l) + p.epsl);
    }
  }
}

// LEAP inner step for T tasks (leap.py:75-102,190-228): one block per task.
struct LeapStepParams {
  int P;
  const float* theta; long long theta_stride;   // theta_k
  float* grad;                                   // [T][P] in: raw grad, out: clipped grad
  float* theta_next;                             // [T][P]
  float* m; float* v;                            // Adam state [T][P] (inner_opt == 1)
  int inner_opt, t_step;                         // t_step = k + 1
  float lr, clip, b1, b2, eps;
};
MPDE_GLOBAL void leap_inner_update_kernel(LeapStepParams p) {
  float* sh = reinterpret_cast<float*>(mpde_smem());
  const int t = blockIdx.x, tid = threadIdx.x, nth = blockDim.x;
  const float* th = p.theta + (size_t)t * p.theta_stride;
  float* g = p.grad + (size_t)t * p.P;
  float* out = p.theta_next + (size_t)t * p.P;
  float ss = 0.f;
  for (int i = tid; i < p.P; i += nth) ss = fmaf(g[i], g[i], ss);
  const float n = sqrtf(block_sum(ss, sh));
  const bool clipped = n > p.clip;
  const float bc1 = 1.f - powf(p.b1, (float)p.t_step), bc2 = 1.f - powf(p.b2, (float)p.t_step);
  for (int i = tid; i < p.P; i += nth) {
    float gi = g[i];
    if (clipped) { gi = p.clip * gi / n; g[i] = gi; }
    if (p.inner_opt == 1) {
      float* m = p.m + (size_t)t * p.P; float* v = p.v + (size_t)t * p.P;
      const float mi = p.b1 * m[i] + (1.f - p.b1) * gi;
      const float vi = p.b2 * v[i] + (1.f - p.b2) * gi * gi;
      m[i] = mi; v[i] = vi;
      out[i] = th[i] - p.lr * (mi / bc1) / (sqrtf(vi / bc2) + p.eps);
    } else {
      out[i] = th[i] - p.lr * gi;
    }
  }
}

struct LeapIncParams {
  int P;
  const float* theta; long long theta_stride;   // theta_k
  const float* theta_next;                       // [T][P]
  const float* grad;                             // clipped grad [T][P]
  const float* loss; long long loss_stride;      // L(theta_k; k1 points)    [T]
  const float* new_loss; long long new_loss_stride;  // L(theta_{k+1}; k2 points) [T]
  float* accum;                                  // [T][P] += increment
  int norm, loss_in_distance, stabilize;
};
MPDE_GLOBAL void leap_increment_kernel(LeapIncParams p) {
  float* sh = reinterpret_cast<float*>(mpde_smem());
  const int t = blockIdx.x, tid = threadIdx.x, nth = blockDim.x;
  const float* th = p.theta + (size_t)t * p.theta_stride;
  const float* tn = p.theta_next + (size_t)t * p.P;
  const float* g = p.grad + (size_t)t * p.P;
  float* acc = p.accum + (size_t)t * p.P;
  float d_loss = p.new_loss[(size_t)t * p.new_loss_stride] - p.loss[(size_t)t * p.loss_stride];
  if (p.stabilize) d_loss = -fabsf(d_loss);
  float nrm = 1.f;
  if (p.norm) {
    float ss = 0.f;
    for (int i = tid; i < p.P; i += nth) { const float d = tn[i] - th[i]; ss = fmaf(d, d, ss); }
    ss = block_sum(ss, sh);
    if (p.loss_in_distance) ss += d_loss * d_loss;
    nrm = sqrtf(ss);
  }
  for (int i = tid; i < p.P; i += nth) {
    float inc = th[i] - tn[i];
    if (p.loss_in_distance) inc -= d_loss * g[i];
    acc[i] += inc / nrm;
  }
}

MPDE_GLOBAL void fill_kernel(float* x, float val, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) x[i] = val;
}
// dst[t*dstride + j] = src[t*sstride]  (scatter per-task scalars into [T][K+1] tables)
MPDE_GLOBAL void scatter_scalar_kernel(float* dst, long long dstride, const float* src, long long sstride, int T) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < T) dst[(size_t)t * dstride] = src[(size_t)t * sstride];
}

}  // namespace mpde
