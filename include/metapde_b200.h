/* metapde_b200.h -- C ABI of libmetapde_b200.so (B200 / sm_100a).
 *
 * The drop-in boundary for meta-pde's data-parallel hot path.  The reference
 * (sunnytqin/meta-pde) is pure Python on JAX/Flax and has NO FFI of its own
 * (SURVEY.md 8(b)); every entry point below therefore cites the reference
 * *Python* interface whose arithmetic it replaces (paths relative to the
 * reference root).  Plain pointers and sizes only: no torch / JAX types.
 *
 * Conventions
 *   - all arrays are float32, device memory, written in full by the callee;
 *   - the caller owns every buffer, including the workspace
 *     (size from the matching *_workspace_bytes); the library keeps no device
 *     allocation between calls;
 *   - every call is stream-ordered on `stream` (a cudaStream_t passed as
 *     void*), never synchronises, never allocates: CUDA-graph capturable;
 *   - return value: MPDE_OK or an error code, never an exception; NaNs
 *     propagate as values (the reference's NaN stop, src/maml_pde.py:260-262,
 *     stays a host decision).
 *
 * Flat parameter layout `theta` (P = mpde_param_count floats), matching the
 * Flax tree of src/nets/field.py:146-199:
 *     for l = 0..L   : Dense_l/kernel [n_in, n_out] row-major, Dense_l/bias [n_out]
 *     if log_scale   : log_input_scale [in_dim], log_output_scale [out_dim]
 *   (L = num_layers hidden layers of layer_size, then the out_dim linear layer.)
 *
 * Point sets and loss terms per PDE (order of `mpde_points.set[]` / `terms[]`):
 *   MPDE_PDE_POISSON          sets: boundary, domain
 *                             terms: boundary_loss, domain_loss            (src/poisson/poisson_common.py:26-39)
 *   MPDE_PDE_TD_BURGERS       sets: left, right, initial, domain
 *                             terms: loss_initial, loss_left, loss_right, loss_domain
 *                                                                          (src/burgers/td_burgers_common.py:70-91)
 *   MPDE_PDE_HYPER_ELASTICITY sets: top, bottom, domain  (left/right/holes are sampled but unused by the loss)
 *                             terms: loss_bottom, loss_top, loss_domain    (src/elasticity/hyper_elasticity_common.py:69-86)
 * pde_params per task (mpde_num_pde_params floats):
 *   poisson:  source_params[2][3], bc_params[5], geo_params[2]   (13)
 *   burgers:  source_params[1] (Reynolds), ic_params[2]          (3)
 *   hyper:    source_params[2], bc_params[2] (E = bc_params[0])  (4)
 * loss = bc_weight * sum(boundary terms) + sum(domain terms)     (src/maml_pde.py:76-85)
 */
#ifndef METAPDE_B200_H_
#define METAPDE_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum mpde_status {
  MPDE_OK = 0,
  MPDE_INVALID_ARG = 1,        /* null pointer, non-positive size, bad enum           */
  MPDE_UNSUPPORTED_NET = 2,    /* width not in {32,64,128,256}, in_dim != 2, out_dim mismatch, unknown nonlinearity */
  MPDE_WORKSPACE_TOO_SMALL = 3,
  MPDE_CUDA_ERROR = 4,
  MPDE_STATIC_SCRATCH_EXCEEDED = 5   /* mpde_field_eval_static: the problem needs more than MPDE_STATIC_SCRATCH_BYTES */
};

enum mpde_pde_kind { MPDE_PDE_POISSON = 0, MPDE_PDE_TD_BURGERS = 1, MPDE_PDE_HYPER_ELASTICITY = 2 };
enum mpde_mode { MPDE_MODE_LOSS = 0, MPDE_MODE_GRAD = 1, MPDE_MODE_HVP = 2 };

typedef enum mpde_nonlinearity { MPDE_ACT_SIN = 0, MPDE_ACT_SWISH = 1 } mpde_nonlinearity;

/* The SIREN constructor arguments of src/maml_pde.py:112-119 that reach the arithmetic. */
typedef struct mpde_net_desc {
  int32_t num_layers;  /* FLAGS.num_layers  (hidden Dense+sin layers)      */
  int32_t layer_size;  /* FLAGS.layer_size                                 */
  int32_t in_dim;      /* 2 (src/maml_pde.py:124 hard-wires (1,2))         */
  int32_t out_dim;     /* 1: NeuralPotential, 2: make_nf_ndim(2)           */
  float omega;         /* FLAGS.siren_omega (applied on every hidden layer, src/nets/field.py:186) */
  int32_t log_scale;   /* FLAGS.log_scale                                  */
  int32_t nonlinearity; /* mpde_nonlinearity: 0 = np.sin (FLAGS.siren, the default), 1 = flax.nn.swish (`--nosiren`,
                           src/maml_pde.py:115, src/nets/field.py:187-188: no omega; FP32-FMA kernels at every width) */
} mpde_net_desc;

typedef struct mpde_pde_desc {
  int32_t kind;         /* mpde_pde_kind                                   */
  float bc_weight;      /* FLAGS.bc_weight                                 */
  int32_t n_points[4];  /* points per task in each set (order above)       */
} mpde_pde_desc;

typedef struct mpde_points {
  const float* set[4];      /* [T][n_points[s]][2]                                      */
  int64_t task_stride[4];   /* floats between consecutive tasks; 0 = one set shared by all tasks */
} mpde_points;

/* ---- sizes ------------------------------------------------------------- */
int mpde_param_count(const mpde_net_desc* net);
int mpde_num_point_sets(int pde_kind);
int mpde_num_terms(int pde_kind);
int mpde_num_pde_params(int pde_kind);
const char* mpde_status_string(int status);
/* workspace for mpde_loss / mpde_loss_grad / mpde_loss_hvp (mode = mpde_mode).
 * Widths 32/64 (FP32-FMA kernels): per-CTA scratch and partial gradients, O(CTAs * P).
 * Widths 128/256 (tensor-core kernels): additionally the activation records and the zbar (zbardot) spills of every
 * point of every task, read once by the weight-gradient contraction: about 0.93 GB per task in GRAD mode and 1.8 GB per
 * task in HVP mode at 5x256 with 16384 + 16384 points (60 / 115 GB at 64 tasks: sized for the 180 GB of a B200; shard
 * the tasks over more GPUs or call with fewer tasks per launch when that does not fit). */
size_t mpde_workspace_bytes(const mpde_net_desc* net, const mpde_pde_desc* pde, int T, int mode);

/* ---- K1: loss only.  Replaces pde.loss_fn + the driver wrapper
 * (src/poisson/poisson_common.py:26-39, src/burgers/td_burgers_common.py:47-91,
 *  src/elasticity/hyper_elasticity_common.py:23-86, src/maml_pde.py:76-85) for T tasks at once
 * (the vmap of src/nets/maml.py:226-228).
 *   theta            [T][P] with theta_task_stride floats between tasks (0: one theta for all tasks)
 *   loss             [T]
 *   terms            [T][mpde_num_terms]            (the aux dict values)
 *   domain_residual  optional [T][n_domain]: per-point residual r (Poisson, Burgers) or energy density psi (hyper) */
int mpde_loss(const mpde_net_desc* net, const mpde_pde_desc* pde, int T, const float* theta,
              int64_t theta_task_stride, const mpde_points* points, const float* pde_params, float* loss,
              float* terms, float* domain_residual, void* workspace, size_t workspace_bytes, void* stream);

/* ---- K2: loss and d loss / d theta.  Replaces jax.value_and_grad(loss)
 * (src/nets/maml.py:66-70, src/nets/leap.py:71-73, src/nn_pde_maml.py:142-144).  grad [T][P]. */
int mpde_loss_grad(const mpde_net_desc* net, const mpde_pde_desc* pde, int T, const float* theta,
                   int64_t theta_task_stride, const mpde_points* points, const float* pde_params, float* loss,
                   float* terms, float* grad, void* workspace, size_t workspace_bytes, void* stream);

/* ---- K3: Hessian-vector product H(theta) v (and optionally the gradient).  Replaces the
 * second-order part of jax.value_and_grad(task_rollout_and_eval) (src/nets/maml.py:195-202).
 *   v [T][P] (v_task_stride floats between tasks), grad optional [T][P], hv [T][P]. */
int mpde_loss_hvp(const mpde_net_desc* net, const mpde_pde_desc* pde, int T, const float* theta,
                  int64_t theta_task_stride, const float* v, int64_t v_task_stride, const mpde_points* points,
                  const float* pde_params, float* loss, float* terms, float* grad, float* hv, void* workspace,
                  size_t workspace_bytes, void* stream);

/* ---- K1f: plain field evaluation model(coords) (src/maml_pde.py:190, src/util/trainer_util.py:508-510).
 *   coords [T][n][2] (coords_task_stride floats between tasks, 0: shared), out [T][n][out_dim]. */
size_t mpde_field_eval_workspace_bytes(const mpde_net_desc* net, int T);
int mpde_field_eval(const mpde_net_desc* net, int T, const float* theta, int64_t theta_task_stride,
                    const float* coords, int64_t coords_task_stride, int n, float* out, void* workspace,
                    size_t workspace_bytes, void* stream);

/* ---- Samplers: jax.random on the GPU.  Replaces the jax.random calls inside the reference's point
 *   samplers (src/poisson/poisson_common.py:69-119, src/burgers/td_burgers_common.py:117-186): jax 0.2.20's threefry2x32
 *   `split` / `_random_bits` / `uniform`, bit-exact with meta_pde_b200/compat/random.py (pinned by the reference's jax.random
 *   known answers).  keys: device uint32 [S][2], one key per sample; every output row s uses keys[s].
 *   mpde_random_bits / _uniform: out [S][n] = random bits / uniform(keys[s], (n,), minval, maxval).
 *   mpde_random_split: out [S][num][2] = split(keys[s], num).
 *   mpde_sample_poisson_candidates: sample_points(keys[s], n, params) of poisson_common.py up to the final selection:
 *     boundary [S][n][2] (final), cand_xy [S][3n][2] and cand_g [S][3n] = the candidates of sample_points_in_domain and their
 *     Gumbel selection keys (+inf inside the hole); the domain points are the n candidates with the smallest cand_g, in
 *     ascending (stable) order -- jax.random.choice(replace=False, p).  geo [S][2] = the task's geometry params (c1, c2). */
int mpde_random_bits(int S, const uint32_t* keys, int64_t n, uint32_t* out, void* stream);
int mpde_random_uniform(int S, const uint32_t* keys, int64_t n, float minval, float maxval, float* out, void* stream);
int mpde_random_split(int S, const uint32_t* keys, int num, uint32_t* out, void* stream);
int mpde_sample_poisson_candidates(int S, const uint32_t* keys, int n, const float* geo, float xmin, float xmax, float ymin,
                                   float ymax, float* boundary, float* cand_xy, float* cand_g, void* stream);

/* ---- K1f without allocation.  NO COUNTERPART IN THE REFERENCE: added for callers
 *   that cannot allocate device memory (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 *   Same arguments and results as mpde_field_eval, minus the workspace: the scratch (theta-hat + per-CTA partial terms,
 *   mpde_field_eval_workspace_bytes) lives in a file-scope `static __device__` array of MPDE_STATIC_SCRATCH_BYTES (8 MiB)
 *   that every GPU loading the library holds until the process exits.  Calls no cudaMalloc / cudaFree / allocating routine.
 *   Limit: mpde_field_eval_workspace_bytes(net, T) <= MPDE_STATIC_SCRATCH_BYTES (about 60 tasks of a 3x64 field, 3 of a
 *   5x256 one), otherwise MPDE_STATIC_SCRATCH_EXCEEDED.  All calls on a device SHARE the array: calls that may overlap on
 *   the GPU (different streams, concurrent graph launches) race on it -- serialise them or use mpde_field_eval.  Nothing in
 *   the array needs clearing between calls (every entry read by a call is written by it first). */
#define MPDE_STATIC_SCRATCH_BYTES (8u << 20)
size_t mpde_static_scratch_bytes(void);
int mpde_field_eval_static(const mpde_net_desc* net, int T, const float* theta, int64_t theta_task_stride,
                           const float* coords, int64_t coords_task_stride, int n, float* out, void* stream);

/* ---- K4: parameter-sized fused ops ---------------------------------------
 * maml_inner_step after the gradient (src/nets/maml.py:72-96):
 *   g' = grad * (softplus ? softplus(alpha) : alpha)   alpha: [P] lr-tree row shared by tasks, or NULL: scalar alpha_scalar
 *   n  = ||g'||_2 per task; if n > clip: g' *= clip / n;  theta_next = theta - inner_lr * g'  (Momentum beta=0)
 *   gnorm [T] (required) receives n.  theta may alias nothing; theta_task_stride 0 = shared theta. */
int mpde_maml_inner_update(int T, int P, const float* theta, int64_t theta_task_stride, const float* grad,
                           const float* alpha, float alpha_scalar, int softplus_lrs, float inner_lr, float clip,
                           float* theta_next, float* gnorm, void* stream);

/* Global-norm clip + Adam (flax.optim.Adam / jax.experimental.optimizers.adam, bias-corrected, eps outside sqrt):
 * used by the outer train steps (src/maml_pde.py:210-225, src/leap_pde.py:180-194, src/nn_pde_maml.py:150-169)
 * and the LEAP inner optimiser (src/leap_pde.py:87).
 *   sumsq = sum(x^2) -> out[0]  (device scalar) */
int mpde_sumsq(const float* x, int64_t n, float* out, void* workspace, size_t workspace_bytes, void* stream);
size_t mpde_sumsq_workspace_bytes(int64_t n);
/*   g_eff = g * (sqrt(*sumsq) > clip ? clip / sqrt(*sumsq) : 1)   (sumsq may be NULL: no clip)
 *   m = b1 m + (1-b1) g_eff ; v = b2 v + (1-b2) g_eff^2 ; x -= lr * (m / (1-b1^t)) / (sqrt(v / (1-b2^t)) + eps) */
int mpde_adam_step(float* x, const float* g, float* m, float* v, int64_t n, int t, float lr, float b1, float b2,
                   float eps, const float* sumsq, float clip, void* stream);
/*   x -= lr * g_eff   (flax.optim.Momentum beta=0 / "sgd", src/util/trainer_util.py:606-608) */
int mpde_sgd_step(float* x, const float* g, int64_t n, float lr, const float* sumsq, float clip, void* stream);

/* out[i] = scale * SUM_t x[t][i] in fixed task order: the task mean of src/nets/maml.py:231 / src/nets/leap.py:186
 * (scale = 1/T_global on every rank, then the ranks all-reduce) and the task SUM of the fine-tune batch loss
 * (src/nn_pde_maml.py:110-124, scale = 1). */
int mpde_task_sum(float* out, const float* x, int T, int64_t P, float scale, void* stream);

/* Fused outer step: one launch replaces the reference's
 *   meta_grad_norm = sqrt(sum(g_theta^2));  (g_theta, g_lrs) *= clip / norm  if norm > clip   (the lr-tree grads are
 *   rescaled by the MODEL-grad norm too: src/maml_pde.py:210-222);  optimizer.apply_gradient(g_theta)
 *   (flax Adam, or Momentum beta 0 when optimizer == 0; src/util/trainer_util.py:593-612);
 *   lr_opt_update(step, g_lrs, lr_opt_state)  (jax.experimental.optimizers.adam, src/maml_pde.py:150,225).
 * g_packed = [g_theta (n_theta) | g_lrs (n_lrs)] is the buffer the ranks all-reduce.  n_lrs = 0: LEAP
 * (src/leap_pde.py:180-194) and the fine-tune step (src/nn_pde_maml.py:150-169).  t_theta / t_lrs are the 1-based
 * Adam step counts.  sumsq_out [1] (optional) receives sum(g_theta^2) before clipping.  Deterministic: every CTA
 * computes the norm in the same fixed order. */
int mpde_outer_step(float* theta, int64_t n_theta, float* lrs, int64_t n_lrs, const float* g_packed, float* m_theta,
                    float* v_theta, float* m_lrs, float* v_lrs, int optimizer, int t_theta, float lr_theta, float b1,
                    float b2, float eps, int t_lrs, float lr_lrs, float b1_lrs, float b2_lrs, float eps_lrs, float clip,
                    float* sumsq_out, void* stream);

/* ---- R: task-batched MAML outer gradient (src/nets/maml.py:100-233) -----
 * Forward sweep of K inner SGD steps per task, reverse sweep of K Hessian-vector products
 * (SURVEY.md 9.4).  Points are pre-sampled by the caller (the plugin's samplers):
 *   inner_points[k], outer_points[k]  k = 0..K-1   (keys k1 / k2 of scan step k, maml.py:134-139)
 *   final_points  (final_key, maml.py:158),  aux_points (outer_loss_key, maml.py:192)
 * Outputs:
 *   meta_grad_theta [P], meta_grad_lrs [K][P]: grad_scale * SUM over the T local tasks
 *       (grad_scale = 1/T_global reproduces the mean of maml.py:231; ranks all-reduce the sums)
 *   losses [T][K+1], meta_losses [T], aux_terms [T][n_terms] (outer_aux of maml.py:192),
 *   theta_final optional [T][P] (theta_K per task).
 *   inner_lrs may be NULL: scalar lr 1.0 per step (maml.py:126-127), no meta_grad_lrs written. */
typedef struct mpde_maml_cfg {
  int32_t inner_steps;      /* K: FLAGS.inner_steps                */
  float inner_lr;           /* FLAGS.inner_lr (Momentum beta=0)    */
  float inner_grad_clip;    /* FLAGS.inner_grad_clip               */
  int32_t softplus_lrs;     /* MamlDef.softplus_lrs                */
  float outer_loss_decay;   /* FLAGS.outer_loss_decay              */
} mpde_maml_cfg;

size_t mpde_maml_workspace_bytes(const mpde_net_desc* net, const mpde_pde_desc* pde_inner,
                                 const mpde_pde_desc* pde_outer, const mpde_maml_cfg* cfg, int T);
int mpde_maml_outer(const mpde_net_desc* net, const mpde_pde_desc* pde_inner, const mpde_pde_desc* pde_outer,
                    const mpde_maml_cfg* cfg, int T, const float* theta0, const float* inner_lrs,
                    const mpde_points* inner_points, const mpde_points* outer_points,
                    const mpde_points* final_points, const mpde_points* aux_points, const float* pde_params,
                    float grad_scale, float* meta_grad_theta, float* meta_grad_lrs, float* losses,
                    float* meta_losses, float* aux_terms, float* theta_final, void* workspace,
                    size_t workspace_bytes, void* stream);

/* First-order rollout only (deployment: src/maml_pde.py:163-193, src/nn_pde_maml.py:209-215):
 * K inner steps, no meta-gradient.  theta_final [T][P], losses [T][K+1]. */
int mpde_maml_adapt(const mpde_net_desc* net, const mpde_pde_desc* pde_inner, const mpde_maml_cfg* cfg, int T,
                    const float* theta0, const float* inner_lrs, const mpde_points* inner_points,
                    const mpde_points* final_points, const float* pde_params, float* theta_final, float* losses,
                    void* workspace, size_t workspace_bytes, void* stream);

/* ---- R: task-batched LEAP meta-gradient (src/nets/leap.py:49-228) --------
 *   loss0_points (loss0_key, leap.py:124-126), step_points[k] (k1), new_points[k] (k2, leap.py:67,93)
 *   meta_grad [P] = grad_scale * SUM over tasks; losses [T][K+1] = [loss0, new_loss_1..K] */
typedef struct mpde_leap_cfg {
  int32_t inner_steps;
  float inner_lr;
  float inner_grad_clip;
  int32_t norm, loss_in_distance, stabilize;  /* LeapDef flags (src/leap_pde.py:94-96) */
  int32_t inner_opt;                          /* 0: SGD, 1: Adam(beta1, beta2) (src/leap_pde.py:87) */
  float beta1, beta2, eps;
} mpde_leap_cfg;

size_t mpde_leap_workspace_bytes(const mpde_net_desc* net, const mpde_pde_desc* pde, const mpde_leap_cfg* cfg, int T);
int mpde_leap_outer(const mpde_net_desc* net, const mpde_pde_desc* pde, const mpde_leap_cfg* cfg, int T,
                    const float* theta0, const mpde_points* loss0_points, const mpde_points* step_points,
                    const mpde_points* new_points, const float* pde_params, float grad_scale, float* meta_grad,
                    float* losses, float* theta_final, void* workspace, size_t workspace_bytes, void* stream);

/* ---- introspection (bench / tests) ---------------------------------------- */
/* number of kernel launches issued by this library since load (a host counter) */
int64_t mpde_launch_count(void);
/* "cuda" for the product build; the test-only emulator build answers "emu" */
const char* mpde_backend(void);
/* Optional device-side timing of the fused jet kernel (bench.py's roofline leg; the reference's
 * counterpart is the wall-clock Timer around train_step, src/util/timer.py:5-27).  While enabled,
 * every launch of the jet kernel is bracketed by a cudaEvent pair on the caller's stream (events
 * are created on first use: do not enable inside a CUDA-graph capture).  mpde_prof_read
 * synchronises on the recorded events of `mode` (mpde_mode), returns their summed duration in
 * milliseconds and their count, and clears them.  mode 3 = the tensor-core path's weight-gradient
 * contraction kernel (tc_wgrad_kernel), all launches of GRAD and HVP evaluations together. */
int mpde_prof_enable(int on);
int mpde_prof_read(int mode, double* total_ms, int64_t* launches);

#ifdef __cplusplus
}
#endif
#endif /* METAPDE_B200_H_ */
