// sampler_kernels.cuh -- jax.random on the GPU for the collocation-point samplers.
//
// The reference draws task points with jax 0.2.20's threefry2x32 PRNG (src/poisson/poisson_common.py:69-119,
// src/burgers/td_burgers_common.py:117-186).  jax is a third-party dependency absent from /root/reference: the published
// Threefry-2x32-20 block (Salmon et al., SC'11: rotations 13/15/26/6, 17/29/16/24, key-schedule constant 0x1BD11BDA) and
// jax 0.2.20's derivations on top of it (split: counts iota(2 num); _random_bits: counts iota(n), odd sizes zero-padded,
// the counter array split in two halves; uniform: mantissa trick on bits >> 9) are restated here exactly as in
// meta_pde_b200/compat/random.py, which is pinned by the reference's three jax.random known answers.  Integer work is
// bit-exact with that host restatement; uniform() uses separately rounded multiply and add (no FMA contraction), so it is
// bit-exact too.  One source, two builds (compat.cuh): the emulator build checks these kernels on the CPU.
#pragma once
#include "compat.cuh"

namespace mpde {

// separately rounded multiply / add: device code must not contract them into an FMA (bit-exactness with numpy)
#if defined(__CUDA_ARCH__)
#define MPDE_FMUL_RN(a, b) __fmul_rn((a), (b))
#define MPDE_FADD_RN(a, b) __fadd_rn((a), (b))
#else
#define MPDE_FMUL_RN(a, b) ((float)((float)(a) * (float)(b)))
#define MPDE_FADD_RN(a, b) ((float)((float)(a) + (float)(b)))
#endif

MPDE_HD uint32_t tf_rotl(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }

// one Threefry-2x32-20 block
MPDE_HD void threefry2x32(uint32_t k0, uint32_t k1, uint32_t& x0, uint32_t& x1) {
  const uint32_t ks[3] = {k0, k1, k0 ^ k1 ^ 0x1BD11BDAu};
  const int rot[2][4] = {{13, 15, 26, 6}, {17, 29, 16, 24}};
  x0 += ks[0];
  x1 += ks[1];
  MPDE_UNROLL
  for (int i = 0; i < 5; ++i) {
    MPDE_UNROLL
    for (int j = 0; j < 4; ++j) {
      x0 += x1;
      x1 = tf_rotl(x1, rot[i & 1][j]);
      x1 ^= x0;
    }
    x0 += ks[(i + 1) % 3];
    x1 += ks[(i + 2) % 3] + (uint32_t)(i + 1);
  }
}

// element i of _threefry_counts(key, iota(n)): counts [0..n) zero-padded to an even length 2h, halves (c[i], c[h + i])
MPDE_HD uint32_t tf_bits_at(uint32_t k0, uint32_t k1, long long n, long long i) {
  const long long h = (n + 1) / 2;
  const long long b = i < h ? i : i - h;                 // block index
  uint32_t x0 = (uint32_t)b;
  uint32_t x1 = (h + b < n) ? (uint32_t)(h + b) : 0u;    // the padding element is zero
  threefry2x32(k0, k1, x0, x1);
  return i < h ? x0 : x1;
}

// key j of jax.random.split(key, num): flat = _threefry_counts(key, iota(2 num)), key j = (flat[2j], flat[2j + 1])
MPDE_HD void tf_split_at(uint32_t k0, uint32_t k1, int num, int j, uint32_t& o0, uint32_t& o1) {
  o0 = tf_bits_at(k0, k1, 2LL * num, 2LL * j);
  o1 = tf_bits_at(k0, k1, 2LL * num, 2LL * j + 1);
}

// jax.random.uniform from 32 random bits: float in [1, 2) from the top 23 bits, minus 1, scaled, clamped from below
MPDE_HD float tf_uniform_from_bits(uint32_t bits, float lo, float hi) {
  union { uint32_t u; float f; } c;
  c.u = (bits >> 9) | 0x3F800000u;
  const float f = MPDE_FADD_RN(c.f, -1.0f);
  const float span = MPDE_FADD_RN(hi, -lo);
  const float v = MPDE_FADD_RN(MPDE_FMUL_RN(f, span), lo);
  return v > lo ? v : lo;
}

// ---------------------------------------------------------------------------------------------------------
// out[s][i] = random bits / uniform(keys[s], (n,), lo, hi)[i].   grid (blocks, S), 256 threads
// ---------------------------------------------------------------------------------------------------------
MPDE_GLOBAL void random_bits_kernel(const uint32_t* keys, long long n, uint32_t* out) {
  const int s = blockIdx.y;
  const uint32_t k0 = keys[2 * s], k1 = keys[2 * s + 1];
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    out[(size_t)s * n + i] = tf_bits_at(k0, k1, n, i);
}
MPDE_GLOBAL void random_uniform_kernel(const uint32_t* keys, long long n, float lo, float hi, float* out) {
  const int s = blockIdx.y;
  const uint32_t k0 = keys[2 * s], k1 = keys[2 * s + 1];
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    out[(size_t)s * n + i] = tf_uniform_from_bits(tf_bits_at(k0, k1, n, i), lo, hi);
}
// out[s][j] = split(keys[s], num)[j].   grid (S), 32 threads
MPDE_GLOBAL void random_split_kernel(const uint32_t* keys, int num, uint32_t* out) {
  const int s = blockIdx.x;
  for (int j = threadIdx.x; j < num; j += blockDim.x) {
    uint32_t a, b;
    tf_split_at(keys[2 * s], keys[2 * s + 1], num, j, a, b);
    out[((size_t)s * num + j) * 2] = a;
    out[((size_t)s * num + j) * 2 + 1] = b;
  }
}

// ---------------------------------------------------------------------------------------------------------
// Poisson sample_points (poisson_common.py:69-119) for S (task, draw) samples at once.
//   k1, k2 = split(key)
//   boundary (:76-88):  theta = linspace(0, 2 pi, n) + uniform(k1, (n,), 0, 2 pi / n);  r0 = 1 + c1 cos 4 theta + c2 cos 8 theta
//   domain (:91-119):   ka, kb, kc = split(k2, 3); xs = uniform(ka, (3n,), xmin, xmax); ys = uniform(kb, (3n,), ymin, ymax);
//                       in_hole = r0(arctan2(xs, ys)) < |xy| + 1e-7   (argument order of the reference kept);
//                       choice(kc, 3n, (n,), replace=False, p = 1 - in_hole) = the n smallest of
//                       g = -gumbel(kc) - log p,  gumbel = -log(-log(uniform(kc, tiny, 1)))
// This kernel writes the boundary points, the 3n candidates and their selection keys g; the caller takes the n smallest
// (stable order) -- a sort, done with the host framework's device sort (data preparation, not the hot path).
// grid (blocks, S), 256 threads.
// ---------------------------------------------------------------------------------------------------------
struct PoissonSampleParams {
  int S, n;
  const uint32_t* keys;      // [S][2]
  const float* geo;          // [S][2]  (c1, c2)
  float xmin, xmax, ymin, ymax;
  float* boundary;           // [S][n][2]
  float* cand_xy;            // [S][3n][2]
  float* cand_g;             // [S][3n]
};
MPDE_GLOBAL void poisson_sample_kernel(PoissonSampleParams p) {
  const int s = blockIdx.y;
  const int n = p.n, m = 3 * p.n;
  const uint32_t k0 = p.keys[2 * s], k1 = p.keys[2 * s + 1];
  const float c1 = p.geo[2 * s], c2 = p.geo[2 * s + 1];
  uint32_t b0, b1, d0, d1;
  tf_split_at(k0, k1, 2, 0, b0, b1);
  tf_split_at(k0, k1, 2, 1, d0, d1);
  uint32_t a0, a1, e0, e1, f0, f1;
  tf_split_at(d0, d1, 3, 0, a0, a1);
  tf_split_at(d0, d1, 3, 1, e0, e1);
  tf_split_at(d0, d1, 3, 2, f0, f1);
  const double two_pi = 6.283185307179586;
  const float umax = (float)(two_pi / (double)n);
  const double step = n > 1 ? two_pi / (double)(n - 1) : 0.0;
  const float tiny = 1.17549435e-38f;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n + m; i += stride) {
    if (i < n) {
      const float lin = (i == n - 1 && n > 1) ? (float)two_pi : (float)((double)i * step);
      const float th = MPDE_FADD_RN(lin, tf_uniform_from_bits(tf_bits_at(b0, b1, n, i), 0.0f, umax));
      const float r0 = MPDE_FADD_RN(MPDE_FADD_RN(1.0f, MPDE_FMUL_RN(c1, cosf(MPDE_FMUL_RN(4.0f, th)))), MPDE_FMUL_RN(c2, cosf(MPDE_FMUL_RN(8.0f, th))));
      p.boundary[((size_t)s * n + i) * 2] = MPDE_FMUL_RN(r0, cosf(th));
      p.boundary[((size_t)s * n + i) * 2 + 1] = MPDE_FMUL_RN(r0, sinf(th));
    } else {
      const long long j = i - n;
      const float x = tf_uniform_from_bits(tf_bits_at(a0, a1, m, j), p.xmin, p.xmax);
      const float y = tf_uniform_from_bits(tf_bits_at(e0, e1, m, j), p.ymin, p.ymax);
      const float th = atan2f(x, y);
      const float r0 = MPDE_FADD_RN(MPDE_FADD_RN(1.0f, MPDE_FMUL_RN(c1, cosf(MPDE_FMUL_RN(4.0f, th)))), MPDE_FMUL_RN(c2, cosf(MPDE_FMUL_RN(8.0f, th))));
      const float len = sqrtf(MPDE_FADD_RN(MPDE_FMUL_RN(x, x), MPDE_FMUL_RN(y, y)));
      const bool in_hole = r0 < MPDE_FADD_RN(len, 1e-7f);
      const float u = tf_uniform_from_bits(tf_bits_at(f0, f1, m, j), tiny, 1.0f);
      // -gumbel - log p:  log(-log u) - log(1) for admissible candidates, +inf for candidates in the hole
      const float g = in_hole ? INFINITY : logf(-logf(u));
      p.cand_xy[((size_t)s * m + j) * 2] = x;
      p.cand_xy[((size_t)s * m + j) * 2 + 1] = y;
      p.cand_g[(size_t)s * m + j] = g;
    }
  }
}

}  // namespace mpde
