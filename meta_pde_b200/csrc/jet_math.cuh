// jet_math.cuh -- per-element Taylor-jet algebra of the SIREN layer and the
// per-point PDE heads, with hand-derived adjoints and tangents.
//
// Replaces the reference's nested autodiff (all paths relative to /root/reference):
//   jacfwd(jacrev(u) * w)  src/nets/field.py:47-72          -> JET_LAP (v, d0, d1, d00+d11)
//   jacfwd(jacfwd(u))[0]   src/burgers/td_burgers_common.py:54-65 -> JET_XX (v, d0, d1, d00)
//   jacfwd(u)              src/elasticity/hyper_elasticity_common.py:23-26 -> JET_G (v, d0, d1)
//   plain u(x)             boundary / initial terms          -> JET_V
// and jax.value_and_grad / the second-order MAML pass through them
// (src/nets/maml.py:66-70,195-202) with closed-form reverse and
// forward-over-reverse rules.
//
// Everything here is elementwise (one point x one neuron, all channels), so the
// same code is used by every kernel mode and can be checked on the host.
#pragma once
#include "compat.cuh"

namespace mpde {

enum JetKind { JET_V = 0, JET_G = 1, JET_LAP = 2, JET_XX = 3 };

template <int JK> struct JetTraits;
template <> struct JetTraits<JET_V>   { static constexpr int C = 1, NF = 2; static constexpr bool Q = false, D0 = false, D1 = false; };
template <> struct JetTraits<JET_G>   { static constexpr int C = 3, NF = 4; static constexpr bool Q = false, D0 = false, D1 = false; };
template <> struct JetTraits<JET_LAP> { static constexpr int C = 4, NF = 5; static constexpr bool Q = true,  D0 = true,  D1 = true;  };
template <> struct JetTraits<JET_XX>  { static constexpr int C = 4, NF = 5; static constexpr bool Q = true,  D0 = true,  D1 = false; };

// saved fields per element: sv = [S, Cc, p0, p1, q] (truncated to NF)
//   S = sin(om*v), Cc = cos(om*v); the pre-activation value v itself is never needed again.

// Activation kinds.  ACT_SIN is the SIREN layer sin(omega * z) (field.py:185-186); ACT_SWISH is the `--nosiren` network
// (nonlinearity = flax.nn.swish, no omega: maml_pde.py:115, field.py:187-188).  The swish jet rules are written in the
// derivatives g0..g4 of the activation at the value channel; the sin rules keep their closed forms.
enum ActKind { ACT_SIN = 0, ACT_SWISH = 1 };

// swish(z) = z * s(z), s = sigmoid: f_n = n * s^(n-1) + z * s^(n).  Saved fields: sv[0] = z, sv[1] = s(z).
template <int NMAX>
MPDE_HD void swish_derivs(float z, float s, float* g) {
  const float s1 = s * (1.f - s);                       // s1 = ds/dz
  g[0] = z * s;
  g[1] = fmaf(z, s1, s);
  if (NMAX >= 2) {
    const float t = 1.f - 2.f * s;
    const float s2 = s1 * t;                            // second derivative of s
    g[2] = fmaf(z, s2, 2.f * s1);
    if (NMAX >= 3) {
      const float u = fmaf(6.f * s, s - 1.f, 1.f);      // 1 - 6 s + 6 s^2
      const float s3 = s1 * u;                          // third
      g[3] = fmaf(z, s3, 3.f * s2);
      if (NMAX >= 4) {
        const float s4 = s2 * u - 6.f * s1 * s1 * t;    // fourth = s2 (1 - 6s + 6s^2) + s1^2 (12 s - 6)
        g[4] = fmaf(z, s4, 4.f * s3);
      }
    }
  }
}
MPDE_HD float swish_sigmoid_f(float z) { return 1.f / (1.f + expf(-z)); }

// a = sin-activation jet of the pre-activation jet z  (field.py:186: sin(a * omega))
template <int JK, int ACT = ACT_SIN>
MPDE_HD void jet_act_fwd(float om, const float* z, float* a, float* sv) {
  if (ACT == ACT_SWISH) {
    typedef JetTraits<JK> T;
    float g[3];
    const float sg = swish_sigmoid_f(z[0]);
    swish_derivs<T::Q ? 2 : 1>(z[0], sg, g);
    sv[0] = z[0]; sv[1] = sg;
    a[0] = g[0];
    if (T::C >= 3) {
      a[1] = g[1] * z[1]; a[2] = g[1] * z[2];
      sv[2] = z[1]; sv[3] = z[2];
      if (T::Q) {
        float pp = 0.f;
        if (T::D0) pp += z[1] * z[1];
        if (T::D1) pp += z[2] * z[2];
        a[3] = g[1] * z[3] + g[2] * pp;
        sv[4] = z[3];
      }
    }
    return;
  }
  typedef JetTraits<JK> T;
  float S, Cc;
  mpde_sincosf(om * z[0], &S, &Cc);
  sv[0] = S; sv[1] = Cc;
  a[0] = S;
  if (T::C >= 3) {
    const float oc = om * Cc;
    a[1] = oc * z[1]; a[2] = oc * z[2];
    sv[2] = z[1]; sv[3] = z[2];
    if (T::Q) {
      float pp = 0.f;
      if (T::D0) pp += z[1] * z[1];
      if (T::D1) pp += z[2] * z[2];
      a[3] = oc * z[3] - om * om * S * pp;
      sv[4] = z[3];
    }
  }
}

// same, from the saved fields (backward pass recomputes h_{l-1} from layer l-1's record)
template <int JK, int ACT = ACT_SIN>
MPDE_HD void jet_act_from_saved(float om, const float* sv, float* a) {
  if (ACT == ACT_SWISH) {
    typedef JetTraits<JK> T;
    float g[3];
    swish_derivs<T::Q ? 2 : 1>(sv[0], sv[1], g);
    a[0] = g[0];
    if (T::C >= 3) {
      a[1] = g[1] * sv[2]; a[2] = g[1] * sv[3];
      if (T::Q) {
        float pp = 0.f;
        if (T::D0) pp += sv[2] * sv[2];
        if (T::D1) pp += sv[3] * sv[3];
        a[3] = g[1] * sv[4] + g[2] * pp;
      }
    }
    return;
  }
  typedef JetTraits<JK> T;
  const float S = sv[0], Cc = sv[1];
  a[0] = S;
  if (T::C >= 3) {
    const float oc = om * Cc;
    a[1] = oc * sv[2]; a[2] = oc * sv[3];
    if (T::Q) {
      float pp = 0.f;
      if (T::D0) pp += sv[2] * sv[2];
      if (T::D1) pp += sv[3] * sv[3];
      a[3] = oc * sv[4] - om * om * S * pp;
    }
  }
}

// zbar = (d a / d z)^T abar
template <int JK, int ACT = ACT_SIN>
MPDE_HD void jet_act_bwd(float om, const float* sv, const float* ab, float* zb) {
  if (ACT == ACT_SWISH) {
    typedef JetTraits<JK> T;
    float g[4];
    swish_derivs<T::Q ? 3 : (T::C >= 3 ? 2 : 1)>(sv[0], sv[1], g);
    float vb = g[1] * ab[0];
    if (T::C >= 3) {
      const float p0 = sv[2], p1 = sv[3];
      vb += g[2] * (p0 * ab[1] + p1 * ab[2]);
      float pb0 = g[1] * ab[1], pb1 = g[1] * ab[2];
      if (T::Q) {
        const float q = sv[4], aq = ab[3];
        float pp = 0.f;
        if (T::D0) { pp += p0 * p0; pb0 += 2.f * g[2] * p0 * aq; }
        if (T::D1) { pp += p1 * p1; pb1 += 2.f * g[2] * p1 * aq; }
        vb += (g[2] * q + g[3] * pp) * aq;
        zb[3] = g[1] * aq;
      }
      zb[1] = pb0; zb[2] = pb1;
    }
    zb[0] = vb;
    return;
  }
  typedef JetTraits<JK> T;
  const float S = sv[0], Cc = sv[1];
  const float oc = om * Cc, o2s = om * om * S;
  float vb = oc * ab[0];
  if (T::C >= 3) {
    const float p0 = sv[2], p1 = sv[3];
    vb -= o2s * (p0 * ab[1] + p1 * ab[2]);
    float pb0 = oc * ab[1], pb1 = oc * ab[2];
    if (T::Q) {
      const float q = sv[4], aq = ab[3];
      float pp = 0.f;
      if (T::D0) { pp += p0 * p0; pb0 -= 2.f * o2s * p0 * aq; }
      if (T::D1) { pp += p1 * p1; pb1 -= 2.f * o2s * p1 * aq; }
      vb -= (o2s * q + om * om * oc * pp) * aq;
      zb[3] = oc * aq;
    }
    zb[1] = pb0; zb[2] = pb1;
  }
  zb[0] = vb;
}

// adot = (d a / d z) zdot
template <int JK, int ACT = ACT_SIN>
MPDE_HD void jet_act_tan(float om, const float* sv, const float* zd, float* ad) {
  if (ACT == ACT_SWISH) {
    typedef JetTraits<JK> T;
    float g[4];
    swish_derivs<T::Q ? 3 : (T::C >= 3 ? 2 : 1)>(sv[0], sv[1], g);
    const float vd = zd[0];
    ad[0] = g[1] * vd;
    if (T::C >= 3) {
      const float p0 = sv[2], p1 = sv[3];
      ad[1] = g[1] * zd[1] + g[2] * p0 * vd;
      ad[2] = g[1] * zd[2] + g[2] * p1 * vd;
      if (T::Q) {
        const float q = sv[4];
        float pp = 0.f, ppd = 0.f;
        if (T::D0) { pp += p0 * p0; ppd += p0 * zd[1]; }
        if (T::D1) { pp += p1 * p1; ppd += p1 * zd[2]; }
        ad[3] = g[1] * zd[3] + g[2] * q * vd + 2.f * g[2] * ppd + g[3] * pp * vd;
      }
    }
    return;
  }
  typedef JetTraits<JK> T;
  const float S = sv[0], Cc = sv[1];
  const float oc = om * Cc, o2s = om * om * S;
  const float vd = zd[0];
  ad[0] = oc * vd;
  if (T::C >= 3) {
    const float p0 = sv[2], p1 = sv[3];
    ad[1] = oc * zd[1] - o2s * p0 * vd;
    ad[2] = oc * zd[2] - o2s * p1 * vd;
    if (T::Q) {
      const float q = sv[4];
      float pp = 0.f, ppd = 0.f;
      if (T::D0) { pp += p0 * p0; ppd += p0 * zd[1]; }
      if (T::D1) { pp += p1 * p1; ppd += p1 * zd[2]; }
      ad[3] = oc * zd[3] - o2s * q * vd - 2.f * o2s * ppd - om * om * oc * pp * vd;
    }
  }
}

// zbardot = d/d eps [ (d a / d z)^T abar ] along (zdot, abardot)
template <int JK, int ACT = ACT_SIN>
MPDE_HD void jet_act_bwd_tan(float om, const float* sv, const float* zd, const float* ab, const float* abd,
                             float* zbd) {
  if (ACT == ACT_SWISH) {
    typedef JetTraits<JK> T;
    float g[5];
    swish_derivs<T::Q ? 4 : (T::C >= 3 ? 3 : 2)>(sv[0], sv[1], g);
    const float vd = zd[0];
    float vbd = g[1] * abd[0] + g[2] * vd * ab[0];
    if (T::C >= 3) {
      const float p0 = sv[2], p1 = sv[3];
      const float pd0 = zd[1], pd1 = zd[2];
      float pbd0 = g[1] * abd[1] + g[2] * vd * ab[1];
      float pbd1 = g[1] * abd[2] + g[2] * vd * ab[2];
      vbd += g[2] * (p0 * abd[1] + p1 * abd[2]) + g[2] * (pd0 * ab[1] + pd1 * ab[2]) +
             g[3] * vd * (p0 * ab[1] + p1 * ab[2]);
      if (T::Q) {
        const float q = sv[4], qd = zd[3], aq = ab[3], aqd = abd[3];
        float pp = 0.f, ppd = 0.f;
        if (T::D0) {
          pp += p0 * p0; ppd += p0 * pd0;
          pbd0 += 2.f * (g[2] * (p0 * aqd + pd0 * aq) + g[3] * vd * p0 * aq);
        }
        if (T::D1) {
          pp += p1 * p1; ppd += p1 * pd1;
          pbd1 += 2.f * (g[2] * (p1 * aqd + pd1 * aq) + g[3] * vd * p1 * aq);
        }
        vbd += (g[2] * q + g[3] * pp) * aqd;
        vbd += (g[2] * qd + g[3] * vd * q + 2.f * g[3] * ppd + g[4] * vd * pp) * aq;
        zbd[3] = g[1] * aqd + g[2] * vd * aq;
      }
      zbd[1] = pbd0; zbd[2] = pbd1;
    }
    zbd[0] = vbd;
    return;
  }
  typedef JetTraits<JK> T;
  const float S = sv[0], Cc = sv[1];
  const float oc = om * Cc, o2s = om * om * S, o3c = om * om * oc, o4s = om * om * o2s;
  const float vd = zd[0];
  // v-bar-dot, value part
  float vbd = oc * abd[0] - o2s * vd * ab[0];
  if (T::C >= 3) {
    const float p0 = sv[2], p1 = sv[3];
    const float pd0 = zd[1], pd1 = zd[2];
    // first-derivative channels
    float pbd0 = oc * abd[1] - o2s * vd * ab[1];
    float pbd1 = oc * abd[2] - o2s * vd * ab[2];
    vbd -= o2s * (p0 * abd[1] + p1 * abd[2]) + o2s * (pd0 * ab[1] + pd1 * ab[2]) +
           o3c * vd * (p0 * ab[1] + p1 * ab[2]);
    if (T::Q) {
      const float q = sv[4], qd = zd[3], aq = ab[3], aqd = abd[3];
      float pp = 0.f, ppd = 0.f;
      if (T::D0) {
        pp += p0 * p0; ppd += p0 * pd0;
        pbd0 -= 2.f * (o2s * (p0 * aqd + pd0 * aq) + o3c * vd * p0 * aq);
      }
      if (T::D1) {
        pp += p1 * p1; ppd += p1 * pd1;
        pbd1 -= 2.f * (o2s * (p1 * aqd + pd1 * aq) + o3c * vd * p1 * aq);
      }
      vbd -= (o2s * q + o3c * pp) * aqd;
      vbd -= (o2s * qd + o3c * vd * q + 2.f * o3c * ppd - o4s * vd * pp) * aq;
      zbd[3] = oc * aqd - o2s * vd * aq;
    }
    zbd[1] = pbd0; zbd[2] = pbd1;
  }
  zbd[0] = vbd;
}

// ============================================================================
// PDE heads.  u[o*C + c] = output jet of output o.  Each head returns the raw
// pointwise value whose mean is the loss term (r^2, (u-g)^2, psi) and the seed
// ubar = coef * d val / d u  (coef = loss_weight * term_scale / n_elems).
// ============================================================================
enum HeadKind {
  HEAD_POISSON_DOMAIN = 0,   // poisson_common.py:35-38
  HEAD_POISSON_BOUNDARY = 1, // poisson_common.py:30-33,122-134
  HEAD_BURGERS_DOMAIN = 2,   // td_burgers_common.py:54-67
  HEAD_BURGERS_IC = 3,       // burgers_formulation/default.py:9-31
  HEAD_HYPER_DOMAIN = 4,     // hyper_elasticity_common.py:34-56
  HEAD_HYPER_TOP = 5,        // hyper_elasticity_common.py:59-61
  HEAD_HYPER_BOTTOM = 6,     // hyper_elasticity_common.py:64-66
  HEAD_EVAL = 7,             // plain model(coords): maml_pde.py:190
  HEAD_EVAL2 = 8
};

// pde_params layouts (floats per task)
//   poisson:          [0..5] source_params (2x3: cx, cy, amp), [6..10] bc_params, [11..12] geo_params
//   td_burgers:       [0] Re (source_params[0]), [1..2] ic_params
//   hyper_elasticity: [0..1] source_params, [2..3] bc_params (E = bc_params[0])
static constexpr int NPRM_POISSON = 13, NPRM_BURGERS = 3, NPRM_HYPER = 4;

MPDE_HD float poisson_source(const float* prm, float x0, float x1) {
  float f = 0.f;
  for (int j = 0; j < 2; ++j) {
    const float dx = x0 - prm[3 * j + 0], dy = x1 - prm[3 * j + 1];
    f += prm[3 * j + 2] * expf(-(dx * dx + dy * dy));
  }
  return f;
}
MPDE_HD float poisson_bc(const float* prm, float x0, float x1) {
  const float* r = prm + 6;
  const float th = atan2f(x1, x0);
  return r[0] + r[1] * 0.25f * cosf(th) + r[2] * 0.25f * sinf(th) + r[3] * 0.25f * cosf(2.f * th) +
         r[4] * 0.25f * sinf(2.f * th);
}
MPDE_HD float burgers_ic(const float* prm, float x0) {
  const float pi = 3.14159265358979323846f;
  return sinf(pi * x0) + prm[1] * sinf(2.0f * pi * x0) + prm[2] * sinf(4.0f * pi * x0);
}

template <int HEAD> struct HeadTraits;
template <> struct HeadTraits<HEAD_POISSON_DOMAIN>   { static constexpr int JK = JET_LAP, OUT = 1; };
template <> struct HeadTraits<HEAD_POISSON_BOUNDARY> { static constexpr int JK = JET_V,   OUT = 1; };
template <> struct HeadTraits<HEAD_BURGERS_DOMAIN>   { static constexpr int JK = JET_XX,  OUT = 1; };
template <> struct HeadTraits<HEAD_BURGERS_IC>       { static constexpr int JK = JET_V,   OUT = 1; };
template <> struct HeadTraits<HEAD_HYPER_DOMAIN>     { static constexpr int JK = JET_G,   OUT = 2; };
template <> struct HeadTraits<HEAD_HYPER_TOP>        { static constexpr int JK = JET_V,   OUT = 2; };
template <> struct HeadTraits<HEAD_HYPER_BOTTOM>     { static constexpr int JK = JET_V,   OUT = 2; };
template <> struct HeadTraits<HEAD_EVAL>             { static constexpr int JK = JET_V,   OUT = 1; };
template <> struct HeadTraits<HEAD_EVAL2>            { static constexpr int JK = JET_V,   OUT = 2; };

// value-target heads: target g[o]
template <int HEAD>
MPDE_HD void head_target(const float* prm, float x0, float x1, float* g) {
  if (HEAD == HEAD_POISSON_BOUNDARY) g[0] = poisson_bc(prm, x0, x1);
  else if (HEAD == HEAD_BURGERS_IC) g[0] = burgers_ic(prm, x0);
  else if (HEAD == HEAD_HYPER_TOP) { g[0] = 0.f; g[1] = -0.12f; }
  else { g[0] = 0.f; g[1] = 0.f; }
}

// Returns val; writes ubar[OUT*C]; *res = per-point residual-like quantity (r, u-g summed, psi).
template <int HEAD>
MPDE_HD float head_eval(const float* u, float x0, float x1, const float* prm, float coef, float* ub, float* res) {
  if (HEAD == HEAD_POISSON_DOMAIN) {
    const float v = u[0], ux = u[1], uy = u[2], uq = u[3];
    const float g2 = ux * ux + uy * uy;
    const float r = (1.f + 0.1f * v * v) * uq + 0.2f * v * g2 - poisson_source(prm, x0, x1);
    const float w = 2.f * coef * r;
    ub[0] = w * (0.2f * v * uq + 0.2f * g2);
    ub[1] = w * 0.4f * v * ux;
    ub[2] = w * 0.4f * v * uy;
    ub[3] = w * (1.f + 0.1f * v * v);
    *res = r;
    return r * r;
  } else if (HEAD == HEAD_BURGERS_DOMAIN) {
    // r = u_t - (u_xx / Re - u * u_x); channel 1 = d/dx, channel 2 = d/dt
    const float v = u[0], ux = u[1], ut = u[2], uq = u[3];
    const float ire = 1.0f / prm[0];
    const float r = ut - (ire * uq - ux * v);
    const float w = 2.f * coef * r;
    ub[0] = w * ux; ub[1] = w * v; ub[2] = w; ub[3] = -w * ire;
    *res = r;
    return r * r;
  } else if (HEAD == HEAD_HYPER_DOMAIN) {
    // u[o*3 + c]; grad u = [[a, b], [c, d]], F = I + grad u, J = det F, Ic = tr(F F^T).
    // psi = (mu/2)(Ic/J - 2) + (kappa/2)(J - 1)^2 is evaluated in the algebraically identical,
    // cancellation-free form  Ic - 2J = (a - d)^2 + (b + c)^2,  J - 1 = a + d + a d - b c
    // (the literal float32 form loses ~4 digits near F = I, where the reference's own f32 result is noise).
    const float E = prm[2];
    const float A = 0.5f * E / (2.f * (1.f + 0.49f)), kap = E / (3.f * (1.f - 2.f * 0.49f));
    const float a = u[1], b = u[2], c = u[4], d = u[5];
    const float m = a + d + a * d - b * c;   // J - 1
    const float J = 1.f + m, iJ = 1.f / J;
    const float e1 = a - d, e2 = b + c;
    const float N = e1 * e1 + e2 * e2;
    const float psi = A * N * iJ + 0.5f * kap * m * m;
    // d psi / dx = A (N_x / J - N J_x / J^2) + kappa m J_x ; J_a = 1+d, J_d = 1+a, J_b = -c, J_c = -b
    const float s = kap * m - A * N * iJ * iJ;
    const float t1 = 2.f * A * iJ;
    ub[0] = 0.f; ub[3] = 0.f;
    ub[1] = coef * (t1 * e1 + s * (1.f + d));
    ub[2] = coef * (t1 * e2 - s * c);
    ub[4] = coef * (t1 * e2 - s * b);
    ub[5] = coef * (-t1 * e1 + s * (1.f + a));
    *res = psi;
    return psi;
  } else if (HEAD == HEAD_EVAL || HEAD == HEAD_EVAL2) {
    for (int o = 0; o < HeadTraits<HEAD>::OUT; ++o) ub[o] = 0.f;
    *res = u[0];
    return 0.f;
  } else {
    float g[2];
    head_target<HEAD>(prm, x0, x1, g);
    float val = 0.f;
    for (int o = 0; o < HeadTraits<HEAD>::OUT; ++o) {
      const float e = u[o] - g[o];
      ub[o] = 2.f * coef * e;
      val += e * e;
    }
    *res = u[0] - g[0];
    return val;
  }
}

// ubardot = d/d eps ubar along udot
template <int HEAD>
MPDE_HD void head_tangent(const float* u, const float* ud, float x0, float x1, const float* prm, float coef,
                          float* ubd) {
  if (HEAD == HEAD_POISSON_DOMAIN) {
    const float v = u[0], ux = u[1], uy = u[2], uq = u[3];
    const float vd = ud[0], uxd = ud[1], uyd = ud[2], uqd = ud[3];
    const float g2 = ux * ux + uy * uy;
    const float r = (1.f + 0.1f * v * v) * uq + 0.2f * v * g2 - poisson_source(prm, x0, x1);
    const float dv = 0.2f * v * uq + 0.2f * g2, dx = 0.4f * v * ux, dy = 0.4f * v * uy, dq = 1.f + 0.1f * v * v;
    const float rd = dv * vd + dx * uxd + dy * uyd + dq * uqd;
    const float gd = ux * uxd + uy * uyd;
    const float c2 = 2.f * coef;
    ubd[0] = c2 * (rd * dv + r * (0.2f * (vd * uq + v * uqd) + 0.4f * gd));
    ubd[1] = c2 * (rd * dx + r * 0.4f * (vd * ux + v * uxd));
    ubd[2] = c2 * (rd * dy + r * 0.4f * (vd * uy + v * uyd));
    ubd[3] = c2 * (rd * dq + r * 0.2f * v * vd);
  } else if (HEAD == HEAD_BURGERS_DOMAIN) {
    const float v = u[0], ux = u[1], ut = u[2], uq = u[3];
    const float vd = ud[0], uxd = ud[1], utd = ud[2], uqd = ud[3];
    const float ire = 1.0f / prm[0];
    const float r = ut - (ire * uq - ux * v);
    const float rd = ux * vd + v * uxd + utd - ire * uqd;
    const float c2 = 2.f * coef;
    ubd[0] = c2 * (rd * ux + r * uxd);
    ubd[1] = c2 * (rd * v + r * vd);
    ubd[2] = c2 * rd;
    ubd[3] = -c2 * rd * ire;
  } else if (HEAD == HEAD_HYPER_DOMAIN) {
    const float E = prm[2];
    const float A = 0.5f * E / (2.f * (1.f + 0.49f)), kap = E / (3.f * (1.f - 2.f * 0.49f));
    const float a = u[1], b = u[2], c = u[4], d = u[5];
    const float ad = ud[1], bd = ud[2], cd = ud[4], dd = ud[5];
    const float m = a + d + a * d - b * c;
    const float J = 1.f + m, iJ = 1.f / J;
    const float e1 = a - d, e2 = b + c, e1d = ad - dd, e2d = bd + cd;
    const float N = e1 * e1 + e2 * e2, Nd = 2.f * (e1 * e1d + e2 * e2d);
    const float Jd = (1.f + d) * ad + (1.f + a) * dd - c * bd - b * cd;
    const float s = kap * m - A * N * iJ * iJ;
    const float sd = kap * Jd - A * (Nd * iJ * iJ - 2.f * N * Jd * iJ * iJ * iJ);
    const float t1 = 2.f * A * iJ, t1d = -2.f * A * Jd * iJ * iJ;
    ubd[0] = 0.f; ubd[3] = 0.f;
    ubd[1] = coef * (t1 * e1d + t1d * e1 + sd * (1.f + d) + s * dd);
    ubd[2] = coef * (t1 * e2d + t1d * e2 - sd * c - s * cd);
    ubd[4] = coef * (t1 * e2d + t1d * e2 - sd * b - s * bd);
    ubd[5] = coef * (-t1 * e1d - t1d * e1 + sd * (1.f + a) + s * ad);
  } else if (HEAD == HEAD_EVAL || HEAD == HEAD_EVAL2) {
    for (int o = 0; o < HeadTraits<HEAD>::OUT; ++o) ubd[o] = 0.f;
  } else {
    for (int o = 0; o < HeadTraits<HEAD>::OUT; ++o) ubd[o] = 2.f * coef * ud[o];
  }
}

}  // namespace mpde
