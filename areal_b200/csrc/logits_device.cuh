// Device-side building blocks shared by the logits kernels (logits_fused.cu, logits_rowres.cu):
// the per-thread online-softmax state, its merge, and the per-vector gradient.
#pragma once
#include "common.cuh"

namespace ab {
namespace dev {

struct RowState {
  float m;   // running max of the raw logits seen by this thread (-inf initially)
  float nb;  // fl(-m * c2): every term below is 2^(x*c2 + nb), i.e. relative to this exact offset
  float s;   // sum e,      e = 2^t,  t = x*c2 + nb  (<= ~0)
  float w;   // sum e * t   (shifted first moment: small whenever the row is peaked, so the
             //              entropy keeps its relative accuracy; no extra per-element instruction)
  float mn;  // running min of the raw logits
};

__device__ __forceinline__ void state_init(RowState& st) {
  st.m = -INFINITY;
  st.nb = 0.f;
  st.s = 0.f;
  st.w = 0.f;
  st.mn = INFINITY;
}

// Move the thread's reference to max(st.m, cm).  Exactly consistent: the old terms are rescaled by
// 2^(nb_new - nb_old), the very offsets they were accumulated with.
__device__ __forceinline__ float state_rebase(RowState& st, float cm, float c2) {
  const float m_new = fmaxf(st.m, cm);
  const float nb_new = -m_new * c2;
  const float delta = (st.m == -INFINITY) ? 0.f : (nb_new - st.nb);   // <= 0
  const float alpha = ex2_approx(delta);
  st.w = alpha * fmaf(delta, st.s, st.w);
  st.s *= alpha;
  st.m = m_new;
  st.nb = nb_new;
  return nb_new;
}

// Canonical form: terms relative to the EXACT max, t = (x - m) * c2.  r = m*c2 + nb is the rounding
// residual of nb, obtained exactly with one FMA.
__device__ __forceinline__ void state_canonical(RowState& st, float c2) {
  if (st.s == 0.f) return;
  const float r = fmaf(st.m, c2, st.nb);
  const float f = ex2_approx(-r);
  st.w = f * fmaf(-r, st.s, st.w);
  st.s *= f;
}

__device__ __forceinline__ uint32_t vec_max_bf16(const uint4& q) {
  return max_bf16x2(max_bf16x2(q.x, q.y), max_bf16x2(q.z, q.w));
}
__device__ __forceinline__ uint32_t vec_min_bf16(const uint4& q) {
  return min_bf16x2(min_bf16x2(q.x, q.y), min_bf16x2(q.z, q.w));
}

#define AB_ACC2(wd, S, W)                        \
  {                                              \
    const float t0 = fmaf(bf16_lo(wd), c2, nb);  \
    const float t1 = fmaf(bf16_hi(wd), c2, nb);  \
    const float e0 = ex2_approx(t0);             \
    const float e1 = ex2_approx(t1);             \
    S##0 += e0;                                  \
    W##0 = fmaf(e0, t0, W##0);                   \
    S##1 += e1;                                  \
    W##1 = fmaf(e1, t1, W##1);                   \
  }

// accumulate one 16-byte vector (8 bf16) against an already rebased reference nb
__device__ __forceinline__ void vec_accumulate(const uint4& q, float c2, float nb, float& s0,
                                               float& s1, float& w0, float& w1) {
  AB_ACC2(q.x, s, w) AB_ACC2(q.y, s, w) AB_ACC2(q.z, s, w) AB_ACC2(q.w, s, w)
}

template <bool kMinMax>
__device__ __forceinline__ void consume_vec(RowState& st, const uint4& q, float c2) {
  const uint32_t mx = vec_max_bf16(q);
  if (kMinMax) {
    const uint32_t mn = vec_min_bf16(q);
    st.mn = fminf(st.mn, fminf(bf16_lo(mn), bf16_hi(mn)));
  }
  const float nb = state_rebase(st, fmaxf(bf16_lo(mx), bf16_hi(mx)), c2);
  float s0 = 0.f, s1 = 0.f, w0 = 0.f, w1 = 0.f;
  vec_accumulate(q, c2, nb, s0, s1, w0, w1);
  st.s += s0 + s1;
  st.w += w0 + w1;
}

// merge two canonical (m, s, w) states: shift each by (m - M) * c2
__device__ __forceinline__ void state_merge(float& m, float& s, float& w, float m2, float s2,
                                            float w2, float c2) {
  const float M = fmaxf(m, m2);
  const float da = (m == -INFINITY) ? 0.f : (m - M) * c2;
  const float db = (m2 == -INFINITY) ? 0.f : (m2 - M) * c2;
  const float fa = (m == -INFINITY) ? 0.f : ex2_approx(da);
  const float fb = (m2 == -INFINITY) ? 0.f : ex2_approx(db);
  w = fa * fmaf(da, s, w) + fb * fmaf(db, s2, w2);
  s = s * fa + s2 * fb;
  m = M;
}

struct RowGrad {
  float k0, k1, gl, m, nl2s;   // nl2s = -log2 S
  bool active;
};

__device__ __forceinline__ uint4 grad_vec(const uint4& q, const RowGrad& g, float c2, int64_t lab_off) {
  float x[8] = {bf16_lo(q.x), bf16_hi(q.x), bf16_lo(q.y), bf16_hi(q.y),
                bf16_lo(q.z), bf16_hi(q.z), bf16_lo(q.w), bf16_hi(q.w)};
  float d[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float pv = ex2_approx(fmaf(x[i] - g.m, c2, g.nl2s));
    d[i] = pv * fmaf(x[i], g.k1, g.k0);
  }
  if (lab_off >= 0 && lab_off < 8) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      if (i == lab_off) d[i] += g.gl;
  }
  uint4 o;
  o.x = pack_bf16x2(d[0], d[1]);
  o.y = pack_bf16x2(d[2], d[3]);
  o.z = pack_bf16x2(d[4], d[5]);
  o.w = pack_bf16x2(d[6], d[7]);
  return o;
}


}  // namespace dev
}  // namespace ab
