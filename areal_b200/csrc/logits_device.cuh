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

// warp-wide max / min of an fp32 value in ONE instruction (redux.sync.f32, sm_100a: CREDUX into a uniform register).
// NaN operands are ignored, like fmaxf / fminf.
__device__ __forceinline__ float warp_max_f32(float v) {
  float r;
  asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
  return r;
}
__device__ __forceinline__ float warp_min_f32(float v) {
  float r;
  asm volatile("redux.sync.min.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
  return r;
}

// (s, w) of a canonical state moved to the reference M >= st.m: one ex2
__device__ __forceinline__ void state_shift_to(const RowState& st, float M, float c2, float& s, float& w) {
  const bool empty = st.m == -INFINITY;
  const float d = empty ? 0.f : (st.m - M) * c2;
  const float f = empty ? 0.f : ex2_approx(d);
  s = st.s * f;
  w = f * fmaf(d, st.s, st.w);
}

// Merge the 32 canonical states of a warp: the row max first (one redux), then every lane rebases its
// (s, w) to it ONCE -- one ex2 per lane instead of two per butterfly round -- and s, w are plain
// butterfly sums, mn one more redux.  All lanes end with the warp's state.
__device__ __forceinline__ void warp_merge_state(RowState& st, float c2) {
  const float M = warp_max_f32(st.m);
  float s, w;
  state_shift_to(st, M, c2, s, w);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_xor_sync(0xffffffffu, s, o);
    w += __shfl_xor_sync(0xffffffffu, w, o);
  }
  st.mn = warp_min_f32(st.mn);
  st.m = M;
  st.s = s;
  st.w = w;
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


// Gradient of one 16-byte vector when only the log-prob carries a gradient (k1 == 0, the fused actor
// step: the entropy is detached): d = -gl * p = -/+ 2^(x*c2 + nbias) with nbias = -m*c2 - log2 S + log2|gl|,
// i.e. the scale rides in the exponent: one FFMA + one MUFU per element, the sign folded into the pack.
template <bool kNeg>
__device__ __forceinline__ uint32_t grad_pair_exp(uint32_t wd, float c2, float nbias) {
  const float e0 = ex2_approx(fmaf(bf16_lo(wd), c2, nbias));
  const float e1 = ex2_approx(fmaf(bf16_hi(wd), c2, nbias));
  const uint32_t o = pack_bf16x2(e0, e1);
  return kNeg ? (o ^ 0x80008000u) : o;
}

template <bool kNeg>
__device__ __forceinline__ uint4 grad_vec_exp(const uint4& q, float c2, float nbias) {
  uint4 o;
  o.x = grad_pair_exp<kNeg>(q.x, c2, nbias);
  o.y = grad_pair_exp<kNeg>(q.y, c2, nbias);
  o.z = grad_pair_exp<kNeg>(q.z, c2, nbias);
  o.w = grad_pair_exp<kNeg>(q.w, c2, nbias);
  return o;
}


// ---- element traits: how a 16-byte vector of logits unpacks to fp32 and packs back -------------------
struct ElemBF16 {
  static constexpr int kN = 8, kSize = 2;
  __device__ static __forceinline__ void unpack(const uint4& q, float (&x)[8]) {
    x[0] = bf16_lo(q.x); x[1] = bf16_hi(q.x); x[2] = bf16_lo(q.y); x[3] = bf16_hi(q.y);
    x[4] = bf16_lo(q.z); x[5] = bf16_hi(q.z); x[6] = bf16_lo(q.w); x[7] = bf16_hi(q.w);
  }
  __device__ static __forceinline__ uint4 pack(const float (&d)[8]) {
    return make_uint4(pack_bf16x2(d[0], d[1]), pack_bf16x2(d[2], d[3]), pack_bf16x2(d[4], d[5]), pack_bf16x2(d[6], d[7]));
  }
};
struct ElemFP16 {
  static constexpr int kN = 8, kSize = 2;
  __device__ static __forceinline__ void unpack(const uint4& q, float (&x)[8]) {
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
      x[2 * i] = f.x;
      x[2 * i + 1] = f.y;
    }
  }
  __device__ static __forceinline__ uint4 pack(const float (&d)[8]) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __half2 h = __floats2half2_rn(d[2 * i], d[2 * i + 1]);
      w[i] = *reinterpret_cast<const uint32_t*>(&h);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
};
struct ElemFP32 {
  static constexpr int kN = 4, kSize = 4;
  __device__ static __forceinline__ void unpack(const uint4& q, float (&x)[4]) {
    x[0] = __uint_as_float(q.x); x[1] = __uint_as_float(q.y); x[2] = __uint_as_float(q.z); x[3] = __uint_as_float(q.w);
  }
  __device__ static __forceinline__ uint4 pack(const float (&d)[4]) {
    return make_uint4(__float_as_uint(d[0]), __float_as_uint(d[1]), __float_as_uint(d[2]), __float_as_uint(d[3]));
  }
};

// statistics of kV vectors at once (one rebase per batch), any element type
template <typename E, bool kMinMax, int kV>
__device__ __forceinline__ void consume_batch(RowState& st, const uint4 (&q)[kV], float c2) {
  float x[kV][E::kN];
  float mx = -INFINITY, mn = INFINITY;
#pragma unroll
  for (int j = 0; j < kV; ++j) {
    E::unpack(q[j], x[j]);
#pragma unroll
    for (int i = 0; i < E::kN; ++i) {
      mx = fmaxf(mx, x[j][i]);
      if (kMinMax) mn = fminf(mn, x[j][i]);
    }
  }
  if (kMinMax) st.mn = fminf(st.mn, mn);
  const float nb = state_rebase(st, mx, c2);
  float s0 = 0.f, s1 = 0.f, w0 = 0.f, w1 = 0.f;
#pragma unroll
  for (int j = 0; j < kV; ++j) {
#pragma unroll
    for (int i = 0; i < E::kN; i += 2) {
      const float t0 = fmaf(x[j][i], c2, nb), t1 = fmaf(x[j][i + 1], c2, nb);
      const float e0 = ex2_approx(t0), e1 = ex2_approx(t1);
      s0 += e0;
      w0 = fmaf(e0, t0, w0);
      s1 += e1;
      w1 = fmaf(e1, t1, w1);
    }
  }
  st.s += s0 + s1;
  st.w += w0 + w1;
}

template <typename E>
__device__ __forceinline__ uint4 grad_vec_t(const uint4& q, const RowGrad& g, float c2, int64_t lab_off) {
  float x[E::kN], d[E::kN];
  E::unpack(q, x);
#pragma unroll
  for (int i = 0; i < E::kN; ++i) {
    const float pv = ex2_approx(fmaf(x[i] - g.m, c2, g.nl2s));
    d[i] = pv * fmaf(x[i], g.k1, g.k0);
  }
  if (lab_off >= 0 && lab_off < E::kN) {
#pragma unroll
    for (int i = 0; i < E::kN; ++i)
      if (i == lab_off) d[i] += g.gl;
  }
  return E::pack(d);
}

}  // namespace dev
}  // namespace ab
