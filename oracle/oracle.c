/* TEST INFRASTRUCTURE ONLY -- plain-C restatement of the integer/byte and sequential-fp32
 * parts of the reference's RL post-processing path, fast enough to check the CUDA path at
 * BASELINE.json's full sizes.  Built by `oracle/build_oracle.py` into oracle/liboracle.so
 * (gcc -O2 -ffp-contract=off: fp32 operation order must stay exactly the reference's).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load it; the product (areal_b200/) never does.
 *
 * Pinned by tests/test_oracle_golden.py against tests/golden/{gae1d,gae_padded,intervals}.npz,
 * which hold outputs of the unmodified reference functions.
 * Citations are relative to the reference root. */
#include <stdint.h>
#include <string.h>

/* pygae1d_nolp_misalign, realhf/impl/model/utils/ppo_functional.py:292-323 (same op order:
 * delta = (r + g*nv) - v  :316 ; last = delta + gl*last  :317 ; ret = last + v  :319).
 * g = (float)gamma, gl = (float)(gamma*lam) computed by the caller in double like Python. */
void oracle_gae1d_nolp_misalign(const float *rewards, const float *values, const int32_t *cu,
                                const uint8_t *bootstrap, int64_t bs, float g, float gl,
                                float *adv, float *ret) {
  for (int64_t i = 0; i < bs; ++i) {
    int64_t rs = cu[i], re = cu[i + 1], vs = rs + i, n = re - rs;
    float last = 0.0f;
    for (int64_t t = n - 1; t >= 0; --t) {
      float nv = values[vs + t + 1];
      if (t == n - 1) nv = nv * (bootstrap[i] ? 1.0f : 0.0f);
      float t1 = g * nv;
      float t2 = rewards[rs + t] + t1;
      float delta = t2 - values[vs + t];
      float t3 = gl * last;
      last = delta + t3;
      adv[rs + t] = last;
      ret[rs + t] = last + values[vs + t];
    }
  }
}

/* PPOActor.compute_advantages GAE loop, areal/engine/ppo/actor.py:131-152. */
void oracle_gae_padded_masked(const float *rewards, const float *values, const float *mask,
                              const uint8_t *no_eos, int64_t B, int64_t S, float g, float gl,
                              float *adv, float *ret) {
  for (int64_t b = 0; b < B; ++b) {
    const float *r = rewards + b * S, *v = values + b * S, *m = mask + b * S;
    float *a = adv + b * S, *o = ret + b * S;
    float last = 0.0f;
    float nv = v[S - 1] * (no_eos[b] ? 1.0f : 0.0f);
    a[S - 1] = 0.0f;
    for (int64_t t = S - 2; t >= 0; --t) {
      float t1 = g * nv;
      float t2 = r[t] + t1;
      float delta = t2 - v[t];
      float t3 = gl * last;
      float nw = delta + t3;
      float mk = m[t], om = 1.0f - mk;
      float x1 = nv * om, x2 = v[t] * mk;
      nv = x1 + x2;
      float y1 = last * om, y2 = nw * mk;
      last = y1 + y2;
      a[t] = last;
    }
    for (int64_t t = 0; t < S; ++t) o[t] = a[t] + v[t];
  }
}

/* _slice_intervals_py, realhf/impl/model/nn/flatten_param.py:65-75 (byte copy, given order). */
void oracle_slice_intervals(const uint8_t *src, uint8_t *packed, const int64_t *iv, int64_t n,
                            int64_t elem_size) {
  int64_t off = 0;
  for (int64_t i = 0; i < n; ++i) {
    int64_t len = (iv[2 * i + 1] - iv[2 * i]) * elem_size;
    memcpy(packed + off, src + iv[2 * i] * elem_size, (size_t)len);
    off += len;
  }
}

/* _set_intervals_py, flatten_param.py:78-90. */
void oracle_set_intervals(const uint8_t *packed, uint8_t *dst, const int64_t *iv, int64_t n,
                          int64_t elem_size) {
  int64_t off = 0;
  for (int64_t i = 0; i < n; ++i) {
    int64_t len = (iv[2 * i + 1] - iv[2 * i]) * elem_size;
    memcpy(dst + iv[2 * i] * elem_size, packed + off, (size_t)len);
    off += len;
  }
}

/* merge_intervals, csrc/interval_op/interval_op.cpp:4-25: neighbours in input order only.
 * Returns the number of merged intervals written to `out` (capacity n). */
int64_t oracle_merge_intervals(const uint64_t *iv, int64_t n, uint64_t *out) {
  int64_t k = 0;
  for (int64_t i = 0; i < n; ++i) {
    if (k > 0 && out[2 * (k - 1) + 1] == iv[2 * i]) {
      out[2 * (k - 1) + 1] = iv[2 * i + 1];
    } else {
      out[2 * k] = iv[2 * i];
      out[2 * k + 1] = iv[2 * i + 1];
      ++k;
    }
  }
  return k;
}
