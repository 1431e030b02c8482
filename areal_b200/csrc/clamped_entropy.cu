// Token log-prob + CLAMPED softmax entropy (AEnt recipe) and their gradient.
//
// Replaces (reference, relative to its root):
//   recipe/AEnt/functional.py:15-36   clamped_softmax_entropy: the k = int(V * clamp) smallest logits of
//       every row are masked out before the softmax entropy.  The reference finds them with a top-k on a
//       CPU copy of the logits (`logits.cpu()`, :18-27) and materialises a [T,V] bool mask and five
//       [T,V] float temporaries per 1024-row chunk.
//   recipe/AEnt/functional.py:39-80   gather_logprobs_clamped_entropy (log-softmax gather + the above)
//
// Design.  Only the VALUE of the k-th smallest logit matters: equal logits are interchangeable for the
// entropy, so a row needs (tau, r) = (the largest removed value, how many elements equal to tau are
// removed).  16-bit logits have 16-bit order-preserving keys, so tau comes from a two-level radix select
// (high byte, then low byte inside the chosen bucket).  The histograms are PRIVATE per thread in shared
// memory ([bin][thread] u16, 128 KB: no contention -- the high byte of a bf16 key is so concentrated
// that a shared histogram would serialise), bumped with result-less shared atomics, summed with a rotated read.  Pass 1 also takes the
// row max; pass 3 accumulates the full and the clamped softmax sums together, so one kernel yields
// log-prob and clamped entropy from three reads of the row, the last two out of L2 (one CTA per row,
// 148 rows = 45 MB in flight).  Elements equal to tau carry the weight f = (n_tau - r) / n_tau in the
// sums and in the gradient: the expected value over the reference's arbitrary tie-breaking (its top-k
// picks SOME r of the n_tau equal elements), identical per class of equal logits.
#include "clamped_entropy.h"

#include "logits_device.cuh"

namespace ab {
namespace {

constexpr int kCeThreads = 256;     // backward kernel
constexpr int kFwdThreads = 512;    // forward kernel: two threads share a histogram slot
constexpr int kSlots = 256;         // private counters per bin
constexpr int kBins = 256;
constexpr size_t kHistBytes = size_t(kBins) * kSlots * sizeof(uint16_t);   // 128 KB

// order-preserving 16-bit key of a bf16 / fp16 bit pattern (sign-magnitude -> unsigned ascending)
__device__ __forceinline__ uint32_t key16(uint32_t b) { return (b & 0x8000u) ? (~b & 0xffffu) : (b | 0x8000u); }
__device__ __forceinline__ uint32_t key_to_bits(uint32_t key) {
  return (key & 0x8000u) ? (key & 0x7fffu) : (~key & 0xffffu);
}

template <bool kIsBF16>
__device__ __forceinline__ float half_to_f32(uint32_t b) {
  if (kIsBF16) return __uint_as_float(b << 16);
  return __half2float(__ushort_as_half(static_cast<unsigned short>(b)));
}

template <int kCeWarps>
__device__ __forceinline__ float block_max(float v, float* red, int warp, int lane) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float r = red[0];
  for (int w = 1; w < kCeWarps; ++w) r = fmaxf(r, red[w]);
  __syncthreads();
  return r;
}
template <int kCeWarps>
__device__ __forceinline__ float block_sum(float v, float* red, int warp, int lane) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float r = 0.f;
  for (int w = 0; w < kCeWarps; ++w) r += red[w];      // fixed order
  __syncthreads();
  return r;
}

// Sum the private histograms (thread b owns bin b: rotated read, conflict-free), then find the bin that
// holds 0-based rank `rank`: returns that bin, the number of elements in lower bins and the bin's count.
// The search is a warp scan: lane l owns bins 8l .. 8l+7.
__device__ __forceinline__ void select_bin(const uint16_t* hist, int* tot, int tid, int rank, int& bin_out,
                                           int& below_out, int& count_out, int* bcast) {
  const uint32_t* h32 = reinterpret_cast<const uint32_t*>(hist);
  if (tid < kBins) {
    int sum = 0;
#pragma unroll 8
    for (int j = 0; j < kSlots / 2; ++j) {
      const uint32_t w = h32[tid * (kSlots / 2) + ((j + tid) & (kSlots / 2 - 1))];
      sum += int(w & 0xffffu) + int(w >> 16);
    }
    tot[tid] = sum;
  }
  __syncthreads();
  if (tid < 32) {
    int c[8], mine = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      c[j] = tot[tid * 8 + j];
      mine += c[j];
    }
    int incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int o = __shfl_up_sync(0xffffffffu, incl, d);
      if (tid >= d) incl += o;
    }
    int acc = incl - mine;                       // elements in bins below this lane's first bin
    const bool here = rank >= acc && rank < incl;
    if (here) {
      int b = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (b == j && acc + c[j] <= rank) {
          acc += c[j];
          b = j + 1;
        }
      }
      b = b > 7 ? 7 : b;
      bcast[0] = tid * 8 + b;
      bcast[1] = acc;
      int cnt = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (j == b) cnt = c[j];
      bcast[2] = cnt;
    }
  }
  __syncthreads();
  bin_out = bcast[0];
  below_out = bcast[1];
  count_out = bcast[2];
  __syncthreads();
}

__device__ __forceinline__ void clear_hist(uint32_t* hist32, int tid) {
  uint4* h = reinterpret_cast<uint4*>(hist32);
  for (int i = tid; i < int(kHistBytes / 16); i += kFwdThreads) h[i] = make_uint4(0u, 0u, 0u, 0u);
  __syncthreads();
}

// result-less add on a 32-bit SHARED address (RED.shared: no generic-address path, nothing to wait for)
__device__ __forceinline__ void red_shared_add(uint32_t saddr, uint32_t v) {
  asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(saddr), "r"(v) : "memory");
}

// f(bits, element index) for every element of a row; 128-bit loads when the row allows them
template <bool kVec, int kNT, typename F>
__device__ __forceinline__ void for_each_elem(const uint16_t* __restrict__ row, int64_t V, int tid, F f) {
  if (kVec) {
    const uint4* rv = reinterpret_cast<const uint4*>(row);
    const int64_t nv = V >> 3;
    constexpr int kU = 8;                         // independent 128-bit loads in flight per thread
    for (int64_t v0 = tid; v0 < nv; v0 += kU * kNT) {
      uint4 q[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u)
        if (v0 + u * kNT < nv) q[u] = __ldg(rv + v0 + u * kNT);
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const int64_t v = v0 + u * kNT;
        if (v < nv) {
          const uint32_t w[4] = {q[u].x, q[u].y, q[u].z, q[u].w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            f(w[j] & 0xffffu, v * 8 + 2 * j);
            f(w[j] >> 16, v * 8 + 2 * j + 1);
          }
        }
      }
    }
  } else {
    for (int64_t i = tid; i < V; i += kNT) f(static_cast<uint32_t>(__ldg(row + i)), i);
  }
}

// per-row results kept for the backward: [0] row max, [1] ln S_full, [2] ln S_kept, [3] H_kept,
// [4] tau key (as float bits), [5] tie weight f, [6] row min, [7] unused
constexpr int kSaved = 8;

template <bool kIsBF16, bool kVec>
__global__ void __launch_bounds__(kFwdThreads, 1)
clamped_entropy_fwd_kernel(const ClampedParams p) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  uint16_t* hist = reinterpret_cast<uint16_t*>(smem_raw);
  // [bin][thread] u16 counters, two threads per 32-bit word; incremented with result-less shared atomics
  // (no read-modify-write chain through shared memory; a thread's count per row stays below 65536)
  const int slot = threadIdx.x & (kSlots - 1);     // threads t and t + 256 share a counter
  const uint32_t my_inc = (slot & 1) ? 65536u : 1u;
  static_assert(kSlots / 2 * 4 == 512, "a bin is 512 bytes of counters");
  const uint32_t my_word = smem_u32(smem_raw) + uint32_t(slot >> 1) * 4u;       // + bin * 512
  __shared__ int tot[kBins];
  __shared__ int bcast[4];
  __shared__ float red[kFwdThreads / 32];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const float inv_t = p.inv_temperature;
  const float c2 = inv_t * kLog2e;
  for (int64_t row = blockIdx.x; row < p.T; row += gridDim.x) {
    if (p.only_flagged && p.saved[row * kSaved + 7] == 0.f) continue;    // done by the one-pass kernel
    const uint16_t* x = reinterpret_cast<const uint16_t*>(reinterpret_cast<const char*>(p.logits) + row * p.row_stride_bytes);
    // ---- pass 1: row max + histogram of the high key byte ----
    clear_hist(reinterpret_cast<uint32_t*>(hist), tid);
    uint32_t kmax = 0, kmin = 0xffffu;             // the key order is the value order: max / min on keys
    for_each_elem<kVec, kFwdThreads>(x, p.V, tid, [&](uint32_t b, int64_t) {
      const uint32_t key = key16(b);
      kmax = max(kmax, key);
      kmin = min(kmin, key);
      red_shared_add(my_word + ((key >> 8) << 9), my_inc);
    });
    const uint32_t bmax = (kmax & 0x8000u) ? (kmax & 0x7fffu) : (~kmax & 0xffffu);
    const float mx = kmax ? half_to_f32<kIsBF16>(bmax) : -INFINITY;
    const uint32_t bmin = (kmin & 0x8000u) ? (kmin & 0x7fffu) : (~kmin & 0xffffu);
    __syncthreads();
    const float m = block_max<kFwdThreads / 32>(mx, red, warp, lane);
    const float row_min = -block_max<kFwdThreads / 32>(-half_to_f32<kIsBF16>(bmin), red, warp, lane);
    uint32_t tau_key = 0;        // keys <= tau_key are candidates for removal
    float f_tie = 1.f;           // weight of elements with key == tau_key
    if (p.k > 0) {
      int hi_bin, below, cnt;
      select_bin(hist, tot, tid, static_cast<int>(p.k - 1), hi_bin, below, cnt, bcast);
      // ---- pass 2: histogram of the low key byte inside that bucket ----
      clear_hist(reinterpret_cast<uint32_t*>(hist), tid);
      for_each_elem<kVec, kFwdThreads>(x, p.V, tid, [&](uint32_t b, int64_t) {
        const uint32_t key = key16(b);
        if (int(key >> 8) == hi_bin) red_shared_add(my_word + ((key & 0xffu) << 9), my_inc);
      });
      __syncthreads();
      int lo_bin, below2, n_tau;
      select_bin(hist, tot, tid, static_cast<int>(p.k - 1) - below, lo_bin, below2, n_tau, bcast);
      tau_key = (uint32_t(hi_bin) << 8) | uint32_t(lo_bin);
      const int removed_at_tau = static_cast<int>(p.k) - (below + below2);        // 1 .. n_tau
      f_tie = float(n_tau - removed_at_tau) / float(n_tau);
    }
    // ---- pass 3: full and clamped softmax sums relative to the row max ----
    const int64_t lab = p.labels[row];
    float s_full = 0.f, s_kept = 0.f, w_kept = 0.f;
    const float xl = (tid == 0 && lab >= 0 && lab < p.V) ? half_to_f32<kIsBF16>(__ldg(x + lab)) : 0.f;
    for_each_elem<kVec, kFwdThreads>(x, p.V, tid, [&](uint32_t b, int64_t) {
      const float xv = half_to_f32<kIsBF16>(b);
      const float t = (xv - m) * c2;                       // <= 0, log2 units
      const float e = ex2_approx(t);
      s_full += e;
      const uint32_t key = key16(b);
      const float wgt = p.k == 0 ? 1.f : (key > tau_key ? 1.f : (key == tau_key ? f_tie : 0.f));
      s_kept = fmaf(wgt, e, s_kept);
      w_kept = fmaf(wgt * e, t, w_kept);
    });
    const float S = block_sum<kFwdThreads / 32>(s_full, red, warp, lane);
    const float SK = block_sum<kFwdThreads / 32>(s_kept, red, warp, lane);
    const float WK = block_sum<kFwdThreads / 32>(w_kept, red, warp, lane);
    const float XL = block_sum<kFwdThreads / 32>(xl, red, warp, lane);
    if (tid == 0) {
      const float ln_s = lg2_approx(S) * kLn2;
      const float ln_sk = lg2_approx(SK) * kLn2;
      const bool has_label = lab >= 0 && lab < p.V;
      p.logp[row] = has_label ? (XL - m) * inv_t - ln_s : __int_as_float(0x7fc00000);
      const float h = ln_sk - kLn2 * (WK / SK);
      p.entropy[row] = h;
      float* sv = p.saved + row * kSaved;
      sv[0] = m;
      sv[1] = ln_s;
      sv[2] = ln_sk;
      sv[3] = h;
      sv[4] = __uint_as_float(p.k == 0 ? 0xffffffffu : tau_key);
      sv[5] = f_tie;
      sv[6] = row_min;
      sv[7] = 0.f;
    }
  }
}

// dlogits_i = [ gl * (1[i = label] - p_i) + gh * w_i * (-q_i * (ln q_i + H)) ] / temp
//   p_i = softmax over the whole row, q_i = softmax over the kept set (w_i = 1, f at tau, 0 removed)
template <bool kIsBF16, bool kVec>
__global__ void __launch_bounds__(kCeThreads)
clamped_entropy_bwd_kernel(const ClampedParams p, const float* __restrict__ g_logp, const float* __restrict__ g_ent,
                           void* __restrict__ dlogits, int64_t d_row_stride_bytes) {
  const int tid = threadIdx.x;
  const float inv_t = p.inv_temperature;
  const float c2 = inv_t * kLog2e;
  for (int64_t row = blockIdx.x; row < p.T; row += gridDim.x) {
    const uint16_t* x = reinterpret_cast<const uint16_t*>(reinterpret_cast<const char*>(p.logits) + row * p.row_stride_bytes);
    uint16_t* d = reinterpret_cast<uint16_t*>(reinterpret_cast<char*>(dlogits) + row * d_row_stride_bytes);
    const float gl = (g_logp ? g_logp[row] : 0.f) * inv_t;
    const float gh = (g_ent ? g_ent[row] : 0.f) * inv_t;
    const float* sv = p.saved + row * kSaved;
    const float m = sv[0], ln_s = sv[1], ln_sk = sv[2], h = sv[3], f_tie = sv[5];
    const uint32_t tau_key = __float_as_uint(sv[4]);
    const bool no_clamp = tau_key == 0xffffffffu;
    const float nl2s = -ln_s * kLog2e, nl2sk = -ln_sk * kLog2e;
    const int64_t lab = p.labels[row];
    auto grad = [&](uint32_t b, int64_t i) -> uint32_t {
      const float xv = half_to_f32<kIsBF16>(b);
      const float t = (xv - m) * c2;
      float g = 0.f;
      if (gl != 0.f) {
        g = -gl * ex2_approx(t + nl2s);
        if (i == lab) g += gl;
      }
      if (gh != 0.f) {
        const uint32_t key = key16(b);
        const float wgt = no_clamp ? 1.f : (key > tau_key ? 1.f : (key == tau_key ? f_tie : 0.f));
        if (wgt != 0.f) {
          const float lq2 = t + nl2sk;                         // log2 q_i
          const float q = ex2_approx(lq2);
          g -= gh * wgt * q * fmaf(lq2, kLn2, h);
        }
      }
      if (kIsBF16) return static_cast<uint32_t>(__bfloat16_as_ushort(__float2bfloat16_rn(g)));
      return static_cast<uint32_t>(__half_as_ushort(__float2half_rn(g)));
    };
    if (gl == 0.f && gh == 0.f) {
      if (kVec) {
        uint4* dv = reinterpret_cast<uint4*>(d);
        for (int64_t v = tid; v < (p.V >> 3); v += kCeThreads) dv[v] = make_uint4(0u, 0u, 0u, 0u);
      } else {
        for (int64_t i = tid; i < p.V; i += kCeThreads) d[i] = 0;
      }
      continue;
    }
    if (kVec) {
      const uint4* rv = reinterpret_cast<const uint4*>(x);
      uint4* dv = reinterpret_cast<uint4*>(d);
      constexpr int kU = 4;
      const int64_t nv = p.V >> 3;
      // Lean path: the row-uniform tests (which gradients exist, clamp or not) and the label test leave the element
      // loop; the threshold compares are float compares against the threshold VALUE, which order elements exactly
      // like the 16-bit keys unless the threshold is a zero (-0 and +0 are different keys: such rows take the
      // general loop below), and both exponents are one FFMA + MUFU.
      const float tau_f = no_clamp ? -INFINITY : half_to_f32<kIsBF16>(key_to_bits(tau_key));
      if (gh != 0.f && (no_clamp || tau_f != 0.f)) {
        const float b1 = fmaf(-m, c2, nl2s), b2 = fmaf(-m, c2, nl2sk);
        const float ngl = -gl, ngh = -gh, ngh_tie = no_clamp ? -gh : -gh * f_tie;
        const int64_t lab_vec = (lab >= 0 && lab < p.V) ? (lab >> 3) : -1;
        auto lean = [&](uint32_t b) -> float {
          const float xv = half_to_f32<kIsBF16>(b);
          const float lq2 = fmaf(xv, c2, b2);
          const float q = ex2_approx(lq2);
          const float e = ex2_approx(fmaf(xv, c2, b1));
          float coef = xv > tau_f ? ngh : 0.f;           // kept / removed; the tie class gets its share
          coef = xv == tau_f ? ngh_tie : coef;
          return fmaf(coef * q, fmaf(lq2, kLn2, h), ngl * e);
        };
        for (int64_t v0 = tid; v0 < nv; v0 += kU * kCeThreads) {
          uint4 q[kU];
#pragma unroll
          for (int u = 0; u < kU; ++u)
            if (v0 + u * kCeThreads < nv) q[u] = __ldg(rv + v0 + u * kCeThreads);
#pragma unroll
          for (int u = 0; u < kU; ++u) {
            const int64_t v = v0 + u * kCeThreads;
            if (v < nv) {
              const uint32_t w[4] = {q[u].x, q[u].y, q[u].z, q[u].w};
              float g8[8];
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                g8[2 * j] = lean(w[j] & 0xffffu);
                g8[2 * j + 1] = lean(w[j] >> 16);
              }
              if (v == lab_vec) {
#pragma unroll
                for (int j = 0; j < 8; ++j)
                  if (j == int(lab & 7)) g8[j] += gl;
              }
              uint32_t o[4];
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                if (kIsBF16) {
                  o[j] = pack_bf16x2(g8[2 * j], g8[2 * j + 1]);
                } else {
                  const __half2 hh = __floats2half2_rn(g8[2 * j], g8[2 * j + 1]);
                  o[j] = *reinterpret_cast<const uint32_t*>(&hh);
                }
              }
              dv[v] = make_uint4(o[0], o[1], o[2], o[3]);
            }
          }
        }
        continue;
      }
      for (int64_t v0 = tid; v0 < nv; v0 += kU * kCeThreads) {
        uint4 q[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u)
          if (v0 + u * kCeThreads < nv) q[u] = __ldg(rv + v0 + u * kCeThreads);
#pragma unroll
        for (int u = 0; u < kU; ++u) {
          const int64_t v = v0 + u * kCeThreads;
          if (v < nv) {
            const uint32_t w[4] = {q[u].x, q[u].y, q[u].z, q[u].w};
            uint32_t o[4];
#pragma unroll
            for (int j = 0; j < 4; ++j)
              o[j] = grad(w[j] & 0xffffu, v * 8 + 2 * j) | (grad(w[j] >> 16, v * 8 + 2 * j + 1) << 16);
            dv[v] = make_uint4(o[0], o[1], o[2], o[3]);
          }
        }
      }
    } else {
      for (int64_t i = tid; i < p.V; i += kCeThreads) d[i] = static_cast<uint16_t>(grad(__ldg(x + i), i));
    }
  }
}

// ---- one-pass forward (16-byte aligned rows) -------------------------------------------------------------
// The three-pass kernel above reads a row three times and bumps a 128 KB private histogram per element.  Here
// the row is read ONCE:
//   1. 2,048 elements sampled evenly over the row (256 vectors) give, by a two-level select on the SAMPLE, two
//      keys that bracket the k-th smallest logit with ~4.6 sigma to spare (sample ranks r +- 96);
//   2. one pass over the row accumulates the full softmax sums online, the sums and the count of everything
//      BELOW the bracket (all of it removed), and an exact histogram of the 16-bit keys INSIDE the bracket
//      (a few thousand elements, spread over up to 1,024 bins: hardly any shared-memory conflicts);
//   3. the threshold key, the tie count and the removed sums inside the bracket follow from that histogram
//      (all elements of one bin have the same value), and S_kept = S_all - S_removed.
// Counts are exact, so (tau, tie weight) are the three-pass kernel's.  A row whose bracket misses the threshold
// (the sample was not representative) or spans more than 1,024 keys is flagged in saved[row][7] and redone by
// the three-pass kernel, launched right after with only_flagged = 1.
constexpr int kF1Threads = 256;
constexpr int kF1Warps = kF1Threads / 32;
constexpr int kSampleVecs = 256;      // sampled 16-byte vectors per row (8 elements each)
constexpr int kBracket = 96;          // half width of the bracket in sample ranks (sigma of the rank is ~21)
constexpr int kWinMax = 1024;         // 16-bit keys a bracket may span

// warp 0: lane l owns bins 8l .. 8l+7 of a 256-bin histogram; bin holding 0-based `rank` -> bcast[0], elements below -> bcast[1]
__device__ __forceinline__ void find_bin256(const int* hist, int rank, int* bcast, int lane) {
  int c[8], mine = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    c[j] = hist[lane * 8 + j];
    mine += c[j];
  }
  int incl = mine;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int o = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += o;
  }
  int acc = incl - mine;
  if (rank >= acc && rank < incl) {
    int b = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (b == j && acc + c[j] <= rank) {
        acc += c[j];
        b = j + 1;
      }
    }
    bcast[0] = lane * 8 + (b > 7 ? 7 : b);
    bcast[1] = acc;
  }
}

// key of the sample element with 0-based rank `rank` (two-level select over samp[2048]); all 256 threads call it
__device__ __forceinline__ uint32_t sample_select(const uint16_t* samp, int* hist, int* bcast, int tid, int rank) {
  hist[tid] = 0;
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 8; ++j) atomicAdd(&hist[samp[tid * 8 + j] >> 8], 1);
  __syncthreads();
  if (tid < 32) find_bin256(hist, rank, bcast, tid);
  __syncthreads();
  const int hb = bcast[0], below = bcast[1];
  __syncthreads();
  hist[tid] = 0;
  __syncthreads();
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const uint32_t key = samp[tid * 8 + j];
    if (int(key >> 8) == hb) atomicAdd(&hist[key & 0xffu], 1);
  }
  __syncthreads();
  if (tid < 32) find_bin256(hist, rank - below, bcast, tid);
  __syncthreads();
  const uint32_t key = (uint32_t(hb) << 8) | uint32_t(bcast[0]);
  __syncthreads();
  return key;
}

__device__ __forceinline__ int block_sum_int(int v, int* red_i, int warp, int lane) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if (lane == 0) red_i[warp] = v;
  __syncthreads();
  int r = 0;
  for (int w = 0; w < kF1Warps; ++w) r += red_i[w];
  __syncthreads();
  return r;
}

template <bool kIsBF16, bool kClamp>        // kClamp = (k > 0): without it the pass is a plain online softmax
__global__ void __launch_bounds__(kF1Threads, 4)
clamped_entropy_fwd1_kernel(const ClampedParams p) {
  __shared__ __align__(16) uint16_t samp[kSampleVecs * 8];
  __shared__ int hist[256];
  __shared__ int win[kWinMax + kF1Threads];   // + one private slot per thread: where elements outside the window count
  __shared__ int bcast[4];
  __shared__ float red[kF1Warps];
  __shared__ int red_i[kF1Warps];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const float inv_t = p.inv_temperature;
  const float c2 = inv_t * kLog2e;
  const int64_t nv = p.V >> 3;
  // 32-bit shared address of the window, produced by volatile asm so that it is computed ONCE (the compiler
  // otherwise rebuilds it -- four uniform instructions -- in front of every atomic)
  uint32_t win_base;
  asm volatile("{\n\t.reg .u64 t;\n\tcvta.to.shared.u64 t, %1;\n\tcvt.u32.u64 %0, t;\n\t}" : "=r"(win_base) : "l"(win));
  for (int64_t row = blockIdx.x; row < p.T; row += gridDim.x) {
    const uint16_t* x = reinterpret_cast<const uint16_t*>(reinterpret_cast<const char*>(p.logits) + row * p.row_stride_bytes);
    const uint4* rv = reinterpret_cast<const uint4*>(x);
    float* sv = p.saved + row * kSaved;
    // ---- 1. bracket from a sample ----
    uint32_t key_lo = 0, key_hi = 0xffffu;
    bool slow = false;
    if (kClamp) {
      {
        const uint4 q = __ldg(rv + (int64_t(tid) * nv) / kSampleVecs);
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          samp[tid * 8 + 2 * j] = static_cast<uint16_t>(key16(w[j] & 0xffffu));
          samp[tid * 8 + 2 * j + 1] = static_cast<uint16_t>(key16(w[j] >> 16));
        }
      }
      __syncthreads();
      const int n_s = kSampleVecs * 8;
      const int r_s = static_cast<int>(((p.k - 1) * int64_t(n_s)) / p.V);
      if (r_s - kBracket >= 0) key_lo = sample_select(samp, hist, bcast, tid, r_s - kBracket);
      if (r_s + kBracket < n_s) key_hi = sample_select(samp, hist, bcast, tid, r_s + kBracket);
      slow = key_hi - key_lo >= uint32_t(kWinMax);
      for (int i = tid; i < kWinMax; i += kF1Threads) win[i] = 0;
      __syncthreads();
    }
    if (slow) {                                      // uniform: the bracket is too wide for the window
      if (tid == 0) sv[7] = 1.f;
      continue;
    }
    const uint32_t span = key_hi - key_lo;           // window = keys key_lo .. key_hi
    // ---- 2. the one pass over the row ----
    float m_t = -INFINITY, s_all = 0.f, w_all = 0.f, s_bel = 0.f, w_bel = 0.f;
    int cnt_bel = 0;
    uint32_t kmin = 0xffffu;
    uint32_t pmin = 0x7f807f80u;                     // bf16: packed (+inf, +inf)
    const uint32_t my_slot = span + 1u + uint32_t(tid);   // (span < kWinMax)
    constexpr int kU = 4;
    for (int64_t v0 = tid; v0 < nv; v0 += kU * kF1Threads) {
      uint4 q[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u)
        if (v0 + u * kF1Threads < nv) q[u] = __ldg(rv + v0 + u * kF1Threads);
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        if (v0 + u * kF1Threads < nv) {
          const uint32_t w[4] = {q[u].x, q[u].y, q[u].z, q[u].w};
          float xv[8];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            xv[2 * j] = half_to_f32<kIsBF16>(w[j] & 0xffffu);
            xv[2 * j + 1] = half_to_f32<kIsBF16>(w[j] >> 16);
          }
          float vmax = fmaxf(fmaxf(fmaxf(xv[0], xv[1]), fmaxf(xv[2], xv[3])), fmaxf(fmaxf(xv[4], xv[5]), fmaxf(xv[6], xv[7])));
          if (kIsBF16) pmin = min_bf16x2(pmin, dev::vec_min_bf16(q[u]));   // row minimum: packed, on the raw words
          if (vmax > m_t) {
            if (m_t != -INFINITY) {
              const float d = (m_t - vmax) * c2;
              const float f = ex2_approx(d);
              w_all = f * fmaf(d, s_all, w_all);
              s_all *= f;
              w_bel = f * fmaf(d, s_bel, w_bel);
              s_bel *= f;
            }
            m_t = vmax;
          }
          const float nb = -m_t * c2;
          // both keys of a 32-bit word at once: key = bits ^ 0x8000 ^ (sign ? 0x7fff : 0)
          uint32_t k2[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) k2[j] = w[j] ^ 0x80008000u ^ (((w[j] >> 15) & 0x00010001u) * 0x7fffu);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t key = (j & 1) ? (k2[j >> 1] >> 16) : (k2[j >> 1] & 0xffffu);
            const float t = fmaf(xv[j], c2, nb);
            const float e = ex2_approx(t);
            s_all += e;
            w_all = fmaf(e, t, w_all);
            if (!kIsBF16) kmin = min(kmin, key);
            if (kClamp) {
              if (key < key_lo) {                      // (predicated adds)
                s_bel += e;
                w_bel = fmaf(e, t, w_bel);
                cnt_bel += 1;
              }
              // every element bumps a counter, unconditionally (ptxas turns a predicated shared atomic into a branch with
              // a convergence barrier: four issue slots): its window bin, or -- outside the window, where key - key_lo
              // wraps or exceeds the span -- this thread's private slot behind the window
              const uint32_t off = key - key_lo;
              const uint32_t idx = min(off, my_slot);           // my_slot > span: bins beyond the span are never read
              asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(win_base + idx * 4u) : "memory");
            }
          }
        }
      }
    }
    // ---- 3. row maximum, sums rebased to it ----
    const float m = block_max<kF1Warps>(m_t, red, warp, lane);
    if (m_t != -INFINITY) {
      const float d = (m_t - m) * c2;
      const float f = ex2_approx(d);
      w_all = f * fmaf(d, s_all, w_all);
      s_all *= f;
      w_bel = f * fmaf(d, s_bel, w_bel);
      s_bel *= f;
    }
    const float S = block_sum<kF1Warps>(s_all, red, warp, lane);
    const float Wa = block_sum<kF1Warps>(w_all, red, warp, lane);
    float s_rem = block_sum<kF1Warps>(s_bel, red, warp, lane);
    float w_rem = block_sum<kF1Warps>(w_bel, red, warp, lane);
    const int below = block_sum_int(cnt_bel, red_i, warp, lane);
    const float tmin = kIsBF16 ? fminf(bf16_lo(pmin), bf16_hi(pmin)) : half_to_f32<kIsBF16>(key_to_bits(kmin));
    const float row_min = -block_max<kF1Warps>(-tmin, red, warp, lane);
    uint32_t tau_key = 0;
    float f_tie = 1.f;
    if (kClamp) {
      // threshold inside the window: thread t owns bins 4t .. 4t+3
      int c[4], mine = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        c[j] = uint32_t(tid * 4 + j) <= span ? win[tid * 4 + j] : 0;   // beyond the span: other elements' overflow slots
        mine += c[j];
      }
      int incl = mine;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += o;
      }
      if (lane == 31) red_i[warp] = incl;
      __syncthreads();
      int base = 0, total = 0;
      for (int w2 = 0; w2 < kF1Warps; ++w2) {
        if (w2 < warp) base += red_i[w2];
        total += red_i[w2];
      }
      const int excl = base + incl - mine;
      const int rank_in = static_cast<int>(p.k - 1) - below;
      if (tid == 0) bcast[2] = (rank_in < 0 || rank_in >= total) ? 1 : 0;
      if (rank_in >= excl && rank_in < excl + mine) {
        int acc = excl, b = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (b == j && acc + c[j] <= rank_in) {
            acc += c[j];
            b = j + 1;
          }
        }
        b = b > 3 ? 3 : b;
        bcast[0] = tid * 4 + b;      // bin of the threshold key
        bcast[1] = acc;              // window elements below that bin
      }
      __syncthreads();
      const bool miss = bcast[2] != 0;
      const int jbin = bcast[0], prefix = bcast[1];
      __syncthreads();
      if (miss) {                                    // uniform: the sample misled the bracket
        if (tid == 0) sv[7] = 1.f;
        continue;
      }
      tau_key = key_lo + uint32_t(jbin);
      const int n_tau = win[jbin];
      const int removed_at_tau = static_cast<int>(p.k) - (below + prefix);       // 1 .. n_tau
      f_tie = float(n_tau - removed_at_tau) / float(n_tau);
      // removed mass inside the window: whole bins below the threshold bin, a share of the threshold bin
      float sr = 0.f, wr = 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int bin = tid * 4 + j;
        if (bin <= jbin && c[j] > 0) {
          const float xvv = half_to_f32<kIsBF16>(key_to_bits(key_lo + uint32_t(bin)));
          const float t = (xvv - m) * c2;
          const float e = ex2_approx(t);
          const float wgt = bin < jbin ? float(c[j]) : float(removed_at_tau);
          sr = fmaf(wgt, e, sr);
          wr = fmaf(wgt * e, t, wr);
        }
      }
      s_rem += block_sum<kF1Warps>(sr, red, warp, lane);
      w_rem += block_sum<kF1Warps>(wr, red, warp, lane);
    } else {
      s_rem = 0.f;
      w_rem = 0.f;
    }
    if (tid == 0) {
      const int64_t lab = p.labels[row];
      const bool has_label = lab >= 0 && lab < p.V;
      const float XL = has_label ? half_to_f32<kIsBF16>(__ldg(x + lab)) : 0.f;
      const float SK = S - s_rem, WK = Wa - w_rem;
      const float ln_s = lg2_approx(S) * kLn2;
      const float ln_sk = lg2_approx(SK) * kLn2;
      p.logp[row] = has_label ? (XL - m) * inv_t - ln_s : __int_as_float(0x7fc00000);
      const float h = ln_sk - kLn2 * (WK / SK);
      p.entropy[row] = h;
      sv[0] = m;
      sv[1] = ln_s;
      sv[2] = ln_sk;
      sv[3] = h;
      sv[4] = __uint_as_float(kClamp ? tau_key : 0xffffffffu);
      sv[5] = f_tie;
      sv[6] = row_min;
      sv[7] = 0.f;
    }
    __syncthreads();
  }
}

bool vec_ok(const void* ptr, int64_t V, int64_t stride_bytes) {
  return (reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && (V % 8) == 0 && (stride_bytes % 16) == 0;
}

}  // namespace

int launch_clamped_entropy_fwd(const ClampedParams& p_in, cudaStream_t stream) {
  if (p_in.T == 0) return kOk;
  const bool vec = vec_ok(p_in.logits, p_in.V, p_in.row_stride_bytes);
  const int64_t cap = num_sms();
  const unsigned int grid = static_cast<unsigned int>(p_in.T < cap ? p_in.T : cap);
  ClampedParams p = p_in;
  p.only_flagged = 0;
  static const bool one_pass = [] {
    const char* e = getenv("AREAL_B200_AENT_ONE_PASS");
    return !(e && e[0] == '0');
  }();
  if (vec && one_pass) {
    // one read of the row; rows it cannot settle are flagged in saved[row][7] and redone by the three-pass kernel
    const int64_t cap1 = int64_t(num_sms()) * 4;
    const unsigned int grid1 = static_cast<unsigned int>(p.T < cap1 ? p.T : cap1);
    if (p.dtype == kBF16) {
      if (p.k > 0) clamped_entropy_fwd1_kernel<true, true><<<grid1, kF1Threads, 0, stream>>>(p);
      else clamped_entropy_fwd1_kernel<true, false><<<grid1, kF1Threads, 0, stream>>>(p);
    } else {
      if (p.k > 0) clamped_entropy_fwd1_kernel<false, true><<<grid1, kF1Threads, 0, stream>>>(p);
      else clamped_entropy_fwd1_kernel<false, false><<<grid1, kF1Threads, 0, stream>>>(p);
    }
    const int rc = check_launch("clamped_entropy_fwd1");
    if (rc) return rc;
    p.only_flagged = 1;
  }
#define AB_CE_FWD(BF, VEC)                                                                                    \
  {                                                                                                           \
    AB_CUDA(cudaFuncSetAttribute(clamped_entropy_fwd_kernel<BF, VEC>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                 int(kHistBytes)));                                                           \
    clamped_entropy_fwd_kernel<BF, VEC><<<grid, kFwdThreads, kHistBytes, stream>>>(p);                         \
  }
  if (p.dtype == kBF16) { if (vec) AB_CE_FWD(true, true) else AB_CE_FWD(true, false) }
  else { if (vec) AB_CE_FWD(false, true) else AB_CE_FWD(false, false) }
#undef AB_CE_FWD
  return check_launch("clamped_entropy_fwd");
}

int launch_clamped_entropy_bwd(const ClampedParams& p, const float* g_logp, const float* g_ent, void* dlogits,
                               int64_t d_row_stride_bytes, cudaStream_t stream) {
  if (p.T == 0) return kOk;
  const bool vec = vec_ok(p.logits, p.V, p.row_stride_bytes) && vec_ok(dlogits, p.V, d_row_stride_bytes);
  const int64_t cap = int64_t(num_sms()) * 8;
  const unsigned int grid = static_cast<unsigned int>(p.T < cap ? p.T : cap);
#define AB_CE_BWD(BF, VEC) \
  clamped_entropy_bwd_kernel<BF, VEC><<<grid, kCeThreads, 0, stream>>>(p, g_logp, g_ent, dlogits, d_row_stride_bytes);
  if (p.dtype == kBF16) { if (vec) AB_CE_BWD(true, true) else AB_CE_BWD(true, false) }
  else { if (vec) AB_CE_BWD(false, true) else AB_CE_BWD(false, false) }
#undef AB_CE_BWD
  return check_launch("clamped_entropy_bwd");
}

}  // namespace ab
