// Fused token log-prob / entropy / (PPO loss) forward and dlogits backward over [T, V] logits.
//
// Replaces (reference, relative to its root):
//   areal/utils/functional.py:20-127   gather_logprobs / gather_logprobs_entropy (+ autograd backward)
//   areal/engine/ppo/actor.py:375-376  logits.min(-1) / logits.max(-1) re-reads
//   areal/utils/mcore/functional.py:8-52 and Megatron vocab_parallel_cross_entropy (per-shard partials)
//   areal/utils/functional.py:171-235  ppo_actor_loss_fn, when the fused epilogue is requested
//
// Design (B200): one streaming pass per row with an online softmax.  Persistent CTAs (grid =
// #SM x resident CTAs) walk rows; a single producer thread stages the row in 16 KB pieces with
// TMA bulk copies (cp.async.bulk -> UBLKCP) into a shared-memory ring guarded by full/empty
// mbarriers; 8 consumer warps read the ring with 128-bit LDS, keep (max, sum e, sum e*x, min) in
// registers and reduce once per row.  No tensor cores: the pass is HBM-bound (2 B/element in,
// 4 B/element in+out for the backward), MUFU.EX2 is the closest secondary limit (1 per element).
// Rows that are not 16-byte aligned / not bf16 take the generic LDG kernel at the end of the file.
#include "common.cuh"
#include "ppo_math.cuh"
#include "logits_fused.h"
#include "logits_device.cuh"

namespace ab {

namespace {

using namespace dev;

constexpr int kConsumerWarps = 8;
constexpr int kConsumers = kConsumerWarps * 32;      // 256
constexpr int kThreads = kConsumers + 32;            // + 1 producer warp
constexpr int kStageBytes = 16384;                   // 8192 bf16 per stage
constexpr int kVecPerThread = kStageBytes / 16 / kConsumers;   // 4 x 16 B per thread per stage
constexpr int kStages = 6;
constexpr int kCtasPerSm = 2;
constexpr size_t kSmemBytes = size_t(kStages) * kStageBytes + 1024;

template <bool kMinMax, int kCons = kConsumers>
__device__ __forceinline__ void consume_full_stage(RowState& st, uint32_t sbase, int tid, float c2) {
  uint4 q[kVecPerThread];
#pragma unroll
  for (int j = 0; j < kVecPerThread; ++j) q[j] = lds_v4(sbase + (tid + j * kCons) * 16);
  uint32_t mx = vec_max_bf16(q[0]);
#pragma unroll
  for (int j = 1; j < kVecPerThread; ++j) mx = max_bf16x2(mx, vec_max_bf16(q[j]));
  if (kMinMax) {
    uint32_t mn = vec_min_bf16(q[0]);
#pragma unroll
    for (int j = 1; j < kVecPerThread; ++j) mn = min_bf16x2(mn, vec_min_bf16(q[j]));
    st.mn = fminf(st.mn, fminf(bf16_lo(mn), bf16_hi(mn)));
  }
  const float nb = state_rebase(st, fmaxf(bf16_lo(mx), bf16_hi(mx)), c2);
  float s0 = 0.f, s1 = 0.f, w0 = 0.f, w1 = 0.f;
#pragma unroll
  for (int j = 0; j < kVecPerThread; ++j) vec_accumulate(q[j], c2, nb, s0, s1, w0, w1);
  st.s += s0 + s1;
  st.w += w0 + w1;
}

// A stage that is not full (the end of a row): the same batch -- one rebase for all the vectors this thread
// has -- with the missing vectors predicated off, instead of one rebase per vector.
template <bool kMinMax, int kCons>
__device__ __forceinline__ void consume_partial_stage(RowState& st, uint32_t sbase, int tid, int nvec, float c2) {
  if (tid >= nvec) return;
  uint4 q[kVecPerThread];
  q[0] = lds_v4(sbase + tid * 16);
  uint32_t mx = vec_max_bf16(q[0]);
  uint32_t mn = kMinMax ? vec_min_bf16(q[0]) : 0u;
#pragma unroll
  for (int j = 1; j < kVecPerThread; ++j) {
    if (tid + j * kCons < nvec) {
      q[j] = lds_v4(sbase + (tid + j * kCons) * 16);
      mx = max_bf16x2(mx, vec_max_bf16(q[j]));
      if (kMinMax) mn = min_bf16x2(mn, vec_min_bf16(q[j]));
    }
  }
  if (kMinMax) st.mn = fminf(st.mn, fminf(bf16_lo(mn), bf16_hi(mn)));
  const float nb = state_rebase(st, fmaxf(bf16_lo(mx), bf16_hi(mx)), c2);
  float s0 = 0.f, s1 = 0.f, w0 = 0.f, w1 = 0.f;
#pragma unroll
  for (int j = 0; j < kVecPerThread; ++j)
    if (tid + j * kCons < nvec) vec_accumulate(q[j], c2, nb, s0, s1, w0, w1);
  st.s += s0 + s1;
  st.w += w0 + w1;
}

struct RowReduceSmem {
  float m[2][kConsumerWarps], s[2][kConsumerWarps], w[2][kConsumerWarps], mn[2][kConsumerWarps];
};

// All consumer threads call; returns true on the single thread (warp 0, lane 0) that holds the row
// totals in st (canonical form).  One named barrier per row (double-buffered by `parity`).
__device__ __forceinline__ bool row_reduce(RowState& st, RowReduceSmem* red, int parity, int warp,
                                           int lane, float c2) {
  state_canonical(st, c2);
  warp_merge_state(st, c2);
  if (lane == 0) {
    red->m[parity][warp] = st.m;
    red->s[parity][warp] = st.s;
    red->w[parity][warp] = st.w;
    red->mn[parity][warp] = st.mn;
  }
  named_bar_sync(1, kConsumers);
  if (warp != 0) return false;
  const int l = lane & (kConsumerWarps - 1);
  st.m = red->m[parity][l];
  st.s = red->s[parity][l];
  st.w = red->w[parity][l];
  st.mn = red->mn[parity][l];
#pragma unroll
  for (int o = kConsumerWarps / 2; o > 0; o >>= 1) {
    const float m2 = __shfl_xor_sync(0xffffffffu, st.m, o);
    const float s2 = __shfl_xor_sync(0xffffffffu, st.s, o);
    const float w2 = __shfl_xor_sync(0xffffffffu, st.w, o);
    const float n2 = __shfl_xor_sync(0xffffffffu, st.mn, o);
    state_merge(st.m, st.s, st.w, m2, s2, w2, c2);
    st.mn = fminf(st.mn, n2);
  }
  return lane == 0;
}

// Rotating variant for the TMA kernel: the finishing work of row number `it` (CTA-local count) belongs to
// warp it % 8.  The other seven warps only ARRIVE on the row's named barrier and go on with the next row;
// the finishing warp waits for them, merges the eight partials and runs the epilogue, so its global loads
// and stores are off the other warps' path and the duty (and its lag) moves from warp to warp.  Eight
// partial slots / barrier ids: warps cannot drift more than kStages (6) rows apart, the ring holds them.
constexpr int kRedSlots = 8;
// What a warp leaves for the finishing warp: its max, its min, and EIGHT partial (s, w) sums relative to its max
// (two butterfly rounds instead of five: the finishing warp's own merge adds up the rest, so the seven warps that
// only pass through a row end spend ~25 instructions there instead of ~60).
template <int kCW>
struct RowReduceRot {
  float2 sw[kRedSlots][kCW][8];
  float m[kRedSlots][kCW], mn[kRedSlots][kCW];
};
template <bool kMinMax, int kCW>
__device__ __forceinline__ bool row_reduce_rot(RowState& st, RowReduceRot<kCW>* red, int64_t it, int warp, int lane,
                                               float c2) {
  constexpr int kConsumerWarps = kCW, kConsumers = kCW * 32;
  state_canonical(st, c2);
  const float M = warp_max_f32(st.m);
  float s, w;
  state_shift_to(st, M, c2, s, w);
  s += __shfl_xor_sync(0xffffffffu, s, 16);
  w += __shfl_xor_sync(0xffffffffu, w, 16);
  s += __shfl_xor_sync(0xffffffffu, s, 8);
  w += __shfl_xor_sync(0xffffffffu, w, 8);
  const float mn = kMinMax ? warp_min_f32(st.mn) : INFINITY;
  const int slot = static_cast<int>(it % kRedSlots);
  if (lane < 8) red->sw[slot][warp][lane] = make_float2(s, w);
  if (lane == 0) {
    red->m[slot][warp] = M;
    if (kMinMax) red->mn[slot][warp] = mn;
  }
  const int fin = static_cast<int>(it % kConsumerWarps);
  if (warp != fin) {
    __syncwarp();
    named_bar_arrive(1 + slot, kConsumers);
    return false;
  }
  named_bar_sync(1 + slot, kConsumers);
  // kCW * 8 partials over 32 lanes
  st.m = -INFINITY;
  st.s = 0.f;
  st.w = 0.f;
#pragma unroll
  for (int e = lane; e < kCW * 8; e += 32) {
    const float2 q = red->sw[slot][e >> 3][e & 7];
    const float m2 = red->m[slot][e >> 3];
    if (e < 32) {
      st.m = m2;
      st.s = q.x;
      st.w = q.y;
    } else {
      state_merge(st.m, st.s, st.w, m2, q.x, q.y, c2);
    }
  }
  st.mn = INFINITY;
  if (kMinMax && lane < kCW) st.mn = red->mn[slot][lane];
  warp_merge_state(st, c2);
  return lane == 0;
}

// ---- push exchange of the vocabulary-sharded forward (NVLink peer memory, plain stores) -----------------
// Every rank owns a symmetric buffer:  u32 rows_in[2] (per slot: how many rows all ranks have delivered so far,
// cumulative over the steps) | u32 pad[2] | u32 ticket[2] (local) | pad to 256 B | float4 parts[2][R][Tmax]  (slot,
// sender, row) = (m, S, W, x_label).
// The row epilogue of the partial kernel stores a row's four numbers into EVERY rank's buffer (one 16-byte store per
// rank, the local one included); when a CTA has finished its rows it fences (system scope) and takes a local ticket, the
// last CTA of the rank tells every rank "T more rows" (one atomic add per rank), and every CTA waits until its own
// buffer's counter says that all R*T rows of the step are in; then the same CTAs combine the R partials of each row -- from LOCAL memory -- in rank order, with the
// arithmetic of vocab_combine_kernel (bit-identical results).  Partial kernel, exchange and combine are one launch and
// the transfer rides under the streaming pass; what is left after the slowest rank's last row is one fence, one
// NVLink round trip and a 16 MB local read.
// A slot is reused two steps later without another handshake: a rank writes slot s of step e+2 only after it has
// seen all rows of step e+1, i.e. after every rank has started step e+1, i.e. after every rank's kernel of step e --
// combine included -- has finished (kernels of one rank run in stream order).
constexpr int64_t kPushHeaderBytes = 256;
__device__ __forceinline__ float4* push_slot_ptr(char* base, int slot, int R, int sender, int64_t tmax, int64_t row) {
  return reinterpret_cast<float4*>(base + kPushHeaderBytes) + (int64_t(slot) * R + sender) * tmax + row;
}
__device__ __forceinline__ void st_weak_v4(float4* q, float a, float b, float c, float d) {
  asm volatile("st.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(q), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
// around L1 (lines of this buffer that L1 may still hold are two steps old)
__device__ __forceinline__ float4 ld_sys_v4(const float4* q) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(q) : "memory");
  return v;
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* q) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(q) : "memory");
  return v;
}
__device__ __forceinline__ void red_add_release_sys(unsigned int* q, unsigned int v) {
  asm volatile("red.release.sys.global.add.u32 [%0], %1;" ::"l"(q), "r"(v) : "memory");
}

// ---- row epilogue (one thread per row) ---------------------------------------------------------
__device__ __forceinline__ float load_logit(const void* base, int dtype, int64_t idx) {
  if (dtype == kBF16) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(base)[idx]);
  if (dtype == kFP16) return __half2float(reinterpret_cast<const __half*>(base)[idx]);
  return reinterpret_cast<const float*>(base)[idx];
}

struct LossAcc {
  float sum;     // sum of masked per-token loss over this CTA's rows (fixed order)
  float count;   // number of masked-in tokens
};

__device__ __forceinline__ void row_epilogue(const LogitsFwdParams& p, int64_t row,
                                             const RowState& st, LossAcc& acc) {
  const float inv_t = p.inv_temperature;
  // label logit (one 2..4-byte gather, normally an L2 hit: the row was just streamed)
  const int64_t lab = p.labels[row] - p.vocab_start;
  const bool local = lab >= 0 && lab < p.V;
  const char* rowp = reinterpret_cast<const char*>(p.logits) + row * p.row_stride_bytes;
  const float xl = local ? load_logit(rowp, p.dtype, lab) : 0.f;
  if (p.raw_min) p.raw_min[row] = st.mn;
  if (p.mode == kModePartial) {
    // (raw max, S, W, raw label logit or 0) for the cross-shard combine
    if (p.push_world > 0) {
      for (int r = 0; r < p.push_world; ++r)
        st_weak_v4(push_slot_ptr(p.push_base[r], p.push_slot, p.push_world, p.push_rank, p.push_tmax, row), st.m,
                   st.s, st.w, xl);
      return;
    }
    p.part_m[row] = st.m;
    p.part_s[row] = st.s;
    p.part_w[row] = st.w;
    p.part_xl[row] = xl;
    return;
  }
  // lse(z) = m/temp + ln S;  E_p[z] = m/temp + ln2 * W/S;  H = lse - E_p[z] = ln S - ln2 * W/S
  const float ln_s = lg2_approx(st.s) * kLn2;
  const float logp = local ? (xl - st.m) * inv_t - ln_s : __int_as_float(0x7fc00000);
  const float ent = ln_s - kLn2 * (st.w / st.s);
  p.logp[row] = logp;
  p.row_max[row] = st.m;
  p.ln_s[row] = ln_s;
  if (p.entropy) p.entropy[row] = ent;
  if (p.mode == kModeFusedActor) {
    const bool mask = p.loss_mask[row] != 0;
    const float prox = p.prox ? p.prox[row] : p.old[row];
    const ActorTokenOut o = actor_token(logp, prox, p.old[row], p.adv[row], mask, p.cfg);
    p.coef[row] = o.coef;
    if (p.stat_loss) p.stat_loss[row] = o.pg;
    if (p.stat_ratio) p.stat_ratio[row] = o.ratio;
    if (p.stat_kl) p.stat_kl[row] = o.kl;
    if (p.clip_mask) p.clip_mask[row] = o.clip;
    if (p.dual_clip_mask) p.dual_clip_mask[row] = o.dual;
    if (p.behave_w) p.behave_w[row] = o.w;
    if (p.behave_kl) p.behave_kl[row] = o.bk;
    if (p.behave_mask) p.behave_mask[row] = o.bmask;
    if (mask) {
      acc.sum += o.pg;
      acc.count += 1.f;
    }
  }
}

// Fixed-order cross-CTA reduction of (loss sum, count): every CTA stores its partial, takes a
// ticket, and the last one sums the partials in index order (deterministic) and resets the ticket.
__device__ __forceinline__ void loss_finalize(const LogitsFwdParams& p, const LossAcc& acc) {
  float2* partials = reinterpret_cast<float2*>(reinterpret_cast<char*>(p.workspace) + 16);
  unsigned int* ticket = reinterpret_cast<unsigned int*>(p.workspace);
  partials[blockIdx.x] = make_float2(acc.sum, acc.count);
  __threadfence();
  const unsigned int t = atomicAdd(ticket, 1u);
  if (t == gridDim.x - 1) {
    __threadfence();
    double sum = 0.0, cnt = 0.0;
    for (unsigned int i = 0; i < gridDim.x; ++i) {
      const volatile float* v = reinterpret_cast<const volatile float*>(&partials[i]);
      sum += v[0];
      cnt += v[1];
    }
    const double denom = cnt > 0.0 ? cnt : 1.0;               // count_nonzero() or 1, functional.py:190
    p.loss_out[0] = static_cast<float>(sum / denom);
    p.inv_count_out[0] = static_cast<float>(1.0 / denom);
    *ticket = 0u;
  }
}

// Second half of the push exchange, run by the consumer threads of every CTA after its last row (see above).
template <int kCons>
__device__ __forceinline__ void push_finish(const LogitsFwdParams& p, unsigned int rows_done, int tid, int* ok_sm) {
  const int R = p.push_world;
  named_bar_sync(1 + kRedSlots, kCons);              // the epilogue stores of all warps of this CTA are issued
  if (tid == 0) {
    __threadfence_system();                          // ... and performed before the ticket that announces them
    // one announcement per RANK, not per CTA (592 CTAs x 8 ranks of same-address atomics over NVLink cost ~25 us at the
    // end of every step): CTAs take a ticket in local memory, the last one tells every rank "T more rows from me"
    unsigned int* ticket = reinterpret_cast<unsigned int*>(p.push_base[p.push_rank]) + 4 + p.push_slot;
    if (atomicAdd(ticket, 1u) == gridDim.x - 1) {
      __threadfence_system();                        // (acquire side of the ticket chain, release side of the adds)
      *ticket = 0u;                                  // next use of this slot: two steps from now
      for (int r = 0; r < R; ++r)
        red_add_release_sys(reinterpret_cast<unsigned int*>(p.push_base[r]) + p.push_slot,
                            static_cast<unsigned int>(p.T));
    }
    const unsigned int* mine = reinterpret_cast<const unsigned int*>(p.push_base[p.push_rank]) + p.push_slot;
    const long long t0 = clock64();
    int ok = 1;
    // counts only grow; the signed difference survives wrap-around
    while (static_cast<int>(ld_acquire_sys(mine) - p.push_target) < 0) {
      if (clock64() - t0 > p.push_timeout) {
        ok = 0;
        break;
      }
      __nanosleep(200);
    }
    if (!ok) atomicExch(p.push_status, 1);
    *ok_sm = ok;
  }
  named_bar_sync(1 + kRedSlots, kCons);
  if (!*ok_sm) return;                               // a rank never delivered: outputs untouched, status says so
  const float c2 = p.inv_temperature * kLog2e;
  const int64_t per = (p.T + gridDim.x - 1) / gridDim.x;
  const int64_t t_end = (blockIdx.x + 1) * per < p.T ? (blockIdx.x + 1) * per : p.T;
  for (int64_t t = blockIdx.x * per + tid; t < t_end; t += kCons) {
    // two passes over the R partials of the row (the second one hits L2): the maximum, then the sums in rank order
    const float4* q0 = push_slot_ptr(p.push_base[p.push_rank], p.push_slot, R, 0, p.push_tmax, t);
    float XL = 0.f, M = -INFINITY;
    for (int r = 0; r < R; ++r) {
      const float4 q = ld_sys_v4(q0 + int64_t(r) * p.push_tmax);
      XL += q.w;
      M = fmaxf(M, q.x);
    }
    float S = 0.f, W = 0.f;
    for (int r = 0; r < R; ++r) {
      const float4 q = ld_sys_v4(q0 + int64_t(r) * p.push_tmax);
      if (q.x != -INFINITY) {
        const float d = (q.x - M) * c2;
        const float f = exp2f(d);
        S = fmaf(q.y, f, S);
        W = fmaf(f, fmaf(d, q.y, q.z), W);
      }
    }
    const float ln_s = logf(S);
    p.logp[t] = (XL - M) * p.inv_temperature - ln_s;
    if (p.entropy) p.entropy[t] = ln_s - kLn2 * (W / S);
    p.row_max[t] = M;
    p.ln_s[t] = ln_s;
  }
}

template <typename E> struct IsBF16 { static constexpr bool value = false; };
template <> struct IsBF16<ElemBF16> { static constexpr bool value = true; };

// ---- TMA-staged forward ------------------------------------------------------------------------
// kCW consumer warps per CTA: 8 (16 KB stages, 2 CTAs/SM) for long rows; 4 (8 KB stages, 4 CTAs/SM) for short
// ones (vocabulary shards), where a row gives each of 256 threads too little to do between two row ends
template <bool kMinMax, typename E, int kCW>
__global__ void __launch_bounds__((kCW + 1) * 32, 16 / kCW)
logits_fwd_tma_kernel(const LogitsFwdParams p) {
  constexpr int kConsumerWarps = kCW, kConsumers = kCW * 32, kStageBytes = kCW * 2048;
  static_assert(kStageBytes / 16 / kConsumers == kVecPerThread, "four vectors per thread and stage");
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[kStages];
  __shared__ __align__(8) uint64_t empty_bar[kStages];
  __shared__ RowReduceRot<kCW> red;
  __shared__ LossAcc acc_sm[kConsumerWarps];

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const uint32_t ring = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int64_t row_bytes = p.V * E::kSize;
  const int nchunk = static_cast<int>((row_bytes + kStageBytes - 1) / kStageBytes);
  const uint32_t tail_bytes = static_cast<uint32_t>(row_bytes - int64_t(nchunk - 1) * kStageBytes);

  if (tid == 0) {
    for (int i = 0; i < kStages; ++i) {
      mbar_init(smem_u32(&full_bar[i]), 1);
      mbar_init(smem_u32(&empty_bar[i]), kConsumerWarps);
    }
    mbar_fence_init();
  }
  __syncthreads();

  if (warp == kConsumerWarps) {
    // ===== producer: one elected lane issues the bulk copies =====
    if (lane == 0) {
      const uint64_t pol = policy_evict_first();
      int stage = 0;
      uint32_t phase = 0;
      for (int64_t row = blockIdx.x; row < p.T; row += gridDim.x) {
        const char* src = reinterpret_cast<const char*>(p.logits) + row * p.row_stride_bytes;
        for (int c = 0; c < nchunk; ++c) {
          mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1u);
          const uint32_t bytes = (c == nchunk - 1) ? tail_bytes : uint32_t(kStageBytes);
          const uint32_t fb = smem_u32(&full_bar[stage]);
          mbar_arrive_expect_tx(fb, bytes);
          tma_load_1d(ring + stage * kStageBytes, src + int64_t(c) * kStageBytes, bytes, fb, pol);
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
    return;
  }

  // ===== consumers =====
  const float c2 = p.inv_temperature * kLog2e;
  int stage = 0;
  uint32_t phase = 0;
  LossAcc acc = {0.f, 0.f};
  int64_t it = 0;
  for (int64_t row = blockIdx.x; row < p.T; row += gridDim.x, ++it) {
    RowState st;
    state_init(st);
    for (int c = 0; c < nchunk; ++c) {
      mbar_wait(smem_u32(&full_bar[stage]), phase);
      const uint32_t sbase = ring + stage * kStageBytes;
      if (c != nchunk - 1 || tail_bytes == kStageBytes) {
        if constexpr (E::kSize == 2 && E::kN == 8 && sizeof(E) == sizeof(ElemBF16) && IsBF16<E>::value) {
          consume_full_stage<kMinMax, kConsumers>(st, sbase, tid, c2);
        } else {
          uint4 q[kVecPerThread];
#pragma unroll
          for (int j = 0; j < kVecPerThread; ++j) q[j] = lds_v4(sbase + (tid + j * kConsumers) * 16);
          consume_batch<E, kMinMax, kVecPerThread>(st, q, c2);
        }
      } else {
        const int nvec = tail_bytes >> 4;
        if constexpr (IsBF16<E>::value) {
          consume_partial_stage<kMinMax, kConsumers>(st, sbase, tid, nvec, c2);
        } else {
#pragma unroll
          for (int j = 0; j < kVecPerThread; ++j) {
            const int v = tid + j * kConsumers;
            if (v < nvec) {
              const uint4 q1[1] = {lds_v4(sbase + v * 16)};
              consume_batch<E, kMinMax, 1>(st, q1, c2);
            }
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&empty_bar[stage]));
      if (++stage == kStages) {
        stage = 0;
        phase ^= 1u;
      }
    }
    if (row_reduce_rot<kMinMax, kCW>(st, &red, it, warp, lane, c2)) row_epilogue(p, row, st, acc);
  }
  if (p.mode == kModePartial && p.push_world > 0) {
    push_finish<kConsumers>(p, static_cast<unsigned int>(it), tid, reinterpret_cast<int*>(acc_sm));
    return;
  }
  if (p.mode == kModeFusedActor) {
    // every warp's lane 0 holds the loss terms of the rows it finished: add them up in warp order
    if (lane == 0) acc_sm[warp] = acc;
    named_bar_sync(1 + kRedSlots, kConsumers);
    if (tid == 0) {
      LossAcc tot = {0.f, 0.f};
      for (int w = 0; w < kConsumerWarps; ++w) {
        tot.sum += acc_sm[w].sum;
        tot.count += acc_sm[w].count;
      }
      loss_finalize(p, tot);
    }
  }
}

// ---- generic forward (any alignment, bf16 / fp16 / fp32): one CTA per row, LDG -----------------
template <typename T>
__device__ __forceinline__ float to_f32(T v);
template <>
__device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <>
__device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }
template <>
__device__ __forceinline__ float to_f32<float>(float v) { return v; }

template <typename T>
__global__ void __launch_bounds__(kConsumers)
logits_fwd_generic_kernel(const LogitsFwdParams p) {
  __shared__ RowReduceSmem red;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const float c2 = p.inv_temperature * kLog2e;
  int parity = 0;
  LossAcc acc = {0.f, 0.f};
  for (int64_t row = blockIdx.x; row < p.T; row += gridDim.x) {
    const T* x = reinterpret_cast<const T*>(reinterpret_cast<const char*>(p.logits) +
                                            row * p.row_stride_bytes);
    RowState st;
    state_init(st);
    // 4 independent elements per iteration, one rebase per group
    int64_t i = tid;
    for (; i + 3 * kConsumers < p.V; i += 4 * kConsumers) {
      const float a = to_f32<T>(x[i]), b = to_f32<T>(x[i + kConsumers]);
      const float c = to_f32<T>(x[i + 2 * kConsumers]), d = to_f32<T>(x[i + 3 * kConsumers]);
      st.mn = fminf(st.mn, fminf(fminf(a, b), fminf(c, d)));
      const float nb = state_rebase(st, fmaxf(fmaxf(a, b), fmaxf(c, d)), c2);
      const float ta = fmaf(a, c2, nb), tb = fmaf(b, c2, nb), tc = fmaf(c, c2, nb), td = fmaf(d, c2, nb);
      const float ea = ex2_approx(ta), eb = ex2_approx(tb), ec = ex2_approx(tc), ed = ex2_approx(td);
      st.s += (ea + eb) + (ec + ed);
      st.w += fmaf(ea, ta, eb * tb) + fmaf(ec, tc, ed * td);
    }
    for (; i < p.V; i += kConsumers) {
      const float a = to_f32<T>(x[i]);
      st.mn = fminf(st.mn, a);
      const float nb = state_rebase(st, a, c2);
      const float ta = fmaf(a, c2, nb);
      const float ea = ex2_approx(ta);
      st.s += ea;
      st.w = fmaf(ea, ta, st.w);
    }
    if (row_reduce(st, &red, parity, warp, lane, c2)) row_epilogue(p, row, st, acc);
    parity ^= 1;
  }
  if (p.mode == kModeFusedActor && tid == 0) loss_finalize(p, acc);
}

// ---- backward ----------------------------------------------------------------------------------
// dlogits[t,v] = p_v * (k0 + k1 * x_v) (+ gl at v == label),  p_v = 2^((x - m)*c2 - log2 S):
// the max is subtracted first (exact in fp32), so 1 - p_label keeps its accuracy on peaked rows.
//   gl = g_logp*scale/temp, gh = g_entropy*scale/temp, lse = m/temp + ln S,
//   k0 = -gl - gh*(H - lse),  k1 = -gh/temp
__device__ __forceinline__ RowGrad row_grad(const LogitsBwdParams& p, int64_t row) {
  RowGrad g;
  float scale = p.inv_temperature;
  if (p.scale0) scale *= p.scale0[0];
  if (p.scale1) scale *= p.scale1[0];
  const float glp = p.g_logp ? p.g_logp[row] : 0.f;
  const float ghp = p.g_entropy ? p.g_entropy[row] : 0.f;
  g.active = (glp != 0.f) || (ghp != 0.f);
  g.m = p.row_max[row];
  const float ln_s = p.ln_s[row];
  g.gl = glp * scale;
  const float gh = ghp * scale;
  const float H = p.g_entropy ? p.entropy[row] : 0.f;
  // k0 + k1*x = -gl - gh*(H - ln S) - (gh/temp)*(x - m)  -> fold m into k0
  g.k1 = -gh * p.inv_temperature;
  g.k0 = -g.gl - gh * (H - ln_s) - g.k1 * g.m;
  g.nl2s = -ln_s * kLog2e;
  return g;
}

template <typename E>
__global__ void __launch_bounds__(kThreads, kCtasPerSm)
logits_bwd_tma_kernel(const LogitsBwdParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[kStages];
  __shared__ __align__(8) uint64_t empty_bar[kStages];

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const uint32_t ring = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int64_t row_bytes = p.V * E::kSize;
  const int nchunk = static_cast<int>((row_bytes + kStageBytes - 1) / kStageBytes);
  const uint32_t tail_bytes = static_cast<uint32_t>(row_bytes - int64_t(nchunk - 1) * kStageBytes);

  if (tid == 0) {
    for (int i = 0; i < kStages; ++i) {
      mbar_init(smem_u32(&full_bar[i]), 1);
      mbar_init(smem_u32(&empty_bar[i]), kConsumerWarps);
    }
    mbar_fence_init();
  }
  __syncthreads();

  if (warp == kConsumerWarps) {
    if (lane == 0) {
      const uint64_t pol = policy_evict_first();
      int stage = 0;
      uint32_t phase = 0;
      for (int64_t row = blockIdx.x; row < p.T; row += gridDim.x) {
        const float glp = p.g_logp ? p.g_logp[row] : 0.f;
        const float ghp = p.g_entropy ? p.g_entropy[row] : 0.f;
        if (glp == 0.f && ghp == 0.f) continue;        // zero row: nothing to read
        const char* src = reinterpret_cast<const char*>(p.logits) + row * p.row_stride_bytes;
        for (int c = 0; c < nchunk; ++c) {
          mbar_wait(smem_u32(&empty_bar[stage]), phase ^ 1u);
          const uint32_t bytes = (c == nchunk - 1) ? tail_bytes : uint32_t(kStageBytes);
          const uint32_t fb = smem_u32(&full_bar[stage]);
          mbar_arrive_expect_tx(fb, bytes);
          tma_load_1d(ring + stage * kStageBytes, src + int64_t(c) * kStageBytes, bytes, fb, pol);
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
    return;
  }

  const float c2 = p.inv_temperature * kLog2e;
  int stage = 0;
  uint32_t phase = 0;
  for (int64_t row = blockIdx.x; row < p.T; row += gridDim.x) {
    const RowGrad g = row_grad(p, row);
    char* dst = reinterpret_cast<char*>(p.dlogits) + row * p.d_row_stride_bytes;
    if (!g.active) {
      const uint4 z = make_uint4(0u, 0u, 0u, 0u);
      for (int64_t off = int64_t(tid) * 16; off < row_bytes; off += int64_t(kConsumers) * 16)
        stg_stream_v4(dst + off, z);
      continue;
    }
    const int64_t lab = p.labels[row] - p.vocab_start;
    for (int c = 0; c < nchunk; ++c) {
      mbar_wait(smem_u32(&full_bar[stage]), phase);
      const uint32_t sbase = ring + stage * kStageBytes;
      const int nvec = ((c == nchunk - 1) ? tail_bytes : uint32_t(kStageBytes)) >> 4;
      uint4 q[kVecPerThread];
#pragma unroll
      for (int j = 0; j < kVecPerThread; ++j) {
        const int v = tid + j * kConsumers;
        if (v < nvec) q[j] = lds_v4(sbase + v * 16);
      }
      // The stage is handed back only after the loaded values have been USED: SYNCS.ARRIVE does not
      // wait on the scoreboards of earlier LDS (ptxas emits no wait mask for it), so an arrive placed
      // right after the loads can overtake their last wavefronts and let the producer's next bulk copy
      // overwrite 128-byte pieces that have not been read yet (seen as rare wrong 64-column runs).
#pragma unroll
      for (int j = 0; j < kVecPerThread; ++j) {
        const int v = tid + j * kConsumers;
        if (v < nvec) {
          const int64_t e0 = int64_t(c) * (kStageBytes / E::kSize) + int64_t(v) * E::kN;
          q[j] = grad_vec_t<E>(q[j], g, c2, lab - e0);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&empty_bar[stage]));
#pragma unroll
      for (int j = 0; j < kVecPerThread; ++j) {
        const int v = tid + j * kConsumers;
        if (v < nvec) {
          const int64_t e0 = int64_t(c) * (kStageBytes / E::kSize) + int64_t(v) * E::kN;
          stg_stream_v4(dst + e0 * E::kSize, q[j]);
        }
      }
      if (++stage == kStages) {
        stage = 0;
        phase ^= 1u;
      }
    }
  }
}

template <typename T>
__device__ __forceinline__ T from_f32(float v);
template <>
__device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <>
__device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }
template <>
__device__ __forceinline__ float from_f32<float>(float v) { return v; }

template <typename T>
__global__ void __launch_bounds__(kConsumers)
logits_bwd_generic_kernel(const LogitsBwdParams p) {
  const int tid = threadIdx.x;
  const float c2 = p.inv_temperature * kLog2e;
  for (int64_t row = blockIdx.x; row < p.T; row += gridDim.x) {
    const RowGrad g = row_grad(p, row);
    const T* x = reinterpret_cast<const T*>(reinterpret_cast<const char*>(p.logits) +
                                            row * p.row_stride_bytes);
    T* d = reinterpret_cast<T*>(reinterpret_cast<char*>(p.dlogits) + row * p.d_row_stride_bytes);
    const int64_t lab = p.labels[row] - p.vocab_start;
    if (!g.active) {
      for (int64_t i = tid; i < p.V; i += kConsumers) d[i] = from_f32<T>(0.f);
      continue;
    }
    for (int64_t i = tid; i < p.V; i += kConsumers) {
      const float xv = to_f32<T>(x[i]);
      const float pv = ex2_approx(fmaf(xv - g.m, c2, g.nl2s));
      float dv = pv * fmaf(xv, g.k1, g.k0);
      if (i == lab) dv += g.gl;
      d[i] = from_f32<T>(dv);
    }
  }
}

// ---- legacy entropy: -sum p * log(p + 1e-7)  (realhf calc_entropy) ----------------------------------
// Second, read-only pass given the row statistics of the forward; one CTA per row.
template <typename T>
__global__ void __launch_bounds__(kConsumers)
legacy_entropy_kernel(const void* __restrict__ logits, int64_t T_rows, int64_t V, int64_t row_stride_bytes,
                      const float* __restrict__ row_max, const float* __restrict__ ln_s, float eps,
                      float* __restrict__ out) {
  __shared__ float red[kConsumerWarps];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int64_t row = blockIdx.x; row < T_rows; row += gridDim.x) {
    const T* x = reinterpret_cast<const T*>(reinterpret_cast<const char*>(logits) + row * row_stride_bytes);
    const float m = row_max[row];
    const float nl2s = -ln_s[row] * kLog2e;
    float acc = 0.f;
    for (int64_t i = tid; i < V; i += kConsumers) {
      const float pv = ex2_approx(fmaf(to_f32<T>(x[i]) - m, kLog2e, nl2s));
      acc = fmaf(pv, lg2_approx(pv + eps), acc);
    }
    acc = warp_sum(acc);
    if (lane == 0) red[warp] = acc;
    __syncthreads();
    if (tid == 0) {
      float t = 0.f;
      for (int w = 0; w < kConsumerWarps; ++w) t += red[w];
      out[row] = -t * kLn2;
    }
    __syncthreads();
  }
}

// ---- cross-shard combine (vocab-sharded path) ----------------------------------------------------
__global__ void vocab_combine_kernel(const float* __restrict__ parts, int R, int64_t T,
                                     float inv_t, float* __restrict__ logp,
                                     float* __restrict__ entropy, float* __restrict__ row_max,
                                     float* __restrict__ ln_s_out) {
  // parts layout: [R][4][T] = (raw max, S, W, raw label logit or 0) per shard
  const int64_t t = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
  if (t >= T) return;
  const float c2 = inv_t * kLog2e;
  float M = -INFINITY;
  for (int r = 0; r < R; ++r) M = fmaxf(M, parts[(int64_t(r) * 4 + 0) * T + t]);
  float S = 0.f, W = 0.f, XL = 0.f;
  for (int r = 0; r < R; ++r) {
    const float m = parts[(int64_t(r) * 4 + 0) * T + t];
    if (m != -INFINITY) {
      const float d = (m - M) * c2;
      const float f = exp2f(d);
      const float sr = parts[(int64_t(r) * 4 + 1) * T + t];
      S = fmaf(sr, f, S);
      W = fmaf(f, fmaf(d, sr, parts[(int64_t(r) * 4 + 2) * T + t]), W);
    }
    XL += parts[(int64_t(r) * 4 + 3) * T + t];
  }
  const float ln_s = logf(S);
  logp[t] = (XL - M) * inv_t - ln_s;
  if (entropy) entropy[t] = ln_s - kLn2 * (W / S);
  row_max[t] = M;
  ln_s_out[t] = ln_s;
}

// ---- cross-shard combine over NVLink peer memory (no all-gather) --------------------------------------
// Every rank owns a symmetric buffer: uint32 flag[kMaxPeers] (256 B) | parts[2][4][Tmax] f32.  The partial
// kernel of step e has written this rank's parts into slot e & 1 (same stream, already complete).  This
// kernel (a) tells every peer "my parts of step e are there" with one release store per peer into THAT
// peer's flag[my rank], (b) waits until all peers have said so here, (c) reads the R partial sets straight
// out of the peers' memory and combines them in rank order -- bit-identical to vocab_combine_kernel.
// Reuse of a slot two steps later needs no second handshake: a rank can only pass (b) of step e+1 after every
// peer has launched its partial kernel of step e+1, i.e. after that peer's combine of step e has finished.
constexpr int kMaxPeers = 16;
struct PeerTable {
  char* base[kMaxPeers];
};
constexpr int64_t kPeerHeaderBytes = 256;

__device__ __forceinline__ unsigned int ld_acquire_sys_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys_u32(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// peer memory is cached in the local L1 only: read around it, or a slot's previous contents come back
__device__ __forceinline__ float ld_peer_f32(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}

__global__ void __launch_bounds__(256)
vocab_combine_p2p_kernel(const PeerTable peers, int rank, int R, int64_t T, int64_t Tmax, int slot,
                         unsigned int epoch, float inv_t, long long timeout_cycles, float* __restrict__ logp,
                         float* __restrict__ entropy, float* __restrict__ row_max, float* __restrict__ ln_s_out,
                         int* __restrict__ status) {
  __shared__ int ok_sm;
  if (blockIdx.x == 0 && threadIdx.x < R) {
    __threadfence_system();
    st_release_sys_u32(reinterpret_cast<unsigned int*>(peers.base[threadIdx.x]) + rank, epoch);
  }
  if (threadIdx.x == 0) {
    const unsigned int* mine = reinterpret_cast<const unsigned int*>(peers.base[rank]);
    const long long t0 = clock64();
    int ok = 1;
    for (int r = 0; r < R && ok; ++r) {
      // epochs only grow; the signed difference survives wrap-around
      while (static_cast<int>(ld_acquire_sys_u32(mine + r) - epoch) < 0) {
        if (clock64() - t0 > timeout_cycles) {
          ok = 0;
          break;
        }
        __nanosleep(100);
      }
    }
    if (!ok) atomicExch(status, 1);
    ok_sm = ok;
  }
  __syncthreads();
  if (!ok_sm) return;                       // a peer never arrived: outputs untouched, status says so
  const int64_t t = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
  if (t >= T) return;
  const float c2 = inv_t * kLog2e;
  float m_r[kMaxPeers], s_r[kMaxPeers], w_r[kMaxPeers];
  float XL = 0.f, M = -INFINITY;
#pragma unroll
  for (int r = 0; r < kMaxPeers; ++r) {
    if (r < R) {
      const float* q = reinterpret_cast<const float*>(peers.base[r] + kPeerHeaderBytes) + int64_t(slot) * 4 * Tmax + t;
      m_r[r] = ld_peer_f32(q);
      s_r[r] = ld_peer_f32(q + T);
      w_r[r] = ld_peer_f32(q + 2 * T);
      XL += ld_peer_f32(q + 3 * T);
      M = fmaxf(M, m_r[r]);
    }
  }
  float S = 0.f, W = 0.f;
#pragma unroll
  for (int r = 0; r < kMaxPeers; ++r) {
    if (r < R && m_r[r] != -INFINITY) {
      const float d = (m_r[r] - M) * c2;
      const float f = exp2f(d);
      S = fmaf(s_r[r], f, S);
      W = fmaf(f, fmaf(d, s_r[r], w_r[r]), W);
    }
  }
  const float ln_s = logf(S);
  logp[t] = (XL - M) * inv_t - ln_s;
  if (entropy) entropy[t] = ln_s - kLn2 * (W / S);
  row_max[t] = M;
  ln_s_out[t] = ln_s;
}

bool tma_ok(const void* ptr, int dtype, int64_t V, int64_t row_stride_bytes) {
  const int64_t es = dtype == kFP32 ? 4 : 2;
  return (reinterpret_cast<uintptr_t>(ptr) & 15) == 0 && ((V * es) % 16) == 0 && (row_stride_bytes % 16) == 0;
}

template <typename K>
int set_smem(K kern) {
  AB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kSmemBytes)));
  return kOk;
}

// cudaFuncSetAttribute is per device: one "already set" flag per device ordinal
constexpr int kMaxDevices = 64;
int current_device_slot() {
  int d = 0;
  cudaGetDevice(&d);
  return d & (kMaxDevices - 1);
}

int grid_for(int64_t T, int per_sm) {
  const int64_t g = int64_t(num_sms()) * per_sm;
  return static_cast<int>(T < g ? T : g);
}

}  // namespace

int64_t logits_loss_workspace_bytes() { return 16 + 8 * int64_t(num_sms()) * 8; }

int launch_logits_fwd(const LogitsFwdParams& p, cudaStream_t stream) {
  if (p.T == 0) return kOk;
  const bool minmax = p.raw_min != nullptr;
  const int dev = current_device_slot();
  if (tma_ok(p.logits, p.dtype, p.V, p.row_stride_bytes)) {
    // rows of at most 64 KB (vocabulary shards): half-width CTAs, twice as many of them
    const bool narrow = p.V * (p.dtype == kFP32 ? 4 : 2) <= 65536;
    // consumer warps of the narrow instance: 4 (8 KB stages), or 2 (4 KB stages, 8 CTAs/SM: a row's ragged last
    // stage wastes half as much and a thread sees twice as many elements between two row ends)
    static const int narrow_w = [] {
      const char* e = getenv("AREAL_B200_FWD_NARROW_W");
      return (e && e[0] == '2') ? 2 : 4;
    }();
#define AB_LAUNCH_FWD_W(MM, E, W)                                                                  \
  {                                                                                                \
    static bool done[kMaxDevices] = {};            /* the attribute is per device */              \
    constexpr size_t smem = size_t(kStages) * (W * 2048) + 1024;                                   \
    if (!done[dev]) {                                                                              \
      AB_CUDA(cudaFuncSetAttribute(logits_fwd_tma_kernel<MM, E, W>,                                \
                                   cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));       \
      done[dev] = true;                                                                            \
    }                                                                                              \
    int per_sm = 16 / W;                                                                           \
    if (p.push_world > 0) {                        /* its CTAs wait for each other: all co-resident */ \
      int occ = 0;                                                                                 \
      AB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, logits_fwd_tma_kernel<MM, E, W>, \
                                                            (W + 1) * 32, smem));                  \
      if (occ < 1) {                                                                               \
        set_error("logits_partial_push: the kernel does not fit an SM");                           \
        return kCudaError;                                                                         \
      }                                                                                            \
      if (occ < per_sm) per_sm = occ;                                                              \
    }                                                                                              \
    logits_fwd_tma_kernel<MM, E, W><<<grid_for(p.T, per_sm), (W + 1) * 32, smem, stream>>>(p);     \
  }
#define AB_LAUNCH_FWD(MM, E)                                                       \
  {                                                                                \
    if (narrow && narrow_w == 2) AB_LAUNCH_FWD_W(MM, E, 2)                         \
    else if (narrow) AB_LAUNCH_FWD_W(MM, E, 4)                                     \
    else AB_LAUNCH_FWD_W(MM, E, 8)                                                 \
  }
    if (p.dtype == kBF16) { if (minmax) AB_LAUNCH_FWD(true, ElemBF16) else AB_LAUNCH_FWD(false, ElemBF16) }
    else if (p.dtype == kFP16) { if (minmax) AB_LAUNCH_FWD(true, ElemFP16) else AB_LAUNCH_FWD(false, ElemFP16) }
    else { if (minmax) AB_LAUNCH_FWD(true, ElemFP32) else AB_LAUNCH_FWD(false, ElemFP32) }
#undef AB_LAUNCH_FWD_W
#undef AB_LAUNCH_FWD
  } else {
    const int grid = grid_for(p.T, 8);
    if (p.dtype == kBF16) logits_fwd_generic_kernel<__nv_bfloat16><<<grid, kConsumers, 0, stream>>>(p);
    else if (p.dtype == kFP16) logits_fwd_generic_kernel<__half><<<grid, kConsumers, 0, stream>>>(p);
    else logits_fwd_generic_kernel<float><<<grid, kConsumers, 0, stream>>>(p);
  }
  return check_launch("logits_fwd");
}

int64_t vocab_push_buffer_bytes(int64_t max_tokens, int world) {
  return kPushHeaderBytes + 2 * int64_t(world) * max_tokens * int64_t(sizeof(float4));
}

int launch_logits_partial_push(const LogitsFwdParams& p, cudaStream_t stream) {
  if (p.T == 0) return kOk;                  // (every rank has the same T: nobody waits for rows of an empty step)
  if (!tma_ok(p.logits, p.dtype, p.V, p.row_stride_bytes)) {
    set_error("logits_partial_push: rows must be 16-byte aligned (base, row width and row stride)");
    return kInvalidArgument;
  }
  return launch_logits_fwd(p, stream);
}

int launch_logits_bwd(const LogitsBwdParams& p, cudaStream_t stream) {
  if (p.T == 0) return kOk;
  const int dev = current_device_slot();
  if (tma_ok(p.logits, p.dtype, p.V, p.row_stride_bytes) &&
      tma_ok(p.dlogits, p.dtype, p.V, p.d_row_stride_bytes)) {
    const int grid = grid_for(p.T, kCtasPerSm);
#define AB_LAUNCH_BWD(E)                                                           \
  {                                                                                \
    static bool done[kMaxDevices] = {};                                            \
    if (!done[dev]) {                                                              \
      const int rc = set_smem(logits_bwd_tma_kernel<E>);                           \
      if (rc) return rc;                                                           \
      done[dev] = true;                                                            \
    }                                                                              \
    logits_bwd_tma_kernel<E><<<grid, kThreads, kSmemBytes, stream>>>(p);           \
  }
    if (p.dtype == kBF16) AB_LAUNCH_BWD(ElemBF16)
    else if (p.dtype == kFP16) AB_LAUNCH_BWD(ElemFP16)
    else AB_LAUNCH_BWD(ElemFP32)
#undef AB_LAUNCH_BWD
  } else {
    const int grid = grid_for(p.T, 8);
    if (p.dtype == kBF16) logits_bwd_generic_kernel<__nv_bfloat16><<<grid, kConsumers, 0, stream>>>(p);
    else if (p.dtype == kFP16) logits_bwd_generic_kernel<__half><<<grid, kConsumers, 0, stream>>>(p);
    else logits_bwd_generic_kernel<float><<<grid, kConsumers, 0, stream>>>(p);
  }
  return check_launch("logits_bwd");
}

int launch_legacy_entropy(const void* logits, int dtype, int64_t T, int64_t V, int64_t row_stride_bytes,
                          const float* row_max, const float* ln_s, float eps, float* out,
                          cudaStream_t stream) {
  if (T == 0) return kOk;
  const int grid = grid_for(T, 8);
  if (dtype == kBF16)
    legacy_entropy_kernel<__nv_bfloat16><<<grid, kConsumers, 0, stream>>>(logits, T, V, row_stride_bytes, row_max, ln_s, eps, out);
  else if (dtype == kFP16)
    legacy_entropy_kernel<__half><<<grid, kConsumers, 0, stream>>>(logits, T, V, row_stride_bytes, row_max, ln_s, eps, out);
  else
    legacy_entropy_kernel<float><<<grid, kConsumers, 0, stream>>>(logits, T, V, row_stride_bytes, row_max, ln_s, eps, out);
  return check_launch("legacy_entropy");
}

int launch_vocab_combine(const float* parts, int R, int64_t T, float inv_t, float* logp,
                         float* entropy, float* row_max, float* ln_s, cudaStream_t stream) {
  if (T == 0) return kOk;
  const int threads = 256;
  vocab_combine_kernel<<<static_cast<int>((T + threads - 1) / threads), threads, 0, stream>>>(
      parts, R, T, inv_t, logp, entropy, row_max, ln_s);
  return check_launch("vocab_combine");
}

int launch_vocab_combine_p2p(void* const* peer_bases, int rank, int R, int64_t T, int64_t Tmax, int slot,
                             unsigned int epoch, float inv_t, long long timeout_cycles, float* logp, float* entropy,
                             float* row_max, float* ln_s, int* status, cudaStream_t stream) {
  PeerTable tab;
  for (int r = 0; r < kMaxPeers; ++r) tab.base[r] = r < R ? static_cast<char*>(peer_bases[r]) : nullptr;
  const int threads = 256;
  const int64_t blocks = T > 0 ? (T + threads - 1) / threads : 1;   // T == 0 still takes part in the handshake
  vocab_combine_p2p_kernel<<<static_cast<unsigned int>(blocks), threads, 0, stream>>>(
      tab, rank, R, T, Tmax, slot, epoch, inv_t, timeout_cycles, logp, entropy, row_max, ln_s, status);
  return check_launch("vocab_combine_p2p");
}

}  // namespace ab
