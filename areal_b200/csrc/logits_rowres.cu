// Single-pass fused actor step with ROW-RESIDENT logits: log-prob, entropy, min/max, the decoupled PPO
// actor loss, its logging vectors AND dlogits from one read of the logits (4*T*V bytes of HBM traffic
// instead of 6*T*V for forward + backward).
//
// Same reference functions as logits_fused.cu (areal/utils/functional.py:20-127, 171-235;
// areal/engine/ppo/actor.py:313-391, 375-376), plus the autograd backward they imply; usable when the
// scalar that multiplies the loss before .backward() (the engine's loss_scale,
// areal/engine/fsdp_engine.py:578-586) is known when the forward runs.
//
// Design (B200): a V=151,936 bf16 row is 304 KB -- more than one SM's shared memory -- so a row is
// held by a thread-block CLUSTER: C CTAs (3 for this V: two 101 KB slices per SM) each TMA-load 1/C of
// the row into their shared memory, reduce it with the online-softmax state of logits_device.cuh,
// exchange the five partial statistics through distributed shared memory (st.shared::cluster +
// remote mbarrier arrives, no cluster-wide barrier in the steady state), and then re-read their
// slice FROM SHARED MEMORY to write the gradient.  Rows are handed to clusters round-robin; inside a
// CTA the work is warp-specialised (producer / statistics group / gradient group), so the gradient
// pass of row r overlaps the statistics pass of row r+1 and the TMA load of row r+2.  The masked-token count that normalises the loss is recomputed by every CTA from the
// mask (T bytes) at kernel start, so the gradient needs no second kernel.
#include "common.cuh"
#include "logits_device.cuh"
#include "logits_fused.h"
#include "ppo_math.cuh"

namespace ab {
namespace {

using namespace dev;

constexpr int kGroupWarps = 8;                       // warps per compute group
constexpr int kGroupThreads = kGroupWarps * 32;      // 256
constexpr int kRRThreads = 2 * kGroupThreads + 64;   // group A (stats) + group B (gradient) + producer + exchange warps
constexpr int kMaxCluster = 8;
constexpr int kTmaPiece = 16384;
constexpr int kMaxBuf = 6;

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_arrive() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
}
__device__ __forceinline__ void cluster_wait() {
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t map_to_rank(const void* local_ptr, uint32_t target_rank) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(local_ptr)), "r"(target_rank));
  return remote;
}
// 4-byte store into another CTA's shared memory that also counts 4 bytes on that CTA's mbarrier
// (no release fence needed: the receiver sees the data once its barrier phase completes)
__device__ __forceinline__ void st_async_f32(uint32_t remote_addr, float v, uint32_t remote_bar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(remote_addr),
               "r"(__float_as_uint(v)), "r"(remote_bar)
               : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAITC_%=:\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONEC_%=;\n\t"
      "bra WAITC_%=;\n\t"
      "DONEC_%=:\n\t"
      "}" ::"r"(bar),
      "r"(parity)
      : "memory");
}

// statistics of kN vectors (kGroupThreads apart) with ONE rebase: packed bf16x2 max / min first
template <int kN>
__device__ __forceinline__ void stats_step(RowState& st, uint32_t sbase, int v, float c2) {
  uint4 q[kN];
#pragma unroll
  for (int j = 0; j < kN; ++j) q[j] = lds_v4(sbase + (v + j * kGroupThreads) * 16);
  uint32_t mx = vec_max_bf16(q[0]), mn = vec_min_bf16(q[0]);
#pragma unroll
  for (int j = 1; j < kN; ++j) {
    mx = max_bf16x2(mx, vec_max_bf16(q[j]));
    mn = min_bf16x2(mn, vec_min_bf16(q[j]));
  }
  st.mn = fminf(st.mn, fminf(bf16_lo(mn), bf16_hi(mn)));
  const float nb = state_rebase(st, fmaxf(bf16_lo(mx), bf16_hi(mx)), c2);
  float s0 = 0.f, s1 = 0.f, w0 = 0.f, w1 = 0.f;
#pragma unroll
  for (int j = 0; j < kN; ++j) vec_accumulate(q[j], c2, nb, s0, s1, w0, w1);
  st.s += s0 + s1;
  st.w += w0 + w1;
}

// lab32: the label's element index inside this CTA's slice, or negative.  The label's element is the
// only one that gets + gl: the thread that stored its vector overwrites those 2 bytes afterwards (same
// thread, same address: program order), so the vector loop carries no label test.  That element is
// gl * (1 - p), a cancellation when the row is peaked (p -> 1): p comes from (x - m) * c2 - log2 S there,
// exact in its leading term, not from the folded exponent (whose 1e-6 relative error would be the result).
template <bool kNeg>
__device__ __forceinline__ void grad_slice(uint32_t sbase, char* dst, int my_nvec, int gt, float c2, float nbias,
                                           float gl, float row_m, float nl2s, int lab32) {
  int v = gt;
  for (; v + kGroupThreads < my_nvec; v += 2 * kGroupThreads) {
    const uint4 q0 = lds_v4(sbase + v * 16);
    const uint4 q1 = lds_v4(sbase + (v + kGroupThreads) * 16);
    stg_stream_v4(dst + int64_t(v) * 16, grad_vec_exp<kNeg>(q0, c2, nbias));
    stg_stream_v4(dst + int64_t(v + kGroupThreads) * 16, grad_vec_exp<kNeg>(q1, c2, nbias));
  }
  if (v < my_nvec) stg_stream_v4(dst + int64_t(v) * 16, grad_vec_exp<kNeg>(lds_v4(sbase + v * 16), c2, nbias));
  if (lab32 >= 0 && ((lab32 >> 3) & (kGroupThreads - 1)) == gt) {
    const float x = __uint_as_float(lds_u16(sbase + uint32_t(lab32) * 2u) << 16);
    const float pv = ex2_approx(fmaf(x - row_m, c2, nl2s));
    const float dd = fmaf(-gl, pv, gl);
    stg_u16(dst + int64_t(lab32) * 2, static_cast<uint16_t>(pack_bf16x2(dd, dd) & 0xffffu));
  }
}

struct RRShared {
  float m[kMaxBuf][kGroupWarps], s[kMaxBuf][kGroupWarps], w[kMaxBuf][kGroupWarps], mn[kMaxBuf][kGroupWarps];   // [slice buffer]
  float xchg[2][kMaxCluster][8];    // [row parity][source CTA rank][m, S, W, xl, mn]
  float rowinfo[kMaxBuf][4];          // per slice buffer, group A -> group B: m, -log2 S, gl
  float count;
};

// Roles inside a CTA (warp-specialised, no CTA-wide barrier in the steady state):
//   producer (warp 16, one lane): keeps both slice buffers filled with TMA bulk loads
//   group A  (warps 0-7): statistics pass over the slice -> 8 warp partials
//   exchange (warp 17): merges them, trades partials with the other CTAs through DSMEM, token epilogue
//   group B  (warps 8-15): gradient pass over the same slice, one row behind group A
__global__ void __launch_bounds__(kRRThreads, 1)
fused_rowres_kernel(const LogitsFwdParams p, void* __restrict__ dlogits, int64_t d_row_stride_bytes,
                    const float* __restrict__ loss_scale, uint32_t buf_stride, int nbuf) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[kMaxBuf];    // slice landed (TMA tx bytes)
  __shared__ __align__(8) uint64_t ready_bar[kMaxBuf];   // row statistics published by group A
  __shared__ __align__(8) uint64_t empty_bar[kMaxBuf];   // group B is done with the slice
  __shared__ __align__(8) uint64_t xchg_bar[2];        // all C partials of a row have arrived
  __shared__ __align__(8) uint64_t part_bar[kMaxBuf];        // group A's 8 warp partials of a row are in shared memory
  __shared__ RRShared sh;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t C = cluster_nctarank(), rank = cluster_ctarank();
  const int64_t cid = blockIdx.x / C, ncl = gridDim.x / C;
  const uint32_t ring = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* ring_ptr = smem_raw + (ring - smem_u32(smem_raw));

  const int64_t nvec = p.V >> 3;
  const int64_t nvq = (nvec + C - 1) / C;
  const int64_t v0 = int64_t(rank) * nvq;
  const int my_nvec = static_cast<int>(nvec - v0 < 0 ? 0 : (nvec - v0 < nvq ? nvec - v0 : nvq));
  const uint32_t my_bytes = uint32_t(my_nvec) * 16u;
  const float c2 = p.inv_temperature * kLog2e;

  // ---- masked-token count (every CTA, fixed order -> identical value everywhere) -------------------
  {
    float c = 0.f;
    for (int64_t t = tid; t < p.T; t += kRRThreads) c += p.loss_mask[t] ? 1.f : 0.f;
    c = warp_sum(c);
    float* tmp = &sh.xchg[0][0][0];          // 17 warps <= 128 floats of scratch
    if (lane == 0) tmp[warp] = c;
    __syncthreads();
    if (tid == 0) {
      float tot = 0.f;
      for (int w = 0; w < kRRThreads / 32; ++w) tot += tmp[w];
      sh.count = tot;
      for (int i = 0; i < kMaxBuf; ++i) {
        mbar_init(smem_u32(&full_bar[i]), 1);
        mbar_init(smem_u32(&ready_bar[i]), 1);
        mbar_init(smem_u32(&empty_bar[i]), kGroupWarps);
        mbar_init(smem_u32(&part_bar[i]), kGroupWarps);
      }
      for (int i = 0; i < 2; ++i) mbar_init(smem_u32(&xchg_bar[i]), 1);
      mbar_fence_init();
    }
    __syncthreads();
  }
  const float count = sh.count;
  const float inv_count = 1.0f / (count > 0.f ? count : 1.0f);           // count_nonzero() or 1
  const float gscale = p.inv_temperature * inv_count * (loss_scale ? loss_scale[0] : 1.0f);
  // every CTA of the cluster must have its barriers initialised before a peer arrives on them
  cluster_arrive();
  cluster_wait();

  if (warp == 2 * kGroupWarps) {
    // ===== producer ===========================================================================
    if (lane == 0 && my_bytes != 0) {
      const uint64_t pol = policy_evict_first();
      int64_t it = 0;
      for (int64_t row = cid; row < p.T; row += ncl, ++it) {
        const int b = static_cast<int>(it % nbuf);
        if (it >= nbuf) mbar_wait(smem_u32(&empty_bar[b]), static_cast<uint32_t>(((it / nbuf) - 1) & 1));
        const char* src = reinterpret_cast<const char*>(p.logits) + row * p.row_stride_bytes + v0 * 16;
        const uint32_t bar = smem_u32(&full_bar[b]);
        mbar_arrive_expect_tx(bar, my_bytes);
        for (uint32_t off = 0; off < my_bytes; off += kTmaPiece) {
          const uint32_t n = my_bytes - off < uint32_t(kTmaPiece) ? my_bytes - off : uint32_t(kTmaPiece);
          tma_load_1d(ring + uint32_t(b) * buf_stride + off, src + off, n, bar, pol);
        }
      }
    }
  } else if (warp < kGroupWarps) {
    // ===== group A: statistics pass only; hands 8 warp partials to the exchange warp ================
    int64_t it = 0;
    for (int64_t row = cid; row < p.T; row += ncl, ++it) {
      const int b = static_cast<int>(it % nbuf);
      const uint32_t ph = static_cast<uint32_t>((it / nbuf) & 1);
      const uint32_t sbase = ring + uint32_t(b) * buf_stride;
      if (my_bytes) mbar_wait(smem_u32(&full_bar[b]), ph);
      RowState st;
      state_init(st);
      {
        int v = tid;
        for (; v + 3 * kGroupThreads < my_nvec; v += 4 * kGroupThreads) stats_step<4>(st, sbase, v, c2);
        if (v + kGroupThreads < my_nvec) {      // fewer than four vectors left for this thread: 2 + 1
          stats_step<2>(st, sbase, v, c2);
          v += 2 * kGroupThreads;
        }
        if (v < my_nvec) stats_step<1>(st, sbase, v, c2);
      }
      state_canonical(st, c2);
      warp_merge_state(st, c2);
      if (lane == 0) {
        sh.m[b][warp] = st.m;
        sh.s[b][warp] = st.s;
        sh.w[b][warp] = st.w;
        sh.mn[b][warp] = st.mn;
        mbar_arrive(smem_u32(&part_bar[b]));   // release: the exchange warp may read my partial
      }
    }
  } else if (warp == 2 * kGroupWarps + 1) {
    // ===== exchange warp: merge, cross-CTA exchange through DSMEM, token epilogue =====================
    float loss_sum = 0.f;   // rank 0, lane 0 only
    int64_t it = 0;
    for (int64_t row = cid; row < p.T; row += ncl, ++it) {
      const int b = static_cast<int>(it % nbuf);
      const int par = static_cast<int>(it & 1);
      const uint32_t ph = static_cast<uint32_t>((it / nbuf) & 1);
      const uint32_t phx = static_cast<uint32_t>((it >> 1) & 1);
      // per-row scalars: issue the global loads now, consume them after the exchange
      const int64_t lab = p.labels[row] - p.vocab_start;
      float t_old = 0.f, t_prox = 0.f, t_adv = 0.f;
      bool t_mask = false;
      if (lane == 0) {
        t_old = p.old[row];
        t_prox = p.prox ? p.prox[row] : t_old;
        t_adv = p.adv[row];
        t_mask = p.loss_mask[row] != 0;
      }
      mbar_wait(smem_u32(&part_bar[b]), ph);
      RowState st;
      {
        const int l = lane & (kGroupWarps - 1);
        st.m = *reinterpret_cast<const volatile float*>(&sh.m[b][l]);
        st.s = *reinterpret_cast<const volatile float*>(&sh.s[b][l]);
        st.w = *reinterpret_cast<const volatile float*>(&sh.w[b][l]);
        st.mn = *reinterpret_cast<const volatile float*>(&sh.mn[b][l]);
#pragma unroll
        for (int o = kGroupWarps / 2; o > 0; o >>= 1) {
          const float m2 = __shfl_xor_sync(0xffffffffu, st.m, o);
          const float s2 = __shfl_xor_sync(0xffffffffu, st.s, o);
          const float w2 = __shfl_xor_sync(0xffffffffu, st.w, o);
          const float n2 = __shfl_xor_sync(0xffffffffu, st.mn, o);
          state_merge(st.m, st.s, st.w, m2, s2, w2, c2);
          st.mn = fminf(st.mn, n2);
        }
      }
      // the label logit, if it lives in my slice (read from shared memory, exact)
      float xl = 0.f;
      const int64_t lv = (lab >> 3) - v0;
      if (lab >= 0 && lab < p.V && lv >= 0 && lv < my_nvec) {
        const uint32_t wd = *reinterpret_cast<const volatile uint16_t*>(ring_ptr + size_t(b) * buf_stride + lv * 16 + (lab & 7) * 2);
        xl = __uint_as_float(wd << 16);
      }
      // lane j < C delivers my five numbers to CTA j, then arrives on CTA j's exchange barrier
      if (lane == 0) mbar_arrive_expect_tx(smem_u32(&xchg_bar[par]), C * 20u);   // 5 floats from each CTA
      if (lane < int(C)) {
        const uint32_t dst = map_to_rank(&sh.xchg[par][rank][0], lane);
        const uint32_t rbar = map_to_rank(&xchg_bar[par], lane);
        st_async_f32(dst + 0, st.m, rbar);
        st_async_f32(dst + 4, st.s, rbar);
        st_async_f32(dst + 8, st.w, rbar);
        st_async_f32(dst + 12, xl, rbar);
        st_async_f32(dst + 16, st.mn, rbar);
      }
      mbar_wait_cluster(smem_u32(&xchg_bar[par]), phx);

      float m = -INFINITY, s = 0.f, w = 0.f, mn = INFINITY;
      xl = 0.f;
      if (lane < int(C)) {
        const volatile float* src = &sh.xchg[par][lane][0];
        m = src[0]; s = src[1]; w = src[2]; xl = src[3]; mn = src[4];
      }
#pragma unroll
      for (int o = kMaxCluster / 2; o > 0; o >>= 1) {
        const float m2 = __shfl_xor_sync(0xffffffffu, m, o);
        const float s2 = __shfl_xor_sync(0xffffffffu, s, o);
        const float w2 = __shfl_xor_sync(0xffffffffu, w, o);
        xl += __shfl_xor_sync(0xffffffffu, xl, o);
        mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
        state_merge(m, s, w, m2, s2, w2, c2);
      }
      if (lane == 0) {
        const bool local = lab >= 0 && lab < p.V;
        const float ln_s = lg2_approx(s) * kLn2;
        const float logp = local ? (xl - m) * p.inv_temperature - ln_s : __int_as_float(0x7fc00000);
        const float ent = ln_s - kLn2 * (w / s);
        const ActorTokenOut o = actor_token(logp, t_prox, t_old, t_adv, t_mask, p.cfg);
        sh.rowinfo[b][0] = m;
        sh.rowinfo[b][1] = -ln_s * kLog2e;
        sh.rowinfo[b][2] = o.coef * gscale;
        mbar_arrive(smem_u32(&ready_bar[b]));          // release: group B may start this row
        if (rank == 0) {
          p.logp[row] = logp;
          p.row_max[row] = m;
          p.ln_s[row] = ln_s;
          if (p.entropy) p.entropy[row] = ent;
          if (p.raw_min) p.raw_min[row] = mn;
          p.coef[row] = o.coef;
          if (p.stat_loss) p.stat_loss[row] = o.pg;
          if (p.stat_ratio) p.stat_ratio[row] = o.ratio;
          if (p.stat_kl) p.stat_kl[row] = o.kl;
          if (p.clip_mask) p.clip_mask[row] = o.clip;
          if (p.dual_clip_mask) p.dual_clip_mask[row] = o.dual;
          if (p.behave_w) p.behave_w[row] = o.w;
          if (p.behave_kl) p.behave_kl[row] = o.bk;
          if (p.behave_mask) p.behave_mask[row] = o.bmask;
          if (t_mask) loss_sum += o.pg;
        }
      }
      __syncwarp();
    }
    // ---- fixed-order loss reduction across clusters (rank-0 CTAs) ---------------------------------------
    if (rank == 0 && lane == 0) {
      float* partials = reinterpret_cast<float*>(reinterpret_cast<char*>(p.workspace) + 16);
      unsigned int* ticket = reinterpret_cast<unsigned int*>(p.workspace);
      partials[cid] = loss_sum;
      __threadfence();
      const unsigned int t = atomicAdd(ticket, 1u);
      if (t == static_cast<unsigned int>(ncl) - 1u) {
        __threadfence();
        double sum = 0.0;
        for (int64_t i = 0; i < ncl; ++i) sum += reinterpret_cast<const volatile float*>(partials)[i];
        p.loss_out[0] = static_cast<float>(sum * double(inv_count));
        p.inv_count_out[0] = inv_count;
        *ticket = 0u;
      }
    }
  } else if (warp < 2 * kGroupWarps) {
    // ===== group B: gradient pass over the resident slice =========================================
    const int gt = tid - kGroupThreads;
    int64_t it = 0;
    for (int64_t row = cid; row < p.T; row += ncl, ++it) {
      const int b = static_cast<int>(it % nbuf);
      const uint32_t ph = static_cast<uint32_t>((it / nbuf) & 1);
      const uint32_t sbase = ring + uint32_t(b) * buf_stride;
      const int64_t lab = p.labels[row] - p.vocab_start;
      mbar_wait(smem_u32(&ready_bar[b]), ph);
      if (my_bytes) mbar_wait(smem_u32(&full_bar[b]), ph);
      const float row_m = *reinterpret_cast<const volatile float*>(&sh.rowinfo[b][0]);
      const float nl2s = *reinterpret_cast<const volatile float*>(&sh.rowinfo[b][1]);   // -log2 S
      const float gl = *reinterpret_cast<const volatile float*>(&sh.rowinfo[b][2]);
      char* dst = reinterpret_cast<char*>(dlogits) + row * d_row_stride_bytes + v0 * 16;
      if (gl != 0.f) {
        // d = gl * (1[label] - p):  -gl * p = -/+ 2^((x - m) * c2 - log2 S + log2 |gl|)
        const float nbias = fmaf(-row_m, c2, nl2s + lg2_approx(fabsf(gl)));
        const int64_t lab_local = lab - v0 * 8;
        const int lab32 = (lab_local >= 0 && lab_local < int64_t(my_nvec) * 8) ? static_cast<int>(lab_local) : -1;
        if (gl > 0.f)
          grad_slice<true>(sbase, dst, my_nvec, gt, c2, nbias, gl, row_m, nl2s, lab32);
        else
          grad_slice<false>(sbase, dst, my_nvec, gt, c2, nbias, gl, row_m, nl2s, lab32);
      } else {
        const uint4 z = make_uint4(0u, 0u, 0u, 0u);
        for (int v = gt; v < my_nvec; v += kGroupThreads) stg_stream_v4(dst + int64_t(v) * 16, z);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(smem_u32(&empty_bar[b]));   // slice and rowinfo may be overwritten
    }
  }
  // nobody leaves while a peer may still write into its shared memory or arrive on its barriers
  __syncwarp();
  cluster_arrive();
  cluster_wait();
}

// dlogits *= (*upstream / *assumed) unless the two scalars are equal (the expected case: nothing to do)
__global__ void rescale_if_needed_kernel(uint4* __restrict__ data, int64_t nvec16, int64_t V8, int64_t stride8,
                                         const float* __restrict__ upstream, const float* __restrict__ assumed) {
  const float a = assumed ? assumed[0] : 1.0f;
  const float u = upstream[0];
  if (u == a) return;
  const float r = u / a;
  for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < nvec16; i += int64_t(gridDim.x) * blockDim.x) {
    const int64_t row = i / V8, col = i - row * V8;
    uint4* ptr = data + row * stride8 + col;
    uint4 q = *ptr;
    q.x = pack_bf16x2(bf16_lo(q.x) * r, bf16_hi(q.x) * r);
    q.y = pack_bf16x2(bf16_lo(q.y) * r, bf16_hi(q.y) * r);
    q.z = pack_bf16x2(bf16_lo(q.z) * r, bf16_hi(q.z) * r);
    q.w = pack_bf16x2(bf16_lo(q.w) * r, bf16_hi(q.w) * r);
    *ptr = q;
  }
}

}  // namespace

int launch_rescale_if_needed(void* dlogits, int64_t T, int64_t V, int64_t row_stride_bytes,
                             const float* upstream, const float* assumed, cudaStream_t stream) {
  if (T == 0) return kOk;
  if ((V % 8) != 0 || (row_stride_bytes % 16) != 0 || (reinterpret_cast<uintptr_t>(dlogits) & 15) != 0) {
    set_error("rescale: bf16 rows must be 16-byte aligned with V %% 8 == 0");
    return kUnsupported;
  }
  rescale_if_needed_kernel<<<num_sms() * 8, 256, 0, stream>>>(reinterpret_cast<uint4*>(dlogits), T * (V / 8), V / 8,
                                                             row_stride_bytes / 16, upstream, assumed);
  return check_launch("rescale_if_needed");
}

// Picks the smallest cluster whose double-buffered slice fits one CTA per SM, sizes the grid with the
// occupancy API, launches.  Returns kUnsupported when the rows cannot take this path.
int launch_fused_rowres(const LogitsFwdParams& p, void* dlogits, int64_t d_row_stride_bytes,
                        const float* loss_scale, int force_cluster, int force_nbuf, cudaStream_t stream) {
  if (p.T == 0) return kOk;
  const bool aligned = p.dtype == kBF16 && (reinterpret_cast<uintptr_t>(p.logits) & 15) == 0 &&
                       (p.V % 8) == 0 && (p.row_stride_bytes % 16) == 0 &&
                       (reinterpret_cast<uintptr_t>(dlogits) & 15) == 0 && (d_row_stride_bytes % 16) == 0;
  if (!aligned) {
    set_error("row-resident path needs bf16 logits / dlogits with 16-byte aligned rows and V %% 8 == 0");
    return kUnsupported;
  }
  const int64_t nvec = p.V / 8;
  // dynamic shared memory one CTA can really ask for: the device's opt-in limit minus the kernel's static
  // shared memory minus the 1 KB of alignment slack added to the request below
  int dev = 0, optin = 0;
  AB_CUDA(cudaGetDevice(&dev));
  AB_CUDA(cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  cudaFuncAttributes fa;
  AB_CUDA(cudaFuncGetAttributes(&fa, fused_rowres_kernel));
  const int64_t kBudget = int64_t(optin) - int64_t(fa.sharedSizeBytes) - 1024;
  auto slice_of = [&](int c) { return (((nvec + c - 1) / c) * 16 + 127) & ~int64_t(127); };
  int C = 0, nbuf = 2;
  if (force_cluster > 0) {
    C = force_cluster;
    nbuf = force_nbuf >= 2 && force_nbuf <= kMaxBuf ? force_nbuf : (3 * slice_of(C) <= kBudget ? 3 : 2);
  } else {
    // prefer three slices in flight (measured: C=4 x 3 beats C=3 x 2 at V=151,936 although a cluster of 4
    // leaves 16 of 148 SMs idle); fall back to two
    for (int c = 1; c <= kMaxCluster && C == 0; ++c)
      if (3 * slice_of(c) <= kBudget) {
        C = c;
        nbuf = 3;
      }
    for (int c = 1; c <= kMaxCluster && C == 0; ++c)
      if (2 * slice_of(c) <= kBudget) {
        C = c;
        nbuf = 2;
      }
  }
  // every CTA of the cluster must own a non-empty slice (the statistics group of an empty rank would
  // never wait on a full barrier and would run ahead of the exchange)
  if (C > 1 && int64_t(C - 1) * ((nvec + C - 1) / C) >= nvec) {
    set_error("row-resident path: cluster_size=%d leaves a CTA without any of the row's %lld 16-byte vectors", C,
              (long long)nvec);
    return kInvalidArgument;
  }
  if (C <= 0 || C > kMaxCluster || nbuf * slice_of(C) > kBudget) {
    set_error("row-resident path: V=%lld does not fit a cluster of at most %d CTAs", (long long)p.V, kMaxCluster);
    return kUnsupported;
  }
  const uint32_t buf_stride = static_cast<uint32_t>(slice_of(C));
  const size_t smem = size_t(buf_stride) * nbuf + 1024;
  AB_CUDA(cudaFuncSetAttribute(fused_rowres_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(kRRThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = C;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cfg.gridDim = dim3(C);
  int max_clusters = 0;
  AB_CUDA(cudaOccupancyMaxActiveClusters(&max_clusters, fused_rowres_kernel, &cfg));
  if (max_clusters <= 0) {
    set_error("row-resident path: no cluster of %d CTAs with %zu bytes of shared memory fits", C, smem);
    return kUnsupported;
  }
  int64_t ncl = p.T < max_clusters ? p.T : max_clusters;
  {
    // experiments: AREAL_B200_ROWRES_MAX_CLUSTERS caps the number of clusters (how the step scales with the SM count:
    // 33 / 29 / 25 / 21 clusters of 4 -> 1.61 / 1.79 / 2.06 / 2.44 ms at T = 16,384, V = 151,936, i.e. time x clusters
    // is constant within 4 %: the kernel is bound by the SMs that work on it, not yet by DRAM)
    static const int cap = [] {
      const char* e = getenv("AREAL_B200_ROWRES_MAX_CLUSTERS");
      return e ? atoi(e) : 0;
    }();
    if (cap > 0 && ncl > cap) ncl = cap;
  }
  cfg.gridDim = dim3(static_cast<unsigned int>(ncl * C));
  AB_CUDA(cudaLaunchKernelEx(&cfg, fused_rowres_kernel, p, dlogits, d_row_stride_bytes, loss_scale, buf_stride, nbuf));
  return check_launch("fused_rowres");
}

}  // namespace ab
