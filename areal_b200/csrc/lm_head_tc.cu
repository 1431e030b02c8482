// LM head + log-prob / entropy WITHOUT the [T, V] logits tensor: a tcgen05 GEMM whose epilogue is the online
// softmax.  Replaces, for the forward-only case, `outputs.logits` followed by `gather_logprobs(logits, labels,
// temperature)` in the actor's compute_logp hook (reference areal/engine/ppo/actor.py:52-70) and
// `gather_logprobs_entropy` (areal/utils/functional.py:84-127) on the same logits.
//
//   logits[t, v] = bf16( sum_h hidden[t, h] * weight[v, h] )        (what F.linear returns in bf16)
//   per row: m = max_v logits, S = sum 2^((x-m)c), W = sum 2^((x-m)c) (x-m)c, x_label   (c = log2(e)/temperature)
//
// Design (sm_100a):
//   * operands: hidden [T, H] and weight [V, H], both K-major bf16, fetched by TMA tensor-map loads
//     (cp.async.bulk.tensor.2d, 64 x 128 boxes, 128-byte swizzle) into a ring of shared-memory stages;
//   * math: tcgen05.mma kind::f16 (bf16 x bf16 -> fp32), accumulators in tensor memory.  cta_group::2: a pair of
//     CTAs on one TPC computes a 256 x 256 tile (each CTA holds 128 rows of A and 128 rows of B per stage, the
//     accumulator rows of a CTA live in its own TMEM); cta_group::1: a 128 x 256 tile per CTA;
//   * two 256-column accumulators per CTA (all 512 TMEM columns): the epilogue of tile i overlaps the MMAs of
//     tile i+1;
//   * warp roles: warp 0 = TMA producer (one lane), warp 1 = MMA issuer (one lane, leader CTA only), warp 2 =
//     TMEM allocation, warps 4-7 = epilogue: tcgen05.ld 32x32b hands every thread 32 consecutive columns of ITS
//     row, so the online-softmax state of a row lives in one thread's registers and needs no cross-thread
//     reduction;
//   * work split: (row block, vocabulary chunk) items, row block fastest, so that the CTAs running at the same
//     time stream the same weight tiles (L2 hits) while every CTA re-reads its small hidden block; each item
//     leaves one partial (m, S, W, x_label) per row, combined by vocab_combine_kernel (the kernel of the
//     vocabulary-sharded path).
// Every mbarrier wait is bounded: a protocol error traps instead of hanging the GPU.
#include <cuda.h>

#include "common.cuh"
#include "lm_head_tc.h"
#include "logits_fused.h"

namespace ab {
namespace {

constexpr int kBlockM = 128;          // rows of A per CTA (= TMEM lanes)
constexpr int kBlockN = 256;          // columns of one accumulator tile (UMMA N)
constexpr int kBlockK = 64;           // bf16 elements per 128-byte swizzle row
constexpr int kUmmaK = 16;
constexpr int kTcThreads = 256;
constexpr int kEpiWarp0 = 4;          // warps 4..7: TMEM lane quarter = warp % 4
constexpr uint32_t kATileBytes = kBlockM * kBlockK * 2;     // 16 KB

template <int CG>
struct Cfg {
  static constexpr int kBRows = kBlockN / CG;                              // rows of B per CTA and stage
  static constexpr uint32_t kBTileBytes = kBRows * kBlockK * 2;
  static constexpr uint32_t kStageBytes = kATileBytes + kBTileBytes;      // 48 KB (CG 1) / 32 KB (CG 2)
  static constexpr int kStages = CG == 1 ? 4 : 6;                          // 192 KB either way
  static constexpr size_t kSmemBytes = size_t(kStages) * kStageBytes + 1024;
};

// ---- PTX wrappers ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// shared::cluster address of `saddr` (a shared::cta address of this CTA) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa(uint32_t saddr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrival on a barrier of another CTA of the cluster.  RELAXED: what has to be ordered before it (tcgen05.ld of the
// accumulator) is ordered by tcgen05.wait::ld + tcgen05.fence::before_thread_sync; a .release.cluster arrive costs a
// cluster-scope memory barrier (~0.7 us with TMA loads in flight: measured, it serialised the whole pipeline)
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_bar) {
  asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_bar) : "memory");
}
// bounded wait: ~seconds of polling, then trap (a broken pipeline must not hang the GPU)
__device__ __forceinline__ void mbar_wait_bounded(uint32_t bar, uint32_t parity) {
  for (uint32_t spin = 0; spin < (1u << 28); ++spin) {
    uint32_t ok;
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t"
        "}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    if (ok) return;
  }
  __trap();
}
template <int CG>
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  if constexpr (CG == 1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(bar)
        : "memory");
  } else {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], "
        "[%4];" ::"r"(dst),
        "l"(reinterpret_cast<uint64_t>(map)), "r"(c0), "r"(c1), "r"(bar)
        : "memory");
  }
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
template <int CG>
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t cols) {
  if constexpr (CG == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  } else {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
}
template <int CG>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
  if constexpr (CG == 1)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
  else
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, bf16 inputs, fp32 accumulate; issued by ONE thread
template <int CG>
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
  if constexpr (CG == 1) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc)
        : "memory");
  } else {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc)
        : "memory");
  }
}
// all MMAs issued so far by this thread -> one arrival on `bar` (of this CTA; CG 2: of both CTAs of the pair)
template <int CG>
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  if constexpr (CG == 1) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
  } else {
    asm volatile(
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
        "h"(uint16_t(3))
        : "memory");
  }
}
// 32 consecutive fp32 columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld_x32(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major, 128-byte swizzled operand tile (rows of 128 bytes, 8-row groups 1024 bytes apart): the layout a
// TMA load with CU_TENSOR_MAP_SWIZZLE_128B and a 64-element inner box writes.  Bits: start address >> 4
// [0,14), leading byte offset >> 4 [16,30) (unused for swizzled K-major, 1), stride byte offset >> 4 [32,46)
// (1024 B between 8-row groups), descriptor version 1 [46,48), layout type 2 = SWIZZLE_128B [61,64).
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr) {
  return uint64_t((saddr >> 4) & 0x3FFFu) | (uint64_t(1) << 16) | (uint64_t(1024 >> 4) << 32) | (uint64_t(1) << 46) |
         (uint64_t(2) << 61);
}
// instruction descriptor, kind::f16: D fp32 (bits [4,6) = 1), A and B bf16 ([7,10) = 1, [10,13) = 1), both
// K-major (bits 15, 16 = 0), N >> 3 at [17,23), M >> 4 at [24,29)
__host__ __device__ constexpr uint32_t make_idesc(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (uint32_t(N >> 3) << 17) | (uint32_t(M >> 4) << 24);
}

struct TcParams {
  int64_t T, V, H;
  const int64_t* labels;
  float c2;                 // log2(e) / temperature
  int tiles_per_chunk;      // vocabulary tiles (of 256 columns) per work item
  int n_chunks;             // R
  int mb_per_group;         // row blocks whose hidden states share L2 at one time (see item_coords)
  float* parts;             // [R][4][T]
  float* part_min;          // [R][T] or null: min of the logits per row and chunk
  int stages;               // pipeline stages in use (<= Cfg::kStages; profiling knob AREAL_B200_LMHEAD_STAGES)
  int want_entropy;         // 0: log-probs only (no W sums)
  int debug;                // profiling only (AREAL_B200_LMHEAD_DEBUG): 1 = epilogue loads TMEM but does not fold, 2 = no TMEM loads, 3 = no TMA loads, 4 = no MMAs
};

struct RowState {
  float m, S, W, xl, mn;
};

// fold 32 consecutive columns [col0, col0 + 32) of one row into its online-softmax state
// kEnt = false: log-probs only (compute_logp's case): the sum W for the entropy is not formed
template <bool kEnt>
__device__ __forceinline__ void fold32(RowState& st, const uint32_t* acc, int64_t col0, int64_t V, int64_t label, float c2,
                                       bool want_min) {
  // the reference rounds the projection to bf16 before the softmax (F.linear returns bf16)
  uint32_t pk[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) pk[j] = pack_bf16x2(__uint_as_float(acc[2 * j]), __uint_as_float(acc[2 * j + 1]));
  if (col0 + 32 <= V) {
    uint32_t mx = pk[0], mi = pk[0];
#pragma unroll
    for (int j = 1; j < 16; ++j) {
      mx = max_bf16x2(mx, pk[j]);
      if (want_min) mi = min_bf16x2(mi, pk[j]);
    }
    const float cmax = fmaxf(bf16_lo(mx), bf16_hi(mx));
    if (want_min) st.mn = fminf(st.mn, fminf(bf16_lo(mi), bf16_hi(mi)));
    if (cmax > st.m) {
      if (st.m != -INFINITY) {
        const float d = (st.m - cmax) * c2;
        const float f = ex2_approx(d);
        if (kEnt) st.W = f * fmaf(d, st.S, st.W);
        st.S *= f;
      }
      st.m = cmax;
    }
    const float nb = -st.m * c2;
    float s0 = 0.f, s1 = 0.f, w0 = 0.f, w1 = 0.f;
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const float t0 = fmaf(bf16_lo(pk[j]), c2, nb), t1 = fmaf(bf16_hi(pk[j]), c2, nb);
      const float e0 = ex2_approx(t0), e1 = ex2_approx(t1);
      s0 += e0;
      s1 += e1;
      if (kEnt) {
        w0 = fmaf(e0, t0, w0);
        w1 = fmaf(e1, t1, w1);
      }
    }
    st.S += s0 + s1;
    if (kEnt) st.W += w0 + w1;
  } else {
    // the chunk straddles the end of the vocabulary: element-wise, columns >= V left out
    for (int j = 0; j < 32; ++j) {
      if (col0 + j < V) {
        const float x = (j & 1) ? bf16_hi(pk[j >> 1]) : bf16_lo(pk[j >> 1]);
        if (want_min) st.mn = fminf(st.mn, x);
        if (x > st.m) {
          if (st.m != -INFINITY) {
            const float d = (st.m - x) * c2;
            const float f = ex2_approx(d);
            if (kEnt) st.W = f * fmaf(d, st.S, st.W);
            st.S *= f;
          }
          st.m = x;
        }
        const float t = (x - st.m) * c2;
        const float e = ex2_approx(t);
        st.S += e;
        if (kEnt) st.W = fmaf(e, t, st.W);
      }
    }
  }
  if (label >= col0 && label < col0 + 32) {
    const int k = int(label - col0);
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (j == k) st.xl = (j & 1) ? bf16_hi(pk[j >> 1]) : bf16_lo(pk[j >> 1]);
  }
}

// Work item -> (row block, vocabulary chunk).  Row blocks are taken in GROUPS small enough for their hidden
// states to stay in L2 (a group's rows are re-read for every vocabulary tile); inside a group the row block runs
// fastest, so the CTAs that run at the same time share the weight tiles they stream.
__device__ __forceinline__ void item_coords(int64_t item, int64_t num_mb, int G, int R, int64_t& mb, int64_t& ch) {
  const int64_t per_group = int64_t(G) * R;
  const int64_t g = item / per_group, r = item - g * per_group;
  const int64_t first = g * G;
  const int64_t size = num_mb - first < G ? num_mb - first : G;      // the last group may be smaller
  ch = r / size;
  mb = first + (r - ch * size);
}

template <int CG>
__global__ void __launch_bounds__(kTcThreads, 1)
lm_head_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, const TcParams p) {
  using C = Cfg<CG>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[C::kStages];     // operands of a stage have landed (TMA tx bytes)
  __shared__ __align__(8) uint64_t empty_bar[C::kStages];    // the MMAs reading a stage have retired
  __shared__ __align__(8) uint64_t tmem_full_bar[2];         // an accumulator is complete
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];        // the epilogue has drained an accumulator
  __shared__ uint32_t tmem_base_sm;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = CG == 1 ? 0u : cluster_ctarank();
  const bool leader = rank == 0;
  const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;

  if (threadIdx.x == 0) {
    for (int i = 0; i < C::kStages; ++i) {
      mbar_init(smem_u32(&full_bar[i]), 1);                   // the leader's producer: arrive.expect_tx for the pair's bytes
      mbar_init(smem_u32(&empty_bar[i]), 1);                  // one tcgen05.commit
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(smem_u32(&tmem_full_bar[i]), 1);              // one tcgen05.commit
      mbar_init(smem_u32(&tmem_empty_bar[i]), 4 * CG);        // one arrival per epilogue warp of the pair
    }
    mbar_fence_init();
  }
  if (warp == 2) tmem_alloc<CG>(smem_u32(&tmem_base_sm), 512);
  tc_fence_before();
  if constexpr (CG == 1) __syncthreads(); else cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_sm;

  const int64_t rows_per_item = int64_t(kBlockM) * CG;
  const int64_t num_mb = (p.T + rows_per_item - 1) / rows_per_item;
  const int64_t n_tiles = (p.V + kBlockN - 1) / kBlockN;
  const int64_t items = num_mb * p.n_chunks;
  const int64_t n_groups = gridDim.x / CG, group = blockIdx.x / CG;
  const int kb_count = int((p.H + kBlockK - 1) / kBlockK);

  if (warp == 0) {
    // ===== TMA producer: one lane, in both CTAs of a pair =====
    if (lane == 0) {
      uint32_t stage = 0, phase = 0;
      for (int64_t item = group; item < items; item += n_groups) {
        int64_t mb, ch;
        item_coords(item, num_mb, p.mb_per_group, p.n_chunks, mb, ch);
        const int m0 = int(mb * rows_per_item + int64_t(rank) * kBlockM);
        const int64_t nt0 = ch * p.tiles_per_chunk;
        const int64_t nt1 = nt0 + p.tiles_per_chunk < n_tiles ? nt0 + p.tiles_per_chunk : n_tiles;
        for (int64_t nt = nt0; nt < nt1; ++nt) {
          const int n0 = int(nt * kBlockN) + int(rank) * C::kBRows;
          for (int kb = 0; kb < kb_count; ++kb) {
            mbar_wait_bounded(smem_u32(&empty_bar[stage]), phase ^ 1u);
            const uint32_t sa = smem0 + stage * C::kStageBytes, sb = sa + kATileBytes;
            // the MMA issuer (leader CTA) waits on ITS full barrier for the bytes of both CTAs
            const uint32_t bar = CG == 1 ? smem_u32(&full_bar[stage]) : mapa(smem_u32(&full_bar[stage]), 0);
            // the leader arms ITS barrier for the bytes of both CTAs; the peer's loads complete on it too and the
            // peer never arrives (it cannot run a phase ahead: its stage is released by the leader's commit)
            uint32_t tx = C::kStageBytes * CG;
            if (p.debug == 3) tx = 0;                        // profiling: no loads, the MMAs run on stale shared memory
            if (p.debug == 5) tx = kATileBytes * CG;         // profiling: A loads only (no MMAs)
            if (p.debug == 6) tx = C::kBTileBytes * CG;      // profiling: B loads only (no MMAs)
            if (leader) mbar_arrive_expect_tx(smem_u32(&full_bar[stage]), tx);
            if (p.debug != 3 && p.debug != 6) tma_load_2d<CG>(sa, &map_a, kb * kBlockK, m0, bar);
            if (p.debug != 3 && p.debug != 5) {
#pragma unroll
              for (int r = 0; r < C::kBRows; r += 128)
                tma_load_2d<CG>(sb + r * (kBlockK * 2), &map_b, kb * kBlockK, n0 + r, bar);
            }
            if (++stage == uint32_t(p.stages)) {
              stage = 0;
              phase ^= 1u;
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: one lane of the leader CTA =====
    if (leader && lane == 0) {
      constexpr uint32_t idesc = make_idesc(kBlockM * CG, kBlockN);
      uint32_t stage = 0, phase = 0, tile = 0;
      for (int64_t item = group; item < items; item += n_groups) {
        int64_t mb, ch;
        item_coords(item, num_mb, p.mb_per_group, p.n_chunks, mb, ch);
        const int64_t nt0 = ch * p.tiles_per_chunk;
        const int64_t nt1 = nt0 + p.tiles_per_chunk < n_tiles ? nt0 + p.tiles_per_chunk : n_tiles;
        for (int64_t nt = nt0; nt < nt1; ++nt, ++tile) {
          const uint32_t acc = tile & 1u, acc_phase = (tile >> 1) & 1u;
          mbar_wait_bounded(smem_u32(&tmem_empty_bar[acc]), acc_phase ^ 1u);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + acc * kBlockN;
          for (int kb = 0; kb < kb_count; ++kb) {
            mbar_wait_bounded(smem_u32(&full_bar[stage]), phase);
            tc_fence_after();
            const uint32_t sa = smem0 + stage * C::kStageBytes, sb = sa + kATileBytes;
            const uint64_t da = smem_desc_sw128(sa), db = smem_desc_sw128(sb);
            if (p.debug < 4) {                               // (4, 5, 6: profiling, loads only)
#pragma unroll
              for (int k = 0; k < kBlockK / kUmmaK; ++k)    // +32 bytes along K inside the swizzle atom: +2 in the address field
                umma_bf16<CG>(d_tmem, da + uint64_t(k * 2), db + uint64_t(k * 2), idesc, uint32_t(kb | k));
            }
            umma_commit<CG>(smem_u32(&empty_bar[stage]));   // the stage may be refilled once these MMAs retire
            if (++stage == uint32_t(p.stages)) {
              stage = 0;
              phase ^= 1u;
            }
          }
          umma_commit<CG>(smem_u32(&tmem_full_bar[acc]));   // accumulator complete -> epilogue (both CTAs)
        }
      }
    }
  } else if (warp >= kEpiWarp0) {
    // ===== epilogue: thread = one row of the tile; online softmax over the item's vocabulary chunk =====
    const int q = warp & 3;                                  // TMEM lane quarter this warp may read
    const bool want_min = p.part_min != nullptr;
    const uint32_t empty_bar0 = CG == 1 ? smem_u32(&tmem_empty_bar[0]) : mapa(smem_u32(&tmem_empty_bar[0]), 0);
    uint32_t tile = 0;
    for (int64_t item = group; item < items; item += n_groups) {
      int64_t mb, ch;
      item_coords(item, num_mb, p.mb_per_group, p.n_chunks, mb, ch);
      const int64_t row = mb * rows_per_item + int64_t(rank) * kBlockM + q * 32 + lane;
      const int64_t label = row < p.T ? p.labels[row] : -1;
      RowState st = {-INFINITY, 0.f, 0.f, 0.f, INFINITY};
      const int64_t nt0 = ch * p.tiles_per_chunk;
      const int64_t nt1 = nt0 + p.tiles_per_chunk < n_tiles ? nt0 + p.tiles_per_chunk : n_tiles;
      for (int64_t nt = nt0; nt < nt1; ++nt, ++tile) {
        const uint32_t acc = tile & 1u, acc_phase = (tile >> 1) & 1u;
        mbar_wait_bounded(smem_u32(&tmem_full_bar[acc]), acc_phase);
        tc_fence_after();
        const uint32_t taddr = tmem_base + acc * kBlockN + (uint32_t(q * 32) << 16);
        const int64_t col_base = nt * kBlockN;
        // 32-column pieces inside the vocabulary (warp-uniform; the last tile may hang over its end)
        int nchunks = int((p.V - col_base + 31) / 32);
        if (nchunks > kBlockN / 32) nchunks = kBlockN / 32;
        if (p.debug == 2) nchunks = 0;
        // two register buffers: the TMEM load of piece c+1 is in flight while piece c is folded
        uint32_t va[32], vb[32];
        if (nchunks > 0) tmem_ld_x32(taddr, va);
#pragma unroll 1
        for (int c = 0; c < nchunks; c += 2) {
          tmem_ld_wait();
          if (c + 1 < nchunks) tmem_ld_x32(taddr + (c + 1) * 32, vb);
          if (p.debug == 1) st.S += __uint_as_float(va[0]) + __uint_as_float(va[31]);
          else if (p.want_entropy) fold32<true>(st, va, col_base + c * 32, p.V, label, p.c2, want_min);
          else fold32<false>(st, va, col_base + c * 32, p.V, label, p.c2, want_min);
          if (c + 1 < nchunks) {
            tmem_ld_wait();
            if (c + 2 < nchunks) tmem_ld_x32(taddr + (c + 2) * 32, va);
            if (p.debug == 1) st.S += __uint_as_float(vb[0]) + __uint_as_float(vb[31]);
            else if (p.want_entropy) fold32<true>(st, vb, col_base + (c + 1) * 32, p.V, label, p.c2, want_min);
            else fold32<false>(st, vb, col_base + (c + 1) * 32, p.V, label, p.c2, want_min);
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(empty_bar0 + acc * 8);   // tmem_empty_bar[acc] of the leader CTA
      }
      if (row < p.T) {
        // a label outside [0, V) has no logit: NaN, like areal_b200_logits_fwd (the reference's gather would raise)
        if (ch == 0 && (label < 0 || label >= p.V)) st.xl = __int_as_float(0x7fc00000);
        float* o = p.parts + int64_t(ch) * 4 * p.T + row;
        o[0] = st.m;
        o[p.T] = st.S;
        o[2 * p.T] = st.W;
        o[3 * p.T] = st.xl;
        if (want_min) p.part_min[int64_t(ch) * p.T + row] = st.mn;
      }
    }
  }
  // ===== teardown: nobody frees TMEM or leaves while the peer may still signal into this CTA =====
  tc_fence_before();
  if constexpr (CG == 1) __syncthreads(); else cluster_sync_all();
  if (warp == 2) tmem_dealloc<CG>(tmem_base, 512);
}

// min over chunks of the per-chunk row minima
__global__ void min_over_chunks_kernel(const float* __restrict__ part_min, int R, int64_t T, float* __restrict__ out) {
  const int64_t t = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
  if (t >= T) return;
  float m = INFINITY;
  for (int r = 0; r < R; ++r) m = fminf(m, part_min[int64_t(r) * T + t]);
  out[t] = m;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int encode_fn(EncodeTiledFn* out) {
  static EncodeTiledFn cached = nullptr;
  if (!cached) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    AB_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    if (qres != cudaDriverEntryPointSuccess || fn == nullptr) {
      set_error("cuTensorMapEncodeTiled is not available from this driver");
      return kCudaError;
    }
    cached = reinterpret_cast<EncodeTiledFn>(fn);
  }
  *out = cached;
  return kOk;
}

// [rows, H] bf16, K-major; box = 64 x 128, 128-byte swizzle, out-of-range elements read as zero
int make_map(CUtensorMap* map, const void* base, int64_t rows, int64_t H, int64_t row_stride_elems) {
  EncodeTiledFn enc;
  const int rc = encode_fn(&enc);
  if (rc) return rc;
  const cuuint64_t dims[2] = {cuuint64_t(H), cuuint64_t(rows)};
  const cuuint64_t strides[1] = {cuuint64_t(row_stride_elems) * 2};
  const cuuint32_t box[2] = {cuuint32_t(kBlockK), 128};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%lld, H=%lld, stride=%lld)", int(r), (long long)rows,
              (long long)H, (long long)row_stride_elems);
    return kCudaError;
  }
  return kOk;
}

}  // namespace

// Work split: R vocabulary chunks so that (row blocks x R) fills the CTA groups of the grid evenly.
static void plan(int64_t T, int64_t V, int cta_group, int n_groups, int* tiles_per_chunk, int* n_chunks) {
  const int64_t rows_per_item = int64_t(kBlockM) * cta_group;
  const int64_t num_mb = (T + rows_per_item - 1) / rows_per_item;
  const int64_t n_tiles = (V + kBlockN - 1) / kBlockN;
  // candidates: R in [1, min(n_tiles, 64)]; score = work / (waves * capacity), ties -> fewer chunks (fewer partials)
  int best_r = 1;
  double best = -1.0;
  const int64_t max_r = n_tiles < 64 ? n_tiles : 64;
  for (int64_t r = 1; r <= max_r; ++r) {
    const int64_t tpc = (n_tiles + r - 1) / r;
    const int64_t chunks = (n_tiles + tpc - 1) / tpc;
    if (chunks != r) continue;
    const int64_t items = num_mb * chunks;
    const int64_t waves = (items + n_groups - 1) / n_groups;
    // time ~ waves * tpc tiles (the last chunk may be shorter; the longest one decides)
    const double eff = double(num_mb * n_tiles) / (double(waves) * double(tpc) * double(n_groups));
    // a chunk shorter than ~8 tiles pays its pipeline fill and the extra partials
    const double penalty = tpc < 8 ? 0.9 : 1.0;
    if (eff * penalty > best + 1e-9) {
      best = eff * penalty;
      best_r = int(chunks);
    }
  }
  const int64_t tpc = (n_tiles + best_r - 1) / best_r;
  *tiles_per_chunk = int(tpc);
  *n_chunks = int((n_tiles + tpc - 1) / tpc);
}

static int cta_group_choice() {
  static int cached = 0;
  if (!cached) {
    const char* e = getenv("AREAL_B200_LMHEAD_CTA_GROUP");
    cached = (e && e[0] == '1') ? 1 : 2;
  }
  return cached;
}

static int grid_groups(int cta_group, int* n_groups_out) {
  // one CTA per SM (192 KB of shared memory, all 512 TMEM columns); pairs need two SMs of one TPC
  int n = num_sms();
  if (cta_group == 2) {
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(kTcThreads);
    cfg.dynamicSmemBytes = Cfg<2>::kSmemBytes;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cfg.gridDim = dim3(2);
    int max_clusters = 0;
    AB_CUDA(cudaFuncSetAttribute(lm_head_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg<2>::kSmemBytes)));
    AB_CUDA(cudaOccupancyMaxActiveClusters(&max_clusters, lm_head_tc_kernel<2>, &cfg));
    if (max_clusters <= 0) {
      set_error("lm_head: no CTA pair with %zu bytes of shared memory fits this device", Cfg<2>::kSmemBytes);
      return kUnsupported;
    }
    *n_groups_out = max_clusters < n / 2 ? max_clusters : n / 2;
  } else {
    AB_CUDA(cudaFuncSetAttribute(lm_head_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(Cfg<1>::kSmemBytes)));
    *n_groups_out = n;
  }
  return kOk;
}

int64_t lm_head_tc_workspace_bytes(int64_t T, int64_t V) {
  // parts [R][4][T] + minima [R][T], R <= 64
  int tpc = 0, r = 0, groups = 0;
  const int cg = cta_group_choice();
  if (grid_groups(cg, &groups) != kOk) groups = num_sms() / cg;
  plan(T < 1 ? 1 : T, V < 1 ? 1 : V, cg, groups, &tpc, &r);
  return int64_t(r) * 5 * (T < 1 ? 1 : T) * 4 + 256;
}

int launch_lm_head_tc(const void* hidden, const void* weight, const int64_t* labels, int64_t T, int64_t H, int64_t V,
                      int64_t hidden_row_stride, int64_t weight_row_stride, float inv_temperature, float* logp,
                      float* entropy, float* row_max, float* ln_s, float* raw_min, void* workspace,
                      int64_t workspace_bytes, cudaStream_t stream) {
  if (T == 0) return kOk;
  AB_REQUIRE((reinterpret_cast<uintptr_t>(hidden) & 15) == 0 && (reinterpret_cast<uintptr_t>(weight) & 15) == 0,
             "lm_head: hidden and weight must be 16-byte aligned");
  AB_REQUIRE(hidden_row_stride % 8 == 0 && weight_row_stride % 8 == 0 && hidden_row_stride >= H && weight_row_stride >= H,
             "lm_head: row strides must be multiples of 8 elements and >= H");
  AB_REQUIRE(H >= 1 && V >= 1 && T < (int64_t(1) << 31) && V < (int64_t(1) << 31), "lm_head: bad sizes");
  const int cg = cta_group_choice();
  int n_groups = 0;
  int rc = grid_groups(cg, &n_groups);
  if (rc) return rc;
  TcParams p;
  p.T = T;
  p.V = V;
  p.H = H;
  p.labels = labels;
  p.c2 = inv_temperature * kLog2e;
  p.want_entropy = entropy != nullptr;
  {
    const char* e = getenv("AREAL_B200_LMHEAD_DEBUG");
    p.debug = e ? atoi(e) : 0;
    const char* st = getenv("AREAL_B200_LMHEAD_STAGES");
    const int max_stages = cg == 1 ? Cfg<1>::kStages : Cfg<2>::kStages;
    p.stages = st ? atoi(st) : max_stages;
    if (p.stages < 1 || p.stages > max_stages) p.stages = max_stages;
  }
  plan(T, V, cg, n_groups, &p.tiles_per_chunk, &p.n_chunks);
  {
    // hidden states of one group of row blocks <= ~48 MB (L2 is 126 MB and also holds the weight tiles in flight)
    const int64_t rows_per_item = int64_t(kBlockM) * cg;
    const int64_t num_mb = (T + rows_per_item - 1) / rows_per_item;
    const int64_t a_bytes = T * H * 2;
    const char* e = getenv("AREAL_B200_LMHEAD_GROUP_MB");
    const int64_t budget = (e ? atoll(e) : 48) << 20;
    const int64_t n_mgroups = a_bytes <= budget + (budget >> 2) ? 1 : (a_bytes + budget - 1) / budget;
    p.mb_per_group = int((num_mb + n_mgroups - 1) / n_mgroups);
    if (p.mb_per_group < 1) p.mb_per_group = 1;
  }
  const int64_t need = int64_t(p.n_chunks) * 5 * T * 4;
  if (workspace == nullptr || workspace_bytes < need) {
    set_error("lm_head: workspace too small (%lld < %lld bytes)", (long long)workspace_bytes, (long long)need);
    return kWorkspaceTooSmall;
  }
  p.parts = static_cast<float*>(workspace);
  p.part_min = raw_min ? p.parts + int64_t(p.n_chunks) * 4 * T : nullptr;
  CUtensorMap map_a, map_b;
  if ((rc = make_map(&map_a, hidden, T, H, hidden_row_stride)) != kOk) return rc;
  if ((rc = make_map(&map_b, weight, V, H, weight_row_stride)) != kOk) return rc;
  const int64_t rows_per_item = int64_t(kBlockM) * cg;
  const int64_t items = ((T + rows_per_item - 1) / rows_per_item) * p.n_chunks;
  const int64_t groups = items < n_groups ? items : n_groups;
  if (cg == 2) {
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(kTcThreads);
    cfg.dynamicSmemBytes = Cfg<2>::kSmemBytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cfg.gridDim = dim3(static_cast<unsigned int>(groups * 2));
    AB_CUDA(cudaLaunchKernelEx(&cfg, lm_head_tc_kernel<2>, map_a, map_b, p));
  } else {
    lm_head_tc_kernel<1><<<static_cast<unsigned int>(groups), kTcThreads, Cfg<1>::kSmemBytes, stream>>>(map_a, map_b, p);
  }
  rc = check_launch("lm_head_tc");
  if (rc) return rc;
  rc = launch_vocab_combine(p.parts, p.n_chunks, T, inv_temperature, logp, entropy, row_max, ln_s, stream);
  if (rc) return rc;
  if (raw_min) {
    min_over_chunks_kernel<<<static_cast<unsigned int>((T + 255) / 256), 256, 0, stream>>>(p.part_min, p.n_chunks, T, raw_min);
    rc = check_launch("lm_head_min");
  }
  return rc;
}

}  // namespace ab
