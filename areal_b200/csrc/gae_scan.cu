// Generalised advantage estimation as a segmented REVERSE scan of affine maps.
//
// Replaces (reference, relative to its root):
//   csrc/cugae/gae.cu:10-60            gae_1d_nolp_misalign  (one thread per sequence, serial)
//   realhf/impl/model/utils/ppo_functional.py:292-323  pygae1d_nolp_misalign (CPU path)
//   areal/engine/ppo/actor.py:131-152  padded [B,S] masked GAE python loop
//   csrc/cugae/gae.cu:62-134           gae_2d_olp (and gae_2d_nolp, whose host wrapper runs the
//                                      olp kernel, gae.cu:199)
//
// A_t = delta_t + a_t * A_{t+1} is the affine map x -> b_t + a_t x with a_t = gamma*lambda inside a
// sequence and a_t = 0 at its last step, so sequence boundaries need no flags: composition
// (a,b)o(a',b') = (a a', b + a b') annihilates the carry there.  Tokens are tiled 2048 per CTA
// (256 threads x 8 consecutive tokens, 128-bit loads/stores); each thread folds its 8 maps, the
// warp does a shuffle suffix-scan, warps combine through shared memory, and tiles chain with a
// single-pass decoupled look-back (tickets issued right-to-left) that stops as soon as the
// accumulated `a` is exactly 0, i.e. at the first sequence end.
//
// Numerics: delta_t is formed in fp32 with the reference's operation order (no FMA contraction),
// the scan itself runs in fp64 so the reassociation of the parallel scan cannot be seen in the
// fp32 result: the output is the correctly rounded recurrence, and its distance to the
// reference is the reference's own sequential fp32 rounding (tests report both against fp64).
#include <mutex>

#include <stdlib.h>

#include "common.cuh"
#include "gae_scan.h"

// Scratch of the allocation-free entry point (areal_b200_gae1d_nolp_misalign_static): the look-back
// header (64 B) plus one 32-byte descriptor per 2048-token tile, for at most 4,194,304 tokens per
// call -> 65,600 bytes.  Lives in the module image: every GPU that loads the library holds these
// 64 KB until the process exits.  This name is a HOST-side stand-in in host code, never a device
// address; the launcher below resolves it with cudaGetSymbolAddress().
#define AB_GAE1D_STATIC_MAX_TOKENS (int64_t(1) << 22)
#define AB_GAE1D_STATIC_SCRATCH_BYTES (64 + (AB_GAE1D_STATIC_MAX_TOKENS / 2048) * 32)
static __device__ __align__(256) unsigned char g_gae1d_static_scratch[AB_GAE1D_STATIC_SCRATCH_BYTES];

namespace ab {
namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kItems = 8;
constexpr int kTile = kThreads * kItems;   // 2048 tokens

struct Affine {
  double a, b;
};
__device__ __forceinline__ Affine ident() { return Affine{1.0, 0.0}; }
// f after g
__device__ __forceinline__ Affine compose(const Affine& f, const Affine& g) {
  return Affine{f.a * g.a, fma(f.a, g.b, f.b)};
}
__device__ __forceinline__ Affine shfl_down(const Affine& v, int d) {
  return Affine{__shfl_down_sync(0xffffffffu, v.a, d), __shfl_down_sync(0xffffffffu, v.b, d)};
}

// Block-wide suffix scan of per-thread maps (thread x owns positions left of thread x+1).
// excl  = F_{x+1} o F_{x+2} o ... o F_{last}   (apply to the tile carry-in to get the value
//         entering thread x from the right);  tile = F_0 o ... o F_last.
// Contains one __syncthreads(); `wsm` may be rewritten after the caller's next barrier.
template <int kNW = kWarps>
__device__ __forceinline__ void block_suffix_scan(const Affine& agg, Affine* wsm, int warp, int lane,
                                                  Affine& excl, Affine& tile) {
  constexpr int kWarps = kNW;
  Affine inc = agg;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const Affine o = shfl_down(inc, d);
    if (lane + d < 32) inc = compose(inc, o);
  }
  if (lane == 0) wsm[warp] = inc;
  __syncthreads();
  Affine W = ident();
  for (int w = kWarps - 1; w > warp; --w) W = compose(wsm[w], W);
  const Affine nxt = shfl_down(inc, 1);
  excl = (lane < 31) ? compose(nxt, W) : W;
  tile = W;
  for (int w = warp; w >= 0; --w) tile = compose(wsm[w], tile);
}

// ---- decoupled look-back descriptors --------------------------------------------------------------
// Three self-validating 8-byte words per tile: a word is "not published yet" while it holds kEmpty
// (all ones -- never produced by arithmetic; canonicalised below), so publishing is ONE relaxed
// 8-byte store per word and needs no fence: there is no separate flag to order against.
struct __align__(16) TileDesc {
  unsigned long long inc;      // bits of A at the first token of the tile, once known
  unsigned long long agg_a;    // bits of the tile map (a, b)
  unsigned long long agg_b;
  int seq0;                    // sequence of the tile's first token (written by the prep kernel)
  int end0;                    // cu[seq0 + 1]: saves the tile one dependent load
};
struct WorkspaceHeader {
  unsigned int pad[16];
};
static_assert(sizeof(TileDesc) == 32, "TileDesc layout");
static_assert(sizeof(WorkspaceHeader) == 64, "header layout");
constexpr unsigned long long kEmpty = ~0ull;

__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_f64(unsigned long long* p, double x) {
  unsigned long long v = static_cast<unsigned long long>(__double_as_longlong(x));
  if (v == kEmpty) v = 0x7ff8000000000000ull;      // canonical NaN instead of the reserved pattern
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// One thread per tile: mark the descriptor empty and look up the sequence that holds the tile's first
// token (smallest i with cu[i+1] > first token).  Replaces a memset node, so it costs no extra launch.
__global__ void gae1d_prep_kernel(const int32_t* __restrict__ cu, int B, int64_t T, int64_t ntiles,
                                  TileDesc* __restrict__ desc, int tile_tokens) {
  // programmatic dependent launch: the scan kernel may start (and prefetch its rewards) while this
  // grid runs; it waits (griddepcontrol.wait) for this grid to COMPLETE before it reads a descriptor
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const int64_t t = blockIdx.x * int64_t(blockDim.x) + threadIdx.x;
  if (t >= ntiles) return;
  const int64_t tok = t * tile_tokens;
  int lo = 0, hi = B - 1;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (int64_t(__ldg(cu + mid + 1)) > tok) hi = mid; else lo = mid + 1;
  }
  TileDesc d;
  d.inc = kEmpty;
  d.agg_a = kEmpty;
  d.agg_b = kEmpty;
  d.seq0 = lo;
  d.end0 = __ldg(cu + lo + 1);
  desc[t] = d;
}

// ---- cluster exchange (small batches: one launch, no global descriptors) ---------------------------
__device__ __forceinline__ uint32_t gae_cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t gae_mapa(uint32_t saddr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
// 16-byte store into another CTA's shared memory that also counts 16 bytes on that CTA's mbarrier: the
// receiver sees the data once its barrier phase completes, no fence and no flag
__device__ __forceinline__ void st_async_affine(uint32_t remote_addr, const double a, const double b, uint32_t remote_bar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.b64 [%0], {%1, %2}, [%3];" ::"r"(remote_addr),
               "l"(__double_as_longlong(a)), "l"(__double_as_longlong(b)), "r"(remote_bar)
               : "memory");
}
__device__ __forceinline__ void mbar_wait_acq_cluster(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAITG_%=:\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONEG_%=;\n\t"
      "bra WAITG_%=;\n\t"
      "DONEG_%=:\n\t"
      "}" ::"r"(bar),
      "r"(parity)
      : "memory");
}
constexpr int kGaeClusterMax = 8;

// ---- 1D packed, misaligned values ------------------------------------------------------------------
// kT threads per CTA: 256 (2048-token tiles, 8 CTAs/SM) or 512 (4096-token tiles, 4 CTAs/SM: half as many
// look-backs and CTA barriers per token; used for large batches)
// kCluster: the whole batch is ONE thread-block cluster (at most 8 tiles): tile aggregates travel right-to-left
// through distributed shared memory (st.async + mbarrier complete_tx), every thread finds its sequence by a
// binary search -- no prep kernel, no global descriptors, one launch.
template <int kT, bool kCluster = false>
__global__ void __launch_bounds__(kT, 2048 / kT)
gae1d_kernel(const Gae1dParams p) {
  constexpr int kThreads = kT, kWarps = kT / 32, kTile = kT * kItems;
  __shared__ Affine wsm[kWarps];
  __shared__ double carry_sm;
  __shared__ __align__(16) Affine inbox[kCluster ? kGaeClusterMax : 1];   // [sender rank]: maps of the tiles to the right
  __shared__ __align__(8) uint64_t xbar;
  // cluster mode: cu_seqlens staged in shared memory by ONE round of parallel loads, so that the per-thread
  // sequence search costs no chain of dependent global loads (each one a DRAM round trip on a cold batch)
  constexpr int kCuSm = kCluster ? 2048 : 1;
  __shared__ int cu_sm[kCuSm + 1];
  // per-token values parked in shared memory ([item][thread]: conflict-free) while the scan runs, so
  // that 8 CTAs fit an SM (32 registers per thread): this kernel lives on memory-level parallelism
  __shared__ float v_sm[kItems][kThreads];
  __shared__ float d_sm[kItems][kThreads];

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int64_t ntiles = (p.T + kTile - 1) / kTile;
  TileDesc* desc = reinterpret_cast<TileDesc*>(reinterpret_cast<char*>(p.workspace) + sizeof(WorkspaceHeader));

  // Tiles are taken right-to-left in blockIdx order: block 0 owns the LAST tile, so a tile only waits
  // on tiles of lower-numbered blocks.  Like CUB's decoupled look-back scan this relies on the
  // hardware dispatching thread blocks in increasing blockIdx order (waiting never blocks a
  // predecessor from being scheduled).
  const int64_t tile = ntiles - 1 - int64_t(blockIdx.x);
  const int64_t tile_start = tile * kTile;
  if (kCluster) {
    // rank == blockIdx.x: rank 0 owns the LAST tile and receives nothing; rank r receives r maps
    if (tid == 0) {
      mbar_init(smem_u32(&xbar), 1);
      mbar_fence_init();
    }
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");   // waited on right before the exchange
    if (p.B <= kCuSm)
      for (int i = tid; i <= int(p.B); i += kThreads) cu_sm[i] = __ldg(p.cu + i);
  }
  const int ntok = static_cast<int>(p.T - tile_start < kTile ? p.T - tile_start : kTile);
  const int B = static_cast<int>(p.B);

  const int p0 = tid * kItems;                      // first token of this thread, relative to the tile
  const int nvalid = p0 >= ntok ? 0 : (ntok - p0 < kItems ? ntok - p0 : kItems);
  const float* rew = p.rewards + tile_start;
  const float* val = p.values + tile_start;         // + sequence id per token ("misaligned")

  // rewards do not depend on the sequence lookup: get them in flight first
  float r[kItems];
  if (p.vec_ok && nvalid == kItems) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(rew + p0));
    const float4 b = __ldg(reinterpret_cast<const float4*>(rew + p0) + 1);
    r[0] = a.x; r[1] = a.y; r[2] = a.z; r[3] = a.w;
    r[4] = b.x; r[5] = b.y; r[6] = b.z; r[7] = b.w;
  } else {
#pragma unroll
    for (int k = 0; k < kItems; ++k) r[k] = (k < nvalid) ? __ldg(rew + p0 + k) : 0.f;
  }

  // everything above is independent of the prep kernel; from here on its descriptors are read
  // (no-op unless launched as a programmatic dependent)
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (kCluster) __syncthreads();                    // cu_sm is complete

  float vcur[kItems], delta[kItems];
  unsigned int last_bits = 0;
  const double gl = double(p.gl);
  if (nvalid > 0) {
    // my sequence: start from the tile's (prep kernel, or sequence 0 for a single tile) and walk the
    // few sequence ends that lie before my first token; cu entries are shared by the CTA -> L1 hits
    int seq = 0;
    int64_t end0;
    if (kCluster) {
      int lo = 0, hi = B - 1;
      const int64_t j0 = tile_start + p0;
      if (B <= kCuSm) {
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          if (int64_t(cu_sm[mid + 1]) > j0) hi = mid; else lo = mid + 1;
        }
        end0 = cu_sm[lo + 1];
      } else {
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          if (int64_t(__ldg(p.cu + mid + 1)) > j0) hi = mid; else lo = mid + 1;
        }
        end0 = __ldg(p.cu + lo + 1);
      }
      seq = lo;
    } else if (ntiles > 1) {
      const int2 se = *reinterpret_cast<const int2*>(&desc[tile].seq0);
      seq = se.x;
      end0 = se.y;
    } else {
      end0 = __ldg(p.cu + 1);
    }
    int next_b = static_cast<int>(end0 - tile_start);   // position right after `seq` ends
    int steps = 0;
    while (next_b <= p0 && seq < B - 1) {
      if (++steps > 16) {           // many tiny sequences: finish with a binary search
        int lo = seq, hi = B - 1;
        const int64_t j0 = tile_start + p0;
        while (lo < hi) {
          const int mid = (lo + hi) >> 1;
          if (int64_t(__ldg(p.cu + mid + 1)) > j0) hi = mid; else lo = mid + 1;
        }
        seq = lo;
        next_b = static_cast<int>(int64_t(__ldg(p.cu + seq + 1)) - tile_start);
        break;
      }
      ++seq;
      next_b = static_cast<int>(int64_t(__ldg(p.cu + seq + 1)) - tile_start);
    }
    if (nvalid == kItems && next_b >= p0 + kItems) {
      // fast path: all 8 tokens in one sequence; only the last one can be a sequence end
      const float* vp = val + p0 + seq;
      float v[kItems + 1];
#pragma unroll
      for (int k = 0; k <= kItems; ++k) v[k] = __ldg(vp + k);
      if (next_b == p0 + kItems) {
        last_bits = 1u << (kItems - 1);
        if (!p.bootstrap[seq]) v[kItems] = 0.f;                  // gae.cu:23
      }
#pragma unroll
      for (int k = 0; k < kItems; ++k) {
        vcur[k] = v[k];
        delta[k] = __fsub_rn(__fadd_rn(r[k], __fmul_rn(p.g, v[k + 1])), v[k]);   // ppo_functional.py:316
      }
    } else {
      // a sequence end inside this thread's tokens (or a ragged last tile): walk token by token
      float vprev = 0.f;
      bool chain = false;
#pragma unroll
      for (int k = 0; k < kItems; ++k) {
        if (k < nvalid) {
          const int pos = p0 + k;
          while (seq < B - 1 && pos >= next_b) {
            ++seq;
            next_b = static_cast<int>(int64_t(__ldg(p.cu + seq + 1)) - tile_start);
            chain = false;
          }
          const bool is_last = (pos + 1 == next_b);
          const float v = chain ? vprev : __ldg(val + pos + seq);
          float nv = __ldg(val + pos + seq + 1);
          vprev = nv;
          chain = !is_last;
          if (is_last) {
            last_bits |= 1u << k;
            if (!p.bootstrap[seq]) nv = 0.f;
          }
          vcur[k] = v;
          delta[k] = __fsub_rn(__fadd_rn(r[k], __fmul_rn(p.g, nv)), v);
        }
      }
    }
  }

  // fold the thread's maps right-to-left, parking the per-token values on the way.  The common case --
  // all 8 tokens inside one sequence, only the last one possibly its end -- has a = gl for 7 of the 8 maps:
  // no per-token selects or guards.
  const bool plain = (nvalid == kItems) && ((last_bits & ((1u << (kItems - 1)) - 1u)) == 0u);
  Affine agg = ident();
  if (plain) {
    double b = double(delta[kItems - 1]);
#pragma unroll
    for (int k = kItems - 2; k >= 0; --k) b = fma(gl, b, double(delta[k]));
    const double gl2 = gl * gl, gl4 = gl2 * gl2;
    agg.a = last_bits ? 0.0 : gl4 * gl4;
    agg.b = b;
#pragma unroll
    for (int k = 0; k < kItems; ++k) {
      v_sm[k][tid] = vcur[k];
      d_sm[k][tid] = delta[k];
    }
  } else {
#pragma unroll
    for (int k = kItems - 1; k >= 0; --k) {
      if (k < nvalid) {
        const double a = (last_bits >> k) & 1u ? 0.0 : gl;
        agg.b = fma(a, agg.b, double(delta[k]));
        agg.a = a * agg.a;
        v_sm[k][tid] = vcur[k];
        d_sm[k][tid] = delta[k];
      }
    }
  }

  // block-wide suffix scan: warp shuffle, then warps through shared memory
  Affine inc = agg, nxt;
  bool by_ballot = false;
  if constexpr (kT == 64 && !kCluster) by_ballot = __all_sync(0xffffffffu, plain);
  if (by_ballot) {
    // Every lane's map is (gl^8, b) or (0, b): the `a` halves need no shuffles.  One ballot says which lanes hold a
    // sequence end; a lane's span of d lanes passes the right-hand part on iff none of them does, with the factor
    // gl^(8d) (squared from round to round); the factor of the lane's whole suffix, gl^(8(32 - lane)), is put together
    // from the same powers by the bits of its exponent.  Half the shuffles and selects of the general scan below.
    const unsigned int mine = __ballot_sync(0xffffffffu, last_bits != 0u) >> lane;   // bit j: lane + j ends a sequence
    const double gl2 = gl * gl, gl4 = gl2 * gl2, g8 = gl4 * gl4;
    double b = agg.b, P = g8, apow = 1.0;           // apow -> gl^(8(31 - lane)): the suffix that starts at lane + 1
    const int e = 31 - lane;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const double ob = __shfl_down_sync(0xffffffffu, b, d);
      if (lane + d < 32 && (mine & ((1u << d) - 1u)) == 0u) b = fma(P, ob, b);
      if (e & d) apow *= P;
      if (d < 16) P *= P;
    }
    inc.a = mine ? 0.0 : apow * g8;
    inc.b = b;
    nxt.a = (mine >> 1) ? 0.0 : apow;
    nxt.b = __shfl_down_sync(0xffffffffu, b, 1);
  } else {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const Affine o = shfl_down(inc, d);
      if (lane + d < 32) inc = compose(inc, o);
    }
    nxt = shfl_down(inc, 1);
  }
  if (lane == 0) wsm[warp] = inc;
  __syncthreads();
  Affine W = ident();
  for (int w = kWarps - 1; w > warp; --w) W = compose(wsm[w], W);
  const Affine excl = (lane < 31) ? compose(nxt, W) : W;

  double carry = 0.0;
  if (kCluster) {
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");   // every CTA's barrier is initialised
    if (ntiles > 1) {
      if (tid == 0) {
        const Affine tagg = compose(inc, W);        // thread 0: whole tile
        const uint32_t me = gae_cluster_ctarank(), n = static_cast<uint32_t>(ntiles);
        for (uint32_t dst = me + 1; dst < n; ++dst)   // every tile to the left needs this tile's map
          st_async_affine(gae_mapa(smem_u32(&inbox[me]), dst), tagg.a, tagg.b, gae_mapa(smem_u32(&xbar), dst));
        if (me > 0) {
          mbar_arrive_expect_tx(smem_u32(&xbar), 16u * me);
          mbar_wait_acq_cluster(smem_u32(&xbar), 0);
          double x = 0.0;                             // A beyond the last token
          for (uint32_t k = 0; k < me; ++k) x = fma(inbox[k].a, x, inbox[k].b);   // rank 0 is the rightmost tile
          carry = x;
        }
        carry_sm = carry;
      }
      __syncthreads();
      carry = carry_sm;
    }
  } else if (ntiles > 1) {
    if (tid == 0) {
      const Affine tagg = compose(inc, W);          // thread 0: whole tile
      TileDesc* me = desc + tile;
      if (tile == ntiles - 1) {
        st_relaxed_f64(&me->inc, tagg.b);           // nothing to the right: A_{T} = 0
      } else {
        st_relaxed_f64(&me->agg_a, tagg.a);
        st_relaxed_f64(&me->agg_b, tagg.b);
        Affine run = ident();        // maps of tiles tile+1 .. k-1
        for (int64_t k = tile + 1;; ++k) {
          if (k >= ntiles || run.a == 0.0) {
            carry = run.b;
            break;
          }
          const TileDesc* d = desc + k;
          unsigned long long vi, va, vb;
          for (;;) {
            vi = ld_relaxed_u64(&d->inc);
            if (vi != kEmpty) break;
            va = ld_relaxed_u64(&d->agg_a);
            vb = ld_relaxed_u64(&d->agg_b);
            if (va != kEmpty && vb != kEmpty) break;
            __nanosleep(20);
          }
          if (vi != kEmpty) {
            carry = fma(run.a, __longlong_as_double(static_cast<long long>(vi)), run.b);
            break;
          }
          run = compose(run, Affine{__longlong_as_double(static_cast<long long>(va)),
                                    __longlong_as_double(static_cast<long long>(vb))});
        }
        st_relaxed_f64(&me->inc, fma(tagg.a, carry, tagg.b));
      }
      carry_sm = carry;
    }
    __syncthreads();
    carry = carry_sm;
  }

  if (nvalid > 0) {
    double A = fma(excl.a, carry, excl.b);   // value entering this thread from the right
    float adv[kItems], ret[kItems];
    if (plain) {
      A = fma(last_bits ? 0.0 : gl, A, double(d_sm[kItems - 1][tid]));
      adv[kItems - 1] = float(A);
      ret[kItems - 1] = __fadd_rn(adv[kItems - 1], v_sm[kItems - 1][tid]);
#pragma unroll
      for (int k = kItems - 2; k >= 0; --k) {
        A = fma(gl, A, double(d_sm[k][tid]));
        adv[k] = float(A);
        ret[k] = __fadd_rn(adv[k], v_sm[k][tid]);                 // ppo_functional.py:319
      }
    } else {
#pragma unroll
      for (int k = kItems - 1; k >= 0; --k) {
        if (k < nvalid) {
          const double a = (last_bits >> k) & 1u ? 0.0 : gl;
          A = fma(a, A, double(d_sm[k][tid]));
          adv[k] = float(A);
          ret[k] = __fadd_rn(adv[k], v_sm[k][tid]);               // ppo_functional.py:319
        }
      }
    }
    float* ao = p.adv + tile_start + p0;
    float* ro = p.ret + tile_start + p0;
    if (p.vec_ok && nvalid == kItems) {
      reinterpret_cast<float4*>(ao)[0] = make_float4(adv[0], adv[1], adv[2], adv[3]);
      reinterpret_cast<float4*>(ao)[1] = make_float4(adv[4], adv[5], adv[6], adv[7]);
      reinterpret_cast<float4*>(ro)[0] = make_float4(ret[0], ret[1], ret[2], ret[3]);
      reinterpret_cast<float4*>(ro)[1] = make_float4(ret[4], ret[5], ret[6], ret[7]);
    } else {
#pragma unroll
      for (int k = 0; k < kItems; ++k)
        if (k < nvalid) {
          ao[k] = adv[k];
          ro[k] = ret[k];
        }
    }
  }
}

// ---- padded [B,S] masked (areal-lite) ----------------------------------------------------------------
__device__ __forceinline__ float mask_as_float(const void* base, int code, int64_t i) {
  switch (code) {
    case 0: return __ldg(reinterpret_cast<const float*>(base) + i);
    case 1: return float(__ldg(reinterpret_cast<const uint8_t*>(base) + i));
    case 2: return float(__ldg(reinterpret_cast<const int32_t*>(base) + i));
    default: return float(__ldg(reinterpret_cast<const int64_t*>(base) + i));
  }
}
__device__ __forceinline__ int64_t mask_as_int(const void* base, int code, int64_t i) {
  switch (code) {
    case 0: return int64_t(__ldg(reinterpret_cast<const float*>(base) + i));
    case 1: return int64_t(__ldg(reinterpret_cast<const uint8_t*>(base) + i));
    case 2: return int64_t(__ldg(reinterpret_cast<const int32_t*>(base) + i));
    default: return __ldg(reinterpret_cast<const int64_t*>(base) + i);
  }
}

// Reward of token t of one row as PPOActor.compute_advantages builds it (actor.py:109-138), with
// the side outputs the reference stores back into the batch.  Returns (tot_reward, mask).
struct RowCtx {
  int64_t row_off, S, zero_col, add_col;
  float add_score;
};
struct TokOut {
  float klr, tot, old, ref;
};
// lm: rolled mask of token t, lp: its (rolled) log-prob, rf: its reference log-prob (unmasked)
__device__ __forceinline__ TokOut fused_math(const AdvFusedParams& f, const RowCtx& c, int64_t t, float lm,
                                             float lp, float rf) {
  TokOut o;
  o.old = __fmul_rn(lp, lm);                                                   // :116, :122
  o.ref = f.ref_logp ? __fmul_rn(rf, lm) : 0.f;                                // :120-121
  float kl = __fsub_rn(o.old, o.ref);                                          // data.py:1344
  if (f.kl_kind == 2) {
    kl = __fmul_rn(__fmul_rn(kl, kl), 0.5f);                                   // :1353
  } else if (f.kl_kind == 3) {
    const float lr = -kl;                                                      // :1359
    kl = __fsub_rn(__fsub_rn(expf(lr), 1.f), lr);                              // :1360
  }
  if (f.kl_clamp) kl = clamp_nan(kl, -10.f, 10.f);                             // :1363 (torch.clamp: NaN stays NaN)
  o.klr = __fmul_rn(f.neg_kl_ctl, kl);                                         // actor.py:128
  o.tot = o.klr;
  if (t == c.zero_col) o.tot = 0.f;                                            // :131
  if (t == c.add_col) o.tot = __fadd_rn(o.tot, c.add_score);                   // :133-138
  return o;
}
// scalar path: loads, math and the side outputs of one token; returns (tot_reward, mask)
__device__ __forceinline__ float2 fused_token(const AdvFusedParams& f, const RowCtx& c, int64_t t) {
  const int64_t tn = (t + 1 == c.S) ? 0 : t + 1;                               // roll(-1), :110
  const float lm = mask_as_float(f.loss_mask, f.loss_mask_code, c.row_off + tn);
  const float lp = __ldg(f.logp + c.row_off + (f.roll_logp ? tn : t));
  const float rf = f.ref_logp ? f.ref_logp[c.row_off + t] : 0.f;
  const TokOut o = fused_math(f, c, t, lm, lp, rf);
  if (f.ref_logp) f.ref_logp[c.row_off + t] = o.ref;                           // in place, :121
  f.kl_rewards[c.row_off + t] = o.klr;
  f.tot_rewards[c.row_off + t] = o.tot;
  f.loss_mask_out[c.row_off + t] = lm;
  f.old_logp_out[c.row_off + t] = o.old;
  return make_float2(o.tot, lm);
}

__device__ __forceinline__ void ld8(const float* p, float* o) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p));
  const float4 b = __ldg(reinterpret_cast<const float4*>(p) + 1);
  o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w;
  o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
}
__device__ __forceinline__ void st8(float* p, const float* v) {
  reinterpret_cast<float4*>(p)[0] = make_float4(v[0], v[1], v[2], v[3]);
  reinterpret_cast<float4*>(p)[1] = make_float4(v[4], v[5], v[6], v[7]);
}
__device__ __forceinline__ void sts8(uint32_t saddr, const float* v) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]) : "memory");
  asm volatile("st.shared.v4.f32 [%0+16], {%1, %2, %3, %4};" ::"r"(saddr), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]) : "memory");
}
// mask elements i .. i+8 as floats (i a multiple of 8 elements from an aligned base)
__device__ __forceinline__ void ld9_mask(const void* base, int code, int64_t i, float* o) {
  if (code == 0) {
    ld8(reinterpret_cast<const float*>(base) + i, o);
    o[8] = __ldg(reinterpret_cast<const float*>(base) + i + 8);
  } else if (code == 1) {
    const uint2 w = __ldg(reinterpret_cast<const uint2*>(reinterpret_cast<const uint8_t*>(base) + i));
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      o[k] = float((w.x >> (8 * k)) & 0xffu);
      o[4 + k] = float((w.y >> (8 * k)) & 0xffu);
    }
    o[8] = float(__ldg(reinterpret_cast<const uint8_t*>(base) + i + 8));
  } else if (code == 2) {
    const int32_t* q = reinterpret_cast<const int32_t*>(base) + i;
    const int4 a = __ldg(reinterpret_cast<const int4*>(q));
    const int4 b = __ldg(reinterpret_cast<const int4*>(q) + 1);
    const int32_t c = __ldg(q + 8);
    o[0] = float(a.x); o[1] = float(a.y); o[2] = float(a.z); o[3] = float(a.w);
    o[4] = float(b.x); o[5] = float(b.y); o[6] = float(b.z); o[7] = float(b.w);
    o[8] = float(c);
  } else {
    const int64_t* q = reinterpret_cast<const int64_t*>(base) + i;
    longlong2 a[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) a[k] = __ldg(reinterpret_cast<const longlong2*>(q) + k);
    const int64_t c = __ldg(q + 8);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      o[2 * k] = float(a[k].x);
      o[2 * k + 1] = float(a[k].y);
    }
    o[8] = float(c);
  }
}

// Partial sum of n mask elements starting at element `off`, for thread `tid` of kNT: the type is chosen OUTSIDE
// the loops, so that the eight loads of a round are independent and in flight together (a per-element switch
// serialises them: one DRAM round trip each); byte masks go sixteen to a load when the start is 16-byte aligned.
template <typename T, int kNT>
__device__ __forceinline__ long long mask_sum_typed(const T* q, int64_t n, int tid) {
  long long sum = 0;
  for (int64_t i = tid; i < n; i += int64_t(kNT) * 8) {
    T x[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int64_t j = i + int64_t(k) * kNT;
      x[k] = j < n ? __ldg(q + j) : T(0);
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) sum += static_cast<long long>(x[k]);
  }
  return sum;
}
template <int kNT>
__device__ __forceinline__ long long mask_partial_sum(const void* base, int code, int64_t off, int64_t n, int tid) {
  switch (code) {
    case 0: return mask_sum_typed<float, kNT>(reinterpret_cast<const float*>(base) + off, n, tid);
    case 1: {
      const uint8_t* q = reinterpret_cast<const uint8_t*>(base) + off;
      if ((reinterpret_cast<uintptr_t>(q) & 15) != 0) return mask_sum_typed<uint8_t, kNT>(q, n, tid);
      const int64_t nvec = n >> 4;
      unsigned int acc = 0;                                          // < 2^32 for any row this library takes
      for (int64_t i = tid; i < nvec; i += int64_t(kNT) * 4) {
        uint4 x[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int64_t j = i + int64_t(k) * kNT;
          x[k] = j < nvec ? __ldg(reinterpret_cast<const uint4*>(q) + j) : make_uint4(0u, 0u, 0u, 0u);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k)
          acc = __dp4a(x[k].x, 0x01010101u, __dp4a(x[k].y, 0x01010101u, __dp4a(x[k].z, 0x01010101u,
                       __dp4a(x[k].w, 0x01010101u, acc))));
      }
      long long sum = acc;
      const int64_t done = nvec << 4;
      if (done + tid < n) sum += __ldg(q + done + tid);             // < 16 elements left
      return sum;
    }
    case 2: return mask_sum_typed<int32_t, kNT>(reinterpret_cast<const int32_t*>(base) + off, n, tid);
    default: return mask_sum_typed<int64_t, kNT>(reinterpret_cast<const int64_t*>(base) + off, n, tid);
  }
}

template <bool kFused, int kT>
__global__ void __launch_bounds__(kT)
gae_padded_kernel(const GaePaddedParams p, const AdvFusedParams f) {
  constexpr int kThreads = kT, kWarps = kT / 32, kTile = kT * kItems;
  __shared__ Affine wsm[kWarps];
  __shared__ float nv_val[kWarps];
  __shared__ int nv_has[kWarps];
  __shared__ long long len_sm[kWarps];
  __shared__ long long resp_sm[kWarps];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const double gl = double(p.gl);
  for (int64_t row = blockIdx.x; row < p.B; row += gridDim.x) {
    const float* r = kFused ? nullptr : p.rewards + row * p.S;
    const float* v = p.values ? p.values + row * p.S : nullptr;  // null (fused only): zeros, :141
    const float* m = kFused ? nullptr : p.mask + row * p.S;
    float* adv = p.adv + row * p.S;
    float* ret = p.ret + row * p.S;
    const int64_t L = p.S - 1;                                  // positions taking part in the loop
    RowCtx ctx;
    bool no_eos;
    if (kFused) {
      // seqlens = attention_mask.sum(-1) (:126)
      long long part = mask_partial_sum<kThreads>(f.attn_mask, f.attn_code, row * p.S, p.S, tid);
      long long resp = f.overlong ? mask_partial_sum<kThreads>(f.loss_mask, f.loss_mask_code, row * p.S, p.S, tid) : 0;
#pragma unroll
      for (int d = 16; d; d >>= 1) {
        part += __shfl_xor_sync(0xffffffffu, part, d);
        resp += __shfl_xor_sync(0xffffffffu, resp, d);
      }
      if (lane == 0) {
        len_sm[warp] = part;
        resp_sm[warp] = resp;
      }
      __syncthreads();
      long long seqlen = 0, resp_len = 0;
      for (int w = 0; w < kWarps; ++w) {
        seqlen += len_sm[w];
        resp_len += resp_sm[w];
      }
      no_eos = seqlen == p.S;                                    // :127
      float sc = f.score[row];
      if (f.overlong) {
        // areal/utils/functional.py:389-394: min(-exceed / overlong_tokens * factor, 0) added to the score
        const float a = __fmul_rn(__fdiv_rn(__ll2float_rn(-(resp_len - f.expected_len)),
                                            __ll2float_rn(f.overlong_tokens)), f.overlong_factor);
        if (!(0.f < a)) sc = __fadd_rn(sc, a);
        __syncthreads();                                         // every thread has read score[row]
        if (tid == 0) f.score[row] = sc;
      }
      if (f.apply_scaling) {
        sc = __fmul_rn(__fadd_rn(sc, f.bias), f.scaling);        // :98
        sc = clamp_nan(sc, -f.clip, f.clip);                     // :99-101
      }
      ctx.row_off = row * p.S;
      ctx.S = p.S;
      ctx.zero_col = seqlen - 1 < 0 ? seqlen - 1 + p.S : seqlen - 1;   // negative index wraps (:131)
      ctx.add_col = seqlen - 2 < 0 ? 0 : seqlen - 2;             // :132
      ctx.add_score = (f.mask_no_eos_with_zero && no_eos) ? 0.f : sc;
      if (ctx.zero_col >= p.S) ctx.zero_col = -1;                // out-of-range sums: the reference raises
      if (ctx.add_col >= p.S) ctx.add_col = -1;
    } else {
      no_eos = p.no_eos[row] != 0;
    }
    const float v_last = v ? v[p.S - 1] : 0.f;
    float nv_carry = no_eos ? v_last : 0.f;                     // actor.py:135 (bool * value)
    double a_carry = 0.0;                                       // adv[:, S-1] = 0
    if (tid == 0) {
      if (kFused) fused_token(f, ctx, p.S - 1);
      adv[p.S - 1] = 0.f;
      ret[p.S - 1] = v_last;                                    // returns = advantages + values, :152
    }
    const int64_t ntiles = (L + kTile - 1) / kTile;
    for (int64_t tile = ntiles - 1; tile >= 0; --tile) {
      const int64_t t0 = tile * kTile + int64_t(tid) * kItems;
      int nvalid = 0;
      if (t0 < L) nvalid = (L - t0) < kItems ? int(L - t0) : kItems;
      float rr[kItems], vv[kItems];
      unsigned int valid_bits = 0;
      const bool vec = p.vec_ok && nvalid == kItems;               // whole 8-token piece, 16-byte aligned
      if (vec) {
        if (kFused) {
          const int64_t g0 = ctx.row_off + t0;
          float m9[kItems + 1], lp9[kItems + 1], rf[kItems], klr[kItems], old[kItems];
          ld9_mask(f.loss_mask, f.loss_mask_code, g0, m9);         // t0+8 <= S-1 here: no wrap
          ld8(f.logp + g0, lp9);
          lp9[kItems] = __ldg(f.logp + g0 + kItems);
          if (f.ref_logp) ld8(f.ref_logp + g0, rf);
          if (v) ld8(v + t0, vv);
#pragma unroll
          for (int k = 0; k < kItems; ++k) {
            if (!v) vv[k] = 0.f;
            const TokOut o = fused_math(f, ctx, t0 + k, m9[k + 1], f.roll_logp ? lp9[k + 1] : lp9[k],
                                        f.ref_logp ? rf[k] : 0.f);
            rr[k] = o.tot;
            klr[k] = o.klr;
            old[k] = o.old;
            rf[k] = o.ref;
            if (m9[k + 1] != 0.f) valid_bits |= 1u << k;
          }
          if (f.ref_logp) st8(f.ref_logp + g0, rf);
          st8(f.kl_rewards + g0, klr);
          st8(f.tot_rewards + g0, rr);
          st8(f.loss_mask_out + g0, m9 + 1);
          st8(f.old_logp_out + g0, old);
        } else {
          float mm[kItems];
          ld8(r + t0, rr);
          ld8(v + t0, vv);
          ld8(m + t0, mm);
#pragma unroll
          for (int k = 0; k < kItems; ++k)
            if (mm[k] != 0.f) valid_bits |= 1u << k;
        }
      } else {
#pragma unroll
        for (int k = 0; k < kItems; ++k) {
          if (k < nvalid) {
            if (kFused) {
              const float2 tm = fused_token(f, ctx, t0 + k);
              rr[k] = tm.x;
              vv[k] = v ? __ldg(v + t0 + k) : 0.f;
              if (tm.y != 0.f) valid_bits |= 1u << k;
            } else {
              rr[k] = __ldg(r + t0 + k);
              vv[k] = __ldg(v + t0 + k);
              if (__ldg(m + t0 + k) != 0.f) valid_bits |= 1u << k;
            }
          }
        }
      }
      // nearest valid value to the right of this thread
      const int first = valid_bits ? (__ffs(valid_bits) - 1) : -1;
      float first_val = 0.f;
#pragma unroll
      for (int k = 0; k < kItems; ++k)
        if (k == first) first_val = vv[k];
      const unsigned int ball = __ballot_sync(0xffffffffu, first >= 0);
      const unsigned int right = (lane == 31) ? 0u : (ball >> (lane + 1));
      const int src = right ? (lane + __ffs(right)) : lane;
      const float wv = __shfl_sync(0xffffffffu, first_val, src);
      const float warp_first = __shfl_sync(0xffffffffu, first_val, ball ? (__ffs(ball) - 1) : 0);
      if (lane == 0) {
        nv_has[warp] = ball != 0u;
        nv_val[warp] = warp_first;
      }
      __syncthreads();                                           // barrier #1
      float nv_in;                                               // value entering this thread
      if (right) {
        nv_in = wv;
      } else {
        nv_in = nv_carry;
        for (int w = kWarps - 1; w > warp; --w)
          if (nv_has[w]) nv_in = nv_val[w];
      }
      float nv_tile = nv_carry;                                  // carry for the next (left) tile
      for (int w = kWarps - 1; w >= 0; --w)
        if (nv_has[w]) nv_tile = nv_val[w];

      float delta[kItems];
      Affine agg = ident();
      {
        float nv = nv_in;
#pragma unroll
        for (int k = kItems - 1; k >= 0; --k) {
          if (k < nvalid && ((valid_bits >> k) & 1u)) {
            delta[k] = __fsub_rn(__fadd_rn(rr[k], __fmul_rn(p.g, nv)), vv[k]);   // actor.py:137
            agg.b = fma(gl, agg.b, double(delta[k]));                             // :138, :143
            agg.a = gl * agg.a;
            nv = vv[k];                                                           // :142
          }
        }
      }
      Affine excl, tagg;
      block_suffix_scan<kWarps>(agg, wsm, warp, lane, excl, tagg);       // barrier #2 inside
      double A = fma(excl.a, a_carry, excl.b);
#pragma unroll
      for (int k = kItems - 1; k >= 0; --k) {
        if (k < nvalid) {
          if ((valid_bits >> k) & 1u) A = fma(gl, A, double(delta[k]));
          const float af = float(A);
          if (vec) {
            rr[k] = af;                                            // rewards are dead: reuse as staging
            vv[k] = __fadd_rn(af, vv[k]);
          } else {
            adv[t0 + k] = af;
            ret[t0 + k] = __fadd_rn(af, vv[k]);
          }
        }
      }
      if (vec) {
        st8(adv + t0, rr);
        st8(ret + t0, vv);
      }
      a_carry = fma(tagg.a, a_carry, tagg.b);
      nv_carry = nv_tile;
    }
    __syncthreads();
  }
}

// ---- padded masked GAE, few LONG rows: one thread-block cluster per row ----------------------------------
// gae_padded_kernel gives a row to ONE CTA, which leaves most SMs idle for a batch like [16, 32768] (long
// chains of thought).  Here a row is cut into C <= 8 tiles of 4096 tokens, one CTA each, the C CTAs forming a
// cluster.  Two things cross tile boundaries, both right-to-left, both through distributed shared memory
// (st.async + mbarrier complete_tx, no global memory, no flags): the "next valid value" (value of the nearest
// token with mask != 0 to the right, actor.py:135,142) and the tile's affine map for the advantage recurrence.
// Same arithmetic as gae_padded_kernel<false, 512>: fp32 deltas in the reference's operation order, fp64 scan.
__device__ __forceinline__ void st_async_b64(uint32_t remote_addr, unsigned long long v, uint32_t remote_bar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b64 [%0], %1, [%2];" ::"r"(remote_addr),
               "l"(v), "r"(remote_bar)
               : "memory");
}

// kFused: the reward construction of PPOActor.compute_advantages in the same launch (what gae_padded_kernel<true>
// does for a row in one CTA).  One more thing crosses tile boundaries then, in BOTH directions and before anything
// else: the row's sequence length (attention-mask sum, actor.py:126) and response length (loss-mask sum, overlong
// penalty), which place the task reward and the zeroed column.  It travels together with the "next valid value"
// summary (which needs only the mask and the values) in ONE all-to-all of 24 bytes per pair of tiles, so the fused
// kernel has two exchanges like the plain one.  The tile's loads are issued before that exchange is waited for; its
// five side outputs are staged in shared memory and leave through bulk copies (direct stores sat in the SM's store
// queue ahead of the exchange messages: 2.6-5 us of the critical path, measured).
struct NvLocal {
  unsigned int right;   // lanes to the right of this one (in the warp) that hold a valid token
  float wv;             // value of the nearest of them
};
// nearest valid value to the right of each thread INSIDE the tile: warp part (ballot + shuffle), warp summaries to smem
__device__ __forceinline__ NvLocal nv_local(unsigned int valid_bits, const float (&vv)[kItems], int warp, int lane,
                                            int* nv_has, float* nv_val) {
  const int first = valid_bits ? (__ffs(valid_bits) - 1) : -1;
  float first_val = 0.f;
#pragma unroll
  for (int k = 0; k < kItems; ++k)
    if (k == first) first_val = vv[k];
  const unsigned int ball = __ballot_sync(0xffffffffu, first >= 0);
  NvLocal o;
  o.right = (lane == 31) ? 0u : (ball >> (lane + 1));
  const int src = o.right ? (lane + __ffs(o.right)) : lane;
  o.wv = __shfl_sync(0xffffffffu, first_val, src);
  const float warp_first = __shfl_sync(0xffffffffu, first_val, ball ? (__ffs(ball) - 1) : 0);
  if (lane == 0) {
    nv_has[warp] = ball != 0u;
    nv_val[warp] = warp_first;
  }
  return o;
}
// (has a valid token << 32) | bits of the value of the tile's LEFTMOST valid token
template <int kW>
__device__ __forceinline__ unsigned long long nv_summary(const int* nv_has, const float* nv_val) {
  int has = 0;
  float val = 0.f;
  for (int w = kW - 1; w >= 0; --w)
    if (nv_has[w]) {
      has = 1;
      val = nv_val[w];
    }
  return (static_cast<unsigned long long>(has) << 32) | __float_as_uint(val);
}
template <int kW>
__device__ __forceinline__ float nv_entering(const NvLocal& l, float carry, int warp, const int* nv_has,
                                             const float* nv_val) {
  if (l.right) return l.wv;
  float nv_in = carry;
  for (int w = kW - 1; w > warp; --w)
    if (nv_has[w]) nv_in = nv_val[w];
  return nv_in;
}

// kCommon: the usual configuration -- both masks one byte per token (torch.bool / uint8), the k1 KL estimator, no
// overlong penalty.  The dtype switches of the mask readers, the other estimators and the penalty fold away, and with
// them a third to a half of the instance's code (the capture of the generic instance shows `no_instruction` --
// instruction fetch -- among its top stalls: every CTA runs this long kernel exactly once).
template <bool kFused, bool kCommon = false>
__global__ void __launch_bounds__(512, 2)
gae_padded_cluster_kernel(const GaePaddedParams p, const AdvFusedParams f_in, int C) {
  AdvFusedParams f = f_in;
  if (kCommon) {
    f.attn_code = 1;
    f.loss_mask_code = 1;
    f.kl_kind = 1;
    f.overlong = 0;
  }
  constexpr int kT = 512, kW = kT / 32, kTileC = kT * kItems;      // 4096 tokens per CTA
  __shared__ Affine wsm[kW];
  __shared__ float nv_val[kW];
  __shared__ int nv_has[kW];
  __shared__ __align__(16) Affine inbox_map[kGaeClusterMax];         // [sender rank]
  __shared__ __align__(8) unsigned long long inbox_nv[kGaeClusterMax];   // (has << 32) | value bits
  __shared__ __align__(16) long long inbox_len[kFused ? kGaeClusterMax : 1][2];   // (attention sum, loss-mask sum)
  __shared__ __align__(8) uint64_t bar_nv, bar_map;
  __shared__ long long len_sm[kFused ? kW : 1], resp_sm[kFused ? kW : 1];
  __shared__ RowCtx ctx_sm;
  __shared__ float nv_carry_sm;
  __shared__ double a_carry_sm;
  extern __shared__ __align__(128) uint8_t stage_raw[];               // kFused: 5 x 4096 floats (side outputs)
  constexpr uint32_t kStageArr = kTileC * 4;
  const uint32_t stage_base = smem_u32(stage_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t rank = gae_cluster_ctarank();                       // rank 0 = RIGHTMOST tile of the row
  const int64_t row = blockIdx.x / C;
  const int64_t tile = C - 1 - int64_t(rank);
  if (tid == 0) {
    mbar_init(smem_u32(&bar_nv), 1);
    mbar_init(smem_u32(&bar_map), 1);
    mbar_fence_init();
  }
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");

  const float* r = kFused ? nullptr : p.rewards + row * p.S;
  const float* v = p.values ? p.values + row * p.S : nullptr;        // null (fused only): zeros, actor.py:141
  const float* m = kFused ? nullptr : p.mask + row * p.S;
  float* adv = p.adv + row * p.S;
  float* ret = p.ret + row * p.S;
  const int64_t L = p.S - 1;
  const double gl = double(p.gl);
  const int64_t t0 = tile * kTileC + int64_t(tid) * kItems;
  int nvalid = 0;
  if (t0 < L) nvalid = (L - t0) < kItems ? int(L - t0) : kItems;
  float rr[kItems], vv[kItems];
  unsigned int valid_bits = 0;
  const bool vec = p.vec_ok && nvalid == kItems;
  const float v_last = v ? v[p.S - 1] : 0.f;
  float nv_in;
  if constexpr (kFused) {
    const int64_t g0 = row * p.S + t0;
    const float sc_early = __ldcv(f.score + row);                    // used by thread 0 only; in flight with the masks
    // partial sums of the two masks over this tile's columns (the rightmost tile takes column S-1 too)
    {
      const int64_t c0 = tile * kTileC;
      const int64_t cn = rank == 0 ? p.S - c0 : int64_t(kTileC);
      long long part = mask_partial_sum<kT>(f.attn_mask, f.attn_code, row * p.S + c0, cn, tid);
      long long resp = f.overlong ? mask_partial_sum<kT>(f.loss_mask, f.loss_mask_code, row * p.S + c0, cn, tid) : 0;
#pragma unroll
      for (int d = 16; d; d >>= 1) {
        part += __shfl_xor_sync(0xffffffffu, part, d);
        resp += __shfl_xor_sync(0xffffffffu, resp, d);
      }
      if (lane == 0) {
        len_sm[warp] = part;
        resp_sm[warp] = resp;
      }
    }
    // the tile's inputs, in flight while the lengths travel; the mask and the values are needed first
    float m9[kItems + 1], lp9[kItems + 1], rf[kItems];
    if (vec) {
      ld9_mask(f.loss_mask, f.loss_mask_code, g0, m9);               // t0+8 <= S-1 here: no wrap
      if (v) ld8(v + t0, vv);
      ld8(f.logp + g0, lp9);
      lp9[kItems] = __ldg(f.logp + g0 + kItems);
      if (f.ref_logp) ld8(f.ref_logp + g0, rf);
#pragma unroll
      for (int k = 0; k < kItems; ++k) {
        if (!v) vv[k] = 0.f;
        if (m9[k + 1] != 0.f) valid_bits |= 1u << k;
      }
    } else {
#pragma unroll
      for (int k = 0; k < kItems; ++k) {
        if (k < nvalid) {                                            // t0+k+1 <= S-1: no wrap
          vv[k] = v ? __ldg(v + t0 + k) : 0.f;
          if (mask_as_float(f.loss_mask, f.loss_mask_code, g0 + k + 1) != 0.f) valid_bits |= 1u << k;
        }
      }
    }
    // column S-1 (side outputs only; its mask and log-prob wrap to column 0): loaded now, finished after the exchange
    float last_lm = 0.f, last_lp = 0.f, last_rf = 0.f;
    const bool do_last = rank == 0 && tid == kT - 1;
    if (do_last) {
      last_lm = mask_as_float(f.loss_mask, f.loss_mask_code, row * p.S);
      last_lp = __ldg(f.logp + row * p.S + (f.roll_logp ? 0 : p.S - 1));
      if (f.ref_logp) last_rf = f.ref_logp[row * p.S + p.S - 1];
    }
    const NvLocal nvl = nv_local(valid_bits, vv, warp, lane, nv_has, nv_val);
    __syncthreads();
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");   // every CTA's barriers are initialised
    // ---- exchange 1 (all-to-all): (attention sum, loss-mask sum, next-valid summary) of every tile to every other
    if (tid == 0) {
      long long seqlen = 0, resp_len = 0;
      for (int w = 0; w < kW; ++w) {
        seqlen += len_sm[w];
        resp_len += resp_sm[w];
      }
      const unsigned long long nvmsg = nv_summary<kW>(nv_has, nv_val);
      // score[row] is read BEFORE this CTA's message leaves (the empty asm makes the message depend on the loaded
      // value), and rewritten (penalty, in place) by rank 0 only after it has received every CTA's
      float sc = sc_early;
      asm volatile("" : "+l"(seqlen) : "f"(sc));
      for (uint32_t dst = 0; dst < uint32_t(C); ++dst)
        if (dst != rank) {
          const uint32_t rb = gae_mapa(smem_u32(&bar_nv), dst);
          st_async_affine(gae_mapa(smem_u32(&inbox_len[rank][0]), dst), __longlong_as_double(seqlen),
                          __longlong_as_double(resp_len), rb);
          st_async_b64(gae_mapa(smem_u32(&inbox_nv[rank]), dst), nvmsg, rb);
        }
      mbar_arrive_expect_tx(smem_u32(&bar_nv), 24u * uint32_t(C - 1));
      mbar_wait_acq_cluster(smem_u32(&bar_nv), 0);
      for (uint32_t k = 0; k < uint32_t(C); ++k)
        if (k != rank) {
          seqlen += inbox_len[k][0];
          resp_len += inbox_len[k][1];
        }
      if (f.overlong) {
        // areal/utils/functional.py:389-394: min(-exceed / overlong_tokens * factor, 0) added to the score
        const float a = __fmul_rn(__fdiv_rn(__ll2float_rn(-(resp_len - f.expected_len)),
                                            __ll2float_rn(f.overlong_tokens)), f.overlong_factor);
        if (!(0.f < a)) sc = __fadd_rn(sc, a);
        if (rank == 0) f.score[row] = sc;
      }
      if (f.apply_scaling) {
        sc = __fmul_rn(__fadd_rn(sc, f.bias), f.scaling);            // actor.py:98
        sc = clamp_nan(sc, -f.clip, f.clip);                         // :99-101
      }
      RowCtx c;
      c.row_off = row * p.S;
      c.S = p.S;
      c.zero_col = seqlen - 1 < 0 ? seqlen - 1 + p.S : seqlen - 1;   // negative index wraps (:131)
      c.add_col = seqlen - 2 < 0 ? 0 : seqlen - 2;                   // :132
      const bool no_eos = seqlen == p.S;                             // :127
      c.add_score = (f.mask_no_eos_with_zero && no_eos) ? 0.f : sc;
      if (c.zero_col >= p.S) c.zero_col = -1;                        // out-of-range sums: the reference raises
      if (c.add_col >= p.S) c.add_col = -1;
      ctx_sm = c;
      float carry = no_eos ? v_last : 0.f;                           // actor.py:135 (bool * value)
      for (uint32_t k = 0; k < rank; ++k)                            // tiles to the right: the smaller k the further
        if (inbox_nv[k] >> 32) carry = __uint_as_float(static_cast<unsigned int>(inbox_nv[k]));
      nv_carry_sm = carry;
    }
    __syncthreads();
    const RowCtx ctx = ctx_sm;
    nv_in = nv_entering<kW>(nvl, nv_carry_sm, warp, nv_has, nv_val);
    if (vec) {
      float klr[kItems], old[kItems];
#pragma unroll
      for (int k = 0; k < kItems; ++k) {
        const TokOut o = fused_math(f, ctx, t0 + k, m9[k + 1], f.roll_logp ? lp9[k + 1] : lp9[k],
                                    f.ref_logp ? rf[k] : 0.f);
        rr[k] = o.tot;
        klr[k] = o.klr;
        old[k] = o.old;
        rf[k] = o.ref;
      }
      const uint32_t so = stage_base + uint32_t(tid) * 32u;
      if (f.ref_logp) sts8(so, rf);
      sts8(so + kStageArr, klr);
      sts8(so + 2 * kStageArr, rr);
      sts8(so + 3 * kStageArr, m9 + 1);
      sts8(so + 4 * kStageArr, old);
      fence_proxy_async();                                           // staged side outputs -> visible to the bulk copies
    } else {
#pragma unroll
      for (int k = 0; k < kItems; ++k)
        if (k < nvalid) rr[k] = fused_token(f, ctx, t0 + k).x;       // (loads, math and direct stores: ragged ends only)
    }
    if (do_last) {                                                   // what fused_token does, loads already here
      const int64_t g = row * p.S + p.S - 1;
      const TokOut o = fused_math(f, ctx, p.S - 1, last_lm, last_lp, last_rf);
      if (f.ref_logp) f.ref_logp[g] = o.ref;
      f.kl_rewards[g] = o.klr;
      f.tot_rewards[g] = o.tot;
      f.loss_mask_out[g] = last_lm;
      f.old_logp_out[g] = o.old;
    }
  } else {
    if (vec) {
      float mm[kItems];
      ld8(r + t0, rr);
      ld8(v + t0, vv);
      ld8(m + t0, mm);
#pragma unroll
      for (int k = 0; k < kItems; ++k)
        if (mm[k] != 0.f) valid_bits |= 1u << k;
    } else {
#pragma unroll
      for (int k = 0; k < kItems; ++k) {
        if (k < nvalid) {
          rr[k] = __ldg(r + t0 + k);
          vv[k] = __ldg(v + t0 + k);
          if (__ldg(m + t0 + k) != 0.f) valid_bits |= 1u << k;
        }
      }
    }
    const NvLocal nvl = nv_local(valid_bits, vv, warp, lane, nv_has, nv_val);
    __syncthreads();
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");   // every CTA's barriers are initialised
    // ---- exchange 1: (has a valid token, value of the tile's LEFTMOST valid token) to every tile on the left
    if (tid == 0) {
      const unsigned long long msg = nv_summary<kW>(nv_has, nv_val);
      for (uint32_t dst = rank + 1; dst < uint32_t(C); ++dst)
        st_async_b64(gae_mapa(smem_u32(&inbox_nv[rank]), dst), msg, gae_mapa(smem_u32(&bar_nv), dst));
      float carry = p.no_eos[row] != 0 ? v_last : 0.f;               // actor.py:135 (bool * value)
      if (rank > 0) {
        mbar_arrive_expect_tx(smem_u32(&bar_nv), 8u * rank);
        mbar_wait_acq_cluster(smem_u32(&bar_nv), 0);
        for (uint32_t k = 0; k < rank; ++k)                          // rank k < rank: further right the smaller k
          if (inbox_nv[k] >> 32) carry = __uint_as_float(static_cast<unsigned int>(inbox_nv[k]));
      }
      nv_carry_sm = carry;
    }
    __syncthreads();
    nv_in = nv_entering<kW>(nvl, nv_carry_sm, warp, nv_has, nv_val);
  }
  if (rank == 0 && tid == 0) {
    adv[p.S - 1] = 0.f;                                              // actor.py:148
    ret[p.S - 1] = v_last;                                           // returns = advantages + values, :152
  }
  float delta[kItems];
  Affine agg = ident();
  {
    float nv = nv_in;
#pragma unroll
    for (int k = kItems - 1; k >= 0; --k) {
      if (k < nvalid && ((valid_bits >> k) & 1u)) {
        delta[k] = __fsub_rn(__fadd_rn(rr[k], __fmul_rn(p.g, nv)), vv[k]);   // actor.py:137
        agg.b = fma(gl, agg.b, double(delta[k]));                             // :138, :143
        agg.a = gl * agg.a;
        nv = vv[k];                                                           // :142
      }
    }
  }
  Affine excl, tagg;
  block_suffix_scan<kW>(agg, wsm, warp, lane, excl, tagg);           // one __syncthreads() inside
  if constexpr (kFused) {
    // whole 8-token pieces of the tile: five bulk copies shared -> global (staged before the barrier above), issued
    // by one thread of ANOTHER warp than the one that runs the exchanges, waited for at the very end
    if (tid == 32 && p.vec_ok) {
      const int64_t left = L - tile * kTileC;
      const uint32_t bytes = uint32_t((left < kTileC ? left : int64_t(kTileC)) / kItems) * (kItems * 4u);
      const int64_t g = row * p.S + tile * kTileC;
      if (bytes) {
        if (f.ref_logp) tma_store_1d(f.ref_logp + g, stage_base, bytes);
        tma_store_1d(f.kl_rewards + g, stage_base + kStageArr, bytes);
        tma_store_1d(f.tot_rewards + g, stage_base + 2 * kStageArr, bytes);
        tma_store_1d(f.loss_mask_out + g, stage_base + 3 * kStageArr, bytes);
        tma_store_1d(f.old_logp_out + g, stage_base + 4 * kStageArr, bytes);
        tma_store_commit();
      }
    }
  }
  // ---- exchange 2: the tile's map to every tile on the left
  if (tid == 0) {
    for (uint32_t dst = rank + 1; dst < uint32_t(C); ++dst)
      st_async_affine(gae_mapa(smem_u32(&inbox_map[rank]), dst), tagg.a, tagg.b, gae_mapa(smem_u32(&bar_map), dst));
    double x = 0.0;                                                  // adv[:, S-1] = 0
    if (rank > 0) {
      mbar_arrive_expect_tx(smem_u32(&bar_map), 16u * rank);
      mbar_wait_acq_cluster(smem_u32(&bar_map), 0);
      for (uint32_t k = 0; k < rank; ++k) x = fma(inbox_map[k].a, x, inbox_map[k].b);
    }
    a_carry_sm = x;
  }
  __syncthreads();
  double A = fma(excl.a, a_carry_sm, excl.b);
#pragma unroll
  for (int k = kItems - 1; k >= 0; --k) {
    if (k < nvalid) {
      if ((valid_bits >> k) & 1u) A = fma(gl, A, double(delta[k]));
      const float af = float(A);
      if (vec) {
        rr[k] = af;
        vv[k] = __fadd_rn(af, vv[k]);
      } else {
        adv[t0 + k] = af;
        ret[t0 + k] = __fadd_rn(af, vv[k]);
      }
    }
  }
  if (vec) {
    st8(adv + t0, rr);
    st8(ret + t0, vv);
  }
  if constexpr (kFused) {
    if (tid == 32) tma_store_wait_read<0>();                         // shared memory stays until the copies have read it
  }
}

// ---- 2D with a segment table (gae_2d_olp semantics) --------------------------------------------------
__global__ void __launch_bounds__(kThreads)
gae2d_olp_kernel(const Gae2dParams p) {
  __shared__ Affine wsm[kWarps];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const double gl = double(p.gl);
  for (int64_t row = blockIdx.x; row < p.B; row += gridDim.x) {
    const float* r = p.rewards + row * p.H;
    const float* v = p.values + row * (p.H + 1);
    const uint8_t* done = p.done + row * (p.H + 1);
    float* adv = p.adv + row * p.H;
    float* ret = p.ret + row * p.H;
    const int64_t s0 = p.cu_num_dones[row], s1 = p.cu_num_dones[row + 1];
    const int32_t* idx = p.done_y_indices + s0;
    const int64_t nd = s1 - s0;
    double a_carry = 0.0;
    const int64_t ntiles = (p.H + kTile - 1) / kTile;
    for (int64_t tile = ntiles - 1; tile >= 0; --tile) {
      const int64_t t0 = tile * kTile + int64_t(tid) * kItems;
      int nvalid = 0;
      if (t0 < p.H) nvalid = (p.H - t0) < kItems ? int(p.H - t0) : kItems;
      float delta[kItems], vv[kItems];
      unsigned int tail_bits = 0, zero_bits = 0;
      if (nvalid > 0) {
        // first table entry with idx >= t0 + 1
        int64_t lo = 0, hi = nd;
        while (lo < hi) {
          const int64_t mid = (lo + hi) >> 1;
          if (int64_t(idx[mid]) >= t0 + 1) hi = mid; else lo = mid + 1;
        }
        int64_t q = lo;
#pragma unroll
        for (int k = 0; k < kItems; ++k) {
          if (k < nvalid) {
            const int64_t t = t0 + k;
            while (q < nd && int64_t(idx[q]) < t + 1) ++q;
            const bool at_done = (q < nd) && (int64_t(idx[q]) == t + 1);
            const bool tail = at_done || (t == p.H - 1);
            const float cur = __ldg(v + t);
            float nv = __ldg(v + t + 1);
            if (tail && done[t + 1]) nv = 0.f;                    // gae.cu:79-81
            vv[k] = cur;
            delta[k] = __fsub_rn(__fadd_rn(__ldg(r + t), __fmul_rn(p.g, nv)), cur);
            if (tail) tail_bits |= 1u << k;
            if (at_done && p.is_truncate[s0 + q]) zero_bits |= 1u << k;   // gae.cu:76-77, :85
          }
        }
      }
      Affine agg = ident();
#pragma unroll
      for (int k = kItems - 1; k >= 0; --k) {
        if (k < nvalid) {
          const double a = (tail_bits >> k) & 1u ? 0.0 : gl;
          const double b = (zero_bits >> k) & 1u ? 0.0 : double(delta[k]);
          agg.b = fma(a, agg.b, b);
          agg.a = a * agg.a;
        }
      }
      Affine excl, tagg;
      block_suffix_scan(agg, wsm, warp, lane, excl, tagg);
      double A = fma(excl.a, a_carry, excl.b);
#pragma unroll
      for (int k = kItems - 1; k >= 0; --k) {
        if (k < nvalid) {
          const double a = (tail_bits >> k) & 1u ? 0.0 : gl;
          const double b = (zero_bits >> k) & 1u ? 0.0 : double(delta[k]);
          A = fma(a, A, b);
          const float af = float(A);
          adv[t0 + k] = af;
          ret[t0 + k] = __fadd_rn(af, vv[k]);
        }
      }
      a_carry = fma(tagg.a, a_carry, tagg.b);
      __syncthreads();   // wsm is rewritten by the next tile
    }
  }
}

}  // namespace

int64_t gae1d_workspace_bytes(int64_t T) {
  const int64_t ntiles = (T + 512 - 1) / 512;        // the smallest tile any variant uses
  return int64_t(sizeof(WorkspaceHeader)) + (ntiles > 1 ? ntiles : 0) * int64_t(sizeof(TileDesc));
}

// The same table written from the sequences' side: one WARP per sequence stores the descriptors of the tiles
// whose first token lies inside it (none for an empty sequence) -- two loads of `cu` and a few coalesced
// stores instead of a 16-step dependent binary search per tile.  Used unless sequences are so long that a
// warp would loop over more than a few hundred tiles (then the per-tile search above is the cheaper one).
__global__ void gae1d_prep_by_sequence_kernel(const int32_t* __restrict__ cu, int B, int64_t ntiles,
                                              TileDesc* __restrict__ desc, int tile_tokens) {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const int64_t seq = (blockIdx.x * int64_t(blockDim.x) + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (seq >= B) return;
  const int64_t s = __ldg(cu + seq), e = __ldg(cu + seq + 1);
  int64_t t0 = (s + tile_tokens - 1) / tile_tokens;           // first tile that starts at or after s
  int64_t t1 = (e + tile_tokens - 1) / tile_tokens;           // first tile that starts at or after e
  // every descriptor gets initialised even if the caller's table does not start at 0 / end at T
  if (seq == 0) t0 = 0;
  if (seq == B - 1 || t1 > ntiles) t1 = ntiles;
  TileDesc d;
  d.inc = kEmpty;
  d.agg_a = kEmpty;
  d.agg_b = kEmpty;
  d.seq0 = static_cast<int>(seq);
  d.end0 = static_cast<int>(e);
  for (int64_t t = t0 + lane; t < t1; t += 32) desc[t] = d;
}

static void launch_gae1d_prep(const Gae1dParams& p, int64_t ntiles, TileDesc* desc, int tile_tokens, cudaStream_t stream) {
  // measured (A/B on one box): 4 M tokens 31.7 -> 27.7 us, 16 M tokens 80.0 -> 76.7 us per call; from 128 M tokens on
  // the B warps cost more to launch than the per-tile search saves (66.3 -> 65.8 %), hence the tile limit
  if (ntiles <= 65536 && ntiles <= 256 * p.B) {       // on average at most 8 rounds of 32 tiles per warp
    const int64_t threads = p.B * 32;
    gae1d_prep_by_sequence_kernel<<<static_cast<unsigned int>((threads + 255) / 256), 256, 0, stream>>>(
        p.cu, static_cast<int>(p.B), ntiles, desc, tile_tokens);
  } else {
    gae1d_prep_kernel<<<static_cast<unsigned int>((ntiles + 255) / 256), 256, 0, stream>>>(
        p.cu, static_cast<int>(p.B), p.T, ntiles, desc, tile_tokens);
  }
}

constexpr int64_t kWideFromTokens = 1 << 20;   // batches of at least this many tokens use 512-thread CTAs

// Batches of up to 8 x 4096 tokens: one cluster launch (no prep kernel, no descriptors, no workspace use).
static int launch_gae1d_cluster(const Gae1dParams& p, int64_t ntiles, cudaStream_t stream) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(static_cast<unsigned int>(ntiles));
  cfg.blockDim = dim3(512);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = static_cast<unsigned int>(ntiles);
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  AB_CUDA(cudaLaunchKernelEx(&cfg, gae1d_kernel<512, true>, p));
  return kOk;
}

int launch_gae1d(const Gae1dParams& p, cudaStream_t stream) {
  if (p.T == 0) return kOk;
  {
    static const bool use_cluster = [] {
      const char* e = getenv("AREAL_B200_GAE_CLUSTER");
      return !(e && e[0] == '0');
    }();
    const int64_t nt = (p.T + 4095) / 4096;
    if (use_cluster && p.T > kTile && nt <= kGaeClusterMax) return launch_gae1d_cluster(p, nt, stream);
  }
  const bool wide = p.T >= kWideFromTokens;
  const int threads = wide ? 64 : kThreads;
  const int tile_tokens = threads * kItems;
  const int64_t ntiles = (p.T + tile_tokens - 1) / tile_tokens;
  if (ntiles > 1) {
    // descriptors start "empty" and carry each tile's first sequence: one tiny kernel, stream-ordered
    TileDesc* desc = reinterpret_cast<TileDesc*>(reinterpret_cast<char*>(p.workspace) + sizeof(WorkspaceHeader));
    launch_gae1d_prep(p, ntiles, desc, tile_tokens, stream);
  }
  if (ntiles > 1) {
    // launched as a programmatic dependent of the prep kernel: its launch latency and the reward
    // loads overlap the prep grid
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(static_cast<unsigned int>(ntiles));
    cfg.blockDim = dim3(threads);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (wide) AB_CUDA(cudaLaunchKernelEx(&cfg, gae1d_kernel<64>, p));
    else AB_CUDA(cudaLaunchKernelEx(&cfg, gae1d_kernel<256>, p));
    return kOk;
  }
  gae1d_kernel<256><<<static_cast<unsigned int>(ntiles), kThreads, 0, stream>>>(p);
  return check_launch("gae1d");
}

// Allocation-free launch: same kernel, look-back descriptors in g_gae1d_static_scratch.
// Set-up happens on the first call per device (with that device current): the symbol is resolved to
// its address on THAT device and cached.  Every call then clears the part of the array it is about
// to use (every tile descriptor must start out "empty", and the array still holds whatever the
// previous call -- possibly a larger or an aborted one -- left there).
int launch_gae1d_static(Gae1dParams p, cudaStream_t stream) {
  static_assert(AB_GAE1D_STATIC_SCRATCH_BYTES ==
                    int64_t(sizeof(WorkspaceHeader)) + (AB_GAE1D_STATIC_MAX_TOKENS / kTile) * int64_t(sizeof(TileDesc)),
                "static scratch size must match the look-back layout");
  if (p.T == 0) return kOk;
  if (p.T > AB_GAE1D_STATIC_MAX_TOKENS) {
    set_error("gae1d (static scratch): %lld tokens exceed the built-in limit of %lld per call; use "
              "areal_b200_gae1d_nolp_misalign with a caller-provided workspace",
              (long long)p.T, (long long)AB_GAE1D_STATIC_MAX_TOKENS);
    return kInvalidArgument;
  }
  constexpr int kMaxDevices = 64;
  static std::mutex mu;
  static void* scratch_of_device[kMaxDevices] = {nullptr};
  int dev = -1;
  AB_CUDA(cudaGetDevice(&dev));
  AB_REQUIRE(dev >= 0 && dev < kMaxDevices, "device ordinal %d out of range", dev);
  void* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lock(mu);
    if (scratch_of_device[dev] == nullptr) {
      void* addr = nullptr;
      const cudaError_t e = cudaGetSymbolAddress(&addr, g_gae1d_static_scratch);
      if (e != cudaSuccess || addr == nullptr) {
        set_error("cudaGetSymbolAddress(g_gae1d_static_scratch) failed on device %d: %s", dev,
                  cudaGetErrorString(e));
        return kCudaError;
      }
      scratch_of_device[dev] = addr;
    }
    scratch = scratch_of_device[dev];
  }
  const int64_t ntiles = (p.T + kTile - 1) / kTile;
  if (ntiles > 1) {
    TileDesc* desc = reinterpret_cast<TileDesc*>(reinterpret_cast<char*>(scratch) + sizeof(WorkspaceHeader));
    launch_gae1d_prep(p, ntiles, desc, kTile, stream);
  }
  p.workspace = scratch;
  gae1d_kernel<256><<<static_cast<unsigned int>(ntiles), kThreads, 0, stream>>>(p);
  return check_launch("gae1d (static scratch)");
}

static bool aligned16(const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; }

// CTA width of the padded kernels (one CTA walks a row's tiles right to left; tile = 8 x width tokens).  Narrow CTAs keep
// more rows in flight per SM and stall fewer warps at the two barriers of a tile; wide CTAs need fewer serial tiles per
// row.  Measured on one box (GPU time, plain / fused): [4096, 4096] 256 threads 68.5 / 201 us, 128: 59.5 / 147, 64: 56.8 /
// 135; [2048, 8192] 68.2 / 178, 60.8 / 144, 54.4 / 160; [1024, 2048] 12.1, 9.9, 10.7; [512, 4096] 11.7 / 44.6, 12.8 / 44.9,
// 15.7 / 42.4; [256, 16384] 27.3 / 55, 35.8 / 68, 51.4 / 94.  Hence: many rows -> narrow, few rows -> wide.
// AREAL_B200_PADDED_THREADS = 64 | 128 | 256 overrides the choice (benchmarks).
static int padded_threads(bool fused, int64_t B, int64_t S) {
  static const int forced = [] {
    const char* e = getenv("AREAL_B200_PADDED_THREADS");
    const int t = e ? atoi(e) : 0;
    return (t == 64 || t == 128 || t == 256) ? t : 0;
  }();
  if (forced) return forced;
  const int64_t sms = num_sms();
  if (fused) {
    if (B >= 10 * sms && S <= 4096) return 64;
    return B >= 4 * sms ? 128 : 256;
  }
  if (B >= 10 * sms) return 64;
  return B >= 5 * sms ? 128 : 256;
}

template <bool kFused>
static void launch_padded_t(const GaePaddedParams& p, const AdvFusedParams& f, cudaStream_t stream) {
  const int t = padded_threads(kFused, p.B, p.S);
  const int64_t cap = int64_t(num_sms()) * (2048 / t);
  const unsigned int grid = static_cast<unsigned int>(p.B < cap ? p.B : cap);
  if (t == 64) gae_padded_kernel<kFused, 64><<<grid, 64, 0, stream>>>(p, f);
  else if (t == 128) gae_padded_kernel<kFused, 128><<<grid, 128, 0, stream>>>(p, f);
  else gae_padded_kernel<kFused, 256><<<grid, 256, 0, stream>>>(p, f);
}

// few long rows: one cluster of C CTAs per row (4096 tokens each) instead of one CTA per row
template <bool kFused>
static bool launch_padded_cluster(const GaePaddedParams& p, const AdvFusedParams& f, cudaStream_t stream, int* status) {
  static const bool use_cluster = [] {
    const char* e = getenv("AREAL_B200_PADDED_CLUSTER");
    return !(e && e[0] == '0');
  }();
  const int64_t C = (p.S - 1 + 4095) / 4096;
  if (!(use_cluster && C >= 2 && C <= kGaeClusterMax && p.B * C <= 2 * int64_t(num_sms()))) return false;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(static_cast<unsigned int>(p.B * C));
  cfg.blockDim = dim3(512);
  cfg.stream = stream;
  if (kFused) {
    constexpr size_t kStage = 5 * 4096 * sizeof(float);              // side outputs of a tile
    // the attribute is per device: one flag per device ordinal
    static bool done[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (!done[dev & 63]) {
      if (cudaFuncSetAttribute(gae_padded_cluster_kernel<kFused, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               int(kStage)) != cudaSuccess ||
          cudaFuncSetAttribute(gae_padded_cluster_kernel<kFused, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               int(kStage)) != cudaSuccess) {
        (void)cudaGetLastError();
        return false;                                                // (one CTA per row instead)
      }
      done[dev & 63] = true;
    }
    cfg.dynamicSmemBytes = kStage;
  }
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = static_cast<unsigned int>(C);
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  static const bool debug = getenv("AREAL_B200_GAE_DEBUG") != nullptr;
  if (debug) {
    int n = -1;
    const cudaError_t q = cudaOccupancyMaxActiveClusters(&n, gae_padded_cluster_kernel<kFused, false>, &cfg);
    fprintf(stderr, "[areal_b200] gae_padded_cluster<%d> B=%lld C=%lld: max active clusters %d (%s)\n", int(kFused),
            static_cast<long long>(p.B), static_cast<long long>(C), n, cudaGetErrorString(q));
  }
  const bool common = kFused && f.attn_code == 1 && f.loss_mask_code == 1 && f.kl_kind == 1 && !f.overlong;
  const cudaError_t e = common ? cudaLaunchKernelEx(&cfg, gae_padded_cluster_kernel<kFused, true>, p, f, int(C))
                                   : cudaLaunchKernelEx(&cfg, gae_padded_cluster_kernel<kFused, false>, p, f, int(C));
  *status = e == cudaSuccess ? kOk : kCudaError;
  if (e != cudaSuccess) set_error("gae_padded_cluster launch failed: %s", cudaGetErrorString(e));
  return true;
}

int launch_gae_padded(const GaePaddedParams& p_in, cudaStream_t stream) {
  GaePaddedParams p = p_in;
  if (p.B == 0 || p.S == 0) return kOk;
  p.vec_ok = p.S % 4 == 0 && aligned16(p.rewards) && aligned16(p.values) && aligned16(p.mask) &&
             aligned16(p.adv) && aligned16(p.ret);
  int st = kOk;
  if (launch_padded_cluster<false>(p, AdvFusedParams{}, stream, &st)) return st;
  launch_padded_t<false>(p, AdvFusedParams{}, stream);
  return check_launch("gae_padded");
}

int launch_advantages_padded(const GaePaddedParams& p_in, const AdvFusedParams& f, cudaStream_t stream) {
  GaePaddedParams p = p_in;
  if (p.B == 0 || p.S == 0) return kOk;
  // S % 8: rows of 1-byte masks start on an 8-byte boundary too
  p.vec_ok = p.S % 8 == 0 && aligned16(p.values) && aligned16(p.adv) && aligned16(p.ret) &&
             aligned16(f.loss_mask) && aligned16(f.logp) && aligned16(f.ref_logp) &&
             aligned16(f.kl_rewards) && aligned16(f.tot_rewards) && aligned16(f.loss_mask_out) &&
             aligned16(f.old_logp_out);
  int st = kOk;
  if (launch_padded_cluster<true>(p, f, stream, &st)) return st;
  launch_padded_t<true>(p, f, stream);
  return check_launch("advantages_padded");
}

int launch_gae2d_olp(const Gae2dParams& p, cudaStream_t stream) {
  if (p.B == 0 || p.H == 0) return kOk;
  const int64_t cap = int64_t(num_sms()) * 8;
  gae2d_olp_kernel<<<static_cast<unsigned int>(p.B < cap ? p.B : cap), kThreads, 0, stream>>>(p);
  return check_launch("gae2d_olp");
}

}  // namespace ab
