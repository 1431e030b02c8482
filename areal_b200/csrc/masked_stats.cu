// Masked sum / count / min / max of up to 32 `[n]` float arrays in ONE launch, optionally folded into
// running totals (two keys of one launch must not share an output slot).
//
// Replaces (reference, relative to its root):
//   areal/utils/stats_tracker.py:210-289  _sum_of / _avg_of / _min_of / _max_of: per key and per recorded
//       micro-batch a `torch.where(d, v, fill)` + reduction (6 kernels per key for AVG_MIN_MAX) at
//       export time, over clones made at record time (:124, `value.detach().clone()`)
//   areal/engine/ppo/actor.py:352-391     the ~15 keys grpo_loss_fn records per micro-batch
//
// Layout: grid (chunks, keys); a CTA reduces 8192 elements of one key (coalesced 4-byte loads, fp64
// sum, exact integer count, fp32 min/max) to one partial in the workspace; the last CTA of a key
// (atomic ticket) folds that key's partials in chunk order, so the result does not depend on the
// schedule, and leaves the ticket at zero for the next call.
#include "masked_stats.h"

namespace ab {
namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;

struct Partial {
  double sum;
  double cnt;
  float mn, mx;
};

// torch.min / torch.max propagate NaN; fminf / fmaxf drop it
__device__ __forceinline__ float nan_min(float a, float b) { return (a != a) ? a : ((b != b) ? b : fminf(a, b)); }
__device__ __forceinline__ float nan_max(float a, float b) { return (a != a) ? a : ((b != b) ? b : fmaxf(a, b)); }
__device__ __forceinline__ Partial combine(const Partial& a, const Partial& b) {
  return Partial{a.sum + b.sum, a.cnt + b.cnt, nan_min(a.mn, b.mn), nan_max(a.mx, b.mx)};
}
__device__ __forceinline__ Partial shfl_down_partial(const Partial& p, int d) {
  Partial o;
  o.sum = __shfl_down_sync(0xffffffffu, p.sum, d);
  o.cnt = __shfl_down_sync(0xffffffffu, p.cnt, d);
  o.mn = __shfl_down_sync(0xffffffffu, p.mn, d);
  o.mx = __shfl_down_sync(0xffffffffu, p.mx, d);
  return o;
}

// workspace: unsigned int ticket[kMaxStatKeys] | pad to 256 B | Partial part[keys][chunks]
__global__ void __launch_bounds__(kThreads)
masked_stats_kernel(const __grid_constant__ StatsTable tab, int64_t n, int chunks, int accumulate,
                    unsigned int* __restrict__ ticket, Partial* __restrict__ parts) {
  __shared__ Partial wsm[kWarps];
  __shared__ bool last_sm;
  const int key = blockIdx.y, chunk = blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float* __restrict__ v = tab.val[key];
  const bool v_u8 = (tab.val_u8_bits >> key) & 1u;
  const uint8_t* __restrict__ m = tab.mask[key];
  const int64_t lo = int64_t(chunk) * kStatChunk;
  const int64_t hi = lo + kStatChunk < n ? lo + kStatChunk : n;
  Partial acc{0.0, 0.0, __int_as_float(0x7f800000), __int_as_float(0xff800000)};
  int cnt = 0;
#pragma unroll 8
  for (int64_t i = lo + tid; i < hi; i += kThreads) {
    const bool on = m ? (__ldg(m + i) != 0) : true;
    const float x = !v ? 1.f : (v_u8 ? float(__ldg(reinterpret_cast<const uint8_t*>(v) + i)) : __ldg(v + i));
    if (on) {                                    // torch.where(d, v, fill): masked-out NaNs never enter
      acc.sum += double(x);
      acc.mn = nan_min(acc.mn, x);
      acc.mx = nan_max(acc.mx, x);
      ++cnt;
    }
  }
  acc.cnt = double(cnt);
#pragma unroll
  for (int d = 16; d; d >>= 1) acc = combine(acc, shfl_down_partial(acc, d));
  if (lane == 0) wsm[warp] = acc;
  __syncthreads();
  if (tid == 0) {
    Partial t = wsm[0];
    for (int w = 1; w < kWarps; ++w) t = combine(t, wsm[w]);
    parts[int64_t(key) * chunks + chunk] = t;
    __threadfence();
    last_sm = atomicAdd(ticket + key, 1u) == static_cast<unsigned int>(chunks - 1);
  }
  __syncthreads();
  if (!last_sm) return;
  __threadfence();
  if (warp == 0) {
    // chunk order: lane l takes chunks l, l+32, ... in order, then a fixed shuffle tree
    Partial t{0.0, 0.0, __int_as_float(0x7f800000), __int_as_float(0xff800000)};
    const volatile Partial* p = parts + int64_t(key) * chunks;
    for (int c = lane; c < chunks; c += 32) {
      Partial q;
      q.sum = p[c].sum;
      q.cnt = p[c].cnt;
      q.mn = p[c].mn;
      q.mx = p[c].mx;
      t = combine(t, q);
    }
#pragma unroll
    for (int d = 16; d; d >>= 1) t = combine(t, shfl_down_partial(t, d));
    if (lane == 0) {
      double* o = tab.out[key];
      if (accumulate) {                          // running totals over micro-batches (stream-ordered)
        o[0] += t.sum;
        o[1] += t.cnt;
        const double mn = double(t.mn), mx = double(t.mx);
        o[2] = (mn != mn || o[2] != o[2]) ? (mn != mn ? mn : o[2]) : fmin(o[2], mn);
        o[3] = (mx != mx || o[3] != o[3]) ? (mx != mx ? mx : o[3]) : fmax(o[3], mx);
      } else {
        o[0] = t.sum;
        o[1] = t.cnt;
        o[2] = double(t.mn);
        o[3] = double(t.mx);
      }
      ticket[key] = 0u;
    }
  }
}

// Per-sequence masked mean of a packed [T] vector: one warp per sequence (fp64 sum, lanes stride the
// sequence, fixed shuffle order).  mean[i] = sum(mask ? v : 0) / count, valid[i] = count > 0 (mean 0 otherwise).
__global__ void __launch_bounds__(kThreads)
segment_masked_mean_kernel(const float* __restrict__ v, const uint8_t* __restrict__ mask, const int32_t* __restrict__ cu,
                           int64_t B, double* __restrict__ mean, uint8_t* __restrict__ valid) {
  const int lane = threadIdx.x & 31;
  const int64_t seq = (blockIdx.x * int64_t(kThreads) + threadIdx.x) >> 5;
  if (seq >= B) return;
  const int64_t s = cu[seq], e = cu[seq + 1];
  double sum = 0.0;
  int cnt = 0;
  for (int64_t i = s + lane; i < e; i += 32) {
    if (!mask || mask[i]) {
      sum += double(v[i]);
      ++cnt;
    }
  }
#pragma unroll
  for (int d = 16; d; d >>= 1) {
    sum += __shfl_down_sync(0xffffffffu, sum, d);
    cnt += __shfl_down_sync(0xffffffffu, cnt, d);
  }
  if (lane == 0) {
    mean[seq] = cnt ? sum / double(cnt) : 0.0;
    valid[seq] = cnt > 0;
  }
}

}  // namespace

int launch_segment_masked_mean(const float* v, const uint8_t* mask, const int32_t* cu, int64_t B, double* mean,
                               uint8_t* valid, cudaStream_t stream) {
  if (B == 0) return kOk;
  const int64_t blocks = (B * 32 + kThreads - 1) / kThreads;
  segment_masked_mean_kernel<<<static_cast<unsigned int>(blocks), kThreads, 0, stream>>>(v, mask, cu, B, mean, valid);
  return check_launch("segment_masked_mean");
}

int64_t masked_stats_workspace_bytes(int64_t n_keys, int64_t n) {
  const int64_t chunks = (n + kStatChunk - 1) / kStatChunk;
  return 256 + n_keys * (chunks > 0 ? chunks : 1) * int64_t(sizeof(Partial));
}

int launch_masked_stats(const StatsTable& tab, int n_keys, int64_t n, int accumulate, void* workspace,
                        cudaStream_t stream) {
  const int64_t chunks = (n + kStatChunk - 1) / kStatChunk;
  AB_REQUIRE(chunks <= 0x7fffffff, "masked_stats: too many elements");
  unsigned int* ticket = reinterpret_cast<unsigned int*>(workspace);
  Partial* parts = reinterpret_cast<Partial*>(reinterpret_cast<char*>(workspace) + 256);
  masked_stats_kernel<<<dim3(static_cast<unsigned int>(chunks), n_keys), kThreads, 0, stream>>>(
      tab, n, static_cast<int>(chunks), accumulate, ticket, parts);
  return check_launch("masked_stats");
}

}  // namespace ab
