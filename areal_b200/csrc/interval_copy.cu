// slice_intervals / set_intervals: gather / scatter of N half-open element ranges between a flat
// parameter buffer and a packed buffer whose offsets are the exclusive cumsum of the range sizes in
// the GIVEN order.
//
// Replaces (reference, relative to its root):
//   csrc/interval_op/interval_op.cu:12-104  copyDataKernel / set_intervals / slice_intervals
//                                           (one THREAD memcpy's up to 2048 elements -> uncoalesced;
//                                            offsets via ATen temporaries; legacy default stream)
//   realhf/impl/model/nn/flatten_param.py:65-90  _slice_intervals_py / _set_intervals_py
//
// Design (B200): (1) one small kernel scans the sizes into packed byte offsets (and validates the
// ranges against the buffer sizes); (2) the copy kernel cuts the PACKED byte space into equal
// contiguous chunks, one per warp, so the load is balanced whatever the interval sizes are; a warp
// binary-searches the interval holding its first byte once, then walks intervals linearly.  Every
// piece is moved with the widest vector the relative alignment of source and destination allows
// (16/8/4/2/1 B), 8 independent 128-bit loads in flight per lane, streaming cache hints; heads and
// tails are byte-granular, so any element size and any alignment is bit-exact.
#include <cstdlib>

#include "common.cuh"
#include "interval_copy.h"

namespace ab {
namespace {

constexpr int kScanThreads = 1024;
constexpr int kCopyThreads = 256;
constexpr int kCopyWarps = kCopyThreads / 32;

constexpr int64_t kShiftMinBytes = 256;   // shorter misaligned pieces keep the narrow-vector path

// workspace (int64 words): status[2] | off[N + 1] (byte offsets into the packed buffer) | any_shift |
//                          scan scratch: ticket, flags, total, pad | one look-back word per scan tile
// any_shift != 0: some interval of >= 256 bytes has flat and packed addresses that disagree modulo 16,
// i.e. the second (funnel-shift) copy kernel has work to do.
//
// The scan runs on ceil(N / 4096) CTAs chained by a decoupled look-back (one self-describing 8-byte word
// per tile: 2 flag bits + 62 value bits, so publishing needs no fence): a table of 200 k intervals takes a
// few microseconds instead of the ~200 us of a single CTA walking it in 49 dependent steps -- which was a
// fifth of the whole copy for the 7B tensor-parallel re-shard (many row-sized pieces).
constexpr int kScanPer = 4;
constexpr int kScanTile = kScanThreads * kScanPer;          // 4096 intervals per CTA
constexpr int kScratchWords = 4;                            // ticket, flags, total, pad
constexpr unsigned long long kFlagAgg = 1ull << 62, kFlagInc = 2ull << 62, kValMask = (1ull << 62) - 1;

__device__ __forceinline__ unsigned long long ld_relaxed(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__global__ void __launch_bounds__(kScanThreads)
interval_scan_kernel(const int64_t* __restrict__ iv, int64_t n, int64_t elem_size,
                     int64_t flat_numel, int64_t packed_numel, int64_t* __restrict__ ws,
                     uint64_t flat_base, uint64_t packed_base) {
  __shared__ int64_t warp_sums[32];
  __shared__ int64_t excl_sm;
  __shared__ int flags_sm;                  // bit 0: invalid interval, bit 1: a long misaligned piece
  int64_t* status = ws;
  int64_t* off = ws + 2;
  unsigned long long* scratch = reinterpret_cast<unsigned long long*>(ws + n + 4);
  unsigned long long* desc = scratch + kScratchWords;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int64_t tile = blockIdx.x, ntiles = gridDim.x;
  if (tid == 0) flags_sm = 0;
  __syncthreads();

  int64_t sz[kScanPer], st0[kScanPer];
  int64_t local = 0;
  bool bad = false;
#pragma unroll
  for (int k = 0; k < kScanPer; ++k) {
    const int64_t i = tile * kScanTile + int64_t(tid) * kScanPer + k;
    sz[k] = 0;
    st0[k] = 0;
    if (i < n) {
      const longlong2 se = reinterpret_cast<const longlong2*>(iv)[i];
      if (se.x < 0 || se.y < se.x || se.y > flat_numel) bad = true;
      else sz[k] = (se.y - se.x) * elem_size;
      st0[k] = se.x * elem_size;
    }
    local += sz[k];
  }
  if (bad) atomicOr(&flags_sm, 1);
  // block exclusive scan of `local`
  int64_t inc = local;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int64_t o = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += o;
  }
  if (lane == 31) warp_sums[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    int64_t w = warp_sums[lane];
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int64_t o = __shfl_up_sync(0xffffffffu, w, d);
      if (lane >= d) w += o;
    }
    warp_sums[lane] = w;   // inclusive over warps
    // ---- decoupled look-back by warp 0: lane j inspects tile (look - j) --------------------------------
    const unsigned long long agg = static_cast<unsigned long long>(__shfl_sync(0xffffffffu, w, 31)) & kValMask;
    if (lane == 0 && tile > 0) st_relaxed(&desc[tile], kFlagAgg | agg);
    int64_t excl = 0;
    for (int64_t look = tile - 1; look >= 0; look -= 32) {
      const int64_t idx = look - lane;
      unsigned long long word = kFlagInc;                    // before tile 0: inclusive prefix 0
      if (idx >= 0) {
        do {
          word = ld_relaxed(&desc[idx]);
        } while ((word >> 62) == 0);
      }
      const unsigned int inc_mask = __ballot_sync(0xffffffffu, (word >> 62) == 2);
      const int first = inc_mask ? __ffs(inc_mask) - 1 : 31;   // nearest tile with a finished prefix
      int64_t v = lane <= first ? static_cast<int64_t>(word & kValMask) : 0;
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
      excl += v;
      if (inc_mask) break;
    }
    if (lane == 0) {
      st_relaxed(&desc[tile], kFlagInc | ((static_cast<unsigned long long>(excl) + agg) & kValMask));
      excl_sm = excl;
    }
  }
  __syncthreads();
  int64_t run = excl_sm + (warp ? warp_sums[warp - 1] : 0) + (inc - local);
  bool shift = false;
#pragma unroll
  for (int k = 0; k < kScanPer; ++k) {
    const int64_t i = tile * kScanTile + int64_t(tid) * kScanPer + k;
    if (i < n) off[i] = run;
    if (sz[k] >= kShiftMinBytes && (((flat_base + uint64_t(st0[k])) - (packed_base + uint64_t(run))) & 15u) != 0)
      shift = true;
    run += sz[k];
  }
  if (shift) atomicOr(&flags_sm, 2);
  __syncthreads();
  if (tid == 0) {
    if (flags_sm) atomicOr(&scratch[1], static_cast<unsigned long long>(flags_sm));
    if (tile == ntiles - 1) st_relaxed(&scratch[2], static_cast<unsigned long long>(excl_sm + warp_sums[31]));
    __threadfence();
    const unsigned long long t = atomicAdd(&scratch[0], 1ull);
    if (t == static_cast<unsigned long long>(ntiles) - 1) {       // last CTA: everything is published
      __threadfence();
      const int64_t total = static_cast<int64_t>(ld_relaxed(&scratch[2]));
      const unsigned long long fl = ld_relaxed(&scratch[1]);
      off[n] = total;
      int64_t st = 0;
      if (fl & 1) st = 1;
      else if (total != packed_numel * elem_size) st = 2;
      status[0] = st;
      status[1] = total;
      off[n + 1] = (fl & 2) ? 1 : 0;
    }
  }
}

template <typename V>
__device__ __forceinline__ V ld_stream(const V* p) { return __ldg(p); }
template <>
__device__ __forceinline__ uint4 ld_stream<uint4>(const uint4* p) { return ldg_stream_v4(p); }
template <typename V>
__device__ __forceinline__ void st_stream(V* p, const V& v) { *p = v; }
template <>
__device__ __forceinline__ void st_stream<uint4>(uint4* p, const uint4& v) { stg_stream_v4(p, v); }

template <typename V>
__device__ __forceinline__ void warp_copy_vec(char* d, const char* s, int64_t n, int lane) {
  constexpr int W = sizeof(V);
  if (W > 1) {
    int64_t head = (W - (reinterpret_cast<uintptr_t>(d) & (W - 1))) & (W - 1);
    if (head > n) head = n;
    if (lane < head) d[lane] = s[lane];
    d += head;
    s += head;
    n -= head;
  }
  const int64_t nv = n / W;
  const V* sv = reinterpret_cast<const V*>(s);
  V* dv = reinterpret_cast<V*>(d);
  // batches of 8 vectors per lane: all loads of a batch are issued before its first store, for
  // short pieces too (a 3.5 KB row is one partially filled batch, not 7 dependent round trips)
  constexpr int kUnroll = 8;
  for (int64_t base = lane; base < nv; base += kUnroll * 32) {
    V t[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u)
      if (base + u * 32 < nv) t[u] = ld_stream<V>(sv + base + u * 32);
#pragma unroll
    for (int u = 0; u < kUnroll; ++u)
      if (base + u * 32 < nv) st_stream<V>(dv + base + u * 32, t[u]);
  }
  if (W > 1) {
    const int64_t done = nv * W;
    const int64_t tail = n - done;
    if (lane < tail) d[done + lane] = s[done + lane];
  }
}

// Source and destination disagree modulo 16: read ALIGNED 16-byte vectors around the source bytes and
// funnel-shift neighbouring pairs into aligned 16-byte stores (instead of falling back to 8/4/2/1-byte
// accesses).  kQ = whole words of the shift (0..3); the byte part r is a run-time funnel-shift amount.
// The aligned reads touch up to 15 bytes outside [s, s+n), always inside a 16-byte block that also
// holds requested bytes (so inside the same allocation).
__device__ __forceinline__ uint4 shift_pair(const uint4& a, const uint4& b, unsigned int q, unsigned int rbits) {
  // v[j] = w[j + q] for the 8 words w = (a, b), j = 0..4; q is warp-uniform
  uint32_t v0 = a.x, v1 = a.y, v2 = a.z, v3 = a.w, v4 = b.x;
  if (q & 1u) { v0 = v1; v1 = v2; v2 = v3; v3 = v4; v4 = b.y; }
  if (q & 2u) {
    const uint32_t n3 = (q & 1u) ? b.z : b.y, n4 = (q & 1u) ? b.w : b.z;
    v0 = v2; v1 = v3; v2 = v4; v3 = n3; v4 = n4;
  }
  uint4 o;
  o.x = __funnelshift_r(v0, v1, rbits);
  o.y = __funnelshift_r(v1, v2, rbits);
  o.z = __funnelshift_r(v2, v3, rbits);
  o.w = __funnelshift_r(v3, v4, rbits);
  return o;
}

__device__ __forceinline__ void warp_copy_shifted(char* d, const char* s, int64_t n, int lane) {
  int64_t head = (16 - (reinterpret_cast<uintptr_t>(d) & 15)) & 15;
  if (head > n) head = n;
  if (lane < head) d[lane] = s[lane];
  d += head;
  s += head;
  n -= head;
  const int64_t nv = n >> 4;
  if (nv > 0) {
    const unsigned int sh = static_cast<unsigned int>(reinterpret_cast<uintptr_t>(s) & 15);   // 1..15 here
    const uint4* sa = reinterpret_cast<const uint4*>(s - sh);
    uint4* dv = reinterpret_cast<uint4*>(d);
    const unsigned int rbits = (sh & 3u) * 8u, q = sh >> 2;
    // each aligned source vector is loaded ONCE: lane l holds vectors i, i+32, i+64, i+96 of a 128-vector
    // batch and gets the successor of each from lane l+1 (lane 31: lane 0's next vector / one extra load)
    constexpr int kU = 8;
#pragma unroll 1
    for (int64_t base = 0; base < nv; base += kU * 32) {
      uint4 a[kU + 1];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        a[u] = make_uint4(0, 0, 0, 0);
        if (base + u * 32 + lane <= nv) a[u] = ld_stream<uint4>(sa + base + u * 32 + lane);
      }
      a[kU] = make_uint4(0, 0, 0, 0);
      if (lane == 0 && base + kU * 32 <= nv) a[kU] = ld_stream<uint4>(sa + base + kU * 32);
      const int src = (lane + 1) & 31;
      uint4 nx[kU + 1];
#pragma unroll
      for (int u = 0; u <= kU; ++u) {
        nx[u].x = __shfl_sync(0xffffffffu, a[u].x, src);
        nx[u].y = __shfl_sync(0xffffffffu, a[u].y, src);
        nx[u].z = __shfl_sync(0xffffffffu, a[u].z, src);
        nx[u].w = __shfl_sync(0xffffffffu, a[u].w, src);
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const int64_t i = base + u * 32 + lane;
        const uint4 b = lane < 31 ? nx[u] : nx[u + 1];       // lane 31's successor is lane 0's next vector
        if (i < nv) stg_stream_v4(dv + i, shift_pair(a[u], b, q, rbits));
      }
    }
  }
  const int64_t done = nv << 4;
  const int64_t tail = n - done;
  if (lane < tail) d[done + lane] = s[done + lane];
}

__device__ __forceinline__ void warp_copy(char* d, const char* s, int64_t n, int lane) {
  const uintptr_t diff = reinterpret_cast<uintptr_t>(d) - reinterpret_cast<uintptr_t>(s);
  if ((diff & 15) == 0) warp_copy_vec<uint4>(d, s, n, lane);
  else if ((diff & 7) == 0) warp_copy_vec<uint2>(d, s, n, lane);
  else if ((diff & 3) == 0) warp_copy_vec<uint32_t>(d, s, n, lane);
  else if ((diff & 1) == 0) warp_copy_vec<uint16_t>(d, s, n, lane);
  else warp_copy_vec<uint8_t>(d, s, n, lane);
}

__device__ __forceinline__ bool takes_shift_path(const char* d, const char* s, int64_t n) {
  return n >= kShiftMinBytes && ((reinterpret_cast<uintptr_t>(d) - reinterpret_cast<uintptr_t>(s)) & 15) != 0;
}

// kSlice: flat -> packed (slice_intervals); otherwise packed -> flat (set_intervals).
// Two launches share the chunking: kShifted = false moves every piece except the long misaligned ones
// (its code and register budget are those of the plain vector copy); kShifted = true moves exactly
// those, and returns at once when the scan found none.
template <bool kSlice, bool kShifted>
__global__ void __launch_bounds__(kCopyThreads)
interval_copy_kernel(const char* __restrict__ flat_in, char* __restrict__ flat_out,
                     const char* __restrict__ packed_in, char* __restrict__ packed_out,
                     const int64_t* __restrict__ iv, int64_t n, int64_t elem_size,
                     const int64_t* __restrict__ ws, int64_t chunk_bytes) {
  const int64_t* status = ws;
  const int64_t* off = ws + 2;
  if (status[0] != 0) return;   // invalid table: touch nothing (status stays readable by the host)
  if (kShifted && off[n + 1] == 0) return;
  const int lane = threadIdx.x & 31;
  const int64_t total = off[n];
  const int64_t gwarp = int64_t(blockIdx.x) * kCopyWarps + (threadIdx.x >> 5);
  int64_t pos = gwarp * chunk_bytes;
  if (pos >= total) return;
  int64_t end = pos + chunk_bytes;
  if (end > total) end = total;
  // last interval whose packed offset is <= pos (skips empty intervals sharing the offset)
  int64_t lo = 0, hi = n;          // first k with off[k] > pos, in [1, n]
  while (lo < hi) {
    const int64_t mid = (lo + hi) >> 1;
    if (off[mid] > pos) hi = mid; else lo = mid + 1;
  }
  int64_t k = lo - 1;
  int64_t k_off = off[k];
  int64_t k_end = off[k + 1];
  while (pos < end) {
    const int64_t piece_end = k_end < end ? k_end : end;
    const int64_t fpos = iv[2 * k] * elem_size + (pos - k_off);
    char* d = kSlice ? packed_out + pos : flat_out + fpos;
    const char* s = kSlice ? flat_in + fpos : packed_in + pos;
    const int64_t len = piece_end - pos;
    if (takes_shift_path(d, s, len) == kShifted) {
      if (kShifted) warp_copy_shifted(d, s, len, lane);
      else warp_copy(d, s, len, lane);
    }
    pos = piece_end;
    while (pos == k_end && k + 1 < n) {
      ++k;
      k_off = k_end;
      k_end = off[k + 1];
    }
  }
}

}  // namespace

int64_t interval_workspace_bytes(int64_t n) {
  const int64_t ntiles = n > 0 ? (n + kScanTile - 1) / kScanTile : 1;
  return (n + 4 + kScratchWords + ntiles) * int64_t(sizeof(int64_t));
}

int launch_interval_copy(bool slice, const void* flat_in, void* flat_out, const void* packed_in,
                         void* packed_out, const int64_t* intervals, int64_t n, int64_t elem_size,
                         int64_t flat_numel, int64_t packed_numel, void* workspace,
                         cudaStream_t stream) {
  int64_t* ws = reinterpret_cast<int64_t*>(workspace);
  const uint64_t flat_base = reinterpret_cast<uint64_t>(slice ? flat_in : flat_out);
  const uint64_t packed_base = reinterpret_cast<uint64_t>(slice ? packed_out : packed_in);
  const int64_t ntiles = n > 0 ? (n + kScanTile - 1) / kScanTile : 1;
  // ticket, flags, total and the look-back words start at zero ("empty")
  AB_CUDA(cudaMemsetAsync(ws + n + 4, 0, size_t(kScratchWords + ntiles) * sizeof(int64_t), stream));
  interval_scan_kernel<<<static_cast<unsigned int>(ntiles), kScanThreads, 0, stream>>>(
      intervals, n, elem_size, flat_numel, packed_numel, ws, flat_base, packed_base);
  int rc = check_launch("interval_scan");
  if (rc) return rc;
  const int64_t total = packed_numel * elem_size;
  if (total == 0 || n == 0) return kOk;
  // equal contiguous chunks of packed bytes, one per warp; >= 4 KB each, 16 B granular
  // AREAL_B200_INTERVAL_CHUNKS_PER_SM (benchmarks): how many chunks of kCopyWarps warps the packed bytes are cut
  // into per SM; the default 8 gives four full waves of 2 resident CTAs
  static const int chunks_per_sm = [] {
    const char* e = getenv("AREAL_B200_INTERVAL_CHUNKS_PER_SM");
    const int v = e ? atoi(e) : 0;
    return v > 0 ? v : 8;
  }();
  const int64_t target_warps = int64_t(num_sms()) * chunks_per_sm * kCopyWarps;
  int64_t chunk = (total + target_warps - 1) / target_warps;
  if (chunk < 4096) chunk = 4096;
  chunk = (chunk + 15) & ~int64_t(15);
  const int64_t nwarps = (total + chunk - 1) / chunk;
  const unsigned int grid = static_cast<unsigned int>((nwarps + kCopyWarps - 1) / kCopyWarps);
  const char* fi = reinterpret_cast<const char*>(flat_in);
  char* fo = reinterpret_cast<char*>(flat_out);
  const char* pi = reinterpret_cast<const char*>(packed_in);
  char* po = reinterpret_cast<char*>(packed_out);
  if (slice) {
    interval_copy_kernel<true, false><<<grid, kCopyThreads, 0, stream>>>(fi, nullptr, nullptr, po, intervals, n,
                                                                         elem_size, ws, chunk);
    interval_copy_kernel<true, true><<<grid, kCopyThreads, 0, stream>>>(fi, nullptr, nullptr, po, intervals, n,
                                                                        elem_size, ws, chunk);
  } else {
    interval_copy_kernel<false, false><<<grid, kCopyThreads, 0, stream>>>(nullptr, fo, pi, nullptr, intervals, n,
                                                                          elem_size, ws, chunk);
    interval_copy_kernel<false, true><<<grid, kCopyThreads, 0, stream>>>(nullptr, fo, pi, nullptr, intervals, n,
                                                                         elem_size, ws, chunk);
  }
  return check_launch("interval_copy");
}

}  // namespace ab
