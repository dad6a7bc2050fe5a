// Single-pass exact median of the distance matrix (K4, tensor-core path).
//
// Replaces torch.median(sq_dists) + bw_median (src/utils/math.py:28-34) without re-reading the
// N^2 matrix three times:
//   1. window_sample : m random entries of the matrix are evaluated directly from the centred fp32
//                      features (difference form) -- they only steer the search, exactness does
//                      not depend on them;
//   2. window_bounds : one block selects two sample order statistics m/2 -+ 5.7 sigma -> [lo, hi],
//                      which brackets the true median except with probability ~1e-8;
//   3. window_select : one cooperative kernel.  Phase A reads the matrix once (right after the
//                      Gram kernel wrote it: L2 hits), counts the entries < lo and compacts the
//                      entries in [lo, hi] (about 5 %) into per-block pools (ballot compaction,
//                      shared-memory fill counters: no global atomics).  Then three radix passes
//                      over the short candidate list select rank k - below.  If the window missed
//                      (or a pool overflowed) the same kernel falls back to radix passes over the
//                      full matrix, so the result is always the exact order statistic.
#include <cooperative_groups.h>

#include "gemm_tc.cuh"

namespace cg = cooperative_groups;

namespace sigsvgd {
namespace {

constexpr int kSample = 8192;       // m
constexpr int kSampleHalfWin = 256; // 5.7 sigma, sigma = sqrt(m)/2 = 45 ranks
constexpr int kBins = kWinBins;

__device__ __forceinline__ uint32_t hash32(uint32_t x) {
  x ^= x >> 16;
  x *= 0x7feb352du;
  x ^= x >> 15;
  x *= 0x846ca68bu;
  x ^= x >> 16;
  return x;
}

__device__ __forceinline__ bool radix_bin(uint32_t bits, int pass, uint32_t prefix, uint32_t* bin) {
  bits &= 0x7fffffffu;
  if (pass == 0) {
    *bin = bits >> 21;
    return true;
  }
  if (pass == 1) {
    *bin = (bits >> 10) & 0x7ffu;
    return (bits >> 21) == (prefix >> 21);
  }
  *bin = bits & 0x3ffu;
  return (bits >> 10) == (prefix >> 10);
}

// warp-aggregated shared-memory histogram increment (most candidates share a few bins)
__device__ __forceinline__ void hist_add(unsigned int* hist, uint32_t bin, bool valid) {
  const unsigned act = __ballot_sync(0xffffffffu, valid);
  if (valid) {
    const unsigned m = __match_any_sync(act, bin);
    if ((threadIdx.x & 31) == (unsigned)(__ffs(m) - 1)) atomicAdd(&hist[bin], (unsigned int)__popc(m));
  }
}

// Block-wide (1024 threads, 2 bins each): find the bucket holding rank k of hist[kBins] and the
// rank inside it.  Parallel prefix sum: warp shuffles + one shared array of 32 warp totals.
__device__ __forceinline__ void block_pick_bucket_vals(unsigned long long h0, unsigned long long h1,
                                                       unsigned long long k, uint32_t* bucket_out,
                                                       unsigned long long* k_out,
                                                       unsigned long long* scratch /*[>= 33]*/,
                                                       unsigned long long* count_out = nullptr) {
  const int t = threadIdx.x, lane = t & 31, wid = t >> 5;
  unsigned long long incl = h0 + h1;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned long long v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  if (lane == 31) scratch[wid] = incl;
  if (t == 0) {
    *bucket_out = kBins - 1;  // only reached if k is past the total (callers never ask for that)
    *k_out = 0ull;
    if (count_out) *count_out = 0ull;
  }
  __syncthreads();
  if (wid == 0) {
    unsigned long long w = scratch[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned long long v = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += v;
    }
    scratch[lane] = w;  // inclusive warp totals
  }
  __syncthreads();
  const unsigned long long excl = (wid > 0 ? scratch[wid - 1] : 0ull) + incl - (h0 + h1);
  if (k >= excl && k < excl + h0 + h1) {
    if (k < excl + h0) {
      *bucket_out = 2 * t;
      *k_out = k - excl;
      if (count_out) *count_out = h0;
    } else {
      *bucket_out = 2 * t + 1;
      *k_out = k - excl - h0;
      if (count_out) *count_out = h1;
    }
  }
  __syncthreads();
}

template <typename CountT>
__device__ __forceinline__ void block_pick_bucket(const CountT* hist, unsigned long long k,
                                                  uint32_t* bucket_out, unsigned long long* k_out,
                                                  unsigned long long* scratch /*[>= 32]*/) {
  block_pick_bucket_vals((unsigned long long)hist[2 * threadIdx.x], (unsigned long long)hist[2 * threadIdx.x + 1], k,
                         bucket_out, k_out, scratch);
}

// exact order statistic `k` of n floats (>= 0) held in shared memory; one block of 1024 threads
__device__ float block_select_smem(const float* vals, int n, unsigned long long k,
                                   unsigned int* hist /*[kBins] shared*/,
                                   unsigned long long* scratch /*[1024] shared*/,
                                   uint32_t* sh_bucket, unsigned long long* sh_k) {
  uint32_t prefix = 0;
  for (int pass = 0; pass < 3; ++pass) {
    for (int i = threadIdx.x; i < kBins; i += blockDim.x) hist[i] = 0u;
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) {  // n is a multiple of blockDim.x
      uint32_t bin;
      const bool in = radix_bin(__float_as_uint(vals[i]), pass, prefix, &bin);
      hist_add(hist, bin, in);
    }
    __syncthreads();
    block_pick_bucket(hist, k, sh_bucket, sh_k, scratch);
    const int shift = pass == 0 ? 21 : pass == 1 ? 10 : 0;
    prefix |= *sh_bucket << shift;
    k = *sh_k;
    __syncthreads();
  }
  return __uint_as_float(prefix);
}

// Window preparation, two launches on the side stream:
//  window_sample_kernel  every warp evaluates two sample entries d = |a_i - b_j|^2 (fp32 difference
//    form; the features need not be centred; all loads of a sample in flight at once).  Small blocks
//    (256 threads): they find room on SMs that the main stream's packing pass is still using, instead of
//    waiting for it to drain and then running beside -- and slowing -- the Gram (which cost it 4-5 us
//    when sampling and bounds were one kernel of 1024-thread blocks);
//  window_bounds_kernel  one block selects the sample order statistics m/2 -+ kSampleHalfWin with two
//    histogram passes (both bounds at once) and sets the window; it also resets the window /
//    histogram / bandwidth records.
constexpr int kPrepThreads = 1024;
constexpr int kSampleThreads = 256;
constexpr int kSamplesPerBlock = 2 * (kSampleThreads / 32);

__global__ void __launch_bounds__(kSampleThreads)
window_sample_kernel(const float* __restrict__ a, int64_t lda, int64_t M, const float* __restrict__ b,
                     int64_t ldb, int64_t N, int64_t D, float* __restrict__ sample) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll 1
  for (int q = 0; q < kSamplesPerBlock / (kSampleThreads / 32); ++q) {
    const int sidx = blockIdx.x * kSamplesPerBlock + q * (kSampleThreads / 32) + wid;
    const int64_t i = (int64_t)(((uint64_t)hash32(2u * sidx + 1u) * (uint64_t)M) >> 32);
    const int64_t j = (int64_t)(((uint64_t)hash32(2u * sidx + 0x9e3779b9u) * (uint64_t)N) >> 32);
    const float* __restrict__ ai = a + i * lda;
    const float* __restrict__ bj = b + j * ldb;
    float acc = 0.f;
    // batches of 8 column pairs: their 16 loads are issued before the first use (one pair per
    // iteration made a sample a chain of D/32 exposed L2 latencies)
    for (int64_t c0 = lane; c0 < D; c0 += 32 * 8) {
      float va[8], vb[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int64_t c = c0 + 32 * u;
        va[u] = c < D ? __ldg(ai + c) : 0.f;
        vb[u] = c < D ? __ldg(bj + c) : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const float t = va[u] - vb[u];
        acc = fmaf(t, t, acc);
      }
    }
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) sample[sidx] = acc;
  }
}

__global__ void __launch_bounds__(kPrepThreads)
window_bounds_kernel(const float* __restrict__ sample, SelWindow* win, unsigned int cap, unsigned int npool,
                     unsigned long long* ghist /*[2*3*kBins]*/, unsigned int* pool_cnt /*[2*kMaxPools + 64]*/,
                     sigsvgd_bw_state* state) {
  extern __shared__ float sv[];  // kSample floats
  __shared__ unsigned int hist[2][kBins];
  __shared__ unsigned long long scratch[1024];
  __shared__ float sh_red[64];
  __shared__ uint32_t sh_bucket[2];
  __shared__ unsigned long long sh_k[2];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < kSample; i += kPrepThreads) sv[i] = __ldcg(sample + i);
  for (int i = threadIdx.x; i < 2 * 3 * kBins; i += kPrepThreads) ghist[i] = 0ull;
  for (int i = threadIdx.x; i < 2 * kMaxPools + 64; i += kPrepThreads) pool_cnt[i] = 0u;
  __syncthreads();
  // Window bounds = sample order statistics m/2 -+ 5.7 sigma, located by two levels of 2048 linear
  // bins (whole range, then the selected bin) and rounded OUTWARD by one fine bin: the bounds only
  // steer the search, so a slightly wider window is harmless.
  float mn = INFINITY, mx = 0.f;
  for (int i = threadIdx.x; i < kSample; i += kPrepThreads) {
    const float v = sv[i];
    mn = fminf(mn, v), mx = fmaxf(mx, v);
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  }
  if (lane == 0) sh_red[wid] = mn, sh_red[32 + wid] = mx;
  __syncthreads();
  mn = sh_red[lane], mx = sh_red[32 + lane];
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  }
  const unsigned long long ranks[2] = {kSample / 2 - kSampleHalfWin, kSample / 2 + kSampleHalfWin};
  const float width0 = fmaxf((mx - mn) * (1.f / kBins), 0.f);
  const float inv_w0 = width0 > 0.f ? 1.f / width0 : 0.f;
  auto bin0_of = [&](float v) { return max(0, min(kBins - 1, (int)((v - mn) * inv_w0))); };
  // level 0: one histogram of the whole range serves both bounds
  for (int i = threadIdx.x; i < 2 * kBins; i += kPrepThreads) (&hist[0][0])[i] = 0u;
  __syncthreads();
  for (int i = threadIdx.x; i < kSample; i += kPrepThreads) atomicAdd(&hist[0][bin0_of(sv[i])], 1u);
  __syncthreads();
  block_pick_bucket(hist[0], ranks[0], &sh_bucket[0], &sh_k[0], scratch);
  block_pick_bucket(hist[0], ranks[1], &sh_bucket[1], &sh_k[1], scratch);
  const uint32_t b0[2] = {sh_bucket[0], sh_bucket[1]};
  const unsigned long long k1[2] = {sh_k[0], sh_k[1]};
  __syncthreads();
  // level 1: the two selected bins, 2048 finer bins each
  const float width1 = width0 * (1.f / kBins);
  const float inv_w1 = width1 > 0.f ? 1.f / width1 : 0.f;
  const float base1[2] = {mn + (float)b0[0] * width0, mn + (float)b0[1] * width0};
  for (int i = threadIdx.x; i < kBins; i += kPrepThreads) hist[0][i] = 0u;
  __syncthreads();
  for (int i = threadIdx.x; i < kSample; i += kPrepThreads) {
    const float v = sv[i];
    const uint32_t bv = (uint32_t)bin0_of(v);
#pragma unroll
    for (int which = 0; which < 2; ++which)
      if (bv == b0[which]) atomicAdd(&hist[which][max(0, min(kBins - 1, (int)((v - base1[which]) * inv_w1)))], 1u);
  }
  __syncthreads();
  block_pick_bucket(hist[0], k1[0], &sh_bucket[0], &sh_k[0], scratch);
  block_pick_bucket(hist[1], k1[1], &sh_bucket[1], &sh_k[1], scratch);
  if (threadIdx.x == 0) {
    const float lo_e = fmaxf(base1[0] + ((float)sh_bucket[0] - 1.f) * width1, 0.f);
    const float hi_e = base1[1] + ((float)sh_bucket[1] + 2.f) * width1;
    win->lo = fminf(lo_e, hi_e);
    win->hi = fmaxf(lo_e, hi_e);
    win->n_cand = 0u;
    win->cap = cap;
    win->below = 0ull;
    win->active = 1u;
    win->fallback = 0u;
    win->npool = npool;
    win->poolcap = cap / npool;
    win->epoch = 0u;
    sigsvgd_bw_state& s = *state;
    s.prefix = 0u;
    s.k_rem = 0ull;
    s.median = 0.f;
    s.sum_kd = 0.0;
    s.n_ties = 0u;
    s.n_ties_all = 0u;
    s.clamped = 0u;
    s.coef = 0.f;
  }
}

constexpr int kShort = 6144;  // capacity of the short list finished by a single block

// cooperative: exact order statistic k_total of the M x N matrix `dist`
//   phase A  one pass over the matrix: count entries < lo, compact entries in [lo, hi] into this
//            block's pool and histogram them on the top 11 bits of rel = bits - lo_bits
//   phase B  entries of the chosen bucket (a few hundred) are gathered into a short list which one
//            block finishes in shared memory; buckets too big for that (massive ties) take up to
//            two more 11-bit radix passes over the pools
//   fallback (window missed / pool overflow): three radix passes over the full matrix
// Histograms / counters are double-buffered on win->epoch so the kernel leaves them clean for the
// next call without a trailing grid barrier.
__global__ void __launch_bounds__(1024)
window_select_kernel(SelWindow* win, float* cand, unsigned int* pool_cnt, float* shortlist,
                     const float* __restrict__ dist, int64_t M, int64_t N, int64_t ldd,
                     unsigned long long k_total, unsigned long long* ghist2, float log_np1,
                     float bw_scale, float tol, sigsvgd_bw_state* state, int symmetric) {
  cg::grid_group grid = cg::this_grid();
  __shared__ unsigned int hist[kBins];
  __shared__ unsigned long long scratch[1024];
  __shared__ float sl[kShort];
  __shared__ uint32_t sh_bucket;
  __shared__ unsigned long long sh_k;
  __shared__ unsigned int sh_ncand, sh_over, sh_fill, sh_blt, sh_last;
  __shared__ unsigned long long sh_below;
  const unsigned int npool = win->npool, poolcap = win->poolcap;
  const unsigned int epoch = win->epoch & 1u;
  unsigned long long* ghist = ghist2 + (size_t)epoch * 3 * kBins;
  unsigned int* short_cnt = pool_cnt + 2 * kMaxPools + epoch;
  // grid totals of phase A (candidates kept, pool overflow flag, entries below the window):
  // accumulated with three atomics per block BEFORE the grid barrier, double-buffered on the epoch
  unsigned int* done_cnt = pool_cnt + 2 * kMaxPools + 16 + epoch;  // ticket of the short-list finish
  unsigned int* tot32 = pool_cnt + 2 * kMaxPools + 8 + 4 * epoch;
  unsigned long long* tot_below = reinterpret_cast<unsigned long long*>(tot32 + 2);
  const int lane = threadIdx.x & 31;
  bool tri_scan = false;
  const uint32_t lo_bits = __float_as_uint(win->lo), span_bits = __float_as_uint(win->hi) - lo_bits;
  // radix digits of rel: s1 = shift of the top digit (<= 11 bits), then s2, then 0
  const int nbits = 32 - __clz(span_bits | 1u);
  const int s1 = max(0, nbits - 11), s2 = max(0, s1 - 11);
  float* __restrict__ pool = cand + (size_t)blockIdx.x * poolcap;

  // ---------------------------------------------------------------- phase A
  if (threadIdx.x == 0) sh_fill = 0u, sh_blt = 0u, sh_ncand = 0u, sh_over = 0u, sh_below = 0ull;
  for (int i = threadIdx.x; i < kBins; i += blockDim.x) hist[i] = 0u;
  if (blockIdx.x == 0) {  // clean the OTHER buffer for the next call
    unsigned long long* other = ghist2 + (size_t)(epoch ^ 1u) * 3 * kBins;
    for (int i = threadIdx.x; i < 3 * kBins; i += blockDim.x) other[i] = 0ull;
    if (threadIdx.x == 0) pool_cnt[2 * kMaxPools + (epoch ^ 1u)] = 0u;
    if (threadIdx.x < 4) pool_cnt[2 * kMaxPools + 8 + 4 * (epoch ^ 1u) + threadIdx.x] = 0u;
    if (threadIdx.x == 4) pool_cnt[2 * kMaxPools + 16 + (epoch ^ 1u)] = 0u;
  }
  __syncthreads();
  {
    const bool vec = (N & 3) == 0 && (ldd & 3) == 0 && (reinterpret_cast<uintptr_t>(dist) & 15u) == 0;
    const int64_t nq = vec ? N / 4 : N;  // items per row
    unsigned int nlt = 0;
    tri_scan = symmetric && vec && M == N && lo_bits > 0u;  // grid-uniform
    // a thread classifies its batch of values locally; the warp then reserves pool space with ONE
    // shared-memory atomic per batch (exclusive scan of the per-thread hit counts)
    constexpr int kRows = 8;  // rows in flight per warp-chunk (independent 16-byte loads per lane)
    // (hits are re-read from L1 instead of being kept in registers across the scan: 8 rows in flight
    // per lane fit the 64-register budget of a 1024-thread block only that way)
    auto flush = [&](uint32_t hitmask, int nval, int64_t r0, int64_t c) {  // warp-collective
      const int hits = __popc(hitmask);
      int incl = hits;
#pragma unroll
      for (int sft = 1; sft < 32; sft <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, sft);
        if (lane >= sft) incl += t;
      }
      const int total = __shfl_sync(0xffffffffu, incl, 31);
      if (total == 0) return;  // warp-uniform
      unsigned int base = 0;
      if (lane == 31) base = atomicAdd(&sh_fill, (unsigned int)total);
      base = __shfl_sync(0xffffffffu, base, 31);
      unsigned int off = base + (unsigned int)(incl - hits);
      while (hitmask) {
        const int e = __ffs(hitmask) - 1;
        hitmask &= hitmask - 1u;
        if (e >= nval) break;
        const float val = vec ? __ldg(dist + (r0 + (e >> 2)) * ldd + 4 * c + (e & 3)) : __ldg(dist + (r0 + e) * ldd + c);
        if (off < poolcap) pool[off] = val;
        atomicAdd(&hist[((__float_as_uint(val) & 0x7fffffffu) - lo_bits) >> s1], 1u);
        ++off;
      }
    };
    // one warp-chunk: rows r0..r0+3, quads c0..c0+31 (128 columns).  tri: only entries strictly
    // above the diagonal (column > row) are classified.
    auto chunk = [&](int64_t r0, int64_t c0, bool tri) {  // warp-collective
      const int64_t c = c0 + lane;
      float v[4 * kRows];
      uint32_t hitmask = 0;
      if (vec && r0 + kRows <= M && c0 + 32 <= nq && (!tri || 4 * c0 > r0 + kRows - 1)) {
        // interior: no validity bookkeeping
#pragma unroll
        for (int j = 0; j < kRows; ++j) {
          const float4 f = __ldg(reinterpret_cast<const float4*>(dist + (r0 + j) * ldd) + c);
          v[4 * j] = f.x, v[4 * j + 1] = f.y, v[4 * j + 2] = f.z, v[4 * j + 3] = f.w;
        }
#pragma unroll
        for (int e = 0; e < 4 * kRows; ++e) {
          const uint32_t bits = __float_as_uint(v[e]) & 0x7fffffffu;
          nlt += bits < lo_bits ? 1u : 0u;
          if (bits - lo_bits <= span_bits) hitmask |= 1u << e;
        }
        flush(hitmask, 4 * kRows, r0, c);
        return;
      }
      const bool cin = c < nq;
      uint32_t validmask = 0;
      if (vec) {
#pragma unroll
        for (int j = 0; j < kRows; ++j) {
          float4 f = make_float4(0.f, 0.f, 0.f, 0.f);
          if (cin && r0 + j < M) {
            f = __ldg(reinterpret_cast<const float4*>(dist + (r0 + j) * ldd) + c);
            uint32_t m4 = 0xfu;
            if (tri) {  // keep columns 4c+u > r0+j
              const int64_t first = r0 + j + 1 - 4 * c;  // first valid component
              m4 = first <= 0 ? 0xfu : first >= 4 ? 0u : (0xfu << first) & 0xfu;
            }
            validmask |= m4 << (4 * j);
          }
          v[4 * j] = f.x, v[4 * j + 1] = f.y, v[4 * j + 2] = f.z, v[4 * j + 3] = f.w;
        }
      } else {
#pragma unroll
        for (int j = 0; j < kRows; ++j) {
          v[j] = 0.f;
          if (cin && r0 + j < M) v[j] = __ldg(dist + (r0 + j) * ldd + c), validmask |= 1u << j;
        }
      }
#pragma unroll
      for (int e = 0; e < 4 * kRows; ++e) {
        if (!vec && e >= kRows) break;
        const uint32_t bits = __float_as_uint(v[e]) & 0x7fffffffu;
        const bool ok_e = (validmask >> e) & 1u;
        nlt += (ok_e && bits < lo_bits) ? 1u : 0u;
        if (ok_e && bits - lo_bits <= span_bits) hitmask |= 1u << e;
      }
      flush(hitmask, vec ? 4 * kRows : kRows, r0, c);
    };
    if (!tri_scan) {
      for (int64_t r0 = (int64_t)blockIdx.x * kRows; r0 < M; r0 += (int64_t)gridDim.x * kRows)
        for (int64_t c0 = threadIdx.x & ~31; c0 < nq; c0 += blockDim.x) chunk(r0, c0, false);
    } else {
      // bit-symmetric matrix with a zero diagonal: only the strict upper triangle is read.  The
      // chunks of this block's row groups form one list that its warps walk round-robin, so the
      // number of sequential chunk steps per warp halves too (not just the bytes).
      const int wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
      int offset = 0;
      // (row groups dealt in serpentine order: a block's groups from the wide top of the triangle
      // pair with groups from its narrow bottom)
      for (int64_t gk = 0;; ++gk) {
        const int64_t g = (gk & 1) ? (gk + 1) * (int64_t)gridDim.x - 1 - blockIdx.x : gk * (int64_t)gridDim.x + blockIdx.x;
        if (gk * (int64_t)gridDim.x * kRows >= M) break;
        const int64_t r0 = g * kRows;
        if (r0 >= M) continue;
        const int64_t q_first = ((r0 + 1) >> 2) & ~(int64_t)31;
        if (q_first >= nq) continue;
        const int nch = (int)((nq - q_first + 31) >> 5);
        int j0 = (wid - offset) % nw;
        if (j0 < 0) j0 += nw;
        for (int j = j0; j < nch; j += nw) chunk(r0, q_first + 32 * (int64_t)j, true);
        offset = (offset + nch) % nw;
      }
    }
    unsigned int tot = nlt;
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
    if (lane == 0 && tot) atomicAdd(&sh_blt, tot);
    __syncthreads();
    if (threadIdx.x == 0) {
      // plain stores of per-block results (nothing to reset between calls)
      pool_cnt[blockIdx.x] = sh_fill;
      pool_cnt[kMaxPools + blockIdx.x] = sh_blt;  // < 2^32: rows are dealt evenly
      atomicAdd(&tot32[0], min(sh_fill, poolcap));
      if (sh_fill > poolcap) atomicOr(&tot32[1], 1u);
      atomicAdd(tot_below, (unsigned long long)sh_blt);
    }
    // blocks finish together: start each one at a different bin so that they do not queue on the
    // same L2 atomic units
    for (int i0 = threadIdx.x; i0 < kBins; i0 += blockDim.x) {
      const int i = (i0 + (int)blockIdx.x * 64) & (kBins - 1);
      if (hist[i]) atomicAdd(&ghist[i], (unsigned long long)hist[i]);
    }
  }
  __threadfence();
  grid.sync();
  // everything the tail needs from global memory is requested here, in ONE round trip: the grid
  // totals, this thread's two bins of the global histogram, and this block's first pool entries
  // (the pool is scanned again once the bucket is known; its values do not depend on the bucket)
  __shared__ unsigned long long sh_bn;
  const unsigned int my_cnt = blockIdx.x < npool ? min(sh_fill, poolcap) : 0u;
  const unsigned long long gh0 = __ldcg(ghist + 2 * threadIdx.x), gh1 = __ldcg(ghist + 2 * threadIdx.x + 1);
  constexpr int kPre = 4;  // pool entries per thread kept in registers (pools hold ~3.5k at N = 4096)
  float pv[kPre];
#pragma unroll
  for (int j = 0; j < kPre; ++j) {
    const unsigned int i = threadIdx.x + j * blockDim.x;
    pv[j] = i < my_cnt ? pool[i] : 0.f;
  }
  if (threadIdx.x == 0) {
    sh_ncand = __ldcg(&tot32[0]);
    sh_over = __ldcg(&tot32[1]);
    sh_below = __ldcg(tot_below);
  }
  __syncthreads();
  const unsigned int n_cand = sh_ncand;
  // triangle scan: every counted entry stands for two, and the M diagonal zeros lie below lo > 0
  const unsigned long long below = tri_scan ? 2ull * sh_below + (unsigned long long)M : sh_below;
  const unsigned long long n_cand_full = tri_scan ? 2ull * n_cand : (unsigned long long)n_cand;
  const bool ok = sh_over == 0u && k_total >= below && k_total < below + n_cand_full;
  uint32_t result_bits = 0;
  bool writer = blockIdx.x == 0;  // who stores the result (the finishing block on the short-list path)
  if (ok) {
    // ---------------------------------------------------------------- window hit
    // (in the doubled candidate list, position k' holds element k'/2 of the single list)
    unsigned long long k = tri_scan ? (k_total - below) >> 1 : k_total - below;
    block_pick_bucket_vals(gh0, gh1, k, &sh_bucket, &sh_k, scratch, &sh_bn);
    uint32_t rel_prefix = sh_bucket << s1;
    k = sh_k;
    const unsigned long long bucket_n = sh_bn;
    __syncthreads();
    if (s1 == 0) {
      result_bits = lo_bits + rel_prefix;
    } else if (bucket_n <= (unsigned long long)kShort) {
      // phase B: gather the bucket into the short list, one block finishes
      writer = false;
      // (matches are gathered in shared memory first: ONE global reservation per block instead of one
      // round trip per warp and step)
      const unsigned int cnt = my_cnt;
      __shared__ unsigned int sh_base;
      __syncthreads();  // (my_cnt was read from sh_fill by every thread)
      if (threadIdx.x == 0) sh_fill = 0u;
      __syncthreads();
      auto keep = [&](float v, bool inb) {  // warp-collective
        const uint32_t rel = (__float_as_uint(v) & 0x7fffffffu) - lo_bits;
        const bool in = inb && (rel >> s1) == (rel_prefix >> s1);
        const unsigned m = __ballot_sync(0xffffffffu, in);
        if (m) {
          unsigned int base = 0;
          if (lane == 0) base = atomicAdd(&sh_fill, (unsigned int)__popc(m));
          base = __shfl_sync(0xffffffffu, base, 0);
          const unsigned int at = base + __popc(m & ((1u << lane) - 1u));
          if (in && at < (unsigned int)kShort) sl[at] = v;  // (a block holds at most bucket_n <= kShort)
        }
      };
#pragma unroll
      for (int j = 0; j < kPre; ++j) keep(pv[j], threadIdx.x + j * blockDim.x < cnt);
      for (unsigned int i0 = kPre * blockDim.x + (threadIdx.x & ~31u); i0 < cnt; i0 += blockDim.x) {
        const unsigned int i = i0 + (unsigned int)lane;
        keep(i < cnt ? pool[i] : 0.f, i < cnt);
      }
      __syncthreads();
      const unsigned int nloc = min(sh_fill, (unsigned int)kShort);
      if (threadIdx.x == 0 && nloc) sh_base = atomicAdd(short_cnt, nloc);
      __syncthreads();
      for (unsigned int i = threadIdx.x; i < nloc; i += blockDim.x) shortlist[sh_base + i] = sl[i];
      // no second grid barrier: the last block to finish its gather (ticket) finishes the select,
      // the others are done
      __threadfence();
      __syncthreads();
      if (threadIdx.x == 0) sh_last = atomicAdd(done_cnt, 1u) == gridDim.x - 1 ? 1u : 0u;
      __syncthreads();
      if (!sh_last) return;
      __threadfence();
      {
        const int n = (int)bucket_n;
        const int npad = (n + 1023) & ~1023;  // <= kShort (a multiple of 1024)
        for (int i = threadIdx.x; i < npad; i += blockDim.x)
          sl[i] = i < n ? __ldcg(shortlist + i) : __uint_as_float(0x7f800000u);
        __syncthreads();
        const float med = block_select_smem(sl, npad, k, hist, scratch, &sh_bucket, &sh_k);
        result_bits = __float_as_uint(med) & 0x7fffffffu;
        writer = true;
      }
    } else {
      // massive ties: keep refining the digits of rel over the pools
      const int shifts[2] = {s2, 0};
      const int prev[2] = {s1, s2};
      for (int pass = 1; pass <= 2; ++pass) {
        const int sp = prev[pass - 1];  // digits above sp are fixed
        if (sp == 0) break;             // nothing left to refine (grid-uniform)
        const int sh = shifts[pass - 1];
        for (int i = threadIdx.x; i < kBins; i += blockDim.x) hist[i] = 0u;
        __syncthreads();
        for (unsigned int pl = blockIdx.x; pl < npool; pl += gridDim.x) {
          const unsigned int cnt = min(pool_cnt[pl], poolcap);
          const float* src = cand + (size_t)pl * poolcap;
          for (unsigned int i0 = threadIdx.x & ~31u; i0 < cnt; i0 += blockDim.x) {
            const unsigned int i = i0 + (unsigned int)lane;
            const bool inb = i < cnt;
            const uint32_t rel = inb ? (__float_as_uint(src[i]) & 0x7fffffffu) - lo_bits : 0u;
            const bool in = inb && (rel >> sp) == (rel_prefix >> sp);
            hist_add(hist, (rel >> sh) & ((1u << (sp - sh)) - 1u), in);
          }
        }
        __syncthreads();
        unsigned long long* g = ghist + pass * kBins;
        for (int i = threadIdx.x; i < kBins; i += blockDim.x)
          if (hist[i]) atomicAdd(&g[i], (unsigned long long)hist[i]);
        __threadfence();
        grid.sync();
        block_pick_bucket(g, k, &sh_bucket, &sh_k, scratch);
        rel_prefix |= sh_bucket << sh;
        k = sh_k;
        __syncthreads();
      }
      result_bits = lo_bits + rel_prefix;
    }
  } else {
    // ---------------------------------------------------------------- fallback: full matrix
    unsigned long long k = k_total;
    uint32_t prefix = 0;
    grid.sync();  // every block has consumed the window attempt's histogram
    if (blockIdx.x == 0)
      for (int i = threadIdx.x; i < kBins; i += blockDim.x) ghist[i] = 0ull;
    __threadfence();
    grid.sync();
    const int64_t count = M * N;
    const int64_t step = (int64_t)gridDim.x * blockDim.x;
    for (int pass = 0; pass < 3; ++pass) {
      for (int i = threadIdx.x; i < kBins; i += blockDim.x) hist[i] = 0u;
      __syncthreads();
      for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31); i0 < count; i0 += step) {
        const int64_t i = i0 + lane;
        float v = 0.f;
        const bool inb = i < count;
        if (inb) {
          const int64_t r = i / N;
          v = __ldg(dist + r * ldd + (i - r * N));
        }
        uint32_t bin;
        const bool in = radix_bin(__float_as_uint(v), pass, prefix, &bin) && inb;
        hist_add(hist, bin, in);
      }
      __syncthreads();
      unsigned long long* g = ghist + pass * kBins;
      for (int i = threadIdx.x; i < kBins; i += blockDim.x)
        if (hist[i]) atomicAdd(&g[i], (unsigned long long)hist[i]);
      __threadfence();
      grid.sync();
      block_pick_bucket(g, k, &sh_bucket, &sh_k, scratch);
      const int shift = pass == 0 ? 21 : pass == 1 ? 10 : 0;
      prefix |= sh_bucket << shift;
      k = sh_k;
      __syncthreads();
    }
    result_bits = prefix;
  }
  if (writer && threadIdx.x == 0) {
    sigsvgd_bw_state& s = *state;
    const float med = __uint_as_float(result_bits);
    float h = bw_scale * sqrtf(med / log_np1);  // math.py:31-34
    const bool clamped = !(h >= tol);
    if (clamped) h = tol;
    s.median = med;
    s.h = h;
    s.gamma = -0.5f / (h * h);  // _kernels.py:110
    s.clamped = clamped ? 1u : 0u;
    s.prefix = result_bits;
    win->n_cand = n_cand;
    win->below = below;
    win->fallback = ok ? 0u : 1u;
    win->epoch = win->epoch + 1u;  // the next call uses the buffers cleaned above
  }
}

}  // namespace

size_t window_workspace_bytes(int64_t M, int64_t N, unsigned int* cap_out) {
  // candidates: ~4.7 % of the entries expected; room for 12 % plus slack
  const double want = 0.12 * (double)M * (double)N + 65536.0;
  const unsigned int cap = (unsigned int)std::min<double>(want, 4.0e9);
  if (cap_out) *cap_out = cap;
  return (size_t)round_up((int64_t)sizeof(SelWindow), 256) + (size_t)round_up(sizeof(float) * kSample, 256) +
         (size_t)round_up(sizeof(unsigned long long) * 2 * 3 * kBins, 256) +
         sizeof(unsigned int) * (2 * kMaxPools + 64) + sizeof(float) * kShort + sizeof(float) * (size_t)cap;
}

// ws layout: [SelWindow | sample | ghist x2 | pool_cnt[kMaxPools] + below_cnt[kMaxPools] + misc[64] |
//             shortlist[kShort] | cand]
static void window_carve(void* ws, SelWindow** win, float** sample, unsigned long long** ghist,
                         unsigned int** pool_cnt, float** cand, float** shortlist = nullptr) {
  char* w = (char*)ws;
  *win = (SelWindow*)w;
  w += round_up((int64_t)sizeof(SelWindow), 256);
  *sample = (float*)w;
  w += round_up(sizeof(float) * kSample, 256);
  *ghist = (unsigned long long*)w;
  w += round_up(sizeof(unsigned long long) * 2 * 3 * kBins, 256);
  *pool_cnt = (unsigned int*)w;
  w += sizeof(unsigned int) * (2 * kMaxPools + 64);
  if (shortlist) *shortlist = (float*)w;
  w += sizeof(float) * kShort;
  *cand = (float*)w;
}

void window_pointers(void* ws, SelWindow** win, float** cand, unsigned int** pool_cnt) {
  float* sample;
  unsigned long long* ghist;
  window_carve(ws, win, &sample, &ghist, pool_cnt, cand);
}

int launch_window_prepare(const float* a, int64_t lda, int64_t M, const float* b, int64_t ldb,
                          int64_t N, int64_t D, void* ws, sigsvgd_bw_state* state, SelWindow** win_out,
                          float** cand_out, cudaStream_t st) {
  SIG_CHECK_ARG(M >= 1 && N >= 1 && D >= 1 && lda >= D && ldb >= D, "window_prepare: bad shape");
  SelWindow* win;
  float *sample, *cand;
  unsigned long long* ghist;
  unsigned int* pool_cnt;
  window_carve(ws, &win, &sample, &ghist, &pool_cnt, &cand);
  unsigned int cap = 0;
  window_workspace_bytes(M, N, &cap);
  static thread_local int configured_dev = -1;
  int dev = 0;
  SIG_CUDA_OK(cudaGetDevice(&dev));
  if (configured_dev != dev) {
    SIG_CUDA_OK(cudaFuncSetAttribute(window_bounds_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)(sizeof(float) * kSample)));
    configured_dev = dev;
  }
  const unsigned int npool = (unsigned int)std::min(num_sms(), kMaxPools);
  static_assert(kSample % kSamplesPerBlock == 0, "whole blocks of samples");
  window_sample_kernel<<<kSample / kSamplesPerBlock, kSampleThreads, 0, st>>>(a, lda, M, b, ldb, N, D, sample);
  SIG_LAUNCH_OK("window_sample_kernel");
  window_bounds_kernel<<<1, kPrepThreads, sizeof(float) * kSample, st>>>(sample, win, cap, npool, ghist, pool_cnt,
                                                                         state);
  SIG_LAUNCH_OK("window_bounds_kernel");
  *win_out = win;
  *cand_out = cand;
  return 0;
}

int launch_window_select(void* ws, const float* dist, int64_t M, int64_t N, int64_t ldd,
                         int64_t n_log, float bw_scale, float tol, sigsvgd_bw_state* state,
                         cudaStream_t st, int symmetric) {
  SIG_CHECK_ARG(n_log >= 1, "window_select: n_log must be >= 1");
  SelWindow* win;
  float *sample, *cand;
  unsigned long long* ghist;
  unsigned int* pool_cnt;
  float* shortlist;
  window_carve(ws, &win, &sample, &ghist, &pool_cnt, &cand, &shortlist);
  const float log_np1 = (float)log((double)(float)((double)n_log + 1.0));
  unsigned long long k_total = (unsigned long long)((M * N - 1) / 2);
  void* args[] = {(void*)&win,   (void*)&cand, (void*)&pool_cnt, (void*)&shortlist, (void*)&dist, (void*)&M,   (void*)&N,    (void*)&ldd,
                  (void*)&k_total, (void*)&ghist, (void*)&log_np1, (void*)&bw_scale, (void*)&tol, (void*)&state, (void*)&symmetric};
  // one block per SM: co-resident by construction (cooperative launch checks it)
  SIG_CUDA_OK(cudaLaunchCooperativeKernel((const void*)window_select_kernel, dim3((unsigned)num_sms()),
                                          dim3(1024), args, 0, st));
  return 0;
}

}  // namespace sigsvgd
