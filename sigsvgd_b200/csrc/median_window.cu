// Single-pass exact median of the distance matrix (K4, tensor-core path).
//
// Replaces torch.median(sq_dists) + bw_median (src/utils/math.py:28-34) without re-reading the
// N^2 matrix three times:
//   1. window_sample : m random entries of the matrix are evaluated directly from the centred fp32
//                      features (difference form) -- they only steer the search, exactness does
//                      not depend on them;
//   2. window_bounds : one block selects two sample order statistics m/2 -+ 6 sigma -> [lo, hi],
//                      which brackets the true median except with probability ~1e-9;
//   3. tc_gram's epilogue counts the entries < lo and appends the entries in [lo, hi]
//                      (about 5 % of the matrix) to a candidate list while it writes the matrix;
//   4. window_select : one cooperative kernel radix-selects rank k - below among the candidates
//                      (3 passes over the short list).  If the window missed (or the list
//                      overflowed) the same kernel falls back to the full matrix, so the result
//                      is always the exact order statistic.
#include <cooperative_groups.h>

#include "gemm_tc.cuh"

namespace cg = cooperative_groups;

namespace sigsvgd {
namespace {

constexpr int kSample = 16384;      // m
constexpr int kSampleHalfWin = 384; // 6 sigma, sigma = sqrt(m)/2 = 64 ranks
constexpr int kBins = 2048;

__device__ __forceinline__ uint32_t hash32(uint32_t x) {
  x ^= x >> 16;
  x *= 0x7feb352du;
  x ^= x >> 15;
  x *= 0x846ca68bu;
  x ^= x >> 16;
  return x;
}

// one warp per sample: d = |a_i - b_j|^2 over D columns (fp32, difference form)
__global__ void __launch_bounds__(256)
window_sample_kernel(const float* __restrict__ a, int64_t lda, int64_t M,
                     const float* __restrict__ b, int64_t ldb, int64_t N, int64_t D,
                     float* __restrict__ sample) {
  const int s = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (s >= kSample) return;
  const int lane = threadIdx.x & 31;
  const int64_t i = (int64_t)(((uint64_t)hash32(2u * s + 1u) * (uint64_t)M) >> 32);
  const int64_t j = (int64_t)(((uint64_t)hash32(2u * s + 0x9e3779b9u) * (uint64_t)N) >> 32);
  const float* __restrict__ ai = a + i * lda;
  const float* __restrict__ bj = b + j * ldb;
  float acc = 0.f;
  for (int64_t c = lane; c < D; c += 32) {
    const float t = __ldg(ai + c) - __ldg(bj + c);
    acc = fmaf(t, t, acc);
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) sample[s] = acc;
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
template <typename CountT>
__device__ __forceinline__ void block_pick_bucket(const CountT* hist, unsigned long long k,
                                                  uint32_t* bucket_out, unsigned long long* k_out,
                                                  unsigned long long* scratch /*[>= 32]*/) {
  const int t = threadIdx.x, lane = t & 31, wid = t >> 5;
  const unsigned long long h0 = (unsigned long long)hist[2 * t];
  const unsigned long long h1 = (unsigned long long)hist[2 * t + 1];
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
    } else {
      *bucket_out = 2 * t + 1;
      *k_out = k - excl - h0;
    }
  }
  __syncthreads();
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

// one block: window bounds from the sample, reset of the window counters, the three global
// histograms used by window_select, and of the bandwidth record
__global__ void __launch_bounds__(1024)
window_bounds_kernel(const float* __restrict__ sample, SelWindow* win, unsigned int cap,
                     unsigned long long* ghist /*[3*kBins]*/, sigsvgd_bw_state* state) {
  extern __shared__ float sv[];  // kSample floats
  __shared__ unsigned int hist[kBins];
  __shared__ unsigned long long scratch[1024];
  __shared__ uint32_t sh_bucket;
  __shared__ unsigned long long sh_k;
  for (int i = threadIdx.x; i < kSample; i += blockDim.x) sv[i] = sample[i];
  for (int i = threadIdx.x; i < 3 * kBins; i += blockDim.x) ghist[i] = 0ull;
  __syncthreads();
  const float lo = block_select_smem(sv, kSample, kSample / 2 - kSampleHalfWin, hist, scratch,
                                     &sh_bucket, &sh_k);
  const float hi = block_select_smem(sv, kSample, kSample / 2 + kSampleHalfWin, hist, scratch,
                                     &sh_bucket, &sh_k);
  if (threadIdx.x == 0) {
    win->lo = lo;
    win->hi = hi;
    win->n_cand = 0u;
    win->cap = cap;
    win->below = 0ull;
    win->active = 1u;
    win->fallback = 0u;
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

// cooperative: exact order statistic k among the candidates (or the whole matrix on fallback)
__global__ void __launch_bounds__(1024)
window_select_kernel(SelWindow* win, const float* __restrict__ cand, const float* __restrict__ dist,
                     int64_t M, int64_t N, int64_t ldd, unsigned long long k_total,
                     unsigned long long* ghist, float log_np1, float bw_scale, float tol,
                     sigsvgd_bw_state* state) {
  cg::grid_group grid = cg::this_grid();
  __shared__ unsigned int hist[kBins];
  __shared__ unsigned long long scratch[1024];
  __shared__ uint32_t sh_bucket;
  __shared__ unsigned long long sh_k;
  const unsigned int n_cand = win->n_cand;
  const unsigned long long below = win->below;
  const bool ok = n_cand <= win->cap && k_total >= below && k_total < below + n_cand;
  unsigned long long k = ok ? k_total - below : k_total;
  const int64_t count = ok ? (int64_t)n_cand : M * N;
  uint32_t prefix = 0;
  for (int pass = 0; pass < 3; ++pass) {
    for (int i = threadIdx.x; i < kBins; i += blockDim.x) hist[i] = 0u;
    __syncthreads();
    const int64_t step = (int64_t)gridDim.x * blockDim.x;
    // whole warps iterate together (hist_add is warp-collective): loop on the warp's first index
    for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31); i0 < count; i0 += step) {
      const int64_t i = i0 + (threadIdx.x & 31);
      float v = 0.f;
      const bool inb = i < count;
      if (inb) {
        if (ok) {
          v = __ldg(cand + i);
        } else {
          const int64_t r = i / N;
          v = __ldg(dist + r * ldd + (i - r * N));
        }
      }
      uint32_t bin;
      const bool in = radix_bin(__float_as_uint(v), pass, prefix, &bin) && inb;
      hist_add(hist, bin, in);
    }
    __syncthreads();
    unsigned long long* g = ghist + pass * kBins;
    for (int i = threadIdx.x; i < kBins; i += blockDim.x)
      if (hist[i]) atomicAdd(&g[i], (unsigned long long)hist[i]);
    grid.sync();
    // every block picks the bucket redundantly from the global histogram (no second barrier)
    block_pick_bucket(g, k, &sh_bucket, &sh_k, scratch);
    const int shift = pass == 0 ? 21 : pass == 1 ? 10 : 0;
    prefix |= sh_bucket << shift;
    k = sh_k;
    __syncthreads();
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    sigsvgd_bw_state& s = *state;
    const float med = __uint_as_float(prefix);
    float h = bw_scale * sqrtf(med / log_np1);  // math.py:31-34
    const bool clamped = !(h >= tol);
    if (clamped) h = tol;
    s.median = med;
    s.h = h;
    s.gamma = -0.5f / (h * h);  // _kernels.py:110
    s.clamped = clamped ? 1u : 0u;
    s.prefix = prefix;
    win->fallback = ok ? 0u : 1u;
  }
}

}  // namespace

size_t window_workspace_bytes(int64_t M, int64_t N, unsigned int* cap_out) {
  // candidates: ~4.7 % of the entries expected; room for 12 % plus slack
  const double want = 0.12 * (double)M * (double)N + 65536.0;
  const unsigned int cap = (unsigned int)std::min<double>(want, 4.0e9);
  if (cap_out) *cap_out = cap;
  return (size_t)round_up((int64_t)sizeof(SelWindow), 256) + (size_t)round_up(sizeof(float) * kSample, 256) +
         (size_t)round_up(sizeof(unsigned long long) * 3 * kBins, 256) + sizeof(float) * (size_t)cap;
}

// ws layout: [SelWindow | sample | ghist | cand]
static void window_carve(void* ws, SelWindow** win, float** sample, unsigned long long** ghist,
                         float** cand) {
  char* w = (char*)ws;
  *win = (SelWindow*)w;
  w += round_up((int64_t)sizeof(SelWindow), 256);
  *sample = (float*)w;
  w += round_up(sizeof(float) * kSample, 256);
  *ghist = (unsigned long long*)w;
  w += round_up(sizeof(unsigned long long) * 3 * kBins, 256);
  *cand = (float*)w;
}

void window_pointers(void* ws, SelWindow** win, float** cand) {
  float* sample;
  unsigned long long* ghist;
  window_carve(ws, win, &sample, &ghist, cand);
}

int launch_window_prepare(const float* a, int64_t lda, int64_t M, const float* b, int64_t ldb,
                          int64_t N, int64_t D, void* ws, sigsvgd_bw_state* state, SelWindow** win_out,
                          float** cand_out, cudaStream_t st) {
  SIG_CHECK_ARG(M >= 1 && N >= 1 && D >= 1 && lda >= D && ldb >= D, "window_prepare: bad shape");
  SelWindow* win;
  float *sample, *cand;
  unsigned long long* ghist;
  window_carve(ws, &win, &sample, &ghist, &cand);
  unsigned int cap = 0;
  window_workspace_bytes(M, N, &cap);
  window_sample_kernel<<<kSample / 8, 256, 0, st>>>(a, lda, M, b, ldb, N, D, sample);
  SIG_LAUNCH_OK("window_sample_kernel");
  static thread_local int configured_dev = -1;
  int dev = 0;
  SIG_CUDA_OK(cudaGetDevice(&dev));
  if (configured_dev != dev) {
    SIG_CUDA_OK(cudaFuncSetAttribute(window_bounds_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)(sizeof(float) * kSample)));
    configured_dev = dev;
  }
  window_bounds_kernel<<<1, 1024, sizeof(float) * kSample, st>>>(sample, win, cap, ghist, state);
  SIG_LAUNCH_OK("window_bounds_kernel");
  *win_out = win;
  *cand_out = cand;
  return 0;
}

int launch_window_select(void* ws, const float* dist, int64_t M, int64_t N, int64_t ldd,
                         int64_t n_log, float bw_scale, float tol, sigsvgd_bw_state* state,
                         cudaStream_t st) {
  SIG_CHECK_ARG(n_log >= 1, "window_select: n_log must be >= 1");
  SelWindow* win;
  float *sample, *cand;
  unsigned long long* ghist;
  window_carve(ws, &win, &sample, &ghist, &cand);
  const float log_np1 = (float)log((double)(float)((double)n_log + 1.0));
  unsigned long long k_total = (unsigned long long)((M * N - 1) / 2);
  const float* cand_c = cand;
  void* args[] = {(void*)&win,   (void*)&cand_c,   (void*)&dist, (void*)&M,   (void*)&N,    (void*)&ldd,
                  (void*)&k_total, (void*)&ghist, (void*)&log_np1, (void*)&bw_scale, (void*)&tol, (void*)&state};
  // one block per SM: co-resident by construction (cooperative launch checks it)
  SIG_CUDA_OK(cudaLaunchCooperativeKernel((const void*)window_select_kernel, dim3((unsigned)num_sms()),
                                          dim3(1024), args, 0, st));
  return 0;
}

}  // namespace sigsvgd
