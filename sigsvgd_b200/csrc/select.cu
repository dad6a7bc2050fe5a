// Exact lower-median selection + median-heuristic bandwidth (K4).
//
// Replaces torch.median(sq_dists) and the arithmetic of bw_median
// (reference: src/utils/math.py:28-34).  torch.median over a flattened tensor with n elements
// returns sorted element (n-1)//2; squared distances are >= 0, so their fp32 bit patterns order
// like uint32 and a 3-pass (11/11/10 bit) radix select finds the exact order statistic.
//
// The passes are separate launches so a row-block-sharded run can all-reduce the 2048-bin
// histogram between `hist` and `scan`; batched problems use blockIdx.y.
#include <mutex>

#include "common.cuh"

namespace sigsvgd {
namespace {

constexpr int kBins = 2048;
constexpr int kHistThreads = 256;

__device__ __forceinline__ bool radix_bin(uint32_t bits, int pass, uint32_t prefix, uint32_t* bin) {
  bits &= 0x7fffffffu;  // -0.0f sorts with +0.0f
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

__global__ void __launch_bounds__(kHistThreads)
select_hist_kernel(const float* __restrict__ dist, int64_t count, int64_t stride, int pass,
                   const sigsvgd_bw_state* __restrict__ state, unsigned long long* __restrict__ hist) {
  __shared__ unsigned int h[kBins];
  const int b = blockIdx.y;
  const float* __restrict__ src = dist + (int64_t)b * stride;
  for (int i = threadIdx.x; i < kBins; i += kHistThreads) h[i] = 0u;
  __syncthreads();
  const uint32_t prefix = state[b].prefix;
  const int64_t step = (int64_t)gridDim.x * kHistThreads;
  // vector body when the segment is 16-byte aligned
  const bool vec = ((reinterpret_cast<uintptr_t>(src) & 15u) == 0);
  const int64_t n4 = vec ? count / 4 : 0;
  for (int64_t i = (int64_t)blockIdx.x * kHistThreads + threadIdx.x; i < n4; i += step) {
    const uint4 q = __ldg(reinterpret_cast<const uint4*>(src) + i);
    uint32_t bin;
    if (radix_bin(q.x, pass, prefix, &bin)) atomicAdd(&h[bin], 1u);
    if (radix_bin(q.y, pass, prefix, &bin)) atomicAdd(&h[bin], 1u);
    if (radix_bin(q.z, pass, prefix, &bin)) atomicAdd(&h[bin], 1u);
    if (radix_bin(q.w, pass, prefix, &bin)) atomicAdd(&h[bin], 1u);
  }
  for (int64_t i = n4 * 4 + (int64_t)blockIdx.x * kHistThreads + threadIdx.x; i < count; i += step) {
    uint32_t bin;
    if (radix_bin(__float_as_uint(__ldg(src + i)), pass, prefix, &bin)) atomicAdd(&h[bin], 1u);
  }
  __syncthreads();
  unsigned long long* __restrict__ g = hist + (int64_t)b * kBins;
  for (int i = threadIdx.x; i < kBins; i += kHistThreads)
    if (h[i]) atomicAdd(&g[i], (unsigned long long)h[i]);
}

// Windowed first pass of the sharded median (dist.py): ONE read of this rank's strip that counts
// the entries below the window [lo, hi] (bounds from a sample of the global distance
// distribution, identical on every rank) and compacts the ~5 % of entries inside it into per-block
// pools.  The exact radix select then runs over the pools only (pool tails are padded with a
// value above every candidate), so the strip is read once instead of three times.
// totals[0] += #below, totals[1] += #candidates kept, totals[2] += #pools that overflowed.
constexpr int kCollectThreads = 512;
constexpr float kCollectPad = 3.0e38f;

__global__ void __launch_bounds__(kCollectThreads)
select_collect_kernel(const float* __restrict__ dist, int64_t count, const float* __restrict__ bounds,
                      float* __restrict__ cand, unsigned int poolcap, int64_t cand_cap,
                      unsigned long long* __restrict__ totals) {
  __shared__ unsigned int sh_fill;
  __shared__ unsigned long long sh_below;
  if (threadIdx.x == 0) sh_fill = 0u, sh_below = 0ull;
  __syncthreads();
  const uint32_t lo = __float_as_uint(__ldg(bounds)) & 0x7fffffffu;
  const uint32_t span = (__float_as_uint(__ldg(bounds + 1)) & 0x7fffffffu) - lo;
  float* __restrict__ pool = cand + (size_t)blockIdx.x * poolcap;
  const int lane = threadIdx.x & 31;
  unsigned int below = 0;
  auto push = [&](float v) {
    const uint32_t bits = __float_as_uint(v) & 0x7fffffffu;
    below += bits < lo ? 1u : 0u;
    const bool hit = bits - lo <= span;
    const unsigned int m = __ballot_sync(0xffffffffu, hit);
    if (m) {
      unsigned int base = 0;
      if (lane == __ffs(m) - 1) base = atomicAdd(&sh_fill, (unsigned int)__popc(m));
      base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
      const unsigned int slot = base + __popc(m & ((1u << lane) - 1u));
      if (hit && slot < poolcap) pool[slot] = v;
    }
  };
  const bool vec = (reinterpret_cast<uintptr_t>(dist) & 15u) == 0;
  const int64_t n4 = vec ? count / 4 : 0;
  const int64_t step = (int64_t)gridDim.x * kCollectThreads;
  // body: a warp takes 128 consecutive float4 (4 coalesced loads in flight per lane), builds a
  // 16-bit hit mask per lane, and reserves pool slots with ONE shared-memory atomic per 512 entries
  const int64_t n_chunks = (n4 + 127) / 128;
  // consecutive chunks go to different blocks, so small strips still spread over all the pools
  const int64_t warp_g = (int64_t)(threadIdx.x >> 5) * gridDim.x + blockIdx.x;
  const int64_t n_warps = (int64_t)gridDim.x * (kCollectThreads / 32);
  for (int64_t c = warp_g; c < n_chunks; c += n_warps) {
    float v[16];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t idx = c * 128 + j * 32 + lane;
      float4 q = make_float4(kCollectPad, kCollectPad, kCollectPad, kCollectPad);
      if (idx < n4) q = __ldg(reinterpret_cast<const float4*>(dist) + idx);
      v[4 * j] = q.x, v[4 * j + 1] = q.y, v[4 * j + 2] = q.z, v[4 * j + 3] = q.w;
    }
    uint32_t mask = 0;
#pragma unroll
    for (int e = 0; e < 16; ++e) {
      const uint32_t bits = __float_as_uint(v[e]) & 0x7fffffffu;
      below += bits < lo ? 1u : 0u;
      if (bits - lo <= span) mask |= 1u << e;
    }
    const int cnt = __popc(mask);
    int incl = cnt;
#pragma unroll
    for (int sft = 1; sft < 32; sft <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, sft);
      if (lane >= sft) incl += t;
    }
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    if (total) {
      unsigned int base = 0;
      if (lane == 31) base = atomicAdd(&sh_fill, (unsigned int)total);
      base = __shfl_sync(0xffffffffu, base, 31);
      unsigned int off = base + (unsigned int)(incl - cnt);
#pragma unroll
      for (int e = 0; e < 16; ++e) {
        if ((mask >> e) & 1u) {
          if (off < poolcap) pool[off] = v[e];
          ++off;
        }
      }
    }
  }
  const int64_t tail0 = n4 * 4, ntail = count - tail0;
  const int64_t ntail_up = (ntail + 31) / 32 * 32;
  for (int64_t i = (int64_t)blockIdx.x * kCollectThreads + threadIdx.x; i < ntail_up; i += step)
    push(i < ntail ? __ldg(dist + tail0 + i) : kCollectPad);
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) below += __shfl_xor_sync(0xffffffffu, below, o);
  if (lane == 0 && below) atomicAdd(&sh_below, (unsigned long long)below);
  __syncthreads();
  const unsigned int fill = sh_fill, kept = min(fill, poolcap);
  for (unsigned int i = kept + threadIdx.x; i < poolcap; i += kCollectThreads) pool[i] = kCollectPad;
  if (blockIdx.x == gridDim.x - 1)  // slots beyond the last pool (cand_cap is not a multiple of the pool size)
    for (int64_t i = (int64_t)gridDim.x * poolcap + threadIdx.x; i < cand_cap; i += kCollectThreads)
      cand[i] = kCollectPad;
  if (threadIdx.x == 0) {
    atomicAdd(&totals[0], sh_below);
    atomicAdd(&totals[1], (unsigned long long)kept);
    if (fill > poolcap) atomicAdd(&totals[2], 1ull);
  }
}

__global__ void select_begin_kernel(sigsvgd_bw_state* state, unsigned long long* hist,
                                    unsigned long long k) {
  const int b = blockIdx.x;
  for (int i = threadIdx.x; i < kBins; i += blockDim.x) hist[(int64_t)b * kBins + i] = 0ull;
  if (threadIdx.x == 0) {
    sigsvgd_bw_state& s = state[b];
    s.prefix = 0u;
    s.k_rem = k;
    s.median = 0.f;
    s.sum_kd = 0.0;
    s.n_ties = 0u;
    s.n_ties_all = 0u;
    s.clamped = 0u;
    s.coef = 0.f;
  }
}

// same, for a select over the candidates of a window (sigsvgd_select_collect): the rank is taken
// relative to the entries below the window, all on the device -- totals = {#below, #kept, #overflowed}
// summed over the ranks; *ok = 1 if the wanted rank lies inside the window and no pool overflowed
__global__ void select_begin_window_kernel(sigsvgd_bw_state* state, unsigned long long* hist,
                                           unsigned long long k_total, const unsigned long long* totals,
                                           int* ok) {
  for (int i = threadIdx.x; i < kBins; i += blockDim.x) hist[i] = 0ull;
  if (threadIdx.x == 0) {
    const unsigned long long below = totals[0], kept = totals[1], over = totals[2];
    const bool hit = over == 0ull && k_total >= below && k_total < below + kept;
    *ok = hit ? 1 : 0;
    sigsvgd_bw_state& s = state[0];
    s.prefix = 0u;
    s.k_rem = hit ? k_total - below : 0ull;
    s.median = 0.f;
    s.sum_kd = 0.0;
    s.n_ties = 0u;
    s.n_ties_all = 0u;
    s.clamped = 0u;
    s.coef = 0.f;
  }
}

// one block of 256 threads per problem
__global__ void __launch_bounds__(256)
select_scan_kernel(int pass, sigsvgd_bw_state* state, unsigned long long* hist) {
  __shared__ unsigned long long part[256];
  __shared__ int chunk_sel;
  __shared__ unsigned long long before_sel;
  const int b = blockIdx.x, t = threadIdx.x;
  unsigned long long* g = hist + (int64_t)b * kBins;
  unsigned long long loc[8], s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    loc[i] = g[t * 8 + i];
    s += loc[i];
  }
  part[t] = s;
  __syncthreads();
  if (t == 0) {
    const unsigned long long k = state[b].k_rem;
    unsigned long long cum = 0;
    int sel = 255;
    for (int c = 0; c < 256; ++c) {
      if (k < cum + part[c]) {
        sel = c;
        break;
      }
      cum += part[c];
    }
    if (sel == 255 && !(k < cum + part[255])) {  // k beyond the total: clamp to the last element
      cum = 0;
      for (int c = 0; c < 255; ++c) cum += part[c];
    }
    chunk_sel = sel;
    before_sel = cum;
  }
  __syncthreads();
  if (t == chunk_sel) {
    const unsigned long long k = state[b].k_rem;
    unsigned long long cum = before_sel;
    int bin = 7;
    for (int i = 0; i < 8; ++i) {
      if (k < cum + loc[i]) {
        bin = i;
        break;
      }
      if (i < 7) cum += loc[i];
    }
    const uint32_t bucket = (uint32_t)(t * 8 + bin);
    const int shift = pass == 0 ? 21 : pass == 1 ? 10 : 0;
    state[b].prefix |= bucket << shift;
    state[b].k_rem = k >= cum ? k - cum : 0ull;
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) g[t * 8 + i] = 0ull;
}

__global__ void select_finish_kernel(float log_np1, float bw_scale, float tol,
                                     sigsvgd_bw_state* state, int batch) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  sigsvgd_bw_state& s = state[b];
  const float med = __uint_as_float(s.prefix);
  // math.py:31-34: h = median / log(N+1); h = bw_scale * sqrt(h); clamp_min(tol)
  float h = bw_scale * sqrtf(med / log_np1);
  const bool clamped = !(h >= tol);
  if (clamped) h = tol;
  s.median = med;
  s.h = h;
  s.gamma = -0.5f / (h * h);  // _kernels.py:110
  s.clamped = clamped ? 1u : 0u;
}

__global__ void fixed_bw_kernel(float h, sigsvgd_bw_state* state, int batch) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  sigsvgd_bw_state& s = state[b];
  s.median = 0.f;
  s.h = h;
  s.gamma = -0.5f / (h * h);
  s.coef = 0.f;
  s.sum_kd = 0.0;
  s.n_ties = 0u;
  s.n_ties_all = 0u;
  s.clamped = 0u;
  s.prefix = 0u;
  s.k_rem = 0ull;
}

// coef = sum_kd / (h^2 * median)  
__global__ void median_coef_kernel(sigsvgd_bw_state* state, const double* sum_kd_all,
                                   const uint32_t* n_ties_all, int enable, int batch) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  sigsvgd_bw_state& s = state[b];
  const double skd = sum_kd_all ? sum_kd_all[b] : s.sum_kd;
  const uint32_t nt = n_ties_all ? n_ties_all[b] : s.n_ties;
  s.n_ties_all = nt;
  const bool on = enable && !s.clamped && s.median > 0.f && nt > 0u;
  s.coef = on ? (float)(skd / ((double)s.h * (double)s.h * (double)s.median)) : 0.f;
}

}  // namespace

size_t select_workspace_bytes(int64_t /*count*/, int64_t batch) {
  return (size_t)batch * kBins * sizeof(unsigned long long);
}

static int hist_grid(int64_t count) {
  const int64_t want = ceil_div(count, (int64_t)kHistThreads * 16);
  const int64_t cap = (int64_t)num_sms() * 8;
  return (int)std::max<int64_t>(1, std::min(want, cap));
}

int launch_select_begin(sigsvgd_bw_state* state, uint64_t* hist, uint64_t k, int64_t batch,
                        cudaStream_t st) {
  select_begin_kernel<<<(unsigned)batch, 256, 0, st>>>(state, (unsigned long long*)hist, k);
  SIG_LAUNCH_OK("select_begin_kernel");
  return 0;
}

int launch_select_begin_window(sigsvgd_bw_state* state, uint64_t* hist, uint64_t k_total,
                               const uint64_t* totals, int* ok, cudaStream_t st) {
  select_begin_window_kernel<<<1, 256, 0, st>>>(state, (unsigned long long*)hist, k_total,
                                                (const unsigned long long*)totals, ok);
  SIG_LAUNCH_OK("select_begin_window_kernel");
  return 0;
}

int launch_select_hist(const float* dist, int64_t count, int64_t batch, int64_t stride, int pass,
                       const sigsvgd_bw_state* state, uint64_t* hist, cudaStream_t st) {
  SIG_CHECK_ARG(pass >= 0 && pass <= 2, "select_hist: pass must be 0..2");
  SIG_CHECK_ARG(batch >= 1 && batch <= 65535, "select_hist: bad batch");
  if (count <= 0) return 0;
  int gx = hist_grid(count);
  if (batch > 1) gx = (int)std::max<int64_t>(1, std::min<int64_t>(gx, ceil_div((int64_t)num_sms() * 8, batch)));
  dim3 grid((unsigned)gx, (unsigned)batch);
  select_hist_kernel<<<grid, kHistThreads, 0, st>>>(dist, count, stride, pass, state,
                                                    (unsigned long long*)hist);
  SIG_LAUNCH_OK("select_hist_kernel");
  return 0;
}

int launch_select_scan(int pass, sigsvgd_bw_state* state, uint64_t* hist, int64_t batch,
                       cudaStream_t st) {
  SIG_CHECK_ARG(pass >= 0 && pass <= 2, "select_scan: pass must be 0..2");
  select_scan_kernel<<<(unsigned)batch, 256, 0, st>>>(pass, state, (unsigned long long*)hist);
  SIG_LAUNCH_OK("select_scan_kernel");
  return 0;
}

int launch_select_finish(int64_t n_log, float bw_scale, float tol, sigsvgd_bw_state* state,
                         int64_t batch, cudaStream_t st) {
  SIG_CHECK_ARG(n_log >= 1, "select_finish: n_log must be >= 1");
  // the reference evaluates ln(N+1) in float32 (math.py:32, a default-dtype 0-dim tensor)
  const float log_np1 = (float)log((double)(float)((double)n_log + 1.0));
  select_finish_kernel<<<(unsigned)ceil_div(batch, 128), 128, 0, st>>>(log_np1, bw_scale, tol,
                                                                       state, (int)batch);
  SIG_LAUNCH_OK("select_finish_kernel");
  return 0;
}

int launch_fixed_bandwidth(float h, sigsvgd_bw_state* state, int64_t batch, cudaStream_t st) {
  SIG_CHECK_ARG(h > 0.f, "fixed_bandwidth: h must be > 0");
  fixed_bw_kernel<<<(unsigned)ceil_div(batch, 128), 128, 0, st>>>(h, state, (int)batch);
  SIG_LAUNCH_OK("fixed_bw_kernel");
  return 0;
}

int launch_median_coef(sigsvgd_bw_state* state, const double* sum_kd_all,
                       const uint32_t* n_ties_all, int enable, int64_t batch, cudaStream_t st) {
  median_coef_kernel<<<(unsigned)ceil_div(batch, 128), 128, 0, st>>>(state, sum_kd_all, n_ties_all,
                                                                     enable, (int)batch);
  SIG_LAUNCH_OK("median_coef_kernel");
  return 0;
}

int launch_median_select(const float* dist, int64_t count, int64_t batch, int64_t stride,
                         int64_t n_log, float bw_scale, float tol, sigsvgd_bw_state* state,
                         void* ws, size_t ws_bytes, cudaStream_t st) {
  SIG_CHECK_ARG(count >= 1, "median_bandwidth: empty distance matrix");
  if (ws_bytes < select_workspace_bytes(count, batch)) {
    set_error("median_bandwidth: workspace too small (%zu < %zu)", ws_bytes,
              select_workspace_bytes(count, batch));
    return SIGSVGD_ERR_WORKSPACE;
  }
  uint64_t* hist = (uint64_t*)ws;
  SIG_PROPAGATE(launch_select_begin(state, hist, (uint64_t)((count - 1) / 2), batch, st));
  for (int pass = 0; pass < 3; ++pass) {
    SIG_PROPAGATE(launch_select_hist(dist, count, batch, stride, pass, state, hist, st));
    SIG_PROPAGATE(launch_select_scan(pass, state, hist, batch, st));
  }
  return launch_select_finish(n_log, bw_scale, tol, state, batch, st);
}

// ------------------------------------------------------------------------------------------
// The same exact median with its histogram scratch in a file-scope device array, for callers that
// may not allocate (sigsvgd_median_bandwidth_static; no counterpart in the reference).
// kStaticMaxBatch records x 2048 bins x 8 bytes = 256 KiB, held by every GPU that loads the library.
// In host code `g_static_hist` is only a stand-in for the symbol: the device address is looked up
// once per device (that device current) and that pointer is what the kernels get.
// ------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned long long g_static_hist[SIGSVGD_STATIC_MAX_BATCH * kBins];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_static_mu;
void* g_static_ptr[kMaxDevices] = {};
}  // namespace

static int static_hist_address(uint64_t** out) {
  int dev = -1;
  SIG_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) {
    set_error("median_bandwidth_static: device ordinal %d outside [0, %d)", dev, kMaxDevices);
    return SIGSVGD_ERR_UNSUPPORTED;
  }
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (g_static_ptr[dev] == nullptr) {
    void* p = nullptr;
    SIG_CUDA_OK(cudaGetSymbolAddress(&p, g_static_hist));  // of the CURRENT device's copy
    if (p == nullptr) {
      set_error("median_bandwidth_static: cudaGetSymbolAddress returned NULL on device %d", dev);
      return SIGSVGD_ERR_CUDA;
    }
    g_static_ptr[dev] = p;
  }
  *out = (uint64_t*)g_static_ptr[dev];
  return 0;
}

int launch_median_select_static(const float* dist, int64_t count, int64_t batch, int64_t stride,
                                int64_t n_log, float bw_scale, float tol, sigsvgd_bw_state* state,
                                cudaStream_t st) {
  SIG_CHECK_ARG(batch >= 1, "median_bandwidth_static: batch must be >= 1");
  if (batch > SIGSVGD_STATIC_MAX_BATCH) {
    set_error("median_bandwidth_static: batch %lld exceeds the static scratch (SIGSVGD_STATIC_MAX_BATCH = %d "
              "records of %d x 8 bytes); use sigsvgd_median_bandwidth with a caller workspace",
              (long long)batch, SIGSVGD_STATIC_MAX_BATCH, kBins);
    return SIGSVGD_ERR_WORKSPACE;
  }
  SIG_CHECK_ARG(batch == 1 || stride >= count, "median_bandwidth_static: stride < count");
  uint64_t* hist = nullptr;
  SIG_PROPAGATE(static_hist_address(&hist));
  // (select_begin zeroes the bins of the records in use: the array keeps the previous call's contents)
  return launch_median_select(dist, count, batch, stride, n_log, bw_scale, tol, state, hist,
                              (size_t)SIGSVGD_STATIC_MAX_BATCH * kBins * sizeof(uint64_t), st);
}

int select_collect_blocks() { return 4 * num_sms(); }

int launch_select_collect(const float* dist, int64_t count, const float* bounds, float* cand,
                          int64_t cand_cap, unsigned long long* totals, cudaStream_t st) {
  const int blocks = select_collect_blocks();
  SIG_CHECK_ARG(count >= 0 && cand_cap >= blocks, "select_collect: candidate buffer too small");
  const unsigned int poolcap = (unsigned int)std::min<int64_t>(cand_cap / blocks, 1ll << 30);
  SIG_CUDA_OK(cudaMemsetAsync(totals, 0, 3 * sizeof(unsigned long long), st));
  select_collect_kernel<<<blocks, kCollectThreads, 0, st>>>(dist, count, bounds, cand, poolcap, cand_cap, totals);
  SIG_LAUNCH_OK("select_collect_kernel");
  return 0;
}

}  // namespace sigsvgd
