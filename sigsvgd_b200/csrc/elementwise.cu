// HBM-bound elementwise / packing kernels around the two contractions.
//
//  rbf_apply : K = exp(gamma * d) in place (+ sum K*d, + median tie detection)
//              reference: src/kernels/_kernels.py:110 ; tie set = the support of torch.median's
//              backward
//  col_mean  : column means of the signature features (for mean-centring before the Gram;
//              distances and the repulsive term are translation invariant)
//  pack_v    : V = [ score | S - mean | 1 ]  -- the right-hand side of the single K.V contraction
//              that yields K.score, K.S and the row sums of K at once
//  sig_grad  : closed-form d(sum K)/dS incl. the gradient through the median bandwidth
#include "common.cuh"

namespace sigsvgd {
namespace {

constexpr int kEwThreads = 256;

__device__ __forceinline__ double block_sum_double(double v, double* sm) {
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) sm[w] = v;
  __syncthreads();
  double s = 0.0;
  if (threadIdx.x == 0)
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) s += sm[i];
  return s;
}

// grid: (col blocks, row blocks, batch); each thread handles 4 consecutive columns
__global__ void __launch_bounds__(kEwThreads)
rbf_apply_kernel(float* __restrict__ k, int64_t ldk, int64_t M, int64_t N, int64_t row_offset,
                 const float* __restrict__ weight, int64_t ldw, sigsvgd_bw_state* __restrict__ state,
                 int64_t stride_k, int64_t stride_w, int rows_per_block) {
  __shared__ double red[kEwThreads / 32];
  const int b = blockIdx.z;
  sigsvgd_bw_state& s = state[b];
  const float gamma = s.gamma;
  const uint32_t med_bits = __float_as_uint(s.median);
  const bool track_ties = s.median > 0.f;
  float* __restrict__ kb = k + (int64_t)b * stride_k;
  const float* __restrict__ wb = weight ? weight + (int64_t)b * stride_w : nullptr;
  const int64_t c0 = ((int64_t)blockIdx.x * kEwThreads + threadIdx.x) * 4;
  const bool vec = ((ldk & 3) == 0) && ((reinterpret_cast<uintptr_t>(kb) & 15u) == 0) && (c0 + 3 < N);
  double acc = 0.0;
  const int64_t r_begin = (int64_t)blockIdx.y * rows_per_block;
  const int64_t r_end = min(M, r_begin + rows_per_block);
  if (c0 < N) {
    for (int64_t r = r_begin; r < r_end; ++r) {
      float d[4], kv[4];
      float* row = kb + r * ldk + c0;
      const int nv = (int)min((int64_t)4, N - c0);
      if (vec) {
        const float4 q = *reinterpret_cast<const float4*>(row);
        d[0] = q.x, d[1] = q.y, d[2] = q.z, d[3] = q.w;
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e) d[e] = e < nv ? row[e] : 0.f;
      }
      float part = 0.f;
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        kv[e] = expf(gamma * d[e]);
        if (wb && e < nv) kv[e] *= __ldg(wb + r * ldw + c0 + e);
        if (e < nv) {
          part = fmaf(kv[e], d[e], part);
          if (track_ties && (__float_as_uint(d[e]) & 0x7fffffffu) == med_bits) {
            const uint32_t slot = atomicAdd(&s.n_ties, 1u);
            if (slot < SIGSVGD_MAX_TIES) {
              s.ties[2 * slot] = (int32_t)(row_offset + r);
              s.ties[2 * slot + 1] = (int32_t)(c0 + e);
            }
          }
        }
      }
      acc += (double)part;
      if (vec) {
        *reinterpret_cast<float4*>(row) = make_float4(kv[0], kv[1], kv[2], kv[3]);
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e)
          if (e < nv) row[e] = kv[e];
      }
    }
  }
  const double tot = block_sum_double(acc, red);
  if (threadIdx.x == 0 && tot != 0.0) atomicAdd(&s.sum_kd, tot);
}

// k_out = exp(gamma * dist), four columns per thread, no statistics (grid: col blocks x row blocks)
__global__ void __launch_bounds__(kEwThreads)
rbf_out_kernel(const float* __restrict__ d, int64_t ldd, int64_t M, int64_t N,
               const sigsvgd_bw_state* __restrict__ state, float* __restrict__ k, int64_t ldk,
               int rows_per_block) {
  const float gamma = state->gamma;
  const int64_t c0 = ((int64_t)blockIdx.x * kEwThreads + threadIdx.x) * 4;
  if (c0 >= N) return;
  const bool vin = ((ldd & 3) == 0) && ((reinterpret_cast<uintptr_t>(d) & 15u) == 0) && (c0 + 3 < N);
  const bool vout = ((ldk & 3) == 0) && ((reinterpret_cast<uintptr_t>(k) & 15u) == 0) && (c0 + 3 < N);
  const int nv = (int)min((int64_t)4, N - c0);
  const int64_t r_end = min(M, ((int64_t)blockIdx.y + 1) * rows_per_block);
  for (int64_t r = (int64_t)blockIdx.y * rows_per_block; r < r_end; ++r) {
    float v[4];
    if (vin) {
      const float4 q = *reinterpret_cast<const float4*>(d + r * ldd + c0);
      v[0] = q.x, v[1] = q.y, v[2] = q.z, v[3] = q.w;
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e) v[e] = e < nv ? d[r * ldd + c0 + e] : 0.f;
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) v[e] = expf(gamma * v[e]);
    if (vout) {
      *reinterpret_cast<float4*>(k + r * ldk + c0) = make_float4(v[0], v[1], v[2], v[3]);
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e)
        if (e < nv) k[r * ldk + c0 + e] = v[e];
    }
  }
}

constexpr int kMeanParts = 64;

// part[b][r][c] = sum over rows n = r, r+64, ... of s[b][n][c]
__global__ void __launch_bounds__(128)
col_part_kernel(const float* __restrict__ s, int64_t lds, int64_t N, int64_t D,
                float* __restrict__ part, int64_t stride_s) {
  const int64_t c = (int64_t)blockIdx.x * 128 + threadIdx.x;
  const int r = blockIdx.y, b = blockIdx.z;
  if (c >= D) return;
  const float* __restrict__ sb = s + (int64_t)b * stride_s;
  float a0 = 0.f, a1 = 0.f;
  int64_t n = r;
  for (; n + kMeanParts < N; n += 2 * kMeanParts) {
    a0 += __ldg(sb + n * lds + c);
    a1 += __ldg(sb + (n + kMeanParts) * lds + c);
  }
  if (n < N) a0 += __ldg(sb + n * lds + c);
  part[((int64_t)b * kMeanParts + r) * D + c] = a0 + a1;
}

__global__ void __launch_bounds__(128)
col_mean_kernel(const float* __restrict__ part, int64_t N, int64_t D, float* __restrict__ mean) {
  const int64_t c = (int64_t)blockIdx.x * 128 + threadIdx.x;
  const int b = blockIdx.z;
  if (c >= D) return;
  float a = 0.f;
  for (int r = 0; r < kMeanParts; ++r) a += part[((int64_t)b * kMeanParts + r) * D + c];
  mean[(int64_t)b * D + c] = a / (float)N;
}

// flat over B*N*C elements
__global__ void __launch_bounds__(kEwThreads)
pack_v_kernel(const float* __restrict__ score, int64_t P, const float* __restrict__ s, int64_t lds,
              const float* __restrict__ mean, int64_t N, int64_t D, float* __restrict__ v,
              int64_t ldv, int64_t C, int64_t rows_total) {
  const int64_t idx = (int64_t)blockIdx.x * kEwThreads + threadIdx.x;
  if (idx >= rows_total * C) return;
  const int64_t row = idx / C, c = idx - row * C;
  float val;
  if (c < P) {
    val = __ldg(score + row * P + c);
  } else if (c < P + D) {
    const int64_t cc = c - P;
    val = __ldg(s + row * lds + cc);
    if (mean) val -= __ldg(mean + (row / N) * D + cc);
  } else {
    val = (c == P + D) ? 1.f : 0.f;
  }
  v[row * ldv + c] = val;
}

// flat over B*M*D elements
__global__ void __launch_bounds__(kEwThreads)
sig_grad_kernel(const float* __restrict__ sc, int64_t lds, const float* __restrict__ sc_all,
                int64_t lds_all, const float* __restrict__ ks, int64_t ldks,
                const float* __restrict__ rowsum, int64_t ldr, int64_t M, int64_t D,
                int64_t n_all, int64_t row_offset, const sigsvgd_bw_state* __restrict__ state,
                float* __restrict__ g, int64_t ldg, int64_t rows_total) {
  const int64_t idx = (int64_t)blockIdx.x * kEwThreads + threadIdx.x;
  if (idx >= rows_total * D) return;
  const int64_t row = idx / D, c = idx - row * D;
  const int64_t b = row / M, i = row - b * M;
  const sigsvgd_bw_state& s = state[b];
  const float si = __ldg(sc + row * lds + c);
  // -(1/h^2) = 2*gamma
  float val = 2.f * s.gamma * (si * __ldg(rowsum + row * ldr) - __ldg(ks + row * ldks + c));
  if (s.coef != 0.f) {
    const uint32_t nl = min(s.n_ties, (uint32_t)SIGSVGD_MAX_TIES);
    const float w = s.coef / (float)s.n_ties_all;
    for (uint32_t t = 0; t < nl; ++t) {
      if ((int64_t)s.ties[2 * t] == row_offset + i) {
        const int64_t jb = s.ties[2 * t + 1];
        val = fmaf(w, si - __ldg(sc_all + (b * n_all + jb) * lds_all + c), val);
      }
    }
  }
  g[row * ldg + c] = val;
}

// ------------------------------------------------------------------------------------------
// Fused optimiser step on the particles: x <- update(x, g) with g = sign * dir
// (dir = phi with sign -1: the reference minimises with X.grad = -phi, svgd.py:82-83, 93-116).
//   SGD      : [buf <- mu*buf + g (buf = g on the first step)]  x -= lr * (buf | g)        torch.optim.SGD
//   ADAGRAD  : inertia += g^2;  g' = g / sqrt(inertia + 1e-12);  x -= lr * g'             svgd.py:110-113
//   ADAM     : m <- m + (g - m)(1 - b1);  v <- b2 v + (1 - b2) g^2;
//              x -= (lr / bc1) * m / (sqrt(v) / sqrt(bc2) + eps)                           torch.optim.Adam
// grad_out (optional) receives what the reference logs as the step's "grad" (g, or g' for Adagrad).
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kEwThreads)
opt_step_kernel(float* __restrict__ x, const float* __restrict__ dir, float sign, float* __restrict__ m,
                float* __restrict__ v, float* __restrict__ grad_out, int64_t n, int kind, float lr, float b1,
                float b2, float omb1, float omb2, float eps, float c1, float c2, int first) {
  for (int64_t i = (int64_t)blockIdx.x * kEwThreads + threadIdx.x; i < n; i += (int64_t)gridDim.x * kEwThreads) {
    const float g = sign * dir[i];
    float xi = x[i], logged = g;
    if (kind == SIGSVGD_OPT_SGD) {
      float step = g;
      if (m != nullptr) {
        step = first ? g : fmaf(b1, m[i], g);
        m[i] = step;
      }
      xi = fmaf(-lr, step, xi);
    } else if (kind == SIGSVGD_OPT_ADAGRAD) {
      const float acc = fmaf(g, g, v[i]);
      v[i] = acc;
      logged = g / sqrtf(acc + 1e-12f);
      xi = fmaf(-lr, logged, xi);
    } else {  // Adam; c1 = lr / (1 - b1^t), c2 = sqrt(1 - b2^t)
      const float mi = fmaf(g - m[i], omb1, m[i]);
      const float vi = fmaf(b2, v[i], omb2 * g * g);
      m[i] = mi;
      v[i] = vi;
      xi -= c1 * (mi / (sqrtf(vi) / c2 + eps));
    }
    x[i] = xi;
    if (grad_out != nullptr) grad_out[i] = logged;
  }
}

}  // namespace

int launch_rbf_apply(float* k, int64_t ldk, int64_t M, int64_t N, int64_t row_offset,
                     const float* weight, int64_t ldw, sigsvgd_bw_state* state, int64_t batch,
                     int64_t stride_k, int64_t stride_w, cudaStream_t st) {
  SIG_CHECK_ARG(M >= 0 && N >= 1 && ldk >= N && batch >= 1 && batch <= 65535,
                "rbf_apply: bad shape M=%lld N=%lld ldk=%lld", (long long)M, (long long)N,
                (long long)ldk);
  SIG_CHECK_ARG(weight == nullptr || ldw >= N, "rbf_apply: ldw < N");
  if (M == 0) return 0;
  const int64_t gx = ceil_div(N, (int64_t)kEwThreads * 4);
  // enough row blocks to fill the machine a few times over, at most 65535
  int64_t gy = std::min<int64_t>(65535, std::max<int64_t>(1, ceil_div((int64_t)num_sms() * 16, gx * batch)));
  gy = std::min(gy, M);
  const int rows_per_block = (int)ceil_div(M, gy);
  gy = ceil_div(M, rows_per_block);
  dim3 grid((unsigned)gx, (unsigned)gy, (unsigned)batch);
  rbf_apply_kernel<<<grid, kEwThreads, 0, st>>>(k, ldk, M, N, row_offset, weight, ldw, state,
                                                stride_k, stride_w, rows_per_block);
  SIG_LAUNCH_OK("rbf_apply_kernel");
  return 0;
}

int launch_opt_step(float* x, const float* dir, float sign, float* m, float* v, float* grad_out, int64_t n,
                    int kind, float lr, double b1, double b2, float eps, float c1, float c2, int first,
                    cudaStream_t st) {
  SIG_CHECK_ARG(kind == SIGSVGD_OPT_SGD || kind == SIGSVGD_OPT_ADAGRAD || kind == SIGSVGD_OPT_ADAM,
                "opt_step: unknown optimiser kind %d", kind);
  SIG_CHECK_ARG(kind != SIGSVGD_OPT_ADAGRAD || v != nullptr, "opt_step: Adagrad needs the inertia buffer v");
  SIG_CHECK_ARG(kind != SIGSVGD_OPT_ADAM || (m != nullptr && v != nullptr), "opt_step: Adam needs m and v");
  if (n <= 0) return 0;
  const int64_t grid = std::min<int64_t>(ceil_div(n, (int64_t)kEwThreads), (int64_t)num_sms() * 16);
  opt_step_kernel<<<(unsigned)grid, kEwThreads, 0, st>>>(
      x, dir, sign, m, v, grad_out, n, kind, lr, (float)b1, (float)b2, (float)(1.0 - b1), (float)(1.0 - b2), eps, c1,
      c2, first);
  SIG_LAUNCH_OK("opt_step_kernel");
  return 0;
}

int launch_rbf_out(const float* d, int64_t ldd, int64_t M, int64_t N, const sigsvgd_bw_state* state,
                   float* k, int64_t ldk, cudaStream_t st) {
  SIG_CHECK_ARG(M >= 0 && N >= 1 && ldd >= N && ldk >= N, "rbf_out: bad shape M=%lld N=%lld", (long long)M,
                (long long)N);
  if (M == 0) return 0;
  const int64_t gx = ceil_div(N, (int64_t)kEwThreads * 4);
  int64_t gy = std::min<int64_t>(65535, std::max<int64_t>(1, ceil_div((int64_t)num_sms() * 16, gx)));
  gy = std::min(gy, M);
  const int rows_per_block = (int)ceil_div(M, gy);
  gy = ceil_div(M, rows_per_block);
  rbf_out_kernel<<<dim3((unsigned)gx, (unsigned)gy), kEwThreads, 0, st>>>(d, ldd, M, N, state, k, ldk,
                                                                        rows_per_block);
  SIG_LAUNCH_OK("rbf_out_kernel");
  return 0;
}

int launch_col_mean(const float* s, int64_t lds, int64_t N, int64_t D, float* part, float* mean,
                    int64_t batch, int64_t stride_s, cudaStream_t st) {
  SIG_CHECK_ARG(N >= 1 && D >= 1 && lds >= D && batch >= 1 && batch <= 65535, "col_mean: bad shape");
  dim3 g1((unsigned)ceil_div(D, 128), kMeanParts, (unsigned)batch);
  col_part_kernel<<<g1, 128, 0, st>>>(s, lds, N, D, part, stride_s);
  SIG_LAUNCH_OK("col_part_kernel");
  dim3 g2((unsigned)ceil_div(D, 128), 1, (unsigned)batch);
  col_mean_kernel<<<g2, 128, 0, st>>>(part, N, D, mean);
  SIG_LAUNCH_OK("col_mean_kernel");
  return 0;
}

int launch_pack_v(const float* score, int64_t P, const float* s, int64_t lds, const float* mean,
                  int64_t N, int64_t D, float* v, int64_t ldv, int64_t C, int64_t batch,
                  cudaStream_t st) {
  SIG_CHECK_ARG(N >= 1 && D >= 1 && P >= 0 && lds >= D && C >= P + D + 1 && ldv >= C,
                "pack_v: bad shape N=%lld D=%lld P=%lld C=%lld", (long long)N, (long long)D,
                (long long)P, (long long)C);
  SIG_CHECK_ARG(P == 0 || score != nullptr, "pack_v: score is NULL but P > 0");
  const int64_t rows = batch * N;
  pack_v_kernel<<<(unsigned)ceil_div(rows * C, kEwThreads), kEwThreads, 0, st>>>(
      score, P, s, lds, mean, N, D, v, ldv, C, rows);
  SIG_LAUNCH_OK("pack_v_kernel");
  return 0;
}

int launch_sig_grad(const float* sc, int64_t lds, const float* sc_all, int64_t lds_all,
                    const float* ks, int64_t ldks, const float* rowsum, int64_t ldr, int64_t M,
                    int64_t D, int64_t n_all, int64_t row_offset, const sigsvgd_bw_state* state,
                    float* g, int64_t ldg, int64_t batch, cudaStream_t st) {
  SIG_CHECK_ARG(M >= 0 && D >= 1 && lds >= D && lds_all >= D && ldks >= D && ldr >= 1 && ldg >= D,
                "sig_grad: bad shape");
  if (M == 0) return 0;
  const int64_t rows = batch * M;
  sig_grad_kernel<<<(unsigned)ceil_div(rows * D, kEwThreads), kEwThreads, 0, st>>>(
      sc, lds, sc_all, lds_all, ks, ldks, rowsum, ldr, M, D, n_all, row_offset, state, g, ldg, rows);
  SIG_LAUNCH_OK("sig_grad_kernel");
  return 0;
}

}  // namespace sigsvgd
