// Signature-PDE (Goursat) kernel: Gram matrix and its gradient w.r.t. the first argument
// (SURVEY.md 8(f)-1).
//
// Replaces sigkernel.SigKernel(static_kernel, dyadic_order).compute_Gram(X.double(), Y.double()) and its
// backward (third party, sigkernel @ 3b2373982e12; reference call sites:
// src/kernels/_traj_kernels.py:198-206 SignatureKernel, src/inference/trajectory_svgd.py:55-69,
// src/inference/score.py:64-76) together with the static kernel the reference feeds it,
// BatchGaussianKernel.Gram_matrix (src/kernels/_traj_kernels.py:176-195: exp(-dist / h) over all point
// pairs, h = bw_median over the [A,B,M,N] distances with ln(A + 1), src/utils/math.py:28-34).
//
//   inc[s,t]  = G[s+1,t+1] + G[s,t] - G[s+1,t] - G[s,t+1]              coarse cell of the pair (a, b)
//   every coarse cell is 2^o x 2^o fine cells of increment p = inc / 4^o
//   K[0,:] = K[:,0] = 1;  K[i+1,j+1] = (K[i,j+1] + K[i+1,j]) (1 + p/2 + p^2/12) - K[i,j] (1 - p^2/12)
//   k(a,b)    = K[MM, NN]
//   backward:   W[s,t] = 4^-o sum_{fine (i,j) in (s,t)} K[i,j] Krev[i+1,j+1],  Krev = the solve on the
//               increments flipped in both directions, flipped back;
//               d k / d x_p = sum_{s,t} W[s,t] d inc[s,t] / d x_p   (analytic, bandwidth held fixed)
//
// Layout: ONE WARP PER PATH PAIR.  The static Gram of the pair is built in the warp's shared memory
// ([M,N] doubles, turned into the increments in place), never in HBM: the reference's [A,B,M,N] fp64
// tensor (2 GB at 256 x 256 x 64 x 64) does not exist here.  The PDE is swept in strips of 32 fine rows:
// lane l owns row 32*strip + l and is l columns behind the lane above (anti-diagonal wavefront); a
// cell needs its left value (own register), the value above (one shuffle) and the value above-left
// (the previous shuffle); the strip's last row is handed to the next strip through shared memory.
// All arithmetic is fp64, as the reference's (`.double()` at every call site).
//
// Bound: fp64 issue / latency of the dependent chain (3 DFMA + 1 shuffle pair per cell and lane);
// algorithmic HBM traffic is only 8 (A M D + B N D + A B) bytes.
#include "common.cuh"

namespace sigsvgd {
namespace {

constexpr int kPdeMaxWarps = 8;
constexpr int kPdeSmemBudget = 200 * 1024;
constexpr int kPdeBins = 2048;

// doubles of the parked reverse solve of one pair: [strips of 32 fine rows][NN + 31 steps][32 lanes]
__host__ __device__ inline size_t pde_r_doubles(int64_t MM, int64_t NN) {
  return (size_t)((MM + 31) / 32) * (size_t)(NN + 31) * 32;
}

struct PdeArgs {
  const double* x;  // [A,M,D]
  const double* y;  // [B,N,D]
  int A, B, M, N, D, order, kind;
  double h;                        // RBF: exp(-d/h); LINEAR: <x,y> * h   (used when state == NULL)
  const sigsvgd_bw_state* state;   // RBF bandwidth record (h) or NULL
};

__device__ __forceinline__ double pde_scale(const PdeArgs& p) {
  if (p.kind == SIGSVGD_PDE_LINEAR) return p.h;
  return 1.0 / (p.state ? (double)p.state->h : p.h);
}

__device__ __forceinline__ double static_value(const double* xs, const double* yt, int D, int kind, double sc) {
  double acc = 0.0;
  if (kind == SIGSVGD_PDE_LINEAR) {
    for (int k = 0; k < D; ++k) acc = fma(xs[k], yt[k], acc);
    return acc * sc;
  }
  for (int k = 0; k < D; ++k) {
    const double df = xs[k] - yt[k];
    acc = fma(df, df, acc);
  }
  return exp(-acc * sc);
}

// G[s,t] of the pair into g[M*N], then the increments (pre-divided by 4^o) in place: g[s*N + t], s < M-1, t < N-1
__device__ __forceinline__ void build_increments(double* g, const double* xs, const double* ys, const PdeArgs& p,
                                                 double sc, int lane) {
  const int M = p.M, N = p.N;
  for (int idx = lane; idx < M * N; idx += 32) {
    const int s = idx / N, t = idx - s * N;
    g[idx] = static_value(xs + s * p.D, ys + t * p.D, p.D, p.kind, sc);
  }
  __syncwarp();
  const double f = 1.0 / (double)(1ll << (2 * p.order));
  for (int s = 0; s + 1 < M; ++s) {
    for (int t0 = 0; t0 < N - 1; t0 += 32) {
      const int t = t0 + lane;
      double v = 0.0;
      if (t < N - 1) v = (g[(s + 1) * N + t + 1] + g[s * N + t] - g[(s + 1) * N + t] - g[s * N + t + 1]) * f;
      __syncwarp();
      if (t < N - 1) g[s * N + t] = v;
      __syncwarp();
    }
  }
}

// One Goursat sweep of the warp's pair.  `flip`: solve on the increments flipped in both directions.
// Visit(r, c, upleft, value) is called for every fine cell (r, c) with K at its lower corner (r, c) and the
// freshly computed K at its upper corner (r+1, c+1).  Returns K[MM, NN] (valid in every lane).
template <class Visit>
__device__ __forceinline__ double goursat_sweep(const double* inc, double* top, const PdeArgs& p, bool flip, int lane,
                                                Visit&& visit) {
  const int o = p.order, N = p.N;
  const int MM = (p.M - 1) << o, NN = (p.N - 1) << o;
  for (int j = lane; j <= NN; j += 32) top[j] = 1.0;
  __syncwarp();
  for (int r0 = 0; r0 < MM; r0 += 32) {
    const int r = r0 + lane;
    const bool row_ok = r < MM;
    const bool last_row = row_ok && (lane == 31 || r == MM - 1);
    const int rs = (flip ? MM - 1 - r : r) >> o;
    double cur = 1.0, up_prev = 1.0;
    for (int tau = 0; tau < NN + 31; ++tau) {
      const int c = tau - lane;
      double up = __shfl_up_sync(0xffffffffu, cur, 1);
      if (lane == 0) up = top[min(max(c + 1, 0), NN)];
      if (row_ok && c >= 0 && c < NN) {
        const int cs = (flip ? NN - 1 - c : c) >> o;
        const double q = inc[rs * N + cs];
        const double q2 = q * q * (1.0 / 12.0);
        const double val = fma(up + cur, 1.0 + 0.5 * q + q2, -up_prev * (1.0 - q2));
        visit(r, c, up_prev, val);
        cur = val;
        if (last_row) top[c + 1] = val;
      }
      up_prev = up;
    }
    __syncwarp();
  }
  return top[NN];
}

__global__ void __launch_bounds__(32 * kPdeMaxWarps)
sigpde_fwd_kernel(PdeArgs p, int warps, double* __restrict__ k_out) {
  extern __shared__ __align__(16) double sm[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int a = blockIdx.y, b = blockIdx.x * warps + w;
  const int NN = (p.N - 1) << p.order;
  double* xs = sm;                                       // [M*D], shared by the block's warps
  const int per_warp = p.N * p.D + p.M * p.N + NN + 1;
  double* ys = sm + p.M * p.D + (size_t)w * per_warp;    // [N*D]
  double* g = ys + p.N * p.D;                            // [M*N]
  double* top = g + p.M * p.N;                           // [NN+1]
  for (int i = threadIdx.x; i < p.M * p.D; i += blockDim.x) xs[i] = p.x[(size_t)a * p.M * p.D + i];
  __syncthreads();
  if (b >= p.B) return;
  for (int i = lane; i < p.N * p.D; i += 32) ys[i] = p.y[(size_t)b * p.N * p.D + i];
  __syncwarp();
  build_increments(g, xs, ys, p, pde_scale(p), lane);
  const double k = goursat_sweep(g, top, p, false, lane, [](int, int, double, double) {});
  if (lane == 0) k_out[(size_t)a * p.B + b] = k;
}


// grad_x[a, :, :] += sum_{b in this block's range} gout[a,b] d k(a,b) / d x^a  (grad_x zeroed by the launcher).
// Grid (b-chunks, A); the block's warps take the b's of the chunk in turn.
// scratch, per warp, in global memory (L2-resident): R = the reverse solve's K at the nodes the forward
// sweep pairs it with, then Wc [(M-1)(N-1)] = the per-coarse-cell sums (red.global.add.f64: one warp owns
// them, no contention).  R is stored in the order the wavefront produces it -- [strip][step][lane], node
// (r+1, c+1) of lane r % 32 at step c + r % 32 -- so that a warp's 32 stores of a step (and the mirrored
// loads of the forward sweep) are one contiguous 256-byte run instead of 32 rows of a matrix.
// Shared memory per warp as in the forward kernel.
__global__ void __launch_bounds__(32 * kPdeMaxWarps)
sigpde_bwd_kernel(PdeArgs p, int warps, int b_per_block, const double* __restrict__ gout,
                  double* __restrict__ scratch, double* __restrict__ grad_x) {
  extern __shared__ __align__(16) double sm[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int a = blockIdx.y;
  const int o = p.order, M = p.M, N = p.N, D = p.D;
  const int MM = (M - 1) << o, NN = (N - 1) << o;
  double* xs = sm;             // [M*D]
  double* gacc = xs + M * D;   // [M*D] gradient of this a, accumulated over the block's pairs
  const int per_warp = N * D + M * N + NN + 1;
  double* ys = gacc + M * D + (size_t)w * per_warp;
  double* g = ys + N * D;   // [M*N] increments
  double* top = g + M * N;  // [NN+1]
  const size_t r_doubles = pde_r_doubles(MM, NN);
  const size_t per_scratch = r_doubles + (size_t)(M - 1) * (N - 1);
  double* R = scratch + (((size_t)blockIdx.y * gridDim.x + blockIdx.x) * warps + w) * per_scratch;
  double* Wc = R + r_doubles;
  const size_t strip_pitch = (size_t)(NN + 31) * 32;
  // Kf at node (rho, kappa) of the reverse solve (1 on its first row / column)
  auto kf_at = [&](int rho, int kappa) -> double {
    if (rho == 0 || kappa == 0) return 1.0;
    const int r = rho - 1, c = kappa - 1;
    return R[(size_t)(r >> 5) * strip_pitch + (size_t)(c + (r & 31)) * 32 + (r & 31)];
  };
  for (int i = threadIdx.x; i < M * D; i += blockDim.x) {
    xs[i] = p.x[(size_t)a * M * D + i];
    gacc[i] = 0.0;
  }
  __syncthreads();
  const double sc = pde_scale(p);
  const int b_lo = blockIdx.x * b_per_block, b_hi = min(p.B, b_lo + b_per_block);
  for (int b = b_lo + w; b < b_hi; b += warps) {
    for (int i = lane; i < N * D; i += 32) ys[i] = p.y[(size_t)b * N * D + i];
    for (int i = lane; i < (M - 1) * (N - 1); i += 32) Wc[i] = 0.0;
    __syncwarp();
    build_increments(g, xs, ys, p, sc, lane);
    // reverse solve: Kf[r+1][c+1], parked in wavefront order
    goursat_sweep(g, top, p, true, lane, [&](int r, int c, double, double val) {
      R[(size_t)(r >> 5) * strip_pitch + (size_t)(c + (r & 31)) * 32 + (r & 31)] = val;
    });
    __syncwarp();
    // forward solve, pairing K at the lower corner of every fine cell with the reverse solve at the
    // mirrored node: GG[i,j] = K[i,j] * Krev[i+1,j+1] = K[i,j] * Kf[MM-1-i][NN-1-j]
    double run = 0.0;
    goursat_sweep(g, top, p, false, lane, [&](int r, int c, double lower, double) {
      run = fma(lower, kf_at(MM - 1 - r, NN - 1 - c), run);
      if ((((c + 1) >> o) != (c >> o)) || c == NN - 1) {  // last fine column of the coarse cell
        atomicAdd(&Wc[(r >> o) * (N - 1) + (c >> o)], run);
        run = 0.0;
      }
    });
    __syncwarp();
    // d inc[s,t]/d x_{s+1} = dG[s+1,t+1] - dG[s+1,t];  d inc[s,t]/d x_s = dG[s,t] - dG[s,t+1]
    // -> per node (q,t): a coefficient from up to four cells, times dG[q,t]/dx_q.  A lane owns the nodes of
    // its rows q = lane, lane + 32, ... and walks t: no cross-lane traffic.
    // (Wc sums the fine cells; the derivative of a fine increment is the coarse one over 4^o)
    const double go = gout[(size_t)a * p.B + b] / (double)(1ll << (2 * o));
    for (int q = lane; q < M; q += 32) {
      for (int k0 = 0; k0 < D; k0 += 8) {
        double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int t = 0; t < N; ++t) {
          double coef = 0.0;
          if (q >= 1) {
            if (t >= 1) coef += Wc[(q - 1) * (N - 1) + t - 1];
            if (t <= N - 2) coef -= Wc[(q - 1) * (N - 1) + t];
          }
          if (q <= M - 2) {
            if (t <= N - 2) coef += Wc[q * (N - 1) + t];
            if (t >= 1) coef -= Wc[q * (N - 1) + t - 1];
          }
          if (p.kind == SIGSVGD_PDE_LINEAR) {
            coef *= sc;
#pragma unroll
            for (int k = 0; k < 8; ++k)
              if (k0 + k < D) acc[k] = fma(coef, ys[t * D + k0 + k], acc[k]);
          } else {
            coef *= static_value(xs + q * D, ys + t * D, D, p.kind, sc) * (-2.0 * sc);
#pragma unroll
            for (int k = 0; k < 8; ++k)
              if (k0 + k < D) acc[k] = fma(coef, xs[q * D + k0 + k] - ys[t * D + k0 + k], acc[k]);
          }
        }
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (k0 + k < D) atomicAdd(&gacc[q * D + k0 + k], go * acc[k]);
      }
    }
    __syncwarp();
  }
  __syncthreads();
  for (int i = threadIdx.x; i < M * D; i += blockDim.x)
    if (gacc[i] != 0.0) atomicAdd(&grad_x[(size_t)a * M * D + i], gacc[i]);
}

// One radix pass of the exact lower median over ALL point-pair squared distances |x^a_s - y^b_t|^2
// (the reference's bw_median on the [A,B,M,N] tensor), recomputed on the fly instead of stored:
// fp64 distance -> fp32 bit pattern -> 11/11/10-bit digit, as select.cu's passes.
__global__ void __launch_bounds__(256)
sigpde_dist_hist_kernel(PdeArgs p, int pass, const sigsvgd_bw_state* __restrict__ state,
                        unsigned long long* __restrict__ hist) {
  __shared__ unsigned int h[kPdeBins];
  for (int i = threadIdx.x; i < kPdeBins; i += 256) h[i] = 0u;
  __syncthreads();
  const uint32_t prefix = state->prefix;
  const int64_t total = (int64_t)p.A * p.B * p.M * p.N;
  for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (int64_t)gridDim.x * 256) {
    const int t = (int)(i % p.N);
    int64_t rest = i / p.N;
    const int s = (int)(rest % p.M);
    rest /= p.M;
    const int b = (int)(rest % p.B), a = (int)(rest / p.B);
    const double* xs = p.x + ((size_t)a * p.M + s) * p.D;
    const double* yt = p.y + ((size_t)b * p.N + t) * p.D;
    double acc = 0.0;
    for (int k = 0; k < p.D; ++k) {
      const double df = xs[k] - yt[k];
      acc = fma(df, df, acc);
    }
    const uint32_t bits = __float_as_uint((float)acc) & 0x7fffffffu;
    uint32_t bin;
    bool in;
    if (pass == 0) {
      bin = bits >> 21, in = true;
    } else if (pass == 1) {
      bin = (bits >> 10) & 0x7ffu, in = (bits >> 21) == (prefix >> 21);
    } else {
      bin = bits & 0x3ffu, in = (bits >> 10) == (prefix >> 10);
    }
    if (in) atomicAdd(&h[bin], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kPdeBins; i += 256)
    if (h[i]) atomicAdd(&hist[i], (unsigned long long)h[i]);
}

int pde_check(const PdeArgs& p, const char* what) {
  SIG_CHECK_ARG(p.x && p.y, "%s: NULL pointer", what);
  SIG_CHECK_ARG(p.A >= 1 && p.B >= 1 && p.M >= 2 && p.N >= 2 && p.D >= 1, "%s: bad shape A=%d B=%d M=%d N=%d D=%d",
                what, p.A, p.B, p.M, p.N, p.D);
  SIG_CHECK_ARG(p.order >= 0 && p.order <= 10, "%s: dyadic order must be in [0, 10]", what);
  SIG_CHECK_ARG(p.kind == SIGSVGD_PDE_RBF || p.kind == SIGSVGD_PDE_LINEAR, "%s: unknown static kernel %d", what, p.kind);
  SIG_CHECK_ARG(p.state != nullptr || p.kind == SIGSVGD_PDE_LINEAR || p.h > 0.0, "%s: bandwidth h must be > 0", what);
  SIG_CHECK_ARG(((int64_t)(p.M - 1) << p.order) < (1 << 20) && ((int64_t)(p.N - 1) << p.order) < (1 << 20),
                "%s: refined grid too large", what);
  return 0;
}

// warps per block such that the block's shared memory fits; 0 if even one warp does not
int pde_warps(int64_t block_doubles, int64_t warp_doubles, int64_t want) {
  int64_t w = std::min<int64_t>(kPdeMaxWarps, std::max<int64_t>(1, want));
  while (w >= 1 && (block_doubles + w * warp_doubles) * 8 > kPdeSmemBudget) --w;
  return (int)w;
}

}  // namespace

static PdeArgs make_args(const double* x, const double* y, int64_t A, int64_t B, int64_t M, int64_t N, int64_t D,
                         int order, int kind, double h, const sigsvgd_bw_state* state) {
  PdeArgs p{};
  p.x = x, p.y = y, p.A = (int)A, p.B = (int)B, p.M = (int)M, p.N = (int)N, p.D = (int)D;
  p.order = order, p.kind = kind, p.h = h, p.state = state;
  return p;
}

int launch_sigpde_bandwidth(const double* x, const double* y, int64_t A, int64_t B, int64_t M, int64_t N,
                            int64_t D, float bw_scale, float tol, sigsvgd_bw_state* state, uint64_t* hist,
                            cudaStream_t st) {
  PdeArgs p = make_args(x, y, A, B, M, N, D, 0, SIGSVGD_PDE_LINEAR, 1.0, nullptr);
  SIG_CHECK_ARG(x && y && state && hist, "sigpde_bandwidth: NULL pointer");
  SIG_CHECK_ARG(A >= 1 && B >= 1 && M >= 1 && N >= 1 && D >= 1 && A < (1 << 30) && B < (1 << 30) && M < (1 << 30) &&
                    N < (1 << 30) && D < (1 << 30),
                "sigpde_bandwidth: bad shape");
  const int64_t total = A * B * M * N;
  SIG_PROPAGATE(launch_select_begin(state, hist, (uint64_t)((total - 1) / 2), 1, st));
  const int grid = (int)std::max<int64_t>(1, std::min<int64_t>(ceil_div(total, 256 * 8), (int64_t)num_sms() * 8));
  for (int pass = 0; pass < 3; ++pass) {
    sigpde_dist_hist_kernel<<<grid, 256, 0, st>>>(p, pass, state, (unsigned long long*)hist);
    SIG_LAUNCH_OK("sigpde_dist_hist_kernel");
    SIG_PROPAGATE(launch_select_scan(pass, state, hist, 1, st));
  }
  return launch_select_finish(A, bw_scale, tol, state, 1, st);  // ln(dist.shape[0] + 1) = ln(A + 1)
}

int launch_sigpde_fwd(const double* x, const double* y, int64_t A, int64_t B, int64_t M, int64_t N, int64_t D,
                      int order, int kind, double h, const sigsvgd_bw_state* state, double* k_out,
                      cudaStream_t st) {
  PdeArgs p = make_args(x, y, A, B, M, N, D, order, kind, h, state);
  SIG_PROPAGATE(pde_check(p, "sigpde_gram"));
  SIG_CHECK_ARG(k_out, "sigpde_gram: NULL output");
  const int64_t NN = (N - 1) << order;
  const int64_t per_warp = N * D + M * N + NN + 1;
  const int warps = pde_warps(M * D, per_warp, B);
  if (warps < 1) {
    set_error("sigpde_gram: M=%lld N=%lld D=%lld order=%d needs %lld bytes of shared memory per path pair (> %d)",
              (long long)M, (long long)N, (long long)D, order, (long long)((M * D + per_warp) * 8), kPdeSmemBudget);
    return SIGSVGD_ERR_UNSUPPORTED;
  }
  const size_t smem = (size_t)(M * D + warps * per_warp) * 8;
  if (smem > 48 * 1024)
    SIG_CUDA_OK(cudaFuncSetAttribute(sigpde_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kPdeSmemBudget));
  SIG_CHECK_ARG(A <= 65535, "sigpde_gram: A > 65535");
  dim3 grid((unsigned)ceil_div(B, warps), (unsigned)A);
  sigpde_fwd_kernel<<<grid, 32 * warps, smem, st>>>(p, warps, k_out);
  SIG_LAUNCH_OK("sigpde_fwd_kernel");
  return 0;
}

struct BwdPlan {
  int warps, b_per_block, chunks;
};

// warps per block (shared memory), and how the B partners of one a are cut over blocks so that the grid
// holds a few blocks per SM
static BwdPlan bwd_plan(int64_t A, int64_t B, int64_t M, int64_t N, int64_t D, int order) {
  const int64_t NN = (N - 1) << order;
  BwdPlan pl{};
  pl.warps = pde_warps(2 * M * D, N * D + M * N + NN + 1, B);
  if (pl.warps < 1) return pl;
  const int64_t want_blocks = 4 * (int64_t)num_sms();
  int64_t chunks = std::max<int64_t>(1, std::min<int64_t>(ceil_div(B, pl.warps), ceil_div(want_blocks, A)));
  // the reverse solve of every resident pair is parked in global memory: keep that scratch under 4 GiB
  const double per = 8.0 * ((double)pde_r_doubles((M - 1) << order, (N - 1) << order) + (double)(M - 1) * (N - 1));
  const double cap = 4.0 * 1024 * 1024 * 1024;
  while (chunks > 1 && per * A * chunks * pl.warps > cap) --chunks;
  while (pl.warps > 1 && per * A * chunks * pl.warps > cap) --pl.warps;
  if (per * A * chunks * pl.warps > 4 * cap) {
    pl.warps = -1;  // refined grid too large for the parked reverse solve
    return pl;
  }
  pl.b_per_block = (int)(ceil_div(ceil_div(B, chunks), pl.warps) * pl.warps);
  pl.chunks = (int)ceil_div(B, pl.b_per_block);
  return pl;
}

size_t sigpde_bwd_workspace_bytes(int64_t A, int64_t B, int64_t M, int64_t N, int64_t D, int order) {
  if (A < 1 || B < 1 || M < 2 || N < 2 || D < 1 || order < 0 || order > 10) return 0;
  const BwdPlan pl = bwd_plan(A, B, M, N, D, order);
  if (pl.warps < 1) return 0;
  const size_t per = pde_r_doubles((M - 1) << order, (N - 1) << order) + (size_t)(M - 1) * (N - 1);
  return (size_t)A * pl.chunks * pl.warps * per * 8;
}

int launch_sigpde_bwd(const double* x, const double* y, int64_t A, int64_t B, int64_t M, int64_t N, int64_t D,
                      int order, int kind, double h, const sigsvgd_bw_state* state, const double* gout,
                      double* grad_x, void* ws, size_t ws_bytes, cudaStream_t st) {
  PdeArgs p = make_args(x, y, A, B, M, N, D, order, kind, h, state);
  SIG_PROPAGATE(pde_check(p, "sigpde_gram_bwd"));
  SIG_CHECK_ARG(gout && grad_x && ws, "sigpde_gram_bwd: NULL pointer");
  SIG_CHECK_ARG(A <= 65535, "sigpde_gram_bwd: A > 65535");
  const BwdPlan pl = bwd_plan(A, B, M, N, D, order);
  if (pl.warps < 1) {
    set_error(pl.warps < 0 ? "sigpde_gram_bwd: A=%lld M=%lld N=%lld order=%d: the parked reverse solves exceed 16 GiB of scratch"
                           : "sigpde_gram_bwd: A=%lld M=%lld N=%lld order=%d does not fit the shared-memory resident pair state",
              (long long)A, (long long)M, (long long)N, order);
    return SIGSVGD_ERR_UNSUPPORTED;
  }
  if (ws_bytes < sigpde_bwd_workspace_bytes(A, B, M, N, D, order)) {
    set_error("sigpde_gram_bwd: workspace too small (%zu < %zu)", ws_bytes,
              sigpde_bwd_workspace_bytes(A, B, M, N, D, order));
    return SIGSVGD_ERR_WORKSPACE;
  }
  const int64_t NN = (N - 1) << order;
  const size_t smem = (size_t)(2 * M * D + pl.warps * (N * D + M * N + NN + 1)) * 8;
  if (smem > 48 * 1024)
    SIG_CUDA_OK(cudaFuncSetAttribute(sigpde_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kPdeSmemBudget));
  SIG_CUDA_OK(cudaMemsetAsync(grad_x, 0, sizeof(double) * A * M * D, st));
  dim3 grid((unsigned)pl.chunks, (unsigned)A);
  sigpde_bwd_kernel<<<grid, 32 * pl.warps, smem, st>>>(p, pl.warps, pl.b_per_block, gout, (double*)ws, grad_x);
  SIG_LAUNCH_OK("sigpde_bwd_kernel");
  return 0;
}

}  // namespace sigsvgd
