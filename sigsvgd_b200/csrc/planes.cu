// Operand preparation for the split-fp16 tensor-core GEMMs (gemm_tc.cu): power-of-two scaling,
// hi/lo fp16 planes, row norms.  (The RBF map is fused into the K.F kernel's operand path, kf3_tc.cu.)
//
//   tc_scales   : alpha_S, alpha_G (powers of two placing the largest |value| in [2^13, 2^14)),
//                 alpha_S^-2 for the distance epilogue, colscale[] for the K.V epilogue
//   tc_pack     : F = [ (S - mean)*alpha_S | 1 | score*alpha_G | 0.. ] as fp16 hi/lo planes,
//                 |F_S|^2 row norms (plane units), and the centred fp32 features
#include "gemm_tc.cuh"

namespace sigsvgd {
namespace {

constexpr int kThreads = 256;
constexpr float kKScale = 16384.f;  // 2^14: K in [0,1] -> [0, 2^14]

__device__ __forceinline__ void split_half(float x, __half& hi, __half& lo) {
  hi = __float2half_rn(x);
  lo = __float2half_rn(x - __half2float(hi));
}

// mx[b*2+0] = max |s - mean| over [N,D], mx[b*2+1] = max |score| over [N,P]   (float bits, >= 0)
__global__ void __launch_bounds__(kThreads)
absmax_kernel(const float* __restrict__ s, int64_t lds, const float* __restrict__ mean, int64_t N,
              int64_t D, const float* __restrict__ score, int64_t lsc, int64_t P,
              unsigned int* __restrict__ mx) {
  // a warp per row, lanes along the columns (coalesced, no index arithmetic per element)
  const int b = blockIdx.y;
  const int lane = threadIdx.x & 31, wpb = kThreads / 32;
  const float* __restrict__ mb = mean ? mean + (int64_t)b * D : nullptr;
  float ms = 0.f, mg = 0.f;
  for (int64_t r = (int64_t)blockIdx.x * wpb + (threadIdx.x >> 5); r < N; r += (int64_t)gridDim.x * wpb) {
    const float* __restrict__ sr = s + ((int64_t)b * N + r) * lds;
    for (int64_t c = lane; c < D; c += 32) {
      float v = __ldg(sr + c);
      if (mb) v -= __ldg(mb + c);
      ms = fmaxf(ms, fabsf(v));
    }
    if (score) {
      const float* __restrict__ gr = score + ((int64_t)b * N + r) * lsc;
      for (int64_t c = lane; c < P; c += 32) mg = fmaxf(mg, fabsf(__ldg(gr + c)));
    }
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    ms = fmaxf(ms, __shfl_xor_sync(0xffffffffu, ms, o));
    mg = fmaxf(mg, __shfl_xor_sync(0xffffffffu, mg, o));
  }
  if (lane == 0) {
    // NaN/inf inputs propagate through the GEMM anyway; the scale only needs finite maxima
    if (ms > 0.f && ms < INFINITY) atomicMax(mx + 2 * b, __float_as_uint(ms));
    if (mg > 0.f && mg < INFINITY) atomicMax(mx + 2 * b + 1, __float_as_uint(mg));
  }
}

__device__ __forceinline__ float pow2_scale(float m) {
  if (!(m > 0.f)) return 1.f;
  int e = ilogbf(m);  // m in [2^e, 2^(e+1))
  e = max(-100, min(100, e));
  return ldexpf(1.f, 13 - e);
}

// tcs[b*4 + {0,1,2,3}] = alpha_S, alpha_S^-2, alpha_G, 0 ; colscale[b*Cp + c]
__global__ void __launch_bounds__(kThreads)
scales_kernel(const unsigned int* __restrict__ mx, int64_t D, int64_t P, int64_t Cp,
              float* __restrict__ tcs, float* __restrict__ colscale) {
  const int b = blockIdx.x;
  const float as = pow2_scale(__uint_as_float(mx[2 * b]));
  const float ag = pow2_scale(__uint_as_float(mx[2 * b + 1]));
  if (threadIdx.x == 0) {
    tcs[4 * b + 0] = as;
    tcs[4 * b + 1] = (1.f / as) * (1.f / as);
    tcs[4 * b + 2] = ag;
    tcs[4 * b + 3] = 0.f;
  }
  for (int64_t c = threadIdx.x; c < Cp; c += kThreads) {
    float v = 0.f;
    if (c < D) v = 1.f / (kKScale * as);
    else if (c == D) v = 1.f / kKScale;
    else if (c < D + 1 + P) v = 1.f / (kKScale * ag);
    colscale[(int64_t)b * Cp + c] = v;
  }
}

// one warp per row
__global__ void __launch_bounds__(kThreads)
pack_planes_kernel(const float* __restrict__ score, int64_t lsc, int64_t P,
                   const float* __restrict__ s, int64_t lds, const float* __restrict__ mean,
                   int64_t N, int64_t D, const float* __restrict__ tcs, __half* __restrict__ fh,
                   __half* __restrict__ fl, int64_t ldp, int64_t Cp, float* __restrict__ norms,
                   float* __restrict__ sc, int64_t ldsc, int64_t rows_total) {
  const int64_t row = (int64_t)blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5);
  if (row >= rows_total) return;
  const int lane = threadIdx.x & 31;
  const int64_t b = row / N;
  const float as = tcs[4 * b], ag = tcs[4 * b + 2];
  const float* __restrict__ mrow = mean ? mean + b * D : nullptr;
  float nrm = 0.f;
  for (int64_t c0 = 2 * lane; c0 < Cp; c0 += 64) {
    float x[2];
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      const int64_t c = c0 + e;
      float v = 0.f;
      if (c < D) {
        v = __ldg(s + row * lds + c);
        if (mrow) v -= __ldg(mrow + c);
        if (sc) sc[row * ldsc + c] = v;
        v *= as;
      } else if (c == D) {
        v = 1.f;
      } else if (c < D + 1 + P) {
        v = __ldg(score + row * lsc + (c - D - 1)) * ag;
      }
      x[e] = v;
    }
    __half h0, l0, h1, l1;
    split_half(x[0], h0, l0);
    split_half(x[1], h1, l1);
    if (c0 < D) {
      const float r0 = __half2float(h0) + __half2float(l0);
      nrm = fmaf(r0, r0, nrm);
    }
    if (c0 + 1 < D) {
      const float r1 = __half2float(h1) + __half2float(l1);
      nrm = fmaf(r1, r1, nrm);
    }
    *reinterpret_cast<__half2*>(fh + row * ldp + c0) = __halves2half2(h0, h1);
    *reinterpret_cast<__half2*>(fl + row * ldp + c0) = __halves2half2(l0, l1);
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) nrm += __shfl_xor_sync(0xffffffffu, nrm, o);
  if (lane == 0 && norms) norms[row] = nrm;
}

// ---------------------------------------------------------------------------------------------
// fused front end of the tensor-core path: one statistics pass + one packing pass
//   stats[0..D)  = column sums of s (float atomics)      stats[D] = max |s| bits
//   stats[D+1]   = max |score| bits                       (zeroed by the launcher)
// The buffer holds kStatSlots such records back to back (slot pitch D + 2): producers that hammer
// it from a thousand blocks (the fused signature forward) spread their atomics over the slots;
// the consumer adds the slots up.
// The pack kernel derives the mean and the scales itself: alpha_S from the bound
// max|s - mean| <= max|s| + max|mean| (the planes keep >= 22 significant bits as long as the
// centred features are not 2^17 times smaller than that bound).
// ---------------------------------------------------------------------------------------------
constexpr int kStatThreads = 512;

__global__ void __launch_bounds__(kStatThreads)
stats_kernel(const float* __restrict__ s, int64_t lds, int64_t N, int64_t D,
             const float* __restrict__ score, int64_t lsc, int64_t P, int rows_per_block,
             int do_s, float* __restrict__ stats) {
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t r1 = min(N, r0 + rows_per_block);
  float ms = 0.f, mg = 0.f;
  for (int64_t c = threadIdx.x; do_s && c < D; c += kStatThreads) {
    float sum = 0.f;
    int64_t r = r0;
    for (; r + 4 <= r1; r += 4) {  // four independent loads in flight
      const float v0 = __ldg(s + r * lds + c), v1 = __ldg(s + (r + 1) * lds + c);
      const float v2 = __ldg(s + (r + 2) * lds + c), v3 = __ldg(s + (r + 3) * lds + c);
      sum += (v0 + v1) + (v2 + v3);
      ms = fmaxf(fmaxf(ms, fmaxf(fabsf(v0), fabsf(v1))), fmaxf(fabsf(v2), fabsf(v3)));
    }
    for (; r < r1; ++r) {
      const float v = __ldg(s + r * lds + c);
      sum += v;
      ms = fmaxf(ms, fabsf(v));
    }
    atomicAdd(stats + c, sum);
  }
  for (int64_t c = threadIdx.x; score && c < P; c += kStatThreads) {
    int64_t r = r0;
    for (; r + 4 <= r1; r += 4) {
      const float v0 = __ldg(score + r * lsc + c), v1 = __ldg(score + (r + 1) * lsc + c);
      const float v2 = __ldg(score + (r + 2) * lsc + c), v3 = __ldg(score + (r + 3) * lsc + c);
      mg = fmaxf(fmaxf(mg, fmaxf(fabsf(v0), fabsf(v1))), fmaxf(fabsf(v2), fabsf(v3)));
    }
    for (; r < r1; ++r) mg = fmaxf(mg, fabsf(__ldg(score + r * lsc + c)));
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    ms = fmaxf(ms, __shfl_xor_sync(0xffffffffu, ms, o));
    mg = fmaxf(mg, __shfl_xor_sync(0xffffffffu, mg, o));
  }
  if ((threadIdx.x & 31) == 0) {
    unsigned int* mx = reinterpret_cast<unsigned int*>(stats + D);
    if (ms > 0.f && ms < INFINITY) atomicMax(mx, __float_as_uint(ms));
    if (mg > 0.f && mg < INFINITY) atomicMax(mx + 1, __float_as_uint(mg));
  }
}

// one warp per row; every block derives mean / scales from `stats` (n_total rows contributed)
__global__ void __launch_bounds__(kThreads)
pack_stats_kernel(const float* __restrict__ score, int64_t lsc, int64_t P,
                  const float* __restrict__ s, int64_t lds, const float* __restrict__ stats,
                  int64_t n_total, int center, int64_t N, int64_t D, float* __restrict__ tcs,
                  float* __restrict__ colscale, __half* __restrict__ fh, __half* __restrict__ fl,
                  int64_t ldp, int64_t Cp, float* __restrict__ norms, float* __restrict__ sc,
                  int64_t ldsc, float* __restrict__ mean_out, int64_t c_lo, int64_t c_hi,
                  float* __restrict__ zero_out, int64_t ld_zero, int mean_in_smem) {
  // columns [c_lo, c_hi) (both even) of the planes are written; a NULL score leaves zeros in its
  // columns (the host-input pipeline packs the feature and the score columns in two launches)
  __shared__ float sh_mm[kThreads / 32];
  extern __shared__ float sh_mean[];  // [D] when mean_in_smem
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const float inv_n = center ? 1.f / (float)n_total : 0.f;
  const int64_t slot_pitch = D + 2;
  auto col_mean_of = [&](int64_t c) {
    float sum = 0.f;
#pragma unroll
    for (int k = 0; k < kStatSlots; ++k) sum += stats[k * slot_pitch + c];
    return sum * inv_n;
  };
  float mm = 0.f;  // max |mean|
  for (int64_t c = threadIdx.x; c < D; c += kThreads) {
    const float m = col_mean_of(c);
    if (mean_in_smem) sh_mean[c] = m;
    mm = fmaxf(mm, fabsf(m));
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) mm = fmaxf(mm, __shfl_xor_sync(0xffffffffu, mm, o));
  if (lane == 0) sh_mm[wid] = mm;
  __syncthreads();
  mm = 0.f;
#pragma unroll
  for (int i = 0; i < kThreads / 32; ++i) mm = fmaxf(mm, sh_mm[i]);
  unsigned int mx[2] = {0u, 0u};
#pragma unroll
  for (int k = 0; k < kStatSlots; ++k) {
    const unsigned int* m = reinterpret_cast<const unsigned int*>(stats + k * slot_pitch + D);
    mx[0] = max(mx[0], m[0]), mx[1] = max(mx[1], m[1]);  // non-negative float bits order like uints
  }
  const float as = pow2_scale(__uint_as_float(mx[0]) + mm);
  const float ag = pow2_scale(__uint_as_float(mx[1]));
  if (blockIdx.x == 0) {
    if (threadIdx.x == 0) {
      tcs[0] = as;
      tcs[1] = (1.f / as) * (1.f / as);
      tcs[2] = ag;
      tcs[3] = 0.f;
    }
    for (int64_t c = threadIdx.x; c < Cp; c += kThreads) {
      float v = 0.f;
      if (c < D) v = 1.f / (kKScale * as);
      else if (c == D) v = 1.f / kKScale;
      else if (c < D + 1 + P) v = 1.f / (kKScale * ag);
      colscale[c] = v;
    }
    if (mean_out)
      for (int64_t c = threadIdx.x; c < D; c += kThreads) mean_out[c] = col_mean_of(c);
  }
  // warps stride over the rows: the per-block prologue above (means of 8 slots, scales) is paid
  // once per ~2 blocks per SM instead of once per 8 rows
  for (int64_t row = (int64_t)blockIdx.x * (kThreads / 32) + wid; row < N;
       row += (int64_t)gridDim.x * (kThreads / 32)) {
  if (zero_out) {  // this row of the K.F result buffer (reduction target): saves a memset node
    float* zr = zero_out + row * ld_zero;
    if (((ld_zero | Cp) & 3) == 0 && (reinterpret_cast<uintptr_t>(zero_out) & 15u) == 0) {
      for (int64_t c = 4 * lane; c < Cp; c += 128) *reinterpret_cast<float4*>(zr + c) = make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
      for (int64_t c = lane; c < Cp; c += 32) zr[c] = 0.f;
    }
  }
  float nrm = 0.f;
  // four column pairs per lane and step: their loads are issued together (a warp owns a whole row,
  // so one pair per step made the row a chain of 14 exposed memory latencies)
  constexpr int kU = 4;
  for (int64_t c00 = c_lo + 2 * lane; c00 < c_hi; c00 += 64 * kU) {
    float raw[kU][2];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int64_t c = c00 + 64 * u + e;
        float v = 0.f;
        if (c00 + 64 * u < c_hi) {
          if (c < D) v = __ldg(s + row * lds + c);
          else if (c > D && c < D + 1 + P && score) v = __ldg(score + row * lsc + (c - D - 1));
        }
        raw[u][e] = v;
      }
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int64_t c0 = c00 + 64 * u;
      if (c0 >= c_hi) break;
      float x[2];
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int64_t c = c0 + e;
        float v = 0.f;
        if (c < D) {
          v = raw[u][e] - (mean_in_smem ? sh_mean[c] : col_mean_of(c));
          if (sc) sc[row * ldsc + c] = v;
          v *= as;
        } else if (c == D) {
          v = 1.f;
        } else if (c < D + 1 + P && score) {
          v = raw[u][e] * ag;
        }
        x[e] = v;
      }
      __half h0, l0, h1, l1;
      split_half(x[0], h0, l0);
      split_half(x[1], h1, l1);
      if (c0 < D) {
        const float r0 = __half2float(h0) + __half2float(l0);
        nrm = fmaf(r0, r0, nrm);
      }
      if (c0 + 1 < D) {
        const float r1 = __half2float(h1) + __half2float(l1);
        nrm = fmaf(r1, r1, nrm);
      }
      *reinterpret_cast<__half2*>(fh + row * ldp + c0) = __halves2half2(h0, h1);
      *reinterpret_cast<__half2*>(fl + row * ldp + c0) = __halves2half2(l0, l1);
    }
  }
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) nrm += __shfl_xor_sync(0xffffffffu, nrm, o);
  if (lane == 0 && norms) norms[row] = nrm;
  }
}

}  // namespace

int launch_tc_scales(const float* s, int64_t lds, const float* mean, int64_t N, int64_t D,
                     const float* score, int64_t lsc, int64_t P, int64_t Cp, unsigned int* mx,
                     float* tcs, float* colscale, int64_t batch, cudaStream_t st) {
  SIG_CHECK_ARG(N >= 1 && D >= 1 && P >= 0 && lds >= D && (P == 0 || (score && lsc >= P)) &&
                    Cp >= D + 1 + P && batch >= 1 && batch <= 65535,
                "tc_scales: bad shape N=%lld D=%lld P=%lld Cp=%lld", (long long)N, (long long)D,
                (long long)P, (long long)Cp);
  SIG_CUDA_OK(cudaMemsetAsync(mx, 0, sizeof(unsigned int) * 2 * batch, st));
  // rows of a scene are dealt to gx blocks of kThreads/32 warps; enough blocks overall to fill the machine
  const int64_t want = ceil_div(N, (int64_t)(kThreads / 32) * 2);
  const int64_t cap = std::max<int64_t>(1, ceil_div((int64_t)num_sms() * 8, batch));
  const int gx = (int)std::max<int64_t>(1, std::min<int64_t>(want, cap));
  absmax_kernel<<<dim3((unsigned)gx, (unsigned)batch), kThreads, 0, st>>>(s, lds, mean, N, D, score,
                                                                        lsc, P, mx);
  SIG_LAUNCH_OK("absmax_kernel");
  scales_kernel<<<(unsigned)batch, kThreads, 0, st>>>(mx, D, P, Cp, tcs, colscale);
  SIG_LAUNCH_OK("scales_kernel");
  return 0;
}

int launch_tc_pack(const float* score, int64_t lsc, int64_t P, const float* s, int64_t lds,
                   const float* mean, int64_t N, int64_t D, const float* tcs, __half* fh, __half* fl,
                   int64_t ldp, int64_t Cp, float* norms, float* sc, int64_t ldsc, int64_t batch,
                   cudaStream_t st) {
  SIG_CHECK_ARG(N >= 1 && D >= 1 && P >= 0 && lds >= D && Cp >= D + 1 + P && ldp >= Cp &&
                    (Cp & 7) == 0 && (ldp & 7) == 0 && (sc == nullptr || ldsc >= D),
                "tc_pack: bad shape N=%lld D=%lld P=%lld Cp=%lld ldp=%lld", (long long)N,
                (long long)D, (long long)P, (long long)Cp, (long long)ldp);
  SIG_CHECK_ARG(P == 0 || (score != nullptr && lsc >= P), "tc_pack: score is NULL but P > 0");
  const int64_t rows = batch * N;
  pack_planes_kernel<<<(unsigned)ceil_div(rows, kThreads / 32), kThreads, 0, st>>>(
      score, lsc, P, s, lds, mean, N, D, tcs, fh, fl, ldp, Cp, norms, sc, ldsc, rows);
  SIG_LAUNCH_OK("pack_planes_kernel");
  return 0;
}

int launch_tc_stats(const float* s, int64_t lds, int64_t N, int64_t D, const float* score,
                    int64_t lsc, int64_t P, float* stats, int zero_first, cudaStream_t st, int do_s) {
  SIG_CHECK_ARG(N >= 1 && D >= 1 && P >= 0 && lds >= D && (P == 0 || score == nullptr || lsc >= P),
                "tc_stats: bad shape N=%lld D=%lld P=%lld", (long long)N, (long long)D, (long long)P);
  if (zero_first) SIG_CUDA_OK(cudaMemsetAsync(stats, 0, sizeof(float) * kStatSlots * (D + 2), st));
  const int blocks = (int)std::max<int64_t>(1, std::min<int64_t>((int64_t)num_sms() * 2, ceil_div(N, 8)));
  const int rows_per_block = (int)ceil_div(N, blocks);
  stats_kernel<<<(unsigned)ceil_div(N, rows_per_block), kStatThreads, 0, st>>>(s, lds, N, D, score, lsc, P,
                                                                            rows_per_block, do_s, stats);
  SIG_LAUNCH_OK("stats_kernel");
  return 0;
}

int launch_tc_pack_stats(const float* score, int64_t lsc, int64_t P, const float* s, int64_t lds,
                         const float* stats, int64_t n_total, int center, int64_t N, int64_t D,
                         float* tcs, float* colscale, __half* fh, __half* fl, int64_t ldp, int64_t Cp,
                         float* norms, float* sc, int64_t ldsc, float* mean_out, cudaStream_t st,
                         int64_t c_lo, int64_t c_hi, float* zero_out, int64_t ld_zero) {
  SIG_CHECK_ARG(zero_out == nullptr || ld_zero >= Cp, "tc_pack_stats: ld_zero < Cp");
  if (c_hi < 0) c_hi = Cp;
  SIG_CHECK_ARG(c_lo >= 0 && c_hi <= Cp && (c_lo & 1) == 0 && (c_hi & 1) == 0, "tc_pack_stats: bad column range");
  SIG_CHECK_ARG(N >= 1 && D >= 1 && P >= 0 && lds >= D && Cp >= D + 1 + P && ldp >= Cp && (Cp & 7) == 0 &&
                    (ldp & 7) == 0 && (sc == nullptr || ldsc >= D) && n_total >= 1,
                "tc_pack_stats: bad shape N=%lld D=%lld P=%lld Cp=%lld ldp=%lld", (long long)N, (long long)D,
                (long long)P, (long long)Cp, (long long)ldp);
  SIG_CHECK_ARG(P == 0 || score == nullptr || lsc >= P, "tc_pack_stats: score pitch < P");
  const int mean_in_smem = D <= 10240 ? 1 : 0;  // 40 KB of dynamic shared memory at most
  const int64_t blocks = std::min<int64_t>(ceil_div(N, kThreads / 32), (int64_t)num_sms() * 4);
  pack_stats_kernel<<<(unsigned)blocks, kThreads, mean_in_smem ? sizeof(float) * D : 0, st>>>(
      score, lsc, P, s, lds, stats, n_total, center, N, D, tcs, colscale, fh, fl, ldp, Cp, norms, sc, ldsc,
      mean_out, c_lo, c_hi, zero_out, ld_zero, mean_in_smem);
  SIG_LAUNCH_OK("pack_stats_kernel");
  return 0;
}

}  // namespace sigsvgd
