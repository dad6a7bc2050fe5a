// Truncated path-signature forward / VJP kernels (K1, K2).
//
// Replaces signatory.signature(path, depth, basepoint=True) and its custom backward
// (reference call sites: src/kernels/_traj_kernels.py:125-126; backward reached from
// torch.autograd.grad at src/inference/svgd.py:43 / src/inference/score.py:55).
//
// Math: S <- S (x) exp(Delta_t), Horner form per level.  Backward is the
// reversible sweep S_{t-1} = S_t (x) exp(-Delta_t) with the adjoint of the Horner chain.
//
// Two implementations:
//  * depth 3, d <= 8 ("sig3"): register-resident.  A group of G = pow2 >= d lanes owns one
//    particle; lane i owns every tensor entry whose FIRST index is i (S1[i], S2[i,:], S3[i,:,:]),
//    so the forward update and all adjoint updates are lane-local; the only cross-lane traffic of
//    the backward is one d-vector reduce-scatter (log2 G shuffle stages) per time step.
//    Waypoints are staged through shared memory with coalesced loads; results leave through
//    shared memory with coalesced stores.
//  * generic (any d, depth <= 8): one warp per particle, state in shared memory, entry-parallel.
//
// Bound: HBM (4*N*(T*d + D) bytes forward, 4*N*(2*T*d + 2*D) backward) but in practice
// latency/issue bound at N = 4096.
#include "common.cuh"

namespace sigsvgd {

namespace {

// One warp per CTA: 32 / G particles.  N = 4096 particles of d = 7 give 1024 CTAs, ~7 per SM, so
// the per-SM load is balanced to 1 % (128-thread CTAs left 40 of the 148 SMs with half the work)
// and every phase of a CTA overlaps with the other resident CTAs' phases.
constexpr int kThreads = 32;
constexpr int kGroupPad = 8;  // floats between the particles' shared-memory blocks: bank spread

// Contiguous global -> shared copy, 16-byte vectors, up to 8 loads in flight per thread.
__device__ __forceinline__ void copy_in_f4(float* dst, const float* __restrict__ src, int n4, int tid) {
  const float4* __restrict__ s4 = reinterpret_cast<const float4*>(src);
  float4* d4 = reinterpret_cast<float4*>(dst);
  int i = tid;
  for (; i + 7 * kThreads < n4; i += 8 * kThreads) {
    float4 v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = __ldg(s4 + i + k * kThreads);
#pragma unroll
    for (int k = 0; k < 8; ++k) d4[i + k * kThreads] = v[k];
  }
  for (; i < n4; i += kThreads) d4[i] = __ldg(s4 + i);
}

__host__ __device__ __forceinline__ bool is_aligned16(const void* p) {
  return (reinterpret_cast<uintptr_t>(p) & 15u) == 0;
}

__host__ __device__ constexpr int group_size(int d) { return d <= 1 ? 1 : d <= 2 ? 2 : d <= 4 ? 4 : 8; }
__host__ __device__ constexpr int row_pad(int d) { return d <= 2 ? d : ((d + 3) & ~3); }

template <int d>
__device__ __forceinline__ void load_row(const float* __restrict__ row, float (&v)[d]) {
  constexpr int DP = row_pad(d);
  if constexpr (DP % 4 == 0) {
#pragma unroll
    for (int q = 0; q < DP / 4; ++q) {
      const float4 f = *reinterpret_cast<const float4*>(row + 4 * q);
      if (4 * q + 0 < d) v[4 * q + 0] = f.x;
      if (4 * q + 1 < d) v[4 * q + 1] = f.y;
      if (4 * q + 2 < d) v[4 * q + 2] = f.z;
      if (4 * q + 3 < d) v[4 * q + 3] = f.w;
    }
  } else if constexpr (DP == 2) {
    const float2 f = *reinterpret_cast<const float2*>(row);
    v[0] = f.x;
    v[1] = f.y;
  } else {
    v[0] = row[0];
  }
}

// Reduce-scatter of a d-vector over the G lanes of a group: lane l returns sum_lanes pd[l].
template <int G, int d>
__device__ __forceinline__ float group_reduce_scatter(const float (&pd)[d], int lane) {
  float v[G];
#pragma unroll
  for (int k = 0; k < G; ++k) v[k] = k < d ? pd[k < d ? k : 0] : 0.f;
#pragma unroll
  for (int w = G / 2; w >= 1; w >>= 1) {
    const bool up = (lane & w) != 0;
#pragma unroll
    for (int k = 0; k < w; ++k) {
      const float send = up ? v[k] : v[k + w];
      const float keep = up ? v[k + w] : v[k];
      v[k] = keep + __shfl_xor_sync(0xffffffffu, send, w);
    }
  }
  return v[0];
}

// ------------------------------------------------------------------------------------------
// depth-3 forward
// ------------------------------------------------------------------------------------------
constexpr int kSeg = 4;  // time segments of the time-parallel backward (sig3_bwd_seg_kernel)
constexpr int kSegThreads = 32 * kSeg;

template <int d>
__global__ void __launch_bounds__(kThreads)
sig3_fwd_kernel(const float* __restrict__ x, int64_t N, int T, int basepoint, int TC,
                float* __restrict__ sig, float* __restrict__ stats, const float* __restrict__ score,
                int64_t ld_score, int P, float* __restrict__ pref, int seg_len) {
  constexpr int G = group_size(d), DP = row_pad(d), PPC = kThreads / G;
  constexpr int D = d + d * d + d * d * d;
  extern __shared__ __align__(16) float smem[];
  const int tid = threadIdx.x, grp = tid / G, lane = tid % G;
  const int il = lane < d ? lane : d - 1;
  const int64_t p0 = (int64_t)blockIdx.x * PPC;
  const int npart = (int)min((int64_t)PPC, N - p0);
  const bool active = grp < npart && lane < d;
  const int gq = grp < npart ? grp : 0;
  const int pstride = TC * DP + kGroupPad;
  // whole paths resident and 16-byte aligned: the CTA's waypoints are ONE contiguous block
  const bool vec = TC >= T && ((T * d) & 3) == 0 && is_aligned16(x);

  float S1 = 0.f, S2[d], S3[d][d], xprev[d], xprev_i = 0.f;
#pragma unroll
  for (int j = 0; j < d; ++j) {
    S2[j] = 0.f;
    xprev[j] = 0.f;
#pragma unroll
    for (int c = 0; c < d; ++c) S3[j][c] = 0.f;
  }
  int next_ck = pref != nullptr ? seg_len : T;  // prefix checkpoints for the time-parallel backward

  for (int c0 = 0; c0 < T; c0 += TC) {
    const int len = min(TC, T - c0);
    __syncthreads();
    if (vec) {
      // all loads of the block in flight at once (batches of 8 x 16 B per thread), then the
      // scatter into the padded rows
      const float4* __restrict__ s4 = reinterpret_cast<const float4*>(x + p0 * (int64_t)T * d);
      const int per4 = (T * d) >> 2, n4 = npart * per4;
      for (int i0 = tid; i0 < n4; i0 += 8 * kThreads) {
        float4 v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (i0 + k * kThreads < n4) v[k] = __ldg(s4 + i0 + k * kThreads);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int i4 = i0 + k * kThreads;
          if (i4 < n4) {
            const int q = i4 / per4, e = (i4 - q * per4) * 4;
            float* dst = smem + q * pstride;
            const float f[4] = {v[k].x, v[k].y, v[k].z, v[k].w};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int t = (e + u) / d, c = (e + u) - t * d;
              dst[t * DP + c] = f[u];
            }
          }
        }
      }
    } else {
      for (int q = 0; q < npart; ++q) {
        const float* __restrict__ src = x + ((p0 + q) * (int64_t)T + c0) * d;
        float* dst = smem + q * pstride;
#pragma unroll 4
        for (int i = tid; i < len * d; i += kThreads) {
          const int t = i / d, c = i - t * d;
          dst[t * DP + c] = __ldg(src + i);
        }
      }
    }
    __syncthreads();
    const float* rows = smem + gq * pstride;
    for (int r = 0; r < len; ++r) {
      float xt[d];
      load_row<d>(rows + r * DP, xt);
      const float xi = rows[r * DP + il];
      float dl[d];
#pragma unroll
      for (int c = 0; c < d; ++c) {
        dl[c] = xt[c] - xprev[c];
        xprev[c] = xt[c];
      }
      const float di = xi - xprev_i;
      xprev_i = xi;
      if (!basepoint && c0 + r == 0) continue;  // first point only seeds xprev
      const float u3h = 0.5f * fmaf(di, 1.f / 3.f, S1);
#pragma unroll
      for (int j = 0; j < d; ++j) {
        // level 3 as packed fp32 FMAs (fma.rn.f32x2): the kernel is issue-bound
        const float b = fmaf(u3h, dl[j], S2[j]);
        const float2 b2 = make_float2(b, b);
#pragma unroll
        for (int c = 0; c + 1 < d; c += 2) {
          const float2 t = __ffma2_rn(b2, make_float2(dl[c], dl[c + 1]), make_float2(S3[j][c], S3[j][c + 1]));
          S3[j][c] = t.x, S3[j][c + 1] = t.y;
        }
        if (d & 1) S3[j][d - 1] = fmaf(b, dl[d - 1], S3[j][d - 1]);
      }
      const float u2 = fmaf(di, 0.5f, S1);
#pragma unroll
      for (int j = 0; j < d; ++j) S2[j] = fmaf(u2, dl[j], S2[j]);
      S1 += di;
      if (c0 + r + 1 == next_ck && next_ck < T) {  // S after time c0+r = end state of a segment
        if (active) {
          // levels 1-2 only: the backward sweep never reads the level-3 state
          float* __restrict__ o =
              pref + ((p0 + grp) * (kSeg - 1) + (next_ck / seg_len - 1)) * (int64_t)(d + d * d);
          o[lane] = S1;
#pragma unroll
          for (int j = 0; j < d; ++j) o[d + lane * d + j] = S2[j];
        }
        next_ck += seg_len;
      }
    }
  }
  // stage the result through shared memory for coalesced stores
  __syncthreads();
  if (active) {
    float* o = smem + grp * D;
    o[lane] = S1;
#pragma unroll
    for (int j = 0; j < d; ++j) {
      o[d + lane * d + j] = S2[j];
#pragma unroll
      for (int c = 0; c < d; ++c) o[d + d * d + (lane * d + j) * d + c] = S3[j][c];
    }
  }
  __syncthreads();
  if (stats != nullptr) {
    // statistics of the tensor-core operand preparation (planes.cu: stats_kernel) while the block's
    // signatures are still in shared memory: column sums, max |s|, and max |score| of its rows
    // (one of the 8 slots of the buffer per block: 1024 blocks on the same 399 addresses cost
    // 9 us of serialised L2 atomics with a single slot)
    stats += (blockIdx.x & 7) * (D + 2);
    float ms = 0.f, mg = 0.f;
    for (int c = tid; c < D; c += kThreads) {
      float sum = 0.f;
      for (int q = 0; q < npart; ++q) {
        const float v = smem[q * D + c];
        sum += v;
        ms = fmaxf(ms, fabsf(v));
      }
      atomicAdd(stats + c, sum);
    }
    if (score != nullptr) {
      for (int q = 0; q < npart; ++q) {
        const float* __restrict__ row = score + (p0 + q) * ld_score;
        for (int i0 = tid; i0 < P; i0 += 8 * kThreads) {
          float v[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) v[k] = i0 + k * kThreads < P ? __ldg(row + i0 + k * kThreads) : 0.f;
#pragma unroll
          for (int k = 0; k < 8; ++k) mg = fmaxf(mg, fabsf(v[k]));
        }
      }
    }
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
      ms = fmaxf(ms, __shfl_xor_sync(0xffffffffu, ms, o));
      mg = fmaxf(mg, __shfl_xor_sync(0xffffffffu, mg, o));
    }
    if ((tid & 31) == 0) {
      unsigned int* mx = reinterpret_cast<unsigned int*>(stats + D);
      if (ms > 0.f && ms < INFINITY) atomicMax(mx, __float_as_uint(ms));
      if (mg > 0.f && mg < INFINITY) atomicMax(mx + 1, __float_as_uint(mg));
    }
  }
  float* __restrict__ dst = sig + p0 * D;
  const int nout = npart * D;
  if ((nout & 3) == 0 && is_aligned16(dst)) {
    float4* __restrict__ d4 = reinterpret_cast<float4*>(dst);
    const float4* s4 = reinterpret_cast<const float4*>(smem);
    for (int i = tid; i < (nout >> 2); i += kThreads) d4[i] = s4[i];
  } else {
    for (int i = tid; i < nout; i += kThreads) dst[i] = smem[i];
  }
}

// ------------------------------------------------------------------------------------------
// depth-3 forward, time-parallel: warp q of the CTA sweeps time segment q of the CTA's particles
// from the identity (R_q = signature of that piece of the path), then Chen's identity stitches
// them: P_1 = R_0 (x) R_1 and Q = R_2 (x) R_3 on two warps at once, S_T = P_1 (x) Q.  The
// level-1-2 checkpoints the time-parallel backward wants (P_0, P_1, P_2 = P_1 (x) R_2) fall out
// of the stitching.  Sequential chain: T/4 steps + two tensor products instead of T steps.
// ------------------------------------------------------------------------------------------
// C = A (x) B in place; lane owns the entries of A whose first index is il; B (all D entries of one
// particle) is read from shared memory.
template <int d>
__device__ __forceinline__ void chen_mul(float& A1, float (&A2)[d], float (&A3)[d][d],
                                         const float* __restrict__ B, int il) {
  constexpr int L12 = d + d * d;
  float B1[d];
#pragma unroll
  for (int k = 0; k < d; ++k) B1[k] = B[k];
#pragma unroll
  for (int j = 0; j < d; ++j) {
#pragma unroll
    for (int k = 0; k < d; ++k)
      A3[j][k] += fmaf(A2[j], B1[k], fmaf(A1, B[d + j * d + k], B[L12 + (il * d + j) * d + k]));
  }
#pragma unroll
  for (int j = 0; j < d; ++j) A2[j] += fmaf(A1, B1[j], B[d + il * d + j]);
  A1 += B[il];
}

template <int d>
__global__ void __launch_bounds__(kSegThreads)
sig3_fwd_seg_kernel(const float* __restrict__ x, int64_t N, int T, int basepoint, int seg_len,
                    float* __restrict__ sig, float* __restrict__ stats, const float* __restrict__ score,
                    int64_t ld_score, int P, float* __restrict__ pref) {
  constexpr int G = group_size(d), DP = row_pad(d), PPC = 32 / G;
  constexpr int D = d + d * d + d * d * d, L12 = d + d * d;
  extern __shared__ __align__(16) float smem[];
  const int tid = threadIdx.x, seg = tid >> 5, l32 = tid & 31, grp = l32 / G, lane = l32 % G;
  const int il = lane < d ? lane : d - 1;
  const int64_t p0 = (int64_t)blockIdx.x * PPC;
  const int npart = (int)min((int64_t)PPC, N - p0);
  const bool active = grp < npart && lane < d;
  const int gq = grp < npart ? grp : 0;
  const int pstride = (T + 1) * DP + kGroupPad;
  float* rbuf = smem;                       // [kSeg][PPC][D]; slot 0 doubles as the output staging
  float* xrows = smem + kSeg * PPC * D;     // [PPC][pstride], row r <-> time r-1 (row 0 = basepoint)

  {
    const float4* __restrict__ s4 = reinterpret_cast<const float4*>(x + p0 * (int64_t)T * d);
    const int per4 = (T * d) >> 2;
    for (int q0 = 0; q0 < npart; q0 += 4) {
      for (int i = tid; i < per4; i += kSegThreads) {
        float4 v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (q0 + k < npart) v[k] = __ldg(s4 + (q0 + k) * per4 + i);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if (q0 + k < npart) {
            float* dst = xrows + (q0 + k) * pstride + DP;
            const float f[4] = {v[k].x, v[k].y, v[k].z, v[k].w};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int t = (i * 4 + u) / d, c = (i * 4 + u) - t * d;
              dst[t * DP + c] = f[u];
            }
          }
        }
      }
    }
    for (int i = tid; i < npart * DP; i += kSegThreads) xrows[(i / DP) * pstride + (i % DP)] = 0.f;
  }
  __syncthreads();

  // ---- R_q: this warp's segment from the identity ----
  float S1 = 0.f, S2[d], S3[d][d];
#pragma unroll
  for (int j = 0; j < d; ++j) {
    S2[j] = 0.f;
#pragma unroll
    for (int c = 0; c < d; ++c) S3[j][c] = 0.f;
  }
  const int t_lo = seg * seg_len, t_hi = min(T, t_lo + seg_len);
  if (t_lo < T) {
    const float* rows = xrows + gq * pstride;
    float xprev[d];
    load_row<d>(rows + t_lo * DP, xprev);
    float xprev_i = rows[t_lo * DP + il];
    for (int t = t_lo; t < t_hi; ++t) {
      float xt[d], dl[d];
      load_row<d>(rows + (t + 1) * DP, xt);
      const float xi = rows[(t + 1) * DP + il];
#pragma unroll
      for (int c = 0; c < d; ++c) {
        dl[c] = xt[c] - xprev[c];
        xprev[c] = xt[c];
      }
      const float di = xi - xprev_i;
      xprev_i = xi;
      if (!basepoint && t == 0) continue;  // first point only seeds xprev
      const float u3 = S1 + di * (1.f / 3.f);
#pragma unroll
      for (int j = 0; j < d; ++j) {
        const float b = S2[j] + u3 * dl[j] * 0.5f;
#pragma unroll
        for (int c = 0; c < d; ++c) S3[j][c] = fmaf(b, dl[c], S3[j][c]);
      }
      const float u2 = S1 + di * 0.5f;
#pragma unroll
      for (int j = 0; j < d; ++j) S2[j] = fmaf(u2, dl[j], S2[j]);
      S1 += di;
    }
  }
  auto put = [&](float* o) {  // this lane's entries of (S1, S2, S3) into a D-vector
    if (active) {
      o[lane] = S1;
#pragma unroll
      for (int j = 0; j < d; ++j) {
        o[d + lane * d + j] = S2[j];
#pragma unroll
        for (int c = 0; c < d; ++c) o[L12 + (lane * d + j) * d + c] = S3[j][c];
      }
    }
  };
  auto put12 = [&](float* __restrict__ o) {  // levels 1-2 only (a checkpoint)
    if (active) {
      o[lane] = S1;
#pragma unroll
      for (int j = 0; j < d; ++j) o[d + lane * d + j] = S2[j];
    }
  };
  float* __restrict__ ck = pref ? pref + (p0 + gq) * (int64_t)(kSeg - 1) * L12 : nullptr;
  if (seg != 0 && seg != 2) put(rbuf + (seg * PPC + gq) * D);  // B operands of the first round
  if (seg == 2) put(rbuf + (2 * PPC + gq) * D);                // R_2: also needed for P_2
  if (seg == 0 && ck && seg_len < T) put12(ck);                // P_0 = R_0
  __syncthreads();
  // ---- first round: P_1 = R_0 (x) R_1 on warp 0, Q = R_2 (x) R_3 on warp 2 ----
  if (seg == 0) {
    chen_mul<d>(S1, S2, S3, rbuf + (1 * PPC + gq) * D, il);
    if (ck && 2 * seg_len < T) put12(ck + L12);                // P_1
  } else if (seg == 2) {
    chen_mul<d>(S1, S2, S3, rbuf + (3 * PPC + gq) * D, il);
    __syncwarp();
    put(rbuf + (3 * PPC + gq) * D);                             // Q replaces R_3
  }
  __syncthreads();
  // ---- second round (warp 0): P_2 = P_1 (x) R_2 (levels 1-2), S_T = P_1 (x) Q ----
  if (seg == 0) {
    if (ck && 3 * seg_len < T && active) {
      const float* R2 = rbuf + (2 * PPC + gq) * D;
      float* __restrict__ o = ck + 2 * L12;
      o[lane] = S1 + R2[lane];
#pragma unroll
      for (int j = 0; j < d; ++j) o[d + lane * d + j] = S2[j] + fmaf(S1, R2[j], R2[d + lane * d + j]);
    }
    chen_mul<d>(S1, S2, S3, rbuf + (3 * PPC + gq) * D, il);
    put(rbuf + gq * D);  // output staging: slot 0, [PPC][D] contiguous
  }
  __syncthreads();

  if (stats != nullptr) {
    // statistics of the operand preparation, as in sig3_fwd_kernel
    stats += (blockIdx.x & 7) * (D + 2);
    float ms = 0.f, mg = 0.f;
    for (int c = tid; c < D; c += kSegThreads) {
      float sum = 0.f;
      for (int q = 0; q < npart; ++q) {
        const float v = rbuf[q * D + c];
        sum += v;
        ms = fmaxf(ms, fabsf(v));
      }
      atomicAdd(stats + c, sum);
    }
    if (score != nullptr) {
      for (int q = 0; q < npart; ++q) {
        const float* __restrict__ row = score + (p0 + q) * ld_score;
        for (int i0 = tid; i0 < P; i0 += 4 * kSegThreads) {
          float v[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) v[k] = i0 + k * kSegThreads < P ? __ldg(row + i0 + k * kSegThreads) : 0.f;
#pragma unroll
          for (int k = 0; k < 4; ++k) mg = fmaxf(mg, fabsf(v[k]));
        }
      }
    }
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
      ms = fmaxf(ms, __shfl_xor_sync(0xffffffffu, ms, o));
      mg = fmaxf(mg, __shfl_xor_sync(0xffffffffu, mg, o));
    }
    if (l32 == 0) {
      unsigned int* mx = reinterpret_cast<unsigned int*>(stats + D);
      if (ms > 0.f && ms < INFINITY) atomicMax(mx, __float_as_uint(ms));
      if (mg > 0.f && mg < INFINITY) atomicMax(mx + 1, __float_as_uint(mg));
    }
  }
  float* __restrict__ dst = sig + p0 * D;
  const int nout = npart * D;
  if ((nout & 3) == 0 && is_aligned16(dst)) {
    float4* __restrict__ d4 = reinterpret_cast<float4*>(dst);
    const float4* s4 = reinterpret_cast<const float4*>(rbuf);
    for (int i = tid; i < (nout >> 2); i += kSegThreads) d4[i] = s4[i];
  } else {
    for (int i = tid; i < nout; i += kSegThreads) dst[i] = rbuf[i];
  }
}

// One reversed time step of the depth-3 VJP for lane i of a particle's group:
//   levels 1-2 of the state go back  S_{t-1} = S_t (x) exp(-dl)  (level 3 feeds nothing),
//   the adjoints (A1, A2; A3 = G3 is constant) go back through  S_t = S_{t-1} (x) exp(dl),
// and the return value is this lane's component of dL/d(dl_t).
template <int G, int d>
__device__ __forceinline__ float sig3_bwd_step(const float (&dl)[d], float di, float& S1, float (&S2)[d],
                                               float& A1, float (&A2)[d], const float (&g3)[d][d],
                                               int lane) {
  // The two 7x7 matrix-vector products with the constant G3 rows dominate the step; they are issued
  // as packed fp32 FMAs (fma.rn.f32x2: two lanes of the pair per instruction) to halve their issue
  // slots -- the kernel is issue-bound.
  constexpr int dh = d / 2;
  const float u2r = fmaf(di, -0.5f, S1);
#pragma unroll
  for (int j = 0; j < d; ++j) S2[j] = fmaf(-u2r, dl[j], S2[j]);
  S1 -= di;
  float hdl[d];
  float2 pd2[dh > 0 ? dh : 1], dl2[dh > 0 ? dh : 1];
  float pd_last = 0.f;
#pragma unroll
  for (int c = 0; c < d; ++c) hdl[c] = 0.5f * dl[c];
#pragma unroll
  for (int c = 0; c < dh; ++c) {
    pd2[c] = make_float2(0.f, 0.f);
    dl2[c] = make_float2(dl[2 * c], dl[2 * c + 1]);
  }
  const float u3 = fmaf(di, 1.f / 3.f, S1), u2 = fmaf(di, 0.5f, S1), u3h = 0.5f * u3;
  float sum_u = 0.f, sum_v = 0.f;  // sum_j rr_j dl_j / 2  and  sum_j A2_j dl_j
  float extra[d];                  // the diagonal terms added to pd[j]
#pragma unroll
  for (int j = 0; j < d; ++j) {
    float2 r2 = make_float2(0.f, 0.f);
#pragma unroll
    for (int c = 0; c < dh; ++c) r2 = __ffma2_rn(make_float2(g3[j][2 * c], g3[j][2 * c + 1]), dl2[c], r2);
    float rr = r2.x + r2.y;
    if (d & 1) rr = fmaf(g3[j][d - 1], dl[d - 1], rr);
    const float B2 = fmaf(u3, hdl[j], S2[j]);
    const float2 B22 = make_float2(B2, B2);
#pragma unroll
    for (int c = 0; c < dh; ++c) pd2[c] = __ffma2_rn(make_float2(g3[j][2 * c], g3[j][2 * c + 1]), B22, pd2[c]);
    if (d & 1) pd_last = fmaf(g3[j][d - 1], B2, pd_last);
    extra[j] = fmaf(rr, u3h, A2[j] * u2);
    sum_u = fmaf(rr, hdl[j], sum_u);
    sum_v = fmaf(A2[j], dl[j], sum_v);
    A2[j] += rr;
  }
  float pd[d];
#pragma unroll
  for (int c = 0; c < dh; ++c) {
    pd[2 * c] = pd2[c].x + extra[2 * c];
    pd[2 * c + 1] = pd2[c].y + extra[2 * c + 1];
  }
  if (d & 1) pd[d - 1] = pd_last + extra[d - 1];
  const float dI = fmaf(sum_u, 1.f / 3.f, fmaf(sum_v, 0.5f, A1));
  A1 += sum_u + sum_v;
  return group_reduce_scatter<G, d>(pd, lane) + dI;
}

// ------------------------------------------------------------------------------------------
// depth-3 backward (VJP) with fused  out = beta * gx + alpha * add
// ------------------------------------------------------------------------------------------
template <int d>
__global__ void __launch_bounds__(kThreads)
sig3_bwd_kernel(const float* __restrict__ x, const float* __restrict__ sig,
                const float* __restrict__ gsig, int64_t N, int T, int basepoint, int TC,
                const float* __restrict__ add, int64_t ld_add, float alpha, float beta,
                float* __restrict__ out, const SigGradSpec gs_all) {
  constexpr int G = group_size(d), DP = row_pad(d), PPC = kThreads / G;
  constexpr int D = d + d * d + d * d * d;
  extern __shared__ __align__(16) float smem[];
  __shared__ float carry[PPC * DP];
  const int tid = threadIdx.x, grp = tid / G, lane = tid % G;
  const int il = lane < d ? lane : d - 1;
  const int64_t p0 = (int64_t)blockIdx.x * PPC;
  const SigGradSpec gs = scene_view(gs_all, p0);  // (a CTA's particles lie in one scene)
  const int npart = (int)min((int64_t)PPC, N - p0);
  const bool active = grp < npart && lane < d;
  const int gq = grp < npart ? grp : 0;
  const int pstride = (TC + 1) * DP + kGroupPad;
  const bool vec = TC >= T && ((T * d) & 3) == 0 && is_aligned16(x) && is_aligned16(out) &&
                   (add == nullptr || (is_aligned16(add) && (ld_add & 3) == 0));

  constexpr int L12 = d + d * d;  // the sweep needs levels 1-2 of the state only (level 3 of S
                                  // feeds nothing: its adjoint is the constant G3)
  float S1 = 0.f, S2[d], A1 = 0.f, A2[d], g3[d][d];
  // ---- load S_T (levels 1-2) and dL/dS_T (coalesced through smem) ----
  {
    const float* __restrict__ src = sig + p0 * D;
    const float* __restrict__ gsrc = gsig + p0 * D;
    float* gstage = smem + PPC * D;
    const int nst = npart * D;
    for (int i = tid; i < npart * L12; i += kThreads) {
      const int q = i / L12;
      smem[i] = __ldg(src + q * D + (i - q * L12));
    }
    if (gs.sc != nullptr) {
      // the cotangent computed in place (see SigGradSpec)
      const float gam2 = 2.f * gs.state->gamma;  // -(1/h^2) = 2*gamma
      for (int i0 = tid; i0 < nst; i0 += 8 * kThreads) {
        float a[8], b[8], r[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int i = i0 + k * kThreads;
          if (i < nst) {
            const int q = i / D, c = i - q * D;
            const int64_t row = p0 + q;
            a[k] = __ldg(gs.sc + row * gs.lds + c);
            b[k] = __ldg(gs.ks + row * gs.ldks + c);
            r[k] = __ldg(gs.rowsum + row * gs.ldr);
          }
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int i = i0 + k * kThreads;
          if (i < nst) gstage[i] = gam2 * (a[k] * r[k] - b[k]);
        }
      }
      uint32_t nt_all;
      const float coef = grad_coef(gs, &nt_all);
      if (coef != 0.f) {
        __syncthreads();
        const uint32_t nl = min(gs.state->n_ties, (uint32_t)SIGSVGD_MAX_TIES);
        const float wt = coef / (float)nt_all;
        for (uint32_t t = 0; t < nl; ++t) {
          const int64_t a_loc = (int64_t)gs.state->ties[2 * t] - gs.row_offset - p0;
          if (a_loc < 0 || a_loc >= npart) continue;  // block-uniform
          const int64_t jb = gs.state->ties[2 * t + 1];
          for (int c = tid; c < D; c += kThreads)
            gstage[a_loc * D + c] = fmaf(wt, __ldg(gs.sc + (p0 + a_loc) * gs.lds + c) -
                                                 __ldg(gs.sc_all + jb * gs.lds_all + c),
                                         gstage[a_loc * D + c]);
          __syncthreads();
        }
      }
    } else if ((nst & 3) == 0 && is_aligned16(gsrc)) {
      copy_in_f4(gstage, gsrc, nst >> 2, tid);
    } else {
      for (int i = tid; i < nst; i += kThreads) gstage[i] = __ldg(gsrc + i);
    }
    __syncthreads();
    const float* o = smem + gq * L12;
    S1 = active ? o[il] : 0.f;
#pragma unroll
    for (int j = 0; j < d; ++j) S2[j] = active ? o[d + il * d + j] : 0.f;
    const float* og = gstage + gq * D;
    A1 = active ? og[il] : 0.f;
#pragma unroll
    for (int j = 0; j < d; ++j) {
      A2[j] = active ? og[d + il * d + j] : 0.f;
#pragma unroll
      for (int c = 0; c < d; ++c) g3[j][c] = active ? og[d + d * d + (il * d + j) * d + c] : 0.f;
    }
  }
  for (int i = tid; i < PPC * DP; i += kThreads) carry[i] = 0.f;

  const int t_first = basepoint ? 0 : 1;
  const int n_chunks = (T + TC - 1) / TC;
  for (int ch = n_chunks - 1; ch >= 0; --ch) {
    const int c0 = ch * TC;
    const int len = min(TC, T - c0);
    __syncthreads();
    // rows r = 0..len  <->  time c0-1+r ; row 0 is zero when c0 == 0 (the basepoint)
    if (vec) {
      const float4* __restrict__ s4 = reinterpret_cast<const float4*>(x + p0 * (int64_t)T * d);
      const int per4 = (T * d) >> 2, n4 = npart * per4;
      for (int i0 = tid; i0 < n4; i0 += 8 * kThreads) {
        float4 v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (i0 + k * kThreads < n4) v[k] = __ldg(s4 + i0 + k * kThreads);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int i4 = i0 + k * kThreads;
          if (i4 < n4) {
            const int q = i4 / per4, e = (i4 - q * per4) * 4;
            float* dst = smem + q * pstride + DP;
            const float f[4] = {v[k].x, v[k].y, v[k].z, v[k].w};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int t = (e + u) / d, c = (e + u) - t * d;
              dst[t * DP + c] = f[u];
            }
          }
        }
      }
      for (int i = tid; i < npart * DP; i += kThreads) smem[(i / DP) * pstride + (i % DP)] = 0.f;
    } else
    for (int q = 0; q < npart; ++q) {
      float* dst = smem + q * pstride;
      const int lo = c0 > 0 ? -1 : 0;  // first time index to read relative to c0
      const float* __restrict__ src = x + ((p0 + q) * (int64_t)T + c0 + lo) * d;
      const int cnt = (len - lo) * d;
#pragma unroll 4
      for (int i = tid; i < cnt; i += kThreads) {
        const int t = i / d, c = i - t * d;
        dst[(t + 1 + lo) * DP + c] = __ldg(src + i);
      }
      if (c0 == 0 && tid < DP) dst[tid] = 0.f;
    }
    __syncthreads();
    float* rows = smem + gq * pstride;
    for (int t = c0 + len - 1; t >= c0; --t) {
      const int r = t - c0 + 1;
      if (t < t_first) {  // no increment ends at point 0 without a basepoint
        __syncwarp();
        if (active) rows[r * DP + lane] = 0.f;
        continue;
      }
      float xt[d], xp[d], dl[d];
      load_row<d>(rows + r * DP, xt);
      load_row<d>(rows + (r - 1) * DP, xp);
      const float di = rows[r * DP + il] - rows[(r - 1) * DP + il];
#pragma unroll
      for (int c = 0; c < d; ++c) dl[c] = xt[c] - xp[c];
      const float tot = sig3_bwd_step<G, d>(dl, di, S1, S2, A1, A2, g3, lane);
      __syncwarp();  // every lane of the group has consumed row r before it is overwritten
      if (active) rows[r * DP + lane] = tot;
    }
    __syncthreads();
    // gx[t] = gDelta[t] - gDelta[t+1]; coalesced store with the fused epilogue
    if (vec) {
      // whole paths resident: one contiguous block of the output, 16-byte stores; the loads of
      // `add` (K.score rows) are issued 4 at a time before they are consumed
      const int per4 = (T * d) >> 2, n4 = npart * per4;
      float4* __restrict__ o4 = reinterpret_cast<float4*>(out + p0 * (int64_t)T * d);
      for (int i0 = tid; i0 < n4; i0 += 4 * kThreads) {
        float4 av[4];
        if (add) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int i4 = i0 + k * kThreads;
            if (i4 < n4) {
              const int q = i4 / per4, e = (i4 - q * per4) * 4;
              av[k] = __ldg(reinterpret_cast<const float4*>(add + (p0 + q) * ld_add + e));
            }
          }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int i4 = i0 + k * kThreads;
          if (i4 < n4) {
            const int q = i4 / per4, e = (i4 - q * per4) * 4;
            const float* src = smem + q * pstride;
            float f[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int t = (e + u) / d, c = (e + u) - t * d;
              const float nxt = (t + 1 < T) ? src[(t + 2) * DP + c] : 0.f;
              f[u] = beta * (src[(t + 1) * DP + c] - nxt);
            }
            if (add) {
              f[0] = fmaf(alpha, av[k].x, f[0]);
              f[1] = fmaf(alpha, av[k].y, f[1]);
              f[2] = fmaf(alpha, av[k].z, f[2]);
              f[3] = fmaf(alpha, av[k].w, f[3]);
            }
            o4[i4] = make_float4(f[0], f[1], f[2], f[3]);
          }
        }
      }
    } else
    for (int q = 0; q < npart; ++q) {
      const float* src = smem + q * pstride;
      const int64_t n = p0 + q;
      float* __restrict__ o = out + (n * (int64_t)T + c0) * d;
      const float* __restrict__ a = add ? add + n * ld_add + (int64_t)c0 * d : nullptr;
#pragma unroll 4
      for (int i = tid; i < len * d; i += kThreads) {
        const int t = i / d, c = i - t * d;
        const float nxt = (t + 1 < len) ? src[(t + 2) * DP + c] : carry[q * DP + c];
        float v = beta * (src[(t + 1) * DP + c] - nxt);
        if (a) v = fmaf(alpha, __ldg(a + i), v);
        o[i] = v;
      }
    }
    __syncthreads();
    for (int i = tid; i < npart * DP; i += kThreads) {
      const int q = i / DP, c = i - q * DP;
      carry[i] = smem[q * pstride + DP + c];
    }
  }
}

// ------------------------------------------------------------------------------------------
// depth-3 backward, time-parallel: the T steps are cut into kSeg segments swept by kSeg warps
// of one CTA at the same time (Chen: S_T = P_q (x) R_q with P_q the signature up to the end of
// segment q -- levels 1-2 checkpointed by the forward kernel -- and R_q the signature of the rest).
//   state at the segment end:    S = P_q
//   adjoint at the segment end:  A3 = G3,  A2 = G2 + G3 . R1,  A1 = G1 + G2 . R1 + G3 : R2
//   with R1 = T1 - P1,  R2 = T2 - P2 - P1 (x) R1   (levels of P_q^-1 (x) S_T; T = S_T)
// The sequential chain per warp is T / kSeg steps instead of T, and there are kSeg times more
// warps to hide latencies with (N = 4096: 4096 warps instead of 1024).
// Needs whole paths resident in shared memory and 16-byte aligned rows (the launcher checks).
// ------------------------------------------------------------------------------------------
template <int d>
__global__ void __launch_bounds__(kSegThreads)
sig3_bwd_seg_kernel(const float* __restrict__ x, const float* __restrict__ sig,
                    const float* __restrict__ pref, const float* __restrict__ gsig, int64_t N, int T,
                    int basepoint, int seg_len, const float* __restrict__ add, int64_t ld_add,
                    float alpha, float beta, float* __restrict__ out, const SigGradSpec gs_all) {
  constexpr int G = group_size(d), DP = row_pad(d), PPC = 32 / G;
  constexpr int D = d + d * d + d * d * d;
  constexpr int L12 = d + d * d;  // levels 1-2: all the sweep needs of the state
  constexpr int RB = DP + d * d;  // R1 (padded) + R2 per (segment, particle)
  extern __shared__ __align__(16) float smem[];
  const int tid = threadIdx.x, seg = tid >> 5, l32 = tid & 31, grp = l32 / G, lane = l32 % G;
  const int il = lane < d ? lane : d - 1;
  const int64_t p0 = (int64_t)blockIdx.x * PPC;
  const SigGradSpec gs = scene_view(gs_all, p0);  // (a CTA's particles lie in one scene)
  const int npart = (int)min((int64_t)PPC, N - p0);
  const bool active = grp < npart && lane < d;
  const int gq = grp < npart ? grp : 0;
  const int pstride = (T + 1) * DP + kGroupPad;
  float* xrows = smem;                                  // [PPC][pstride]  x, row r <-> time r-1
  float* regb = smem + PPC * pstride;                   // staging [2][PPC*D], later gDelta rows
  float* rbuf = regb + max(PPC * pstride, 2 * PPC * D);  // [kSeg][PPC][RB]

  // ---- cooperative loads: waypoints, S_T, cotangent ----
  {
    // particle-major so that no index needs a run-time division; 4 loads in flight per thread
    const float4* __restrict__ s4 = reinterpret_cast<const float4*>(x + p0 * (int64_t)T * d);
    const int per4 = (T * d) >> 2;
    for (int q0 = 0; q0 < npart; q0 += 4) {
      for (int i = tid; i < per4; i += kSegThreads) {
        float4 v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (q0 + k < npart) v[k] = __ldg(s4 + (q0 + k) * per4 + i);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if (q0 + k < npart) {
            float* dst = xrows + (q0 + k) * pstride + DP;
            const float f[4] = {v[k].x, v[k].y, v[k].z, v[k].w};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int t = (i * 4 + u) / d, c = (i * 4 + u) - t * d;
              dst[t * DP + c] = f[u];
            }
          }
        }
      }
    }
    for (int i = tid; i < npart * DP; i += kSegThreads) xrows[(i / DP) * pstride + (i % DP)] = 0.f;
    const float* __restrict__ src = sig + p0 * D;
    float* gstage = regb + PPC * D;
    const int nst = npart * D;
    for (int i = tid; i < npart * L12; i += kSegThreads) {
      const int q = i / L12;
      regb[i] = __ldg(src + q * D + (i - q * L12));
    }
    if (gs.sc != nullptr) {
      const float gam2 = 2.f * gs.state->gamma;  // -(1/h^2) = 2*gamma
      for (int i0 = tid; i0 < nst; i0 += 4 * kSegThreads) {
        float a[4], b[4], r[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int i = i0 + k * kSegThreads;
          if (i < nst) {
            const int q = i / D, c = i - q * D;
            const int64_t row = p0 + q;
            a[k] = __ldg(gs.sc + row * gs.lds + c);
            b[k] = __ldg(gs.ks + row * gs.ldks + c);
            r[k] = __ldg(gs.rowsum + row * gs.ldr);
          }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int i = i0 + k * kSegThreads;
          if (i < nst) gstage[i] = gam2 * (a[k] * r[k] - b[k]);
        }
      }
      uint32_t nt_all;
      const float coef = grad_coef(gs, &nt_all);
      if (coef != 0.f) {
        __syncthreads();
        const uint32_t nl = min(gs.state->n_ties, (uint32_t)SIGSVGD_MAX_TIES);
        const float wt = coef / (float)nt_all;
        for (uint32_t t = 0; t < nl; ++t) {
          const int64_t a_loc = (int64_t)gs.state->ties[2 * t] - gs.row_offset - p0;
          if (a_loc < 0 || a_loc >= npart) continue;  // block-uniform
          const int64_t jb = gs.state->ties[2 * t + 1];
          for (int c = tid; c < D; c += kSegThreads)
            gstage[a_loc * D + c] = fmaf(wt, __ldg(gs.sc + (p0 + a_loc) * gs.lds + c) -
                                                 __ldg(gs.sc_all + jb * gs.lds_all + c),
                                         gstage[a_loc * D + c]);
          __syncthreads();
        }
      }
    } else {
      const float* __restrict__ gsrc = gsig + p0 * D;
      for (int i = tid; i < nst; i += kSegThreads) gstage[i] = __ldg(gsrc + i);
    }
  }
  __syncthreads();

  // ---- this warp's segment: state and adjoint at its end ----
  const int t_lo = seg * seg_len, t_hi = min(T, t_lo + seg_len);
  const bool nonempty = t_lo < T;
  const bool last = t_hi == T;
  float S1 = 0.f, S2[d], A1 = 0.f, A2[d], g3[d][d];
  {
    const float* o = regb + gq * L12;           // S_T, levels 1-2
    const float* og = regb + PPC * D + gq * D;  // G
    A1 = active ? og[il] : 0.f;
#pragma unroll
    for (int j = 0; j < d; ++j) {
      A2[j] = active ? og[d + il * d + j] : 0.f;
#pragma unroll
      for (int c = 0; c < d; ++c) g3[j][c] = active ? og[d + d * d + (il * d + j) * d + c] : 0.f;
    }
    if (last || !nonempty) {
      S1 = active ? o[il] : 0.f;
#pragma unroll
      for (int j = 0; j < d; ++j) S2[j] = active ? o[d + il * d + j] : 0.f;
    } else {
      const float* __restrict__ pq = pref + ((p0 + gq) * (kSeg - 1) + seg) * (int64_t)L12;
      S1 = active ? __ldg(pq + il) : 0.f;
#pragma unroll
      for (int j = 0; j < d; ++j) S2[j] = active ? __ldg(pq + d + il * d + j) : 0.f;
      // R1, R2 of the remaining path, shared inside the group through shared memory
      float* rb = rbuf + (seg * PPC + gq) * RB;
      const float r1 = (active ? o[il] : 0.f) - S1;
      __syncwarp();
      if (active) rb[lane] = r1;
      __syncwarp();
      float R1[d];
#pragma unroll
      for (int k = 0; k < d; ++k) R1[k] = rb[k];
      if (active) {
#pragma unroll
        for (int k = 0; k < d; ++k) rb[DP + lane * d + k] = o[d + il * d + k] - S2[k] - S1 * R1[k];
      }
      __syncwarp();
      float acc1 = 0.f;
#pragma unroll
      for (int j = 0; j < d; ++j) {
        float row = 0.f, c2 = 0.f;
#pragma unroll
        for (int k = 0; k < d; ++k) {
          row = fmaf(g3[j][k], R1[k], row);
          c2 = fmaf(g3[j][k], rb[DP + j * d + k], c2);
        }
        acc1 += A2[j] * R1[j] + c2;  // original G2 here, before it is advanced below
        A2[j] += row;
      }
      A1 += acc1;
    }
  }
  __syncthreads();  // staging consumed: region B becomes the gDelta rows

  // ---- reversible sweep over the segment ----
  const int t_first = basepoint ? 0 : 1;
  const float* rows = xrows + gq * pstride;
  float* grow = regb + gq * pstride;
  if (nonempty) {
    float xt[d];
    load_row<d>(rows + t_hi * DP, xt);
    float xi = rows[t_hi * DP + il];
    for (int t = t_hi - 1; t >= t_lo; --t) {
      const int r = t + 1;
      if (t < t_first) {  // no increment ends at point 0 without a basepoint
        if (active) grow[r * DP + lane] = 0.f;
        continue;
      }
      float xp[d], dl[d];
      load_row<d>(rows + (r - 1) * DP, xp);
      const float xpi = rows[(r - 1) * DP + il];
      const float di = xi - xpi;
      xi = xpi;
#pragma unroll
      for (int c = 0; c < d; ++c) {
        dl[c] = xt[c] - xp[c];
        xt[c] = xp[c];
      }
      const float tot = sig3_bwd_step<G, d>(dl, di, S1, S2, A1, A2, g3, lane);
      if (active) grow[r * DP + lane] = tot;
    }
  }
  __syncthreads();

  // ---- gx[t] = gDelta[t] - gDelta[t+1], fused epilogue, 16-byte stores ----
  {
    const int per4 = (T * d) >> 2;
    float4* __restrict__ o4 = reinterpret_cast<float4*>(out + p0 * (int64_t)T * d);
    for (int q0 = 0; q0 < npart; q0 += 4) {
      for (int i = tid; i < per4; i += kSegThreads) {
        float4 av[4];
        if (add) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            if (q0 + k < npart)
              av[k] = __ldg(reinterpret_cast<const float4*>(add + (p0 + q0 + k) * ld_add) + i);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if (q0 + k < npart) {
            const float* src = regb + (q0 + k) * pstride;
            float f[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int t = (i * 4 + u) / d, c = (i * 4 + u) - t * d;
              const float nxt = (t + 1 < T) ? src[(t + 2) * DP + c] : 0.f;
              f[u] = beta * (src[(t + 1) * DP + c] - nxt);
            }
            if (add) {
              f[0] = fmaf(alpha, av[k].x, f[0]);
              f[1] = fmaf(alpha, av[k].y, f[1]);
              f[2] = fmaf(alpha, av[k].z, f[2]);
              f[3] = fmaf(alpha, av[k].w, f[3]);
            }
            o4[(q0 + k) * per4 + i] = make_float4(f[0], f[1], f[2], f[3]);
          }
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------------
// generic kernels: one warp per particle, state in shared memory
// ------------------------------------------------------------------------------------------
struct LevelInfo {
  int off[10];  // off[k] = offset of level k (1-based) in the D-vector
  int pw[10];   // pw[k] = d^k
};

// S <- S (x) exp(sgn*dl); dl[] in shared memory; warp-synchronous.
__device__ __forceinline__ void generic_mul_exp(float* S, const float* dl, float sgn, int d, int m,
                                                const LevelInfo& L, int lane) {
  for (int k = m; k >= 1; --k) {
    float* Sk = S + L.off[k];
    for (int e = lane; e < L.pw[k]; e += 32) {
      float b = 1.f;
      for (int j = 1; j <= k - 1; ++j) {
        const int pre = e / L.pw[k - j];  // index of (u_1..u_j) inside level j
        const int uj = pre % d;
        b = S[L.off[j] + pre] + b * (sgn * dl[uj]) / (float)(k - j + 1);
      }
      Sk[e] += b * (sgn * dl[e % d]);
    }
    __syncwarp();
  }
}

__global__ void __launch_bounds__(32)
sig_generic_fwd_kernel(const float* __restrict__ x, int64_t N, int T, int d, int m, int basepoint,
                       LevelInfo L, int D, float* __restrict__ sig) {
  extern __shared__ __align__(16) float smem[];
  float* S = smem;       // D
  float* dl = smem + D;  // d
  const int lane = threadIdx.x;
  const int64_t n = blockIdx.x;
  const float* xn = x + n * (int64_t)T * d;
  for (int i = lane; i < D; i += 32) S[i] = 0.f;
  __syncwarp();
  for (int t = basepoint ? 0 : 1; t < T; ++t) {
    for (int c = lane; c < d; c += 32)
      dl[c] = __ldg(xn + (int64_t)t * d + c) - (t > 0 ? __ldg(xn + (int64_t)(t - 1) * d + c) : 0.f);
    __syncwarp();
    generic_mul_exp(S, dl, 1.f, d, m, L, lane);
  }
  for (int i = lane; i < D; i += 32) sig[n * D + i] = S[i];
}

__global__ void __launch_bounds__(32)
sig_generic_bwd_kernel(const float* __restrict__ x, const float* __restrict__ sig,
                       const float* __restrict__ gsig, int64_t N, int T, int d, int m,
                       int basepoint, LevelInfo L, int D, const float* __restrict__ add,
                       int64_t ld_add, float alpha, float beta, float* __restrict__ out) {
  extern __shared__ __align__(16) float smem[];
  float* S = smem;            // D
  float* A = S + D;           // D
  float* dl = A + D;          // d
  float* gprev = dl + d;      // d : gDelta of step t+1
  float* acc = gprev + d;     // d * 33 per-lane partial gDelta
  const int lane = threadIdx.x;
  const int64_t n = blockIdx.x;
  const float* xn = x + n * (int64_t)T * d;
  for (int i = lane; i < D; i += 32) {
    S[i] = __ldg(sig + n * D + i);
    A[i] = __ldg(gsig + n * D + i);
  }
  for (int c = lane; c < d; c += 32) gprev[c] = 0.f;
  __syncwarp();
  const int t_first = basepoint ? 0 : 1;
  for (int t = T - 1; t >= 0; --t) {
    if (t < t_first) {  // point 0 without basepoint: gx[0] = -gDelta[1]
      for (int c = lane; c < d; c += 32) {
        float v = beta * (0.f - gprev[c]);
        if (add) v = fmaf(alpha, __ldg(add + n * ld_add + (int64_t)t * d + c), v);
        out[(n * (int64_t)T + t) * d + c] = v;
      }
      break;
    }
    for (int c = lane; c < d; c += 32)
      dl[c] = __ldg(xn + (int64_t)t * d + c) - (t > 0 ? __ldg(xn + (int64_t)(t - 1) * d + c) : 0.f);
    for (int i = lane; i < d * 33; i += 32) acc[i] = 0.f;
    __syncwarp();
    generic_mul_exp(S, dl, -1.f, d, m, L, lane);  // S = S_{t-1}
    // (a) gDelta contributions, entry-parallel over dL/dS_t (A still holds the "new" adjoint)
    for (int k = 1; k <= m; ++k) {
      for (int e = lane; e < L.pw[k]; e += 32) {
        const float a = A[L.off[k] + e];
        // suffix products suf_p = prod_{q>p} dl[u_q]; walk p = k..1 keeping Lp via recomputation
        // L_1 = 1/k! ; L_{p+1} = L_p*dl[u_p] + P^p[u_1..u_p]/(k-p)!
        float fact_inv[9];
        {
          float f = 1.f;
          fact_inv[0] = 1.f;
          for (int q = 1; q <= k; ++q) {
            f *= (float)q;
            fact_inv[q] = 1.f / f;
          }
        }
        float Lp = fact_inv[k];
        // forward over p computing L_p and multiplying by suffix computed on the fly
        for (int p = 1; p <= k; ++p) {
          const int pre = e / L.pw[k - p];
          const int up = pre % d;
          float suf = 1.f;
          for (int q = p + 1; q <= k; ++q) suf *= dl[(e / L.pw[k - q]) % d];
          acc[up * 33 + lane] += a * Lp * suf;
          if (p < k) Lp = Lp * dl[up] + S[L.off[p] + pre] * fact_inv[k - p];
        }
      }
    }
    __syncwarp();
    // (b) adjoint update, ascending levels, in place:
    //     A^i[u] += sum_{j>=1} sum_{|v|=j} A^{i+j}[u,v] prod dl[v] / j!
    for (int i = 1; i <= m - 1; ++i) {
      for (int e = lane; e < L.pw[i]; e += 32) {
        float s = 0.f, fj = 1.f;
        for (int j = 1; j <= m - i; ++j) {
          fj *= (float)j;
          const float* Ak = A + L.off[i + j] + (int64_t)e * L.pw[j];
          float sj = 0.f;
          for (int v = 0; v < L.pw[j]; ++v) {
            float w = 1.f;
            int vv = v;
            for (int q = 0; q < j; ++q) {
              w *= dl[vv % d];
              vv /= d;
            }
            sj = fmaf(Ak[v], w, sj);
          }
          s += sj / fj;
        }
        A[L.off[i] + e] += s;
      }
      __syncwarp();
    }
    // (c) reduce per-lane partials, emit gx[t] = gDelta[t] - gDelta[t+1]
    for (int c = lane; c < d; c += 32) {
      float s = 0.f;
      for (int l = 0; l < 32; ++l) s += acc[c * 33 + l];
      float v = beta * (s - gprev[c]);
      if (add) v = fmaf(alpha, __ldg(add + n * ld_add + (int64_t)t * d + c), v);
      out[(n * (int64_t)T + t) * d + c] = v;
      gprev[c] = s;
    }
    __syncwarp();
  }
}

size_t seg_bwd_smem(int64_t T, int64_t d);

template <int d>
int launch_sig3_fwd(const float* x, int64_t N, int64_t T, int basepoint, float* sig,
                    cudaStream_t st, float* stats, const float* score, int64_t ld_score, int64_t P,
                    float* pref) {
  constexpr int G = group_size(d), DP = row_pad(d), PPC = kThreads / G;
  constexpr int D = d + d * d + d * d * d;
  constexpr int kBudget = 40 * 1024;
  static_assert(PPC * D * 4 <= kBudget, "output staging must fit");
  int TC = (int)std::min<int64_t>(T, (kBudget / 4 / PPC - kGroupPad) / DP);
  size_t smem = (size_t)std::max(PPC * (TC * DP + kGroupPad), PPC * D) * 4;
  const int64_t grid = ceil_div(N, PPC);
  // the time-parallel forward pays (Chen products, four times the warps) only while the plain
  // kernel leaves the SMs short of warps: measured 16 -> 10 us at N=256, 16 -> 13 us at N=1024, but
  // 20 -> 25 us at N=4096 where both are issue-bound
  if (pref != nullptr && is_aligned16(x) && ((T * d) & 3) == 0 && grid <= 4 * (int64_t)num_sms()) {
    const size_t smem_seg = (size_t)(kSeg * PPC * D + PPC * ((T + 1) * DP + kGroupPad)) * 4;
    if (smem_seg <= 96 * 1024) {
      if (smem_seg > 48 * 1024)
        SIG_CUDA_OK(cudaFuncSetAttribute(sig3_fwd_seg_kernel<d>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         96 * 1024));
      sig3_fwd_seg_kernel<d><<<(unsigned)grid, kSegThreads, smem_seg, st>>>(
          x, N, (int)T, basepoint, (int)ceil_div(T, kSeg), sig, stats, score, ld_score, (int)P, pref);
      SIG_LAUNCH_OK("sig3_fwd_seg_kernel");
      return 0;
    }
  }
  sig3_fwd_kernel<d><<<(unsigned)grid, kThreads, smem, st>>>(x, N, (int)T, basepoint, TC, sig, stats,
                                                              score, ld_score, (int)P, pref,
                                                              (int)ceil_div(T, kSeg));
  SIG_LAUNCH_OK("sig3_fwd_kernel");
  return 0;
}

template <int d>
int launch_sig3_bwd(const float* x, const float* sig, const float* gsig, int64_t N, int64_t T,
                    int basepoint, const float* add, int64_t ld_add, float alpha, float beta,
                    float* out, cudaStream_t st, const SigGradSpec& gs, const float* pref) {
  constexpr int G = group_size(d), DP = row_pad(d), PPC = kThreads / G;
  constexpr int D = d + d * d + d * d * d;
  constexpr int kBudget = 40 * 1024;
  static_assert(2 * PPC * D * 4 <= kBudget, "state staging must fit");
  if (pref != nullptr && is_aligned16(x) && is_aligned16(out) &&
      (add == nullptr || (is_aligned16(add) && (ld_add & 3) == 0))) {
    // time-parallel kernel: whole paths resident, kSeg warps per group of particles
    const size_t smem = seg_bwd_smem(T, d);
    if (smem > 48 * 1024)
      SIG_CUDA_OK(cudaFuncSetAttribute(sig3_bwd_seg_kernel<d>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       96 * 1024));
    sig3_bwd_seg_kernel<d><<<(unsigned)ceil_div(N, PPC), kSegThreads, smem, st>>>(
        x, sig, pref, gsig, N, (int)T, basepoint, (int)ceil_div(T, kSeg), add, ld_add, alpha, beta, out, gs);
    SIG_LAUNCH_OK("sig3_bwd_seg_kernel");
    return 0;
  }
  int TC = (int)std::min<int64_t>(T, (kBudget / 4 / PPC - kGroupPad) / DP - 1);
  size_t smem = (size_t)std::max(PPC * ((TC + 1) * DP + kGroupPad), 2 * PPC * D) * 4;
  const int64_t grid = ceil_div(N, PPC);
  sig3_bwd_kernel<d><<<(unsigned)grid, kThreads, smem, st>>>(x, sig, gsig, N, (int)T, basepoint,
                                                              TC, add, ld_add, alpha, beta, out, gs);
  SIG_LAUNCH_OK("sig3_bwd_kernel");
  return 0;
}

int make_levels(int64_t d, int depth, LevelInfo* L, int64_t* D) {
  int64_t off = 0, p = 1;
  L->pw[0] = 1;
  for (int k = 1; k <= depth; ++k) {
    p *= d;
    if (p > (1 << 24)) return 1;
    L->off[k] = (int)off;
    L->pw[k] = (int)p;
    off += p;
  }
  *D = off;
  return 0;
}

}  // namespace

bool sig_fwd_fuses_stats(int64_t d, int depth) { return depth == 3 && d >= 1 && d <= 8; }

// The time-parallel backward needs: depth 3, d <= 8, whole paths resident in both kernels,
// 16-byte rows, and segments of at least two steps.
namespace {
size_t seg_bwd_smem(int64_t T, int64_t d) {
  const int64_t PPC = 32 / group_size((int)d), DP = row_pad((int)d), D = d + d * d + d * d * d;
  const int64_t pstride = (T + 1) * DP + kGroupPad;
  return (size_t)(PPC * pstride + std::max(PPC * pstride, 2 * PPC * D) + kSeg * PPC * (DP + d * d)) * 4;
}
}  // namespace
bool sig_bwd_uses_prefix(int64_t T, int64_t d, int depth) {
  if (depth != 3 || d < 1 || d > 8 || T < 2 * kSeg || T > 4096 || ((T * d) & 3) != 0) return false;
  return seg_bwd_smem(T, d) <= 96 * 1024;
}
int64_t sig_prefix_floats(int64_t N, int64_t T, int64_t d, int depth) {
  return sig_bwd_uses_prefix(T, d, depth) ? N * (kSeg - 1) * (d + d * d) : 0;
}

int launch_sig_fwd(const float* x, int64_t N, int64_t T, int64_t d, int depth, int basepoint,
                   float* sig, cudaStream_t st, float* stats, const float* score, int64_t ld_score,
                   int64_t P, float* pref) {
  SIG_CHECK_ARG(pref == nullptr || sig_bwd_uses_prefix(T, d, depth),
                "signature_fwd: prefix checkpoints are not available for this shape");
  SIG_CHECK_ARG(stats == nullptr || sig_fwd_fuses_stats(d, depth),
                "signature_fwd: fused statistics need depth 3 and d <= 8");
  SIG_CHECK_ARG(N >= 0 && T >= 1 && d >= 1 && depth >= 1 && depth <= 8,
                "signature_fwd: bad shape N=%lld T=%lld d=%lld depth=%d", (long long)N,
                (long long)T, (long long)d, depth);
  SIG_CHECK_ARG(basepoint || T >= 2, "signature_fwd: need >= 2 points without a basepoint");
  SIG_CHECK_ARG(T < (1 << 24), "signature_fwd: T too large");
  if (N == 0) return 0;
  if (depth == 3) {
    switch (d) {
      case 1: return launch_sig3_fwd<1>(x, N, T, basepoint, sig, st, stats, score, ld_score, P, pref);
      case 2: return launch_sig3_fwd<2>(x, N, T, basepoint, sig, st, stats, score, ld_score, P, pref);
      case 3: return launch_sig3_fwd<3>(x, N, T, basepoint, sig, st, stats, score, ld_score, P, pref);
      case 4: return launch_sig3_fwd<4>(x, N, T, basepoint, sig, st, stats, score, ld_score, P, pref);
      case 5: return launch_sig3_fwd<5>(x, N, T, basepoint, sig, st, stats, score, ld_score, P, pref);
      case 6: return launch_sig3_fwd<6>(x, N, T, basepoint, sig, st, stats, score, ld_score, P, pref);
      case 7: return launch_sig3_fwd<7>(x, N, T, basepoint, sig, st, stats, score, ld_score, P, pref);
      case 8: return launch_sig3_fwd<8>(x, N, T, basepoint, sig, st, stats, score, ld_score, P, pref);
      default: break;
    }
  }
  if (depth == 4 && sig4_available(d)) return launch_sig4_fwd(x, N, T, d, basepoint, sig, st);
  LevelInfo L;
  int64_t D;
  if (make_levels(d, depth, &L, &D) != 0 || (D + d) * 4 > 200 * 1024) {
    set_error("signature_fwd: d=%lld depth=%d (D=%lld) exceeds the shared-memory resident state",
              (long long)d, depth, (long long)sig_channels(d, depth));
    return SIGSVGD_ERR_UNSUPPORTED;
  }
  const size_t smem = (size_t)(D + d) * 4;
  if (smem > 48 * 1024)
    SIG_CUDA_OK(cudaFuncSetAttribute(sig_generic_fwd_kernel,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  sig_generic_fwd_kernel<<<(unsigned)N, 32, smem, st>>>(x, N, (int)T, (int)d, depth, basepoint, L,
                                                        (int)D, sig);
  SIG_LAUNCH_OK("sig_generic_fwd_kernel");
  return 0;
}

bool sig_bwd_fuses_grad(int64_t d, int depth) {
  return (depth == 3 && d >= 1 && d <= 8) || (depth == 4 && sig4_available(d));
}

int launch_sig_bwd(const float* x, const float* sig, const float* gsig, int64_t N, int64_t T,
                   int64_t d, int depth, int basepoint, const float* add, int64_t ld_add,
                   float alpha, float beta, float* out, cudaStream_t st, const SigGradSpec* spec,
                   const float* pref) {
  SIG_CHECK_ARG(pref == nullptr || sig_bwd_uses_prefix(T, d, depth),
                "signature_bwd: prefix checkpoints are not available for this shape");
  SigGradSpec gs{};
  if (spec) {
    gs = *spec;
    SIG_CHECK_ARG(sig_bwd_fuses_grad(d, depth) && gs.sc && gs.sc_all && gs.ks && gs.rowsum && gs.state,
                  "signature_bwd: fused gradient needs depth 3 (d <= 8) or depth 4 (2 <= d <= 8) and a complete spec");
    SIG_CHECK_ARG(gs.scene_rows == 0 || (gs.scene_rows % 8 == 0 && gs.coef_mode == 0 && N % gs.scene_rows == 0),
                  "signature_bwd: scenes of a fused gradient need a multiple of 8 rows and coef_mode 0");
  }
  SIG_CHECK_ARG(N >= 0 && T >= 1 && d >= 1 && depth >= 1 && depth <= 8,
                "signature_bwd: bad shape N=%lld T=%lld d=%lld depth=%d", (long long)N,
                (long long)T, (long long)d, depth);
  SIG_CHECK_ARG(basepoint || T >= 2, "signature_bwd: need >= 2 points without a basepoint");
  SIG_CHECK_ARG(T < (1 << 24), "signature_bwd: T too large");
  SIG_CHECK_ARG(add == nullptr || ld_add >= T * d, "signature_bwd: ld_add < T*d");
  if (N == 0) return 0;
  if (depth == 3) {
    switch (d) {
#define SIG3_BWD_CASE(dd) \
  case dd:                \
    return launch_sig3_bwd<dd>(x, sig, gsig, N, T, basepoint, add, ld_add, alpha, beta, out, st, gs, pref);
      SIG3_BWD_CASE(1)
      SIG3_BWD_CASE(2)
      SIG3_BWD_CASE(3)
      SIG3_BWD_CASE(4)
      SIG3_BWD_CASE(5)
      SIG3_BWD_CASE(6)
      SIG3_BWD_CASE(7)
      SIG3_BWD_CASE(8)
#undef SIG3_BWD_CASE
      default: break;
    }
  }
  if (depth == 4 && sig4_available(d))
    return launch_sig4_bwd(x, sig, gsig, N, T, d, basepoint, add, ld_add, alpha, beta, out, st, spec ? &gs : nullptr);
  LevelInfo L;
  int64_t D;
  if (make_levels(d, depth, &L, &D) != 0 || (2 * D + 2 * d + 33 * d) * 4 > 200 * 1024) {
    set_error("signature_bwd: d=%lld depth=%d (D=%lld) exceeds the shared-memory resident state",
              (long long)d, depth, (long long)sig_channels(d, depth));
    return SIGSVGD_ERR_UNSUPPORTED;
  }
  const size_t smem = (size_t)(2 * D + 2 * d + 33 * d) * 4;
  if (smem > 48 * 1024)
    SIG_CUDA_OK(cudaFuncSetAttribute(sig_generic_bwd_kernel,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  sig_generic_bwd_kernel<<<(unsigned)N, 32, smem, st>>>(x, sig, gsig, N, (int)T, (int)d, depth,
                                                        basepoint, L, (int)D, add, ld_add, alpha,
                                                        beta, out);
  SIG_LAUNCH_OK("sig_generic_bwd_kernel");
  return 0;
}

}  // namespace sigsvgd
