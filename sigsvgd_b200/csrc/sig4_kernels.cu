// Depth-4 truncated path-signature forward / VJP kernels (BASELINE.json configs[3], SURVEY.md 8 row +N2).
//
// Same operator as sig_kernels.cu (signatory.signature(path, 4, basepoint=True) and its backward;
// reference call sites src/kernels/_traj_kernels.py:125-126, src/inference/svgd.py:43), but level 4
// has d^4 entries per particle (2401 for d = 7) and no longer fits the register layout of the depth-3
// kernels, nor does one warp per particle walking a shared-memory state (the generic kernels) come
// anywhere near the machine.
//
// Layout: d^2 threads per particle.  Thread (i, j) owns every tensor entry whose first two indices
// are (i, j): S4[i,j,:,:] (d^2 registers), S3[i,j,:] (d registers), S2[i,j], and its own copy of
// S1[i].  With the Horner form
//     S4 += [S3 + (S2 + (S1 + D/4) (x) D/3) (x) D/2] (x) D,   S3 += [S2 + (S1 + D/3) (x) D/2] (x) D, ...
// every factor a thread needs carries its own (i, j) prefix, so the FORWARD has no cross-thread
// traffic at all: 2 d^2 FMAs per thread and step, waypoints broadcast from shared memory.
//
// BACKWARD: reversible sweep (levels 1-3 of S_{t-1} = S_t (x) exp(-D_t); level 4 of S is never
// needed), level-4 adjoint constant in registers (d^2 per thread), A3[i,j,:], A2[i,j] thread-local.
// Per step a thread produces a d-vector of contributions to dL/dD_t plus one contribution to the
// level-1 adjoint; these are parked in shared memory and reduced over the particle's d^2 threads
// once per chunk of kChunk steps (the level-1 adjoint only feeds dL/dD_t additively, so it is
// carried by the chunk epilogue instead of the step loop).
//
// Bound: issue (fp32 FMA): 2*d^4 (forward) / ~5*d^4 (backward) FMAs per particle and step; HBM
// traffic is 4*N*(T*d + D) bytes forward, 4*N*(2*T*d + 2*D) backward, D = d + d^2 + d^3 + d^4.
#include "common.cuh"

namespace sigsvgd {

namespace {

template <int d>
struct Sig4Cfg {
  static constexpr int DD = d * d;
  static constexpr int PPC = (256 / DD) < 8 ? (256 / DD) : 8;  // particles per CTA
  static constexpr int NT = ((PPC * DD + 31) / 32) * 32;       // threads per CTA
  static constexpr int D = d + DD + DD * d + DD * DD;
  static constexpr int O2 = d, O3 = d + DD, O4 = d + DD + DD * d;  // level offsets in the D-vector
  static constexpr int kChunk = 8;                                  // time steps per staging chunk
  static constexpr int DP = (d + 3) & ~3;                           // padded row of staged increments
  static constexpr int XS = PPC * kChunk * DP;                      // one buffer of staged increments (floats)
  static constexpr int kPre = (PPC * kChunk * d + NT - 1) / NT;     // increments a thread prefetches per chunk
  // backward: rows of per-thread contributions per step: d (gDelta, free index), r (level-1 adjoint),
  // gj (target j, stored transposed), gi (target i)
  static constexpr int kRows = d + 3;
  // particle pitch of the contribution buffer: congruent to d^2 mod 32, so that a warp straddling two
  // particles still maps its 32 lanes to 32 different banks
  static constexpr int CP0 = kChunk * kRows * DD;
  static constexpr int CP = CP0 + ((DD - CP0) % 32 + 32) % 32;
  static constexpr int PP = kChunk * kRows * d;                     // reduced sums per particle
};

__host__ __device__ __forceinline__ bool aligned16(const void* p) {
  return (reinterpret_cast<uintptr_t>(p) & 15u) == 0;
}

// Increments Delta_t = x_t - x_{t-1} (x_{-1} = basepoint = 0), t = t0 .. t0+nt-1, of every particle of the
// CTA, through registers: issue() starts the global loads of a chunk, store() parks the differences in
// shared memory as xs[p][s][DP] -- called one chunk ahead of its use, so the load latency is covered
// by the previous chunk's arithmetic.
template <int d>
struct DeltaPrefetch {
  using C = Sig4Cfg<d>;
  float cur[C::kPre], prev[C::kPre];
  // element e = threadIdx.x + u*NT of the [PPC][kChunk][d] grid: thread constants, computed once
  int src[C::kPre];   // offset of (particle, row 0, c) inside x, or -1 for "no such particle/element"
  int dst[C::kPre];   // offset inside one xs buffer
  int row[C::kPre];
  __device__ __forceinline__ void init(int64_t N, int T, int64_t n0) {
#pragma unroll
    for (int u = 0; u < C::kPre; ++u) {
      const int e = threadIdx.x + u * C::NT;
      const int pp = e / (C::kChunk * d), r = e - pp * (C::kChunk * d);
      row[u] = r / d;
      const int c = r - row[u] * d;
      const bool ok = e < C::PPC * C::kChunk * d && n0 + pp < N;
      src[u] = ok ? (pp * T + row[u]) * d + c : -1;
      dst[u] = (pp * C::kChunk + row[u]) * C::DP + c;
    }
  }
  __device__ __forceinline__ void issue(const float* __restrict__ xc, int t0, int nt) {  // xc = x + n0*T*d
#pragma unroll
    for (int u = 0; u < C::kPre; ++u) {
      cur[u] = prev[u] = 0.f;
      if (src[u] >= 0 && row[u] < nt) {
        const float* px = xc + (int64_t)t0 * d + src[u];
        cur[u] = __ldg(px);
        if (t0 + row[u] > 0) prev[u] = __ldg(px - d);
      }
    }
  }
  __device__ __forceinline__ void store(float* xs, int nt) const {
#pragma unroll
    for (int u = 0; u < C::kPre; ++u)
      if (src[u] >= 0 && row[u] < nt) xs[dst[u]] = cur[u] - prev[u];
  }
};

// one staged increment row -> registers (16-byte shared loads; the row is a broadcast within a particle)
template <int d>
__device__ __forceinline__ void load_delta(const float* r0, float (&dl)[d]) {
  if constexpr (d >= 4) {
    float tmp[Sig4Cfg<d>::DP];
#pragma unroll
    for (int c = 0; c < Sig4Cfg<d>::DP; c += 4) {
      const float4 v = *reinterpret_cast<const float4*>(r0 + c);
      tmp[c] = v.x, tmp[c + 1] = v.y, tmp[c + 2] = v.z, tmp[c + 3] = v.w;
    }
#pragma unroll
    for (int c = 0; c < d; ++c) dl[c] = tmp[c];
  } else {
#pragma unroll
    for (int c = 0; c < d; ++c) dl[c] = r0[c];
  }
}

template <int d>
__global__ void __launch_bounds__(Sig4Cfg<d>::NT, d <= 7 ? 3 : 2)
sig4_fwd_kernel(const float* __restrict__ x, int64_t N, int T, int basepoint, float* __restrict__ sig) {
  using C = Sig4Cfg<d>;
  constexpr int DD = C::DD, D = C::D;
  extern __shared__ __align__(16) float smem[];
  float* xs = smem;                 // [2][PPC][kChunk][DP] increments, double-buffered
  float* stage = smem + 2 * C::XS;  // [PPC][D]
  const int tid = threadIdx.x;
  const int p = tid / DD, ij = tid - p * DD, i = ij / d, j = ij - i * d;
  const int64_t n0 = (int64_t)blockIdx.x * C::PPC;
  const bool active = p < C::PPC && n0 + p < N;

  float S4[d][d], S3[d], s2 = 0.f, s1 = 0.f;
#pragma unroll
  for (int k = 0; k < d; ++k) {
    S3[k] = 0.f;
#pragma unroll
    for (int l = 0; l < d; ++l) S4[k][l] = 0.f;
  }
  DeltaPrefetch<d> pf;
  pf.init(N, T, n0);
  const float* xc = x + n0 * T * d;
  const int t_first = basepoint ? 0 : 1;
  if (t_first < T) {
    const int nt0 = min(C::kChunk, T - t_first);
    pf.issue(xc, t_first, nt0);
    pf.store(xs, nt0);
  }
  __syncthreads();
  int buf = 0;
  for (int t0 = t_first; t0 < T; t0 += C::kChunk, buf ^= 1) {
    const int nt = min(C::kChunk, T - t0);
    const int nt_next = min(C::kChunk, T - t0 - C::kChunk);
    if (nt_next > 0) pf.issue(xc, t0 + C::kChunk, nt_next);
    if (active) {
      for (int s = 0; s < nt; ++s) {
        const float* r0 = xs + buf * C::XS + (p * C::kChunk + s) * C::DP;
        float dl[d];
        load_delta<d>(r0, dl);
        const float di = r0[i], dj = r0[j];
        const float b2 = fmaf(fmaf(di, 0.25f, s1), dj * (1.f / 3.f), s2);
#pragma unroll
        for (int k = 0; k < d; ++k) {
          const float b3 = fmaf(b2, 0.5f * dl[k], S3[k]);
#pragma unroll
          for (int l = 0; l < d; ++l) S4[k][l] = fmaf(b3, dl[l], S4[k][l]);
        }
        const float c2 = fmaf(fmaf(di, 1.f / 3.f, s1), 0.5f * dj, s2);
#pragma unroll
        for (int k = 0; k < d; ++k) S3[k] = fmaf(c2, dl[k], S3[k]);
        s2 = fmaf(fmaf(di, 0.5f, s1), dj, s2);
        s1 += di;
      }
    }
    if (nt_next > 0) pf.store(xs + (buf ^ 1) * C::XS, nt_next);
    __syncthreads();
  }
  __syncthreads();
  if (active) {
    float* so = stage + p * D;
    if (j == 0) so[i] = s1;
    so[C::O2 + ij] = s2;
#pragma unroll
    for (int k = 0; k < d; ++k) so[C::O3 + ij * d + k] = S3[k];
#pragma unroll
    for (int k = 0; k < d; ++k)
#pragma unroll
      for (int l = 0; l < d; ++l) so[C::O4 + ij * DD + k * d + l] = S4[k][l];
  }
  __syncthreads();
  const int64_t np = min((int64_t)C::PPC, N - n0);
  const int64_t total = np * D;
  float* dst = sig + n0 * D;
  if (aligned16(dst) && (total & 3) == 0) {
    const float4* s4 = reinterpret_cast<const float4*>(stage);
    float4* d4 = reinterpret_cast<float4*>(dst);
    for (int64_t e = tid; e < total / 4; e += C::NT) d4[e] = s4[e];
  } else {
    for (int64_t e = tid; e < total; e += C::NT) dst[e] = stage[e];
  }
}


template <int d>
__global__ void __launch_bounds__(Sig4Cfg<d>::NT, 2)
sig4_bwd_kernel(const float* __restrict__ x, const float* __restrict__ sig, const float* __restrict__ gsig,
                int64_t N, int T, int basepoint, const float* __restrict__ add, int64_t ld_add, float alpha,
                float beta, float* __restrict__ out, const SigGradSpec gs_all) {
  using C = Sig4Cfg<d>;
  constexpr int DD = C::DD, D = C::D, CH = C::kChunk;
  extern __shared__ __align__(16) float smem[];
  constexpr int kRows = C::kRows;
  float* xs = smem;                     // [2][PPC][CH][DP] increments, double-buffered
  float* cbuf = smem + 2 * C::XS;       // [PPC] x pitch CP: [CH][kRows][DD] per-thread contributions
  float* part = cbuf + C::PPC * C::CP;  // [PPC][CH][kRows][d] sums over groups of d entries of the rows above
  const int tid = threadIdx.x;
  const int p = tid / DD, ij = tid - p * DD, i = ij / d, j = ij - i * d;
  const int64_t n0 = (int64_t)blockIdx.x * C::PPC;
  const int64_t n = n0 + p;
  const bool active = p < C::PPC && n < N;

  float A4[d][d], A3[d], S3[d], a2 = 0.f, s2 = 0.f, s1 = 0.f;
  float a1 = 0.f, gprev = 0.f;  // carried by the threads ij < d (component m = ij)
  if (active) {
    const float* sn = sig + n * D;
#pragma unroll
    for (int k = 0; k < d; ++k) S3[k] = __ldg(sn + C::O3 + ij * d + k);
    s2 = __ldg(sn + C::O2 + ij);
    s1 = __ldg(sn + i);
    const SigGradSpec gs = scene_view(gs_all, n);
    if (gs.sc != nullptr) {
      // the cotangent formed while it is loaded (SigGradSpec): g = -(1/h^2)(sc*rowsum - K.Sc) [+ tie term]
      const float gam2 = 2.f * gs.state->gamma;  // -(1/h^2) = 2*gamma
      const float* sc = gs.sc + n * gs.lds;
      const float* ks = gs.ks + n * gs.ldks;
      const float rs = __ldg(gs.rowsum + n * gs.ldr);
      auto cot = [&](int e) { return gam2 * (__ldg(sc + e) * rs - __ldg(ks + e)); };
#pragma unroll
      for (int k = 0; k < d; ++k) {
#pragma unroll
        for (int l = 0; l < d; ++l) A4[k][l] = cot(C::O4 + ij * DD + k * d + l);
        A3[k] = cot(C::O3 + ij * d + k);
      }
      a2 = cot(C::O2 + ij);
      if (ij < d) a1 = cot(ij);
      uint32_t nt_all;
      const float coef = grad_coef(gs, &nt_all);
      if (coef != 0.f) {
        const uint32_t nl = min(gs.state->n_ties, (uint32_t)SIGSVGD_MAX_TIES);
        const float wt = coef / (float)nt_all;
        for (uint32_t t = 0; t < nl; ++t) {
          if ((int64_t)gs.state->ties[2 * t] != gs.row_offset + n) continue;
          const float* sb = gs.sc_all + (int64_t)gs.state->ties[2 * t + 1] * gs.lds_all;
          auto tie = [&](int e) { return wt * (__ldg(sc + e) - __ldg(sb + e)); };
#pragma unroll
          for (int k = 0; k < d; ++k) {
#pragma unroll
            for (int l = 0; l < d; ++l) A4[k][l] += tie(C::O4 + ij * DD + k * d + l);
            A3[k] += tie(C::O3 + ij * d + k);
          }
          a2 += tie(C::O2 + ij);
          if (ij < d) a1 += tie(ij);
        }
      }
    } else {
      const float* gn = gsig + n * D;
#pragma unroll
      for (int k = 0; k < d; ++k) {
#pragma unroll
        for (int l = 0; l < d; ++l) A4[k][l] = __ldg(gn + C::O4 + ij * DD + k * d + l);
        A3[k] = __ldg(gn + C::O3 + ij * d + k);
      }
      a2 = __ldg(gn + C::O2 + ij);
      if (ij < d) a1 = __ldg(gn + ij);
    }
  }
  const int t_first = basepoint ? 0 : 1;
  DeltaPrefetch<d> pf;
  pf.init(N, T, n0);
  const float* xc = x + n0 * T * d;
  if (T - 1 >= t_first) {
    const int t_lo0 = max(t_first, T - CH);
    pf.issue(xc, t_lo0, T - t_lo0);
    pf.store(xs, T - t_lo0);
  }
  __syncthreads();
  int buf = 0;
  for (int t_hi = T - 1; t_hi >= t_first; t_hi -= CH, buf ^= 1) {
    const int t_lo = max(t_first, t_hi - CH + 1), nt = t_hi - t_lo + 1;
    const int t_hi_next = t_hi - CH, t_lo_next = max(t_first, t_hi_next - CH + 1);
    const int nt_next = t_hi_next >= t_first ? t_hi_next - t_lo_next + 1 : 0;
    if (nt_next > 0) pf.issue(xc, t_lo_next, nt_next);
    if (active) {
      for (int s = nt - 1; s >= 0; --s) {
        const float* r0 = xs + buf * C::XS + (p * CH + s) * C::DP;
        float dl[d];
        load_delta<d>(r0, dl);
        const float di = r0[i], dj = r0[j];
        // levels 1-3 of S_{t-1} = S_t (x) exp(-Delta_t)
        {
          const float c2 = fmaf(fmaf(di, -1.f / 3.f, s1), -0.5f * dj, s2);
#pragma unroll
          for (int k = 0; k < d; ++k) S3[k] = fmaf(c2, -dl[k], S3[k]);
          s2 = fmaf(fmaf(di, -0.5f, s1), -dj, s2);
          s1 -= di;
        }
        // weights of the prefix (W_{n,q} of DESIGN.md's depth-4 section), all functions of (i, j)
        const float q43 = fmaf(dj, fmaf(di, 1.f / 24.f, s1 * (1.f / 6.f)), 0.5f * s2);
        const float w33 = fmaf(dj, fmaf(di, 1.f / 6.f, 0.5f * s1), s2);
        const float w42 = fmaf(di, 1.f / 24.f, s1 * (1.f / 6.f));
        const float w32 = fmaf(di, 1.f / 6.f, 0.5f * s1);
        const float w22 = fmaf(di, 0.5f, s1);
        float C3[d], g7[d];
#pragma unroll
        for (int k = 0; k < d; ++k) {
          float c = 0.f;
#pragma unroll
          for (int l = 0; l < d; ++l) c = fmaf(A4[k][l], dl[l], c);
          C3[k] = c;
        }
#pragma unroll
        for (int m = 0; m < d; ++m) g7[m] = fmaf(C3[m], q43, A3[m] * w33);
#pragma unroll
        for (int k = 0; k < d; ++k) {
          const float w44 = fmaf(dl[k], q43, S3[k]);
#pragma unroll
          for (int m = 0; m < d; ++m) g7[m] = fmaf(A4[k][m], w44, g7[m]);
        }
        float u = 0.f, v = 0.f, r = 0.f, a2n = 0.f;
#pragma unroll
        for (int k = 0; k < d; ++k) {
          u = fmaf(dl[k], fmaf(C3[k], w42, A3[k] * w32), u);
          v = fmaf(dl[k], fmaf(C3[k], 1.f / 24.f, A3[k] * (1.f / 6.f)), v);
          r = fmaf(dl[k], fmaf(C3[k], 1.f / 6.f, 0.5f * A3[k]), r);
          a2n = fmaf(dl[k], fmaf(C3[k], 0.5f, A3[k]), a2n);
        }
        float* cb = cbuf + p * C::CP + s * kRows * DD;
#pragma unroll
        for (int m = 0; m < d; ++m) cb[m * DD + ij] = g7[m];
        cb[d * DD + ij] = dj * (a2 + r);                      // -> level-1 adjoint, first index i
        cb[(d + 1) * DD + j * d + i] = fmaf(a2, w22, u);      // -> gDelta[j] (transposed: groups of fixed j)
        cb[(d + 2) * DD + ij] = dj * fmaf(a2, 0.5f, v);       // -> gDelta[i]
        a2 += a2n;
#pragma unroll
        for (int k = 0; k < d; ++k) A3[k] += C3[k];
      }
    }
    if (nt_next > 0) pf.store(xs + (buf ^ 1) * C::XS, nt_next);
    __syncthreads();
    if (active) {
      // stage 1, all d^2 threads: every row of d^2 entries summed in groups of d consecutive ones
      constexpr int per_s = kRows * d;
      for (int task = ij; task < nt * per_s; task += DD) {
        const int s = task / per_s, vb = task - s * per_s;
        const float* src = cbuf + p * C::CP + s * kRows * DD + vb * d;
        float acc = 0.f;
#pragma unroll
        for (int e = 0; e < d; ++e) acc += src[e];
        part[(p * CH + s) * per_s + vb] = acc;
      }
    }
    __syncthreads();
    if (active && ij < d) {
      // stage 2, one thread per component m: finish the sums, carry the level-1 adjoint, emit gx
      constexpr int per_s = kRows * d;
      float addv[CH];
#pragma unroll
      for (int s = 0; s < CH; ++s)
        addv[s] = (add && s < nt) ? __ldg(add + n * ld_add + (int64_t)(t_lo + s) * d + ij) : 0.f;
#pragma unroll
      for (int s = CH - 1; s >= 0; --s) {
        if (s < nt) {
          const float* ps = part + (p * CH + s) * per_s;
          float gd = a1 + ps[(d + 1) * d + ij] + ps[(d + 2) * d + ij];
#pragma unroll
          for (int b = 0; b < d; ++b) gd += ps[ij * d + b];
          a1 += ps[d * d + ij];
          out[(n * (int64_t)T + t_lo + s) * d + ij] = fmaf(alpha, addv[s], beta * (gd - gprev));
          gprev = gd;
        }
      }
    }
  }
  if (!basepoint && active && ij < d) {  // point 0 without a basepoint: gx[0] = -gDelta[1]
    float o = beta * (0.f - gprev);
    if (add) o = fmaf(alpha, __ldg(add + n * ld_add + ij), o);
    out[n * (int64_t)T * d + ij] = o;
  }
}

template <int d>
size_t sig4_fwd_smem() {
  using C = Sig4Cfg<d>;
  return (size_t)(2 * C::XS + C::PPC * C::D) * 4;
}
template <int d>
size_t sig4_bwd_smem() {
  using C = Sig4Cfg<d>;
  return (size_t)(2 * C::XS + C::PPC * C::CP + C::PPC * C::PP) * 4;
}

template <int d>
int run_sig4_fwd(const float* x, int64_t N, int64_t T, int basepoint, float* sig, cudaStream_t st) {
  using C = Sig4Cfg<d>;
  const size_t smem = sig4_fwd_smem<d>();
  if (smem > 48 * 1024)
    SIG_CUDA_OK(cudaFuncSetAttribute(sig4_fwd_kernel<d>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  sig4_fwd_kernel<d><<<(unsigned)ceil_div(N, (int64_t)C::PPC), C::NT, smem, st>>>(x, N, (int)T, basepoint, sig);
  SIG_LAUNCH_OK("sig4_fwd_kernel");
  return 0;
}

template <int d>
int run_sig4_bwd(const float* x, const float* sig, const float* gsig, int64_t N, int64_t T, int basepoint,
                 const float* add, int64_t ld_add, float alpha, float beta, float* out, cudaStream_t st,
                 const SigGradSpec& gs) {
  using C = Sig4Cfg<d>;
  const size_t smem = sig4_bwd_smem<d>();
  if (smem > 48 * 1024)
    SIG_CUDA_OK(cudaFuncSetAttribute(sig4_bwd_kernel<d>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  sig4_bwd_kernel<d><<<(unsigned)ceil_div(N, (int64_t)C::PPC), C::NT, smem, st>>>(
      x, sig, gsig, N, (int)T, basepoint, add, ld_add, alpha, beta, out, gs);
  SIG_LAUNCH_OK("sig4_bwd_kernel");
  return 0;
}

}  // namespace

bool sig4_available(int64_t d) { return d >= 2 && d <= 8; }

int launch_sig4_fwd(const float* x, int64_t N, int64_t T, int64_t d, int basepoint, float* sig, cudaStream_t st) {
  switch (d) {
    case 2: return run_sig4_fwd<2>(x, N, T, basepoint, sig, st);
    case 3: return run_sig4_fwd<3>(x, N, T, basepoint, sig, st);
    case 4: return run_sig4_fwd<4>(x, N, T, basepoint, sig, st);
    case 5: return run_sig4_fwd<5>(x, N, T, basepoint, sig, st);
    case 6: return run_sig4_fwd<6>(x, N, T, basepoint, sig, st);
    case 7: return run_sig4_fwd<7>(x, N, T, basepoint, sig, st);
    case 8: return run_sig4_fwd<8>(x, N, T, basepoint, sig, st);
    default: break;
  }
  set_error("signature_fwd: the depth-4 kernels cover 2 <= d <= 8 (got %lld)", (long long)d);
  return SIGSVGD_ERR_UNSUPPORTED;
}

int launch_sig4_bwd(const float* x, const float* sig, const float* gsig, int64_t N, int64_t T, int64_t d,
                    int basepoint, const float* add, int64_t ld_add, float alpha, float beta, float* out,
                    cudaStream_t st, const SigGradSpec* spec) {
  const SigGradSpec gs = spec ? *spec : SigGradSpec{};
#define SIG4_BWD_CASE(dd) \
  case dd:                \
    return run_sig4_bwd<dd>(x, sig, gsig, N, T, basepoint, add, ld_add, alpha, beta, out, st, gs);
  switch (d) {
    SIG4_BWD_CASE(2)
    SIG4_BWD_CASE(3)
    SIG4_BWD_CASE(4)
    SIG4_BWD_CASE(5)
    SIG4_BWD_CASE(6)
    SIG4_BWD_CASE(7)
    SIG4_BWD_CASE(8)
    default: break;
  }
#undef SIG4_BWD_CASE
  set_error("signature_bwd: the depth-4 kernels cover 2 <= d <= 8 (got %lld)", (long long)d);
  return SIGSVGD_ERR_UNSUPPORTED;
}

}  // namespace sigsvgd
