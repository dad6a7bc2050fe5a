// fp32 SIMT (FFMA) GEMMs: the exact-fp32 path used for small swarms (N below the tcgen05
// threshold: MPC with 35 policies, planning with 20-256 particles, batched 128-particle scenes)
// and as the on-device cross-check of the tensor-core path.
//
//  sqdist_nt : dist[i,j] = max(0, |a_i|^2 + |b_j|^2 - 2 a_i.b_j)        (src/utils/math.py:69-86)
//  gemm_nn   : out[i,c]  = sum_j a[i,j] v[j,c]                           (src/inference/svgd.py:82)
//
// 64x64x16 tiles, 256 threads, 4x4 outputs per thread laid out with stride 16 so shared-memory
// reads are conflict free.  Batched through blockIdx.z.
#include "common.cuh"

namespace sigsvgd {
namespace {

constexpr int BM = 64, BN = 64, BK = 16, NT = 256;

// A tile: As[r][k] = a[i0+r][k0+k]   (rows x BK, padded)
__device__ __forceinline__ void load_rows_tile(const float* __restrict__ a, int64_t lda, int64_t i0,
                                               int64_t nrows, int64_t k0, int64_t K,
                                               float (*As)[BK + 1], int tid) {
#pragma unroll
  for (int l = 0; l < (BM * BK) / NT; ++l) {
    const int idx = tid + l * NT;
    const int r = idx / BK, c = idx % BK;
    const int64_t gi = i0 + r, gk = k0 + c;
    As[r][c] = (gi < nrows && gk < K) ? __ldg(a + gi * lda + gk) : 0.f;
  }
}

__global__ void __launch_bounds__(NT)
sqdist_nt_kernel(const float* __restrict__ a, int64_t lda, const float* __restrict__ b, int64_t ldb,
                 int64_t M, int64_t N, int64_t K, float* __restrict__ dist, int64_t ldd,
                 int64_t stride_a, int64_t stride_b, int64_t stride_d) {
  __shared__ float As[BM][BK + 1];
  __shared__ float Bs[BN][BK + 1];
  __shared__ float na[BM], nb[BN];
  a += blockIdx.z * stride_a;
  b += blockIdx.z * stride_b;
  dist += blockIdx.z * stride_d;
  const int tid = threadIdx.x, tx = tid % 16, ty = tid / 16;
  const int64_t i0 = (int64_t)blockIdx.y * BM, j0 = (int64_t)blockIdx.x * BN;
  float acc[4][4] = {};
  float nrm = 0.f;  // threads 0..63: |a_row|^2 ; threads 64..127: |b_row|^2
  for (int64_t k0 = 0; k0 < K; k0 += BK) {
    load_rows_tile(a, lda, i0, M, k0, K, As, tid);
    load_rows_tile(b, ldb, j0, N, k0, K, Bs, tid);
    __syncthreads();
    if (tid < 128) {
      const float(*T)[BK + 1] = tid < 64 ? As : Bs;
      const int r = tid & 63;
#pragma unroll
      for (int k = 0; k < BK; ++k) nrm = fmaf(T[r][k], T[r][k], nrm);
    }
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float av[4], bv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) av[i] = As[ty + 16 * i][k];
#pragma unroll
      for (int j = 0; j < 4; ++j) bv[j] = Bs[tx + 16 * j][k];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
  if (tid < 64) na[tid] = nrm;
  else if (tid < 128) nb[tid - 64] = nrm;
  __syncthreads();
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t gi = i0 + ty + 16 * i;
    if (gi >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t gj = j0 + tx + 16 * j;
      if (gj >= N) continue;
      // (na + nb) first: commutative, so d_ij == d_ji bit-for-bit when a == b (math.py:83-85 adds
      // the norms one after the other, which is what makes the reference's matrix asymmetric)
      const float v = fmaf(-2.f, acc[i][j], na[ty + 16 * i] + nb[tx + 16 * j]);
      dist[gi * ldd + gj] = fmaxf(v, 0.f);
    }
  }
}

__global__ void __launch_bounds__(NT)
gemm_nn_kernel(const float* __restrict__ a, int64_t lda, const float* __restrict__ v, int64_t ldv,
               int64_t M, int64_t K, int64_t C, float* __restrict__ out, int64_t ldo,
               int64_t stride_a, int64_t stride_v, int64_t stride_o) {
  __shared__ float As[BM][BK + 1];
  __shared__ float Vs[BK][BN];
  a += blockIdx.z * stride_a;
  v += blockIdx.z * stride_v;
  out += blockIdx.z * stride_o;
  const int tid = threadIdx.x, tx = tid % 16, ty = tid / 16;
  const int64_t i0 = (int64_t)blockIdx.y * BM, c0 = (int64_t)blockIdx.x * BN;
  float acc[4][4] = {};
  for (int64_t k0 = 0; k0 < K; k0 += BK) {
    load_rows_tile(a, lda, i0, M, k0, K, As, tid);
#pragma unroll
    for (int l = 0; l < (BK * BN) / NT; ++l) {
      const int idx = tid + l * NT;
      const int r = idx / BN, c = idx % BN;
      const int64_t gk = k0 + r, gc = c0 + c;
      Vs[r][c] = (gk < K && gc < C) ? __ldg(v + gk * ldv + gc) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float av[4], bv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) av[i] = As[ty + 16 * i][k];
#pragma unroll
      for (int j = 0; j < 4; ++j) bv[j] = Vs[k][tx + 16 * j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int64_t gi = i0 + ty + 16 * i;
    if (gi >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int64_t gc = c0 + tx + 16 * j;
      if (gc < C) out[gi * ldo + gc] = acc[i][j];
    }
  }
}

}  // namespace

int launch_sqdist_simt(const float* a, int64_t lda, const float* b, int64_t ldb, int64_t M,
                       int64_t N, int64_t D, float* dist, int64_t ldd, int64_t batch,
                       int64_t stride_a, int64_t stride_b, int64_t stride_d, cudaStream_t st) {
  SIG_CHECK_ARG(M >= 0 && N >= 0 && D >= 1 && lda >= D && ldb >= D && ldd >= N && batch >= 1,
                "sqdist: bad shape M=%lld N=%lld D=%lld", (long long)M, (long long)N, (long long)D);
  if (M == 0 || N == 0) return 0;
  const int64_t gx = ceil_div(N, BN), gy = ceil_div(M, BM);
  SIG_CHECK_ARG(gy <= 65535 && batch <= 65535, "sqdist: grid too large (M=%lld, batch=%lld)",
                (long long)M, (long long)batch);
  dim3 grid((unsigned)gx, (unsigned)gy, (unsigned)batch);
  sqdist_nt_kernel<<<grid, NT, 0, st>>>(a, lda, b, ldb, M, N, D, dist, ldd, stride_a, stride_b,
                                        stride_d);
  SIG_LAUNCH_OK("sqdist_nt_kernel");
  return 0;
}

int launch_gemm_nn_simt(const float* a, int64_t lda, const float* v, int64_t ldv, int64_t M,
                        int64_t N, int64_t C, float* out, int64_t ldo, int64_t batch,
                        int64_t stride_a, int64_t stride_v, int64_t stride_o, cudaStream_t st) {
  SIG_CHECK_ARG(M >= 0 && N >= 1 && C >= 1 && lda >= N && ldv >= C && ldo >= C && batch >= 1,
                "matmul_nn: bad shape M=%lld N=%lld C=%lld", (long long)M, (long long)N,
                (long long)C);
  if (M == 0) return 0;
  const int64_t gx = ceil_div(C, BN), gy = ceil_div(M, BM);
  SIG_CHECK_ARG(gy <= 65535 && batch <= 65535, "matmul_nn: grid too large");
  dim3 grid((unsigned)gx, (unsigned)gy, (unsigned)batch);
  gemm_nn_kernel<<<grid, NT, 0, st>>>(a, lda, v, ldv, M, N, C, out, ldo, stride_a, stride_v,
                                      stride_o);
  SIG_LAUNCH_OK("gemm_nn_kernel");
  return 0;
}

}  // namespace sigsvgd
