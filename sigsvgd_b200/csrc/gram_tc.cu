// Pairwise squared distances of signature features on the tensor cores (K3).
// Reference: pw_dist_sq, src/utils/math.py:69-86.
//
// Output tile 128 x 128.  Per accumulator stage two TMEM buffers: `main` takes the hi.hi products,
// `corr` the hi.lo + lo.hi products (2^-11 smaller), and the epilogue adds them in fp32 RN.  The
// TMEM accumulator adds with truncation, so keeping the long full-magnitude chain to one MMA per
// K step (25 for D = 399) keeps the bias of the distances at the level of fp32 FMA rounding.
// Symmetric mode computes the upper-triangular tiles only and mirrors them in the epilogue.
#include <cuda.h>

#include "gemm_tc.cuh"
#include "tc_ptx.cuh"

namespace sigsvgd {
namespace {

constexpr int GM = 128, GN = 128, GK = 64;
constexpr int kStages = 3;
constexpr int kThreads = 256;
constexpr uint32_t kPlaneBytes = GM * GK * 2;            // 16 KB: one plane of one operand tile
constexpr uint32_t kStageBytes = 4 * kPlaneBytes;        // A hi/lo + B hi/lo = 64 KB
constexpr uint32_t kBarBytes = 256;
constexpr uint32_t kOutTileBytes = 32 * 32 * 4;          // one 32 x 32 fp32 staging tile (TMA store box)
constexpr uint32_t kColNormBytes = 4 * 32 * 4;           // scaled column norms of one chunk, per epilogue warp
// [operand stages | 4 block tiles | 4 mirror tiles | barriers | column norms] + alignment slack
constexpr uint32_t kSmemBytes = kStages * kStageBytes + 8 * kOutTileBytes + kBarBytes + kColNormBytes + 1024;
static_assert(kSmemBytes <= 232448, "exceeds the 227 KB shared-memory limit of sm_100");
constexpr uint32_t kTmemCols = 512;                      // 2 stages x (main 128 + corr 128)

struct GramParams {
  int64_t M, N;
  int num_k16;
  int tiles_m, tiles_n;  // symmetric: tiles_m == tiles_n, work = upper triangle
  int num_work;
  float* out;
  int64_t ldo;
  const float* na;
  const float* nb;
  const float* scal;
  // batch of independent problems stacked along the rows of A, B and out (batch == 1: one problem);
  // scal + b * scal_stride is the scale record of problem b
  int batch;
  int64_t scal_stride;
};

__device__ __forceinline__ void decode_sym(int t, int T, int& tm, int& tn) {
  // row-wise enumeration of the upper triangle: row m starts at m*T - m(m-1)/2
  const float b = 2.f * T + 1.f;
  int m = (int)((b - sqrtf(fmaxf(b * b - 8.f * t, 0.f))) * 0.5f);
  m = max(0, min(T - 1, m));
  auto start = [&](int r) { return r * T - (r * (r - 1)) / 2; };
  while (m + 1 < T && start(m + 1) <= t) ++m;
  while (m > 0 && start(m) > t) --m;
  tm = m;
  tn = m + (t - start(m));
}

template <bool kSym, bool kTma>
__global__ void __launch_bounds__(kThreads, 1)
gram_kernel(const __grid_constant__ CUtensorMap tmAh, const __grid_constant__ CUtensorMap tmAl,
            const __grid_constant__ CUtensorMap tmBh, const __grid_constant__ CUtensorMap tmBl,
            const __grid_constant__ CUtensorMap tmOut, const GramParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t sbase = (raw + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (sbase - raw);
  const uint32_t bar0 = sbase + kStages * kStageBytes + 8 * kOutTileBytes;
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (kStages + s); };
  auto tfull_bar = [&](int a) { return bar0 + 8u * (2 * kStages + a); };
  auto tempty_bar = [&](int a) { return bar0 + 8u * (2 * kStages + 2 + a); };
  volatile uint32_t* tmem_slot =
      reinterpret_cast<volatile uint32_t*>(smem + kStages * kStageBytes + 8 * kOutTileBytes + 8 * (2 * kStages + 4));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    ptx::tma_prefetch_desc(&tmAh);
    ptx::tma_prefetch_desc(&tmAl);
    ptx::tma_prefetch_desc(&tmBh);
    ptx::tma_prefetch_desc(&tmBl);
    if (kTma) ptx::tma_prefetch_desc(&tmOut);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      ptx::mbar_init(full_bar(s), 1);
      ptx::mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      ptx::mbar_init(tfull_bar(a), 1);
      ptx::mbar_init(tempty_bar(a), 4);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 2) ptx::tmem_alloc(ptx::smem_u32((const void*)tmem_slot), kTmemCols);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int num_kb = (p.num_k16 + 3) / 4;
  const int work_per_problem = p.num_work / p.batch;
  auto decode = [&](int w, int& tm, int& tn) {
    w %= work_per_problem;
    if constexpr (kSym) {
      decode_sym(w, p.tiles_m, tm, tn);
    } else {
      tm = w / p.tiles_n;
      tn = w - tm * p.tiles_n;
    }
  };

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      uint32_t it = 0;
      for (int w = blockIdx.x; w < p.num_work; w += gridDim.x) {
        int tm, tn;
        decode(w, tm, tn);
        const int b = w / work_per_problem;
        const int ra = (int)(b * p.M) + tm * GM, rb = (int)(b * p.N) + tn * GN;
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          const int s = it % kStages;
          const uint32_t ph = (it / kStages) & 1u;
          ptx::mbar_wait(empty_bar(s), ph ^ 1u);
          const uint32_t st = sbase + s * kStageBytes;
          ptx::mbar_expect_tx(full_bar(s), kStageBytes);
          ptx::tma_load_2d(st, &tmAh, full_bar(s), kb * GK, ra);
          ptx::tma_load_2d(st + kPlaneBytes, &tmAl, full_bar(s), kb * GK, ra);
          ptx::tma_load_2d(st + 2 * kPlaneBytes, &tmBh, full_bar(s), kb * GK, rb);
          ptx::tma_load_2d(st + 3 * kPlaneBytes, &tmBl, full_bar(s), kb * GK, rb);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = ptx::make_idesc_f16(GM, GN, false, false);
      uint32_t it = 0, tcount = 0;
      for (int w = blockIdx.x; w < p.num_work; w += gridDim.x, ++tcount) {
        const int a = tcount & 1;
        const uint32_t aph = (tcount >> 1) & 1u;
        ptx::mbar_wait(tempty_bar(a), aph ^ 1u);
        ptx::tc_fence_after();
        const uint32_t t_main = tmem_base + a * 256, t_corr = t_main + 128;
        uint32_t acc = 0;
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          const int s = it % kStages;
          const uint32_t ph = (it / kStages) & 1u;
          ptx::mbar_wait(full_bar(s), ph);
          ptx::tc_fence_after();
          const uint32_t st = sbase + s * kStageBytes;
          const int steps = min(4, p.num_k16 - kb * 4);
          for (int k = 0; k < steps; ++k) {
            const uint64_t ah = ptx::make_smem_desc(st + k * 32, 16, 1024, ptx::kLayoutSw128);
            const uint64_t al = ptx::make_smem_desc(st + kPlaneBytes + k * 32, 16, 1024, ptx::kLayoutSw128);
            const uint64_t bh = ptx::make_smem_desc(st + 2 * kPlaneBytes + k * 32, 16, 1024, ptx::kLayoutSw128);
            const uint64_t bl = ptx::make_smem_desc(st + 3 * kPlaneBytes + k * 32, 16, 1024, ptx::kLayoutSw128);
            ptx::umma_f16(t_main, ah, bh, idesc, acc);
            ptx::umma_f16(t_corr, ah, bl, idesc, acc);
            ptx::umma_f16(t_corr, al, bh, idesc, 1u);
            acc = 1u;
          }
          ptx::umma_commit(empty_bar(s));
        }
        ptx::umma_commit(tfull_bar(a));
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------------ epilogue
    // A thread owns one row (TMEM lane) of the tile and walks it in 4 chunks of 32 columns.
    //  fast path (16-byte aligned output): the chunk is staged in shared memory in the
    //    128-byte-swizzled box layout (two staging tiles per warp, alternating) and leaves through
    //    a TMA store; in symmetric mode its transpose is stored straight from the registers (lanes
    //    hold consecutive rows = consecutive addresses of the mirrored block);
    //  slow path (unaligned output): per-element guarded stores.
    const int q = warp & 3;
    uint32_t tcount = 0;
    uint8_t* const stg0 = smem + kStages * kStageBytes + q * (2 * kOutTileBytes);
    const uint32_t stg0_u = sbase + kStages * kStageBytes + q * (2 * kOutTileBytes);
    float* xp = reinterpret_cast<float*>(stg0);  // slow path: 32 x 32 transpose tile, XOR-swizzled
    float* cn = reinterpret_cast<float*>(smem + kStages * kStageBytes + 8 * kOutTileBytes + kBarBytes) + q * 32;
    uint32_t nstore = 0;  // TMA stores committed by lane 0 (staging tile = nstore & 1)

    for (int w = blockIdx.x; w < p.num_work; w += gridDim.x, ++tcount) {
      int tm, tn;
      decode(w, tm, tn);
      const int64_t pb = w / work_per_problem;  // problem of this tile
      const float scal0 = __ldg(p.scal + pb * p.scal_stride);  // alpha^-2, a power of two: scaling commutes with rounding
      const float m2s = -2.f * scal0;
      const float* const na_b = p.na + pb * p.M;
      const float* const nb_b = p.nb + pb * p.N;
      float* const out_b = p.out + pb * p.M * p.ldo;
      const int a = tcount & 1;
      const uint32_t aph = (tcount >> 1) & 1u;
      const int64_t row0 = (int64_t)tm * GM + q * 32;  // first row of this warp (within the problem)
      const int64_t row = row0 + lane;
      const bool row_ok = row < p.M;
      const bool diag = kSym && tm == tn;
      // norms of this thread's row and of the tile's columns (lane-strided), fetched while the
      // accumulator is still being produced
      const float rvs = row_ok ? __ldg(na_b + row) * scal0 : 0.f;
      float cnv[GN / 32];
#pragma unroll
      for (int k = 0; k < GN / 32; ++k) {
        const int64_t col = (int64_t)tn * GN + k * 32 + lane;
        cnv[k] = col < p.N ? __ldg(nb_b + col) * scal0 : 0.f;
      }
      ptx::mbar_wait(tfull_bar(a), aph);
      ptx::tc_fence_after();
      if (kTma) {
        // ------------------------------------------------------------ fast path
#pragma unroll
        for (int c = 0; c < GN / 32; ++c) {
          const int col0 = tn * GN + c * 32;
          if (col0 >= p.N) break;  // warp-uniform
          // 32 x 32 blocks of a diagonal tile: above the diagonal -> like any other block (stored
          // with its mirror image); on the diagonal -> symmetrised inside the warp; below -> skipped
          // (it is the mirror image of a block another warp stores)
          const int blk = !diag ? 0 : (c > q ? 0 : (c == q ? 1 : 2));
          if (blk == 2) continue;
          uint32_t v[32], u[32];
          const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + a * 256 + c * 32;
          ptx::tmem_ld_32x32(taddr, v);
          ptx::tmem_ld_32x32(taddr + 128, u);
          __syncwarp();
          cn[lane] = cnv[c];
          __syncwarp();
          float cnr[32];
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const float4 f = *reinterpret_cast<const float4*>(cn + 4 * k);
            cnr[4 * k] = f.x, cnr[4 * k + 1] = f.y, cnr[4 * k + 2] = f.z, cnr[4 * k + 3] = f.w;
          }
          const int ncol = (int)min((int64_t)32, p.N - col0);
          ptx::tmem_ld_wait();
          float o[32];
#pragma unroll
          for (int e = 0; e < 32; ++e) {
            const float acc = __uint_as_float(v[e]) + __uint_as_float(u[e]);
            o[e] = fmaxf(fmaf(m2s, acc, rvs + cnr[e]), 0.f);
          }
          // staging tile of this chunk; the TMA store issued two chunks ago read the same tile
          uint8_t* const stg = stg0 + (nstore & 1u) * kOutTileBytes;
          const uint32_t stg_u = stg0_u + (nstore & 1u) * kOutTileBytes;
          if (nstore >= 2) {
            if (lane == 0) ptx::bulk_wait_read1();
            __syncwarp();
          }
          if (kSym && blk == 1) {
            // diagonal block: (r, c) and (c, r) were accumulated in different orders; keep the upper
            // triangle, take the lower one from the transposed tile, exact zeros on the diagonal
#pragma unroll
            for (int e = 0; e < 32; ++e)
              *reinterpret_cast<float*>(stg + e * 128 + ((((lane >> 2) ^ (e & 7)) << 4) | ((lane & 3) << 2))) = o[e];
            __syncwarp();
#pragma unroll
            for (int k = 0; k < 8; ++k) {
              const float4 f = *reinterpret_cast<const float4*>(stg + lane * 128 + ((k ^ (lane & 7)) << 4));
              if (4 * k + 0 < lane) o[4 * k + 0] = f.x;
              if (4 * k + 1 < lane) o[4 * k + 1] = f.y;
              if (4 * k + 2 < lane) o[4 * k + 2] = f.z;
              if (4 * k + 3 < lane) o[4 * k + 3] = f.w;
            }
#pragma unroll
            for (int e = 0; e < 32; ++e)
              if (e == lane) o[e] = 0.f;
            __syncwarp();
          }
          // block: row `lane`, 16-byte chunk k at (k ^ (lane & 7)) -- the SWIZZLE_128B box layout
#pragma unroll
          for (int k = 0; k < 8; ++k)
            *reinterpret_cast<float4*>(stg + lane * 128 + ((k ^ (lane & 7)) << 4)) =
                make_float4(o[4 * k], o[4 * k + 1], o[4 * k + 2], o[4 * k + 3]);
          ptx::fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            ptx::tma_store_2d(&tmOut, stg_u, col0, (int)(pb * p.M + row0));  // clipped at the matrix edges
            ptx::bulk_commit();
          }
          ++nstore;
          if (kSym && blk == 0) {
            // mirrored block: entry (row, col0 + e) -> out[col0 + e][row]; one coalesced line per e
            float* mp = out_b + (int64_t)col0 * p.ldo + row;
            if (ncol == 32 && row0 + 32 <= p.M) {
#pragma unroll
              for (int e = 0; e < 32; ++e) mp[(int64_t)e * p.ldo] = o[e];
            } else if (row_ok) {
#pragma unroll
              for (int e = 0; e < 32; ++e)
                if (e < ncol) mp[(int64_t)e * p.ldo] = o[e];
            }
          }
        }
      } else {
        // ------------------------------------------------------------ slow path
        // symmetric mode: an entry above the diagonal stands for two matrix entries (it is stored
        // twice and counted twice), the diagonal for one, and entries below the diagonal (they only
        // occur inside diagonal tiles) are dropped in favour of their mirror image
#pragma unroll 1
        for (int c = 0; c < GN / 32; ++c) {
          const int64_t col0 = (int64_t)tn * GN + c * 32;
          if (col0 >= p.N) break;  // warp-uniform
          uint32_t v[32], u[32];
          const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + a * 256 + c * 32;
          ptx::tmem_ld_32x32(taddr, v);
          ptx::tmem_ld_32x32(taddr + 128, u);
          const float cvl = c == 0 ? cnv[0] : c == 1 ? cnv[1] : c == 2 ? cnv[2] : cnv[3];
          // bit e set: column col0+e is inside the matrix and (symmetric) strictly above the diagonal
          const int64_t ncol = min((int64_t)32, p.N - col0);
          const uint32_t inmask = ncol >= 32 ? 0xffffffffu : ((1u << ncol) - 1u);
          uint32_t upmask = 0u, dgmask = 0u;  // mirrored entries / plain entries of this thread
          if (row_ok) {
            if (!kSym) {
              dgmask = inmask;
            } else if (!diag) {
              upmask = inmask;
            } else {
              const int64_t rel = row - col0;  // column index of the diagonal inside this chunk
              if (rel < 0) upmask = inmask;
              else if (rel < 32) {
                upmask = inmask & ~((2u << rel) - 1u);  // 2u << 31 == 0 -> all lower: mask 0 ok
                if (rel == 31) upmask = 0u;
                dgmask = inmask & (1u << rel);
              }
            }
          }
          ptx::tmem_ld_wait();
          float o[32];
#pragma unroll
          for (int e = 0; e < 32; ++e) {
            const float cv = __shfl_sync(0xffffffffu, cvl, e);
            const float acc = __uint_as_float(v[e]) + __uint_as_float(u[e]);
            float dv = fmaxf(fmaf(m2s, acc, rvs + cv), 0.f);
            if (kSym && ((dgmask >> e) & 1u)) dv = 0.f;
            o[e] = dv;
          }
          // transposed block straight from registers: lanes hold consecutive rows, so every store
          // instruction writes one coalesced line of the mirrored block
          if (kSym && upmask) {
#pragma unroll
            for (int e = 0; e < 32; ++e)
              if ((upmask >> e) & 1u) out_b[(col0 + e) * p.ldo + row] = o[e];
          }
          // direct block through a 32 x 32 shared-memory transpose so that each store instruction
          // writes 32 consecutive columns of one row (a thread owns a row in TMEM)
#pragma unroll
          for (int e = 0; e < 32; ++e) xp[lane * 32 + (e ^ lane)] = o[e];
          __syncwarp();
          const int64_t col = col0 + lane;
          const bool col_ok = col < p.N;
#pragma unroll 8
          for (int r = 0; r < 32; ++r) {
            const int64_t rr = row0 + r;
            if (rr < p.M && col_ok && (!diag || col >= rr)) out_b[rr * p.ldo + col] = xp[r * 32 + (lane ^ r)];
          }
          __syncwarp();
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(tempty_bar(a));
    }
    if (nstore && lane == 0) ptx::bulk_wait0();
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, kTmemCols);
  }
}

}  // namespace

int make_tmap_f16(CUtensorMap* tm, const __half* base, int64_t inner, int64_t outer, int64_t ld,
                  uint32_t box_inner, uint32_t box_outer);  // gemm_tc.cu
int make_tmap_f32(CUtensorMap* tm, const float* base, int64_t inner, int64_t outer, int64_t ld,
                  uint32_t box_inner, uint32_t box_outer);  // gemm_tc.cu

int tc_gram(TcPlanes A, TcPlanes B, int64_t M, int64_t N, int64_t D, const float* na,
            const float* nb, const float* scal, int symmetric, float* dist, int64_t ldd,
            cudaStream_t st, int64_t batch, int64_t scal_stride) {
  SIG_CHECK_ARG(batch >= 1 && (batch == 1 || (M % GM == 0 && N % GN == 0)),
                "tc_gram: batched problems need M and N to be multiples of 128 (M=%lld N=%lld)", (long long)M,
                (long long)N);
  SIG_CHECK_ARG(M >= 1 && N >= 1 && D >= 1 && A.ld >= D && B.ld >= D && ldd >= N,
                "tc_gram: bad shape M=%lld N=%lld D=%lld", (long long)M, (long long)N, (long long)D);
  SIG_CHECK_ARG(M < (1ll << 30) && N < (1ll << 30), "tc_gram: too many rows");
  SIG_CHECK_ARG(!symmetric || (M == N && A.hi == B.hi && A.lo == B.lo && na == nb),
                "tc_gram: symmetric mode needs identical operands");
  CUtensorMap ah, al, bh, bl;
  SIG_PROPAGATE(make_tmap_f16(&ah, A.hi, D, batch * M, A.ld, GK, GM));
  SIG_PROPAGATE(make_tmap_f16(&al, A.lo, D, batch * M, A.ld, GK, GM));
  SIG_PROPAGATE(make_tmap_f16(&bh, B.hi, D, batch * N, B.ld, GK, GN));
  SIG_PROPAGATE(make_tmap_f16(&bl, B.lo, D, batch * N, B.ld, GK, GN));
  GramParams p;
  p.M = M, p.N = N;
  p.num_k16 = (int)ceil_div(D, 16);
  p.tiles_m = (int)ceil_div(M, GM), p.tiles_n = (int)ceil_div(N, GN);
  const int64_t work =
      batch * (symmetric ? (int64_t)p.tiles_m * (p.tiles_m + 1) / 2 : (int64_t)p.tiles_m * p.tiles_n);
  SIG_CHECK_ARG(work < (1ll << 31) && (!symmetric || p.tiles_m <= 2048), "tc_gram: too many tiles");
  p.num_work = (int)work;
  p.out = dist, p.ldo = ldd;
  p.na = na, p.nb = nb, p.scal = scal;
  p.batch = (int)batch, p.scal_stride = scal_stride;
  // TMA-store epilogue needs a 16-byte aligned output with a 16-byte row pitch
  const bool tma_out = (reinterpret_cast<uintptr_t>(dist) & 15u) == 0 && (ldd & 3) == 0;
  CUtensorMap to = ah;  // placeholder when unused
  SIG_CHECK_ARG(batch == 1 || tma_out, "tc_gram: batched problems need a 16-byte aligned output with ldd %% 4 == 0");
  if (tma_out) SIG_PROPAGATE(make_tmap_f32(&to, dist, N, batch * M, ldd, 32, 32));
  static thread_local int configured_dev = -1;
  int dev = 0;
  SIG_CUDA_OK(cudaGetDevice(&dev));
  if (configured_dev != dev) {
    SIG_CUDA_OK(cudaFuncSetAttribute(gram_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
    SIG_CUDA_OK(cudaFuncSetAttribute(gram_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
    SIG_CUDA_OK(cudaFuncSetAttribute(gram_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
    SIG_CUDA_OK(cudaFuncSetAttribute(gram_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
    configured_dev = dev;
  }
  const int grid = (int)std::min<int64_t>(work, num_sms());
  if (symmetric && tma_out)
    gram_kernel<true, true><<<grid, kThreads, kSmemBytes, st>>>(ah, al, bh, bl, to, p);
  else if (symmetric)
    gram_kernel<true, false><<<grid, kThreads, kSmemBytes, st>>>(ah, al, bh, bl, to, p);
  else if (tma_out)
    gram_kernel<false, true><<<grid, kThreads, kSmemBytes, st>>>(ah, al, bh, bl, to, p);
  else
    gram_kernel<false, false><<<grid, kThreads, kSmemBytes, st>>>(ah, al, bh, bl, to, p);
  SIG_LAUNCH_OK("gram_kernel");
  return 0;
}

}  // namespace sigsvgd
