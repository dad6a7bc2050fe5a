// Split-fp16 (3-MMA) tensor-core GEMM for sm_100a: TMA-staged operand tiles, tcgen05.mma with
// fp32 accumulators in TMEM, warp-specialised persistent CTAs.
//
//   tc_sqdist    : pairwise squared distances of signature features   (src/utils/math.py:69-86)
//   tc_matmul_nn : K . [S | 1 | score]      (src/inference/svgd.py:82 and the K.S / row-sum terms
//                  of autograd.grad(k_xx.sum(), X), svgd.py:43)
//
// CTA = 256 threads: warp 0 TMA producer, warp 1 MMA issuer (one elected lane each), warp 2 TMEM
// allocator, warps 4-7 epilogue (TMEM -> registers -> global).  Output tile 128 x 256 (TMEM lanes
// x columns), two accumulator stages (all 512 TMEM columns) so the epilogue of tile t overlaps
// the MMAs of tile t+1; K is consumed 64 elements per smem stage (one 128-byte swizzle atom).
#include <cuda.h>

#include "gemm_tc.cuh"
#include "tc_ptx.cuh"

namespace sigsvgd {
namespace {

constexpr int BM = 128, BN = 256, BK = 64;
constexpr int kStages = 2;
constexpr int kGemmThreads = 256;
constexpr uint32_t kABytes = BM * BK * 2;   // 16 KB per plane
constexpr uint32_t kBBytes = BN * BK * 2;   // 32 KB per plane
constexpr uint32_t kStageBytes = 2 * kABytes + 2 * kBBytes;  // hi+lo of A and B: 96 KB
constexpr uint32_t kBarBytes = 256;
constexpr int kXposePitch = 33;                               // 32 x 32 transpose tile per epilogue warp
constexpr uint32_t kXposeBytes = 4 * 32 * kXposePitch * 4;
constexpr uint32_t kSmemBytes = kStages * kStageBytes + kBarBytes + kXposeBytes + 1024;  // + alignment slack
constexpr uint32_t kTmemCols = 512;
constexpr int kMaxChainKb = 16;

enum { EPI_SQDIST = 0, EPI_COLSCALE = 1 };

struct GemmParams {
  int64_t M, N;        // output rows / columns
  int num_k16;         // ceil(Kdim / 16): number of UMMA K steps
  int tiles_m, tiles_n;
  int nsplit;          // K is cut into nsplit chunks of kb_per_split 64-wide blocks; chunks of one
  int kb_per_split;    // tile are separate work items whose results are summed with red.add
                       // (bounds the length of the truncating TMEM accumulation chain)
  float* out;
  int64_t ldo;
  const float* row_vec;  // EPI_SQDIST: na[M]
  const float* col_vec;  // EPI_SQDIST: nb[N] ; EPI_COLSCALE: colscale[N]
  const float* scal;     // EPI_SQDIST: scal[0] = alpha^-2 (device)
};

template <bool kBMN, int kEpi>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm3x_kernel(const __grid_constant__ CUtensorMap tmAh, const __grid_constant__ CUtensorMap tmAl,
              const __grid_constant__ CUtensorMap tmBh, const __grid_constant__ CUtensorMap tmBl,
              const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t sbase = (raw + 1023u) & ~1023u;  // SWIZZLE_128B atoms need 1024-byte alignment
  uint8_t* smem = smem_raw + (sbase - raw);
  const uint32_t bar0 = sbase + kStages * kStageBytes;
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (kStages + s); };
  auto tfull_bar = [&](int a) { return bar0 + 8u * (2 * kStages + a); };
  auto tempty_bar = [&](int a) { return bar0 + 8u * (2 * kStages + 2 + a); };
  volatile uint32_t* tmem_slot =
      reinterpret_cast<volatile uint32_t*>(smem + kStages * kStageBytes + 8 * (2 * kStages + 4));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    ptx::tma_prefetch_desc(&tmAh);
    ptx::tma_prefetch_desc(&tmAl);
    ptx::tma_prefetch_desc(&tmBh);
    ptx::tma_prefetch_desc(&tmBl);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      ptx::mbar_init(full_bar(s), 1);
      ptx::mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      ptx::mbar_init(tfull_bar(a), 1);
      ptx::mbar_init(tempty_bar(a), 4);  // one arrive per epilogue warp
    }
    ptx::fence_mbar_init();
  }
  if (warp == 2) ptx::tmem_alloc(ptx::smem_u32((const void*)tmem_slot), kTmemCols);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // work item w -> (column tile, row tile, K chunk); the column tile is the slowest index so the
  // (possibly narrow) last column tile is scheduled last
  const int num_work = p.tiles_m * p.tiles_n * p.nsplit;
  const int num_kb_total = (p.num_k16 + 3) / 4;
  struct Work {
    int tm, tn, kb0, kb1, ncols;
  };
  auto decode = [&](int w) {
    Work k;
    k.tn = w / (p.tiles_m * p.nsplit);
    const int rem = w - k.tn * (p.tiles_m * p.nsplit);
    k.tm = rem / p.nsplit;
    const int sp = rem - k.tm * p.nsplit;
    k.kb0 = sp * p.kb_per_split;
    k.kb1 = min(num_kb_total, k.kb0 + p.kb_per_split);
    // MMA N of this tile: multiple of 16, at most BN
    k.ncols = (int)min((int64_t)BN, ((p.N - (int64_t)k.tn * BN) + 15) / 16 * 16);
    return k;
  };

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      uint32_t it = 0;
      for (int w = blockIdx.x; w < num_work; w += gridDim.x) {
        const Work wk = decode(w);
        const int tm = wk.tm, tn = wk.tn;
        const int nbox = (wk.ncols + 63) / 64;  // MN-major B: 64-column boxes actually needed
        for (int kb = wk.kb0; kb < wk.kb1; ++kb, ++it) {
          const int s = it % kStages;
          const uint32_t ph = (it / kStages) & 1u;
          ptx::mbar_wait(empty_bar(s), ph ^ 1u);
          const uint32_t st = sbase + s * kStageBytes;
          ptx::mbar_expect_tx(full_bar(s), kBMN ? 2 * kABytes + 2 * nbox * (BK * 128) : kStageBytes);
          ptx::tma_load_2d(st, &tmAh, full_bar(s), kb * BK, tm * BM);
          ptx::tma_load_2d(st + kABytes, &tmAl, full_bar(s), kb * BK, tm * BM);
          if constexpr (!kBMN) {
            ptx::tma_load_2d(st + 2 * kABytes, &tmBh, full_bar(s), kb * BK, tn * BN);
            ptx::tma_load_2d(st + 2 * kABytes + kBBytes, &tmBl, full_bar(s), kb * BK, tn * BN);
          } else {
            // MN-major B: four [64 columns x BK rows] boxes per plane, 8 KB each
#pragma unroll
            for (int q = 0; q < BN / 64; ++q) {
              if (q >= nbox) break;
              ptx::tma_load_2d(st + 2 * kABytes + q * (BK * 128), &tmBh, full_bar(s),
                               tn * BN + q * 64, kb * BK);
              ptx::tma_load_2d(st + 2 * kABytes + kBBytes + q * (BK * 128), &tmBl, full_bar(s),
                               tn * BN + q * 64, kb * BK);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      uint32_t it = 0, tcount = 0;
      for (int w = blockIdx.x; w < num_work; w += gridDim.x, ++tcount) {
        const Work wk = decode(w);
        const uint32_t idesc = ptx::make_idesc_f16(BM, wk.ncols, false, kBMN);
        const int a = tcount & 1;
        const uint32_t aph = (tcount >> 1) & 1u;
        ptx::mbar_wait(tempty_bar(a), aph ^ 1u);
        ptx::tc_fence_after();
        const uint32_t tacc = tmem_base + a * BN;
        uint32_t accumulate = 0;
        for (int kb = wk.kb0; kb < wk.kb1; ++kb, ++it) {
          const int s = it % kStages;
          const uint32_t ph = (it / kStages) & 1u;
          ptx::mbar_wait(full_bar(s), ph);
          ptx::tc_fence_after();
          const uint32_t st = sbase + s * kStageBytes;
          const int steps = min(4, p.num_k16 - kb * 4);
          for (int k = 0; k < steps; ++k) {
            // K-major SWIZZLE_128B: 8-row groups 1024 B apart, 16 K-elements = 32 bytes
            const uint64_t ah = ptx::make_smem_desc(st + k * 32, 16, 1024, ptx::kLayoutSw128);
            const uint64_t al = ptx::make_smem_desc(st + kABytes + k * 32, 16, 1024, ptx::kLayoutSw128);
            uint64_t bh, bl;
            if constexpr (!kBMN) {
              bh = ptx::make_smem_desc(st + 2 * kABytes + k * 32, 16, 1024, ptx::kLayoutSw128);
              bl = ptx::make_smem_desc(st + 2 * kABytes + kBBytes + k * 32, 16, 1024, ptx::kLayoutSw128);
            } else {
              // MN-major SWIZZLE_128B: 64-column groups BK*128 B apart (LBO), 8-row K groups
              // 1024 B apart (SBO); 16 K-rows = 2048 bytes
              bh = ptx::make_smem_desc(st + 2 * kABytes + k * 2048, BK * 128, 1024, ptx::kLayoutSw128);
              bl = ptx::make_smem_desc(st + 2 * kABytes + kBBytes + k * 2048, BK * 128, 1024,
                                       ptx::kLayoutSw128);
            }
            ptx::umma_f16(tacc, ah, bh, idesc, accumulate);
            ptx::umma_f16(tacc, ah, bl, idesc, 1u);
            ptx::umma_f16(tacc, al, bh, idesc, 1u);
            accumulate = 1u;
          }
          ptx::umma_commit(empty_bar(s));  // smem stage reusable once these MMAs retire
        }
        ptx::umma_commit(tfull_bar(a));  // accumulator complete
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------------ epilogue
    const int q = warp & 3;  // TMEM lane quadrant this warp may read
    uint32_t tcount = 0;
    float scal0 = 1.f;
    if constexpr (kEpi == EPI_SQDIST) scal0 = __ldg(p.scal);
    const bool accumulate_out = p.nsplit > 1;
    for (int w = blockIdx.x; w < num_work; w += gridDim.x, ++tcount) {
      const Work wk = decode(w);
      const int tm = wk.tm, tn = wk.tn;
      const int a = tcount & 1;
      const uint32_t aph = (tcount >> 1) & 1u;
      ptx::mbar_wait(tfull_bar(a), aph);
      ptx::tc_fence_after();
      const int64_t row0 = (int64_t)tm * BM + q * 32;
      const int64_t row = row0 + lane;
      const bool row_ok = row < p.M;
      float rv = 0.f;
      if constexpr (kEpi == EPI_SQDIST) rv = row_ok ? __ldg(p.row_vec + row) : 0.f;
      float* xp = reinterpret_cast<float*>(smem + kStages * kStageBytes + kBarBytes) + q * (32 * kXposePitch);
#pragma unroll 1
      for (int c = 0; c < BN / 32; ++c) {
        const int64_t col0 = (int64_t)tn * BN + c * 32;
        if (c * 32 >= wk.ncols || col0 >= p.N) break;  // warp-uniform
        uint32_t v[32];
        ptx::tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + a * BN + c * 32, v);
        const bool col_ok = col0 + lane < p.N;
        const float cvl = col_ok ? __ldg(p.col_vec + col0 + lane) : 0.f;
        ptx::tmem_ld_wait();
        // a thread owns one row in TMEM; transpose 32 x 32 through shared memory so that each
        // store / red instruction covers 32 consecutive columns of one row (one 128-byte line)
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          const float acc = __uint_as_float(v[e]);
          if constexpr (kEpi == EPI_SQDIST) {
            const float cv = __shfl_sync(0xffffffffu, cvl, e);
            xp[lane * kXposePitch + e] = fmaxf(fmaf(-2.f, acc, rv + cv) * scal0, 0.f);
          } else {
            xp[lane * kXposePitch + e] = acc;
          }
        }
        __syncwarp();
#pragma unroll 8
        for (int r = 0; r < 32; ++r) {
          if (row0 + r < p.M && col_ok) {
            float val = xp[r * kXposePitch + lane];
            if constexpr (kEpi == EPI_COLSCALE) val *= cvl;
            float* dst = p.out + (row0 + r) * p.ldo + col0 + lane;
            if (accumulate_out) atomicAdd(dst, val);
            else *dst = val;
          }
        }
        __syncwarp();
      }
      // all TMEM reads of this accumulator stage have completed (wait::ld above)
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(tempty_bar(a));
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, kTmemCols);
  }
}

// ---------------------------------------------------------------------------------------------
// host side: TMA descriptors + launch
// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = []() -> EncodeTiledFn {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      return nullptr;
    return reinterpret_cast<EncodeTiledFn>(p);
  }();
  return fn;
}

}  // namespace

// row-major matrix [outer x inner] of 2-byte (fp16) or 4-byte (fp32) elements with row pitch ld
// (elements); box = [box_outer x box_inner], 128-byte swizzle (box_inner * elem_bytes == 128)
static int make_tmap_any(CUtensorMap* tm, const void* base, int elem_bytes, int64_t inner,
                         int64_t outer, int64_t ld, uint32_t box_inner, uint32_t box_outer) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from this driver");
    return SIGSVGD_ERR_CUDA;
  }
  if ((reinterpret_cast<uintptr_t>(base) & 15u) || ((ld * elem_bytes) & 15)) {
    set_error("TMA operands must be 16-byte aligned with a row pitch that is a multiple of 16 bytes");
    return SIGSVGD_ERR_INVALID;
  }
  cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t strides[1] = {(cuuint64_t)ld * (cuuint64_t)elem_bytes};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult rc = fn(tm, elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32,
                   2, const_cast<void*>(base), dims, strides,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (rc != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d) inner=%lld outer=%lld ld=%lld", (int)rc,
              (long long)inner, (long long)outer, (long long)ld);
    return SIGSVGD_ERR_CUDA;
  }
  return 0;
}

int make_tmap_f16(CUtensorMap* tm, const __half* base, int64_t inner, int64_t outer, int64_t ld,
                  uint32_t box_inner, uint32_t box_outer) {
  return make_tmap_any(tm, base, 2, inner, outer, ld, box_inner, box_outer);
}
int make_tmap_f32(CUtensorMap* tm, const float* base, int64_t inner, int64_t outer, int64_t ld,
                  uint32_t box_inner, uint32_t box_outer) {
  return make_tmap_any(tm, base, 4, inner, outer, ld, box_inner, box_outer);
}

namespace {

int make_tmap(CUtensorMap* tm, const __half* base, int64_t inner, int64_t outer, int64_t ld,
              uint32_t box_inner, uint32_t box_outer) {
  return make_tmap_f16(tm, base, inner, outer, ld, box_inner, box_outer);
}

template <bool kBMN, int kEpi>
int launch_gemm3x(const CUtensorMap& ah, const CUtensorMap& al, const CUtensorMap& bh,
                  const CUtensorMap& bl, const GemmParams& p, cudaStream_t st) {
  static thread_local int configured_dev = -1;
  int dev = 0;
  SIG_CUDA_OK(cudaGetDevice(&dev));
  if (configured_dev != dev) {
    SIG_CUDA_OK(cudaFuncSetAttribute(gemm3x_kernel<kBMN, kEpi>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
    configured_dev = dev;
  }
  const int grid = std::min(p.tiles_m * p.tiles_n * p.nsplit, num_sms());
  gemm3x_kernel<kBMN, kEpi><<<grid, kGemmThreads, kSmemBytes, st>>>(ah, al, bh, bl, p);
  SIG_LAUNCH_OK("gemm3x_kernel");
  return 0;
}

}  // namespace

bool tc_available() { return encode_fn() != nullptr; }

int tc_sqdist(TcPlanes A, TcPlanes B, int64_t M, int64_t N, int64_t D, const float* na,
              const float* nb, const float* scal, float* dist, int64_t ldd, cudaStream_t st) {
  SIG_CHECK_ARG(M >= 1 && N >= 1 && D >= 1 && A.ld >= D && B.ld >= D && ldd >= N,
                "tc_sqdist: bad shape M=%lld N=%lld D=%lld", (long long)M, (long long)N, (long long)D);
  SIG_CHECK_ARG(M < (1ll << 31) && N < (1ll << 31), "tc_sqdist: too many rows");
  CUtensorMap ah, al, bh, bl;
  SIG_PROPAGATE(make_tmap(&ah, A.hi, D, M, A.ld, BK, BM));
  SIG_PROPAGATE(make_tmap(&al, A.lo, D, M, A.ld, BK, BM));
  SIG_PROPAGATE(make_tmap(&bh, B.hi, D, N, B.ld, BK, BN));
  SIG_PROPAGATE(make_tmap(&bl, B.lo, D, N, B.ld, BK, BN));
  GemmParams p;
  p.M = M, p.N = N;
  p.num_k16 = (int)ceil_div(D, 16);
  p.tiles_m = (int)ceil_div(M, BM), p.tiles_n = (int)ceil_div(N, BN);
  p.nsplit = 1, p.kb_per_split = (p.num_k16 + 3) / 4;
  p.out = dist, p.ldo = ldd;
  p.row_vec = na, p.col_vec = nb, p.scal = scal;
  return launch_gemm3x<false, EPI_SQDIST>(ah, al, bh, bl, p, st);
}

int tc_matmul_nn(TcPlanes A, TcPlanes V, int64_t M, int64_t K, int64_t C, const float* colscale,
                 float* out, int64_t ldo, cudaStream_t st) {
  SIG_CHECK_ARG(M >= 1 && K >= 1 && C >= 1 && A.ld >= K && V.ld >= C && ldo >= C,
                "tc_matmul_nn: bad shape M=%lld K=%lld C=%lld", (long long)M, (long long)K, (long long)C);
  SIG_CHECK_ARG(M < (1ll << 31) && K < (1ll << 31), "tc_matmul_nn: too many rows");
  CUtensorMap ah, al, bh, bl;
  SIG_PROPAGATE(make_tmap(&ah, A.hi, K, M, A.ld, BK, BM));
  SIG_PROPAGATE(make_tmap(&al, A.lo, K, M, A.ld, BK, BM));
  SIG_PROPAGATE(make_tmap(&bh, V.hi, C, K, V.ld, 64, BK));
  SIG_PROPAGATE(make_tmap(&bl, V.lo, C, K, V.ld, 64, BK));
  GemmParams p;
  p.M = M, p.N = C;
  p.num_k16 = (int)ceil_div(K, 16);
  p.tiles_m = (int)ceil_div(M, BM), p.tiles_n = (int)ceil_div(C, BN);
  // K chunks of at most 16 blocks (1024 contraction indices = 192 chained MMAs): the TMEM
  // accumulator adds with truncation, so the chain length bounds the rounding bias
  const int num_kb = (p.num_k16 + 3) / 4;
  p.kb_per_split = std::min(num_kb, kMaxChainKb);
  p.nsplit = (int)ceil_div(num_kb, p.kb_per_split);
  p.kb_per_split = (int)ceil_div(num_kb, p.nsplit);  // balance the chunks
  p.out = out, p.ldo = ldo;
  p.row_vec = nullptr, p.col_vec = colscale, p.scal = nullptr;
  if (p.nsplit > 1)
    SIG_CUDA_OK(cudaMemset2DAsync(out, sizeof(float) * ldo, 0, sizeof(float) * C, M, st));
  return launch_gemm3x<true, EPI_COLSCALE>(ah, al, bh, bl, p, st);
}

}  // namespace sigsvgd
