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
#include <stdlib.h>

#include "gemm_tc.cuh"
#include "tc_ptx.cuh"

namespace sigsvgd {
namespace {

constexpr int BM = 128, BN = 256;
constexpr int kGemmThreads = 256;
// Pipeline shape per operand layout.  K-major B (tc_sqdist): 64-wide K blocks (one 128-byte
// swizzle atom), 2 stages of 96 KB.  MN-major B (tc_matmul_nn, the dominant kernel): 32-wide K
// blocks, 4 stages of 48 KB -- the same 192 KB, but a stage's TMA round trip now has three MMA
// stage times to complete in instead of one (with 2 x 96 KB the tensor pipe idled ~45 % waiting
// for the refill; profiles/r01e).  A (K-major) then has 64-byte rows: SWIZZLE_64B.
template <bool kBMN>
struct Pipe {
  static constexpr int BK = kBMN ? 32 : 64;
  static constexpr int kStages = kBMN ? 4 : 2;
  static constexpr uint32_t kABytes = BM * BK * 2;  // per plane
  static constexpr uint32_t kBBytes = BN * BK * 2;  // per plane
  static constexpr uint32_t kStageBytes = 2 * kABytes + 2 * kBBytes;  // hi+lo of A and B
  static constexpr int kK16 = BK / 16;               // UMMA K steps per stage
  static constexpr int kMaxChainKb = 1024 / BK;      // K blocks per accumulation chain
};
constexpr uint32_t kPipeBytes = 192 * 1024;
static_assert(Pipe<true>::kStages * Pipe<true>::kStageBytes == kPipeBytes, "");
static_assert(Pipe<false>::kStages * Pipe<false>::kStageBytes == kPipeBytes, "");
constexpr uint32_t kBarBytes = 256;
constexpr int kXposePitch = 33;                               // 32 x 32 transpose tile per epilogue warp
constexpr uint32_t kXposeBytes = 4 * 32 * kXposePitch * 4;
constexpr uint32_t kSmemBytes = kPipeBytes + kBarBytes + kXposeBytes + 1024;  // + alignment slack
constexpr uint32_t kTmemCols = 512;

enum { EPI_SQDIST = 0, EPI_COLSCALE = 1 };

struct GemmParams {
  int64_t M, N;        // output rows / columns
  int num_k16;         // ceil(Kdim / 16): number of UMMA K steps
  int tiles_m, tiles_n;
  int nsplit;          // K is cut into nsplit chunks of kb_per_split BK-wide blocks; chunks of one
  int dbg;             // experiments: 1 = scalar red, 2 = no epilogue stores
  int streamk;         // 1: equal-time spans of the (tile, K block) sequence (see the kernel)
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
  using P = Pipe<kBMN>;
  constexpr int BK = P::BK, kStages = P::kStages, kK16 = P::kK16;
  constexpr uint32_t kABytes = P::kABytes, kBBytes = P::kBBytes, kStageBytes = P::kStageBytes;
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

  // Work decomposition.  Static (p.streamk == 0): item w -> (column tile, row tile, K chunk), dealt
  // round-robin.  Stream-K (p.streamk == 1, the K.F contraction): the (tile, K block) sequence --
  // column tile slowest, each K block weighted by the tile's MMA width -- is cut into gridDim.x
  // spans of equal tensor time; a CTA walks its span as chains of <= p.kb_per_split K blocks that
  // never cross a tile.  With 512 static items of unequal width the busiest CTAs carried 3.3
  // chains against a mean of 2.8 (profiles/r01e: 55 % tensor-pipe activity).
  const int num_work = p.tiles_m * p.tiles_n * p.nsplit;
  const int num_kb_total = (p.num_k16 + kK16 - 1) / kK16;
  struct Work {
    int tm, tn, kb0, kb1, ncols;
  };
  struct WorkIter {
    int w;                     // static: next item
    int t, kb, t_end, kb_end;  // stream-K: position and end of the span (tile, K block)
  };
  const int num_tiles = p.tiles_m * p.tiles_n;
  const int w_last = (int)((min((int64_t)BN, p.N - (int64_t)(p.tiles_n - 1) * BN) + 15) / 16);
  const long long u_full_tile = (long long)num_kb_total * (BN / 16);
  const long long u_last_tile = (long long)num_kb_total * w_last;
  const long long u_full = (long long)p.tiles_m * (p.tiles_n - 1) * u_full_tile;
  auto locate = [&](long long u, int& t, int& kb) {
    if (u < u_full) {
      t = (int)(u / u_full_tile);
      kb = (int)((u - (long long)t * u_full_tile) / (BN / 16));
    } else {
      const long long v = u - u_full;
      const int tl = (int)(v / u_last_tile);
      t = p.tiles_m * (p.tiles_n - 1) + tl;
      kb = (int)((v - (long long)tl * u_last_tile) / w_last);
    }
  };
  auto iter_init = [&]() {
    WorkIter it;
    it.w = blockIdx.x;
    it.t = it.kb = it.t_end = it.kb_end = 0;
    if (p.streamk) {
      const long long U = u_full + (long long)p.tiles_m * u_last_tile;
      locate(U * blockIdx.x / gridDim.x, it.t, it.kb);
      if (blockIdx.x + 1 == gridDim.x) it.t_end = num_tiles, it.kb_end = 0;
      else locate(U * (blockIdx.x + 1) / gridDim.x, it.t_end, it.kb_end);
    }
    return it;
  };
  auto iter_next = [&](WorkIter& it, Work& k) {
    if (!p.streamk) {
      // Items are dealt round-robin in (column tile, row tile, K chunk) order.  CTAs that run
      // side by side then walk the SAME K blocks of the same column tile in phase, so their
      // requests for the shared B lines coincide and L2 serves them once.  Two schedules with
      // better-balanced tensor time -- stream-K spans (157 us) and dealing the narrow last
      // column tile to the lighter CTAs (105 us) -- both lost to this one (93 us at N = 4096)
      // because they break that phase alignment.
      if (it.w >= num_work) return false;
      const int w = it.w;
      it.w += gridDim.x;
      // row tile slowest, then column tile, then K chunk: the CTAs running side by side share
      // A row tiles as well as B column tiles (2 % faster than column-tile-slowest)
      k.tm = w / (p.tiles_n * p.nsplit);
      const int rem = w - k.tm * (p.tiles_n * p.nsplit);
      k.tn = rem / p.nsplit;
      const int sp = rem - k.tn * p.nsplit;
      k.kb0 = sp * p.kb_per_split;
      k.kb1 = min(num_kb_total, k.kb0 + p.kb_per_split);
    } else {
      if (it.t > it.t_end || (it.t == it.t_end && it.kb >= it.kb_end)) return false;
      k.tn = it.t / p.tiles_m;
      k.tm = it.t - k.tn * p.tiles_m;
      k.kb0 = it.kb;
      k.kb1 = min(num_kb_total, it.kb + p.kb_per_split);
      if (it.t == it.t_end) k.kb1 = min(k.kb1, it.kb_end);
      it.kb = k.kb1;
      if (it.kb >= num_kb_total) it.t += 1, it.kb = 0;
    }
    // MMA N of this tile: multiple of 16, at most BN
    k.ncols = (int)min((int64_t)BN, ((p.N - (int64_t)k.tn * BN) + 15) / 16 * 16);
    return true;
  };

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      uint32_t it = 0;
      Work wk;
      for (WorkIter wi = iter_init(); iter_next(wi, wk);) {
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
      Work wk;
      for (WorkIter wi = iter_init(); iter_next(wi, wk); ++tcount) {
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
          const int steps = min(kK16, p.num_k16 - kb * kK16);
          for (int k = 0; k < steps; ++k) {
            // K-major A, rows of BK fp16 = one swizzle span (128 B or 64 B): 8-row groups
            // 8 * span bytes apart, 16 K-elements = 32 bytes
            constexpr uint32_t kASbo = 8 * BK * 2;
            constexpr uint32_t kALayout = BK == 64 ? ptx::kLayoutSw128 : ptx::kLayoutSw64;
            const uint64_t ah = ptx::make_smem_desc(st + k * 32, 16, kASbo, kALayout);
            const uint64_t al = ptx::make_smem_desc(st + kABytes + k * 32, 16, kASbo, kALayout);
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
    const bool accumulate_out = p.nsplit > 1 || p.streamk;
    // 16-byte reductions (red.global.add.v4.f32): a quarter of the L2 atomic operations
    const bool vec_red = accumulate_out && !(p.dbg & 1) && (p.ldo & 3) == 0 && (p.N & 3) == 0 &&
                         (reinterpret_cast<uintptr_t>(p.out) & 15u) == 0;
    Work wk;
    for (WorkIter wi = iter_init(); iter_next(wi, wk); ++tcount) {
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
        if (p.dbg & 2) {
        } else if (vec_red) {
          // lane -> (row r0 + lane / 8, columns 4 * (lane % 8) ..+3): 4 rows x 128 bytes per instruction
          const int cg = (lane & 7) * 4;
          const float s0 = __shfl_sync(0xffffffffu, cvl, cg), s1 = __shfl_sync(0xffffffffu, cvl, cg + 1);
          const float s2 = __shfl_sync(0xffffffffu, cvl, cg + 2), s3 = __shfl_sync(0xffffffffu, cvl, cg + 3);
          if (col0 + cg < p.N) {
#pragma unroll
            for (int rr = 0; rr < 32; rr += 4) {
              const int r = rr + (lane >> 3);
              if (row0 + r < p.M) {
                const float* src = xp + r * kXposePitch + cg;
                ptx::red_add_v4(p.out + (row0 + r) * p.ldo + col0 + cg, src[0] * s0, src[1] * s1,
                                src[2] * s2, src[3] * s3);
              }
            }
          }
        } else
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
  // swizzle span = bytes of one box row (128 or 64)
  const CUtensorMapSwizzle swz =
      box_inner * elem_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B;
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
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
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
  const int grid = (int)std::min<int64_t>((int64_t)p.tiles_m * p.tiles_n * p.nsplit, num_sms());
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
  constexpr int BK = Pipe<false>::BK;
  CUtensorMap ah, al, bh, bl;
  SIG_PROPAGATE(make_tmap(&ah, A.hi, D, M, A.ld, BK, BM));
  SIG_PROPAGATE(make_tmap(&al, A.lo, D, M, A.ld, BK, BM));
  SIG_PROPAGATE(make_tmap(&bh, B.hi, D, N, B.ld, BK, BN));
  SIG_PROPAGATE(make_tmap(&bl, B.lo, D, N, B.ld, BK, BN));
  GemmParams p;
  p.M = M, p.N = N;
  p.num_k16 = (int)ceil_div(D, 16);
  p.tiles_m = (int)ceil_div(M, BM), p.tiles_n = (int)ceil_div(N, BN);
  p.dbg = 0;
  p.nsplit = 1, p.streamk = 0, p.kb_per_split = (p.num_k16 + Pipe<false>::kK16 - 1) / Pipe<false>::kK16;
  p.out = dist, p.ldo = ldd;
  p.row_vec = na, p.col_vec = nb, p.scal = scal;
  return launch_gemm3x<false, EPI_SQDIST>(ah, al, bh, bl, p, st);
}

int tc_matmul_nn(TcPlanes A, TcPlanes V, int64_t M, int64_t K, int64_t C, const float* colscale,
                 float* out, int64_t ldo, cudaStream_t st, bool out_zeroed) {
  SIG_CHECK_ARG(M >= 1 && K >= 1 && C >= 1 && A.ld >= K && V.ld >= C && ldo >= C,
                "tc_matmul_nn: bad shape M=%lld K=%lld C=%lld", (long long)M, (long long)K, (long long)C);
  SIG_CHECK_ARG(M < (1ll << 31) && K < (1ll << 31), "tc_matmul_nn: too many rows");
  // CTA pairs (kf2_tc.cu) when the device can co-schedule them and there are 256-row tiles to fill
  if (M >= 256 && tc_matmul_pair_available())
    return tc_matmul_pair(A, V, M, K, C, colscale, out, ldo, st, out_zeroed);
  using P = Pipe<true>;
  constexpr int BK = P::BK;
  CUtensorMap ah, al, bh, bl;
  SIG_PROPAGATE(make_tmap(&ah, A.hi, K, M, A.ld, BK, BM));  // 64-byte rows: SWIZZLE_64B
  SIG_PROPAGATE(make_tmap(&al, A.lo, K, M, A.ld, BK, BM));
  SIG_PROPAGATE(make_tmap(&bh, V.hi, C, K, V.ld, 64, BK));
  SIG_PROPAGATE(make_tmap(&bl, V.lo, C, K, V.ld, 64, BK));
  GemmParams p;
  p.M = M, p.N = C;
  p.num_k16 = (int)ceil_div(K, 16);
  p.tiles_m = (int)ceil_div(M, BM), p.tiles_n = (int)ceil_div(C, BN);
  // K chunks of at most 1024 contraction indices (192 chained MMAs): the TMEM accumulator adds
  // with truncation, so the chain length bounds the rounding bias
  const int num_kb = (p.num_k16 + P::kK16 - 1) / P::kK16;
  p.kb_per_split = std::min(num_kb, P::kMaxChainKb);
  p.nsplit = (int)ceil_div(num_kb, p.kb_per_split);
  p.kb_per_split = (int)ceil_div(num_kb, p.nsplit);  // balance the chunks
  // stream-K measured 1.7x SLOWER than the static deal (157 vs 93 us at N = 4096): CTAs that walk
  // the K range out of phase never request the same operand lines at the same time, and the
  // L2 -> SM path then has to deliver every byte separately.  Kept as an experiment switch.
  p.streamk = getenv("SIGSVGD_KF_STREAMK") ? 1 : 0;
  p.dbg = getenv("SIGSVGD_KF_DBG") ? atoi(getenv("SIGSVGD_KF_DBG")) : 0;
  if (p.streamk) p.kb_per_split = P::kMaxChainKb;
  p.out = out, p.ldo = ldo;
  p.row_vec = nullptr, p.col_vec = colscale, p.scal = nullptr;
  if ((p.nsplit > 1 || p.streamk) && !out_zeroed)
    SIG_CUDA_OK(cudaMemset2DAsync(out, sizeof(float) * ldo, 0, sizeof(float) * C, M, st));
  return launch_gemm3x<true, EPI_COLSCALE>(ah, al, bh, bl, p, st);
}

}  // namespace sigsvgd
