// K . [Sc | 1 | score] on CTA PAIRS (tcgen05 cta_group::2): src/inference/svgd.py:82 and the
// K.S / row-sum terms of autograd.grad(k_xx.sum(), X), svgd.py:43.
//
// Same arithmetic as gemm3x_kernel<MN-major B> (gemm_tc.cu): split-fp16 operands, 3 MMAs per fp32
// product, accumulation chains of <= 1024 contraction indices combined with 16-byte reductions.
// What changes is who fetches what: two CTAs of a cluster (one TPC) own a 256 x 256 output tile;
// each CTA stages ITS 128 rows of A and ITS half of B's columns, and one tcgen05.mma issued by the
// even CTA runs on both SMs' tensor cores reading both shared memories.  Per 32-wide K block a CTA
// receives 32 KB instead of 48 KB (the single-CTA kernel was bound by operand delivery L2 -> SM,
// DESIGN.md 5), there is room for 6 stages, and each CTA keeps two 128 x 256 accumulators in its
// TMEM so the epilogue of one chain still overlaps the MMAs of the next.
//
// CTA = 256 threads: warp 0 TMA producer (both CTAs), warp 1 MMA issuer (even CTA only), warp 2
// TMEM allocator, warps 4-7 epilogue (each CTA drains its own 128 rows).
#include <cuda.h>
#include <stdlib.h>

#include <algorithm>

#include "gemm_tc.cuh"
#include "tc_ptx.cuh"

namespace sigsvgd {
namespace {

constexpr int PM = 128;   // rows per CTA (256 per pair)
constexpr int PN = 256;   // columns per pair tile (each CTA stages half)
constexpr int PK = 32;    // K block
constexpr int kPStages = 6;
constexpr int kPThreads = 256;
constexpr uint32_t kPABytes = PM * PK * 2;        // 8 KB per plane (64-byte rows, SWIZZLE_64B)
constexpr uint32_t kPBBytes = (PN / 2) * PK * 2;  // 8 KB per plane (two [64 col x 32 row] boxes)
constexpr uint32_t kPStageBytes = 2 * kPABytes + 2 * kPBBytes;  // 32 KB
constexpr uint32_t kPBarBytes = 256;
constexpr int kPPitch = 33;
constexpr uint32_t kPXposeBytes = 4 * 32 * kPPitch * 4;
constexpr uint32_t kPSmemBytes = kPStages * kPStageBytes + kPBarBytes + kPXposeBytes + 1024;
constexpr int kPMaxChainKb = 1024 / PK;

struct PairParams {
  int64_t M, N;  // output rows, columns
  int num_k16;
  int tiles_m, tiles_n;  // tiles of 256 x tile_n
  int tile_n;            // columns per pair tile (<= PN, multiple of 16)
  int nsplit, kb_per_split;
  int order;  // 0: row tile slowest, 1: column tile slowest (wide tiles first, the narrow one last)
  float* out;
  int64_t ldo;
  const float* colscale;
};

__global__ void __launch_bounds__(kPThreads, 1)
kf_pair_kernel(const __grid_constant__ CUtensorMap tmAh, const __grid_constant__ CUtensorMap tmAl,
               const __grid_constant__ CUtensorMap tmBh, const __grid_constant__ CUtensorMap tmBl,
               const PairParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t sbase = (raw + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (sbase - raw);
  const uint32_t bar0 = sbase + kPStages * kPStageBytes;
  auto full_bar = [&](int s) { return bar0 + 8u * s; };                  // used in the even CTA
  auto empty_bar = [&](int s) { return bar0 + 8u * (kPStages + s); };    // both CTAs (multicast commit)
  auto tfull_bar = [&](int a) { return bar0 + 8u * (2 * kPStages + a); };       // both CTAs
  auto tempty_bar = [&](int a) { return bar0 + 8u * (2 * kPStages + 2 + a); };  // even CTA, 8 arrivals
  volatile uint32_t* tmem_slot =
      reinterpret_cast<volatile uint32_t*>(smem + kPStages * kPStageBytes + 8 * (2 * kPStages + 4));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = ptx::cluster_ctarank();
  const bool leader = rank == 0;
  const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;
  if (warp == 0 && lane == 0) {
    ptx::tma_prefetch_desc(&tmAh);
    ptx::tma_prefetch_desc(&tmAl);
    ptx::tma_prefetch_desc(&tmBh);
    ptx::tma_prefetch_desc(&tmBl);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kPStages; ++s) {
      ptx::mbar_init(full_bar(s), 1);
      ptx::mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      ptx::mbar_init(tfull_bar(a), 1);
      ptx::mbar_init(tempty_bar(a), 8);  // 4 epilogue warps of each CTA
    }
    ptx::fence_mbar_init();
  }
  if (warp == 2) ptx::tmem_alloc_pair(ptx::smem_u32((const void*)tmem_slot), 512);
  ptx::tc_fence_before();
  ptx::cluster_sync();  // barrier inits and the allocation are visible to the peer
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // pair-level work item w -> (row tile, column tile, K chunk), row tile slowest; dealt round-robin
  // over the pairs (see gemm_tc.cu for why the in-phase round-robin deal wins)
  const int num_work = p.tiles_m * p.tiles_n * p.nsplit;
  const int num_kb_total = (p.num_k16 + 1) / 2;
  struct Work {
    int tm, tn, kb0, kb1, ncols, nhalf;
  };
  auto decode = [&](int w) {
    Work k;
    int sp;
    if (p.order == 0) {
      k.tm = w / (p.tiles_n * p.nsplit);
      const int rem = w - k.tm * (p.tiles_n * p.nsplit);
      k.tn = rem / p.nsplit;
      sp = rem - k.tn * p.nsplit;
    } else {
      k.tn = w / (p.tiles_m * p.nsplit);
      const int rem = w - k.tn * (p.tiles_m * p.nsplit);
      k.tm = rem / p.nsplit;
      sp = rem - k.tm * p.nsplit;
    }
    k.kb0 = sp * p.kb_per_split;
    k.kb1 = min(num_kb_total, k.kb0 + p.kb_per_split);
    k.ncols = (int)min((int64_t)p.tile_n, ((p.N - (int64_t)k.tn * p.tile_n) + 15) / 16 * 16);
    k.nhalf = k.ncols / 2;  // columns of B staged by each CTA (multiple of 8)
    return k;
  };

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (both CTAs)
    if (lane == 0) {
      uint32_t it = 0;
      for (int w = pair; w < num_work; w += npairs) {
        const Work wk = decode(w);
        const int nbox = (wk.nhalf + 63) / 64;
        const int row = wk.tm * (2 * PM) + (int)rank * PM;
        const int col = wk.tn * p.tile_n + (int)rank * wk.nhalf;
        for (int kb = wk.kb0; kb < wk.kb1; ++kb, ++it) {
          const int s = it % kPStages;
          const uint32_t ph = (it / kPStages) & 1u;
          ptx::mbar_wait(empty_bar(s), ph ^ 1u);
          const uint32_t st = sbase + s * kPStageBytes;
          // the even CTA's barrier counts the bytes of BOTH CTAs
          if (leader) ptx::mbar_expect_tx(full_bar(s), 2 * (2 * kPABytes + 2 * nbox * (PK * 128)));
          ptx::tma_load_2d_pair(st, &tmAh, full_bar(s), kb * PK, row);
          ptx::tma_load_2d_pair(st + kPABytes, &tmAl, full_bar(s), kb * PK, row);
#pragma unroll
          for (int q = 0; q < PN / 128; ++q) {
            if (q >= nbox) break;
            ptx::tma_load_2d_pair(st + 2 * kPABytes + q * (PK * 128), &tmBh, full_bar(s), col + q * 64,
                                  kb * PK);
            ptx::tma_load_2d_pair(st + 2 * kPABytes + kPBBytes + q * (PK * 128), &tmBl, full_bar(s),
                                  col + q * 64, kb * PK);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (even CTA)
    if (leader && lane == 0) {
      uint32_t it = 0, tcount = 0;
      for (int w = pair; w < num_work; w += npairs, ++tcount) {
        const Work wk = decode(w);
        const uint32_t idesc = ptx::make_idesc_f16(2 * PM, wk.ncols, false, true);
        const int a = tcount & 1;
        const uint32_t aph = (tcount >> 1) & 1u;
        ptx::mbar_wait(tempty_bar(a), aph ^ 1u);  // both CTAs drained this accumulator stage
        ptx::tc_fence_after();
        const uint32_t tacc = tmem_base + a * PN;
        uint32_t accumulate = 0;
        for (int kb = wk.kb0; kb < wk.kb1; ++kb, ++it) {
          const int s = it % kPStages;
          const uint32_t ph = (it / kPStages) & 1u;
          ptx::mbar_wait(full_bar(s), ph);
          ptx::tc_fence_after();
          const uint32_t st = sbase + s * kPStageBytes;
          const int steps = min(2, p.num_k16 - kb * 2);
          for (int k = 0; k < steps; ++k) {
            const uint64_t ah = ptx::make_smem_desc(st + k * 32, 16, 512, ptx::kLayoutSw64);
            const uint64_t al = ptx::make_smem_desc(st + kPABytes + k * 32, 16, 512, ptx::kLayoutSw64);
            const uint64_t bh =
                ptx::make_smem_desc(st + 2 * kPABytes + k * 2048, PK * 128, 1024, ptx::kLayoutSw128);
            const uint64_t bl = ptx::make_smem_desc(st + 2 * kPABytes + kPBBytes + k * 2048, PK * 128,
                                                    1024, ptx::kLayoutSw128);
            ptx::umma_f16_pair(tacc, ah, bh, idesc, accumulate);
            ptx::umma_f16_pair(tacc, ah, bl, idesc, 1u);
            ptx::umma_f16_pair(tacc, al, bh, idesc, 1u);
            accumulate = 1u;
          }
          ptx::umma_commit_pair(empty_bar(s));  // frees the stage in both CTAs
        }
        ptx::umma_commit_pair(tfull_bar(a));  // accumulators of both CTAs complete
      }
    }
  } else if (warp >= 4) {
    // ------------------------------------------------------------------ epilogue (own 128 rows)
    const int q = warp & 3;
    uint32_t tcount = 0;
    const bool accumulate_out = p.nsplit > 1;
    const bool vec_red = accumulate_out && (p.ldo & 3) == 0 && (p.N & 3) == 0 &&
                         (reinterpret_cast<uintptr_t>(p.out) & 15u) == 0;
    float* xp = reinterpret_cast<float*>(smem + kPStages * kPStageBytes + kPBarBytes) + q * (32 * kPPitch);
    for (int w = pair; w < num_work; w += npairs, ++tcount) {
      const Work wk = decode(w);
      const int a = tcount & 1;
      const uint32_t aph = (tcount >> 1) & 1u;
      ptx::mbar_wait(tfull_bar(a), aph);
      ptx::tc_fence_after();
      const int64_t row0 = (int64_t)wk.tm * (2 * PM) + (int64_t)rank * PM + q * 32;
#pragma unroll 1
      for (int c = 0; c < PN / 32; ++c) {
        const int64_t col0 = (int64_t)wk.tn * p.tile_n + c * 32;
        if (c * 32 >= wk.ncols || col0 >= p.N) break;  // warp-uniform
        uint32_t v[32];
        ptx::tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + a * PN + c * 32, v);
        const bool col_ok = col0 + lane < p.N;
        const float cvl = col_ok ? __ldg(p.colscale + col0 + lane) : 0.f;
        ptx::tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 32; ++e) xp[lane * kPPitch + e] = __uint_as_float(v[e]);
        __syncwarp();
        if (vec_red) {
          const int cg = (lane & 7) * 4;
          const float s0 = __shfl_sync(0xffffffffu, cvl, cg), s1 = __shfl_sync(0xffffffffu, cvl, cg + 1);
          const float s2 = __shfl_sync(0xffffffffu, cvl, cg + 2), s3 = __shfl_sync(0xffffffffu, cvl, cg + 3);
          if (col0 + cg < p.N) {
#pragma unroll
            for (int rr = 0; rr < 32; rr += 4) {
              const int r = rr + (lane >> 3);
              if (row0 + r < p.M) {
                const float* src = xp + r * kPPitch + cg;
                ptx::red_add_v4(p.out + (row0 + r) * p.ldo + col0 + cg, src[0] * s0, src[1] * s1,
                                src[2] * s2, src[3] * s3);
              }
            }
          }
        } else {
#pragma unroll 8
          for (int r = 0; r < 32; ++r) {
            if (row0 + r < p.M && col_ok) {
              const float val = xp[r * kPPitch + lane] * cvl;
              float* dst = p.out + (row0 + r) * p.ldo + col0 + lane;
              if (accumulate_out) atomicAdd(dst, val);
              else *dst = val;
            }
          }
        }
        __syncwarp();
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster(ptx::mapa(tempty_bar(a), 0));  // the even CTA's barrier
    }
  }
  ptx::tc_fence_before();
  ptx::cluster_sync();  // nobody frees TMEM / exits while the peer's MMAs or arrivals are in flight
  if (warp == 2) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc_pair(tmem_base, 512);
  }
}

}  // namespace

// defined in gemm_tc.cu
int make_tmap_f16(CUtensorMap* tm, const __half* base, int64_t inner, int64_t outer, int64_t ld,
                  uint32_t box_inner, uint32_t box_outer);

// Pairs that can be co-resident (one 2-CTA cluster per TPC); 0 if cluster launch is unavailable.
static int max_pairs() {
  static thread_local int cached_dev = -1, cached = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (dev == cached_dev) return cached;
  cached_dev = dev, cached = 0;
  if (cudaFuncSetAttribute(kf_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)kPSmemBytes) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * (unsigned)num_sms());
  cfg.blockDim = dim3(kPThreads);
  cfg.dynamicSmemBytes = kPSmemBytes;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2, attr[0].val.clusterDim.y = 1, attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr, cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, kf_pair_kernel, &cfg) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  cached = std::max(0, n);
  return cached;
}

bool tc_matmul_pair_available() { return max_pairs() > 0 && !getenv("SIGSVGD_KF_NO_PAIRS"); }

int tc_matmul_pair(TcPlanes A, TcPlanes V, int64_t M, int64_t K, int64_t C, const float* colscale,
                   float* out, int64_t ldo, cudaStream_t st, bool out_zeroed) {
  const int pairs_max = max_pairs();
  if (pairs_max <= 0) {
    set_error("tc_matmul_pair: 2-CTA clusters cannot be launched on this device");
    return SIGSVGD_ERR_UNSUPPORTED;
  }
  CUtensorMap ah, al, bh, bl;
  SIG_PROPAGATE(make_tmap_f16(&ah, A.hi, K, M, A.ld, PK, PM));  // 64-byte rows: SWIZZLE_64B
  SIG_PROPAGATE(make_tmap_f16(&al, A.lo, K, M, A.ld, PK, PM));
  SIG_PROPAGATE(make_tmap_f16(&bh, V.hi, C, K, V.ld, 64, PK));
  SIG_PROPAGATE(make_tmap_f16(&bl, V.lo, C, K, V.ld, 64, PK));
  PairParams p;
  p.M = M, p.N = C;
  p.num_k16 = (int)ceil_div(K, 16);
  p.tiles_m = (int)ceil_div(M, 2 * PM), p.tiles_n = (int)ceil_div(C, PN);
  p.tile_n = PN;
  {
    // experiment knob: column tiles of equal width (848 = 3 x 224 + 176 instead of 3 x 256 + 80)
    static const int equal_tiles = [] {
      const char* e = getenv("SIGSVGD_KF_EQUAL_TILES");
      return e ? atoi(e) : 0;
    }();
    if (equal_tiles) p.tile_n = (int)std::min<int64_t>(PN, round_up(ceil_div(C, p.tiles_n), 32));
  }
  const int num_kb = (p.num_k16 + 1) / 2;
  p.kb_per_split = std::min(num_kb, kPMaxChainKb);
  p.nsplit = (int)ceil_div(num_kb, p.kb_per_split);
  // shorter chains while there are fewer than ~6 items per pair (N = 4096: 8 chunks of 512 indices,
  // 79 us against 81-85 us with 4-6 chunks); never below 8 K blocks per chain
  while ((int64_t)p.tiles_m * p.tiles_n * p.nsplit < 6ll * pairs_max && num_kb / (2 * p.nsplit) >= 8)
    p.nsplit *= 2;
  p.kb_per_split = (int)ceil_div(num_kb, p.nsplit);
  p.nsplit = (int)ceil_div(num_kb, p.kb_per_split);
  p.order = 0;  // row tile slowest (column-tile-slowest measured 12 % slower at N = 8192)
  p.out = out, p.ldo = ldo, p.colscale = colscale;
  if (p.nsplit > 1 && !out_zeroed)
    SIG_CUDA_OK(cudaMemset2DAsync(out, sizeof(float) * ldo, 0, sizeof(float) * C, M, st));
  const int64_t items = (int64_t)p.tiles_m * p.tiles_n * p.nsplit;
  const int pairs = (int)std::min<int64_t>(items, pairs_max);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * (unsigned)pairs);
  cfg.blockDim = dim3(kPThreads);
  cfg.dynamicSmemBytes = kPSmemBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2, attr[0].val.clusterDim.y = 1, attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr, cfg.numAttrs = 1;
  SIG_CUDA_OK(cudaLaunchKernelEx(&cfg, kf_pair_kernel, ah, al, bh, bl, p));
  SIG_LAUNCH_OK("kf_pair_kernel");
  return 0;
}

}  // namespace sigsvgd
