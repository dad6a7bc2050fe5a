// out = f(A) . [Sc | 1 | score]  with the RBF map fused into the A-operand path (K5).
// Reference: K = exp(-0.5/h^2 sq) (src/kernels/_kernels.py:110), k_xx @ score (src/inference/svgd.py:82)
// and the K.S / row-sum terms of autograd.grad(k_xx.sum(), X) (svgd.py:43).
//
// The kernel matrix is never written as fp16 planes (and, when the caller does not ask for k_xx,
// never written at all): the A operand arrives as the fp32 SQUARED DISTANCES (or an fp32 kernel
// matrix), one [128 rows x 32 columns] TMA box per stage and CTA, and eight converter warps (two
// groups of four taking alternate stages) turn that box IN PLACE into the split-fp16 operand the
// tensor core reads:
//
//   fp32 row of 32 values = 128 bytes  ->  [hi k0..15 | lo k0..15 | hi k16..31 | lo k16..31] fp16
//
// i.e. the same 128 bytes, K-major SWIZZLE_128B, so that the UMMA descriptor of (plane, k16 step)
// is "row base + 32 bytes x {0: hi step 0, 1: lo step 0, 2: hi step 1, 3: lo step 1}".  A thread owns
// a row and converts it as two halves (64 bytes in, the same 64 bytes out), so the conversion needs
// no extra shared memory and no block-level synchronisation.  While they hold the values the converters of the first
// column tile also accumulate sum K*d and record the entries equal to the median (the support of
// torch.median's backward, src/utils/math.py:31).
//
// Everything else follows kf_pair (round 1): split-fp16 B operand [Sc | 1 | score] as MN-major planes
// staged by TMA, 3 MMAs per fp32 product, chains of <= 1024 contraction indices combined with 16-byte
// reductions, CTA pairs (tcgen05 cta_group::2, 256 x 256 tile per SM pair) with two accumulator
// stages so the epilogue of one chain overlaps the MMAs of the next.  kPair = false is the
// single-CTA form (128-row tiles) for strips shorter than 256 rows and for batches of small scenes.
//
// CTA = 512 threads: warps 0-7 converters, warps 8-11 epilogue, warp 12 TMEM allocator, warp 13 TMA
// producer, warp 14 MMA issuer (even CTA of a pair).
#include <cuda.h>

#include <algorithm>

#include "gemm_tc.cuh"
#include "tc_ptx.cuh"

namespace sigsvgd {
namespace {

constexpr int FM = 128;    // rows per CTA
constexpr int FN = 256;    // columns per output tile
constexpr int FK = 32;     // contraction indices per stage
constexpr int kFThreads = 512;
constexpr int kConvWarps = 8;
constexpr int kConvGroups = 2;  // groups of converter warps taking alternate stages
constexpr uint32_t kFABytes = FM * FK * 4;  // 16 KB: fp32 box, converted in place
constexpr uint32_t kFBarBytes = 256;
constexpr int kFPitch = 33;
constexpr uint32_t kFXposeBytes = 4 * 32 * kFPitch * 4;
constexpr uint32_t kFRedBytes = 128;
constexpr int kFMaxChainKb = 1024 / FK;
constexpr float kKScale = 16384.f;  // K in [0,1] -> [0, 2^14] (planes.cu uses the same scaling)

// kWide (pairs only): an item is up to 512 columns wide -- both halves of TMEM as ONE accumulator set
// (columns 0-255 and 256-511, two MMAs per pass sharing the converted A stage), so every distance tile is
// fetched and converted once per 512 output columns instead of once per 256; the price is that the
// epilogue of an item no longer overlaps the MMAs of the next.
template <bool kPair, bool kWide = false>
struct FCfg {
  static_assert(kPair || !kWide, "wide items need CTA pairs");
  static constexpr int kTileCols = kWide ? 2 * FN : FN;          // columns per item (upper bound)
  static constexpr int kStages = kWide ? 4 : (kPair ? 6 : 4);
  static constexpr int kBCols = kPair ? kTileCols / 2 : kTileCols;  // columns of B staged per CTA
  static constexpr uint32_t kBBytes = kBCols * FK * 2;           // per plane
  static constexpr uint32_t kStageBytes = kFABytes + 2 * kBBytes;
  static constexpr uint32_t kSmemBytes = kStages * kStageBytes + kFBarBytes + kFXposeBytes + kFRedBytes + 1024;
  static constexpr int kCtas = kPair ? 2 : 1;
};
static_assert(FCfg<true>::kSmemBytes <= 232448 && FCfg<false>::kSmemBytes <= 232448 &&
                  FCfg<true, true>::kSmemBytes <= 232448,
              "shared memory");

struct FusedParams {
  int64_t M, N;          // output rows, columns
  int64_t row_offset;    // global index of row 0 (tie records)
  int num_k16;
  int tiles_m, tiles_n;
  int tile_cols;         // columns per item (multiple of 32; <= 256, or <= 512 for the wide kernel)
  int nsplit, kb_per_split;
  int mode;              // SIGSVGD_KF_EXP: A = exp(gamma * a) ; SIGSVGD_KF_IDENTITY: A = a
  int stats;             // accumulate state->sum_kd / ties (first column tile only)
  float* out;
  int64_t ldo;
  const float* colscale;
  sigsvgd_bw_state* state;
  // batch of independent problems stacked along the rows of every operand (batch == 1: one problem)
  int batch;
  int64_t a_batch_rows;    // rows of A per problem (= M)
  int64_t v_batch_rows;    // rows of V per problem (= K)
  int64_t colscale_stride; // floats between the colscale vectors of consecutive problems
};

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <bool kPair, bool kWide>
__global__ void __launch_bounds__(kFThreads, 1)
kf_fused_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmBh,
                const __grid_constant__ CUtensorMap tmBl, const FusedParams p) {
  using Cfg = FCfg<kPair, kWide>;
  constexpr int kStages = Cfg::kStages;
  constexpr uint32_t kBBytes = Cfg::kBBytes, kStageBytes = Cfg::kStageBytes;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t sbase = (raw + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (sbase - raw);
  const uint32_t bar0 = sbase + kStages * kStageBytes;
  auto raw_bar = [&](int s) { return bar0 + 8u * s; };                   // own CTA: fp32 box landed
  auto full_bar = [&](int s) { return bar0 + 8u * (kStages + s); };      // leader: B landed + A converted
  auto empty_bar = [&](int s) { return bar0 + 8u * (2 * kStages + s); }; // every CTA (multicast commit)
  auto tfull_bar = [&](int a) { return bar0 + 8u * (3 * kStages + a); };
  auto tempty_bar = [&](int a) { return bar0 + 8u * (3 * kStages + 2 + a); };
  volatile uint32_t* tmem_slot =
      reinterpret_cast<volatile uint32_t*>(smem + kStages * kStageBytes + 8 * (3 * kStages + 4));
  static_assert(8 * (3 * 6 + 4) + 8 <= kFBarBytes, "barrier block");

  // Roles by warp: the scheduler favours high warp ids, so the two single-thread roles whose latency
  // paces everything (MMA issue, TMA issue) sit above the converter warps that would otherwise
  // take their issue slots.
  constexpr int kWarpConv0 = 0, kWarpEpi0 = 8, kWarpAlloc = 12, kWarpTma = 13, kWarpMma = 14;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = kPair ? ptx::cluster_ctarank() : 0u;
  const bool leader = rank == 0;
  const int unit = kPair ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;   // pair (or CTA) index
  const int nunits = kPair ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  if (warp == kWarpTma && lane == 0) {
    ptx::tma_prefetch_desc(&tmA);
    ptx::tma_prefetch_desc(&tmBh);
    ptx::tma_prefetch_desc(&tmBl);
  }
  if (warp == kWarpMma && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      ptx::mbar_init(raw_bar(s), 1);
      ptx::mbar_init(full_bar(s), 1 + (kConvWarps / kConvGroups) * Cfg::kCtas);  // producer + one group's warps
      ptx::mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      ptx::mbar_init(tfull_bar(a), 1);
      ptx::mbar_init(tempty_bar(a), 4 * Cfg::kCtas);
    }
    ptx::fence_mbar_init();
  }
  if (warp == kWarpAlloc) {
    if constexpr (kPair) ptx::tmem_alloc_pair(ptx::smem_u32((const void*)tmem_slot), 512);
    else ptx::tmem_alloc(ptx::smem_u32((const void*)tmem_slot), 512);
  }
  ptx::tc_fence_before();
  if constexpr (kPair) ptx::cluster_sync();
  else __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // work item w -> (row tile, column tile, K chunk), row tile slowest; dealt round-robin (units that
  // run side by side then walk the same K blocks in phase and L2 merges their operand requests)
  const int work_per_problem = p.tiles_m * p.tiles_n * p.nsplit;
  const int num_work = p.batch * work_per_problem;
  const int num_kb_total = (p.num_k16 + 1) / 2;
  struct Work {
    int b, tm, tn, kb0, kb1, ncols;
  };
  auto decode = [&](int w) {
    Work k;
    k.b = w / work_per_problem;
    w -= k.b * work_per_problem;
    k.tm = w / (p.tiles_n * p.nsplit);
    const int rem = w - k.tm * (p.tiles_n * p.nsplit);
    k.tn = rem / p.nsplit;
    const int sp = rem - k.tn * p.nsplit;
    k.kb0 = sp * p.kb_per_split;
    k.kb1 = min(num_kb_total, k.kb0 + p.kb_per_split);
    k.ncols = (int)min((int64_t)p.tile_cols, ((p.N - (int64_t)k.tn * p.tile_cols) + 15) / 16 * 16);
    return k;
  };
  constexpr int kRowsPerUnit = FM * Cfg::kCtas;
  // j-th item of this unit.  One problem: round-robin over all items (see above).  Batches of problems:
  // a unit takes whole problems (every item of problem unit, unit + nunits, ...), so that the per-problem
  // work of the converters -- bandwidth record, statistics flush -- happens once per problem, not per item
  auto work_at = [&](int j) -> int {
    if (p.batch == 1) {
      const int64_t w = (int64_t)unit + (int64_t)j * nunits;
      return w < num_work ? (int)w : -1;
    }
    const int64_t prob = (int64_t)unit + (int64_t)(j / work_per_problem) * nunits;
    return prob < p.batch ? (int)(prob * work_per_problem + j % work_per_problem) : -1;
  };

  if (warp == kWarpTma) {
    // ------------------------------------------------------------------ TMA producer (every CTA)
    if (lane == 0) {
      uint32_t it = 0;
      for (int j = 0, w; (w = work_at(j)) >= 0; ++j) {
        const Work wk = decode(w);
        // columns of B this CTA stages: its half of the item, or (wide) its halves of the item's two parts
        const int n0 = min(wk.ncols, FN), n1 = wk.ncols - n0;
        const int ncta_cols = kPair ? n0 / 2 : n0;
        const int nbox = (ncta_cols + 63) / 64;
        const int nbox1 = kWide ? (n1 / 2 + 63) / 64 : 0;
        const int row = (int)(wk.b * p.a_batch_rows) + wk.tm * kRowsPerUnit + (int)rank * FM;
        const int col = wk.tn * p.tile_cols + (int)rank * ncta_cols;
        const int col1 = wk.tn * p.tile_cols + FN + (int)rank * (n1 / 2);
        const int vrow0 = (int)(wk.b * p.v_batch_rows);
        for (int kb = wk.kb0; kb < wk.kb1; ++kb, ++it) {
          const int s = it % kStages;
          const uint32_t ph = (it / kStages) & 1u;
          ptx::mbar_wait(empty_bar(s), ph ^ 1u);
          const uint32_t st = sbase + s * kStageBytes;
          // A: this CTA's 128 rows x 32 fp32, completion on its OWN barrier (its converters wait there)
          ptx::mbar_expect_tx(raw_bar(s), kFABytes);
          ptx::tma_load_2d(st, &tmA, raw_bar(s), kb * FK, row);
          // B: completion counted on the leader's barrier (bytes of every CTA of the unit)
          if (leader) ptx::mbar_expect_tx(full_bar(s), Cfg::kCtas * 2 * (nbox + nbox1) * (FK * 128));
          if constexpr (kWide) {
            // second part of the item (columns 256..ncols): boxes 2, 3 of the stage
            for (int q = 0; q < nbox1; ++q) {
              const uint32_t dh = st + kFABytes + (2 + q) * (FK * 128), dl = dh + kBBytes;
              ptx::tma_load_2d_pair(dh, &tmBh, full_bar(s), col1 + q * 64, vrow0 + kb * FK);
              ptx::tma_load_2d_pair(dl, &tmBl, full_bar(s), col1 + q * 64, vrow0 + kb * FK);
            }
          }
#pragma unroll
          for (int q = 0; q < (kWide ? 2 : Cfg::kBCols / 64); ++q) {
            if (q >= nbox) break;
            const uint32_t dh = st + kFABytes + q * (FK * 128), dl = dh + kBBytes;
            if constexpr (kPair) {
              ptx::tma_load_2d_pair(dh, &tmBh, full_bar(s), col + q * 64, vrow0 + kb * FK);
              ptx::tma_load_2d_pair(dl, &tmBl, full_bar(s), col + q * 64, vrow0 + kb * FK);
            } else {
              ptx::tma_load_2d(dh, &tmBh, full_bar(s), col + q * 64, vrow0 + kb * FK);
              ptx::tma_load_2d(dl, &tmBl, full_bar(s), col + q * 64, vrow0 + kb * FK);
            }
          }
        }
      }
    }
  } else if (warp == kWarpMma) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA)
    if (leader && lane == 0) {
      uint32_t it = 0, tcount = 0;
      for (int j = 0, w; (w = work_at(j)) >= 0; ++j, ++tcount) {
        const Work wk = decode(w);
        const int n0 = min(wk.ncols, FN), n1 = wk.ncols - n0;
        const uint32_t idesc = ptx::make_idesc_f16(kRowsPerUnit, n0, false, true);
        const uint32_t idesc1 = ptx::make_idesc_f16(kRowsPerUnit, max(n1, 16), false, true);
        const int a = kWide ? 0 : (int)(tcount & 1);
        const uint32_t aph = kWide ? (tcount & 1u) : ((tcount >> 1) & 1u);
        ptx::mbar_wait(tempty_bar(a), aph ^ 1u);
        ptx::tc_fence_after();
        const uint32_t tacc = tmem_base + a * FN;
        uint32_t accumulate = 0;
        for (int kb = wk.kb0; kb < wk.kb1; ++kb, ++it) {
          const int s = it % kStages;
          const uint32_t ph = (it / kStages) & 1u;
          ptx::mbar_wait(full_bar(s), ph);
          ptx::tc_fence_after();
          const uint32_t st = sbase + s * kStageBytes;
          const int steps = min(2, p.num_k16 - kb * 2);
          for (int k = 0; k < steps; ++k) {
            // A: K-major SWIZZLE_128B rows of [hi0 | lo0 | hi1 | lo1] (32 bytes each)
            const uint64_t ah = ptx::make_smem_desc(st + (2 * k) * 32, 16, 1024, ptx::kLayoutSw128);
            const uint64_t al = ptx::make_smem_desc(st + (2 * k + 1) * 32, 16, 1024, ptx::kLayoutSw128);
            // B: MN-major SWIZZLE_128B, 64-column groups FK*128 bytes apart, 16 K-rows = 2048 bytes
            const uint64_t bh = ptx::make_smem_desc(st + kFABytes + k * 2048, FK * 128, 1024, ptx::kLayoutSw128);
            const uint64_t bl =
                ptx::make_smem_desc(st + kFABytes + kBBytes + k * 2048, FK * 128, 1024, ptx::kLayoutSw128);
            if constexpr (kWide) {
              ptx::umma_f16_pair(tacc, ah, bh, idesc, accumulate);
              ptx::umma_f16_pair(tacc, ah, bl, idesc, 1u);
              ptx::umma_f16_pair(tacc, al, bh, idesc, 1u);
              if (n1 > 0) {
                const uint64_t bh1 = ptx::make_smem_desc(st + kFABytes + 2 * (FK * 128) + k * 2048, FK * 128, 1024,
                                                         ptx::kLayoutSw128);
                const uint64_t bl1 = ptx::make_smem_desc(st + kFABytes + kBBytes + 2 * (FK * 128) + k * 2048, FK * 128,
                                                         1024, ptx::kLayoutSw128);
                ptx::umma_f16_pair(tacc + FN, ah, bh1, idesc1, accumulate);
                ptx::umma_f16_pair(tacc + FN, ah, bl1, idesc1, 1u);
                ptx::umma_f16_pair(tacc + FN, al, bh1, idesc1, 1u);
              }
            } else if constexpr (kPair) {
              ptx::umma_f16_pair(tacc, ah, bh, idesc, accumulate);
              ptx::umma_f16_pair(tacc, ah, bl, idesc, 1u);
              ptx::umma_f16_pair(tacc, al, bh, idesc, 1u);
            } else {
              ptx::umma_f16(tacc, ah, bh, idesc, accumulate);
              ptx::umma_f16(tacc, ah, bl, idesc, 1u);
              ptx::umma_f16(tacc, al, bh, idesc, 1u);
            }
            accumulate = 1u;
          }
          if constexpr (kPair) ptx::umma_commit_pair(empty_bar(s));
          else ptx::umma_commit(empty_bar(s));
        }
        if constexpr (kPair) ptx::umma_commit_pair(tfull_bar(a));
        else ptx::umma_commit(tfull_bar(a));
      }
    }
  } else if (warp < kWarpConv0 + kConvWarps) {
    // ------------------------------------------------------------------ converters (own 128 rows)
    // thread -> (row, half): 16 fp32 in (64 bytes), [8 hi | 8 hi | 8 lo | 8 lo] fp16 out (the same 64
    // bytes).  Eight lanes of a quarter warp touch eight different 16-byte columns of the
    // 128-byte swizzle atom: conflict-free.
    // Two groups of four warps take alternate stages (a thread converts a whole row, as two halves):
    // a stage's fixed latencies (barrier wake-up, shared-memory round trip, proxy fence, arrive) then
    // overlap with the other group's stage instead of adding up.
    const int ct = threadIdx.x - 32 * kWarpConv0;
    static_assert(kConvWarps * 32 == kConvGroups * FM, "one thread per row in every group");
    const int group = ct / FM, row = ct % FM;
    const uint32_t sw = (uint32_t)(row & 7);
    const uint32_t row_off = (uint32_t)row * 128u;
    float g2 = 0.f;
    uint32_t med_bits = 0xffffffffu;  // never matches a non-negative finite distance
    int cur_b = -1;                   // problem whose bandwidth record is loaded
    double acc = 0.0;                 // sum K*d of problem cur_b gathered by this thread
    double* red = reinterpret_cast<double*>(smem + kStages * kStageBytes + kFBarBytes + kFXposeBytes);
    // sum K*d of one problem, this CTA's converter threads -> one atomic (all 256 threads call it)
    auto flush_stats = [&](int b) {
#pragma unroll
      for (int o = 16; o >= 1; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      if (lane == 0) red[warp - kWarpConv0] = acc;
      asm volatile("bar.sync 1, 256;" ::: "memory");  // the eight converter warps
      if (ct == 0) {
        double tot = 0.0;
#pragma unroll
        for (int i = 0; i < kConvWarps; ++i) tot += red[i];
        if (tot != 0.0) atomicAdd(&p.state[b].sum_kd, tot * (1.0 / (double)kKScale));
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
      acc = 0.0;
    };
    uint32_t it = 0;
    const uint32_t leader_full0 = kPair ? ptx::mapa(full_bar(0), 0) : full_bar(0);
    for (int j = 0, w; (w = work_at(j)) >= 0; ++j) {
      const Work wk = decode(w);
      if (wk.b != cur_b) {  // CTA-uniform: every converter thread walks the same items
        if (p.stats && cur_b >= 0) flush_stats(cur_b);
        cur_b = wk.b;
        if (p.mode == SIGSVGD_KF_EXP) {
          g2 = p.state[cur_b].gamma * 1.4426950408889634f;  // log2(e)
          med_bits = (p.stats && p.state[cur_b].median > 0.f) ? __float_as_uint(p.state[cur_b].median) : 0xffffffffu;
        }
      }
      sigsvgd_bw_state* const st_b = p.state + (p.state ? cur_b : 0);
      const int64_t grow = p.row_offset + (int64_t)wk.tm * kRowsPerUnit + (int64_t)rank * FM + row;
      float part = 0.f;
      for (int kb = wk.kb0; kb < wk.kb1; ++kb, ++it) {
        if ((int)(it % (uint32_t)kConvGroups) != group) continue;
        const int s = it % kStages;
        const uint32_t ph = (it / kStages) & 1u;
        ptx::mbar_wait(raw_bar(s), ph);
        // the statistics of K block kb are taken by the column tile kb % tiles_n: every item pays for a
        // 1/tiles_n share instead of the first column tile's items paying for everything
        const bool do_stats = p.stats != 0 && (kb % p.tiles_n) == wk.tn;
        uint8_t* rp = smem + s * kStageBytes + row_off;
#pragma unroll 1
        for (int half = 0; half < 2; ++half) {
          float d[16];
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const float4 f = *reinterpret_cast<const float4*>(rp + ((((uint32_t)(4 * half + c)) ^ sw) << 4));
            d[4 * c] = f.x, d[4 * c + 1] = f.y, d[4 * c + 2] = f.z, d[4 * c + 3] = f.w;
          }
          uint32_t hi[8], lo[8];
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            float k0, k1;
            if (p.mode == SIGSVGD_KF_EXP) {
              // ex2.approx of (gamma log2e) d: relative error 2^-22 + 1.2e-7 |gamma d| (< 5e-8 absolute on K)
              k0 = ex2_approx(g2 * d[2 * q]) * kKScale;
              k1 = ex2_approx(g2 * d[2 * q + 1]) * kKScale;
            } else {
              k0 = d[2 * q] * kKScale;
              k1 = d[2 * q + 1] * kKScale;
            }
            const __half2 h2 = __floats2half2_rn(k0, k1);
            const float2 hf = __half22float2(h2);
            const __half2 l2 = __floats2half2_rn(k0 - hf.x, k1 - hf.y);
            hi[q] = *reinterpret_cast<const uint32_t*>(&h2);
            lo[q] = *reinterpret_cast<const uint32_t*>(&l2);
            if (do_stats) {
              part = fmaf(k0, d[2 * q], part);
              part = fmaf(k1, d[2 * q + 1], part);
            }
          }
          if (do_stats) {
            // entries equal to the median (rare: normally the one symmetric pair).  Distances are >= 0,
            // so the bit patterns compare as integers; one test per half row, details only on a hit
            uint32_t near = 0xffffffffu;
#pragma unroll
            for (int e = 0; e < 16; ++e) near = min(near, __float_as_uint(d[e]) ^ med_bits);
            if (near == 0u) {
#pragma unroll
              for (int e = 0; e < 16; ++e) {
                if (__float_as_uint(d[e]) == med_bits) {
                  const uint32_t slot = atomicAdd(&st_b->n_ties, 1u);
                  if (slot < SIGSVGD_MAX_TIES) {
                    st_b->ties[2 * slot] = (int32_t)grow;
                    st_b->ties[2 * slot + 1] = (int32_t)(kb * FK + 16 * half + e);
                  }
                }
              }
            }
          }
          *reinterpret_cast<uint4*>(rp + ((((uint32_t)(4 * half + 0)) ^ sw) << 4)) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
          *reinterpret_cast<uint4*>(rp + ((((uint32_t)(4 * half + 1)) ^ sw) << 4)) = make_uint4(hi[4], hi[5], hi[6], hi[7]);
          *reinterpret_cast<uint4*>(rp + ((((uint32_t)(4 * half + 2)) ^ sw) << 4)) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
          *reinterpret_cast<uint4*>(rp + ((((uint32_t)(4 * half + 3)) ^ sw) << 4)) = make_uint4(lo[4], lo[5], lo[6], lo[7]);
        }
        ptx::fence_proxy_async_smem();  // generic-proxy stores -> visible to the tensor core's reads
        __syncwarp();
        if (lane == 0) {
          if constexpr (kPair) ptx::mbar_arrive_remote(leader_full0 + 8u * s);
          else ptx::mbar_arrive(full_bar(s));
        }
      }
      acc += (double)part;
    }
    if (p.stats && cur_b >= 0) flush_stats(cur_b);
  } else if (warp >= kWarpEpi0 && warp < kWarpEpi0 + 4) {
    // ------------------------------------------------------------------ epilogue (own 128 rows)
    const int q = warp & 3;
    uint32_t tcount = 0;
    const bool accumulate_out = p.nsplit > 1;
    // 16-byte reductions (several chains per output) when the target allows; the single-CTA kernel (batches
    // of small problems: one chain per output) also has 16-byte plain stores and fetches the column scales
    // one chunk ahead -- measured: scenes 391 -> 250 us; the same two changes in the pair kernels cost the
    // N = 4096 update 8 us (97 vs 89 us), so they are compiled out there
    const bool vec_ok = (p.ldo & 3) == 0 && (p.N & 3) == 0 && (reinterpret_cast<uintptr_t>(p.out) & 15u) == 0;
    const bool vec_io = vec_ok && (accumulate_out || !kPair);
    float* xp = reinterpret_cast<float*>(smem + kStages * kStageBytes + kFBarBytes) + q * (32 * kFPitch);
    const uint32_t leader_tempty0 = kPair ? ptx::mapa(tempty_bar(0), 0) : tempty_bar(0);
    for (int j = 0, w; (w = work_at(j)) >= 0; ++j, ++tcount) {
      const Work wk = decode(w);
      const int a = kWide ? 0 : (int)(tcount & 1);
      const uint32_t aph = kWide ? (tcount & 1u) : ((tcount >> 1) & 1u);
      ptx::mbar_wait(tfull_bar(a), aph);
      ptx::tc_fence_after();
      const int64_t row0 = (int64_t)wk.tm * kRowsPerUnit + (int64_t)rank * FM + q * 32;  // within the problem
      float* const outb = p.out + (int64_t)wk.b * p.M * p.ldo;
      const float* const csb = p.colscale + (int64_t)wk.b * p.colscale_stride;
      // column scales one chunk ahead (a global load per chunk on the drain's critical path otherwise)
      auto load_scale = [&](int c) {
        const int64_t col = (int64_t)wk.tn * p.tile_cols + c * 32 + lane;
        return (c * 32 < wk.ncols && col < p.N) ? __ldg(csb + col) : 0.f;
      };
      float cvl_next = kPair ? 0.f : load_scale(0);
#pragma unroll 1
      for (int c = 0; c < Cfg::kTileCols / 32; ++c) {
        const int64_t col0 = (int64_t)wk.tn * p.tile_cols + c * 32;
        if (c * 32 >= wk.ncols || col0 >= p.N) break;  // warp-uniform
        uint32_t v[32];
        ptx::tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + a * FN + c * 32, v);
        const bool col_ok = col0 + lane < p.N;
        float cvl;
        if constexpr (kPair) {
          cvl = col_ok ? __ldg(csb + col0 + lane) : 0.f;
        } else {
          cvl = cvl_next;
          cvl_next = load_scale(c + 1);
        }
        ptx::tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 32; ++e) xp[lane * kFPitch + e] = __uint_as_float(v[e]);
        __syncwarp();
        if (vec_io) {
          const int cg = (lane & 7) * 4;
          const float s0 = __shfl_sync(0xffffffffu, cvl, cg), s1 = __shfl_sync(0xffffffffu, cvl, cg + 1);
          const float s2 = __shfl_sync(0xffffffffu, cvl, cg + 2), s3 = __shfl_sync(0xffffffffu, cvl, cg + 3);
          if (col0 + cg < p.N) {
#pragma unroll
            for (int rr = 0; rr < 32; rr += 4) {
              const int r = rr + (lane >> 3);
              if (row0 + r < p.M) {
                const float* src = xp + r * kFPitch + cg;
                float* dst = outb + (row0 + r) * p.ldo + col0 + cg;
                if (kPair || accumulate_out) ptx::red_add_v4(dst, src[0] * s0, src[1] * s1, src[2] * s2, src[3] * s3);
                else *reinterpret_cast<float4*>(dst) = make_float4(src[0] * s0, src[1] * s1, src[2] * s2, src[3] * s3);
              }
            }
          }
        } else {
#pragma unroll 8
          for (int r = 0; r < 32; ++r) {
            if (row0 + r < p.M && col_ok) {
              const float val = xp[r * kFPitch + lane] * cvl;
              float* dst = outb + (row0 + r) * p.ldo + col0 + lane;
              if (accumulate_out) atomicAdd(dst, val);
              else *dst = val;
            }
          }
        }
        __syncwarp();
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (kPair) ptx::mbar_arrive_cluster(leader_tempty0 + 8u * a);
        else ptx::mbar_arrive(tempty_bar(a));
      }
    }
  }
  ptx::tc_fence_before();
  if constexpr (kPair) ptx::cluster_sync();  // nobody frees TMEM / exits while the peer's MMAs or arrivals are in flight
  else __syncthreads();
  if (warp == kWarpAlloc) {
    ptx::tc_fence_after();
    if constexpr (kPair) ptx::tmem_dealloc_pair(tmem_base, 512);
    else ptx::tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace

// defined in gemm_tc.cu
int make_tmap_f16(CUtensorMap* tm, const __half* base, int64_t inner, int64_t outer, int64_t ld,
                  uint32_t box_inner, uint32_t box_outer);
int make_tmap_f32(CUtensorMap* tm, const float* base, int64_t inner, int64_t outer, int64_t ld,
                  uint32_t box_inner, uint32_t box_outer);

namespace {

// Pairs that can be co-resident (one 2-CTA cluster per TPC); 0 if cluster launch is unavailable.
int max_pairs() {
  static thread_local int cached_dev = -1, cached = 0;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (dev == cached_dev) return cached;
  cached_dev = dev, cached = 0;
  if (cudaFuncSetAttribute(kf_fused_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)FCfg<true>::kSmemBytes) != cudaSuccess ||
      cudaFuncSetAttribute(kf_fused_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)FCfg<true, true>::kSmemBytes) != cudaSuccess ||
      cudaFuncSetAttribute(kf_fused_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)FCfg<false>::kSmemBytes) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * (unsigned)num_sms());
  cfg.blockDim = dim3(kFThreads);
  cfg.dynamicSmemBytes = FCfg<true>::kSmemBytes;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2, attr[0].val.clusterDim.y = 1, attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr, cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, kf_fused_kernel<true, false>, &cfg) != cudaSuccess) {
    cudaGetLastError();
    n = 0;
  }
  cached = std::max(0, n);
  return cached;
}

}  // namespace

int tc_kf_fused(const float* a, int64_t lda, int64_t M, int64_t K, int mode, sigsvgd_bw_state* state,
                int stats, int64_t row_offset, TcPlanes V, int64_t C, const float* colscale, float* out,
                int64_t ldo, cudaStream_t st, bool out_zeroed, int64_t batch, int64_t colscale_stride) {
  SIG_CHECK_ARG(M >= 1 && K >= 1 && C >= 1 && lda >= K && V.ld >= C && ldo >= C,
                "tc_kf: bad shape M=%lld K=%lld C=%lld", (long long)M, (long long)K, (long long)C);
  SIG_CHECK_ARG(M < (1ll << 31) && K < (1ll << 31), "tc_kf: too many rows");
  SIG_CHECK_ARG(mode == SIGSVGD_KF_EXP || mode == SIGSVGD_KF_IDENTITY, "tc_kf: bad mode %d", mode);
  SIG_CHECK_ARG(state != nullptr || (mode == SIGSVGD_KF_IDENTITY && !stats), "tc_kf: state is NULL");
  SIG_CHECK_ARG((lda & 3) == 0 && (reinterpret_cast<uintptr_t>(a) & 15u) == 0,
                "tc_kf: the fp32 operand must be 16-byte aligned with a row pitch that is a multiple of 4");
  // a batch stacks the problems along the rows of a, V and out; tiles must not straddle two problems
  SIG_CHECK_ARG(batch >= 1 && (batch == 1 || (M % FM == 0 && K % FK == 0)),
                "tc_kf: batched problems need M %% 128 == 0 and K %% 32 == 0 (M=%lld K=%lld)", (long long)M,
                (long long)K);
  SIG_CHECK_ARG(batch * M < (1ll << 31) && batch * K < (1ll << 31), "tc_kf: batch too large");
  const int pairs_max = max_pairs();
  const bool pair = M > FM && pairs_max > 0 && (batch == 1 || M % (2 * FM) == 0);
  CUtensorMap ta, bh, bl;
  SIG_PROPAGATE(make_tmap_f32(&ta, a, K, batch * M, lda, FK, FM));  // 128-byte rows: SWIZZLE_128B
  SIG_PROPAGATE(make_tmap_f16(&bh, V.hi, C, batch * K, V.ld, 64, FK));
  SIG_PROPAGATE(make_tmap_f16(&bl, V.lo, C, batch * K, V.ld, 64, FK));
  FusedParams p;
  p.M = M, p.N = C, p.row_offset = row_offset;
  p.num_k16 = (int)ceil_div(K, 16);
  // wide items (<= 512 columns, pairs only) when there is more than one 256-column tile: the fp32 operand is
  // fetched and converted once per item, so halving the number of column tiles halves that traffic.
  // Columns are dealt evenly over the tiles (widths a multiple of 32).
  // Their epilogue does not overlap the next item's MMAs, so they only pay while there are few, long
  // chains: enough items of full (1024-index) chains to occupy every pair.  Measured at N = 4096, C = 848
  // (us, L2 flushed): narrow 101; wide with chains of 256 / 512 / 1024 indices 132 / 112 / 93
  // (chains of 2048: 83, but twice the truncation error of the TMEM accumulation).
  const int units_max = pair ? pairs_max : num_sms();
  const int num_kb = (int)ceil_div(ceil_div(K, 16), 2);
  // Only for up to four 256-column tiles (C <= 1024: the update's 848 columns): with many column tiles
  // the exposed epilogues cost more than the saved conversions (depth 4, C = 3249: 355 vs 296 us).
  const bool wide = pair && batch == 1 && C > FN && C <= 4 * FN &&
                    ceil_div(M, 2 * FM) * ceil_div(C, 2 * FN) * ceil_div(num_kb, kFMaxChainKb) >= units_max;
  const int64_t max_cols = wide ? 2 * FN : FN;
  p.tiles_m = (int)ceil_div(M, pair ? 2 * FM : FM), p.tiles_n = (int)ceil_div(C, max_cols);
  p.tile_cols = wide ? (int)round_up(ceil_div(C, p.tiles_n), 32) : FN;
  // chains of at most 1024 contraction indices (the TMEM accumulator adds with truncation); shorter
  // ones while there are fewer than ~6 items per unit, never below 8 K blocks
  p.kb_per_split = std::min(num_kb, kFMaxChainKb);
  p.nsplit = (int)ceil_div(num_kb, p.kb_per_split);
  while (!wide && batch * p.tiles_m * p.tiles_n * p.nsplit < 6ll * units_max && num_kb / (2 * p.nsplit) >= 8)
    p.nsplit *= 2;
  p.kb_per_split = (int)ceil_div(num_kb, p.nsplit);
  p.nsplit = (int)ceil_div(num_kb, p.kb_per_split);
  p.mode = mode, p.stats = stats ? 1 : 0;
  p.out = out, p.ldo = ldo, p.colscale = colscale, p.state = state;
  p.batch = (int)batch, p.a_batch_rows = M, p.v_batch_rows = K, p.colscale_stride = colscale_stride;
  if (p.nsplit > 1 && !out_zeroed)
    SIG_CUDA_OK(cudaMemset2DAsync(out, sizeof(float) * ldo, 0, sizeof(float) * C, batch * M, st));
  const int64_t items = batch * p.tiles_m * p.tiles_n * p.nsplit;
  SIG_CHECK_ARG(items < (1ll << 31), "tc_kf: too many work items");
  // (batches: units take whole problems)
  const int units = (int)std::min<int64_t>(batch > 1 ? batch : items, units_max);
  if (!pair) {
    if (pairs_max <= 0) {  // attributes were not set by max_pairs()
      SIG_CUDA_OK(cudaFuncSetAttribute(kf_fused_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)FCfg<false>::kSmemBytes));
    }
    kf_fused_kernel<false, false><<<units, kFThreads, FCfg<false>::kSmemBytes, st>>>(ta, bh, bl, p);
    SIG_LAUNCH_OK("kf_fused_kernel<single>");
    return 0;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * (unsigned)units);
  cfg.blockDim = dim3(kFThreads);
  cfg.dynamicSmemBytes = wide ? FCfg<true, true>::kSmemBytes : FCfg<true>::kSmemBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2, attr[0].val.clusterDim.y = 1, attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr, cfg.numAttrs = 1;
  if (wide) SIG_CUDA_OK(cudaLaunchKernelEx(&cfg, kf_fused_kernel<true, true>, ta, bh, bl, p));
  else SIG_CUDA_OK(cudaLaunchKernelEx(&cfg, kf_fused_kernel<true, false>, ta, bh, bl, p));
  SIG_LAUNCH_OK("kf_fused_kernel<pair>");
  return 0;
}

}  // namespace sigsvgd
