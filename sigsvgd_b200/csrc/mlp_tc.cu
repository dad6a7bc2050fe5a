// Cost side of robot planning: the learned occupancy / self-collision networks
//   Linear(in, H) ReLU  [Linear(H, H) ReLU] x n_mid  Linear(H, 1) Sigmoid
// (reference: src/models/robot_learning/continuous_occupancy_map.py:10-35 and
// continuous_self_collision_pred.py:11-36; H = 200, n_mid = 3, in = 3 or the robot's dof), queried at
// batch x timesteps x body-points positions per cost evaluation (examples/script_planning_robot.py:300-318)
// and differentiated with respect to the query points (the cost's autograd.grad, :107).
//
// One kernel does the whole query: probability AND d probability / d point.  A CTA owns 128 query points
// (one TMEM lane each).  The first layer (K = in <= 8) and the last (N = 1) are SIMT; every H x H layer --
// forward, and again transposed for the input gradient -- is a 128 x 208 x 208 tcgen05 GEMM:
//   A = the activations (or the back-propagated rows) of the 128 points as packed fp16 pairs IN TENSOR MEMORY
//       (row = lane, K = 208 -> 104 columns per plane), written by the epilogue threads with tcgen05.st: the
//       MMA takes its A operand from there (the .ts form), shared memory feeds B only;
//   B = the layer's fp16 weights (or their transpose), streamed by TMA through a ring of [208 rows x 64
//       columns] stages that runs ahead across layers and tiles (the weights never depend on the data);
//   D = fp32 in tensor memory; the epilogue (bias, ReLU, ReLU mask bits, fp16 pack) reads it with
//       tcgen05.ld.  Nothing but the points, the probabilities and the gradients touches HBM.
// The backward pass propagates d logit / d h; the sigmoid's p (1 - p) multiplies the final 1 x in row.
//
// Arithmetic, two modes:
//   precise  every fp32 operand is held as two fp16 planes, x = hi + lo, and a product is accumulated as
//            hi.hi + lo.hi + hi.lo (3 MMAs, fp32 accumulator; the dropped lo.lo term is 2^-22 relative): the
//            reference's fp32 GEMMs to fp32 accuracy, so that the ReLU patterns -- and with them the
//            piecewise-constant Jacobian -- agree with the reference's except where a pre-activation lies
//            within fp32 rounding of zero (where the reference itself is arbitrary);
//   fast     hi planes only (1 MMA): probabilities to ~5e-4; about 1 point in 20 has some pre-activation
//            within the fp16 rounding error of zero and takes the neighbouring linear piece's gradient.
//
// Measured (B200, 3.07 M points, 3 -> 200 x 4 -> 1, probabilities and gradients): precise 5.3 ms, fast 2.85 ms,
// torch fp32 statements + autograd 56 ms.  What was tried on the way (thread 0's cycle counters,
// sigsvgd_mlp_profile_enable, say where a tile's time goes):
//  * the GEMM phase is ~60 % of a tile; only 4-7 % of the tile is spent waiting for weight stages.  Cutting the
//    weight traffic changed nothing: clusters of two CTAs multicasting each other half of every stage, and
//    CTA pairs (cta_group::2, each CTA staging half the weight rows) both ran green at the same speed
//    (tools/experiments/mlp_cluster_multicast.cu.txt, mlp_pair_cta_group2.cu.txt); rings of twice as many
//    half-size stages were 1.5x slower (TMA request count);
//  * the MMA issue path mattered: issued by `if (tid == 0)`, every operand went through a register-to-uniform
//    broadcast loop per MMA (R2UR / ELECT / BRA.U.ANY in the SASS); with the whole warp walking the jobs and
//    one elected lane issuing, the operands live in uniform registers and the MMAs of a stage issue back to
//    back (precise 6.9 -> 5.3 ms with A in tensor memory; A in shared memory had measured 6.0 ms);
//  * an MMA still takes ~165 cycles (floor 128 x 208 / 256 = 104) and that is its throughput, not a latency:
//    splitting the 208 columns into two accumulators (112 + 96) issued alternately was 3 % slower; the serial
//    epilogue (25 %) and first layer (10 %) are the rest.  A second tile in flight would hide them, but two tiles' accumulators and operands
//    (2 x 312 columns) do not fit the 512 columns of tensor memory.
#include <cuda.h>
#include <cuda_fp16.h>

#include <algorithm>
#include <atomic>

#include "common.cuh"
#include "gemm_tc.cuh"
#include "tc_ptx.cuh"

namespace sigsvgd {

int make_tmap_f16(CUtensorMap* tm, const __half* base, int64_t inner, int64_t outer, int64_t ld,
                  uint32_t box_inner, uint32_t box_outer);  // tmap.cu

namespace {

constexpr int kHP = 208;             // padded hidden width (13 k16 steps; N of the GEMM)
constexpr int kMaxIn = 8;            // first-layer fan-in held in registers
constexpr int kMaxMid = 4;           // H x H layers
constexpr int kRows = 128;           // query points per CTA
constexpr int kMlpThreads = 256;     // two threads per point: columns [0, 96) and [96, 208)
constexpr int kSplitCol = 96;
constexpr uint32_t kTmemCols = 512;
// tensor-memory columns: the accumulator [0, 208), the operand A as packed fp16 pairs (K = 208 -> 104 columns):
// hi plane at 256, lo plane (precise) at 384
constexpr uint32_t kTmemAHi = 256, kTmemALo = 384;
// A stage of the weight ring = [208 rows x 64 fp16 columns] of one plane, SWIZZLE_128B.  precise: 8 stage
// loads per GEMM (4 column chunks x 2 planes), 3 stages beside the two planes of the operand tile; fast: 4
// loads per GEMM, 4 stages (a whole layer in flight).
template <bool kPrecise>
struct MlpCfg {
  static constexpr int kStages = 6;
  static constexpr int kJobs = kPrecise ? 8 : 4;
  static constexpr uint32_t kStageBytes = kHP * 128;
};

// fp32 parameter block (floats): W1 [kHP][kMaxIn] | b1 [kHP] | bmid [n_mid][kHP] | wlast [kHP] | blast [4]
__host__ __device__ constexpr int param_floats(int n_mid) { return kHP * (kMaxIn + 2 + n_mid) + 4; }
constexpr int kMaskWords = 4;  // per thread and hidden layer: 12 or 14 groups of 8 columns -> <= 128 bits

struct MlpSmem {
  uint32_t ring, params, parts, bars, total;
};
__host__ __device__ inline MlpSmem mlp_smem(int n_mid, bool precise) {
  MlpSmem s;
  s.ring = 0;
  s.params = s.ring + (precise ? MlpCfg<true>::kStages * MlpCfg<true>::kStageBytes
                               : MlpCfg<false>::kStages * MlpCfg<false>::kStageBytes);
  s.parts = s.params + (uint32_t)((param_floats(n_mid) * 4 + 15) & ~15);
  s.bars = s.parts + (2 + 2 * kMaxIn) * kRows * 4;
  s.total = s.bars + 256 + 1024;  // barriers; + slack for the 1024-byte alignment of the operand tiles
  return s;
}

// Development aid, off unless sigsvgd_mlp_profile_enable(1) was called: thread 0 of every CTA adds the cycles
// it spent per phase (first layer | barrier | MMA issue incl. waits for full stages | wait for the MMAs |
// epilogue | output | of the third: waiting for weights) -- read with sigsvgd_mlp_profile_cycles
// (tools/occupancy_time.py --profile).  It changes no result.
__device__ unsigned long long g_mlp_prof[8];
std::atomic<int> g_profile_enabled{0};

struct MlpArgs {
  const float* params;  // fp32 block of the packed weights
  const float* x;       // [P, in]
  float* prob;          // [P]
  float* grad;          // [P, in] or nullptr
  int64_t P;
  int in_dim, hidden, n_mid, num_tiles, profile;
};

__device__ __forceinline__ void tmem_ld_32x16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32"
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}

// 4 words (8 packed fp16) to this thread's lane of tensor memory, columns [taddr, taddr + 4)
__device__ __forceinline__ void tmem_st_32x4(uint32_t taddr, const uint32_t (&w)[4]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr), "r"(w[0]), "r"(w[1]),
               "r"(w[2]), "r"(w[3])
               : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// D[tmem] (+)= A[tmem] . B[smem]: the A operand read from tensor memory (row = lane, two fp16 per column)
__device__ __forceinline__ void umma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t"
      "}\n"
      :
      : "r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// 8 operand values of this thread's row, columns [col, col + 8) -> the hi plane (and, kLo, the lo plane = the
// rounding remainder) of the operand in tensor memory.  Warp-collective (every lane stores to its own lane).
template <bool kLo>
__device__ __forceinline__ void store_operand8(const float (&h)[8], uint32_t tmem_row, int col) {
  __half2 hh[4];
  uint32_t w[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    hh[k] = __floats2half2_rn(h[2 * k], h[2 * k + 1]);
    w[k] = *reinterpret_cast<const uint32_t*>(&hh[k]);
  }
  tmem_st_32x4(tmem_row + kTmemAHi + (uint32_t)(col >> 1), w);
  if constexpr (kLo) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float2 f = __half22float2(hh[k]);
      const __half2 l = __floats2half2_rn(h[2 * k] - f.x, h[2 * k + 1] - f.y);
      w[k] = *reinterpret_cast<const uint32_t*>(&l);
    }
    tmem_st_32x4(tmem_row + kTmemALo + (uint32_t)(col >> 1), w);
  }
}

enum { kFwd = 0, kFwdLast = 1, kBwd = 2, kBwdLast = 3 };

// One group of 8 accumulator columns [col, col + 8) of this thread's row.
//   kFwd      h = relu(acc + bias)             -> A, mask bits
//   kFwdLast  h = relu(acc + bias); z += h . wlast;  (d logit / d pre) = wlast where h > 0 -> A
//   kBwd      v = acc where the layer's input was active (mask)  -> A
//   kBwdLast  v = acc where h1 was active; gx[i] += v . W1[:, i]
template <int kMode, bool kLo>
__device__ __forceinline__ void group8(const uint32_t* v, int col, int row, uint32_t a_tmem, const float* bias,
                                       const float* wlast, const float* w1, uint32_t& mask_byte, float& z,
                                       float (&gx)[kMaxIn], bool write_a, bool wide_in) {
  float h[8];
  if constexpr (kMode == kFwd || kMode == kFwdLast) {
    const float4 b0 = *reinterpret_cast<const float4*>(bias + col);
    const float4 b1 = *reinterpret_cast<const float4*>(bias + col + 4);
    const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
    uint32_t m = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      h[k] = fminf(fmaxf(__uint_as_float(v[k]) + bb[k], 0.f), 65504.f);
      m |= (h[k] > 0.f ? 1u : 0u) << k;
    }
    mask_byte = m;
    if constexpr (kMode == kFwdLast) {
      const float4 w0 = *reinterpret_cast<const float4*>(wlast + col);
      const float4 w1v = *reinterpret_cast<const float4*>(wlast + col + 4);
      const float ww[8] = {w0.x, w0.y, w0.z, w0.w, w1v.x, w1v.y, w1v.z, w1v.w};
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        z = fmaf(h[k], ww[k], z);
        h[k] = (m >> k) & 1u ? ww[k] : 0.f;
      }
    }
  } else {
    const uint32_t m = mask_byte;
#pragma unroll
    for (int k = 0; k < 8; ++k) h[k] = (m >> k) & 1u ? __uint_as_float(v[k]) : 0.f;
    if constexpr (kMode == kBwdLast) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float4 r0 = *reinterpret_cast<const float4*>(w1 + (col + k) * kMaxIn);
        gx[0] = fmaf(h[k], r0.x, gx[0]), gx[1] = fmaf(h[k], r0.y, gx[1]);
        gx[2] = fmaf(h[k], r0.z, gx[2]), gx[3] = fmaf(h[k], r0.w, gx[3]);
        if (wide_in) {  // (uniform) more than four inputs
          const float4 r1 = *reinterpret_cast<const float4*>(w1 + (col + k) * kMaxIn + 4);
          gx[4] = fmaf(h[k], r1.x, gx[4]), gx[5] = fmaf(h[k], r1.y, gx[5]);
          gx[6] = fmaf(h[k], r1.z, gx[6]), gx[7] = fmaf(h[k], r1.w, gx[7]);
        }
      }
    }
  }
  if (kMode != kBwdLast && write_a) store_operand8<kLo>(h, a_tmem, col);
}

// 32 (or 16) accumulator columns starting at `col`: one mask word (or half of one)
template <int kMode, bool kLo, int kCols>
__device__ __forceinline__ void block_cols(const uint32_t* v, int col, int row, uint32_t a_tmem, const float* bias,
                                           const float* wlast, const float* w1, uint32_t& mask_word, float& z,
                                           float (&gx)[kMaxIn], bool write_a, bool wide_in) {
  uint32_t out_word = 0;
#pragma unroll
  for (int g = 0; g < kCols / 8; ++g) {
    uint32_t mb = (mask_word >> (8 * g)) & 0xffu;
    group8<kMode, kLo>(v + 8 * g, col + 8 * g, row, a_tmem, bias, wlast, w1, mb, z, gx, write_a, wide_in);
    out_word |= mb << (8 * g);
  }
  if constexpr (kMode == kFwd || kMode == kFwdLast) mask_word = out_word;
}

// The epilogue of one GEMM for this thread: its row, columns [0, 96) (half 0) or [96, 208) (half 1).
// mask: this thread's kMaskWords words of the layer (read in the backward modes, written in kFwd).
template <int kMode, bool kLo>
__device__ __forceinline__ void epilogue(uint32_t tmem_row, int half, int row, uint32_t a_tmem, const float* bias,
                                         const float* wlast, const float* w1, uint32_t* mask, float& z,
                                         float (&gx)[kMaxIn], bool write_a, bool wide_in) {
  const int col0 = half ? kSplitCol : 0;
  constexpr bool kReads = kMode == kBwd || kMode == kBwdLast;
  constexpr bool kWrites = kMode == kFwd;
  uint32_t va[32], vb[32];
  ptx::tmem_ld_32x32(tmem_row + col0, va);
  ptx::tmem_ld_32x32(tmem_row + col0 + 32, vb);
  uint32_t m0 = kReads ? mask[0] : 0u, m1 = kReads ? mask[1] : 0u;
  ptx::tmem_ld_wait();
  block_cols<kMode, kLo, 32>(va, col0, row, a_tmem, bias, wlast, w1, m0, z, gx, write_a, wide_in);
  block_cols<kMode, kLo, 32>(vb, col0 + 32, row, a_tmem, bias, wlast, w1, m1, z, gx, write_a, wide_in);
  if (kWrites) mask[0] = m0, mask[1] = m1;
  ptx::tmem_ld_32x32(tmem_row + col0 + 64, va);
  uint32_t vc[16];
  if (half) tmem_ld_32x16(tmem_row + col0 + 96, vc);  // columns 192 .. 207
  uint32_t m2 = kReads ? mask[2] : 0u, m3 = (kReads && half) ? mask[3] : 0u;
  ptx::tmem_ld_wait();
  block_cols<kMode, kLo, 32>(va, col0 + 64, row, a_tmem, bias, wlast, w1, m2, z, gx, write_a, wide_in);
  if (kWrites) mask[2] = m2;
  if (half) {
    block_cols<kMode, kLo, 16>(vc, col0 + 96, row, a_tmem, bias, wlast, w1, m3, z, gx, write_a, wide_in);
    if (kWrites) mask[3] = m3;
  }
}

template <bool kPrecise>
__global__ void __launch_bounds__(kMlpThreads, 1)
mlp_query_kernel(const __grid_constant__ CUtensorMap tmW, const MlpArgs p) {
  using Cfg = MlpCfg<kPrecise>;
  constexpr int NS = Cfg::kStages, J = Cfg::kJobs;
  constexpr uint32_t kStageBytes = Cfg::kStageBytes;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t sbase = (raw + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (sbase - raw);
  const MlpSmem L = mlp_smem(p.n_mid, kPrecise);
  float* const prm = reinterpret_cast<float*>(smem + L.params);
  const float* const w1 = prm;                               // [kHP][kMaxIn]
  const float* const b1 = prm + kHP * kMaxIn;                // [kHP]
  const float* const bmid = b1 + kHP;                        // [n_mid][kHP]
  const float* const wlast = bmid + p.n_mid * kHP;           // [kHP]
  const float* const blast = wlast + kHP;
  // partial logits / gradients of the two threads of a row meet here
  float* const zpart = reinterpret_cast<float*>(smem + L.parts);   // [2][kRows]
  float* const gpart = zpart + 2 * kRows;                          // [2][kMaxIn][kRows]
  uint32_t masks[kMaxMid][kMaskWords];  // ReLU patterns of this thread's columns of h1 .. h_{n_mid} (thread-local)
  const uint32_t bar0 = sbase + L.bars;
  auto full_bar = [&](int st) { return bar0 + 8u * st; };
  auto empty_bar = [&](int st) { return bar0 + 8u * (NS + st); };
  const uint32_t mma_done = bar0 + 8u * (2 * NS);
  volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(smem + L.bars + 8 * (2 * NS + 1));

  const int tid = threadIdx.x, warp = tid >> 5;
  const int half = tid / kRows, row = tid % kRows;
  const bool want_grad = p.grad != nullptr;
  const bool wide_in = p.in_dim > 4;
  const int S = want_grad ? 2 * p.n_mid : p.n_mid;  // GEMMs per tile

  for (int i = tid; i < param_floats(p.n_mid); i += kMlpThreads) prm[i] = __ldg(p.params + i);
  if (tid == 0) {
    ptx::tma_prefetch_desc(&tmW);
    for (int st = 0; st < NS; ++st) {
      ptx::mbar_init(full_bar(st), 1);
      ptx::mbar_init(empty_bar(st), 1);
    }
    ptx::mbar_init(mma_done, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 1) ptx::tmem_alloc(ptx::smem_u32((const void*)tmem_slot), kTmemCols);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_row = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);

  // ---- the weight stream.  Stage load q of this CTA: GEMM (q / J) % S of its (q / (J S))-th tile,
  // job q % J = (column chunk c, plane).  GEMM s multiplies by layer s (s < n_mid) or by the
  // transpose of layer 2 n_mid - 1 - s.
  const int my_tiles = (int)blockIdx.x < p.num_tiles ? (p.num_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  const uint32_t total_loads = (uint32_t)my_tiles * (uint32_t)S * (uint32_t)J;
  auto job_of = [&](int j, int& c, int& plane) {
    if constexpr (kPrecise) c = j >> 1, plane = j & 1;
    else c = j, plane = 0;
  };
  auto issue_load = [&](uint32_t q) {
    const int s = (int)((q / J) % (uint32_t)S), j = (int)(q % J);
    int c, plane;
    job_of(j, c, plane);
    const int l2 = s < p.n_mid ? s : p.n_mid + (2 * p.n_mid - 1 - s);
    const int st = (int)(q % NS);
    if (q >= (uint32_t)NS) ptx::mbar_wait(empty_bar(st), ((q / NS) - 1u) & 1u);  // the MMAs of load q - NS are done
    const int row0 = (plane * 2 * p.n_mid + l2) * kHP;
    ptx::mbar_expect_tx(full_bar(st), kStageBytes);
    ptx::tma_load_2d(sbase + L.ring + (uint32_t)st * kStageBytes, &tmW, full_bar(st), c * 64, row0);
  };
  // Two single-thread roles during a GEMM, both epilogue threads otherwise: thread 0 issues the MMAs (it waits
  // for full stages only), thread 32 refills the ring NS loads ahead (it waits for the MMAs that free a stage).
  uint32_t head = 0;    // next stage load to issue (thread 32)
  uint32_t gemm_q = 0;  // first stage load of the current GEMM (uniform)
  if (tid == 32) {
    while (head < total_loads && head < (uint32_t)NS) issue_load(head++);
  }
  uint32_t mphase = 0;
  long long prof[7] = {0, 0, 0, 0, 0, 0, 0}, t_prev = 0;
  const bool profiling = p.profile && tid == 0;
  auto lap = [&](int k) {
    if (profiling) {
      const long long t = clock64();
      prof[k] += t - t_prev;
      t_prev = t;
    }
  };
  if (profiling) t_prev = clock64();
  constexpr uint32_t idesc = ptx::make_idesc_f16(kRows, kHP, false, false);

  for (int tile = blockIdx.x; tile < p.num_tiles; tile += gridDim.x) {
    const int64_t r = (int64_t)tile * kRows + row;
    const bool row_ok = r < p.P;
    // ---------------------------------------------------------------- first layer (SIMT)
    float x[kMaxIn];
#pragma unroll
    for (int i = 0; i < kMaxIn; ++i) x[i] = (row_ok && i < p.in_dim) ? __ldg(p.x + r * p.in_dim + i) : 0.f;
    {
      const int col0 = half ? kSplitCol : 0, ngroups = half ? 14 : 12;
      uint32_t word = 0;
      for (int g = 0; g < ngroups; ++g) {
        const int col = col0 + 8 * g;
        float h[8];
        uint32_t m = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float4 r0 = *reinterpret_cast<const float4*>(w1 + (col + k) * kMaxIn);
          float acc = b1[col + k];
          acc = fmaf(r0.x, x[0], acc), acc = fmaf(r0.y, x[1], acc), acc = fmaf(r0.z, x[2], acc);
          acc = fmaf(r0.w, x[3], acc);
          if (wide_in) {  // (uniform) more than four inputs
            const float4 r1 = *reinterpret_cast<const float4*>(w1 + (col + k) * kMaxIn + 4);
            acc = fmaf(r1.x, x[4], acc), acc = fmaf(r1.y, x[5], acc);
            acc = fmaf(r1.z, x[6], acc), acc = fmaf(r1.w, x[7], acc);
          }
          h[k] = fminf(fmaxf(acc, 0.f), 65504.f);
          m |= (h[k] > 0.f ? 1u : 0u) << k;
        }
        store_operand8<kPrecise>(h, tmem_row, col);
        word |= m << (8 * (g & 3));
        if ((g & 3) == 3 || g == ngroups - 1) {
          masks[0][g >> 2] = word;  // mask layer 0 = h1
          word = 0;
        }
      }
    }
    lap(0);
    float z = 0.f;
    float gx[kMaxIn];
#pragma unroll
    for (int i = 0; i < kMaxIn; ++i) gx[i] = 0.f;

    for (int s = 0; s < S; ++s) {
      tmem_st_wait();  // this thread's operand stores have landed in tensor memory
      ptx::tc_fence_before();
      __syncthreads();
      lap(1);
      if (tid == 32) {
        // (not the slot of this GEMM's last job: waiting for it would hold this thread -- and, at the
        //  warp-collective tcgen05.ld, its warp -- past the end of the GEMM; it is refilled first thing in the
        //  next one, three jobs before it is needed)
        const uint32_t lim = min(total_loads, gemm_q + (uint32_t)(J + NS - 1));
        while (head < lim) issue_load(head++);
      } else if (warp == 0) {
        // The whole warp walks the jobs (so that every operand of the MMAs is a warp-uniform value, which the
        // compiler keeps in uniform registers: computed by thread 0 alone they went through per-MMA
        // register-to-uniform broadcasts, and the issue path, not the tensor pipe, set the pace); one elected
        // lane issues.
        ptx::tc_fence_after();
        const uint32_t ring_u = __shfl_sync(0xffffffffu, sbase + L.ring, 0);
        const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
        for (int j = 0; j < J; ++j) {
          const uint32_t q = gemm_q + (uint32_t)j;
          int c, plane;
          job_of(j, c, plane);
          const uint32_t st = q % (uint32_t)NS;
          const long long tw0 = profiling ? clock64() : 0;
          ptx::mbar_wait(full_bar((int)st), (q / NS) & 1u);
          ptx::tc_fence_after();
          if (profiling) prof[6] += clock64() - tw0;  // (part of phase 2) waiting for the weights
          const uint32_t bst = ring_u + st * kStageBytes;
          const uint32_t acol = tmem_u + kTmemAHi + (uint32_t)c * 32u;  // 64 fp16 = 32 columns per chunk
          const int ksteps = c < 3 ? 4 : 1;  // columns 192 .. 207 only in the last chunk
          if (ptx::elect_one()) {
            for (int kk = 0; kk < ksteps; ++kk) {
              const uint64_t bd = ptx::make_smem_desc(bst + kk * 32, 16, 1024, ptx::kLayoutSw128);
              const uint32_t ah = acol + (uint32_t)kk * 8u;  // 16 fp16 = 8 columns per K step
              if (plane == 0) {
                umma_f16_ts(tmem_u, ah, bd, idesc, (c | kk) ? 1u : 0u);  // hi . hi opens the accumulator
                if constexpr (kPrecise) {
                  umma_f16_ts(tmem_u, ah + (kTmemALo - kTmemAHi), bd, idesc, 1u);  // lo . hi
                }
              } else {
                umma_f16_ts(tmem_u, ah, bd, idesc, 1u);  // hi . lo
              }
            }
            ptx::umma_commit(empty_bar((int)st));
          }
          __syncwarp();
        }
        if (ptx::elect_one()) ptx::umma_commit(mma_done);
        __syncwarp();
      }
      gemm_q += (uint32_t)J;
      lap(2);
      ptx::mbar_wait(mma_done, mphase);
      mphase ^= 1u;
      ptx::tc_fence_after();
      lap(3);
      if (s < p.n_mid - 1) {  // hidden layer s+2; its mask is layer s+1
        epilogue<kFwd, kPrecise>(tmem_row, half, row, tmem_row, bmid + s * kHP, wlast, w1,
                                 masks[s + 1], z, gx, true, wide_in);
      } else if (s == p.n_mid - 1) {
        epilogue<kFwdLast, kPrecise>(tmem_row, half, row, tmem_row, bmid + s * kHP, wlast, w1, nullptr, z, gx, want_grad, wide_in);
      } else {
        const int l = 2 * p.n_mid - 1 - s;  // transposed layer l: the result is masked by h_{l+1} = mask layer l
        uint32_t* const mk = masks[l];
        if (l > 0) epilogue<kBwd, kPrecise>(tmem_row, half, row, tmem_row, nullptr, wlast, w1, mk, z, gx, true, wide_in);
        else epilogue<kBwdLast, kPrecise>(tmem_row, half, row, tmem_row, nullptr, wlast, w1, mk, z, gx, false, wide_in);
      }
      lap(4);
    }
    zpart[half * kRows + row] = z;
    if (want_grad) {
#pragma unroll
      for (int i = 0; i < kMaxIn; ++i) gpart[(half * kMaxIn + i) * kRows + row] = gx[i];
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (half == 0 && row_ok) {
      const float logit = zpart[row] + zpart[kRows + row] + blast[0];
      const float pr = 1.f / (1.f + __expf(-logit));
      p.prob[r] = pr;
      if (want_grad) {
        const float gz = pr * (1.f - pr);
        for (int i = 0; i < p.in_dim; ++i)
          p.grad[r * p.in_dim + i] = gz * (gpart[i * kRows + row] + gpart[(kMaxIn + i) * kRows + row]);
      }
    }
    lap(5);
  }
  if (profiling) {
    for (int k = 0; k < 7; ++k) atomicAdd(&g_mlp_prof[k], (unsigned long long)prof[k]);
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem_base, kTmemCols);
}

// fp16 weight planes [hi | lo][2 n_mid][kHP][kHP] (layer l, then the transposes; lo = the rounding remainder
// of hi) and the fp32 parameter block
__global__ void mlp_pack_kernel(const float* __restrict__ w_first, const float* __restrict__ b_first,
                                const float* __restrict__ w_mid, const float* __restrict__ b_mid,
                                const float* __restrict__ w_last, const float* __restrict__ b_last, int in_dim, int H,
                                int n_mid, __half* __restrict__ wh, float* __restrict__ prm) {
  const int64_t n16 = (int64_t)2 * n_mid * kHP * kHP;
  const int64_t n32 = param_floats(n_mid);
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n16 + n32; i += (int64_t)gridDim.x * blockDim.x) {
    if (i < n16) {
      const int l2 = (int)(i / (kHP * kHP)), n = (int)(i / kHP) % kHP, k = (int)(i % kHP);
      float v = 0.f;
      if (n < H && k < H) v = l2 < n_mid ? w_mid[((int64_t)l2 * H + n) * H + k] : w_mid[((int64_t)(l2 - n_mid) * H + k) * H + n];
      const __half hi = __float2half_rn(v);
      wh[i] = hi;
      wh[n16 + i] = __float2half_rn(v - __half2float(hi));
    } else {
      const int j = (int)(i - n16);
      float v = 0.f;
      if (j < kHP * kMaxIn) {
        const int u = j / kMaxIn, c = j % kMaxIn;
        if (u < H && c < in_dim) v = w_first[u * in_dim + c];
      } else if (j < kHP * (kMaxIn + 1)) {
        const int u = j - kHP * kMaxIn;
        if (u < H) v = b_first[u];
      } else if (j < kHP * (kMaxIn + 1 + n_mid)) {
        const int q = j - kHP * (kMaxIn + 1), l = q / kHP, u = q % kHP;
        if (u < H) v = b_mid[l * H + u];
      } else if (j < kHP * (kMaxIn + 2 + n_mid)) {
        const int u = j - kHP * (kMaxIn + 1 + n_mid);
        if (u < H) v = w_last[u];
      } else if (j == kHP * (kMaxIn + 2 + n_mid)) {
        v = b_last[0];
      }
      prm[j] = v;
    }
  }
}

int mlp_check(int in_dim, int hidden, int n_mid) {
  SIG_CHECK_ARG(in_dim >= 1 && in_dim <= kMaxIn, "mlp: in_dim must be in [1, %d] (got %d)", kMaxIn, in_dim);
  SIG_CHECK_ARG(hidden >= 1 && hidden <= kHP, "mlp: hidden width must be in [1, %d] (got %d)", kHP, hidden);
  if (n_mid < 1 || n_mid > kMaxMid) {
    set_error("mlp: %d hidden-to-hidden layers; this kernel takes 1 .. %d", n_mid, kMaxMid);
    return SIGSVGD_ERR_UNSUPPORTED;
  }
  return 0;
}

size_t plane_halfs(int n_mid) { return (size_t)2 * n_mid * kHP * kHP; }

template <bool kPrecise>
int launch_query(const CUtensorMap& tw, const MlpArgs& a, cudaStream_t st) {
  static thread_local int configured_dev = -1;
  int dev = 0;
  SIG_CUDA_OK(cudaGetDevice(&dev));
  if (configured_dev != dev) {
    SIG_CUDA_OK(cudaFuncSetAttribute(mlp_query_kernel<kPrecise>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)mlp_smem(kMaxMid, kPrecise).total));
    configured_dev = dev;
  }
  const int grid = std::min(a.num_tiles, num_sms());
  mlp_query_kernel<kPrecise><<<grid, kMlpThreads, mlp_smem(a.n_mid, kPrecise).total, st>>>(tw, a);
  SIG_LAUNCH_OK("mlp_query_kernel");
  return 0;
}

}  // namespace

int mlp_profile_enable(int on) {
  g_profile_enabled.store(on ? 1 : 0, std::memory_order_relaxed);
  return 0;
}

int mlp_profile_cycles(unsigned long long* out8, int reset) {
  SIG_CUDA_OK(cudaMemcpyFromSymbol(out8, g_mlp_prof, sizeof(g_mlp_prof)));
  if (reset) {
    unsigned long long zero[8] = {0};
    SIG_CUDA_OK(cudaMemcpyToSymbol(g_mlp_prof, zero, sizeof(zero)));
  }
  return 0;
}

size_t mlp_packed_bytes(int in_dim, int hidden, int n_mid) {
  if (mlp_check(in_dim, hidden, n_mid) != 0) return 0;
  return 2 * plane_halfs(n_mid) * sizeof(__half) + (size_t)param_floats(n_mid) * sizeof(float);
}

int launch_mlp_pack(const float* w_first, const float* b_first, const float* w_mid, const float* b_mid,
                    const float* w_last, const float* b_last, int in_dim, int hidden, int n_mid, void* packed,
                    size_t packed_bytes, cudaStream_t st) {
  SIG_PROPAGATE(mlp_check(in_dim, hidden, n_mid));
  SIG_CHECK_ARG(w_first && b_first && w_mid && b_mid && w_last && b_last && packed, "mlp_pack: NULL pointer");
  SIG_CHECK_ARG((reinterpret_cast<uintptr_t>(packed) & 15u) == 0, "mlp_pack: the packed buffer must be 16-byte aligned");
  if (packed_bytes < mlp_packed_bytes(in_dim, hidden, n_mid)) {
    set_error("mlp_pack: packed buffer of %zu bytes, %zu needed", packed_bytes, mlp_packed_bytes(in_dim, hidden, n_mid));
    return SIGSVGD_ERR_WORKSPACE;
  }
  __half* wh = reinterpret_cast<__half*>(packed);
  float* prm = reinterpret_cast<float*>(wh + 2 * plane_halfs(n_mid));
  mlp_pack_kernel<<<num_sms() * 2, 256, 0, st>>>(w_first, b_first, w_mid, b_mid, w_last, b_last, in_dim, hidden, n_mid,
                                                 wh, prm);
  SIG_LAUNCH_OK("mlp_pack_kernel");
  return 0;
}

int launch_mlp_query(const void* packed, int in_dim, int hidden, int n_mid, int precise, const float* x, int64_t P,
                     float* prob, float* grad, cudaStream_t st) {
  SIG_PROPAGATE(mlp_check(in_dim, hidden, n_mid));
  SIG_CHECK_ARG(P >= 0 && P < (1ll << 31) * kRows, "mlp_query: bad number of points");
  if (P == 0) return 0;
  SIG_CHECK_ARG(packed && x && prob, "mlp_query: NULL pointer");
  if (!tc_available()) {
    set_error("mlp_query: the driver does not export cuTensorMapEncodeTiled (no TMA): no fallback exists");
    return SIGSVGD_ERR_UNSUPPORTED;
  }
  const __half* wh = reinterpret_cast<const __half*>(packed);
  CUtensorMap tw;  // boxes of [208 rows x 64 columns]: one stage
  SIG_PROPAGATE(make_tmap_f16(&tw, wh, kHP, (int64_t)4 * n_mid * kHP, kHP, 64, kHP));
  MlpArgs a;
  a.params = reinterpret_cast<const float*>(wh + 2 * plane_halfs(n_mid));
  a.x = x, a.prob = prob, a.grad = grad, a.P = P;
  a.in_dim = in_dim, a.hidden = hidden, a.n_mid = n_mid;
  a.num_tiles = (int)ceil_div(P, kRows);
  a.profile = g_profile_enabled.load(std::memory_order_relaxed);
  return precise ? launch_query<true>(tw, a, st) : launch_query<false>(tw, a, st);
}

}  // namespace sigsvgd
