// Pairwise squared distances of signature features on the tensor cores (K3), with the median
// bookkeeping fused into the epilogue.  Reference: pw_dist_sq, src/utils/math.py:69-86.
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
constexpr int kXposePitch = 33;                          // floats per row of a 32 x 32 transpose tile
constexpr uint32_t kXposeBytes = 4 * 32 * kXposePitch * 4;  // one tile per epilogue warp
constexpr int kCandStage = 1024;                         // staged median candidates per epilogue warp
constexpr uint32_t kCandBytes = 4 * kCandStage * 4;
constexpr uint32_t kSmemBytes = kStages * kStageBytes + kBarBytes + kXposeBytes + kCandBytes + 1024;
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
  SelWindow* win;
  float* cand;
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

template <bool kSym>
__global__ void __launch_bounds__(kThreads, 1)
gram_kernel(const __grid_constant__ CUtensorMap tmAh, const __grid_constant__ CUtensorMap tmAl,
            const __grid_constant__ CUtensorMap tmBh, const __grid_constant__ CUtensorMap tmBl,
            const GramParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = ptx::smem_u32(smem_raw);
  const uint32_t sbase = (raw + 1023u) & ~1023u;
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
  auto decode = [&](int w, int& tm, int& tn) {
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
        for (int kb = 0; kb < num_kb; ++kb, ++it) {
          const int s = it % kStages;
          const uint32_t ph = (it / kStages) & 1u;
          ptx::mbar_wait(empty_bar(s), ph ^ 1u);
          const uint32_t st = sbase + s * kStageBytes;
          ptx::mbar_expect_tx(full_bar(s), kStageBytes);
          ptx::tma_load_2d(st, &tmAh, full_bar(s), kb * GK, tm * GM);
          ptx::tma_load_2d(st + kPlaneBytes, &tmAl, full_bar(s), kb * GK, tm * GM);
          ptx::tma_load_2d(st + 2 * kPlaneBytes, &tmBh, full_bar(s), kb * GK, tn * GN);
          ptx::tma_load_2d(st + 3 * kPlaneBytes, &tmBl, full_bar(s), kb * GK, tn * GN);
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
    const int q = warp & 3;
    uint32_t tcount = 0;
    const float scal0 = __ldg(p.scal);
    const bool use_win = p.win != nullptr && p.win->active != 0;
    float wlo = 0.f, whi = 0.f;
    unsigned int cap = 0;
    if (use_win) wlo = p.win->lo, whi = p.win->hi, cap = p.win->cap;
    float* xp = reinterpret_cast<float*>(smem + kStages * kStageBytes + kBarBytes) + q * (32 * kXposePitch);
    float* cstage = reinterpret_cast<float*>(smem + kStages * kStageBytes + kBarBytes + kXposeBytes) + q * kCandStage;
    int cfill = 0;                     // staged candidates of this warp (warp-uniform)
    unsigned long long below = 0ull;   // per-thread count of entries < lo (x multiplicity)
    // flush the warp's staged candidates: one global atomic, coalesced copy
    auto flush_cands = [&]() {
      if (cfill == 0) return;
      unsigned int base = 0;
      if (lane == 0) base = atomicAdd(&p.win->n_cand, (unsigned int)cfill);
      base = __shfl_sync(0xffffffffu, base, 0);
      for (int i = lane; i < cfill; i += 32)
        if (base + i < cap) p.cand[base + i] = cstage[i];
      __syncwarp();
      cfill = 0;
    };
    for (int w = blockIdx.x; w < p.num_work; w += gridDim.x, ++tcount) {
      int tm, tn;
      decode(w, tm, tn);
      const int a = tcount & 1;
      const uint32_t aph = (tcount >> 1) & 1u;
      ptx::mbar_wait(tfull_bar(a), aph);
      ptx::tc_fence_after();
      const int64_t row0 = (int64_t)tm * GM + q * 32;  // first row of this warp
      const int64_t row = row0 + lane;
      const bool row_ok = row < p.M;
      const float rv = row_ok ? __ldg(p.na + row) : 0.f;
      // symmetric mode: an entry above the diagonal stands for two matrix entries (it is stored
      // twice and counted twice), the diagonal for one, and entries below the diagonal (they only
      // occur inside diagonal tiles) are dropped in favour of their mirror image
      const bool diag = kSym && tm == tn;
      unsigned int below1 = 0, below2 = 0;
#pragma unroll 1
      for (int c = 0; c < GN / 32; ++c) {
        const int64_t col0 = (int64_t)tn * GN + c * 32;
        if (col0 >= p.N) break;  // warp-uniform
        uint32_t v[32], u[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + a * 256 + c * 32;
        ptx::tmem_ld_32x32(taddr, v);
        ptx::tmem_ld_32x32(taddr + 128, u);
        // column norms of this chunk: one coalesced load, broadcast by shuffle
        const float cvl = (col0 + lane < p.N) ? __ldg(p.nb + col0 + lane) : 0.f;
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
        uint32_t hitmask = 0;
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          const float cv = __shfl_sync(0xffffffffu, cvl, e);
          const float acc = __uint_as_float(v[e]) + __uint_as_float(u[e]);
          float dv = fmaxf(fmaf(-2.f, acc, rv + cv) * scal0, 0.f);
          if (kSym && ((dgmask >> e) & 1u)) dv = 0.f;
          o[e] = dv;
          if (use_win) {
            if (dv < wlo) below2 += (upmask >> e) & 1u, below1 += (dgmask >> e) & 1u;
            if (dv >= wlo && dv <= whi) hitmask |= 1u << e;
          }
        }
        if (use_win) {
          const uint32_t hit2 = hitmask & upmask, hit1 = hitmask & dgmask;
          const int hits = 2 * __popc(hit2) + __popc(hit1);
          int incl = hits;
#pragma unroll
          for (int sft = 1; sft < 32; sft <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, sft);
            if (lane >= sft) incl += t;
          }
          const int total = __shfl_sync(0xffffffffu, incl, 31);
          if (total > 0) {  // warp-uniform
            if (cfill + total > kCandStage) flush_cands();
            if (total <= kCandStage) {
              int off = cfill + incl - hits;
#pragma unroll
              for (int e = 0; e < 32; ++e) {
                if ((hit2 >> e) & 1u) {
                  cstage[off] = o[e];
                  cstage[off + 1] = o[e];
                  off += 2;
                } else if ((hit1 >> e) & 1u) {
                  cstage[off++] = o[e];
                }
              }
              cfill += total;
              __syncwarp();
            } else {
              // dense window (more hits than the stage holds): straight to global
              unsigned int base = 0;
              if (lane == 31) base = atomicAdd(&p.win->n_cand, (unsigned int)total);
              base = __shfl_sync(0xffffffffu, base, 31);
              unsigned int off = base + (unsigned int)(incl - hits);
#pragma unroll
              for (int e = 0; e < 32; ++e) {
                if ((hit2 >> e) & 1u) {
                  if (off < cap) p.cand[off] = o[e];
                  if (off + 1 < cap) p.cand[off + 1] = o[e];
                  off += 2;
                } else if ((hit1 >> e) & 1u) {
                  if (off < cap) p.cand[off] = o[e];
                  off += 1;
                }
              }
            }
          }
        }
        // transposed block straight from registers: lanes hold consecutive rows, so every store
        // instruction writes one coalesced line of the mirrored block
        if (kSym && upmask) {
#pragma unroll
          for (int e = 0; e < 32; ++e)
            if ((upmask >> e) & 1u) p.out[(col0 + e) * p.ldo + row] = o[e];
        }
        // direct block through a 32 x 32 shared-memory transpose so that each store instruction
        // writes 32 consecutive columns of one row (a thread owns a row in TMEM)
#pragma unroll
        for (int e = 0; e < 32; ++e) xp[lane * kXposePitch + e] = o[e];
        __syncwarp();
        const int64_t col = col0 + lane;
        const bool col_ok = col < p.N;
#pragma unroll 8
        for (int r = 0; r < 32; ++r) {
          const int64_t rr = row0 + r;
          if (rr < p.M && col_ok && (!diag || col >= rr)) p.out[rr * p.ldo + col] = xp[r * kXposePitch + lane];
        }
        __syncwarp();
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(tempty_bar(a));
      below += (unsigned long long)below1 + 2ull * (unsigned long long)below2;
    }
    if (use_win) {
      flush_cands();
#pragma unroll
      for (int sft = 16; sft >= 1; sft >>= 1) below += __shfl_xor_sync(0xffffffffu, below, sft);
      if (lane == 0 && below) atomicAdd(&p.win->below, below);
    }
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

int tc_gram(TcPlanes A, TcPlanes B, int64_t M, int64_t N, int64_t D, const float* na,
            const float* nb, const float* scal, int symmetric, SelWindow* win, float* cand,
            float* dist, int64_t ldd, cudaStream_t st) {
  SIG_CHECK_ARG(M >= 1 && N >= 1 && D >= 1 && A.ld >= D && B.ld >= D && ldd >= N,
                "tc_gram: bad shape M=%lld N=%lld D=%lld", (long long)M, (long long)N, (long long)D);
  SIG_CHECK_ARG(M < (1ll << 30) && N < (1ll << 30), "tc_gram: too many rows");
  SIG_CHECK_ARG(!symmetric || (M == N && A.hi == B.hi && A.lo == B.lo && na == nb),
                "tc_gram: symmetric mode needs identical operands");
  SIG_CHECK_ARG((win == nullptr) == (cand == nullptr), "tc_gram: win and cand go together");
  CUtensorMap ah, al, bh, bl;
  SIG_PROPAGATE(make_tmap_f16(&ah, A.hi, D, M, A.ld, GK, GM));
  SIG_PROPAGATE(make_tmap_f16(&al, A.lo, D, M, A.ld, GK, GM));
  SIG_PROPAGATE(make_tmap_f16(&bh, B.hi, D, N, B.ld, GK, GN));
  SIG_PROPAGATE(make_tmap_f16(&bl, B.lo, D, N, B.ld, GK, GN));
  GramParams p;
  p.M = M, p.N = N;
  p.num_k16 = (int)ceil_div(D, 16);
  p.tiles_m = (int)ceil_div(M, GM), p.tiles_n = (int)ceil_div(N, GN);
  const int64_t work = symmetric ? (int64_t)p.tiles_m * (p.tiles_m + 1) / 2 : (int64_t)p.tiles_m * p.tiles_n;
  SIG_CHECK_ARG(work < (1ll << 31) && (!symmetric || p.tiles_m <= 2048), "tc_gram: too many tiles");
  p.num_work = (int)work;
  p.out = dist, p.ldo = ldd;
  p.na = na, p.nb = nb, p.scal = scal;
  p.win = win, p.cand = cand;
  static thread_local int configured_dev = -1;
  int dev = 0;
  SIG_CUDA_OK(cudaGetDevice(&dev));
  if (configured_dev != dev) {
    SIG_CUDA_OK(cudaFuncSetAttribute(gram_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)kSmemBytes));
    SIG_CUDA_OK(cudaFuncSetAttribute(gram_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)kSmemBytes));
    configured_dev = dev;
  }
  const int grid = (int)std::min<int64_t>(work, num_sms());
  if (symmetric)
    gram_kernel<true><<<grid, kThreads, kSmemBytes, st>>>(ah, al, bh, bl, p);
  else
    gram_kernel<false><<<grid, kThreads, kSmemBytes, st>>>(ah, al, bh, bl, p);
  SIG_LAUNCH_OK("gram_kernel");
  return 0;
}

}  // namespace sigsvgd
