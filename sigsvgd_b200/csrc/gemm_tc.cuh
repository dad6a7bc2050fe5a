// tcgen05 / TMA tensor-core path for the two N^2 contractions (declared here, gemm_tc.cu).
//
// fp32-grade accuracy on the fp16 tensor pipe: every fp32 operand x is held as two fp16 "planes"
//   x * alpha = hi + lo ,  hi = fp16(x*alpha), lo = fp16(x*alpha - hi)       (alpha = power of two)
// and a product is accumulated as  hi.hi + hi.lo + lo.hi  in fp32 TMEM accumulators (3 MMAs of
// kind::f16 per fp32 product; the dropped lo.lo term is 2^-22 relative).  See DESIGN.md.
#pragma once
#include <cuda_fp16.h>

#include "common.cuh"

namespace sigsvgd {

// fp16 hi/lo planes of a row-major matrix: element (r, c) at hi[r*ld + c], lo[r*ld + c].
// ld (elements) must be a multiple of 8 and both pointers 16-byte aligned (TMA constraints).
struct TcPlanes {
  const __half* hi;
  const __half* lo;
  int64_t ld;
};

// Candidate window of the single-pass exact median (median_window.cu): in ONE pass over the
// freshly written distance matrix, entries below `lo` are counted and entries inside [lo, hi] are
// compacted into candidate pools; the order statistic is then selected among the candidates only.
struct SelWindow {
  float lo, hi;              // window bounds (sample quantiles bracketing the median)
  unsigned int n_cand;       // total candidates (filled in by the select kernel)
  unsigned int cap;          // capacity of the candidate buffer (all pools)
  unsigned long long below;  // matrix entries < lo
  unsigned int active;       // 0: no window (fixed bandwidth) -> epilogue skips the bookkeeping
  unsigned int fallback;     // set by the select kernel when the window missed the median
  // The candidate buffer is cut into `npool` pools of `poolcap` entries; block b of the select
  // kernel appends to pool b through a shared-memory counter (no global atomics) and leaves its
  // fill count in pool_cnt[b] (a count above poolcap = overflow -> fallback).
  unsigned int npool, poolcap;
  unsigned int epoch;        // parity selects the histogram / counter buffers of the next select
};
constexpr int kMaxPools = 1024;
// Counter block behind the pools: pool_cnt[0..kMaxPools) fill of pool b, [kMaxPools..2*kMaxPools)
// entries below the window seen by block b, then 64 words `misc`:
//   misc[epoch]            fill of the short list          misc[2]            (unused)
//   misc[8 + 4*epoch + 0]  candidates kept (all pools)     misc[8 + 4*epoch + 1]  pool-overflow flag
//   misc[8 + 4*epoch + 2..3]  entries below the window (u64)
//   misc[16 + epoch]       ticket of the short-list finish
// (epoch = SelWindow::epoch & 1 double-buffers what a select leaves behind for the next one)
constexpr int kWinBins = 2048;      // histogram bins of the select kernel (top 11 bits of bits - lo_bits)

// dist[i,j] = max(0, (na[i] + nb[j] - 2 * sum_c A[i,c] B[j,c]) * scal[0])   i < M, j < N, c < D
//   A, B: planes [M x >=D], [N x >=D] (features along the row: K-major for both operands)
//   na, nb: squared row norms in plane units; scal: DEVICE pointer, scal[0] = alpha^-2
// 128 x 128 tiles, separate TMEM accumulators for the hi.hi and the
// correction products, TMA-store epilogue.
// symmetric != 0: A and B are the same planes (M == N); only tiles on or above the diagonal are
// computed, each is stored together with its transpose, and the diagonal is exactly 0.
// batch > 1: `batch` independent problems stacked along the rows of A, B, na, nb and dist
// [batch*M x N] (needs M, N multiples of 128), scale record of problem b at scal + b * scal_stride.
int tc_gram(TcPlanes A, TcPlanes B, int64_t M, int64_t N, int64_t D, const float* na,
            const float* nb, const float* scal, int symmetric, float* dist, int64_t ldd,
            cudaStream_t st, int64_t batch = 1, int64_t scal_stride = 0);

// out[i,c] = colscale[c] * sum_j f(a[i,j]) V[j,c]   (kf3_tc.cu): a = fp32 [M x K] (row pitch lda, a
// multiple of 4; 16-byte aligned), f = exp(state->gamma * .) (SIGSVGD_KF_EXP) or the identity
// (SIGSVGD_KF_IDENTITY); V planes [K x >=C] as for tc_matmul_nn.  stats != 0: also state->sum_kd +=
// sum f(a) a and the entries equal to state->median are recorded (global row = row_offset + i).
// batch > 1: `batch` independent problems stacked along the rows of a [batch*M x K], V [batch*K x C]
// and out [batch*M x C] (needs M % 128 == 0, K % 32 == 0), state[b] and colscale + b * colscale_stride
// per problem.
int tc_kf_fused(const float* a, int64_t lda, int64_t M, int64_t K, int mode, sigsvgd_bw_state* state,
                int stats, int64_t row_offset, TcPlanes V, int64_t C, const float* colscale, float* out,
                int64_t ldo, cudaStream_t st, bool out_zeroed = false, int64_t batch = 1,
                int64_t colscale_stride = 0);

// true when this process can create TMA descriptors (driver entry point resolved)
bool tc_available();

// ---- single-pass exact median (median_window.cu) ---------------------------------------------
// ws: window_workspace_bytes(M, N) bytes = [SelWindow | sample | histograms | candidates]
size_t window_workspace_bytes(int64_t M, int64_t N, unsigned int* cap_out);
void window_pointers(void* ws, SelWindow** win, float** cand, unsigned int** pool_cnt);
// sample the matrix (a, b: centred fp32 features of its rows / columns), set the window, reset
// the counters and the selection fields of `state`
int launch_window_prepare(const float* a, int64_t lda, int64_t M, const float* b, int64_t ldb,
                          int64_t N, int64_t D, void* ws, sigsvgd_bw_state* state, SelWindow** win_out,
                          float** cand_out, cudaStream_t st);
// after tc_gram has filled dist + candidates: exact lower median -> state->{median,h,gamma,clamped}
int launch_window_select(void* ws, const float* dist, int64_t M, int64_t N, int64_t ldd,
                         int64_t n_log, float bw_scale, float tol, sigsvgd_bw_state* state,
                         cudaStream_t st, int symmetric = 0);

// ---- operand preparation (planes.cu) ---------------------------------------------------------
// per problem b: tcs[4b..] = {alpha_S, alpha_S^-2, alpha_G, 0}; colscale[b*Cp + c]; mx: 2 uints
int launch_tc_scales(const float* s, int64_t lds, const float* mean, int64_t N, int64_t D,
                     const float* score, int64_t lsc, int64_t P, int64_t Cp, unsigned int* mx,
                     float* tcs, float* colscale, int64_t batch, cudaStream_t st);
// F = [ (s - mean)*alpha_S | 1 | score*alpha_G | 0.. ] -> planes fh/fl [rows x ldp];
// norms[row] = |F_S|^2 (plane units, may be NULL); sc (optional) = centred fp32 features
int launch_tc_pack(const float* score, int64_t lsc, int64_t P, const float* s, int64_t lds,
                   const float* mean, int64_t N, int64_t D, const float* tcs, __half* fh, __half* fl,
                   int64_t ldp, int64_t Cp, float* norms, float* sc, int64_t ldsc, int64_t batch,
                   cudaStream_t st);
constexpr int kStatSlots = 8;  // `stats` buffers hold kStatSlots records of D + 2 floats (planes.cu)
// fused front end: statistics pass (column sums, max |s|, max |score| -> stats[D + 2]) ...
int launch_tc_stats(const float* s, int64_t lds, int64_t N, int64_t D, const float* score,
                    int64_t lsc, int64_t P, float* stats, int zero_first, cudaStream_t st,
                    int do_s = 1);  // do_s = 0: only max |score| (score == NULL: only the s part)
// ... and packing pass deriving mean (= colsum / n_total, 0 if !center) and scales from `stats`;
// tcs[4], colscale[Cp], planes, norms, centred features (sc) and the mean (mean_out) are outputs
int launch_tc_pack_stats(const float* score, int64_t lsc, int64_t P, const float* s, int64_t lds,
                         const float* stats, int64_t n_total, int center, int64_t N, int64_t D,
                         float* tcs, float* colscale, __half* fh, __half* fl, int64_t ldp, int64_t Cp,
                         float* norms, float* sc, int64_t ldsc, float* mean_out, cudaStream_t st,
                         int64_t c_lo = 0, int64_t c_hi = -1,  // plane columns [c_lo, c_hi) only
                         float* zero_out = nullptr, int64_t ld_zero = 0);  // also zero [N, Cp] of this buffer

}  // namespace sigsvgd
