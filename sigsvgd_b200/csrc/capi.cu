// extern "C" entry points of libsigsvgd_b200.so (declared in include/sigsvgd_b200.h) and the
// fused single-GPU phi pipeline.
#include <stdarg.h>
#include <string.h>

#include <mutex>

#include "common.cuh"
#include "gemm_tc.cuh"

namespace sigsvgd {

static_assert(kStatSlots == 8, "sig3_fwd_kernel (sig_kernels.cu) spreads its statistics over 8 slots");

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* get_error() { return g_err; }

int num_sms() {
  static thread_local int cached_dev = -1, cached = 148;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev != cached_dev) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0)
      cached = v;
    cached_dev = dev;
  }
  return cached;
}

// Per-device side stream + fork/join events: independent pieces of the fused update (the median
// window preparation) run beside the main pipeline.  Event record / wait are asynchronous and
// CUDA-graph capturable (the side stream joins a capture through the fork event).
struct SideStream {
  cudaStream_t stream = nullptr;
  cudaEvent_t fork = nullptr, join = nullptr;
};
static int side_stream(SideStream** out) {
  static SideStream table[64];
  static std::mutex mu;
  int dev = 0;
  SIG_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) {
    set_error("device index %d out of range", dev);
    return SIGSVGD_ERR_UNSUPPORTED;
  }
  std::lock_guard<std::mutex> lock(mu);
  SideStream& s = table[dev];
  if (!s.stream) {
    // highest priority: the side work (the median's sampling pass) is small and latency-bound; it should
    // take its SMs as soon as it is ready instead of queueing behind the main stream's wide kernels and
    // then running beside -- and slowing -- the Gram
    int prio_lo = 0, prio_hi = 0;
    SIG_CUDA_OK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    SIG_CUDA_OK(cudaStreamCreateWithPriority(&s.stream, cudaStreamNonBlocking, prio_hi));
    SIG_CUDA_OK(cudaEventCreateWithFlags(&s.fork, cudaEventDisableTiming));
    SIG_CUDA_OK(cudaEventCreateWithFlags(&s.join, cudaEventDisableTiming));
  }
  *out = &s;
  return 0;
}

// Host-input pipeline (sigsvgd_phi_host): the score's H2D copy runs on its own stream beside the
// front half of the update (which needs only X), and phi leaves in row chunks whose D2H copies
// overlap the signature backward of the next chunk.
constexpr int kHostChunks = 4;
struct HostPipe {
  cudaStream_t h2d = nullptr, d2h = nullptr;
  cudaEvent_t start = nullptr, score_ready = nullptr, done = nullptr;
  cudaEvent_t chunk[kHostChunks] = {};
  float* phi_host = nullptr;  // per call
};
static int host_pipe(HostPipe** out) {
  static HostPipe table[64];
  static std::mutex mu;
  int dev = 0;
  SIG_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) {
    set_error("device index %d out of range", dev);
    return SIGSVGD_ERR_UNSUPPORTED;
  }
  std::lock_guard<std::mutex> lock(mu);
  HostPipe& h = table[dev];
  if (!h.h2d) {
    SIG_CUDA_OK(cudaStreamCreateWithFlags(&h.h2d, cudaStreamNonBlocking));
    SIG_CUDA_OK(cudaStreamCreateWithFlags(&h.d2h, cudaStreamNonBlocking));
    SIG_CUDA_OK(cudaEventCreateWithFlags(&h.start, cudaEventDisableTiming));
    SIG_CUDA_OK(cudaEventCreateWithFlags(&h.score_ready, cudaEventDisableTiming));
    SIG_CUDA_OK(cudaEventCreateWithFlags(&h.done, cudaEventDisableTiming));
    for (int i = 0; i < kHostChunks; ++i)
      SIG_CUDA_OK(cudaEventCreateWithFlags(&h.chunk[i], cudaEventDisableTiming));
  }
  *out = &h;
  return 0;
}

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

struct PhiLayout {
  int64_t D, P, C, ldk;
  size_t off_sig, off_v, off_k, off_out, off_g, off_mean, off_part, off_sel, off_pref, total;
  // tensor-core path (B == 1): fp16 planes of F [N x C], norms, scales, median window
  size_t off_fh, off_fl, off_norms, off_scales, off_win, off_stats;
};

static size_t bump(size_t* cur, size_t bytes) {
  const size_t o = *cur;
  *cur = (size_t)round_up((int64_t)(o + bytes), 256);
  return o;
}

// batches of independent scenes take the tcgen05 kernels when no tile straddles two scenes
static bool tc_batch_shape(int64_t B, int64_t N) { return B > 1 && N >= 128 && N % 128 == 0; }

static PhiLayout phi_layout(int64_t B, int64_t N, int64_t T, int64_t d, int depth, bool want_sig,
                            bool want_k) {
  PhiLayout L;
  L.D = sig_channels(d, depth);
  L.P = T * d;
  // columns of V / F = row pitch of the planes: 64 fp16 = 128 bytes, so every 128-byte TMA row
  // segment is one aligned L2 line (misaligned rows double the L2 sector traffic)
  L.C = round_up(L.P + L.D + 1, 64);
  size_t cur = 0;
  L.off_sig = want_sig ? 0 : bump(&cur, sizeof(float) * B * N * L.D);
  L.off_v = bump(&cur, sizeof(float) * B * N * L.C);
  // squared distances.  One swarm: always in the workspace, rows padded to whole 128-byte lines (the
  // fused K.F kernel reads them by TMA; k_xx, if wanted, is a separate output).  Batches (SIMT): the
  // kernel matrix is formed in place, straight in k_out when the caller wants it.
  L.ldk = B == 1 ? round_up(N, 32) : N;
  L.off_k = (want_k && B != 1) ? 0 : bump(&cur, sizeof(float) * B * N * L.ldk);
  L.off_out = bump(&cur, sizeof(float) * B * N * L.C);
  L.off_g = bump(&cur, sizeof(float) * B * N * L.D);
  L.off_mean = bump(&cur, sizeof(float) * B * L.D);
  L.off_part = bump(&cur, sizeof(float) * B * 64 * L.D);
  L.off_sel = bump(&cur, select_workspace_bytes(N * N, B));
  // prefix checkpoints of the time-parallel signature backward (0 floats when not applicable)
  L.off_pref = bump(&cur, sizeof(float) * sig_prefix_floats(B * N, T, d, depth));
  L.off_fh = L.off_fl = L.off_norms = L.off_scales = L.off_win = L.off_stats = 0;
  if (B == 1 || tc_batch_shape(B, N)) {
    L.off_fh = bump(&cur, sizeof(__half) * B * N * L.C);
    L.off_fl = bump(&cur, sizeof(__half) * B * N * L.C);
    L.off_norms = bump(&cur, sizeof(float) * B * N);
    // per problem: {alpha_S, alpha_S^-2, alpha_G, 0} | 4 words of scratch ; then the colscale vectors
    L.off_scales = bump(&cur, sizeof(float) * B * (8 + L.C));
  }
  if (B == 1) {
    L.off_win = bump(&cur, window_workspace_bytes(N, N, nullptr));
    L.off_stats = bump(&cur, sizeof(float) * kStatSlots * (L.D + 2));
  }
  L.total = cur;
  return L;
}

// The tcgen05 path serves one swarm (B == 1) whose N^2 work dwarfs the fixed costs, and batches of
// scenes whose particle count is a multiple of the 128-row tile (one grouped Gram / K.F launch for
// all scenes); small swarms and odd-sized scenes stay on the exact-fp32 SIMT kernels.
static bool tc_path_enabled(int64_t B, int64_t N, uint32_t flags) {
  if (flags & SIGSVGD_FLAG_FORCE_SIMT) return false;
  if (!tc_available()) return false;
  if (B != 1) return tc_batch_shape(B, N);
  if (flags & SIGSVGD_FLAG_FORCE_TC) return true;
  return N >= 256;  // measured: 129 us (tcgen05) vs 175 us (SIMT) per update at N = 256
}

}  // namespace sigsvgd

using namespace sigsvgd;

extern "C" {

const char* sigsvgd_last_error(void) { return get_error(); }
const char* sigsvgd_version(void) { return "sigsvgd_b200 0.1 (sm_100a)"; }

int64_t sigsvgd_sig_channels(int64_t d, int depth) {
  if (d < 1 || depth < 1 || depth > 8) return -1;
  return sig_channels(d, depth);
}

int sigsvgd_signature_fwd(const float* x, int64_t N, int64_t T, int64_t d, int depth,
                          int basepoint, float* sig, void* stream) {
  SIG_CHECK_ARG(N == 0 || (x && sig), "signature_fwd: NULL pointer");
  return launch_sig_fwd(x, N, T, d, depth, basepoint, sig, (cudaStream_t)stream);
}

int sigsvgd_signature_bwd(const float* x, const float* sig, const float* gsig, int64_t N,
                          int64_t T, int64_t d, int depth, int basepoint, const float* add,
                          int64_t ld_add, float alpha, float beta, float* out, void* stream) {
  SIG_CHECK_ARG(N == 0 || (x && sig && gsig && out), "signature_bwd: NULL pointer");
  return launch_sig_bwd(x, sig, gsig, N, T, d, depth, basepoint, add, ld_add, alpha, beta, out,
                        (cudaStream_t)stream);
}

int64_t sigsvgd_signature_prefix_floats(int64_t N, int64_t T, int64_t d, int depth) {
  if (N < 0 || T < 1 || d < 1 || depth < 1 || depth > 8) return -1;
  return sig_prefix_floats(N, T, d, depth);
}

int sigsvgd_signature_fwd_prefix(const float* x, int64_t N, int64_t T, int64_t d, int depth,
                                 int basepoint, float* sig, float* pref, void* stream) {
  SIG_CHECK_ARG(N == 0 || (x && sig && pref), "signature_fwd_prefix: NULL pointer");
  return launch_sig_fwd(x, N, T, d, depth, basepoint, sig, (cudaStream_t)stream, nullptr, nullptr, 0, 0,
                        pref);
}

int sigsvgd_signature_bwd_prefix(const float* x, const float* sig, const float* pref,
                                 const float* gsig, int64_t N, int64_t T, int64_t d, int depth,
                                 int basepoint, const float* add, int64_t ld_add, float alpha,
                                 float beta, float* out, void* stream) {
  SIG_CHECK_ARG(N == 0 || (x && sig && pref && gsig && out), "signature_bwd_prefix: NULL pointer");
  return launch_sig_bwd(x, sig, gsig, N, T, d, depth, basepoint, add, ld_add, alpha, beta, out,
                        (cudaStream_t)stream, nullptr, pref);
}

int sigsvgd_sqdist(const float* a, int64_t lda, const float* b, int64_t ldb, int64_t M, int64_t N,
                   int64_t D, float* dist, int64_t ldd, uint32_t flags, void* stream) {
  SIG_CHECK_ARG((M == 0 || N == 0) || (a && b && dist), "sqdist: NULL pointer");
  if (flags & SIGSVGD_FLAG_FORCE_TC) {
    set_error("sqdist: the tcgen05 path takes fp16 planes; use sigsvgd_tc_gram");
    return SIGSVGD_ERR_UNSUPPORTED;
  }
  return launch_sqdist_simt(a, lda, b, ldb, M, N, D, dist, ldd, 1, 0, 0, 0, (cudaStream_t)stream);
}

size_t sigsvgd_select_workspace_bytes(int64_t count) { return select_workspace_bytes(count, 1); }

int sigsvgd_median_bandwidth(const float* dist, int64_t count, int64_t n_log, float bw_scale,
                             float tol, sigsvgd_bw_state* state, void* ws, size_t ws_bytes,
                             void* stream) {
  SIG_CHECK_ARG(dist && state && ws, "median_bandwidth: NULL pointer");
  return launch_median_select(dist, count, 1, 0, n_log, bw_scale, tol, state, ws, ws_bytes,
                              (cudaStream_t)stream);
}

int sigsvgd_median_bandwidth_static(const float* dist, int64_t count, int64_t batch, int64_t stride,
                                    int64_t n_log, float bw_scale, float tol, sigsvgd_bw_state* state,
                                    void* stream) {
  SIG_CHECK_ARG(dist && state, "median_bandwidth_static: NULL pointer");
  return launch_median_select_static(dist, count, batch, stride, n_log, bw_scale, tol, state,
                                     (cudaStream_t)stream);
}

int sigsvgd_fixed_bandwidth(float h, sigsvgd_bw_state* state, void* stream) {
  SIG_CHECK_ARG(state, "fixed_bandwidth: NULL pointer");
  return launch_fixed_bandwidth(h, state, 1, (cudaStream_t)stream);
}

int sigsvgd_select_begin(sigsvgd_bw_state* state, uint64_t* hist, uint64_t k, void* stream) {
  SIG_CHECK_ARG(state && hist, "select_begin: NULL pointer");
  return launch_select_begin(state, hist, k, 1, (cudaStream_t)stream);
}

int sigsvgd_select_begin_window(sigsvgd_bw_state* state, uint64_t* hist, uint64_t k_total,
                                const uint64_t* totals, int32_t* ok, void* stream) {
  SIG_CHECK_ARG(state && hist && totals && ok, "select_begin_window: NULL pointer");
  return launch_select_begin_window(state, hist, k_total, totals, ok, (cudaStream_t)stream);
}

int sigsvgd_select_hist(const float* dist, int64_t count, int pass, const sigsvgd_bw_state* state,
                        uint64_t* hist, void* stream) {
  SIG_CHECK_ARG(state && hist && (count == 0 || dist), "select_hist: NULL pointer");
  return launch_select_hist(dist, count, 1, 0, pass, state, hist, (cudaStream_t)stream);
}

int sigsvgd_select_scan(int pass, sigsvgd_bw_state* state, uint64_t* hist, void* stream) {
  SIG_CHECK_ARG(state && hist, "select_scan: NULL pointer");
  return launch_select_scan(pass, state, hist, 1, (cudaStream_t)stream);
}

int sigsvgd_select_finish(int64_t n_log, float bw_scale, float tol, sigsvgd_bw_state* state,
                          void* stream) {
  SIG_CHECK_ARG(state, "select_finish: NULL pointer");
  return launch_select_finish(n_log, bw_scale, tol, state, 1, (cudaStream_t)stream);
}

size_t sigsvgd_select_window_workspace_bytes(void) { return window_workspace_bytes(1, 1, nullptr); }

int sigsvgd_select_window(const float* feat, int64_t ld, int64_t N, int64_t D, void* ws,
                          size_t ws_bytes, sigsvgd_bw_state* state, void* stream) {
  SIG_CHECK_ARG(feat && ws && state && N >= 1 && D >= 1 && ld >= D, "select_window: bad argument");
  if (ws_bytes < window_workspace_bytes(1, 1, nullptr)) {
    set_error("select_window: workspace too small");
    return SIGSVGD_ERR_WORKSPACE;
  }
  // only the header of the window workspace is touched (bounds at its start; no candidate pool)
  SelWindow* win;
  float* cand;
  return launch_window_prepare(feat, ld, N, feat, ld, N, D, ws, state, &win, &cand,
                               (cudaStream_t)stream);
}

int64_t sigsvgd_select_collect_min_capacity(void) { return select_collect_blocks(); }

int sigsvgd_select_collect(const float* dist, int64_t count, const float* bounds, float* cand,
                           int64_t cand_cap, uint64_t* totals, void* stream) {
  SIG_CHECK_ARG(bounds && cand && totals && (count == 0 || dist), "select_collect: NULL pointer");
  return launch_select_collect(dist, count, bounds, cand, cand_cap, (unsigned long long*)totals,
                               (cudaStream_t)stream);
}

int sigsvgd_rbf_apply(float* k, int64_t ldk, int64_t M, int64_t N, int64_t row_offset,
                      const float* weight, int64_t ldw, sigsvgd_bw_state* state, void* stream) {
  SIG_CHECK_ARG(state && (M == 0 || k), "rbf_apply: NULL pointer");
  return launch_rbf_apply(k, ldk, M, N, row_offset, weight, ldw, state, 1, 0, 0,
                          (cudaStream_t)stream);
}

int sigsvgd_median_coef(sigsvgd_bw_state* state, const double* sum_kd_all,
                        const uint32_t* n_ties_all, int enable, void* stream) {
  SIG_CHECK_ARG(state, "median_coef: NULL pointer");
  return launch_median_coef(state, sum_kd_all, n_ties_all, enable, 1, (cudaStream_t)stream);
}

int sigsvgd_matmul_nn(const float* a, int64_t lda, const float* v, int64_t ldv, int64_t M,
                      int64_t N, int64_t C, float* out, int64_t ldo, uint32_t flags, void* stream) {
  SIG_CHECK_ARG(M == 0 || (a && v && out), "matmul_nn: NULL pointer");
  if (flags & SIGSVGD_FLAG_FORCE_TC) {
    set_error("matmul_nn: the tcgen05 path takes fp16 planes; use sigsvgd_tc_kf");
    return SIGSVGD_ERR_UNSUPPORTED;
  }
  return launch_gemm_nn_simt(a, lda, v, ldv, M, N, C, out, ldo, 1, 0, 0, 0, (cudaStream_t)stream);
}

int sigsvgd_col_mean(const float* s, int64_t lds, int64_t N, int64_t D, float* part, float* mean,
                     void* stream) {
  SIG_CHECK_ARG(s && part && mean, "col_mean: NULL pointer");
  return launch_col_mean(s, lds, N, D, part, mean, 1, 0, (cudaStream_t)stream);
}

int sigsvgd_pack_v(const float* score, int64_t P, const float* s, int64_t lds, const float* mean,
                   int64_t N, int64_t D, float* v, int64_t ldv, int64_t C, void* stream) {
  SIG_CHECK_ARG(s && v, "pack_v: NULL pointer");
  return launch_pack_v(score, P, s, lds, mean, N, D, v, ldv, C, 1, (cudaStream_t)stream);
}

int sigsvgd_sig_grad(const float* sc, int64_t lds, const float* sc_all, int64_t lds_all,
                     const float* ks, int64_t ldks, const float* rowsum, int64_t ldr, int64_t M,
                     int64_t D, int64_t row_offset, const sigsvgd_bw_state* state, float* g,
                     int64_t ldg, void* stream) {
  SIG_CHECK_ARG(state && (M == 0 || (sc && sc_all && ks && rowsum && g)), "sig_grad: NULL pointer");
  return launch_sig_grad(sc, lds, sc_all, lds_all, ks, ldks, rowsum, ldr, M, D, 0, row_offset,
                         state, g, ldg, 1, (cudaStream_t)stream);
}

int sigsvgd_tc_available(void) { return tc_available() ? 1 : 0; }

int sigsvgd_tc_scales(const float* s, int64_t lds, const float* mean, int64_t N, int64_t D,
                      const float* score, int64_t ld_score, int64_t P, int64_t Cp, float* scales,
                      void* stream) {
  SIG_CHECK_ARG(s && scales, "tc_scales: NULL pointer");
  return launch_tc_scales(s, lds, mean, N, D, score, ld_score, P, Cp, (unsigned int*)(scales + 4),
                          scales, scales + 8, 1, (cudaStream_t)stream);
}

int sigsvgd_tc_pack(const float* score, int64_t ld_score, int64_t P, const float* s, int64_t lds,
                    const float* mean, int64_t N, int64_t D, const float* scales, void* f_hi,
                    void* f_lo, int64_t ldp, int64_t Cp, float* norms, float* s_centred,
                    int64_t ld_sc, void* stream) {
  SIG_CHECK_ARG(s && scales && f_hi && f_lo, "tc_pack: NULL pointer");
  SIG_CHECK_ARG(aligned16(f_hi) && aligned16(f_lo), "tc_pack: planes must be 16-byte aligned");
  return launch_tc_pack(score, ld_score, P, s, lds, mean, N, D, scales, (__half*)f_hi, (__half*)f_lo,
                        ldp, Cp, norms, s_centred, ld_sc, 1, (cudaStream_t)stream);
}

int sigsvgd_tc_pack_fused(const float* score, int64_t ld_score, int64_t P, const float* s, int64_t lds,
                          int64_t N, int64_t D, int center, float* scales, void* f_hi, void* f_lo,
                          int64_t ldp, int64_t Cp, float* norms, float* s_centred, int64_t ld_sc,
                          float* mean_out, float* stats, void* stream) {
  SIG_CHECK_ARG(s && scales && f_hi && f_lo && stats, "tc_pack_fused: NULL pointer");
  SIG_CHECK_ARG(aligned16(f_hi) && aligned16(f_lo), "tc_pack_fused: planes must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  SIG_PROPAGATE(launch_tc_stats(s, lds, N, D, score, ld_score, P, stats, 1, st));
  return launch_tc_pack_stats(score, ld_score, P, s, lds, stats, N, center, N, D, scales, scales + 8,
                              (__half*)f_hi, (__half*)f_lo, ldp, Cp, norms, s_centred, ld_sc, mean_out,
                              st);
}

/* The two passes of sigsvgd_tc_pack_fused as separate calls, for rows that are one shard of a larger
 * swarm: the statistics of every shard are combined by the caller (column sums add, maxima take the
 * max) into the 8-slot record, and the packing pass derives the mean (with n_total = all particles)
 * and the scales from it -- identical on every shard. */
int sigsvgd_tc_stats(const float* s, int64_t lds, int64_t N, int64_t D, const float* score,
                     int64_t ld_score, int64_t P, float* stats, void* stream) {
  SIG_CHECK_ARG(s && stats, "tc_stats: NULL pointer");
  return launch_tc_stats(s, lds, N, D, score, ld_score, P, stats, 1, (cudaStream_t)stream);
}

int sigsvgd_tc_pack_from_stats(const float* score, int64_t ld_score, int64_t P, const float* s,
                               int64_t lds, const float* stats, int64_t n_total, int64_t N, int64_t D,
                               int center, float* scales, void* f_hi, void* f_lo, int64_t ldp,
                               int64_t Cp, float* norms, float* s_centred, int64_t ld_sc,
                               float* mean_out, void* stream) {
  SIG_CHECK_ARG(s && scales && f_hi && f_lo && stats, "tc_pack_from_stats: NULL pointer");
  SIG_CHECK_ARG(aligned16(f_hi) && aligned16(f_lo), "tc_pack_from_stats: planes must be 16-byte aligned");
  SIG_CHECK_ARG(n_total >= N && N >= 0, "tc_pack_from_stats: n_total < N");
  return launch_tc_pack_stats(score, ld_score, P, s, lds, stats, n_total, center, N, D, scales, scales + 8,
                              (__half*)f_hi, (__half*)f_lo, ldp, Cp, norms, s_centred, ld_sc, mean_out,
                              (cudaStream_t)stream);
}

int sigsvgd_tc_gram(const void* a_hi, const void* a_lo, int64_t lda, const float* na,
                    const void* b_hi, const void* b_lo, int64_t ldb, const float* nb, int64_t M,
                    int64_t N, int64_t D, const float* scales, int symmetric, float* dist,
                    int64_t ldd, void* stream) {
  SIG_CHECK_ARG(a_hi && a_lo && b_hi && b_lo && na && nb && scales && dist, "tc_gram: NULL pointer");
  return tc_gram(TcPlanes{(const __half*)a_hi, (const __half*)a_lo, lda},
                 TcPlanes{(const __half*)b_hi, (const __half*)b_lo, ldb}, M, N, D, na, nb, scales + 1,
                 symmetric, dist, ldd, (cudaStream_t)stream);
}

size_t sigsvgd_tc_median_workspace_bytes(int64_t M, int64_t N) {
  if (M < 1 || N < 1) return 0;
  return window_workspace_bytes(M, N, nullptr);
}

int sigsvgd_tc_median_prepare(const float* a, int64_t lda, int64_t M, const float* b, int64_t ldb,
                              int64_t N, int64_t D, void* median_ws, size_t ws_bytes,
                              sigsvgd_bw_state* state, void* stream) {
  SIG_CHECK_ARG(a && b && median_ws && state, "tc_median_prepare: NULL pointer");
  if (ws_bytes < window_workspace_bytes(M, N, nullptr)) {
    set_error("tc_median_prepare: workspace too small");
    return SIGSVGD_ERR_WORKSPACE;
  }
  SelWindow* win;
  float* cand;
  return launch_window_prepare(a, lda, M, b, ldb, N, D, median_ws, state, &win, &cand,
                               (cudaStream_t)stream);
}

int sigsvgd_tc_median_select(void* median_ws, const float* dist, int64_t M, int64_t N, int64_t ldd,
                             int symmetric, int64_t n_log, float bw_scale, float tol,
                             sigsvgd_bw_state* state, void* stream) {
  SIG_CHECK_ARG(median_ws && dist && state, "tc_median_select: NULL pointer");
  return launch_window_select(median_ws, dist, M, N, ldd, n_log, bw_scale, tol, state,
                              (cudaStream_t)stream, symmetric);
}

int sigsvgd_tc_kf(const float* a, int64_t lda, int64_t M, int64_t K, int mode, sigsvgd_bw_state* state,
                  int stats, int64_t row_offset, const void* f_hi, const void* f_lo, int64_t ldf,
                  int64_t C, const float* scales, float* out, int64_t ldo, void* stream) {
  SIG_CHECK_ARG(a && f_hi && f_lo && scales && out, "tc_kf: NULL pointer");
  return tc_kf_fused(a, lda, M, K, mode, state, stats, row_offset,
                     TcPlanes{(const __half*)f_hi, (const __half*)f_lo, ldf}, C, scales + 8, out, ldo,
                     (cudaStream_t)stream);
}

int sigsvgd_rbf_out(const float* dist, int64_t ldd, int64_t M, int64_t N, const sigsvgd_bw_state* state,
                    float* k_out, int64_t ldk, void* stream) {
  SIG_CHECK_ARG(state && (M == 0 || (dist && k_out)), "rbf_out: NULL pointer");
  return launch_rbf_out(dist, ldd, M, N, state, k_out, ldk, (cudaStream_t)stream);
}

int sigsvgd_sigpde_bandwidth(const double* x, const double* y, int64_t A, int64_t B, int64_t M, int64_t N,
                             int64_t D, float bw_scale, float tol, sigsvgd_bw_state* state, uint64_t* hist,
                             void* stream) {
  return launch_sigpde_bandwidth(x, y, A, B, M, N, D, bw_scale, tol, state, hist, (cudaStream_t)stream);
}

int sigsvgd_sigpde_gram(const double* x, const double* y, int64_t A, int64_t B, int64_t M, int64_t N, int64_t D,
                        int order, int kind, double h, const sigsvgd_bw_state* state, double* k_out,
                        void* stream) {
  return launch_sigpde_fwd(x, y, A, B, M, N, D, order, kind, h, state, k_out, (cudaStream_t)stream);
}

size_t sigsvgd_sigpde_bwd_workspace_bytes(int64_t A, int64_t B, int64_t M, int64_t N, int64_t D, int order) {
  return sigpde_bwd_workspace_bytes(A, B, M, N, D, order);
}

int sigsvgd_sigpde_gram_bwd(const double* x, const double* y, int64_t A, int64_t B, int64_t M, int64_t N,
                            int64_t D, int order, int kind, double h, const sigsvgd_bw_state* state,
                            const double* gout, double* grad_x, void* ws, size_t ws_bytes, void* stream) {
  return launch_sigpde_bwd(x, y, A, B, M, N, D, order, kind, h, state, gout, grad_x, ws, ws_bytes,
                           (cudaStream_t)stream);
}

int sigsvgd_particle_rollout(const float* init_state, const float* actions, const float* mass, float mass_scalar,
                             int64_t R, int64_t H, float dt, float max_acc, float max_speed, const uint8_t* grid,
                             int64_t gx, int64_t gy, float inv_cell, float off_x, float off_y, int can_crash,
                             float* states, uint8_t* flags, void* stream) {
  return launch_particle_rollout(init_state, actions, mass, mass_scalar, R, H, dt, max_acc, max_speed, grid, gx, gy,
                                 inv_cell, off_x, off_y, can_crash, states, flags, (cudaStream_t)stream);
}

int sigsvgd_particle_rollout_bwd(const float* mass, float mass_scalar, int64_t R, int64_t H, float dt,
                                 const float* grad_states, const uint8_t* flags, float* grad_actions, void* stream) {
  return launch_particle_rollout_bwd(mass, mass_scalar, R, H, dt, grad_states, flags, grad_actions,
                                     (cudaStream_t)stream);
}

int sigsvgd_chain_fk(const float* xyz, const float* rot, const float* axis, const int* dof, int J, int n_dof,
                     const float* q, int64_t B, float* xs, void* stream) {
  return launch_chain_fk(xyz, rot, axis, dof, J, n_dof, q, B, xs, (cudaStream_t)stream);
}

int sigsvgd_chain_fk_bwd(const float* xyz, const float* rot, const float* axis, const int* dof, int J, int n_dof,
                         const float* q, int64_t B, const float* gxs, float* gq, void* stream) {
  return launch_chain_fk_bwd(xyz, rot, axis, dof, J, n_dof, q, B, gxs, gq, (cudaStream_t)stream);
}

int sigsvgd_mlp_profile_enable(int on) { return mlp_profile_enable(on); }

int sigsvgd_mlp_profile_cycles(unsigned long long* out8, int reset) { return mlp_profile_cycles(out8, reset); }

size_t sigsvgd_mlp_packed_bytes(int in_dim, int hidden, int n_mid) { return mlp_packed_bytes(in_dim, hidden, n_mid); }

int sigsvgd_mlp_pack(const float* w_first, const float* b_first, const float* w_mid, const float* b_mid,
                     const float* w_last, const float* b_last, int in_dim, int hidden, int n_mid, void* packed,
                     size_t packed_bytes, void* stream) {
  return launch_mlp_pack(w_first, b_first, w_mid, b_mid, w_last, b_last, in_dim, hidden, n_mid, packed, packed_bytes,
                         (cudaStream_t)stream);
}

int sigsvgd_mlp_query(const void* packed, int in_dim, int hidden, int n_mid, int precise, const float* x, int64_t P,
                      float* prob, float* grad, void* stream) {
  return launch_mlp_query(packed, in_dim, hidden, n_mid, precise, x, P, prob, grad, (cudaStream_t)stream);
}

int sigsvgd_opt_step(float* x, const float* dir, float sign, float* m, float* v, float* grad_out, int64_t n,
                     int kind, float lr, double beta1, double beta2, float eps, float c1, float c2, int first,
                     void* stream) {
  SIG_CHECK_ARG(n == 0 || (x && dir), "opt_step: NULL pointer");
  return launch_opt_step(x, dir, sign, m, v, grad_out, n, kind, lr, beta1, beta2, eps, c1, c2, first,
                         (cudaStream_t)stream);
}

size_t sigsvgd_phi_workspace_bytes(int64_t B, int64_t N, int64_t T, int64_t d, int depth,
                                   int want_k) {
  if (B < 1 || N < 1 || T < 1 || d < 1 || depth < 1 || depth > 8) return 0;
  // sized for the worst case: signatures kept in the workspace
  return phi_layout(B, N, T, d, depth, false, want_k != 0).total;
}

}  // extern "C"

// The fused update.  hp == nullptr: everything is device resident (sigsvgd_phi).  hp != nullptr:
// `score` is still in flight on hp->h2d (event hp->score_ready) and phi is also copied to
// hp->phi_host chunk by chunk (sigsvgd_phi_host).
static int phi_pipeline(const float* x, const float* score, int64_t B, int64_t N, int64_t T, int64_t d,
                        int depth, int bw_mode, float h_fixed, float bw_scale, float tol, uint32_t flags,
                        float* phi, float* k_out, float* gradk_out, float* sig_out,
                        sigsvgd_bw_state* state, void* ws, size_t ws_bytes, void* stream,
                        HostPipe* hp) {
  cudaStream_t st = (cudaStream_t)stream;
  SIG_CHECK_ARG(B >= 1 && N >= 1 && T >= 1 && d >= 1 && depth >= 1 && depth <= 8,
                "phi: bad shape B=%lld N=%lld T=%lld d=%lld depth=%d", (long long)B, (long long)N,
                (long long)T, (long long)d, depth);
  SIG_CHECK_ARG(x && score && phi && state && ws, "phi: NULL pointer");
  SIG_CHECK_ARG(bw_mode == SIGSVGD_BW_MEDIAN || bw_mode == SIGSVGD_BW_FIXED, "phi: bad bw_mode %d",
                bw_mode);
  SIG_CHECK_ARG(aligned16(x) && aligned16(score) && aligned16(phi) && aligned16(ws) &&
                    aligned16(k_out) && aligned16(gradk_out) && aligned16(sig_out),
                "phi: pointers must be 16-byte aligned");
  SIG_CHECK_ARG(N <= (int64_t)1 << 30, "phi: N too large");
  const PhiLayout L = phi_layout(B, N, T, d, depth, sig_out != nullptr, k_out != nullptr);
  if (ws_bytes < L.total) {
    set_error("phi: workspace too small (%zu < %zu bytes)", ws_bytes, L.total);
    return SIGSVGD_ERR_WORKSPACE;
  }
  char* w = (char*)ws;
  float* sig = sig_out ? sig_out : (float*)(w + L.off_sig);
  float* V = (float*)(w + L.off_v);
  const int64_t ldk = L.ldk;
  float* out = (float*)(w + L.off_out);
  float* g = (float*)(w + L.off_g);
  float* mean = (float*)(w + L.off_mean);
  float* part = (float*)(w + L.off_part);
  void* selws = (void*)(w + L.off_sel);
  // time-parallel signature pair (prefix checkpoints) while the machine is short of warps; once every SM is
  // full the sequential kernels execute fewer instructions (measured, d = 7: backward 38.8 vs 43.0 us at
  // 4096 particles, 822 vs 770 us at 131072)
  float* pref = (sig_bwd_uses_prefix(T, d, depth) && B * N <= 16384) ? (float*)(w + L.off_pref) : nullptr;
  const int64_t D = L.D, P = L.P, C = L.C, R = B * N;
  const bool center = !(flags & SIGSVGD_FLAG_NO_CENTER);
  const bool use_tc = tc_path_enabled(B, N, flags);
  // squared distances: tcgen05 path -> workspace, pitch ldk (k_xx is a separate output); SIMT path ->
  // pitch N, turned into the kernel matrix in place, directly in k_out when the caller wants it
  float* K = (k_out && !(use_tc && B == 1)) ? k_out : (float*)(w + L.off_k);
  if ((flags & SIGSVGD_FLAG_FORCE_TC) && !use_tc) {
    set_error("phi: tcgen05 path not available for B=%lld N=%lld", (long long)B, (long long)N);
    return SIGSVGD_ERR_UNSUPPORTED;
  }
  __half* Fh = (__half*)(w + L.off_fh);
  __half* Fl = (__half*)(w + L.off_fl);
  float* norms = (float*)(w + L.off_norms);
  float* scales = (float*)(w + L.off_scales);

  // K1: signatures of all B*N particles; on the tensor-core path (depth 3) the same kernel also
  // accumulates the statistics the operand preparation needs (column sums, max |s|, max |score|)
  const bool fwd_stats = use_tc && B == 1 && sig_fwd_fuses_stats(d, depth);
  if (fwd_stats) {
    float* stats0 = (float*)(w + L.off_stats);
    SIG_CUDA_OK(cudaMemsetAsync(stats0, 0, sizeof(float) * kStatSlots * (D + 2), st));
    SIG_PROPAGATE(launch_sig_fwd(x, R, T, d, depth, 1, sig, st, stats0, hp ? nullptr : score, P, P, pref));
  } else {
    SIG_PROPAGATE(launch_sig_fwd(x, R, T, d, depth, 1, sig, st, nullptr, nullptr, 0, 0, pref));
  }
  // feature block (centred), its row pitch, and the [K.S | rowsum | K.score] blocks of `out`
  const float* feat;
  int64_t ldf;
  const float *out_ks, *out_rowsum, *out_kg;
  SideStream* side = nullptr;
  const bool median_tc = use_tc && B == 1 && bw_mode == SIGSVGD_BW_MEDIAN;
  const bool tc_batch = use_tc && B > 1;
  if (tc_batch) {
    // grouped path: every scene is one 128-row tile row of the stacked operands
    float* tcs = scales;
    unsigned int* mx = (unsigned int*)(scales + 4 * B);
    float* colscale = scales + 8 * B;
    if (center) SIG_PROPAGATE(launch_col_mean(sig, D, N, D, part, mean, B, N * D, st));
    SIG_PROPAGATE(launch_tc_scales(sig, D, center ? mean : nullptr, N, D, score, P, P, C, mx, tcs, colscale, B, st));
    SIG_PROPAGATE(launch_tc_pack(score, P, P, sig, D, center ? mean : nullptr, N, D, tcs, Fh, Fl, C, C, norms, V, D,
                                 B, st));
    feat = V, ldf = D;
    out_ks = out, out_rowsum = out + D, out_kg = out + D + 1;
    SIG_PROPAGATE(tc_gram(TcPlanes{Fh, Fl, C}, TcPlanes{Fh, Fl, C}, N, N, D, norms, norms, tcs + 1, 1, K, N, st, B,
                          4));
  } else if (use_tc) {
    if (median_tc) {
      // K4a on the side stream: sample the distance matrix straight from the raw signatures
      // (distances are translation invariant) and set the candidate window of the median
      SIG_PROPAGATE(side_stream(&side));
      SIG_CUDA_OK(cudaEventRecord(side->fork, st));
      SIG_CUDA_OK(cudaStreamWaitEvent(side->stream, side->fork, 0));
      SelWindow* win = nullptr;
      float* cand = nullptr;
      SIG_PROPAGATE(launch_window_prepare(sig, D, N, sig, D, N, D, w + L.off_win, state, &win, &cand,
                                          side->stream));
      SIG_CUDA_OK(cudaEventRecord(side->join, side->stream));
    }
    // statistics pass (column sums, maxima), then planes F = [S - mean | 1 | score] (fp16 hi/lo)
    // with the centred fp32 features going to the V buffer
    float* stats = (float*)(w + L.off_stats);
    // the packing pass also zeroes `out`, the target of the K.F reductions
    if (!hp) {
      if (!fwd_stats) SIG_PROPAGATE(launch_tc_stats(sig, D, N, D, score, P, P, stats, 1, st));
      SIG_PROPAGATE(launch_tc_pack_stats(score, P, P, sig, D, stats, N, center ? 1 : 0, N, D, scales,
                                         scales + 8, Fh, Fl, C, C, norms, V, D, mean, st, 0, -1, out, C));
    } else {
      // feature columns only; the score columns follow once its copy has landed
      if (!fwd_stats) SIG_PROPAGATE(launch_tc_stats(sig, D, N, D, nullptr, P, P, stats, 1, st));
      SIG_PROPAGATE(launch_tc_pack_stats(nullptr, P, P, sig, D, stats, N, center ? 1 : 0, N, D, scales,
                                         scales + 8, Fh, Fl, C, C, norms, V, D, mean, st, 0,
                                         round_up(D + 1, 2), out, C));
    }
    feat = V, ldf = D;
    out_ks = out, out_rowsum = out + D, out_kg = out + D + 1;
    // K3: squared distances (tcgen05, 3 MMAs per product; upper triangle mirrored)
    SIG_PROPAGATE(tc_gram(TcPlanes{Fh, Fl, C}, TcPlanes{Fh, Fl, C}, N, N, D, norms, norms, scales + 1, 1, K,
                          ldk, st));
    if (median_tc) SIG_CUDA_OK(cudaStreamWaitEvent(st, side->join, 0));
  } else {
    if (hp) SIG_CUDA_OK(cudaStreamWaitEvent(st, hp->score_ready, 0));
    if (center) SIG_PROPAGATE(launch_col_mean(sig, D, N, D, part, mean, B, N * D, st));
    SIG_PROPAGATE(launch_pack_v(score, P, sig, D, center ? mean : nullptr, N, D, V, C, C, B, st));
    feat = V + P, ldf = C;
    out_kg = out, out_ks = out + P, out_rowsum = out + P + D;
    SIG_PROPAGATE(launch_sqdist_simt(V + P, C, V + P, C, N, N, D, K, N, B, N * C, N * C, N * N, st));
  }
  // K4: bandwidth
  if (bw_mode == SIGSVGD_BW_MEDIAN && use_tc && B == 1) {
    // K3 wrote a bit-symmetric matrix with an exact-zero diagonal: the select reads the upper
    // triangle only
    SIG_PROPAGATE(launch_window_select(w + L.off_win, K, N, N, ldk, N, bw_scale, tol, state, st, 1));
  } else if (bw_mode == SIGSVGD_BW_MEDIAN) {
    SIG_PROPAGATE(launch_median_select(K, N * N, B, N * N, N, bw_scale, tol, state, selws,
                                       select_workspace_bytes(N * N, B), st));
  } else {
    SIG_PROPAGATE(launch_fixed_bandwidth(h_fixed, state, B, st));
  }
  // K5: one contraction exp(gamma d).[S | 1 | score]; on the tensor-core path the RBF map runs inside
  // the contraction's operand path and the kernel matrix is never stored
  const int median_grad =
      (bw_mode == SIGSVGD_BW_MEDIAN) && (flags & SIGSVGD_FLAG_MEDIAN_GRAD) ? 1 : 0;
  if (tc_batch) {
    SIG_PROPAGATE(tc_kf_fused(K, N, N, N, SIGSVGD_KF_EXP, state, /*stats=*/1, 0, TcPlanes{Fh, Fl, C}, D + 1 + P,
                              scales + 8 * B, out, C, st, /*out_zeroed=*/false, B, C));
    SIG_PROPAGATE(launch_median_coef(state, nullptr, nullptr, median_grad, B, st));
    if (k_out)  // the distances were formed in k_out: turn them into the kernel values in place
      for (int64_t b = 0; b < B; ++b)
        SIG_PROPAGATE(launch_rbf_out(K + b * N * N, N, N, N, state + b, K + b * N * N, N, st));
  } else if (use_tc) {
    if (hp) {
      float* stats = (float*)(w + L.off_stats);
      SIG_CUDA_OK(cudaStreamWaitEvent(st, hp->score_ready, 0));
      SIG_PROPAGATE(launch_tc_stats(sig, D, N, D, score, P, P, stats, 0, st, 0));
      SIG_PROPAGATE(launch_tc_pack_stats(score, P, P, sig, D, stats, N, center ? 1 : 0, N, D, scales,
                                         scales + 8, Fh, Fl, C, C, nullptr, nullptr, D, nullptr, st,
                                         (D + 1) & ~(int64_t)1, C));
    }
    // (Host-input pipeline: running the [Sc | 1] columns as a launch of their own, beside the score's
    // H2D copy, was measured and dropped: the front half ends just as the score lands, so nothing
    // overlaps, and two launches of the contraction cost 27 us more than one.)
    SIG_PROPAGATE(tc_kf_fused(K, ldk, N, N, SIGSVGD_KF_EXP, state, /*stats=*/1, 0, TcPlanes{Fh, Fl, C},
                              D + 1 + P, scales + 8, out, C, st, /*out_zeroed=*/true));
    // the coefficient of the median-gradient term needs sum(k*d), which the contraction has just
    // accumulated; at depth 3 the signature backward derives it itself while forming the cotangent
    // (SigGradSpec::coef_mode)
    if (!(B == 1 && sig_bwd_fuses_grad(d, depth)))
      SIG_PROPAGATE(launch_median_coef(state, nullptr, nullptr, median_grad, 1, st));
    if (k_out) SIG_PROPAGATE(launch_rbf_out(K, ldk, N, N, state, k_out, N, st));  // k_xx for the caller
  } else {
    SIG_PROPAGATE(launch_rbf_apply(K, N, N, N, 0, nullptr, 0, state, B, N * N, 0, st));
    SIG_PROPAGATE(launch_median_coef(state, nullptr, nullptr, median_grad, B, st));
    SIG_PROPAGATE(launch_gemm_nn_simt(K, N, V, C, N, N, C, out, C, B, N * N, N * C, N * C, st));
  }
  // sig-space gradient, then K2 with the fused epilogue phi = (K.score - grad_k)/N
  // (one swarm, depth 3: the backward kernel forms the cotangent itself while staging)
  // (batches of scenes: the same, each scene with its own record; scenes of a multiple of 8 particles)
  const bool fuse_grad = sig_bwd_fuses_grad(d, depth) && (B == 1 || N % 8 == 0);
  SigGradSpec gs{feat, ldf, feat, ldf, out_ks, C, out_rowsum, C, 0, state};
  gs.coef_mode = (use_tc && B == 1) ? (median_grad ? 1 : 2) : 0;
  gs.scene_rows = B > 1 ? N : 0;
  const SigGradSpec* spec = fuse_grad ? &gs : nullptr;
  if (!fuse_grad)
    SIG_PROPAGATE(launch_sig_grad(feat, ldf, feat, ldf, out_ks, C, out_rowsum, C, N, D, N, 0, state,
                                  g, D, B, st));
  if (gradk_out)
    SIG_PROPAGATE(launch_sig_bwd(x, sig, g, R, T, d, depth, 1, nullptr, 0, 0.f, 1.f, gradk_out, st, spec,
                                 pref));
  const float inv_n = 1.f / (float)N;
  if (!hp) {
    SIG_PROPAGATE(launch_sig_bwd(x, sig, g, R, T, d, depth, 1, out_kg, C, inv_n, -inv_n, phi, st, spec,
                                 pref));
    return 0;
  }
  // K2 in row chunks; each chunk's D2H copy overlaps the next chunk's backward sweep.  The copies are
  // the long pole (PCIe), so the first chunk is small -- phi starts to leave early -- and the later
  // ones grow (1/8, 1/4, 5/8 of the rows): each finishes before the copy in front of it does.
  const int64_t unit = round_up(ceil_div(R, 8), 32);
  const int64_t sizes[kHostChunks] = {unit, 2 * unit, 5 * unit, R};
  int n = 0;
  for (int64_t r0 = 0; r0 < R && n < kHostChunks; ++n) {
    const int64_t per = n + 1 < kHostChunks ? sizes[n] : R - r0;
    const int64_t rows = std::min(per, R - r0);
    SigGradSpec gc = gs;
    gc.sc += r0 * ldf, gc.ks += r0 * C, gc.rowsum += r0 * C, gc.row_offset = r0;
    SIG_PROPAGATE(launch_sig_bwd(x + r0 * T * d, sig + r0 * D, g + r0 * D, rows, T, d, depth, 1,
                                 out_kg + r0 * C, C, inv_n, -inv_n, phi + r0 * T * d, st,
                                 fuse_grad ? &gc : nullptr, pref ? pref + sig_prefix_floats(r0, T, d, depth) : nullptr));
    SIG_CUDA_OK(cudaEventRecord(hp->chunk[n], st));
    SIG_CUDA_OK(cudaStreamWaitEvent(hp->d2h, hp->chunk[n], 0));
    SIG_CUDA_OK(cudaMemcpyAsync(hp->phi_host + r0 * T * d, phi + r0 * T * d,
                                sizeof(float) * rows * T * d, cudaMemcpyDeviceToHost, hp->d2h));
    r0 += rows;
  }
  SIG_CUDA_OK(cudaEventRecord(hp->done, hp->d2h));
  SIG_CUDA_OK(cudaStreamWaitEvent(st, hp->done, 0));
  return 0;
}

// sigsvgd_phi / sigsvgd_phi_host fork work onto ONE side stream / copy-stream set per device and order it
// with one set of events: the record / wait pairs of a call must not interleave with another host
// thread's, so the enqueue phase of these two entry points is serialised per device.
static std::mutex& enqueue_mutex() {
  static std::mutex table[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) dev = 0;
  return table[dev];
}

extern "C" {

int sigsvgd_phi(const float* x, const float* score, int64_t B, int64_t N, int64_t T, int64_t d,
                int depth, int bw_mode, float h_fixed, float bw_scale, float tol, uint32_t flags,
                float* phi, float* k_out, float* gradk_out, float* sig_out,
                sigsvgd_bw_state* state, void* ws, size_t ws_bytes, void* stream) {
  std::lock_guard<std::mutex> lock(enqueue_mutex());
  return phi_pipeline(x, score, B, N, T, d, depth, bw_mode, h_fixed, bw_scale, tol, flags, phi, k_out,
                      gradk_out, sig_out, state, ws, ws_bytes, stream, nullptr);
}

size_t sigsvgd_phi_host_workspace_bytes(int64_t N, int64_t T, int64_t d, int depth) {
  if (N < 1 || T < 1 || d < 1 || depth < 1 || depth > 8) return 0;
  const size_t io = (size_t)round_up((int64_t)sizeof(float) * N * T * d, 256);
  return 3 * io + phi_layout(1, N, T, d, depth, false, false).total;
}

int sigsvgd_phi_host(const float* x_host, const float* score_host, int64_t N, int64_t T, int64_t d,
                     int depth, int bw_mode, float h_fixed, float bw_scale, float tol, uint32_t flags,
                     float* phi_host, sigsvgd_bw_state* state, void* ws, size_t ws_bytes,
                     void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  SIG_CHECK_ARG(N >= 1 && T >= 1 && d >= 1 && depth >= 1 && depth <= 8,
                "phi_host: bad shape N=%lld T=%lld d=%lld depth=%d", (long long)N, (long long)T,
                (long long)d, depth);
  SIG_CHECK_ARG(x_host && score_host && phi_host && state && ws, "phi_host: NULL pointer");
  SIG_CHECK_ARG(aligned16(ws), "phi_host: workspace must be 16-byte aligned");
  if (ws_bytes < sigsvgd_phi_host_workspace_bytes(N, T, d, depth)) {
    set_error("phi_host: workspace too small (%zu < %zu bytes)", ws_bytes,
              sigsvgd_phi_host_workspace_bytes(N, T, d, depth));
    return SIGSVGD_ERR_WORKSPACE;
  }
  std::lock_guard<std::mutex> lock(enqueue_mutex());
  HostPipe* hp = nullptr;
  SIG_PROPAGATE(host_pipe(&hp));
  const size_t bytes = sizeof(float) * (size_t)(N * T * d);
  const size_t io = (size_t)round_up((int64_t)bytes, 256);
  char* w = (char*)ws;
  float* x_dev = (float*)w;
  float* score_dev = (float*)(w + io);
  float* phi_dev = (float*)(w + 2 * io);
  hp->phi_host = phi_host;
  // the copy streams join the caller's stream here (earlier work on it may still use the buffers)
  // X first (the front half needs nothing else); the score's copy queues behind it on its own stream
  SIG_CUDA_OK(cudaMemcpyAsync(x_dev, x_host, bytes, cudaMemcpyHostToDevice, st));
  SIG_CUDA_OK(cudaEventRecord(hp->start, st));
  SIG_CUDA_OK(cudaStreamWaitEvent(hp->h2d, hp->start, 0));
  SIG_CUDA_OK(cudaStreamWaitEvent(hp->d2h, hp->start, 0));
  SIG_CUDA_OK(cudaMemcpyAsync(score_dev, score_host, bytes, cudaMemcpyHostToDevice, hp->h2d));
  SIG_CUDA_OK(cudaEventRecord(hp->score_ready, hp->h2d));
  return phi_pipeline(x_dev, score_dev, 1, N, T, d, depth, bw_mode, h_fixed, bw_scale, tol, flags,
                      phi_dev, nullptr, nullptr, nullptr, state, w + 3 * io, ws_bytes - 3 * io, stream,
                      hp);
}

}  // extern "C"
