/*
 * sigsvgd_b200.h -- C ABI of libsigsvgd_b200.so: the B200 (sm_100a) implementation of the
 * signature-kernel SVGD update of lubaroli/sigsvgd.
 *
 * The reference is pure Python; its FFI boundary for this path is the third-party `signatory`
 * torch extension plus torch library calls.  Each entry point below cites the reference
 * statement(s) it replaces (paths relative to the reference root).  A maintainer binds these
 * with ctypes; sigsvgd_b200/_lib.py is that binding.
 *
 * Conventions
 *  - all pointers are DEVICE pointers to caller-owned fp32 row-major buffers unless stated;
 *    16-byte aligned; inputs are never written; nothing is retained after the call returns;
 *  - every call is asynchronous on `stream` (a cudaStream_t passed as void*), performs no host
 *    synchronisation and no allocation, and is CUDA-graph capturable;
 *  - return value 0 = success; otherwise an SIGSVGD_ERR_* code, message via
 *    sigsvgd_last_error() (thread-local).  There is no CPU fallback: unsupported shapes fail.
 *  - host threads: every entry point may be called from several threads.  sigsvgd_phi and
 *    sigsvgd_phi_host share one internal side stream / copy-stream set (and its events) per device, so
 *    their ENQUEUE phase is serialised per device inside the library (a mutex held for the ~50 us of
 *    host work); the GPU work of concurrent calls on different caller streams is ordered correctly but
 *    passes through that one side stream.  sigsvgd_median_bandwidth_static shares its scratch (see there).
 */
#ifndef SIGSVGD_B200_H_
#define SIGSVGD_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SIGSVGD_OK 0
#define SIGSVGD_ERR_INVALID 1     /* bad shape / argument / alignment            */
#define SIGSVGD_ERR_UNSUPPORTED 2 /* valid but not implemented on this path      */
#define SIGSVGD_ERR_WORKSPACE 3   /* workspace too small                         */
#define SIGSVGD_ERR_CUDA 4        /* CUDA runtime error (launch / attribute)     */

/* Entries equal to the median that are RECORDED (the support of torch.median's backward).  A generic
 * swarm has 2 (the symmetric pair); duplicated / lattice-placed particles can have more.  If more than
 * SIGSVGD_MAX_TIES entries tie, n_ties / n_ties_all still hold the true count, every recorded entry
 * keeps the reference's weight coef / n_ties_all, and the entries beyond the cap (which ones is
 * unspecified) get no median-gradient term: a documented truncation of a term that is itself an
 * artefact of differentiating through the median; callers detect it by
 * n_ties_all > SIGSVGD_MAX_TIES. */
#define SIGSVGD_MAX_TIES 64

/* bandwidth modes (src/kernels/_kernels.py:106-109: h = get_bandwidth(sq_dists) | float(h)) */
#define SIGSVGD_BW_MEDIAN 0 /* src/utils/math.py:28-34 bw_median                        */
#define SIGSVGD_BW_FIXED 1  /* caller-supplied constant h                               */

/* flags for sigsvgd_phi / sigsvgd_gram */
#define SIGSVGD_FLAG_MEDIAN_GRAD 1u /* differentiate through the median (autograd parity)  */
#define SIGSVGD_FLAG_NO_CENTER 2u   /* do not mean-centre signature features before the Gram */
#define SIGSVGD_FLAG_FORCE_SIMT 4u  /* never take the tcgen05 path                          */
#define SIGSVGD_FLAG_FORCE_TC 8u    /* fail instead of falling back to SIMT                 */

/* A-operand map of sigsvgd_tc_kf */
#define SIGSVGD_KF_EXP 0      /* a = squared distances: K = exp(state->gamma * a)            */
#define SIGSVGD_KF_IDENTITY 1 /* a = kernel values already (e.g. a weighted kernel matrix)   */

/* Device-resident result/scratch record of one bandwidth selection (one per problem). */
typedef struct sigsvgd_bw_state {
  float median;        /* lower median of all squared distances (element (n-1)//2 of the sort) */
  float h;             /* bandwidth: max(tol, bw_scale*sqrt(median/ln(n_log+1)))  or fixed h   */
  float gamma;         /* -0.5 / h^2                                                           */
  float coef;          /* median-gradient coefficient  sum_kd / (h^2*median); 0 if inactive    */
  double sum_kd;       /* sum_ij K_ij d_ij over the rows given to sigsvgd_rbf_apply            */
  uint32_t n_ties;     /* entries with d == median seen by sigsvgd_rbf_apply (this rank)       */
  uint32_t n_ties_all; /* the same over all ranks (set by sigsvgd_median_coef)                 */
  uint32_t clamped;    /* 1 if h hit tol (no median gradient)                                  */
  uint32_t prefix;     /* radix-select scratch: bits of the median found so far                */
  uint64_t k_rem;      /* radix-select scratch: remaining rank                                 */
  int32_t ties[2 * SIGSVGD_MAX_TIES]; /* (row, col) global indices of the first ties           */
} sigsvgd_bw_state;

const char* sigsvgd_last_error(void);
const char* sigsvgd_version(void);

/* D = sum_{k=1..depth} d^k ; returns -1 on bad arguments. */
int64_t sigsvgd_sig_channels(int64_t d, int depth);

/* ---------------------------------------------------------------------------------------------
 * Signature transform.  Replaces signatory.signature(path, depth, basepoint=True)
 * (src/kernels/_traj_kernels.py:125-126).  x [N,T,d] -> sig [N,D]; levels 1..depth, level-major,
 * each level row-major, no scalar term; basepoint!=0 prepends a zero point.
 * ------------------------------------------------------------------------------------------- */
int sigsvgd_signature_fwd(const float* x, int64_t N, int64_t T, int64_t d, int depth,
                          int basepoint, float* sig, void* stream);

/* Vector-Jacobian product of the above (signatory's custom backward; reached from
 * torch.autograd.grad(k_xx.sum(), X) at src/inference/svgd.py:43 and score.py:55).
 *   out[N,T,d] = beta * (dS/dx)^T gsig  +  alpha * add        (add may be NULL; row stride ld_add)
 * With add = K.G rows, alpha = 1/N, beta = -1/N this is the Stein direction phi
 * (src/inference/svgd.py:82).  sig = forward output for the same x (reversible recompute). */
int sigsvgd_signature_bwd(const float* x, const float* sig, const float* gsig, int64_t N,
                          int64_t T, int64_t d, int depth, int basepoint, const float* add,
                          int64_t ld_add, float alpha, float beta, float* out, void* stream);

/* Time-parallel variant of the same pair (same reference statements: _traj_kernels.py:125-126 and
 * the backward reached from svgd.py:43).  The forward also writes levels 1-2 of the signature of each path up
 * to the end of time segments 0..2 of ceil(T/4) steps into pref[N][3][d+d*d]; the backward with that
 * buffer sweeps the 4 segments of every path at once (Chen: S_T = P_q (x) R_q), a 4x shorter
 * sequential chain.  Available when sigsvgd_signature_prefix_floats(N,T,d,depth) > 0 (depth 3,
 * d <= 8, T >= 8, T*d a multiple of 4); that many floats is the size of pref.  Results equal the
 * plain pair up to fp32 rounding. */
int64_t sigsvgd_signature_prefix_floats(int64_t N, int64_t T, int64_t d, int depth);
int sigsvgd_signature_fwd_prefix(const float* x, int64_t N, int64_t T, int64_t d, int depth,
                                 int basepoint, float* sig, float* pref, void* stream);
int sigsvgd_signature_bwd_prefix(const float* x, const float* sig, const float* pref,
                                 const float* gsig, int64_t N, int64_t T, int64_t d, int depth,
                                 int basepoint, const float* add, int64_t ld_add, float alpha,
                                 float beta, float* out, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Pairwise squared distances.  Replaces pw_dist_sq (src/utils/math.py:69-86):
 *   dist[i,j] = max(0, |a_i|^2 + |b_j|^2 - 2 a_i.b_j),  a [M,D] (row stride lda), b [N,D] (ldb).
 * ------------------------------------------------------------------------------------------- */
int sigsvgd_sqdist(const float* a, int64_t lda, const float* b, int64_t ldb, int64_t M,
                   int64_t N, int64_t D, float* dist, int64_t ldd, uint32_t flags, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Median-heuristic bandwidth.  Replaces bw_median (src/utils/math.py:28-34) incl. torch.median's
 * lower-median convention.  Exact order statistic of the `count` floats (all >= 0) at `dist`.
 * Single-GPU form: selects rank (count-1)/2 and fills state->{median,h,gamma,clamped}.
 * n_log = the N of ln(N+1) (sq_dists.shape[0]).  ws from sigsvgd_select_workspace_bytes(count).
 * ------------------------------------------------------------------------------------------- */
size_t sigsvgd_select_workspace_bytes(int64_t count);
int sigsvgd_median_bandwidth(const float* dist, int64_t count, int64_t n_log, float bw_scale,
                             float tol, sigsvgd_bw_state* state, void* ws, size_t ws_bytes,
                             void* stream);
/* The same exact median for callers that may not allocate (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc): NO counterpart in the reference -- added at the
 * request of the operators of the run that produced this library.  The 2048-bin u64 histograms that
 * sigsvgd_median_bandwidth takes as `ws` live in a file-scope device array inside the library
 * (SIGSVGD_STATIC_MAX_BATCH x 16 KiB = 256 KiB, held by every GPU that loads the library until exit);
 * the call allocates nothing (no cudaMalloc / cudaFree / allocating library routine).
 *   dist: `batch` arrays of `count` floats >= 0, array b at dist + b * stride (stride ignored for
 *   batch 1); state: `batch` records; each record gets the exact lower median (rank (count-1)/2) of its
 *   array and h, gamma, clamped as sigsvgd_median_bandwidth fills them.
 * batch > SIGSVGD_STATIC_MAX_BATCH -> SIGSVGD_ERR_WORKSPACE (use the workspace form).  Every call on a
 * device shares the one array (its bins are re-zeroed at the start of each call): calls that overlap
 * on the GPU -- different streams, or graphs replayed concurrently -- RACE; serialise them. */
#define SIGSVGD_STATIC_MAX_BATCH 16
int sigsvgd_median_bandwidth_static(const float* dist, int64_t count, int64_t batch, int64_t stride,
                                    int64_t n_log, float bw_scale, float tol, sigsvgd_bw_state* state,
                                    void* stream);
/* Fixed bandwidth: fills state for a caller-chosen h (bandwidth_fn=lambda _: const). */
int sigsvgd_fixed_bandwidth(float h, sigsvgd_bw_state* state, void* stream);

/* Multi-rank building blocks (row-block sharding; the 2048-bin histogram is summed across
 * ranks by the caller, e.g. torch.distributed.all_reduce, between _hist and _scan):
 *   k = rank of the wanted order statistic among ALL ranks' elements, (total-1)/2 for the median;
 *   pass p in {0,1,2} refines 11/11/10 bits of the fp32 pattern;
 *   hist: uint64[2048], zeroed by _begin and by every _scan. */
int sigsvgd_select_begin(sigsvgd_bw_state* state, uint64_t* hist, uint64_t k, void* stream);
int sigsvgd_select_hist(const float* dist, int64_t count, int pass, const sigsvgd_bw_state* state,
                        uint64_t* hist, void* stream);
int sigsvgd_select_scan(int pass, sigsvgd_bw_state* state, uint64_t* hist, void* stream);
/* Windowed first pass for the sharded median: `sigsvgd_select_window` draws a fixed pseudo-random
 * sample of pair distances from the centred features of ALL particles (identical input on every
 * rank -> identical result) and writes bounds[0..1] = [lo, hi] bracketing the median at the start
 * of `ws`; `sigsvgd_select_collect` reads one strip once, counts the entries below lo and compacts
 * the entries inside [lo, hi] into `cand` (cand_cap floats; unused slots are padded with a value
 * above every candidate).  totals[3] (device, zeroed by the call) += {#below, #kept, #overflowed
 * pools}.  After summing totals over the ranks the caller runs select_begin/hist/scan/finish on
 * `cand` with k - #below, or falls back to the strip if the window missed or a pool overflowed. */
size_t sigsvgd_select_window_workspace_bytes(void);
int sigsvgd_select_window(const float* feat, int64_t ld, int64_t N, int64_t D, void* ws,
                          size_t ws_bytes, sigsvgd_bw_state* state, void* stream);
int64_t sigsvgd_select_collect_min_capacity(void);
int sigsvgd_select_collect(const float* dist, int64_t count, const float* bounds, float* cand,
                           int64_t cand_cap, uint64_t* totals, void* stream);
/* select_begin for a select over the candidates, with the bookkeeping kept on the device (no host
 * read of totals): rank = k_total - totals[0]; *ok (device int32) = 1 if the window holds the rank and
 * no pool overflowed, else 0 (the caller then repeats the select over the full strips). */
int sigsvgd_select_begin_window(sigsvgd_bw_state* state, uint64_t* hist, uint64_t k_total,
                                const uint64_t* totals, int32_t* ok, void* stream);
int sigsvgd_select_finish(int64_t n_log, float bw_scale, float tol, sigsvgd_bw_state* state,
                          void* stream);

/* ---------------------------------------------------------------------------------------------
 * RBF map, in place.  Replaces K = exp(-0.5/h^2 * sq_dists) (src/kernels/_kernels.py:110).
 *   k[i,j] = exp(state->gamma * k[i,j]) for the M x N strip (row stride ldk);
 * accumulates state->sum_kd and records entries equal to state->median in state->ties
 * (global row index = row_offset + i).  weight (optional, [M,N], row stride ldw): the stored value
 * and the sum_kd statistic become w[i,j]*K[i,j] (generic-cotangent backward).
 * ------------------------------------------------------------------------------------------- */
int sigsvgd_rbf_apply(float* k, int64_t ldk, int64_t M, int64_t N, int64_t row_offset,
                      const float* weight, int64_t ldw, sigsvgd_bw_state* state, void* stream);

/* Median-gradient coefficient (the d h / d median factor of torch.median's backward):
 *   state->coef = sum_kd / (h^2 * median)  if enable && !clamped && median > 0, else 0.
 * sum_kd_all / n_ties_all: optional DEVICE scalars holding the all-rank totals (NULL = use the
 * values accumulated by sigsvgd_rbf_apply on this rank). */
int sigsvgd_median_coef(sigsvgd_bw_state* state, const double* sum_kd_all,
                        const uint32_t* n_ties_all, int enable, void* stream);

/* out[M,C] = a[M,N] . v[N,C]  -- replaces k_xx @ score (src/inference/svgd.py:82) and the K.S /
 * row-sum contractions inside autograd.grad(k_xx.sum(), X) (svgd.py:43). */
int sigsvgd_matmul_nn(const float* a, int64_t lda, const float* v, int64_t ldv, int64_t M,
                      int64_t N, int64_t C, float* out, int64_t ldo, uint32_t flags, void* stream);

/* Column means of s [N,D] (row stride lds) -> mean [D].  part: scratch of 64*D floats. */
int sigsvgd_col_mean(const float* s, int64_t lds, int64_t N, int64_t D, float* part, float* mean,
                     void* stream);
/* v[N,C] (row stride ldv) = [ score[N,P] | s - mean [N,D] | 1 | 0.. ],  C >= P + D + 1.
 * score may be NULL when P == 0; mean may be NULL (no centring). */
int sigsvgd_pack_v(const float* score, int64_t P, const float* s, int64_t lds, const float* mean,
                   int64_t N, int64_t D, float* v, int64_t ldv, int64_t C, void* stream);

/* Signature-space gradient of sum(K) wrt the first argument (closed form of the autograd route
 * src/inference/svgd.py:43 through _kernels.py:110, math.py:28-34 and math.py:81-86):
 *   g[i,:] = -(1/h^2) ( sc[i,:] * rowsum[i] - ks[i,:] )
 *            + sum_{(a,b) in ties, a == row_offset+i} coef/n_ties_all * ( sc[i,:] - sc_all[b,:] )
 * sc = centred features of the local rows [M,D] (ld lds), sc_all = those of all columns. */
int sigsvgd_sig_grad(const float* sc, int64_t lds, const float* sc_all, int64_t lds_all,
                     const float* ks, int64_t ldks, const float* rowsum, int64_t ldr, int64_t M,
                     int64_t D, int64_t row_offset, const sigsvgd_bw_state* state, float* g,
                     int64_t ldg, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Tensor-core building blocks (tcgen05 + TMA, split-fp16: every fp32 operand is two fp16 planes
 * hi/lo of value*alpha, a product is hi.hi + hi.lo + lo.hi accumulated in fp32 TMEM).  They replace
 * the same reference statements as sigsvgd_sqdist / sigsvgd_rbf_apply / sigsvgd_matmul_nn and are
 * what sigsvgd_phi runs for large swarms; exported so a row-block-sharded caller can compose them.
 *   scales : device buffer of 8 + Cp floats: {alpha_S, alpha_S^-2, alpha_G, 0, <4 words scratch>,
 *            colscale[Cp]}, filled by sigsvgd_tc_scales
 *   planes : fp16 row-major, row pitch (elements) a multiple of 8, 16-byte aligned
 *   F      : [ (S - mean)*alpha_S | 1 | score*alpha_G | 0.. ]   Cp = round_up(D + 1 + P, 8) columns
 * ------------------------------------------------------------------------------------------- */
int sigsvgd_tc_available(void);
int sigsvgd_tc_scales(const float* s, int64_t lds, const float* mean, int64_t N, int64_t D,
                      const float* score, int64_t ld_score, int64_t P, int64_t Cp, float* scales,
                      void* stream);
int sigsvgd_tc_pack(const float* score, int64_t ld_score, int64_t P, const float* s, int64_t lds,
                    const float* mean, int64_t N, int64_t D, const float* scales, void* f_hi,
                    void* f_lo, int64_t ldp, int64_t Cp, float* norms, float* s_centred,
                    int64_t ld_sc, void* stream);
/* tc_scales + tc_pack (and the column means) in two passes over the data instead of four: one
 * statistics pass (column sums, max |s|, max |score| into stats[8 * (D + 2)], scratch) and one packing pass that
 * derives the mean (center != 0) and the scales itself.  mean_out [D] / s_centred / norms may be NULL. */
int sigsvgd_tc_pack_fused(const float* score, int64_t ld_score, int64_t P, const float* s, int64_t lds,
                          int64_t N, int64_t D, int center, float* scales, void* f_hi, void* f_lo,
                          int64_t ldp, int64_t Cp, float* norms, float* s_centred, int64_t ld_sc,
                          float* mean_out, float* stats, void* stream);
/* The two passes of sigsvgd_tc_pack_fused as separate calls, for rows that are ONE SHARD of a larger
 * swarm (row-block sharded update): sigsvgd_tc_stats fills stats[8 * (D + 2)] (8 slots
 * of: D column sums, max |s|, max |score|; the slots add / max together) for this shard's N rows; the
 * caller combines the shards' records (sums add, maxima max) into one 8-slot record, and
 * sigsvgd_tc_pack_from_stats packs the shard's rows with the mean (n_total = particles of the whole
 * swarm) and the power-of-two scales derived from it -- bit-identical on every shard. */
int sigsvgd_tc_stats(const float* s, int64_t lds, int64_t N, int64_t D, const float* score,
                     int64_t ld_score, int64_t P, float* stats, void* stream);
int sigsvgd_tc_pack_from_stats(const float* score, int64_t ld_score, int64_t P, const float* s,
                               int64_t lds, const float* stats, int64_t n_total, int64_t N, int64_t D,
                               int center, float* scales, void* f_hi, void* f_lo, int64_t ldp,
                               int64_t Cp, float* norms, float* s_centred, int64_t ld_sc,
                               float* mean_out, void* stream);
/* dist[i,j] = max(0, |a_i|^2 + |b_j|^2 - 2 a_i.b_j) over the first D plane columns, in the
 * units of the unscaled features (math.py:69-86); na/nb = norms from sigsvgd_tc_pack.
 * 128x128 tiles with separate TMEM accumulators for the hi.hi and correction
 * products and a TMA-store epilogue (needs dist 16-byte aligned and ldd % 4 == 0, else a slow
 * per-element epilogue runs); symmetric != 0 (same planes on both sides, M == N) computes the
 * upper triangle and mirrors it: the result is bit-symmetric with an exactly zero diagonal. */
int sigsvgd_tc_gram(const void* a_hi, const void* a_lo, int64_t lda, const float* na,
                    const void* b_hi, const void* b_lo, int64_t ldb, const float* nb, int64_t M,
                    int64_t N, int64_t D, const float* scales, int symmetric, float* dist,
                    int64_t ldd, void* stream);
size_t sigsvgd_tc_median_workspace_bytes(int64_t M, int64_t N);
/* Single-pass exact median (math.py:28-34).  prepare: a/b = centred fp32 features of the rows /
 * columns (s_centred of sigsvgd_tc_pack); samples 8192 entries of the matrix, sets a candidate
 * window [lo, hi] around the sample median, resets the selection fields of *state.  It does not
 * need the matrix and may run before or after the Gram kernel. */
int sigsvgd_tc_median_prepare(const float* a, int64_t lda, int64_t M, const float* b, int64_t ldb,
                              int64_t N, int64_t D, void* median_ws, size_t ws_bytes,
                              sigsvgd_bw_state* state, void* stream);
/* select (after the matrix is written): one pass over dist counts the entries below the window
 * and collects those inside it, then the exact lower median of the M*N distances is selected
 * among the candidates -> state->{median,h,gamma,clamped}.  Falls back to radix passes over the
 * full matrix if the window missed, so the result is always exact. */
int sigsvgd_tc_median_select(void* median_ws, const float* dist, int64_t M, int64_t N, int64_t ldd,
                             int symmetric /* dist is bit-symmetric with an exact-zero diagonal
                                              (sigsvgd_tc_gram, symmetric = 1): only the upper
                                              triangle is read */,
                             int64_t n_log, float bw_scale, float tol, sigsvgd_bw_state* state,
                             void* stream);
/* out[M,C] = f(a)[M,K] . F[K,C] in unscaled units: [K.(S-mean) | rowsum(K) | K.score] (svgd.py:82, :43)
 * with the RBF map fused into the operand path: a = fp32 [M,K] (row pitch lda, a multiple of 4;
 * 16-byte aligned) holds the squared distances (mode SIGSVGD_KF_EXP: f = exp(state->gamma * .),
 * _kernels.py:110) or kernel values (SIGSVGD_KF_IDENTITY); the kernel matrix is never written.
 * stats != 0 (EXP mode): also accumulates state->sum_kd and records the entries equal to
 * state->median like sigsvgd_rbf_apply (global row index = row_offset + i).  F planes / scales as
 * produced by sigsvgd_tc_pack*.  out (row pitch ldo >= C) is overwritten. */
int sigsvgd_tc_kf(const float* a, int64_t lda, int64_t M, int64_t K, int mode, sigsvgd_bw_state* state,
                  int stats, int64_t row_offset, const void* f_hi, const void* f_lo, int64_t ldf,
                  int64_t C, const float* scales, float* out, int64_t ldo, void* stream);
/* k_out[M,N] (row pitch ldk) = exp(state->gamma * dist[M,N]) (_kernels.py:110): the kernel matrix for
 * callers that ask for k_xx; no statistics.  k_out may alias dist. */
int sigsvgd_rbf_out(const float* dist, int64_t ldd, int64_t M, int64_t N, const sigsvgd_bw_state* state,
                    float* k_out, int64_t ldk, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Signature-PDE (Goursat) kernel, SURVEY.md 8(f)-1.  Replaces
 *   sigkernel.SigKernel(static_kernel, dyadic_order).compute_Gram(X.double(), Y.double())
 * (third party sigkernel @ 3b2373982e12, absent; call sites src/kernels/_traj_kernels.py:198-206,
 * src/inference/trajectory_svgd.py:55-69, src/inference/score.py:64-76) with the reference's static kernel
 * BatchGaussianKernel.Gram_matrix = exp(-|x_s - y_t|^2 / h) (src/kernels/_traj_kernels.py:176-195; note
 * "/ h") or sigkernel.LinearKernel (<x_s, y_t> * h), and its backward w.r.t. the FIRST argument.
 * x [A,M,D], y [B,N,D] fp64 row-major; k_out [A,B] fp64.  order = dyadic refinement (2^order fine cells
 * per segment and side).  The bandwidth is state->h when state != NULL (RBF), else the argument h.
 * One warp per path pair with the pair's [M,N] static Gram in shared memory: shapes whose pair state
 * exceeds 200 KB (roughly M*N > 12000) return SIGSVGD_ERR_UNSUPPORTED.
 *
 * sigsvgd_sigpde_bandwidth: the reference's default bandwidth for that static kernel -- bw_median
 * (src/utils/math.py:28-34) over ALL A*B*M*N point-pair distances, h = sqrt(median / ln(A + 1)) -- by the
 * exact radix select with the distances recomputed on the fly (no [A,B,M,N] tensor); fills
 * state->{median,h,clamped}; hist: 2048 x 8 bytes of scratch.
 * sigsvgd_sigpde_gram_bwd: grad_x [A,M,D] = sum_b gout[a,b] d k[a,b] / d x^a by the package's formula
 * (reverse-direction solve, products of the two solutions summed per coarse cell, analytic derivative of
 * the increments at a fixed bandwidth); ws from sigsvgd_sigpde_bwd_workspace_bytes.
 * ------------------------------------------------------------------------------------------- */
#define SIGSVGD_PDE_RBF 0
#define SIGSVGD_PDE_LINEAR 1
int sigsvgd_sigpde_bandwidth(const double* x, const double* y, int64_t A, int64_t B, int64_t M, int64_t N,
                             int64_t D, float bw_scale, float tol, sigsvgd_bw_state* state, uint64_t* hist,
                             void* stream);
int sigsvgd_sigpde_gram(const double* x, const double* y, int64_t A, int64_t B, int64_t M, int64_t N, int64_t D,
                        int order, int kind, double h, const sigsvgd_bw_state* state, double* k_out,
                        void* stream);
size_t sigsvgd_sigpde_bwd_workspace_bytes(int64_t A, int64_t B, int64_t M, int64_t N, int64_t D, int order);
int sigsvgd_sigpde_gram_bwd(const double* x, const double* y, int64_t A, int64_t B, int64_t M, int64_t N,
                            int64_t D, int order, int kind, double h, const sigsvgd_bw_state* state,
                            const double* gout, double* grad_x, void* ws, size_t ws_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Stein-MPC rollout of the particle model, SURVEY.md 8(f)-3.  Replaces the Python loop of DuSt._rollout
 * (src/controllers/dust.py:293-300) over ParticleModel.step (src/models/particle.py:123-173; control_type
 * "acceleration", deterministic) and the autograd graph torch records through it:
 *   a = clamp(act / m, +-max_acc);  mask = can_crash ? grid[floor(x * inv_cell + off_x), floor(y * inv_cell +
 *   off_y)] (indices clamped; ObstacleMap.get_collisions, src/utils/obstacle_map.py:68-97) : 0;
 *   s' = s + ((vx, vy, ax, ay) * dt) * (1 - mask);  s'.v = clamp(s'.v, +-max_speed)
 * init_state [4]; actions [R,H,2]; mass [R] or NULL (then mass_scalar); grid [gx,gy] bytes (0 / 1) or NULL;
 * max_acc / max_speed may be +inf.  states [R,H+1,4] (bit-identical to the reference's fp32 CPU result: same
 * operation order, no fused multiply-adds); flags [R,H] bytes (crash / active clamps) for the backward.
 * sigsvgd_particle_rollout_bwd: grad_actions [R,H,2] for the cotangent grad_states [R,H+1,4] (the mask and
 * the clamps are piecewise constant, as in torch's autograd).
 * ------------------------------------------------------------------------------------------- */
int sigsvgd_particle_rollout(const float* init_state, const float* actions, const float* mass, float mass_scalar,
                             int64_t R, int64_t H, float dt, float max_acc, float max_speed, const uint8_t* grid,
                             int64_t gx, int64_t gy, float inv_cell, float off_x, float off_y, int can_crash,
                             float* states, uint8_t* flags, void* stream);
int sigsvgd_particle_rollout_bwd(const float* mass, float mass_scalar, int64_t R, int64_t H, float dt,
                                 const float* grad_states, const uint8_t* flags, float* grad_actions, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Forward kinematics of a serial chain, SURVEY.md 8(f)-4.  Replaces Robot.qs_to_joints_xs
 * (src/models/robot/robot_simulator.py:138-170: differentiable_robot_model's
 * compute_forward_kinematics_all_links, positions only) and autograd through it.
 * The chain is described in HOST memory, URDF semantics per joint j (parent -> child):
 *   T_child = T_parent . [rot_j | xyz_j] . Rot(axis_j, q[dof_j])       (dof_j = -1: fixed joint)
 * xyz [J,3]; rot [J,9] (R_rpy of the joint origin, row-major); axis [J,3] (unit); dof [J]; J, n_dof <= 16.
 * q [B,n_dof] -> xs [J,B,3]: the position of every child-link frame (the caller picks its target links).
 * sigsvgd_chain_fk_bwd: gq [B,n_dof] = d (sum gxs . xs) / d q, geometric Jacobian, nothing saved by the forward.
 * ------------------------------------------------------------------------------------------- */
int sigsvgd_chain_fk(const float* xyz, const float* rot, const float* axis, const int* dof, int J, int n_dof,
                     const float* q, int64_t B, float* xs, void* stream);
int sigsvgd_chain_fk_bwd(const float* xyz, const float* rot, const float* axis, const int* dof, int J, int n_dof,
                         const float* q, int64_t B, const float* gxs, float* gq, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Learned occupancy / self-collision networks of the robot-planning cost, SURVEY.md 8(f)-4.  Replaces the
 * forward of ContinuousOccupancyMap (src/models/robot_learning/continuous_occupancy_map.py:10-35) and
 * ContinuousSelfCollisionPredictor (continuous_self_collision_pred.py:11-36),
 *   Linear(in, H) ReLU  [Linear(H, H) ReLU] x n_mid  Linear(H, 1) Sigmoid        (H = 200, n_mid = 3)
 * as called by batch_cost_function (examples/script_planning_robot.py:300-318) on batch x timesteps x
 * body-point positions, AND the input gradient torch's autograd forms for it (:107): one kernel, the H x H
 * layers as tcgen05 GEMMs, activations never leave the SM.
 * precise != 0: operands as fp16 hi + lo planes, 3 MMAs per product: the reference's fp32 GEMMs to fp32 accuracy
 * (the ReLU patterns, hence the piecewise-constant Jacobians, agree with the reference's);  precise == 0: hi
 * planes only, 1 MMA: probabilities to ~5e-4, and about 1 point in 20 has a pre-activation within fp16 rounding
 * of zero and takes the neighbouring linear piece's gradient.
 * Limits: 1 <= in_dim <= 8, 1 <= hidden <= 208, 1 <= n_mid <= 4 (else SIGSVGD_ERR_INVALID / _UNSUPPORTED).
 * sigsvgd_mlp_pack: weights in torch's nn.Linear layout -- w_first [H,in], b_first [H], w_mid [n_mid,H,H]
 * (out, in), b_mid [n_mid,H], w_last [H], b_last [1] -> `packed` (sigsvgd_mlp_packed_bytes, 16-byte aligned);
 * repeat after the weights change.
 * sigsvgd_mlp_query: x [P,in] -> prob [P] and, unless grad == NULL, grad [P,in] = d prob_p / d x_p.
 * ------------------------------------------------------------------------------------------- */
size_t sigsvgd_mlp_packed_bytes(int in_dim, int hidden, int n_mid);
/* development aid, off by default and without effect on any result: after sigsvgd_mlp_profile_enable(1) thread 0 of
 * every CTA of sigsvgd_mlp_query adds the cycles it spent per phase to eight counters, read (and reset) here */
int sigsvgd_mlp_profile_enable(int on);
int sigsvgd_mlp_profile_cycles(unsigned long long* out8, int reset);
int sigsvgd_mlp_pack(const float* w_first, const float* b_first, const float* w_mid, const float* b_mid,
                     const float* w_last, const float* b_last, int in_dim, int hidden, int n_mid, void* packed,
                     size_t packed_bytes, void* stream);
int sigsvgd_mlp_query(const void* packed, int in_dim, int hidden, int n_mid, int precise, const float* x, int64_t P,
                      float* prob, float* grad, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Fused optimiser step on the particles.  Replaces the update statements of
 * SVGD.step (src/inference/svgd.py:93-116): `optimizer.step()` of torch.optim.SGD / torch.optim.Adam
 * (default flags: no weight decay, no amsgrad, no nesterov, dampening 0) and the hand-written Adagrad
 * (svgd.py:110-113), as ONE elementwise launch over the n = N*T*d particle coordinates.
 *   g = sign * dir[i]   (dir = phi with sign = -1: the reference minimises with X.grad = -phi)
 *   SGD:     m != NULL: buf = first ? g : beta1*buf + g (beta1 = momentum; buf in m);  x -= lr * (buf | g)
 *   ADAGRAD: v += g^2;  g' = g / sqrt(v + 1e-12);  x -= lr * g'
 *   ADAM:    m += (g - m)(1 - beta1);  v = beta2*v + (1 - beta2) g^2;  x -= c1 * m / (sqrt(v)/c2 + eps)
 *            with c1 = lr / (1 - beta1^t), c2 = sqrt(1 - beta2^t) formed by the caller for step t; the betas are
 *            doubles so that 1 - beta is formed as torch forms it (in double, then rounded to fp32)
 * grad_out (may be NULL) receives what the reference logs as the step's "grad" (g; g' for Adagrad). */
#define SIGSVGD_OPT_SGD 0
#define SIGSVGD_OPT_ADAGRAD 1
#define SIGSVGD_OPT_ADAM 2
int sigsvgd_opt_step(float* x, const float* dir, float sign, float* m, float* v, float* grad_out, int64_t n,
                     int kind, float lr, double beta1, double beta2, float eps, float c1, float c2, int first,
                     void* stream);

/* ---------------------------------------------------------------------------------------------
 * Fused single-GPU Stein update.  Replaces SVGD._compute_kernel + the arithmetic of
 * SVGD._velocity (src/inference/svgd.py:36-44, 82-83) for kernel = PathSigKernel(GaussianKernel):
 *   phi = ( K . score - grad_k ) / N ,   the reference's velocity is -phi.
 * x [B,N,T,d], score [B,N,T*d]  ->  phi [B,N,T,d]; B independent problems (B = 1 for one swarm),
 * each with its own bandwidth.  Optional outputs (NULL to skip): k_out [B,N,N] (k_xx),
 * gradk_out [B,N,T*d] (grad_k), sig_out [B,N,D].  state: B records (device).
 * ------------------------------------------------------------------------------------------- */
size_t sigsvgd_phi_workspace_bytes(int64_t B, int64_t N, int64_t T, int64_t d, int depth,
                                   int want_k);
int sigsvgd_phi(const float* x, const float* score, int64_t B, int64_t N, int64_t T, int64_t d,
                int depth, int bw_mode, float h_fixed, float bw_scale, float tol, uint32_t flags,
                float* phi, float* k_out, float* gradk_out, float* sig_out,
                sigsvgd_bw_state* state, void* ws, size_t ws_bytes, void* stream);

/* ---------------------------------------------------------------------------------------------
 * The same update for HOST-resident particles and scores (one swarm).  The reference keeps its
 * per-step tensors on the CPU side of the loop (src/inference/svgd.py:85-90 copies k_xx, grad_k and
 * every tensor kwarg to the host each step; grad_log_p arrives from CPU cost functions in
 * examples/script_planning_obstacle_field.py); this entry takes x_host [N,T,d], score_host
 * [N,T*d] (pinned host memory for asynchronous copies) and writes phi_host [N,T,d].  X is copied
 * first; the score's H2D copy runs on an internal stream beside the front half of the update
 * (signatures, Gram, median, RBF need only X); phi is produced in row chunks whose D2H copies
 * overlap the remaining signature backward.  phi_host is complete when `stream` reaches the end of
 * the call.  state: one device record.  The workspace also holds the device copies.
 * ------------------------------------------------------------------------------------------- */
size_t sigsvgd_phi_host_workspace_bytes(int64_t N, int64_t T, int64_t d, int depth);
int sigsvgd_phi_host(const float* x_host, const float* score_host, int64_t N, int64_t T, int64_t d,
                     int depth, int bw_mode, float h_fixed, float bw_scale, float tol, uint32_t flags,
                     float* phi_host, sigsvgd_bw_state* state, void* ws, size_t ws_bytes,
                     void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SIGSVGD_B200_H_ */
