// Shared helpers for the sigsvgd_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <algorithm>

#include "../../include/sigsvgd_b200.h"

namespace sigsvgd {

// thread-local last-error string (C ABI: sigsvgd_last_error)
void set_error(const char* fmt, ...);
const char* get_error();

#define SIG_CHECK_ARG(cond, ...)             \
  do {                                       \
    if (!(cond)) {                           \
      ::sigsvgd::set_error(__VA_ARGS__);     \
      return SIGSVGD_ERR_INVALID;            \
    }                                        \
  } while (0)

#define SIG_CUDA_OK(expr)                                                                   \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess) {                                                                \
      ::sigsvgd::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                           __LINE__);                                                       \
      return SIGSVGD_ERR_CUDA;                                                              \
    }                                                                                       \
  } while (0)

#define SIG_LAUNCH_OK(name)                                                              \
  do {                                                                                   \
    cudaError_t _e = cudaGetLastError();                                                 \
    if (_e != cudaSuccess) {                                                             \
      ::sigsvgd::set_error("launch of %s failed: %s", name, cudaGetErrorString(_e));     \
      return SIGSVGD_ERR_CUDA;                                                           \
    }                                                                                    \
  } while (0)

#define SIG_PROPAGATE(expr)  \
  do {                       \
    int _rc = (expr);        \
    if (_rc != 0) return _rc; \
  } while (0)

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }
static inline int64_t round_up(int64_t a, int64_t b) { return ceil_div(a, b) * b; }

static inline int64_t sig_channels(int64_t d, int depth) {
  int64_t D = 0, p = 1;
  for (int k = 1; k <= depth; ++k) {
    p *= d;
    D += p;
  }
  return D;
}

int num_sms();  // cached cudaDevAttrMultiProcessorCount of the current device

// ---- launchers implemented in the individual .cu files ------------------------------------
// signature (sig_kernels.cu)
// stats != NULL (depth 3, d <= 8): also accumulates stats[0..D) += column sums of sig,
// stats[D] = max |sig| bits, stats[D+1] = max |score| bits (score may be NULL); caller zeroes stats
bool sig_fwd_fuses_stats(int64_t d, int depth);
int launch_sig_fwd(const float* x, int64_t N, int64_t T, int64_t d, int depth, int basepoint,
                   float* sig, cudaStream_t st, float* stats = nullptr, const float* score = nullptr,
                   int64_t ld_score = 0, int64_t P = 0, float* pref = nullptr);
// Prefix checkpoints for the time-parallel backward (sig3_bwd_seg_kernel): pref[N][3][d+d*d] holds
// levels 1-2 of the signature of each path up to the end of time segments 0..2 of ceil(T/4) steps.  The forward
// writes them when pref != NULL, the backward with the same pref sweeps the 4 segments at once.
bool sig_bwd_uses_prefix(int64_t T, int64_t d, int depth);
int64_t sig_prefix_floats(int64_t N, int64_t T, int64_t d, int depth);  // 0 when not available
// Optional producer of the cotangent: instead of reading gsig, the depth-3 backward kernel forms
//   g_i = -(1/h^2) (sc_i * rowsum_i - ks_i) + median-gradient term on the tie rows
// (the closed form of autograd.grad(k_xx.sum(), X) in signature space) while it
// stages its operands -- what sig_grad_kernel computes, without the [N,D] round trip through HBM.
struct SigGradSpec {
  const float* sc;       // centred features of these rows, pitch lds
  int64_t lds;
  const float* sc_all;   // centred features of ALL particles (tie partners), pitch lds_all
  int64_t lds_all;
  const float* ks;       // K . Sc rows, pitch ldks
  int64_t ldks;
  const float* rowsum;   // K row sums, element pitch ldr
  int64_t ldr;
  int64_t row_offset;    // global index of row 0 (ties hold global indices)
  const sigsvgd_bw_state* state;  // one record
  int coef_mode = 0;     // 0: state->coef / n_ties_all as set by median_coef; 1: derive them from the
                         // record's own sum_kd / n_ties (and store them); 2: same, gradient term off
  // batches of independent scenes stacked along the rows (0 = one problem): scene b = row / scene_rows has
  // its own record state[b], its tie indices count from the scene's first row, and its partners are
  // the scene's own rows of sc_all.  coef_mode must be 0 then.
  int64_t scene_rows = 0;
};
// the spec as the scene of global row `row` sees it
__device__ __forceinline__ SigGradSpec scene_view(SigGradSpec gs, int64_t row) {
  if (gs.scene_rows > 0 && gs.sc != nullptr) {
    const int64_t scene = row / gs.scene_rows, r0 = scene * gs.scene_rows;
    gs.state += scene;
    gs.sc_all += r0 * gs.lds_all;
    gs.row_offset = -r0;
  }
  return gs;
}
// Median-gradient coefficient for the fused cotangent: either the record's (coef_mode 0: set by
// median_coef_kernel, e.g. from all-reduced sums on the sharded path) or derived here from the
// record's own sums, coef = sum_kd / (h^2 median) -- one launch less
// on the one-swarm path.  The first block of the first row chunk also stores it in the record.
__device__ __forceinline__ float grad_coef(const SigGradSpec& gs, uint32_t* n_ties_all) {
  const sigsvgd_bw_state& st = *gs.state;
  if (gs.coef_mode == 0) {
    *n_ties_all = st.n_ties_all;
    return st.coef;
  }
  const uint32_t nt = st.n_ties;
  const bool on = gs.coef_mode == 1 && !st.clamped && st.median > 0.f && nt > 0u;
  const float coef = on ? (float)st.sum_kd / (st.h * st.h * st.median) : 0.f;
  if (blockIdx.x == 0 && threadIdx.x == 0 && gs.row_offset == 0) {
    sigsvgd_bw_state* w = const_cast<sigsvgd_bw_state*>(gs.state);
    w->coef = coef;
    w->n_ties_all = nt;
  }
  *n_ties_all = nt;
  return coef;
}

int select_collect_blocks();
int launch_select_collect(const float* dist, int64_t count, const float* bounds, float* cand,
                          int64_t cand_cap, unsigned long long* totals, cudaStream_t st);
bool sig_bwd_fuses_grad(int64_t d, int depth);
int launch_sig_bwd(const float* x, const float* sig, const float* gsig, int64_t N, int64_t T,
                   int64_t d, int depth, int basepoint, const float* add, int64_t ld_add,
                   float alpha, float beta, float* out, cudaStream_t st,
                   const SigGradSpec* spec = nullptr, const float* pref = nullptr);

// depth 4, 2 <= d <= 8 (sig4_kernels.cu): d^2 threads per particle, level 4 in registers
bool sig4_available(int64_t d);
int launch_sig4_fwd(const float* x, int64_t N, int64_t T, int64_t d, int basepoint, float* sig, cudaStream_t st);
int launch_sig4_bwd(const float* x, const float* sig, const float* gsig, int64_t N, int64_t T, int64_t d,
                    int basepoint, const float* add, int64_t ld_add, float alpha, float beta, float* out,
                    cudaStream_t st, const SigGradSpec* spec = nullptr);

// signature-PDE kernel (sigpde.cu)
int launch_sigpde_bandwidth(const double* x, const double* y, int64_t A, int64_t B, int64_t M, int64_t N,
                            int64_t D, float bw_scale, float tol, sigsvgd_bw_state* state, uint64_t* hist,
                            cudaStream_t st);
int launch_sigpde_fwd(const double* x, const double* y, int64_t A, int64_t B, int64_t M, int64_t N, int64_t D,
                      int order, int kind, double h, const sigsvgd_bw_state* state, double* k_out,
                      cudaStream_t st);
size_t sigpde_bwd_workspace_bytes(int64_t A, int64_t B, int64_t M, int64_t N, int64_t D, int order);
int launch_sigpde_bwd(const double* x, const double* y, int64_t A, int64_t B, int64_t M, int64_t N, int64_t D,
                      int order, int kind, double h, const sigsvgd_bw_state* state, const double* gout,
                      double* grad_x, void* ws, size_t ws_bytes, cudaStream_t st);

// Stein-MPC particle-model rollout (rollout.cu)
int launch_particle_rollout(const float* init, const float* actions, const float* mass, float mass_scalar, int64_t R,
                            int64_t H, float dt, float max_acc, float max_speed, const uint8_t* grid, int64_t gx,
                            int64_t gy, float inv_cell, float off_x, float off_y, int can_crash, float* states,
                            uint8_t* flags, cudaStream_t st);
int launch_particle_rollout_bwd(const float* mass, float mass_scalar, int64_t R, int64_t H, float dt,
                                const float* grad_states, const uint8_t* flags, float* grad_actions,
                                cudaStream_t st);

// serial-chain forward kinematics (fk.cu)
int launch_chain_fk(const float* xyz, const float* rot, const float* axis, const int* dof, int J, int n_dof,
                    const float* q, int64_t B, float* xs, cudaStream_t st);
int launch_chain_fk_bwd(const float* xyz, const float* rot, const float* axis, const int* dof, int J, int n_dof,
                        const float* q, int64_t B, const float* gxs, float* gq, cudaStream_t st);

// occupancy / self-collision network query with its input gradient (mlp_tc.cu)
size_t mlp_packed_bytes(int in_dim, int hidden, int n_mid);
int mlp_profile_enable(int on);
int mlp_profile_cycles(unsigned long long* out8, int reset);
int launch_mlp_pack(const float* w_first, const float* b_first, const float* w_mid, const float* b_mid,
                    const float* w_last, const float* b_last, int in_dim, int hidden, int n_mid, void* packed,
                    size_t packed_bytes, cudaStream_t st);
int launch_mlp_query(const void* packed, int in_dim, int hidden, int n_mid, int precise, const float* x, int64_t P,
                     float* prob, float* grad, cudaStream_t st);

// SIMT GEMMs (gemm_simt.cu)
int launch_sqdist_simt(const float* a, int64_t lda, const float* b, int64_t ldb, int64_t M,
                       int64_t N, int64_t D, float* dist, int64_t ldd, int64_t batch,
                       int64_t stride_a, int64_t stride_b, int64_t stride_d, cudaStream_t st);
int launch_gemm_nn_simt(const float* a, int64_t lda, const float* v, int64_t ldv, int64_t M,
                        int64_t N, int64_t C, float* out, int64_t ldo, int64_t batch,
                        int64_t stride_a, int64_t stride_v, int64_t stride_o, cudaStream_t st);

// selection / bandwidth (select.cu)
size_t select_workspace_bytes(int64_t count, int64_t batch);
int launch_select_begin(sigsvgd_bw_state* state, uint64_t* hist, uint64_t k, int64_t batch,
                        cudaStream_t st);
int launch_select_begin_window(sigsvgd_bw_state* state, uint64_t* hist, uint64_t k_total,
                               const uint64_t* totals, int* ok, cudaStream_t st);
int launch_select_hist(const float* dist, int64_t count, int64_t batch, int64_t stride, int pass,
                       const sigsvgd_bw_state* state, uint64_t* hist, cudaStream_t st);
int launch_select_scan(int pass, sigsvgd_bw_state* state, uint64_t* hist, int64_t batch,
                       cudaStream_t st);
int launch_select_finish(int64_t n_log, float bw_scale, float tol, sigsvgd_bw_state* state,
                         int64_t batch, cudaStream_t st);
int launch_fixed_bandwidth(float h, sigsvgd_bw_state* state, int64_t batch, cudaStream_t st);
int launch_median_coef(sigsvgd_bw_state* state, const double* sum_kd_all,
                       const uint32_t* n_ties_all, int enable, int64_t batch, cudaStream_t st);
int launch_median_select(const float* dist, int64_t count, int64_t batch, int64_t stride,
                         int64_t n_log, float bw_scale, float tol, sigsvgd_bw_state* state,
                         void* ws, size_t ws_bytes, cudaStream_t st);
// same, scratch in a file-scope device array (select.cu); batch <= SIGSVGD_STATIC_MAX_BATCH
int launch_median_select_static(const float* dist, int64_t count, int64_t batch, int64_t stride,
                                int64_t n_log, float bw_scale, float tol, sigsvgd_bw_state* state,
                                cudaStream_t st);

// elementwise / packing (elementwise.cu)
int launch_rbf_apply(float* k, int64_t ldk, int64_t M, int64_t N, int64_t row_offset,
                     const float* weight, int64_t ldw, sigsvgd_bw_state* state, int64_t batch,
                     int64_t stride_k, int64_t stride_w, cudaStream_t st);
int launch_opt_step(float* x, const float* dir, float sign, float* m, float* v, float* grad_out, int64_t n,
                    int kind, float lr, double b1, double b2, float eps, float c1, float c2, int first,
                    cudaStream_t st);
int launch_rbf_out(const float* d, int64_t ldd, int64_t M, int64_t N, const sigsvgd_bw_state* state,
                   float* k, int64_t ldk, cudaStream_t st);
int launch_col_mean(const float* s, int64_t lds, int64_t N, int64_t D, float* part, float* mean,
                    int64_t batch, int64_t stride_s, cudaStream_t st);
int launch_pack_v(const float* score, int64_t P, const float* s, int64_t lds, const float* mean,
                  int64_t N, int64_t D, float* v, int64_t ldv, int64_t C, int64_t batch,
                  cudaStream_t st);
int launch_sig_grad(const float* sc, int64_t lds, const float* sc_all, int64_t lds_all,
                    const float* ks, int64_t ldks, const float* rowsum, int64_t ldr, int64_t M,
                    int64_t D, int64_t n_all, int64_t row_offset, const sigsvgd_bw_state* state,
                    float* g, int64_t ldg, int64_t batch, cudaStream_t st);

}  // namespace sigsvgd
