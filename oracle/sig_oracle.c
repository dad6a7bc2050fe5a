/*
 * oracle/sig_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT.
 *
 * CPU restatement of the truncated path-signature transform that the reference
 * obtains from the third-party extension  signatory==1.2.6.1.9.0
 * (pinned at /root/reference/setup.py:58, environment.yaml:22; call sites
 * /root/reference/src/kernels/_traj_kernels.py:125-126:
 *     signatory.signature(X, depth, basepoint=True) ).
 * signatory's sources are NOT under /root/reference and the package cannot be
 * installed here, so this file restates its *published* algorithm:
 *   forward  : S <- S (x) exp(Delta_t) with the fused Horner update
 *              ("mult_fused_restricted_exp"), levels 1..depth, level-major,
 *              each level flattened row-major, no scalar term;
 *   backward : reversible sweep S_{t-1} = S_t (x) exp(-Delta_t), adjoint of the
 *              Horner chain, gradient wrt increments -> gradient wrt points.
 *
 * PARITY STATUS: "parity unpinned" for this transform -- the reference holds no
 * golden vectors for it (its tests/ have no assertions) and signatory itself is
 * absent. It is pinned only by closed-form known-answer tests
 * (tests/test_oracle_signature.py: level-1 = last point, level-2 closed form,
 * shuffle identities, straight-line exp, Chen, re-parametrisation invariance)
 * and by fp64 finite differences for the backward.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference leg may load this library.
 *
 * Build: see oracle/Makefile  (gcc -O3 -fopenmp -shared -fPIC).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define MAX_DEPTH 8

/* number of output channels  D = sum_{k=1..m} d^k */
int64_t sig_oracle_channels(int64_t d, int depth) {
  int64_t D = 0, p = 1;
  for (int k = 1; k <= depth; ++k) {
    p *= d;
    D += p;
  }
  return D;
}

int sig_oracle_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

#define DEFINE_ORACLE(real, SFX)                                                               \
  /* S <- S (x) exp(dl), in place. lvl[k] points at level k+1 (k = 0..m-1).            */      \
  /* scratch: two buffers of d^{m-1} reals.                                            */      \
  static void fused_exp_mul_##SFX(real *S, const real *dl, int d, int m, const int64_t *off,   \
                                  const int64_t *pw, real *sa, real *sb) {                     \
    for (int k = m; k >= 2; --k) {                                                             \
      real *cur = sa, *nxt = sb;                                                               \
      const real rk = (real)1 / (real)k;                                                       \
      for (int i = 0; i < d; ++i) cur[i] = S[off[1] + i] + dl[i] * rk;                         \
      for (int j = 2; j <= k - 1; ++j) {                                                       \
        const real r = (real)1 / (real)(k - j + 1);                                            \
        const real *Sj = S + off[j];                                                           \
        for (int64_t u = 0; u < pw[j - 1]; ++u) {                                              \
          const real cu = cur[u] * r;                                                          \
          for (int c = 0; c < d; ++c) nxt[u * d + c] = Sj[u * d + c] + cu * dl[c];             \
        }                                                                                      \
        real *t = cur;                                                                         \
        cur = nxt;                                                                             \
        nxt = t;                                                                               \
      }                                                                                        \
      real *Sk = S + off[k];                                                                   \
      for (int64_t u = 0; u < pw[k - 1]; ++u) {                                                \
        const real cu = cur[u];                                                                \
        for (int c = 0; c < d; ++c) Sk[u * d + c] += cu * dl[c];                               \
      }                                                                                        \
    }                                                                                          \
    for (int i = 0; i < d; ++i) S[off[1] + i] += dl[i];                                        \
  }                                                                                            \
                                                                                               \
  /* Adjoint of  Snew = P (x) exp(dl).  A holds dL/dSnew on entry, dL/dP on exit;      */      \
  /* gdl += dL/d dl.  B: scratch of D reals (Horner intermediates), bb/bc: d^{m-1}.    */      \
  static void fused_exp_mul_bwd_##SFX(const real *P, real *A, const real *dl, real *gdl,       \
                                      int d, int m, const int64_t *off, const int64_t *pw,     \
                                      real *B, real *bb, real *bc) {                           \
    for (int c = 0; c < d; ++c) gdl[c] += A[off[1] + c]; /* Snew^1 = P^1 + dl */               \
    for (int k = 2; k <= m; ++k) { /* ascending: A^k is still dL/dSnew^k */                    \
      const real rk = (real)1 / (real)k;                                                       \
      /* forward intermediates B_1 .. B_{k-1} (B_j stored at B+off[j]) */                      \
      for (int i = 0; i < d; ++i) B[off[1] + i] = P[off[1] + i] + dl[i] * rk;                  \
      for (int j = 2; j <= k - 1; ++j) {                                                       \
        const real r = (real)1 / (real)(k - j + 1);                                            \
        for (int64_t u = 0; u < pw[j - 1]; ++u) {                                              \
          const real cu = B[off[j - 1] + u] * r;                                               \
          for (int c = 0; c < d; ++c)                                                          \
            B[off[j] + u * d + c] = P[off[j] + u * d + c] + cu * dl[c];                        \
        }                                                                                      \
      }                                                                                        \
      /* top: Snew^k = P^k + B_{k-1} (x) dl */                                                 \
      real *gb = bb, *gn = bc;                                                                 \
      for (int64_t u = 0; u < pw[k - 1]; ++u) {                                                \
        real s = 0;                                                                            \
        const real bu = B[off[k - 1] + u];                                                     \
        for (int c = 0; c < d; ++c) {                                                          \
          const real a = A[off[k] + u * d + c];                                                \
          s += a * dl[c];                                                                      \
          gdl[c] += a * bu;                                                                    \
        }                                                                                      \
        gb[u] = s;                                                                             \
      }                                                                                        \
      for (int j = k - 1; j >= 1; --j) { /* B_j = P^j + B_{j-1} (x) dl / (k-j+1) */            \
        for (int64_t u = 0; u < pw[j]; ++u) A[off[j] + u] += gb[u];                            \
        const real r = (real)1 / (real)(k - j + 1);                                            \
        if (j >= 2) {                                                                          \
          for (int64_t u = 0; u < pw[j - 1]; ++u) {                                            \
            real s = 0;                                                                        \
            const real bu = B[off[j - 1] + u] * r;                                             \
            for (int c = 0; c < d; ++c) {                                                      \
              const real g = gb[u * d + c];                                                    \
              s += g * dl[c];                                                                  \
              gdl[c] += g * bu;                                                                \
            }                                                                                  \
            gn[u] = s * r;                                                                     \
          }                                                                                    \
          real *t = gb;                                                                        \
          gb = gn;                                                                             \
          gn = t;                                                                              \
        } else {                                                                               \
          for (int c = 0; c < d; ++c) gdl[c] += gb[c] * r; /* B_0 = 1 */                       \
        }                                                                                      \
      }                                                                                        \
    }                                                                                          \
  }                                                                                            \
                                                                                               \
  /* x: [N,T,d] row-major; out: [N,D]. basepoint!=0 prepends a zero point. */                  \
  int sig_oracle_fwd_##SFX(const real *x, int64_t N, int64_t T, int64_t d, int depth,          \
                           int basepoint, real *out) {                                         \
    if (depth < 1 || depth > MAX_DEPTH || d < 1 || T < 1) return 1;                            \
    if (!basepoint && T < 2) return 1;                                                         \
    int64_t off[MAX_DEPTH + 2], pw[MAX_DEPTH + 2];                                             \
    pw[0] = 1;                                                                                 \
    off[1] = 0;                                                                                \
    for (int k = 1; k <= depth; ++k) {                                                         \
      pw[k] = pw[k - 1] * d;                                                                   \
      off[k + 1] = off[k] + pw[k];                                                             \
    }                                                                                          \
    const int64_t D = off[depth + 1];                                                          \
    const int64_t sc = pw[depth - 1] > d ? pw[depth - 1] : d;                                  \
    _Pragma("omp parallel") {                                                                  \
      real *sa = (real *)malloc(sizeof(real) * (size_t)(2 * sc + d));                          \
      real *sb = sa + sc, *dl = sb + sc;                                                       \
      _Pragma("omp for schedule(static)") for (int64_t n = 0; n < N; ++n) {                    \
        const real *xn = x + n * T * d;                                                        \
        real *S = out + n * D;                                                                 \
        memset(S, 0, sizeof(real) * (size_t)D);                                                \
        for (int64_t t = basepoint ? 0 : 1; t < T; ++t) {                                      \
          for (int c = 0; c < d; ++c)                                                          \
            dl[c] = xn[t * d + c] - (t > 0 ? xn[(t - 1) * d + c] : (real)0);                   \
          fused_exp_mul_##SFX(S, dl, (int)d, depth, off, pw, sa, sb);                          \
        }                                                                                      \
      }                                                                                        \
      free(sa);                                                                                \
    }                                                                                          \
    return 0;                                                                                  \
  }                                                                                            \
                                                                                               \
  /* Reversible backward: sig = S_T [N,D], gsig = dL/dS_T [N,D] -> gx [N,T,d]. */              \
  int sig_oracle_bwd_##SFX(const real *x, const real *sig, const real *gsig, int64_t N,        \
                           int64_t T, int64_t d, int depth, int basepoint, real *gx) {         \
    if (depth < 1 || depth > MAX_DEPTH || d < 1 || T < 1) return 1;                            \
    if (!basepoint && T < 2) return 1;                                                         \
    int64_t off[MAX_DEPTH + 2], pw[MAX_DEPTH + 2];                                             \
    pw[0] = 1;                                                                                 \
    off[1] = 0;                                                                                \
    for (int k = 1; k <= depth; ++k) {                                                         \
      pw[k] = pw[k - 1] * d;                                                                   \
      off[k + 1] = off[k] + pw[k];                                                             \
    }                                                                                          \
    const int64_t D = off[depth + 1];                                                          \
    const int64_t sc = pw[depth - 1] > d ? pw[depth - 1] : d;                                  \
    _Pragma("omp parallel") {                                                                  \
      real *buf = (real *)malloc(sizeof(real) * (size_t)(3 * D + 4 * sc + 3 * d));             \
      real *S = buf, *A = S + D, *B = A + D, *sa = B + D, *sb = sa + sc, *bb = sb + sc,        \
           *bc = bb + sc, *dl = bc + sc, *ndl = dl + d, *gdl = ndl + d;                        \
      _Pragma("omp for schedule(static)") for (int64_t n = 0; n < N; ++n) {                    \
        const real *xn = x + n * T * d;                                                        \
        real *gxn = gx + n * T * d;                                                            \
        memcpy(S, sig + n * D, sizeof(real) * (size_t)D);                                      \
        memcpy(A, gsig + n * D, sizeof(real) * (size_t)D);                                     \
        memset(gxn, 0, sizeof(real) * (size_t)(T * d));                                        \
        for (int64_t t = T - 1; t >= (basepoint ? 0 : 1); --t) {                               \
          for (int c = 0; c < d; ++c) {                                                        \
            dl[c] = xn[t * d + c] - (t > 0 ? xn[(t - 1) * d + c] : (real)0);                   \
            ndl[c] = -dl[c];                                                                   \
            gdl[c] = 0;                                                                        \
          }                                                                                    \
          fused_exp_mul_##SFX(S, ndl, (int)d, depth, off, pw, sa, sb); /* S_{t-1} */           \
          fused_exp_mul_bwd_##SFX(S, A, dl, gdl, (int)d, depth, off, pw, B, bb, bc);           \
          for (int c = 0; c < d; ++c) {                                                        \
            gxn[t * d + c] += gdl[c];                                                          \
            if (t > 0) gxn[(t - 1) * d + c] -= gdl[c];                                         \
          }                                                                                    \
        }                                                                                      \
      }                                                                                        \
      free(buf);                                                                               \
    }                                                                                          \
    return 0;                                                                                  \
  }

DEFINE_ORACLE(float, f32)
DEFINE_ORACLE(double, f64)
