"""CPU restatement (numpy, float64) of the signature-PDE kernel the reference's example scripts use
(SURVEY.md 8(f)-1).  TEST INFRASTRUCTURE ONLY -- nothing under ``sigsvgd_b200/`` imports this.

Reference call sites (relative to the reference root):

* ``src/kernels/_traj_kernels.py:147-195`` ``BatchGaussianKernel``: ``Gram_matrix(X, Y)`` =
  ``exp(-dist / h)`` over ALL point pairs ``[A, B, M, N]`` (note ``/ h``, not ``/ 2h^2``), ``h`` from
  ``bandwidth_fn(dist)``; the default ``bw_median`` (src/utils/math.py:28-34) takes the lower median of
  all A*B*M*N entries and divides by ``ln(dist.shape[0] + 1) = ln(A + 1)``;
* ``src/kernels/_traj_kernels.py:198-206`` ``SignatureKernel``:
  ``sigkernel.SigKernel(static_kernel, dyadic_order=depth).compute_Gram(X.double(), Y.double())``;
* ``src/inference/trajectory_svgd.py:55-69``, ``src/inference/score.py:64-76``: the Gram matrix and
  ``autograd.grad(k_xx.sum(), x)``.

The arithmetic lives in the third-party package ``sigkernel`` pinned at commit
``3b2373982e12b3d499a80228311a04debcc1bea1`` (reference ``setup.py:71``), which is ABSENT here (not under
/root/reference, no network): **parity unpinned**.  What is restated is its published algorithm
(Salvi et al., "The Signature Kernel is the solution of a Goursat PDE", 2021, and the package's
``_SigKernelGram``):

  forward   inc[s,t] = G[s+1,t+1] + G[s,t] - G[s+1,t] - G[s,t+1]          (coarse cell of the two paths)
            every coarse cell is split into 2^o x 2^o fine cells of increment inc / 4^o
            K[0,:] = K[:,0] = 1
            K[i+1,j+1] = (K[i,j+1] + K[i+1,j]) (1 + inc/2 + inc^2/12) - K[i,j] (1 - inc^2/12)
            k(x, y) = K[-1,-1]
  backward  (w.r.t. the FIRST argument only, as the package returns it)
            K_rev = the same solve on the increments flipped in both directions, flipped back
            W[s,t] = 4^-o * sum over the fine cells (i,j) of coarse cell (s,t) of K[i,j] K_rev[i+1,j+1]
            d k / d x_p = sum_{s,t} W[s,t] * d inc[s,t] / d x_p
  The package forms ``d inc / d x`` by a forward difference with step 1e-9 through
  ``static_kernel.Gram_matrix`` (which, with the reference's BatchGaussianKernel, re-evaluates the median
  bandwidth on the perturbed points: an O(1) artefact of that differencing).  Here -- and in the CUDA
  kernels -- it is the analytic derivative at a FIXED bandwidth.

Pinned by known answers instead (tests/test_oracle_sigpde.py): with the linear static kernel and two
straight lines the Goursat solution is ``sum_n <u,v>^n / (n!)^2 = I0(2 sqrt(<u,v>))``; symmetry;
invariance of the limit under re-sampling; the backward against the exact derivative of the discrete
scheme (torch autograd through the same recursion), which it approaches as the grid is refined.
"""
from __future__ import annotations

import numpy as np

RBF, LINEAR = 0, 1


def static_gram(X, Y, kind=RBF, h=None):
    """G [A,B,M,N] and the squared distances; RBF: exp(-|x_s - y_t|^2 / h); LINEAR: <x_s, y_t> * h (h = scale)."""
    X, Y = np.asarray(X, np.float64), np.asarray(Y, np.float64)
    diff = X[:, None, :, None, :] - Y[None, :, None, :, :]
    dist = (diff * diff).sum(-1)
    if kind == LINEAR:
        return np.einsum("ipk,jqk->ijpq", X, Y) * (1.0 if h is None else h), dist
    return np.exp(-dist / h), dist


def bw_median_points(dist, bw_scale=1.0, tol=1e-8):
    """``bw_median`` on the [A,B,M,N] tensor: lower median of all entries, ln(A + 1) (math.py:28-34)."""
    flat = np.sort(dist.reshape(-1))
    med = flat[(flat.size - 1) // 2]
    # (the reference evaluates ln(A + 1) on a default-dtype, i.e. float32, 0-dim tensor: math.py:32; torch's
    # float32 log itself is used here, it can differ from libm's by an ulp)
    import torch

    log_a1 = float(torch.tensor(dist.shape[0] + 1.0, dtype=torch.float32).log())
    return max(bw_scale * np.sqrt(med / log_a1), tol), med


def _solve(inc, order):
    """Goursat solve on the 2^order-refined grid for increments inc [..., M-1, N-1] -> K [..., MM+1, NN+1]."""
    f = 1 << order
    fine = np.repeat(np.repeat(inc, f, axis=-2), f, axis=-1) / (f * f)
    MM, NN = fine.shape[-2:]
    K = np.ones(fine.shape[:-2] + (MM + 1, NN + 1))
    for i in range(MM):
        for j in range(NN):
            p = fine[..., i, j]
            K[..., i + 1, j + 1] = (K[..., i, j + 1] + K[..., i + 1, j]) * (1.0 + 0.5 * p + p * p / 12.0) \
                - K[..., i, j] * (1.0 - p * p / 12.0)
    return K


def increments(G):
    return G[..., 1:, 1:] + G[..., :-1, :-1] - G[..., 1:, :-1] - G[..., :-1, 1:]


def sigpde_gram(X, Y, order=0, kind=RBF, h=None, return_all=False):
    """k [A,B] (and, with return_all, the intermediate tensors)."""
    G, dist = static_gram(X, Y, kind, h)
    inc = increments(G)
    K = _solve(inc, order)
    if return_all:
        return K[..., -1, -1], dict(G=G, dist=dist, inc=inc, K=K)
    return K[..., -1, -1]


def sigpde_weights(inc, order):
    """W [..., M-1, N-1] of the backward (see the header)."""
    f = 1 << order
    K = _solve(inc, order)
    Kr = _solve(inc[..., ::-1, ::-1], order)[..., ::-1, ::-1]
    GG = K[..., :-1, :-1] * Kr[..., 1:, 1:]
    sh = GG.shape[:-2] + (GG.shape[-2] // f, f, GG.shape[-1] // f, f)
    return GG.reshape(sh).sum(axis=(-3, -1)) / (f * f)


def sigpde_gram_bwd(X, Y, gout, order=0, kind=RBF, h=None):
    """d (sum_ab gout[a,b] k[a,b]) / d X  [A,M,D], first argument only, fixed bandwidth."""
    X, Y, gout = np.asarray(X, np.float64), np.asarray(Y, np.float64), np.asarray(gout, np.float64)
    A, M, D = X.shape
    G, _ = static_gram(X, Y, kind, h)
    W = sigpde_weights(increments(G), order) * gout[:, :, None, None]  # [A,B,M-1,N-1]
    if kind == LINEAR:
        # dG[s,t]/dx_s = y_t * scale
        dG = np.broadcast_to(Y[None, :, None, :, :] * (1.0 if h is None else h), G.shape + (D,))
    else:
        dG = G[..., None] * (-2.0 / h) * (X[:, None, :, None, :] - Y[None, :, None, :, :])  # [A,B,M,N,D]
    # inc[s,t] = G[s+1,t+1] - G[s+1,t] (through x_{s+1})  +  G[s,t] - G[s,t+1] (through x_s)
    hi = dG[:, :, 1:, 1:] - dG[:, :, 1:, :-1]   # d inc[s,t] / d x_{s+1}
    lo = dG[:, :, :-1, :-1] - dG[:, :, :-1, 1:]  # d inc[s,t] / d x_s
    grad = np.zeros((A, M, D))
    grad[:, 1:] += (W[..., None] * hi).sum(axis=(1, 3))
    grad[:, :-1] += (W[..., None] * lo).sum(axis=(1, 3))
    return grad


def sigpde_gram_torch(X, Y, order=0, kind=RBF, h=None):
    """The forward recursion in torch (float64, autograd-able): the EXACT derivative of the discrete
    scheme, against which the package's continuous-limit backward formula is compared in the tests."""
    import torch

    diff = X[:, None, :, None, :] - Y[None, :, None, :, :]
    dist = (diff * diff).sum(-1)
    G = torch.einsum("ipk,jqk->ijpq", X, Y) * (1.0 if h is None else h) if kind == LINEAR else torch.exp(-dist / h)
    inc = G[..., 1:, 1:] + G[..., :-1, :-1] - G[..., 1:, :-1] - G[..., :-1, 1:]
    f = 1 << order
    fine = inc.repeat_interleave(f, -2).repeat_interleave(f, -1) / (f * f)
    MM, NN = fine.shape[-2:]
    rows = [[torch.ones(fine.shape[:-2], dtype=X.dtype) for _ in range(NN + 1)]]
    for i in range(MM):
        row = [torch.ones(fine.shape[:-2], dtype=X.dtype)]
        for j in range(NN):
            p = fine[..., i, j]
            row.append((rows[i][j + 1] + row[j]) * (1.0 + 0.5 * p + p * p / 12.0) - rows[i][j] * (1.0 - p * p / 12.0))
        rows.append(row)
    return rows[-1][-1]
