"""CPU *checker* backend for sigsvgd_b200.dist.phi_sharded -- TEST INFRASTRUCTURE.

Implements the backend interface (one method per C-ABI entry point, see CudaBackend) with
numpy/torch on the CPU, using the oracle for the signature transform, so that the world_size-2
``gloo`` tests can drive the *orchestration* of the row-block-sharded update (all-gather,
histogram all-reduces of the exact global median, tie/sum_kd reduction, row offsets) without a
GPU.  It is never importable from the product package."""
from __future__ import annotations

import math

import numpy as np
import torch

import oracle

_SHIFT = (21, 10, 0)
_BITS = (11, 11, 10)


class _State(dict):
    """Stands in for the device-resident sigsvgd_bw_state record."""


class CheckerBackend:
    name = "checker"

    def sig_fwd(self, x, depth):
        return torch.from_numpy(oracle.signature_np(x.numpy().astype(np.float32), depth, True))

    def pack(self, s_all, g_all, center, r0, M):
        from sigsvgd_b200.dist import Features

        N, D = s_all.shape
        P = g_all.shape[1]
        C = (P + D + 1 + 7) // 8 * 8
        v = torch.zeros(N, C, dtype=torch.float32)
        v[:, :P] = g_all
        v[:, P:P + D] = s_all - (s_all.mean(0) if center else 0.0)
        v[:, P + D] = 1.0
        return Features(v[:, P:], v[r0:r0 + M, P:], (v, D, P, r0, M))

    def gram(self, f):
        v, D, P, r0, M = f.data
        a, b = v[r0:r0 + M, P:P + D].double(), v[:, P:P + D].double()
        d = (a * a).sum(1)[:, None] + (b * b).sum(1)[None, :] - 2.0 * a @ b.T
        return d.clamp_(min=0).float().contiguous()

    def new_state(self, device):
        return _State(median=0.0, h=1.0, gamma=-0.5, coef=0.0, sum_kd=0.0, n_ties=0, n_ties_all=0, clamped=0,
                      prefix=0, k_rem=0, ties=[])

    def new_hist(self, device):
        return torch.zeros(2048, dtype=torch.int64)

    def fixed_bandwidth(self, h, state):
        state.update(h=float(h), gamma=-0.5 / float(h) ** 2, median=0.0, clamped=0)

    # ---- three-pass radix select on the fp32 bit patterns (11 / 11 / 10 bits) ---------------
    def select_begin(self, state, hist, k):
        state.update(prefix=0, k_rem=int(k))
        hist.zero_()

    def select_begin_window(self, state, hist, k_total, totals):
        below, kept, over = (int(v) for v in totals.tolist())
        hit = over == 0 and below <= k_total < below + kept
        self.select_begin(state, hist, k_total - below if hit else 0)
        return torch.tensor([1 if hit else 0], dtype=torch.int32)

    def select_hist(self, dist_loc, p, state, hist):
        bits = dist_loc.contiguous().view(torch.int32).reshape(-1).to(torch.int64)
        if p > 0:
            hi_shift = _SHIFT[p - 1]
            bits = bits[(bits >> hi_shift) == (state["prefix"] >> hi_shift)]
        bins = (bits >> _SHIFT[p]) & ((1 << _BITS[p]) - 1)
        hist += torch.bincount(bins, minlength=2048)[:2048]

    def select_scan(self, p, state, hist):
        cum = torch.cumsum(hist, 0)
        b = int(torch.searchsorted(cum, torch.tensor(state["k_rem"]), right=True))
        state["k_rem"] -= int(cum[b - 1]) if b > 0 else 0
        state["prefix"] |= b << _SHIFT[p]
        hist.zero_()

    def select_window(self, sc_all, state):
        a = sc_all.double()
        idx = torch.arange(0, a.shape[0], max(1, a.shape[0] // 16))
        d = ((a[idx, None, :] - a[None, :, :]) ** 2).sum(-1).reshape(-1)
        lo, hi = torch.quantile(d, 0.35), torch.quantile(d, 0.65)
        return torch.tensor([float(lo), float(hi)], dtype=torch.float32)

    def select_collect(self, dist_loc, bounds):
        flat = dist_loc.reshape(-1)
        lo, hi = float(bounds[0]), float(bounds[1])
        inside = flat[(flat >= lo) & (flat <= hi)]
        cap = max(64, flat.numel() // 2)
        kept = min(cap, inside.numel())
        cand = torch.full((cap,), 3.0e38, dtype=torch.float32)
        cand[:kept] = inside[:kept]
        totals = torch.tensor([int((flat < lo).sum()), kept, int(inside.numel() > cap)], dtype=torch.int64)
        return cand, totals

    def select_finish(self, n_log, bw_scale, tol, state):
        med = float(np.array([state["prefix"]], dtype=np.uint32).view(np.float32)[0])
        h = bw_scale * math.sqrt(med / math.log(n_log + 1.0))
        state.update(median=med, clamped=int(h < tol))
        h = max(h, tol)
        state.update(h=h, gamma=-0.5 / (h * h))

    def rbf(self, k, state, row_offset, f, want_k):
        med = np.float32(state["median"])
        d = k.clone()
        k.copy_(torch.exp(np.float32(state["gamma"]) * d))
        state["sum_kd"] = float((k.double() * d.double()).sum())
        idx = torch.nonzero(d == float(med)) if state["median"] > 0 else torch.zeros(0, 2, dtype=torch.int64)
        state["n_ties"] = int(idx.shape[0])
        state["ties"] = [(int(i) + row_offset, int(j)) for i, j in idx[:64]]
        return k, (k if want_k else None)

    def local_stats(self, state):
        return torch.tensor([state["sum_kd"], float(state["n_ties"])], dtype=torch.float64)

    def median_coef(self, state, enable, stats):
        state["n_ties_all"] = int(stats[1])
        ok = enable and not state["clamped"] and state["median"] > 0
        state["coef"] = float(stats[0]) / (state["h"] ** 2 * state["median"]) if ok else 0.0

    def contract(self, kmat, f, state=None, row_offset=0):
        v, D, P, r0, M = f.data
        out = (kmat.double() @ v.double()).float()
        return out[:, P:P + D], out[:, P + D], out[:, :P]

    def sig_grad(self, sc, sc_all, ks, rowsum, state, row_offset):
        D = ks.shape[1]  # sc / sc_all are views whose first D columns are the centred features
        sc, sc_all = sc[:, :D], sc_all[:, :D]
        g = -(sc.double() * rowsum.double()[:, None] - ks.double()) / state["h"] ** 2
        if state["coef"] != 0.0:
            for a, b in state["ties"]:
                i = a - row_offset
                if 0 <= i < g.shape[0]:
                    g[i] += state["coef"] / state["n_ties_all"] * (sc[i].double() - sc_all[b].double())
        return g.float()

    def sig_bwd(self, x, sig, g, depth, add, alpha, beta):
        vjp = oracle.signature_bwd_np(x.numpy().astype(np.float64), sig.numpy().astype(np.float64),
                                      g.numpy().astype(np.float64), depth, True)
        M, T, d = x.shape
        out = beta * vjp
        if add is not None:
            out = out + alpha * add[:, :T * d].double().numpy().reshape(M, T, d)
        return torch.from_numpy(out.astype(np.float32))
