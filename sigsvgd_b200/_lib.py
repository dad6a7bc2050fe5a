"""ctypes binding of ``libsigsvgd_b200.so`` (C ABI in ``include/sigsvgd_b200.h``).

This is the stub a maintainer of the reference would add: plain pointers and
sizes, no torch types cross the boundary.  There is NO fallback: if the shared library is missing
or a call fails, a ``RuntimeError`` is raised.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_double, c_float, c_int, c_int64, c_size_t, c_uint32, c_uint64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsigsvgd_b200.so")

MAX_TIES = 64
BW_MEDIAN, BW_FIXED = 0, 1
FLAG_MEDIAN_GRAD, FLAG_NO_CENTER, FLAG_FORCE_SIMT, FLAG_FORCE_TC = 1, 2, 4, 8
OPT_SGD, OPT_ADAGRAD, OPT_ADAM = 0, 1, 2
PDE_RBF, PDE_LINEAR = 0, 1
KF_EXP, KF_IDENTITY = 0, 1


class BwState(ctypes.Structure):
    """Mirror of ``sigsvgd_bw_state`` (device-resident; this layout is used to decode copies)."""

    _fields_ = [
        ("median", c_float),
        ("h", c_float),
        ("gamma", c_float),
        ("coef", c_float),
        ("sum_kd", c_double),
        ("n_ties", c_uint32),
        ("n_ties_all", c_uint32),
        ("clamped", c_uint32),
        ("prefix", c_uint32),
        ("k_rem", c_uint64),
        ("ties", ctypes.c_int32 * (2 * MAX_TIES)),
    ]


STATE_BYTES = ctypes.sizeof(BwState)
assert STATE_BYTES == 48 + 8 * MAX_TIES, STATE_BYTES
# byte offsets used by the host code to address fields of a device copy
OFF_MEDIAN, OFF_H, OFF_GAMMA, OFF_COEF, OFF_SUM_KD, OFF_N_TIES, OFF_N_TIES_ALL, OFF_CLAMPED = 0, 4, 8, 12, 16, 24, 28, 32
OFF_TIES = BwState.ties.offset

_SIGNATURES = {
    # name: (restype, argtypes)
    "sigsvgd_last_error": (ctypes.c_char_p, []),
    "sigsvgd_version": (ctypes.c_char_p, []),
    "sigsvgd_sig_channels": (c_int64, [c_int64, c_int]),
    "sigsvgd_signature_fwd": (c_int, [c_void_p, c_int64, c_int64, c_int64, c_int, c_int, c_void_p, c_void_p]),
    "sigsvgd_signature_bwd": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int, c_int, c_void_p, c_int64, c_float, c_float,
         c_void_p, c_void_p],
    ),
    "sigsvgd_signature_prefix_floats": (c_int64, [c_int64, c_int64, c_int64, c_int]),
    "sigsvgd_signature_fwd_prefix": (
        c_int, [c_void_p, c_int64, c_int64, c_int64, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "sigsvgd_signature_bwd_prefix": (
        c_int,
        [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int, c_int, c_void_p, c_int64, c_float,
         c_float, c_void_p, c_void_p],
    ),
    "sigsvgd_sqdist": (
        c_int,
        [c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int64, c_int64, c_void_p, c_int64, c_uint32, c_void_p],
    ),
    "sigsvgd_select_workspace_bytes": (c_size_t, [c_int64]),
    "sigsvgd_median_bandwidth": (
        c_int, [c_void_p, c_int64, c_int64, c_float, c_float, c_void_p, c_void_p, c_size_t, c_void_p]),
    "sigsvgd_median_bandwidth_static": (
        c_int, [c_void_p, c_int64, c_int64, c_int64, c_int64, c_float, c_float, c_void_p, c_void_p]),
    "sigsvgd_fixed_bandwidth": (c_int, [c_float, c_void_p, c_void_p]),
    "sigsvgd_select_begin": (c_int, [c_void_p, c_void_p, c_uint64, c_void_p]),
    "sigsvgd_select_begin_window": (c_int, [c_void_p, c_void_p, c_uint64, c_void_p, c_void_p, c_void_p]),
    "sigsvgd_select_hist": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_void_p, c_void_p]),
    "sigsvgd_select_scan": (c_int, [c_int, c_void_p, c_void_p, c_void_p]),
    "sigsvgd_select_finish": (c_int, [c_int64, c_float, c_float, c_void_p, c_void_p]),
    "sigsvgd_rbf_apply": (
        c_int, [c_void_p, c_int64, c_int64, c_int64, c_int64, c_void_p, c_int64, c_void_p, c_void_p]),
    "sigsvgd_median_coef": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "sigsvgd_matmul_nn": (
        c_int,
        [c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int64, c_int64, c_void_p, c_int64, c_uint32, c_void_p],
    ),
    "sigsvgd_col_mean": (c_int, [c_void_p, c_int64, c_int64, c_int64, c_void_p, c_void_p, c_void_p]),
    "sigsvgd_pack_v": (
        c_int,
        [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int64, c_void_p, c_int64, c_int64, c_void_p],
    ),
    "sigsvgd_sig_grad": (
        c_int,
        [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int64, c_int64,
         c_void_p, c_void_p, c_int64, c_void_p],
    ),
    "sigsvgd_tc_available": (c_int, []),
    "sigsvgd_tc_scales": (
        c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int64, c_void_p, c_int64, c_int64, c_int64, c_void_p, c_void_p]),
    "sigsvgd_tc_pack": (
        c_int,
        [c_void_p, c_int64, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p,
         c_int64, c_int64, c_void_p, c_void_p, c_int64, c_void_p],
    ),
    "sigsvgd_tc_gram": (
        c_int,
        [c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int64,
         c_void_p, c_int, c_void_p, c_int64, c_void_p],
    ),
    "sigsvgd_tc_median_workspace_bytes": (c_size_t, [c_int64, c_int64]),
    "sigsvgd_tc_median_prepare": (
        c_int, [c_void_p, c_int64, c_int64, c_void_p, c_int64, c_int64, c_int64, c_void_p, c_size_t, c_void_p, c_void_p]),
    "sigsvgd_tc_median_select": (
        c_int,
        [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int, c_int64, c_float, c_float, c_void_p, c_void_p]),
    "sigsvgd_tc_kf": (
        c_int,
        [c_void_p, c_int64, c_int64, c_int64, c_int, c_void_p, c_int, c_int64, c_void_p, c_void_p, c_int64, c_int64,
         c_void_p, c_void_p, c_int64, c_void_p],
    ),
    "sigsvgd_rbf_out": (c_int, [c_void_p, c_int64, c_int64, c_int64, c_void_p, c_void_p, c_int64, c_void_p]),
    "sigsvgd_phi_workspace_bytes": (c_size_t, [c_int64, c_int64, c_int64, c_int64, c_int, c_int]),
    "sigsvgd_phi": (
        c_int,
        [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_int, c_int, c_float, c_float, c_float, c_uint32,
         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
    "sigsvgd_tc_pack_fused": (
        c_int,
        [c_void_p, c_int64, c_int64, c_void_p, c_int64, c_int64, c_int64, c_int, c_void_p, c_void_p, c_void_p, c_int64,
         c_int64, c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_void_p],
    ),
    "sigsvgd_tc_stats": (c_int, [c_void_p, c_int64, c_int64, c_int64, c_void_p, c_int64, c_int64, c_void_p, c_void_p]),
    "sigsvgd_tc_pack_from_stats": (
        c_int,
        [c_void_p, c_int64, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int64, c_int, c_void_p, c_void_p,
         c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_int64, c_void_p, c_void_p],
    ),
    "sigsvgd_sigpde_bandwidth": (
        c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_int64, c_float, c_float, c_void_p, c_void_p,
                c_void_p]),
    "sigsvgd_sigpde_gram": (
        c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_int64, c_int, c_int, c_double, c_void_p,
                c_void_p, c_void_p]),
    "sigsvgd_sigpde_bwd_workspace_bytes": (c_size_t, [c_int64, c_int64, c_int64, c_int64, c_int64, c_int]),
    "sigsvgd_sigpde_gram_bwd": (
        c_int, [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int64, c_int64, c_int, c_int, c_double, c_void_p,
                c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "sigsvgd_particle_rollout": (
        c_int, [c_void_p, c_void_p, c_void_p, c_float, c_int64, c_int64, c_float, c_float, c_float, c_void_p, c_int64,
                c_int64, c_float, c_float, c_float, c_int, c_void_p, c_void_p, c_void_p]),
    "sigsvgd_particle_rollout_bwd": (
        c_int, [c_void_p, c_float, c_int64, c_int64, c_float, c_void_p, c_void_p, c_void_p, c_void_p]),
    "sigsvgd_chain_fk": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_int64, c_void_p, c_void_p]),
    "sigsvgd_chain_fk_bwd": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_int64, c_void_p, c_void_p, c_void_p]),
    "sigsvgd_mlp_packed_bytes": (c_size_t, [c_int, c_int, c_int]),
    "sigsvgd_mlp_profile_enable": (c_int, [c_int]),
    "sigsvgd_mlp_profile_cycles": (c_int, [c_void_p, c_int]),
    "sigsvgd_mlp_pack": (
        c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_size_t,
                c_void_p]),
    "sigsvgd_mlp_query": (
        c_int, [c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_int64, c_void_p, c_void_p, c_void_p]),
    "sigsvgd_opt_step": (
        c_int, [c_void_p, c_void_p, c_float, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_float, c_double, c_double,
                c_float, c_float, c_float, c_int, c_void_p]),
    "sigsvgd_select_window_workspace_bytes": (c_size_t, []),
    "sigsvgd_select_window": (c_int, [c_void_p, c_int64, c_int64, c_int64, c_void_p, c_size_t, c_void_p, c_void_p]),
    "sigsvgd_select_collect_min_capacity": (c_int64, []),
    "sigsvgd_select_collect": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_void_p, c_void_p]),
    "sigsvgd_phi_host_workspace_bytes": (c_size_t, [c_int64, c_int64, c_int64, c_int]),
    "sigsvgd_phi_host": (
        c_int,
        [c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int, c_int, c_float, c_float, c_float, c_uint32,
         c_void_p, c_void_p, c_void_p, c_size_t, c_void_p],
    ),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)

_lib = None


def load() -> ctypes.CDLL:
    """Load the shared library (once).  Raises RuntimeError if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a).  sigsvgd_b200 has no CPU or eager-PyTorch fallback."
            )
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError if the .so does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def last_error() -> str:
    return load().sigsvgd_last_error().decode()


def check(rc: int, what: str) -> None:
    if rc != 0:
        kind = {1: "invalid argument", 2: "unsupported", 3: "workspace too small", 4: "CUDA error"}.get(rc, "error")
        msg = f"{what}: {kind}: {last_error()}"
        if rc == 1:
            raise ValueError(msg)
        raise RuntimeError(msg)


def call(name: str, *args) -> None:
    check(getattr(load(), name)(*args), name)
