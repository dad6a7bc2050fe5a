"""The drop-in boundary without a GPU: the shared library loads, exports every symbol that
include/sigsvgd_b200.h declares (and nothing in the header is missing from the ctypes binding),
argument validation returns error codes before any CUDA work, and the host mirror of the
reference interface keeps the reference's names / exceptions and refuses CPU tensors (there is no
CPU fallback).  No compute is launched here."""
import ctypes
import os
import re

import pytest
import torch

from sigsvgd_b200 import _build, _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "sigsvgd_b200.h")


@pytest.fixture(scope="module")
def lib():
    _build.build()  # nvcc cross-compiles for sm_100a without a GPU; no-op when up to date
    return _lib.load()


def _declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sigsvgd_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_agree(lib):
    declared = _declared_symbols()
    assert len(declared) >= 30
    assert sorted(_lib.EXPORTED_SYMBOLS) == declared
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in the header but not exported"


def test_header_cites_reference_for_every_group():
    text = open(HEADER).read()
    for cite in ("_traj_kernels.py:125-126", "math.py:69-86", "math.py:28-34", "_kernels.py:110", "svgd.py:82",
                 "svgd.py:36-44"):
        assert cite in text


def test_state_struct_layout_matches_header():
    text = open(HEADER).read()
    assert f"#define SIGSVGD_MAX_TIES {_lib.MAX_TIES}" in text and _lib.MAX_TIES == 64
    assert _lib.STATE_BYTES == 48 + 8 * _lib.MAX_TIES and _lib.OFF_SUM_KD == 16 and _lib.OFF_TIES == 48


def test_channels_and_version(lib):
    assert lib.sigsvgd_sig_channels(7, 3) == 399
    assert lib.sigsvgd_sig_channels(7, 4) == 2800
    assert lib.sigsvgd_sig_channels(2, 3) == 14
    assert lib.sigsvgd_sig_channels(0, 3) == -1 and lib.sigsvgd_sig_channels(3, 9) == -1
    assert b"sm_100a" in lib.sigsvgd_version()


def test_invalid_arguments_return_codes_not_crashes(lib):
    # NULL pointers / bad shapes are rejected on the host, before any CUDA call
    rc = lib.sigsvgd_phi(None, None, 1, 8, 4, 2, 3, 0, 1.0, 1.0, 1e-8, 0, None, None, None, None, None, None, 0, None)
    assert rc == 1 and b"NULL" in lib.sigsvgd_last_error()
    buf = ctypes.create_string_buffer(64)
    p = ctypes.addressof(buf)
    rc = lib.sigsvgd_phi(p, p, 0, 8, 4, 2, 3, 0, 1.0, 1.0, 1e-8, 0, p, None, None, None, p, p, 64, None)
    assert rc == 1 and b"bad shape" in lib.sigsvgd_last_error()
    rc = lib.sigsvgd_phi(p, p, 1, 8, 4, 2, 3, 7, 1.0, 1.0, 1e-8, 0, p, None, None, None, p, p, 64, None)
    assert rc == 1 and b"bw_mode" in lib.sigsvgd_last_error()
    assert lib.sigsvgd_signature_fwd(None, 4, 4, 2, 3, 1, None, None) == 1
    assert lib.sigsvgd_phi_workspace_bytes(0, 8, 4, 2, 3, 0) == 0
    need = lib.sigsvgd_phi_workspace_bytes(1, 4096, 64, 7, 3, 0)
    assert need > 4096 * 4096 * 4  # holds the distance matrix
    with pytest.raises(ValueError):
        _lib.check(1, "x")
    with pytest.raises(RuntimeError):
        _lib.check(4, "x")


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "libsigsvgd_b200.so"))
    with pytest.raises(RuntimeError, match="no CPU or eager-PyTorch fallback"):
        _lib.load()


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "sigsvgd_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert not re.search(r"^\s*(import|from)\s+oracle\b", src, flags=re.M), fn
    for fn in os.listdir(os.path.join(pkg, "csrc")):
        assert "oracle/" not in open(os.path.join(pkg, "csrc", fn)).read() or fn.endswith(".md")


# ---------------------------------------------------------------------------------------------
# host mirror of the reference interface (src/kernels, src/inference)
# ---------------------------------------------------------------------------------------------
def test_kernel_classes_keep_reference_conventions():
    from sigsvgd_b200.kernels import BaseKernel, GaussianKernel, PathSigKernel, bw_median

    k = PathSigKernel()
    assert isinstance(k, BaseKernel) and k.analytic_grad is False  # _traj_kernels.py:92
    assert isinstance(k.static_kernel, GaussianKernel) and k.static_kernel.analytic_grad is True
    assert k.static_kernel.get_bandwidth is bw_median  # _kernels.py:34-35
    with pytest.raises(ValueError, match="callable scalar function"):  # _kernels.py:38-42
        GaussianKernel(bandwidth_fn=3.0)
    X = torch.zeros(4, 5, 2)
    with pytest.raises(AssertionError, match="same dimensions"):  # _traj_kernels.py:120
        k(X, torch.zeros(3, 5, 2))
    with pytest.raises(AssertionError, match="same dimensions"):  # _kernels.py:102
        GaussianKernel()(torch.zeros(4, 3), torch.zeros(5, 3))


def test_cpu_tensors_are_refused_not_emulated():
    from sigsvgd_b200 import ops
    from sigsvgd_b200.inference import make_fused
    from tests import ref_mirror

    SVGD = make_fused(ref_mirror.SVGD)
    from sigsvgd_b200.kernels import PathSigKernel

    X, G = torch.zeros(4, 5, 2), torch.zeros(4, 10)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        PathSigKernel()(X, X, compute_grad=False)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.phi(X, G)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.signature(X, 3, basepoint=True)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        SVGD(kernel=PathSigKernel(), optimizer_class=None, lr=0.1)._velocity(X, G)
    with pytest.raises(ValueError, match="log probability"):  # svgd.py:49-53
        SVGD(kernel=PathSigKernel(), optimizer_class=None)._velocity(X, None)


def test_bandwidth_spec_and_fused_dispatch():
    from sigsvgd_b200 import ops
    from sigsvgd_b200.inference import FixedBandwidth, _fused_spec
    from tests.ref_mirror import ScoreEstimator
    from sigsvgd_b200.kernels import GaussianKernel, PathSigKernel

    assert _fused_spec(PathSigKernel()).mode == _lib.BW_MEDIAN
    fx = _fused_spec(PathSigKernel(static_kernel=GaussianKernel(bandwidth_fn=FixedBandwidth(1.5))))
    assert fx.mode == _lib.BW_FIXED and fx.h == 1.5 and fx.median_grad is False
    assert _fused_spec(PathSigKernel(static_kernel=GaussianKernel(bandwidth_fn=lambda _: 1.5))) is None
    assert _fused_spec(GaussianKernel()) is None
    with pytest.raises(ValueError):
        ops.Bandwidth.fixed(0.0)
    # ScoreEstimator dispatch (score.py:19-25)
    est = ScoreEstimator(PathSigKernel(), lambda x: (x.sum(), {}), {})
    assert est.score == est._svgd_ag_score
    est = ScoreEstimator(GaussianKernel(), lambda x: (x.sum(), {}), {})
    assert est.score == est._svgd_score


def test_compat_install_registers_stand_ins():
    import sys

    from sigsvgd_b200 import compat

    saved = {k: sys.modules.get(k) for k in ("signatory", "sigkernel")}
    try:
        assert set(compat.install(force=True)) == {"signatory"}  # `sigkernel` is opt-in: different mathematics
        with pytest.warns(UserWarning, match="NOT the Goursat-PDE"):
            done = compat.install(force=True, sigkernel_truncated=True)
        assert set(done) == {"signatory", "sigkernel"}
        import signatory
        import sigkernel

        assert signatory.signature_channels(7, 3) == 399
        k = sigkernel.SigKernel(sigkernel.RBFKernel(sigma=1.0), dyadic_order=0)
        assert isinstance(k, compat.SigKernel)
        with pytest.raises(NotImplementedError):
            signatory.signature(torch.zeros(1, 2, 2), 3, stream=True)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
