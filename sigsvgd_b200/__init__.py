"""sigsvgd_b200 -- B200 (sm_100a) implementation of lubaroli/sigsvgd's signature-kernel SVGD
update behind the reference's own Python API.

    from sigsvgd_b200.kernels import PathSigKernel, GaussianKernel     # mirrors src/kernels
    from sigsvgd_b200.inference import SVGD, ScoreEstimator            # mirrors src/inference
    import sigsvgd_b200.compat as compat; compat.install()            # signatory / sigkernel stand-ins

Importing the package does not load the CUDA library; the first operator call does, and raises
``RuntimeError`` if ``libsigsvgd_b200.so`` has not been built (there is no fallback path).
"""
__version__ = "0.1.0"

from . import _lib  # noqa: F401


def build(verbose: bool = False) -> str:
    """Compile ``libsigsvgd_b200.so`` in-tree with nvcc for sm_100a.  Returns the path."""
    from ._build import build as _b

    return _b(verbose=verbose)


def lib_path() -> str:
    return _lib.LIB_PATH
