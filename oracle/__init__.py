"""CPU oracle for the signature-kernel SVGD phi-update.  TEST INFRASTRUCTURE ONLY.

Nothing under ``sigsvgd_b200/`` imports this package.  Allowed users: ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs.

Parity status
-------------
* signature transform (``signatory.signature``, third party, absent): **parity unpinned** --
  pinned only by closed-form known-answer tests, see ``sig_oracle.c`` header.
* everything downstream of the signature (``pw_dist_sq``, ``bw_median``, ``GaussianKernel``,
  ``PathSigKernel``, ``SVGD._compute_kernel/_velocity``): pinned by golden vectors generated in
  the build container by running the reference's own *unmodified* Python
  (``oracle/make_golden.py`` -> ``tests/golden/*.npz``).
* signature-PDE kernel (``sigkernel`` @ 3b23739, third party, absent): **parity unpinned** for the Goursat
  solver and its backward (closed-form known answers, ``sigpde_oracle.py``); the static kernel and the
  bandwidth convention are the reference's own code and are pinned by ``tests/golden/sigpde_static.npz``
  (``oracle/make_golden_sigpde.py``).
"""
from .signature import signature_np, signature_bwd_np, sig_channels, OracleSignature, signature_torch  # noqa: F401
from .phi_oracle import (  # noqa: F401
    pw_dist_sq,
    bw_median,
    lower_median,
    gaussian_gram,
    sig_space_grad,
    phi_closed_form,
    phi_torch_literal,
)
