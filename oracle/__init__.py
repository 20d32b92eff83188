"""CPU oracle for the distillation tail (TEST INFRASTRUCTURE, not product code).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import this package.  The shipped path
(``mdistiller_ddp_b200``) never does: it calls the sm_100a kernels through the
C-ABI and raises if the library is missing.

Parity status: the reference ships no tests or golden vectors for this path
(SURVEY.md section 4), so the oracle is pinned against outputs of the reference
itself, generated in the build container by ``tests/golden/make_golden.py``
(which imports ``/root/reference`` through :mod:`oracle.ref_loader`) and
committed under ``tests/golden/``.
"""

from .tail_oracle import (  # noqa: F401
    dkd_loss,
    kd_loss,
    ce_loss,
    loss_and_grad,
    dkd_closed_form,
    kd_closed_form,
    ce_closed_form,
    dot_lists_step,
    DotListOracle,
    dot_flat_step,
    sgd_flat_step,
    adam_flat_step,
    clip_coefficient,
    max_rel_err,
)
