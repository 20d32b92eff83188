"""Property tests of the CPU oracle itself (hypothesis; CPU only): the two independent statements of
the loss -- torch restatement under autograd vs numpy closed forms -- agree on random shapes and
hyper-parameters, and the losses obey the identities the kernels are later checked against."""

import numpy as np
import torch
from hypothesis import given, settings, strategies as st

import oracle


@settings(max_examples=40, deadline=None)
@given(B=st.integers(1, 9), C=st.integers(2, 33), T=st.sampled_from([1.0, 2.0, 4.0]),
       alpha=st.floats(0.1, 2.0), beta=st.floats(0.1, 8.0), seed=st.integers(0, 10_000), sigma=st.sampled_from([0.5, 3.0, 8.0]))
def test_restatement_and_closed_forms_agree(B, C, T, alpha, beta, seed, sigma):
    g = torch.Generator().manual_seed(seed)
    zs = (torch.randn(B, C, generator=g) * sigma).double()
    zt = (torch.randn(B, C, generator=g) * sigma).double()
    y = torch.randint(0, C, (B,), generator=g)
    r = oracle.loss_and_grad("dkd", zs, zt, y, alpha=alpha, beta=beta, temperature=T, dtype=torch.float64)
    l, gr = oracle.dkd_closed_form(zs.numpy(), zt.numpy(), y.numpy(), alpha, beta, T)
    assert abs(r["loss_kd"].item() - l) <= 1e-9 * max(1.0, abs(l))
    assert np.abs(r["grad_kd"].numpy() - gr).max() <= 1e-9 * max(1e-12, np.abs(gr).max())
    r = oracle.loss_and_grad("kd", zs, zt, y, temperature=T, dtype=torch.float64)
    l, gr = oracle.kd_closed_form(zs.numpy(), zt.numpy(), T)
    assert abs(r["loss_kd"].item() - l) <= 1e-9 * max(1.0, abs(l))
    assert np.abs(r["grad_kd"].numpy() - gr).max() <= 1e-9 * max(1e-12, np.abs(gr).max())
    l, gr = oracle.ce_closed_form(zs.numpy(), y.numpy())
    assert abs(r["loss_ce"].item() - l) <= 1e-9 * max(1.0, abs(l))
    assert np.abs(r["grad_ce"].numpy() - gr).max() <= 1e-9 * np.abs(gr).max()


@settings(max_examples=25, deadline=None)
@given(B=st.integers(1, 6), C=st.integers(2, 20), T=st.sampled_from([1.0, 4.0]), seed=st.integers(0, 10_000))
def test_loss_identities(B, C, T, seed):
    g = torch.Generator().manual_seed(seed)
    zs = (torch.randn(B, C, generator=g) * 3).double().numpy()
    zt = (torch.randn(B, C, generator=g) * 3).double().numpy()
    y = torch.randint(0, C, (B,), generator=g).numpy()
    # gradient rows sum to zero; a per-row shift of the student logits changes nothing
    l, gr = oracle.dkd_closed_form(zs, zt, y, 1.0, 8.0, T)
    assert np.abs(gr.sum(1)).max() <= 1e-12
    l2, gr2 = oracle.dkd_closed_form(zs + np.arange(B)[:, None], zt, y, 1.0, 8.0, T)
    assert abs(l - l2) <= 1e-10 * max(1.0, abs(l)) and np.abs(gr - gr2).max() <= 1e-10
    # identical nets: zero loss and gradient; KL is non-negative
    l0, g0 = oracle.kd_closed_form(zs, zs, T)
    assert abs(l0) <= 1e-12 and np.abs(g0).max() <= 1e-15
    assert oracle.kd_closed_form(zs, zt, T)[0] >= -1e-12
    # with two classes the decoupled loss at alpha = 1 is Hinton's KD (NCKD vanishes)
    if C == 2:
        assert abs(oracle.dkd_closed_form(zs, zt, y, 1.0, 5.0, T)[0] - oracle.kd_closed_form(zs, zt, T)[0]) <= 1e-10


@settings(max_examples=20, deadline=None)
@given(n=st.integers(1, 50), steps=st.integers(1, 6), wd=st.sampled_from([0.0, 5e-4]), damp=st.sampled_from([0.0, 0.1]),
       seed=st.integers(0, 1000))
def test_flat_recurrence_equals_list_form_fp64(n, steps, wd, damp, seed):
    g = torch.Generator().manual_seed(seed)
    p = torch.randn(n, generator=g, dtype=torch.float64)
    flat, vt, vk = p.clone(), torch.zeros(n, dtype=torch.float64), torch.zeros(n, dtype=torch.float64)
    opt = oracle.DotListOracle([p], lr=0.05, momentum=0.825, momentum_kd=0.975, dampening=damp, weight_decay=wd)
    for s in range(steps):
        gk, gt = torch.randn(n, generator=g, dtype=torch.float64), torch.randn(n, generator=g, dtype=torch.float64)
        opt.step_kd([gk])
        opt.step([gt])
        oracle.dot_flat_step(flat, gt, gk, vt, vk, lr=0.05, m_task=0.825, m_kd=0.975, weight_decay=wd,
                             dampening=damp, first_step=(s == 0))
    assert torch.equal(flat, p)
