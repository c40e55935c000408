"""GPU parity of conv2's data gradient with conv1's weight gradient taken in the same kernel
(pdt_conv2_bwd_data_conv1_bwd_weight) against the two separate launches and against the oracle.
Reference: autograd of model/model.py:16-17 as run by trainer/trainer.py:57."""
import pytest
import torch

from oracle import ref_step as R
from util_gpu import close, grad_matches, make_model, max_err

pytestmark = pytest.mark.gpu


def _grads(n, seed, fuse, robust=False):
    from model.loss import nll_loss
    from pdt import ops
    params = R.init_params(0)
    x, t, k2, k1 = (R.robust_batch if robust else R.synthetic_batch)(n, seed)
    model = make_model(params)
    model.train()
    model.inject_dropout_masks([(k2, k1)])
    old = ops.FUSE_DGRAD_WGRAD1
    ops.FUSE_DGRAD_WGRAD1 = fuse
    try:
        loss = nll_loss(model(x.cuda()), t.cuda())
        loss.backward()
        torch.cuda.synchronize()
    finally:
        ops.FUSE_DGRAD_WGRAD1 = old
    return [p.grad.detach().clone() for p in model._ordered], (params, x, t, k2, k1)


@pytest.mark.parametrize("n", [1, 31, 37, 64, 257, 1024, 4099])
def test_fused_equals_separate_kernels(n):
    g_sep, _ = _grads(n, 700 + n, False)
    g_fus, _ = _grads(n, 700 + n, True)
    # conv1 weight/bias: same products, summed by quadrant instead of by row band
    for a, b in zip(g_sep[:2], g_fus[:2]):
        assert close(b, a, rtol=2e-6), max_err(b, a)
    # everything else does not go through the fused part
    for a, b in zip(g_sep[2:], g_fus[2:]):
        assert torch.equal(a, b)


@pytest.mark.parametrize("n", [37, 128, 256, 1024, 4096])     # 128/1024/4096: BASELINE.json configs
def test_fused_matches_oracle(n):
    g_fus, (params, x, t, k2, k1) = _grads(n, 800 + n, True, robust=True)
    _, _, ref_g = R.loss_and_grads(params, x, t, k2, k1)
    p64 = {k: v.double() for k, v in params.items()}
    _, _, g64 = R.loss_and_grads(p64, x.double(), t, k2, k1)
    for (k, _), g in zip(R.PARAM_SHAPES, g_fus):
        ok, msg = grad_matches(g, ref_g[k], g64[k])
        assert ok, (k, msg)
        assert max_err(g, g64[k]) <= 2 * max_err(ref_g[k], g64[k]) + 1e-6, k
