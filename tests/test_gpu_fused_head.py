"""GPU parity of the head kernels with the criterion folded in (pdt_head_fwd_nll / pdt_head_bwd_nll) against
the separate launches they replace (pdt_head_fwd + pdt_nll_metrics_fwd, pdt_nll_bwd + pdt_head_bwd) and against
the oracle.  Reference call sites: trainer/trainer.py:49-50,57; model/loss.py:4-5; model/metric.py:4-20."""
import pytest
import torch

from oracle import ref_step as R
from util_gpu import close, make_model, max_err

pytestmark = pytest.mark.gpu


def _grads(model):
    return [p.grad.detach().clone() for p in model._ordered]


def _run(model, x, t, masks, fused, extra=None):
    from model.loss import nll_loss
    model.train()
    model.zero_grad()
    model.inject_dropout_masks([masks])
    if fused:
        model.expect_target(t)
    out = model(x)
    assert (getattr(out, "_pdt_fused", None) is not None) == fused
    loss = nll_loss(out, t)
    total = loss if extra is None else loss + extra(out)
    total.backward()
    stats = loss._pdt_stats.cpu().clone()
    return out.detach().clone(), loss.detach().clone(), stats, _grads(model)


@pytest.mark.parametrize("n", [1, 31, 37, 64, 257, 1024, 4099])
def test_fused_equals_separate_kernels(n):
    params = R.init_params(0)
    x, t, k2, k1 = R.synthetic_batch(n, 900 + n)
    t = t.clone()
    if n > 8:
        t[3] = -100            # an ignored row: counted by the metrics' denominators, not by the loss
        t[n - 1] = -100
    xd, td = x.cuda(), t.cuda()
    model = make_model(params)
    out_a, loss_a, stats_a, g_a = _run(model, xd, td, (k2, k1), fused=False)
    out_b, loss_b, stats_b, g_b = _run(model, xd, td, (k2, k1), fused=True)
    assert torch.equal(out_a, out_b)
    # integer record (top1, topk, n_valid, n_rows) bit-exact; the loss differs only by summation order
    assert torch.equal(stats_a[2:6], stats_b[2:6])
    assert abs(loss_a.item() - loss_b.item()) <= 2e-6 * abs(loss_a.item())
    for a, b in zip(g_a, g_b):
        assert torch.equal(a, b)


@pytest.mark.parametrize("n", [37, 256])
def test_fused_matches_oracle(n):
    params = R.init_params(0)
    x, t, k2, k1 = R.synthetic_batch(n, 950 + n)
    model = make_model(params)
    _, loss, stats, grads = _run(model, x.cuda(), t.cuda(), (k2, k1), fused=True)
    ref_loss, ref_out, ref_g = R.loss_and_grads(params, x, t, k2, k1)
    assert abs(loss.item() - float(ref_loss)) <= 1e-5 * abs(float(ref_loss))
    top1, top3 = R.accuracy_count(ref_out, t), R.top_k_count(ref_out, t, 3)
    assert int(stats[2]) == top1 and int(stats[3]) == top3 and int(stats[4]) == n and int(stats[5]) == n
    for (k, _), g in zip(R.PARAM_SHAPES, grads):
        assert close(g, ref_g[k]), (k, max_err(g, ref_g[k]))


def test_output_with_a_second_consumer():
    """The announced contract (the loss is the output's only consumer) is not silently relied on: when the output
    also feeds something else, backward forms the criterion's gradient explicitly and adds it."""
    n = 64
    params = R.init_params(0)
    x, t, k2, k1 = R.synthetic_batch(n, 77)
    xd, td = x.cuda(), t.cuda()
    model = make_model(params)
    extra = lambda out: 0.01 * (out * out).sum()
    _, _, _, g_a = _run(model, xd, td, (k2, k1), fused=False, extra=extra)
    _, _, _, g_b = _run(model, xd, td, (k2, k1), fused=True, extra=extra)
    for a, b in zip(g_a, g_b):
        assert close(b, a, rtol=1e-6)


def test_announcement_is_one_shot_and_checked():
    from model.loss import nll_loss
    params = R.init_params(0)
    x, t, k2, k1 = R.synthetic_batch(32, 5)
    model = make_model(params)
    model.eval()
    td = t.cuda()
    model.expect_target(td)
    with torch.no_grad():
        out = model(x.cuda())
        assert getattr(out, "_pdt_fused", None) is not None
        loss = nll_loss(out, td)
        assert loss._pdt_stats is out._pdt_fused[1]          # evaluated by the head kernel, no loss kernel
        out2 = model(x.cuda())                       # the announcement was consumed by the first forward
        assert getattr(out2, "_pdt_fused", None) is None
        ref = nll_loss(out2, t.cuda())
        assert abs(loss.item() - ref.item()) <= 2e-6 * abs(ref.item())
        # a different target object than the one announced: the separate loss kernel runs on the real target
        model.expect_target(t.cuda())
        out3 = model(x.cuda())
        other = torch.zeros_like(t).cuda()
        l3 = nll_loss(out3, other)
        want = -out3[:, 0].mean()
        assert abs(l3.item() - want.item()) <= 1e-5 * abs(want.item())
        # a target of the wrong length is ignored
        model.expect_target(t.cuda()[:7].contiguous())
        out4 = model(x.cuda())
        assert getattr(out4, "_pdt_fused", None) is None
