"""GPU parity of the fused middle of the training step against the separate launches it replaces and against
the oracle.  Two levels: "nll" = the head kernels with the criterion folded in (pdt_head_fwd_nll /
pdt_head_bwd_nll; same arithmetic order as the separate kernels, so bit-exact), "train" = conv2 + head forward +
criterion + head backward as one cluster kernel (pdt_conv2_head_train; fc1 is summed per CTA feature set, so
equal within fp32 rounding).  Reference call sites: trainer/trainer.py:49-50,57; model/loss.py:4-5;
model/metric.py:4-20."""
import pytest
import torch

from oracle import ref_step as R
from util_gpu import close, grad_matches, make_model, max_err

pytestmark = pytest.mark.gpu


def _grads(model):
    return [p.grad.detach().clone() for p in model._ordered]


def _run(model, x, t, masks, fused, extra=None, root=None):
    """fused: False = separate kernels, "nll" = criterion inside the head kernels, "train" = one cluster kernel."""
    from model.loss import nll_loss
    from pdt import ops
    model.train()
    model.zero_grad()
    model.inject_dropout_masks([masks])
    old, old_tc = ops.FUSE_HEAD_TRAIN, ops.TRAIN_TC
    ops.FUSE_HEAD_TRAIN = fused in ("train", "train_tc")
    ops.TRAIN_TC = fused == "train_tc"           # the training kernel with conv2 on the tensor cores
    try:
        if fused:
            model.expect_target(t)
        out = model(x)
        assert (getattr(out, "_pdt_fused", None) is not None) == bool(fused)
        loss = nll_loss(out, t)
        total = loss if extra is None else loss + extra(out)
        total.backward(root)
        torch.cuda.synchronize()
    finally:
        ops.FUSE_HEAD_TRAIN, ops.TRAIN_TC = old, old_tc
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
    out_b, loss_b, stats_b, g_b = _run(model, xd, td, (k2, k1), fused="nll")
    assert torch.equal(out_a, out_b)
    # integer record (top1, topk, n_valid, n_rows) bit-exact; the loss differs only by summation order
    assert torch.equal(stats_a[2:6], stats_b[2:6])
    assert abs(loss_a.item() - loss_b.item()) <= 2e-6 * abs(loss_a.item())
    for a, b in zip(g_a, g_b):
        assert torch.equal(a, b)
    # the one-kernel path: fc1 sums its 320 products in a different grouping
    out_c, loss_c, stats_c, g_c = _run(model, xd, td, (k2, k1), fused="train")
    assert close(out_c, out_a, rtol=2e-6), max_err(out_c, out_a)
    assert torch.equal(stats_a[2:6], stats_c[2:6])
    assert abs(loss_a.item() - loss_c.item()) <= 2e-6 * abs(loss_a.item())
    for a, c in zip(g_a, g_c):
        assert close(c, a, rtol=1e-5), max_err(c, a)
    # the same kernel with conv2 on the tensor cores (split TF32: equal within fp32 rounding)
    out_d, loss_d, stats_d, g_d = _run(model, xd, td, (k2, k1), fused="train_tc")
    assert close(out_d, out_a, rtol=1e-5), max_err(out_d, out_a)
    assert abs(loss_a.item() - loss_d.item()) <= 1e-5 * abs(loss_a.item())
    assert torch.equal(stats_a[2:6], stats_d[2:6])
    # (its gradients are compared on well-conditioned inputs in test_fused_matches_oracle: these batches are not
    # screened, and a pooling arg-max decided by less than the 1e-6 the two conv2 arithmetics differ by may flip)


@pytest.mark.parametrize("n", [37, 128, 256, 1024, 4096])     # 128/1024/4096: BASELINE.json configs
@pytest.mark.parametrize("level", ["nll", "train", "train_tc"])
def test_fused_matches_oracle(n, level):
    params = R.init_params(0)
    # no relu / arg-max decision within rounding noise of flipping (the split-TF32 conv2 is ~1e-6 of max|p2| from the
    # fp32 one: a wider margin for it)
    x, t, k2, k1 = R.robust_batch(n, 950 + n, margin=2e-5 if level == "train_tc" else 2e-6)
    model = make_model(params)
    _, loss, stats, grads = _run(model, x.cuda(), t.cuda(), (k2, k1), fused=level)
    ref_loss, ref_out, ref_g = R.loss_and_grads(params, x, t, k2, k1)
    p64 = {k: v.double() for k, v in params.items()}
    _, _, g64 = R.loss_and_grads(p64, x.double(), t, k2, k1)
    for (k, _), g in zip(R.PARAM_SHAPES, grads):
        assert max_err(g, g64[k]) <= 2 * max_err(ref_g[k], g64[k]) + 1e-6, k
    assert abs(loss.item() - float(ref_loss)) <= 1e-5 * abs(float(ref_loss))
    top1, top3 = R.accuracy_count(ref_out, t), R.top_k_count(ref_out, t, 3)
    assert int(stats[2]) == top1 and int(stats[3]) == top3 and int(stats[4]) == n and int(stats[5]) == n
    for (k, _), g in zip(R.PARAM_SHAPES, grads):
        ok, msg = grad_matches(g, ref_g[k], g64[k])
        assert ok, (k, msg)


@pytest.mark.parametrize("level", ["nll", "train"])
def test_output_with_a_second_consumer(level):
    """The announced contract (the loss is the output's only consumer) is not silently relied on: when the output
    also feeds something else, backward forms the criterion's gradient explicitly and adds it."""
    n = 64
    params = R.init_params(0)
    x, t, k2, k1 = R.synthetic_batch(n, 77)
    xd, td = x.cuda(), t.cuda()
    model = make_model(params)
    extra = lambda out: 0.01 * (out * out).sum()
    _, _, _, g_a = _run(model, xd, td, (k2, k1), fused=False, extra=extra)
    _, _, _, g_b = _run(model, xd, td, (k2, k1), fused=level, extra=extra)
    for a, b in zip(g_a, g_b):
        assert close(b, a, rtol=1e-5), max_err(b, a)


@pytest.mark.parametrize("level", ["nll", "train"])
def test_root_gradient_other_than_one(level):
    """loss.backward(g): the one-kernel path forms its partials for g = 1 and the reduction applies g."""
    n = 100
    params = R.init_params(0)
    x, t, k2, k1 = R.synthetic_batch(n, 78)
    xd, td = x.cuda(), t.cuda()
    model = make_model(params)
    root = torch.tensor(-2.5, device="cuda")
    _, _, _, g_a = _run(model, xd, td, (k2, k1), fused=False, root=root)
    _, _, _, g_b = _run(model, xd, td, (k2, k1), fused=level, root=root)
    _, _, _, g_1 = _run(model, xd, td, (k2, k1), fused=False)
    for a, b, one in zip(g_a, g_b, g_1):
        assert close(b, a, rtol=1e-5), max_err(b, a)
        assert close(a, -2.5 * one, rtol=1e-5)


def test_two_forwards_before_backward():
    """The one-kernel path leaves its head backward in a per-batch-size workspace: a second forward of the same
    size overwrites it, and each backward must notice and recompute instead of consuming the other's result."""
    from model.loss import nll_loss
    n = 64
    params = R.init_params(0)
    xa, ta, k2a, k1a = R.synthetic_batch(n, 81)
    xb, tb, k2b, k1b = R.synthetic_batch(n, 82)
    model = make_model(params)
    _, _, _, g_a = _run(model, xa.cuda(), ta.cuda(), (k2a, k1a), fused=False)
    _, _, _, g_b = _run(model, xb.cuda(), tb.cuda(), (k2b, k1b), fused=False)
    model.train()
    model.zero_grad()
    model.inject_dropout_masks([(k2a, k1a), (k2b, k1b)])
    tad, tbd = ta.cuda(), tb.cuda()
    model.expect_target(tad)
    out_a = model(xa.cuda())
    model.expect_target(tbd)
    out_b = model(xb.cuda())
    la, lb = nll_loss(out_a, tad), nll_loss(out_b, tbd)
    la.backward()
    lb.backward()
    for p, a, b in zip(model._ordered, g_a, g_b):
        assert close(p.grad, a + b, rtol=1e-5), max_err(p.grad, a + b)


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


@pytest.mark.parametrize("level", ["nll", "train"])
def test_eval_mode_with_gradients(level):
    """model.eval() (no dropout) with gradients enabled through the fused paths: the mask pointers are NULL."""
    from model.loss import nll_loss
    from pdt import ops
    n = 70
    params = R.init_params(0)
    x, t, _, _ = R.synthetic_batch(n, 83)
    xd, td = x.cuda(), t.cuda()
    grads = {}
    for fused in (False, level):
        model = make_model(params)
        model.eval()
        old = ops.FUSE_HEAD_TRAIN
        ops.FUSE_HEAD_TRAIN = fused == "train"
        try:
            if fused:
                model.expect_target(td)
            loss = nll_loss(model(xd), td)
            loss.backward()
            torch.cuda.synchronize()
        finally:
            ops.FUSE_HEAD_TRAIN = old
        grads[fused] = [p.grad.detach().clone() for p in model._ordered]
    ref_loss, _, ref_g = R.loss_and_grads(params, x, t, None, None)
    for (k, _), a, b in zip(R.PARAM_SHAPES, grads[False], grads[level]):
        assert close(b, a, rtol=1e-5), (k, max_err(b, a))
        assert close(b, ref_g[k]), (k, max_err(b, ref_g[k]))
