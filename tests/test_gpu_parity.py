"""GPU parity: the CUDA path (through the C ABI) against the oracle and the golden fixtures.
Tolerance (north_star): metric counts bit-exact; loss, gradients, weights within rtol 1e-5 with
atol = 1e-5 * max|ref| per tensor."""
import os

import numpy as np
import pytest
import torch

from oracle import ref_step as R
from util_gpu import close, flat_from_params, grad_matches, lane_to_rows, make_model, max_err, params_from_flat

pytestmark = pytest.mark.gpu


def _dev():
    return torch.device("cuda")


@pytest.mark.parametrize("n", [1, 31, 32, 37, 64, 100, 128, 257, 1024, 4096])     # 128/1024/4096: BASELINE.json configs
@pytest.mark.parametrize("train", [False, True])
def test_forward_matches_oracle(n, train):
    from pdt import ops
    params = R.init_params(0)
    x, t, k2, k1 = R.synthetic_batch(n, 100 + n)
    flat = flat_from_params(params, _dev())
    kk = (k2.cuda(), k1.cuda()) if train else (None, None)
    logp, saved = ops.mnist_forward_raw(x.cuda(), flat, kk[0], kk[1], True)
    ref = R.mnist_forward(params, x, k2 if train else None, k1 if train else None)
    assert close(logp, ref), max_err(logp, ref)
    assert tuple(logp.shape) == (n, 10)


@pytest.mark.parametrize("n", [1, 37, 128, 1024])
def test_loss_and_metric_counts(n):
    from pdt import ops
    g = torch.Generator().manual_seed(n)
    out = torch.log_softmax(torch.randn(n, 10, generator=g), dim=1)
    t = torch.randint(0, 10, (n,), generator=g)
    assert R.rows_are_tie_free(out, 5)
    loss, stats = ops.nll_loss_with_stats(out.cuda(), t.cuda(), topk=3)
    st = ops.unpack_stats(stats)
    assert st["top1"] == R.accuracy_count(out, t)
    assert st["topk"] == R.top_k_count(out, t, 3)
    assert st["n_valid"] == n and st["n_rows"] == n
    ref = float(R.nll_loss(out, t))
    assert abs(loss.item() - ref) <= 1e-5 * abs(ref)


def test_loss_ignore_index_and_ties():
    from pdt import ops
    from model.metric import accuracy, top_k_acc
    g = torch.Generator().manual_seed(5)
    out = torch.log_softmax(torch.randn(50, 10, generator=g), dim=1)
    t = torch.randint(0, 10, (50,), generator=g)
    t[::7] = -100
    loss, stats = ops.nll_loss_with_stats(out.cuda(), t.cuda())
    ref = float(R.nll_loss(out, t))
    assert abs(loss.item() - ref) <= 1e-5 * abs(ref)
    assert ops.unpack_stats(stats)["n_valid"] == int((t != -100).sum())
    # tie rules
    tie = torch.tensor([[1., 3., 3., 0., 3., 3., 0., 0., 0., 0.], [0.] * 10]).cuda()
    assert accuracy(tie, torch.tensor([1, 0]).cuda()) == 1.0
    assert accuracy(tie, torch.tensor([2, 1]).cuda()) == 0.0
    assert top_k_acc(tie, torch.tensor([4, 2]).cuda(), k=3) == 1.0
    assert top_k_acc(tie, torch.tensor([5, 3]).cuda(), k=3) == 0.0


def test_metrics_match_golden(golden_dir):
    from model.metric import accuracy, top_k_acc
    from model.loss import nll_loss
    g = np.load(os.path.join(golden_dir, "ref_eval_b50.npz"))
    out, t = torch.from_numpy(g["out"]).cuda(), torch.from_numpy(g["t"]).cuda()
    assert accuracy(out, t) == float(g["accuracy"])
    assert top_k_acc(out, t) == float(g["top_k_acc"])
    assert top_k_acc(out, t, k=5) == float(g["top_k_acc_k5"])
    assert abs(nll_loss(out, t).item() - float(g["loss"])) < 1e-5 * abs(float(g["loss"]))
    t_ign = torch.from_numpy(g["t_ign"]).cuda()
    assert abs(nll_loss(out, t_ign).item() - float(g["loss_ign"])) < 1e-5 * abs(float(g["loss_ign"]))


@pytest.mark.parametrize("n", [1, 37, 64, 128, 257, 1024, 4096])     # 128/1024/4096: BASELINE.json configs
def test_backward_matches_oracle(n):
    """Gradients against the fp32 oracle, and both against the fp64 truth."""
    from model.loss import nll_loss
    params = R.init_params(0)
    x, t, k2, k1 = R.robust_batch(n, 300 + n)        # no relu / arg-max decision within rounding noise of flipping
    model = make_model(params)
    model.train()
    model.inject_dropout_masks([(k2, k1)])
    out = model(x.cuda())
    loss = nll_loss(out, t.cuda())
    loss.backward()
    ref_loss, _, ref_g = R.loss_and_grads(params, x, t, k2, k1)
    p64 = {k: v.double() for k, v in params.items()}
    _, _, g64 = R.loss_and_grads(p64, x.double(), t, k2, k1)
    assert abs(loss.item() - float(ref_loss)) <= 1e-5 * abs(float(ref_loss))
    for (k, _), p in zip(R.PARAM_SHAPES, model._ordered):
        ok, msg = grad_matches(p.grad, ref_g[k], g64[k])
        assert ok, (k, msg)
        # ours is at least as close to the truth as 2x the oracle's own fp32 error + 1e-6
        assert max_err(p.grad, g64[k]) <= 2 * max_err(ref_g[k], g64[k]) + 1e-6, k


def check_adam_weights(sd, g, steps, opt64, prefix_keys=("x", "t", "keep2d", "keep1d")):
    """Adam divides by sqrt(v)+eps (eps 1e-8), so an element whose gradient is ~1e-7 in some step
    turns a 1e-9 absolute gradient difference into a ~1e-6 weight difference: the REFERENCE's own
    fp32 run sits 1.9e-5 (relative to max|w|) from the fp64 truth on such elements (measured, see
    DESIGN.md "Tolerances").  Hence:
    (1) rtol 1e-5 (atol 1e-5*max|ref|) on every element whose gradient was exactly 0 or above 1e-6 in
        every step (99.8 % of them);
    (2) the remaining, ill-conditioned elements are few (< 1 %), and on them this run must be as close to
        the fp64 truth as the reference's own fp32 run is: per tensor, mean|ours - fp64| <= 4 x mean|ref_fp32 -
        fp64| (+ the same 1e-5*max|w| floor as (1)).  The bound is per tensor, not per element, because the two
        fp32 runs' rounding noise on such an element is independent: where the reference happens to land on the
        truth, a per-element multiple of its error would demand exactness."""
    xk, tk, k2k, k1k = prefix_keys
    p64 = {k: torch.from_numpy(g["init/" + k]).double() for k, _ in R.PARAM_SHAPES}
    well = {k: torch.ones_like(v, dtype=torch.bool) for k, v in p64.items()}
    for s in range(steps):
        _, _, g64 = R.loss_and_grads(p64, torch.from_numpy(g[f"{xk}/{s}"]).double(), torch.from_numpy(g[f"{tk}/{s}"]),
                                     torch.from_numpy(g[f"{k2k}/{s}"]), torch.from_numpy(g[f"{k1k}/{s}"]))
        for k in well:
            well[k] &= (g64[k].abs() > 1e-6) | (g64[k] == 0)
        p64 = opt64.step(dict(p64), g64)
    n_ill = n_all = 0
    for k, _ in R.PARAM_SHAPES:
        ours, gold = sd[k].detach().cpu().double(), torch.from_numpy(g["final/" + k]).double()
        scale = float(gold.abs().max())
        m = well[k]
        diff = (ours - gold).abs()
        assert float(diff[m].max()) <= 1e-5 * scale + 1e-5 * float(gold[m].abs().max()), (k, "well-conditioned")
        if (~m).any():
            ours64 = float((ours - p64[k]).abs()[~m].mean())
            ref64 = float((gold - p64[k]).abs()[~m].mean())
            assert ours64 <= 4 * ref64 + 1e-5 * scale, (k, "ill-conditioned", ours64, ref64)
        n_ill += int((~m).sum())
        n_all += m.numel()
    assert n_ill < 0.01 * n_all, (n_ill, n_all)


def test_grad_accumulates_without_zero_grad():
    from model.loss import nll_loss
    params = R.init_params(0)
    x, t, k2, k1 = R.synthetic_batch(40, 7)
    model = make_model(params)
    model.train()
    model.inject_dropout_masks([(k2, k1), (k2, k1)])
    nll_loss(model(x.cuda()), t.cuda()).backward()
    g1 = model.flat.grad[:21840].clone()
    nll_loss(model(x.cuda()), t.cuda()).backward()
    assert close(model.flat.grad[:21840], 2 * g1)


@pytest.mark.parametrize("name,make", [
    ("ref_sgd_b37", lambda p, O: O.SGD(p, lr=0.01)),
    ("ref_sgdm_b32", lambda p, O: O.SGD(p, lr=0.05, momentum=0.9, weight_decay=1e-4, nesterov=True)),
    ("ref_adam_b64", lambda p, O: O.Adam(p, lr=1e-3, weight_decay=0, amsgrad=True)),
])
def test_train_steps_match_golden(golden_dir, name, make):
    """N steps of trainer.py:48-58 through the plug-in surface against the reference's own run."""
    from model.loss import nll_loss
    from pdt import optim as O
    g = np.load(os.path.join(golden_dir, name + ".npz"))
    params = {k: torch.from_numpy(g["init/" + k]) for k, _ in R.PARAM_SHAPES}
    model = make_model(params)
    opt = make(model.parameters(), O)
    model.train()
    steps = int(g["steps"])
    for s in range(steps):
        model.inject_dropout_masks([(torch.from_numpy(g[f"keep2d/{s}"]), torch.from_numpy(g[f"keep1d/{s}"]))])
        opt.zero_grad()
        out = model(torch.from_numpy(g[f"x/{s}"]).cuda())
        loss = nll_loss(out, torch.from_numpy(g[f"t/{s}"]).cuda())
        loss.backward()
        assert close(out, g[f"out/{s}"]), (s, max_err(out, g[f"out/{s}"]))
        assert abs(loss.item() - float(g[f"loss/{s}"])) <= 1e-5 * abs(float(g[f"loss/{s}"]))
        for (k, _), p in zip(R.PARAM_SHAPES, model._ordered):
            assert close(p.grad, g[f"grad/{s}/{k}"]), (s, k, max_err(p.grad, g[f"grad/{s}/{k}"]))
        opt.step()
    sd = model.state_dict()
    if "adam" in name:
        check_adam_weights(sd, g, steps, R.Adam(lr=1e-3, weight_decay=0.0, amsgrad=True))
    else:
        for k, _ in R.PARAM_SHAPES:
            assert close(sd[k], g["final/" + k]), (k, max_err(sd[k], g["final/" + k]))
    assert opt.device_step() == steps


@pytest.mark.parametrize("kind", ["sgd", "sgdm", "adam", "adam_ams_wd"])
def test_optimizer_kernels_match_oracle(kind):
    from pdt import optim as O
    torch.manual_seed(3)
    w = [torch.nn.Parameter(torch.randn(1000, device="cuda")), torch.nn.Parameter(torch.randn(37, 5, device="cuda"))]
    ref = {str(i): p.detach().cpu().clone() for i, p in enumerate(w)}
    if kind == "sgd":
        opt, ro = O.SGD(w, lr=0.1), R.SGD(lr=0.1)
    elif kind == "sgdm":
        opt, ro = O.SGD(w, lr=0.1, momentum=0.8, dampening=0.1, weight_decay=0.01), \
            R.SGD(lr=0.1, momentum=0.8, dampening=0.1, weight_decay=0.01)
    elif kind == "adam":
        opt, ro = O.Adam(w, lr=1e-2), R.Adam(lr=1e-2)
    else:
        opt, ro = O.Adam(w, lr=1e-2, weight_decay=0.01, amsgrad=True), R.Adam(lr=1e-2, weight_decay=0.01, amsgrad=True)
    for step in range(4):
        grads = {str(i): torch.randn(p.shape, generator=torch.Generator().manual_seed(10 * step + i))
                 for i, p in enumerate(w)}
        for i, p in enumerate(w):
            p.grad.copy_(grads[str(i)])
        opt.step()
        ref = ro.step(dict(ref), grads)
    for i, p in enumerate(w):
        assert close(p, ref[str(i)]), (kind, i, max_err(p, ref[str(i)]))


def test_lr_scheduler_reaches_the_kernel():
    from pdt import optim as O
    w = [torch.nn.Parameter(torch.ones(64, device="cuda"))]
    opt = O.SGD(w, lr=1.0)
    sched = torch.optim.lr_scheduler.StepLR(opt, step_size=1, gamma=0.1)
    w[0].grad.fill_(1.0)
    opt.step()
    sched.step()
    opt.step()
    assert torch.allclose(w[0].detach().cpu(), torch.full((64,), 1.0 - 1.0 - 0.1))


@pytest.mark.parametrize("count", [4, 21844, 1000003])
def test_flat_scale_cast(count):
    """pdt_flat_scale_cast (the DDP bucket's divide-by-world, reference train.py:46-52): fp32 in place and out
    of place, and fp32 -> bf16, bit-exact against torch (x * s, then round-to-nearest-even to bf16)."""
    from pdt.native import call, stream
    g = torch.Generator().manual_seed(count)
    src = (torch.randn(count, generator=g) * 3).cuda()
    for scale in (1.0 / 8, 0.3):
        ref32 = src * scale
        out = torch.empty_like(src)
        call("pdt_flat_scale_cast", src.data_ptr(), out.data_ptr(), count, scale, 0, stream())
        assert torch.equal(out, ref32)
        inplace = src.clone()
        call("pdt_flat_scale_cast", inplace.data_ptr(), inplace.data_ptr(), count, scale, 0, stream())
        assert torch.equal(inplace, ref32)
        out16 = torch.empty(count, dtype=torch.bfloat16, device="cuda")
        call("pdt_flat_scale_cast", src.data_ptr(), out16.data_ptr(), count, scale, 1, stream())
        assert torch.equal(out16, ref32.to(torch.bfloat16))
    # special values survive the bf16 rounding path
    edge = torch.tensor([0.0, -0.0, float("inf"), -float("inf"), 1.0 + 2 ** -8, 1.0 + 3 * 2 ** -9, 3.3895e38, 1e-40],
                        device="cuda")
    out16 = torch.empty(8, dtype=torch.bfloat16, device="cuda")
    call("pdt_flat_scale_cast", edge.data_ptr(), out16.data_ptr(), 8, 1.0, 1, stream())
    assert torch.equal(out16.view(torch.int16), edge.to(torch.bfloat16).view(torch.int16))


def test_dropout_mask_statistics():
    """Native RNG: keep-rate 0.5, Dropout2d acts per (sample, channel), fresh masks per call."""
    model = make_model()
    model.train()
    x = torch.randn(4096, 1, 28, 28, device="cuda")
    model(x)
    k2a, k1a = [m.clone() for m in model.last_masks]
    model(x)
    k2b, k1b = model.last_masks
    for k in (k2a, k1a):
        assert set(k.unique().tolist()) <= {0, 1}
        assert abs(k.float().mean().item() - 0.5) < 0.01
    assert (k2a != k2b).float().mean().item() > 0.4 and (k1a != k1b).float().mean().item() > 0.4
    model.eval()
    a, b = model(x[:64]), model(x[:64])
    assert torch.equal(a, b)


def test_state_dict_roundtrip_and_names():
    model = make_model()
    sd = model.state_dict()
    assert list(sd.keys()) == [k for k, _ in R.PARAM_SHAPES]
    other = make_model(seed=5)
    other.load_state_dict(sd)
    assert torch.equal(other.flat.data[:21840], model.flat.data[:21840])
    for (k, shape), p in zip(R.PARAM_SHAPES, other._ordered):
        assert tuple(p.shape) == shape and p.data_ptr() >= other.flat.data.data_ptr()


@pytest.mark.parametrize("n", [1024, 4096])
def test_full_size_properties(n):
    """BASELINE sizes: gradient of a batch == mean of the gradients of its two halves (linearity),
    metric counts add up, and the fp32 oracle agrees on the loss."""
    from model.loss import nll_loss
    params = R.init_params(0)
    x, t, k2, k1 = R.synthetic_batch(n, 555)
    model = make_model(params)
    model.train()
    h = n // 2

    def run(sl):
        model.inject_dropout_masks([(k2[sl], k1[sl])])
        model.flat.zero_grad()
        out = model(x[sl].cuda())
        loss = nll_loss(out, t[sl].cuda())
        loss.backward()
        from pdt import ops
        return loss.item(), model.flat.grad[:21840].clone(), ops.unpack_stats(loss._pdt_stats)

    l_all, g_all, s_all = run(slice(0, n))
    l_a, g_a, s_a = run(slice(0, h))
    l_b, g_b, s_b = run(slice(h, n))
    assert close(g_all, 0.5 * (g_a + g_b)), max_err(g_all, 0.5 * (g_a + g_b))
    assert abs(l_all - 0.5 * (l_a + l_b)) < 1e-5 * abs(l_all)
    assert s_all["top1"] == s_a["top1"] + s_b["top1"] and s_all["topk"] == s_a["topk"] + s_b["topk"]
    ref = float(R.nll_loss(R.mnist_forward(params, x, k2, k1), t))
    assert abs(l_all - ref) < 1e-5 * abs(ref)
