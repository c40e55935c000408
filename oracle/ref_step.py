"""CPU oracle for the data-parallel training step.  TEST INFRASTRUCTURE ONLY.

This file restates, in plain torch CPU ops, the arithmetic of the reference's hot path
(`Trainer._train_epoch` inner loop, /root/reference/trainer/trainer.py:45-72).  It exists
so that the CUDA path can be checked on the GPU box, where /root/reference is absent.
Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference`
legs may import it; the product (the package next to this directory) never does.

Where the arithmetic lives: the reference delegates every FLOP to the third-party
dependency `torch` (requirements.txt:1, `torch>=1.1`, unpinned; installed 2.11.0).  The
restatement therefore calls the same torch CPU primitives in the order the reference's
call sites do, and adds the one thing the reference cannot offer: *injected* dropout
masks, so both sides of a parity test drop the same units.

Pinning: tests/test_oracle_golden.py checks every function here against fixtures in
tests/golden/ that oracle/gen_golden.py produced by running the UNMODIFIED reference
classes (model.model.MnistModel, model.loss.nll_loss, model.metric.*, torch.optim) imported
from /root/reference under torch 2.11.0 CPU.  The reference ships no tests or golden
vectors of its own, so that run is the pin.
"""
import math

import torch
import torch.nn.functional as F

PARAM_SHAPES = (
    ("conv1.weight", (10, 1, 5, 5)),
    ("conv1.bias", (10,)),
    ("conv2.weight", (20, 10, 5, 5)),
    ("conv2.bias", (20,)),
    ("fc1.weight", (50, 320)),
    ("fc1.bias", (50,)),
    ("fc2.weight", (10, 50)),
    ("fc2.bias", (10,)),
)
NUM_PARAMS = sum(math.prod(s) for _, s in PARAM_SHAPES)  # 21,840


def init_params(seed=0, dtype=torch.float32):
    """Same initial weights as `torch.manual_seed(seed); MnistModel()` in the reference
    (model/model.py:7-13): the five layers are built in the reference's order so the
    Kaiming-uniform draws consume the global RNG identically."""
    torch.manual_seed(seed)
    conv1 = torch.nn.Conv2d(1, 10, kernel_size=5)
    conv2 = torch.nn.Conv2d(10, 20, kernel_size=5)
    torch.nn.Dropout2d()
    fc1 = torch.nn.Linear(320, 50)
    fc2 = torch.nn.Linear(50, 10)
    mods = {"conv1": conv1, "conv2": conv2, "fc1": fc1, "fc2": fc2}
    out = {}
    for name, _ in PARAM_SHAPES:
        m, attr = name.split(".")
        out[name] = getattr(mods[m], attr).detach().clone().to(dtype)
    return out


def mnist_forward(params, x, keep2d=None, keep1d=None):
    """model/model.py:15-22, op for op.  `keep2d` [N,20] and `keep1d` [N,50] are 0/1 keep
    masks standing in for Dropout2d (model.py:11,17) and F.dropout (model.py:20), both
    p=0.5, i.e. kept units are scaled by 2.  None means eval mode (identity)."""
    p = params
    h = F.relu(F.max_pool2d(F.conv2d(x, p["conv1.weight"], p["conv1.bias"]), 2))
    h = F.conv2d(h, p["conv2.weight"], p["conv2.bias"])
    if keep2d is not None:
        h = h * (keep2d.to(h.dtype) * 2.0)[:, :, None, None]
    h = F.relu(F.max_pool2d(h, 2))
    h = h.view(-1, 320)
    h = F.relu(F.linear(h, p["fc1.weight"], p["fc1.bias"]))
    if keep1d is not None:
        h = h * (keep1d.to(h.dtype) * 2.0)
    h = F.linear(h, p["fc2.weight"], p["fc2.bias"])
    return F.log_softmax(h, dim=1)


def nll_loss(output, target, ignore_index=-100):
    """model/loss.py:4-5 -> F.nll_loss defaults: mean over rows whose target is not
    `ignore_index`."""
    valid = target != ignore_index
    safe = torch.where(valid, target, torch.zeros_like(target))
    picked = output.gather(1, safe[:, None])[:, 0]
    picked = torch.where(valid, picked, torch.zeros_like(picked))
    return -(picked.sum() / valid.sum().to(output.dtype))


def accuracy_count(output, target):
    """model/metric.py:4-10 without the final divide: number of rows whose argmax (first
    maximal index on ties) equals the target."""
    pred = torch.argmax(output, dim=1)
    return int(torch.sum(pred == target).item())


def top_k_count(output, target, k=3):
    """model/metric.py:13-20 without the final divide, in the tie-free form SURVEY.md A.3
    pins: a row is correct iff fewer than k classes rank ahead of the target, where class j
    ranks ahead when o_j > o_t, or o_j == o_t and j < t.  (torch.topk's own tie order is
    implementation defined; parity inputs are asserted tie-free.)"""
    n, c = output.shape
    t = target.clamp(0, c - 1)
    ot = output.gather(1, t[:, None])
    j = torch.arange(c)[None, :]
    ahead = (output > ot) | ((output == ot) & (j < t[:, None]))
    ok = (ahead.sum(1) < k) & (target >= 0) & (target < c)
    return int(ok.sum().item())


def accuracy(output, target):
    return accuracy_count(output, target) / len(target)


def top_k_acc(output, target, k=3):
    return top_k_count(output, target, k) / len(target)


def rows_are_tie_free(output, k=3):
    """True when no row has a duplicate among its top-(k+1) values."""
    top = torch.topk(output, min(k + 1, output.shape[1]), dim=1)[0]
    return bool((top[:, 1:] != top[:, :-1]).all())


def loss_and_grads(params, x, target, keep2d=None, keep1d=None):
    """trainer/trainer.py:49-50,57: forward, criterion, backward.  Returns
    (loss, output, grads) with grads keyed like PARAM_SHAPES."""
    leaf = {k: v.detach().clone().requires_grad_(True) for k, v in params.items()}
    out = mnist_forward(leaf, x, keep2d, keep1d)
    loss = nll_loss(out, target)
    loss.backward()
    return loss.detach(), out.detach(), {k: v.grad for k, v in leaf.items()}


class SGD:
    """torch/optim/sgd.py (_single_tensor_sgd): weight decay, momentum, dampening,
    nesterov.  The reference selects it by name (train.py:42)."""

    def __init__(self, lr, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False):
        self.lr, self.momentum, self.dampening = lr, momentum, dampening
        self.weight_decay, self.nesterov = weight_decay, nesterov
        self.buf = {}

    def step(self, params, grads):
        for k, p in params.items():
            g = grads[k]
            if self.weight_decay != 0:
                g = g + self.weight_decay * p
            if self.momentum != 0:
                if k not in self.buf:
                    self.buf[k] = g.clone()
                else:
                    self.buf[k] = self.buf[k] * self.momentum + (1 - self.dampening) * g
                g = g + self.momentum * self.buf[k] if self.nesterov else self.buf[k]
            params[k] = p - self.lr * g
        return params


class Adam:
    """torch/optim/adam.py (_single_tensor_adam), incl. amsgrad -- the template default
    (config/config.json:38-45: lr 1e-3, weight_decay 0, amsgrad true)."""

    def __init__(self, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, amsgrad=False):
        self.lr, self.betas, self.eps = lr, betas, eps
        self.weight_decay, self.amsgrad = weight_decay, amsgrad
        self.t = 0
        self.m, self.v, self.vmax = {}, {}, {}

    def step(self, params, grads):
        b1, b2 = self.betas
        self.t += 1
        bc1 = 1 - b1 ** self.t
        bc2 = 1 - b2 ** self.t
        step_size = self.lr / bc1
        bc2_sqrt = math.sqrt(bc2)
        for k, p in params.items():
            g = grads[k]
            if self.weight_decay != 0:
                g = g + self.weight_decay * p
            if k not in self.m:
                self.m[k] = torch.zeros_like(p)
                self.v[k] = torch.zeros_like(p)
                self.vmax[k] = torch.zeros_like(p)
            self.m[k] = self.m[k] + (g - self.m[k]) * (1 - b1)  # lerp_
            self.v[k] = self.v[k] * b2 + (1 - b2) * g * g
            if self.amsgrad:
                self.vmax[k] = torch.maximum(self.vmax[k], self.v[k])
                denom = self.vmax[k].sqrt() / bc2_sqrt + self.eps
            else:
                denom = self.v[k].sqrt() / bc2_sqrt + self.eps
            params[k] = p - step_size * (self.m[k] / denom)
        return params


def reduce_loss(local_losses):
    """base/base_trainer.py:165-174 as seen on rank 0: sum over ranks / world_size."""
    return sum(local_losses) / len(local_losses)


def ddp_average(per_rank_grads):
    """DDP (train.py:46-52): every gradient is divided by world_size and summed."""
    w = len(per_rank_grads)
    keys = per_rank_grads[0].keys()
    return {k: sum(g[k] / w for g in per_rank_grads) for k in keys}


def train_steps(params, batches, opt, world_size=1):
    """N iterations of trainer/trainer.py:48-58.  `batches` is a list of steps; each step
    is a list (one entry per rank) of (x, target, keep2d, keep1d).  Returns the updated
    params, the per-step rank-0 reduced loss, and the last step's averaged grads."""
    losses, grads = [], None
    for step in batches:
        assert len(step) == world_size
        per_rank, local = [], []
        for (x, t, k2, k1) in step:
            loss, _, g = loss_and_grads(params, x, t, k2, k1)
            per_rank.append(g)
            local.append(loss)
        grads = ddp_average(per_rank)
        losses.append(reduce_loss(local))
        params = opt.step(dict(params), grads)
    return params, losses, grads


def synthetic_batch(n, seed, dtype=torch.float32):
    """x ~ N(0,1) [n,1,28,28], t ~ U{0..9} int64, plus keep
    masks drawn from the same generator."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, 1, 28, 28, generator=g, dtype=torch.float32).to(dtype)
    t = torch.randint(0, 10, (n,), generator=g, dtype=torch.int64)
    k2 = (torch.rand(n, 20, generator=g) < 0.5).to(torch.uint8)
    k1 = (torch.rand(n, 50, generator=g) < 0.5).to(torch.uint8)
    return x, t, k2, k1


def decision_margins(params, x, keep2d=None, keep1d=None):
    """How far every DISCRETE decision of the forward pass is from flipping, per sample (fp64, inputs as given):
    the gap between the two largest values of every 2x2 pooling window (arg-max routing of max_pool2d's
    backward, model/model.py:16-17), the distance from zero of every value entering a relu (:16,17,19) and the
    gaps between the four largest log-probabilities (arg-max / top-k of model/metric.py).  A decision closer to
    flipping than the fp32 rounding noise of the sum behind it (~1e-7) may legitimately come out differently in
    two correct fp32 implementations -- and then one whole gradient path of that sample differs (measured:
    sample 110 of synthetic_batch(128, 1078) has a pooled conv1 value of 8.9e-8; flipping its relu moves
    conv1.weight's gradient by 3e-3 of its maximum).  Parity inputs are therefore drawn so that no decision
    is within `margin` of flipping (robust_batch); that is a statement about the INPUTS, like the tie-free rows
    SURVEY.md A.3 asks for, not a tolerance on the result."""
    p = {k: v.double() for k, v in params.items()}
    x = x.double()
    n = x.shape[0]
    big = torch.full((n,), float("inf"), dtype=torch.float64)

    def pool_margin(c, alive=None):
        nb, ch, hh, ww = c.shape
        win = c.view(nb, ch, hh // 2, 2, ww // 2, 2).permute(0, 1, 2, 4, 3, 5).reshape(nb, ch, hh // 2, ww // 2, 4)
        top = win.topk(2, dim=-1)[0]
        m = torch.minimum(top[..., 0] - top[..., 1], top[..., 0].abs())
        if alive is not None:
            m = torch.where(alive[:, :, None, None], m, torch.full_like(m, float("inf")))
        return m.reshape(nb, -1).min(dim=1)[0]

    c1 = F.conv2d(x, p["conv1.weight"], p["conv1.bias"])
    m = torch.minimum(big, pool_margin(c1))
    h = F.relu(F.max_pool2d(c1, 2))
    c2 = F.conv2d(h, p["conv2.weight"], p["conv2.bias"])
    alive = None
    if keep2d is not None:
        c2 = c2 * (keep2d.double() * 2.0)[:, :, None, None]
        alive = keep2d.bool()
    m = torch.minimum(m, pool_margin(c2, alive))
    h = F.relu(F.max_pool2d(c2, 2)).view(-1, 320)
    z = F.linear(h, p["fc1.weight"], p["fc1.bias"])
    m = torch.minimum(m, z.abs().min(dim=1)[0])
    h = F.relu(z)
    if keep1d is not None:
        h = h * (keep1d.double() * 2.0)
    out = F.log_softmax(F.linear(h, p["fc2.weight"], p["fc2.bias"]), dim=1)
    top = out.topk(4, dim=1)[0]
    m = torch.minimum(m, (top[:, :-1] - top[:, 1:]).min(dim=1)[0])
    return m


def robust_batch(n, seed, params=None, margin=2e-6, dtype=torch.float32, modes=("train",)):
    """synthetic_batch(n, seed) with every sample whose decision margin (decision_margins, at `params`, default
    init_params(0)) is below `margin` replaced by a fresh draw, until none is left.  Deterministic.
    modes: "train" checks the forward with the batch's dropout masks, "eval" the forward without dropout."""
    params = init_params(0) if params is None else params
    x, t, k2, k1 = synthetic_batch(n, seed, dtype)
    for attempt in range(1, 64):
        m = torch.full((x.shape[0],), float("inf"), dtype=torch.float64)
        if "train" in modes:
            m = torch.minimum(m, decision_margins(params, x, k2, k1))
        if "eval" in modes:
            m = torch.minimum(m, decision_margins(params, x))
        bad = (m < margin).nonzero()[:, 0]
        if bad.numel() == 0:
            return x, t, k2, k1
        fx, ft, fk2, fk1 = synthetic_batch(int(bad.numel()), seed * 1000003 + attempt, dtype)
        x, t, k2, k1 = x.clone(), t.clone(), k2.clone(), k1.clone()
        x[bad], t[bad], k2[bad], k1[bad] = fx, ft, fk2, fk1
    raise RuntimeError("robust_batch: could not draw a well-conditioned batch")
