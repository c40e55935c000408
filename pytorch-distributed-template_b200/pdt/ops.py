"""autograd glue between the plug-in surface and the C ABI.

Mirrors reference call sites: model/model.py:15-22 (forward), model/loss.py:4-5,
model/metric.py:4-20, trainer/trainer.py:57 (backward).

Two levels: every stage as its own kernel for any caller, and -- when the Trainer has
announced the criterion's target to the model and the batch is a single wave of CTAs -- the fused kernels
(pdt_conv2_head_train, pdt_conv2_bwd_data_conv1_bwd_weight), with fallbacks back to the general kernels whenever
the announced contract does not hold (another consumer of the output, a second forward before the backward).
"""
import os
import struct
import weakref

import torch

from . import native
from .native import call, ptr, stream, num_tiles

P1F, P2F, HF = 1440, 320, 50
IDX1, IDX2 = 360, 320
STATS_BYTES = 32
# backward: conv1's weight gradient inside conv2's data-gradient kernel (PDT_FUSE_WG1=0: the two separate kernels)
FUSE_DGRAD_WGRAD1 = os.environ.get("PDT_FUSE_WG1", "1") != "0"
# training forward with an announced target: conv2 + head forward + criterion + head backward as ONE kernel
# (pdt_conv2_head_train); PDT_FUSE_TRAIN=0: conv2_fwd, head_fwd_nll and (in backward) head_bwd_nll
FUSE_HEAD_TRAIN = os.environ.get("PDT_FUSE_TRAIN", "1") != "0"
TRAIN_FUSED_MAX_ROWS = 8192
# Both fused kernels trade occupancy (one CTA per SM) for fewer kernel boundaries: that pays while a kernel is
# a single wave of CTAs (4 per tile of 32 samples on 148 SMs), not when the batch is large enough to be
# throughput-bound (measured at 4096: 344 us fused, 330 us with the separate kernels).
# The limit is one wave of the fused kernels' 4-CTA clusters: (SMs // 4) tiles of 32 samples -- 37 * 32 = 1,184 on a
# B200's 148 SMs -- taken from the device at the first call (fused_rows); PDT_FUSE_ROWS, or assigning FUSED_ROWS,
# overrides it.
FUSED_ROWS = int(os.environ["PDT_FUSE_ROWS"]) if "PDT_FUSE_ROWS" in os.environ else None
_FUSED_ROWS_BY_DEVICE = {}


def fused_rows(device):
    """Largest batch for which the fused (one CTA per SM) kernels are used on `device`."""
    if FUSED_ROWS is not None:
        return FUSED_ROWS
    key = device.index if device.index is not None else torch.cuda.current_device()
    rows = _FUSED_ROWS_BY_DEVICE.get(key)
    if rows is None:
        rows = _FUSED_ROWS_BY_DEVICE[key] = torch.cuda.get_device_properties(key).multi_processor_count // 4 * 32
    return rows
# conv2's forward outside the fused training kernel: on the tensor cores (tcgen05.mma, split-TF32: 13.5 us against
# 25.0 us on the CUDA cores at batch 1024, equal within fp32 rounding); PDT_CONV2_TC=0 selects the CUDA-core kernel
CONV2_TC = os.environ.get("PDT_CONV2_TC", "1") != "0"
# the fused training kernel with its conv2 on the tensor cores (pdt_conv2_head_train_tc: 28.2 us against 29.8 us at batch
# 1024, 91.5 against 92.4 us per step); PDT_TRAIN_TC=0: the CUDA-core version
TRAIN_TC = os.environ.get("PDT_TRAIN_TC", "1") != "0"


def _carve(device, sizes):
    """One allocation, 256-byte aligned slices (sizes in bytes)."""
    offs, total = [], 0
    for s in sizes:
        offs.append(total)
        total += (s + 255) // 256 * 256
    buf = torch.empty(total, dtype=torch.uint8, device=device)
    return buf, [buf[o:o + s] for o, s in zip(offs, sizes)]


class _DeviceState:
    """Per-device persistent scratch: backward temporaries, nll partials, RNG state."""

    _by_device = {}

    def __init__(self, device):
        self.device = device
        self.bwd = {}
        self.bwd_owner = {}         # n -> token of the forward whose head backward currently sits in the workspace
        self.nll = {}
        self.rng = None
        self.side = None
        self.events = None

    @classmethod
    def get(cls, device):
        key = (device.type, device.index if device.index is not None else torch.cuda.current_device())
        st = cls._by_device.get(key)
        if st is None:
            st = cls._by_device[key] = cls(device)
        return st

    def bwd_workspace(self, n):
        ws = self.bwd.get(n)
        if ws is None:
            T = num_tiles(n)
            scratch_bytes = native.load().pdt_mnist_scratch_bytes(n)
            _, (dq2, dq1, scratch) = _carve(self.device, [T * P2F * 32 * 4, T * P1F * 32 * 4, scratch_bytes])
            ws = self.bwd[n] = (dq2.view(torch.float32), dq1.view(torch.float32), scratch)
        return ws

    def side_stream(self, which=0):
        """A second stream + (fork, join) event pairs for running independent kernels of one step side by
        side (pair 0: backward's conv2 weight gradient; pair 1: forward's dropout masks)."""
        if self.side is None:
            self.side = torch.cuda.Stream(self.device)
            self.events = [(torch.cuda.Event(), torch.cuda.Event()) for _ in range(2)]
        return self.side, self.events[which]

    def nll_scratch(self, n):
        need = native.load().pdt_nll_scratch_bytes(n)
        buf = self.nll.get(need)
        if buf is None:
            buf = self.nll[need] = torch.zeros(need, dtype=torch.uint8, device=self.device)
        return buf

    def head_nll_scratch(self, n):
        need = native.load().pdt_head_nll_scratch_bytes(n)
        buf = self.nll.get(("head", need))
        if buf is None:
            buf = self.nll[("head", need)] = torch.zeros(need, dtype=torch.uint8, device=self.device)
        return buf

    def rng_state(self, seed=None):
        if self.rng is None or seed is not None:
            if seed is None:
                seed = torch.initial_seed()
            host = torch.tensor([seed & 0x7FFFFFFFFFFFFFFF, 0, 0], dtype=torch.int64)
            self.rng = host.to(self.device)
        return self.rng


def seed_dropout(seed, device=None):
    """Seed the dropout mask generator of `device` (default: current CUDA device)."""
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    _DeviceState.get(device).rng_state(seed)


# Normalize((0.1307,), (0.3081,)) of the reference's MnistDataLoader (data_loader/data_loaders.py:13-16), applied
# inside conv1's load when the batch arrives as raw uint8 pixels
INPUT_NORM = (0.1307, 0.3081)


def mnist_forward_raw(x, flat, keep2d, keep1d, save, mask_rng=None, loss=None, norm=INPUT_NORM):
    """Run the three forward kernels.  flat: fp32 vector of the 21,840 parameters.
    x: [n,1,28,28] fp32, or uint8 raw pixels normalised on load with norm = (mean, std).
    Returns (logp, saved) where saved = (p1, idx1, p2, idx2, hd) when `save`.
    loss = (target int64 [n], stats int32[8], topk, ignore_index): the head kernel also evaluates the
    criterion (pdt_head_fwd_nll) and fills `stats`."""
    n = x.shape[0]
    T = num_tiles(n)
    dev = x.device
    _, (p1, p2, hd, idx1, idx2) = _carve(dev, [T * P1F * 128, T * P2F * 128, T * HF * 128, T * IDX1 * 32, T * IDX2 * 32])
    p1, p2, hd = p1.view(torch.float32), p2.view(torch.float32), hd.view(torch.float32)
    logp = torch.empty((n, 10), dtype=torch.float32, device=dev)
    base = flat.data_ptr()
    s = stream()
    u8 = x.dtype == torch.uint8
    if mask_rng is not None:     # training with the native generator: masks are drawn inside conv1's launch
        if u8:
            call("pdt_conv1_fwd_masks_u8", ptr(x), norm[0], norm[1], base, base + 4 * 250, ptr(p1), ptr(idx1), ptr(keep2d),
                 ptr(keep1d), 0.5, ptr(mask_rng), n, s)
        else:
            call("pdt_conv1_fwd_masks", ptr(x), base, base + 4 * 250, ptr(p1), ptr(idx1), ptr(keep2d), ptr(keep1d), 0.5,
                 ptr(mask_rng), n, s)
    elif u8:
        call("pdt_conv1_pool_relu_fwd_u8", ptr(x), norm[0], norm[1], base, base + 4 * 250, ptr(p1), ptr(idx1), n, s)
    else:
        call("pdt_conv1_pool_relu_fwd", ptr(x), base, base + 4 * 250, ptr(p1), ptr(idx1), n, s)
    if loss is not None and save and FUSE_HEAD_TRAIN and n <= min(fused_rows(dev), TRAIN_FUSED_MAX_ROWS):
        target, stats, topk, ignore_index = loss
        sink = _LOSS_SINK if (_LOSS_SINK is not None and _LOSS_SINK[0].device == dev) else (None, None, 0)
        st = _DeviceState.get(dev)
        dq2, _, scratch = st.bwd_workspace(n)
        sc = 2.0
        call("pdt_conv2_head_train_tc" if TRAIN_TC else "pdt_conv2_head_train", ptr(p1), base + 4 * 260, base + 4 * 5260, ptr(keep2d), sc, ptr(p2), ptr(idx2),
             base + 4 * 5280, base + 4 * 21280, base + 4 * 21330, base + 4 * 21830, ptr(keep1d), sc, ptr(hd), ptr(logp),
             ptr(target), int(topk), int(ignore_index), ptr(stats), ptr(st.head_nll_scratch(n)), ptr(sink[0]),
             ptr(sink[1]), sink[2], ptr(dq2), ptr(scratch), n, s)
        token = st.bwd_owner[n] = object()
        return logp, (p1, idx1, p2, idx2, hd, token)
    call("pdt_conv2_drop_pool_relu_fwd_tc" if CONV2_TC else "pdt_conv2_drop_pool_relu_fwd", ptr(p1), base + 4 * 260,
         base + 4 * 5260, ptr(keep2d), 2.0, ptr(p2), ptr(idx2), n, s)
    if loss is not None:
        target, stats, topk, ignore_index = loss
        sink = _LOSS_SINK if (_LOSS_SINK is not None and _LOSS_SINK[0].device == dev) else (None, None, 0)
        call("pdt_head_fwd_nll", ptr(p2), base + 4 * 5280, base + 4 * 21280, base + 4 * 21330, base + 4 * 21830,
             ptr(keep1d), 2.0, ptr(hd), ptr(logp), n, ptr(target), int(topk), int(ignore_index), ptr(stats),
             ptr(_DeviceState.get(dev).head_nll_scratch(n)), ptr(sink[0]), ptr(sink[1]), sink[2], s)
    else:
        call("pdt_head_fwd", ptr(p2), base + 4 * 5280, base + 4 * 21280, base + 4 * 21330, base + 4 * 21830,
             ptr(keep1d), 2.0, ptr(hd), ptr(logp), n, s)
    return logp, ((p1, idx1, p2, idx2, hd, None) if save else None)


def head_backward_done(x, saved):
    """True when the forward ran the training kernel AND its head backward still sits in the workspace (no
    other forward of the same batch size has overwritten it since)."""
    token = saved[5]
    return token is not None and _DeviceState.get(x.device).bwd_owner.get(x.shape[0]) is token


def mnist_backward_raw(x, flat, saved, logp, dlogp, training, flat_grad, beta, scale, stats, reduce=True,
                       nll=None, head_done=False, norm=INPUT_NORM):
    """Run the backward kernels and reduce the partials into flat_grad (+ tail if stats).
    nll = (target, stats, grad_loss or None, ignore_index) instead of dlogp: the loss gradient is formed
    inside the head kernel (pdt_head_bwd_nll)."""
    n = x.shape[0]
    p1, idx1, p2, idx2, hd = saved[:5]
    st = _DeviceState.get(x.device)
    dq2, dq1, scratch = st.bwd_workspace(n)
    base = flat.data_ptr()
    s = stream()
    sc = 2.0 if training else 1.0
    if head_done:
        pass        # dq2 and the fc1/fc2 partials were written by pdt_conv2_head_train in the forward
    else:
        st.bwd_owner[n] = None      # this backward overwrites the workspace: no forward's head backward is in it now
    if head_done:
        pass
    elif nll is not None:
        target, nll_stats, grad_loss, ignore_index = nll
        call("pdt_head_bwd_nll", ptr(target), ptr(nll_stats), ptr(grad_loss), int(ignore_index), ptr(logp), ptr(hd),
             ptr(p2), base + 4 * 5280, base + 4 * 21330, sc, sc, ptr(dq2), ptr(scratch), n, s)
    else:
        call("pdt_head_bwd", ptr(dlogp), ptr(logp), ptr(hd), ptr(p2), base + 4 * 5280, base + 4 * 21330,
             sc, sc, ptr(dq2), ptr(scratch), n, s)
    # conv2's weight gradient does not depend on its data gradient: fork it onto a side stream so the
    # two (and conv1's weight gradient after the data gradient) share the SMs instead of queueing.
    # Inside a CUDA-graph capture the fork/join becomes two parallel branches of the graph.
    main = torch.cuda.current_stream()
    side, (forked, joined) = st.side_stream()
    forked.record(main)
    # the data-gradient kernel goes first (128 CTAs, one per SM); the weight-gradient kernel then
    # co-resides with it on the same SMs instead of taking every slot for itself
    fuse_wg1 = FUSE_DGRAD_WGRAD1 and n <= fused_rows(x.device)
    u8 = x.dtype == torch.uint8
    if fuse_wg1 and u8:
        call("pdt_conv2_bwd_data_conv1_bwd_weight_u8", ptr(dq2), ptr(idx2), base + 4 * 260, ptr(p1), ptr(x), norm[0], norm[1],
             ptr(idx1), ptr(scratch), n, s)
    elif fuse_wg1:
        # conv1's weight gradient is taken inside the data-gradient kernel (dq1 never reaches memory)
        call("pdt_conv2_bwd_data_conv1_bwd_weight", ptr(dq2), ptr(idx2), base + 4 * 260, ptr(p1), ptr(x), ptr(idx1),
             ptr(scratch), n, s)
    else:
        call("pdt_conv2_bwd_data", ptr(dq2), ptr(idx2), base + 4 * 260, ptr(p1), ptr(dq1), n, s)
    side.wait_event(forked)
    call("pdt_conv2_bwd_weight", ptr(dq2), ptr(idx2), ptr(p1), ptr(scratch), n, side.cuda_stream)
    joined.record(side)
    if not fuse_wg1 and u8:
        call("pdt_conv1_bwd_weight_u8", ptr(dq1), ptr(idx1), ptr(x), norm[0], norm[1], ptr(scratch), n, s)
    elif not fuse_wg1:
        call("pdt_conv1_bwd_weight", ptr(dq1), ptr(idx1), ptr(x), ptr(scratch), n, s)
    main.wait_event(joined)
    if not reduce:
        return scratch
    call("pdt_grad_reduce", ptr(scratch), n, float(beta), float(scale), ptr(stats), ptr(flat_grad),
         ptr(nll[2]) if head_done else None, s)
    return None


class MnistFunction(torch.autograd.Function):
    """forward = model/model.py:15-22; backward writes straight into the flat gradient buffer
    (and launches the data-parallel all-reduce when the owner is wrapped), returning no
    per-parameter gradients to autograd."""

    @staticmethod
    def forward(ctx, x, owner_ref, keep2d, keep1d, *params):
        owner = owner_ref()
        x = x.contiguous()
        native.require_cuda(x)
        need_grad = any(ctx.needs_input_grad[4:])
        rng, owner._mask_rng = owner._mask_rng, None
        loss, owner._fused_fwd = owner._fused_fwd, None
        logp, saved = mnist_forward_raw(x, owner.flat.data, keep2d, keep1d, need_grad, rng, loss, owner.input_norm)
        if need_grad:
            ctx.owner_ref = owner_ref
            ctx.x = x
            ctx.saved = saved
            ctx.logp = logp.detach()  # not the output object itself: that would be a reference cycle
            ctx.training = keep2d is not None
            ctx.fused_key = None if loss is None else loss[1].data_ptr()
        ctx.set_materialize_grads(False)
        return logp

    @staticmethod
    def backward(ctx, dlogp):
        n_in = 4 + 8
        owner = ctx.owner_ref()
        # left by FusedNllFunction.backward for THIS forward (keyed by its stats record)
        nll = owner._fused_bwd.pop(ctx.fused_key, None) if ctx.fused_key is not None else None
        if nll is not None and dlogp is not None:
            # the output had another consumer besides the fused criterion: form the criterion's share explicitly
            target, stats, grad_loss, ignore_index = nll
            own = torch.empty_like(dlogp)
            call("pdt_nll_bwd", ptr(target), ptr(stats), ptr(grad_loss), ptr(own), own.shape[0], own.shape[1],
                 int(ignore_index), stream())
            dlogp, nll = dlogp + own, None
        if dlogp is None and nll is None:
            return (None,) * n_in
        head_done = nll is not None and head_backward_done(ctx.x, ctx.saved)
        owner._native_backward(ctx.x, ctx.saved, ctx.logp, None if dlogp is None else dlogp.contiguous(),
                               ctx.training, nll, head_done)
        return (None,) * n_in


# (ring tensor, counter tensor, capacity) while a trainer wants every criterion call logged on the device
_LOSS_SINK = None


def set_loss_sink(ring=None, counter=None, capacity=0):
    """Trainer hook: while set, nll_loss appends its value to `ring` inside the loss kernel itself."""
    global _LOSS_SINK
    _LOSS_SINK = None if ring is None else (ring, counter, int(capacity))


class NllFunction(torch.autograd.Function):
    """model/loss.py:4-5 with the metric counts of model/metric.py as a by-product."""

    @staticmethod
    def forward(ctx, output, target, stats, topk, ignore_index):
        native.require_cuda(output, target)
        output = output.contiguous()
        target = target.contiguous()
        if target.dtype != torch.int64:
            raise RuntimeError("nll_loss expects int64 targets (as F.nll_loss does)")
        n, classes = output.shape
        st = _DeviceState.get(output.device)
        sink = _LOSS_SINK
        if sink is not None and sink[0].device == output.device:
            call("pdt_nll_metrics_fwd_log", ptr(output), ptr(target), n, classes, topk, ignore_index, ptr(stats),
                 ptr(st.nll_scratch(n)), ptr(sink[0]), ptr(sink[1]), sink[2], stream())
        else:
            call("pdt_nll_metrics_fwd", ptr(output), ptr(target), n, classes, topk, ignore_index, ptr(stats),
                 ptr(st.nll_scratch(n)), stream())
        ctx.save_for_backward(target)
        ctx.stats = stats
        ctx.shape = (n, classes)
        ctx.ignore_index = ignore_index
        return stats[:1].view(torch.float32).reshape(())

    @staticmethod
    def backward(ctx, grad_loss):
        (target,) = ctx.saved_tensors
        n, classes = ctx.shape
        dlogp = torch.empty((n, classes), dtype=torch.float32, device=target.device)
        g = None if grad_loss is None else grad_loss.contiguous()
        call("pdt_nll_bwd", ptr(target), ptr(ctx.stats), ptr(g), ptr(dlogp), n, classes, ctx.ignore_index, stream())
        return dlogp, None, None, None, None


class FusedNllFunction(torch.autograd.Function):
    """The criterion when the model's head kernel has already evaluated it (MnistModel.expect_target):
    forward launches nothing; backward hands (target, stats, grad_loss) to the model's backward, which forms
    the log-probability gradient inside its first kernel, and returns no tensor of its own."""

    @staticmethod
    def forward(ctx, output, target, stats, owner_ref, ignore_index):
        ctx.owner_ref = owner_ref
        ctx.target = target
        ctx.stats = stats
        ctx.ignore_index = ignore_index
        ctx.set_materialize_grads(False)
        return stats[:1].view(torch.float32).reshape(())

    @staticmethod
    def backward(ctx, grad_loss):
        owner = ctx.owner_ref()
        if grad_loss is not None and owner is not None:
            owner._fused_bwd[ctx.stats.data_ptr()] = (ctx.target, ctx.stats, grad_loss.contiguous(), ctx.ignore_index)
        return None, None, None, None, None


def nll_loss_with_stats(output, target, topk=3, ignore_index=-100):
    """Returns (loss 0-dim tensor, stats int32[8] tensor laid out as pdt_stats)."""
    owner_ref = getattr(output, "_pdt_owner", None)
    owner = owner_ref() if owner_ref is not None else None
    fused = getattr(output, "_pdt_fused", None)
    if (fused is not None and owner is not None and fused[0] is target and fused[2] == topk
            and fused[3] == ignore_index):
        stats = fused[1]
        loss = FusedNllFunction.apply(output, target, stats, owner_ref, ignore_index)
    else:
        stats = torch.empty(8, dtype=torch.int32, device=output.device)
        loss = NllFunction.apply(output, target, stats, topk, ignore_index)
    if owner is not None:
        owner._last_stats = stats
    loss._pdt_stats = stats
    return loss, stats


def metric_counts(output, target, k):
    """(top1, topk, n_rows) as Python ints; one kernel + one 32-byte D2H read."""
    native.require_cuda(output, target)
    with torch.no_grad():
        output = output.detach().contiguous().float()
        target = target.contiguous().to(torch.int64)
        n, classes = output.shape
        st = _DeviceState.get(output.device)
        stats = torch.empty(8, dtype=torch.int32, device=output.device)
        call("pdt_nll_metrics_fwd", ptr(output), ptr(target), n, classes, int(k), -100, ptr(stats),
             ptr(st.nll_scratch(n)), stream())
        host = stats.cpu()
    return int(host[2]), int(host[3]), n


def stats_async(output, target, k=3, ignore_index=-100):
    """The loss/metric record of one batch as a device int32[8] tensor (pdt_stats layout); no host sync."""
    native.require_cuda(output, target)
    with torch.no_grad():
        output = output.detach().contiguous().float()
        target = target.contiguous().to(torch.int64)
        n, classes = output.shape
        st = _DeviceState.get(output.device)
        stats = torch.empty(8, dtype=torch.int32, device=output.device)
        call("pdt_nll_metrics_fwd", ptr(output), ptr(target), n, classes, int(k), ignore_index, ptr(stats),
             ptr(st.nll_scratch(n)), stream())
    return stats


def unpack_stats(stats):
    raw = stats.cpu().numpy().tobytes()
    loss_mean, loss_sum, top1, topk, n_valid, n_rows = struct.unpack("ffiiii", raw[:24])
    return dict(loss_mean=loss_mean, loss_sum=loss_sum, top1=top1, topk=topk, n_valid=n_valid, n_rows=n_rows)
