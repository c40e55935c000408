"""Flat parameter / gradient buffers.

The reference keeps one tensor per parameter and lets DDP copy gradients into 25 MB buckets and
back (train.py:46-52, gradient_as_bucket_view=False).  Here every parameter is a view into ONE
fp32 buffer and every .grad a view into a second one, followed by a 4-float tail
[loss, top1, topk, n_valid]; the optimizer and the all-reduce then run over the whole buffer in
one launch each and no copies are needed.
"""
import torch

from . import native

TAIL = 4
_REGISTRY = []  # live FlatBuffers, looked up by storage pointer


class FlatBuffer:
    def __init__(self, params, pad_to=4):
        params = [p for p in params]
        assert params, "no parameters"
        dev, dt = params[0].device, params[0].dtype
        assert dt == torch.float32, "flat buffers are fp32 (master weights)"
        self.params = params
        self.offsets = []
        off = 0
        for p in params:
            assert p.device == dev and p.dtype == dt
            self.offsets.append(off)
            off += p.numel()
        self.numel = off
        total = off + TAIL
        total = (total + pad_to - 1) // pad_to * pad_to
        self.data = torch.zeros(total, device=dev, dtype=dt)
        self.grad = torch.zeros(total, device=dev, dtype=dt)
        with torch.no_grad():
            for p, o in zip(params, self.offsets):
                self.data[o:o + p.numel()].copy_(p.detach().reshape(-1))
        self.attach()
        self.zeroed = True  # gradients hold nothing yet: the next backward may overwrite
        # Deferred reduction (native model, single GPU): backward may leave the per-tile partial gradients
        # in scratch and let the optimizer's fused kernel sum them (pdt_adam_step_reduce).  `pending` holds
        # what that kernel needs; materialize() runs the plain reduction for anyone who needs .grad earlier.
        self.defer_reduce = False
        self.fused_step_ready = False
        self.pending = None
        self.pending_dp = None      # the data-parallel wrapper whose all-reduce is deferred along with the reduction
        self.dead = None            # set (to the reason) when the owner replaced this buffer: see check_alive()
        _REGISTRY.append(self)

    def check_alive(self):
        if self.dead:
            raise native.PdtError("this flat parameter buffer is no longer the model's: %s -- rebuild it after "
                                  "model.to(...)" % self.dead)

    def release(self):
        if self in _REGISTRY:
            _REGISTRY.remove(self)

    def attach(self):
        """(Re)point every parameter's .data and .grad at its slice of the flat buffers."""
        for p, o in zip(self.params, self.offsets):
            p.data = self.data[o:o + p.numel()].view(p.shape)
            p.grad = self.grad[o:o + p.numel()].view(p.shape)

    def params_attached(self):
        """True while every parameter's storage is still its slice of this buffer (same device and dtype)."""
        base = self.data.data_ptr()
        for p, o in zip(self.params, self.offsets):
            if p.device != self.data.device or p.dtype != self.data.dtype or p.data_ptr() != base + 4 * o:
                return False
        return True

    def grads_attached(self):
        for p, o in zip(self.params, self.offsets):
            g = p.grad
            if g is None or g.data_ptr() != self.grad.data_ptr() + 4 * o:
                return False
        return True

    def materialize(self):
        """Make .grad valid now if a deferred reduction is pending."""
        self.check_alive()
        if self.pending is not None:
            scratch, n, beta, scale, stats, gscale = self.pending
            self.pending = None
            native.call("pdt_grad_reduce", scratch.data_ptr(), n, float(beta), float(scale), native.ptr(stats),
                        self.grad.data_ptr(), native.ptr(gscale), native.stream())
            dp, self.pending_dp = self.pending_dp, None
            if dp is not None:
                dp.allreduce_native()

    @property
    def tail(self):
        return self.grad[self.numel:self.numel + TAIL]

    def zero_grad(self):
        """Nothing is written: the native backward overwrites (beta = 0) when this flag is set;
        generic autograd accumulation needs real zeros, see FlatDataParallel / optim."""
        self.zeroed = True

    def zero_grad_memset(self):
        self.grad.zero_()
        self.zeroed = True


def lookup(params):
    """Return (FlatBuffer, start, count) if `params` are a contiguous run of one live buffer."""
    params = list(params)
    for fb in _REGISTRY:
        base = fb.data.data_ptr()
        end = base + 4 * fb.numel
        if not all(base <= p.data_ptr() < end and p.device == fb.data.device for p in params):
            continue
        start = (params[0].data_ptr() - base) // 4
        pos = start
        ok = True
        for p in params:
            if (p.data_ptr() - base) // 4 != pos:
                ok = False
                break
            pos += p.numel()
        if ok:
            return fb, start, pos - start
    return None


def ensure(params):
    """Flatten `params` unless they already are a contiguous run of a live FlatBuffer."""
    params = list(params)
    native.require_cuda(*params)
    hit = lookup(params)
    if hit is not None:
        return hit
    fb = FlatBuffer(params)
    return fb, 0, fb.numel
