"""Fused flat-buffer optimizers with torch.optim's constructor surface.

The reference builds its optimizer by name from torch.optim (train.py:42,
config/config.json:38-45: Adam, lr 1e-3, weight_decay 0, amsgrad true) and steps it at
trainer/trainer.py:58.  These classes subclass torch.optim.Optimizer (so lr schedulers and
state_dict work, train.py:43, base/base_trainer.py:122,161) but update every parameter of a
group in ONE kernel launch over the flat buffer (pdt_sgd_step_flat / pdt_adam_step_flat).
Hyper-parameters are mirrored into a small device struct that the kernels read, so an lr
scheduler's change reaches a captured CUDA graph without re-capture: step() uploads the struct when
it differs from the last upload, and a caller that REPLAYS a captured step (which does not pass
through step()) calls sync_hyper() before the replay, as the Trainer does.
"""
import struct

import torch

from . import flat as _flat
from . import native
from .native import call, ptr, stream


# torch.optim keyword arguments that only select among torch's own implementations of the same update: a config
# written for the reference may carry them (config.json "optimizer.args"); they are accepted and have no effect
# here (there is one implementation: a single fused launch).  Anything that changes the arithmetic is refused.
_IMPLEMENTATION_ONLY = {"foreach": (None, True, False), "fused": (None, True, False), "capturable": (True, False),
                        "differentiable": (False,), "maximize": (False,)}


def _check_torch_kwargs(name, kwargs):
    for key, value in kwargs.items():
        allowed = _IMPLEMENTATION_ONLY.get(key)
        if allowed is None:
            raise TypeError("%s.__init__() got an unexpected keyword argument %r" % (name, key))
        if value not in allowed:
            raise ValueError("pdt.optim.%s does not support %s=%r" % (name, key, value))


class _FlatOptimizer(torch.optim.Optimizer):
    def _setup(self):
        self._groups = []
        for group in self.param_groups:
            params = [p for p in group["params"]]
            fb, start, count = _flat.ensure(params)
            dev = fb.data.device
            hyper = torch.zeros(64, dtype=torch.uint8, device=dev)
            self._groups.append(dict(fb=fb, start=start, count=count, hyper=hyper, cached=None, dev=dev))
        self.grad_scale = 1.0

    def zero_grad(self, set_to_none=False):
        """Keeps .grad views alive (the reference's default set_to_none=True would orphan the flat
        buffer).  Native models overwrite on the next backward, so nothing is
        written here; for any other module the buffer is cleared with one memset."""
        for g in self._groups:
            fb = g["fb"]
            if not fb.grads_attached():
                fb.attach()
            if getattr(fb, "native_backward", False):
                fb.zero_grad()
            else:
                fb.zero_grad_memset()

    def _upload(self, g, packed):
        if g["cached"] != packed:
            host = torch.frombuffer(bytearray(packed), dtype=torch.uint8)
            g["hyper"][:len(packed)].copy_(host, non_blocking=False)
            g["cached"] = packed

    def sync_hyper(self):
        """Upload the hyper-parameters of `param_groups` to the device structs the kernels read, if they changed
        since the last upload (a bytes compare; a change costs one 44-byte copy on the current stream).  step()
        does this itself; a step replayed from a CUDA graph does NOT pass through step(), so whoever replays
        (Trainer._run_staged) calls this first -- that is how an lr scheduler's or a user's edit of
        param_groups (reference trainer.py:90-91) reaches a captured step.  Must not be called during capture."""
        for g, group in zip(self._groups, self.param_groups):
            self._upload(g, self._pack(group))

    def device_step(self, gi=0):
        """Number of updates applied so far (reads the device counter; synchronises)."""
        raw = self._groups[gi]["hyper"].cpu().numpy().tobytes()
        return struct.unpack_from("i", raw, self._STEP_OFF)[0]

    def _set_device_step(self, gi, step):
        g = self._groups[gi]
        host = torch.frombuffer(bytearray(struct.pack("iI", int(step), 0)), dtype=torch.uint8)
        g["hyper"][self._STEP_OFF:self._STEP_OFF + 8].copy_(host)


class SGD(_FlatOptimizer):
    """torch.optim.SGD semantics (torch/optim/sgd.py): weight_decay, momentum, dampening, nesterov."""
    _STEP_OFF = 36

    def __init__(self, params, lr=1e-3, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False, **torch_kwargs):
        _check_torch_kwargs("SGD", torch_kwargs)
        if nesterov and (momentum <= 0 or dampening != 0):
            raise ValueError("Nesterov momentum requires a momentum and zero dampening")
        defaults = dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay, nesterov=nesterov)
        super().__init__(params, defaults)
        self._setup()
        for g, group in zip(self._groups, self.param_groups):
            g["buf"] = None

    @staticmethod
    def _pack(group):
        return struct.pack("ddddi", group["lr"], group["momentum"], group["dampening"], group["weight_decay"],
                           int(group["nesterov"]))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for g, group in zip(self._groups, self.param_groups):
            fb, start, count = g["fb"], g["start"], g["count"]
            fb.check_alive()
            self._upload(g, self._pack(group))
            fb.materialize()
            if group["momentum"] != 0 and g["buf"] is None:
                g["buf"] = torch.zeros(count, dtype=torch.float32, device=g["dev"])
                self._link_state(g, group)
            call("pdt_sgd_step_flat", fb.data.data_ptr() + 4 * start, fb.grad.data_ptr() + 4 * start,
                 ptr(g["buf"]), count, float(self.grad_scale), ptr(g["hyper"]), stream())
        return loss

    def _link_state(self, g, group):
        off = 0
        for p in group["params"]:
            self.state[p]["momentum_buffer"] = g["buf"][off:off + p.numel()].view(p.shape)
            off += p.numel()

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        for gi, (g, group) in enumerate(zip(self._groups, self.param_groups)):
            bufs = [self.state[p].get("momentum_buffer") for p in group["params"]]
            if all(b is not None for b in bufs) and bufs:
                g["buf"] = torch.cat([b.reshape(-1).to(g["dev"], torch.float32) for b in bufs])
                self._link_state(g, group)
                self._set_device_step(gi, 1)
            g["cached"] = None


class Adam(_FlatOptimizer):
    """torch.optim.Adam semantics incl. amsgrad (torch/optim/adam.py, _single_tensor_adam)."""
    _STEP_OFF = 44

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, amsgrad=False, **torch_kwargs):
        _check_torch_kwargs("Adam", torch_kwargs)
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=amsgrad)
        super().__init__(params, defaults)
        self._setup()
        for g, group in zip(self._groups, self.param_groups):
            n = g["count"]
            g["m"] = torch.zeros(n, dtype=torch.float32, device=g["dev"])
            g["v"] = torch.zeros(n, dtype=torch.float32, device=g["dev"])
            g["vmax"] = torch.zeros(n, dtype=torch.float32, device=g["dev"]) if group["amsgrad"] else None
            self._link_state(g, group)
            if getattr(g["fb"], "native_backward", False) and g["start"] == 0 and g["count"] == g["fb"].numel:
                g["fb"].fused_step_ready = True

    @staticmethod
    def _pack(group):
        b1, b2 = group["betas"]
        return struct.pack("dddddi", group["lr"], b1, b2, group["eps"], group["weight_decay"], int(group["amsgrad"]))

    def _link_state(self, g, group):
        off = 0
        for p in group["params"]:
            st = self.state[p]
            k = p.numel()
            st["exp_avg"] = g["m"][off:off + k].view(p.shape)
            st["exp_avg_sq"] = g["v"][off:off + k].view(p.shape)
            if g["vmax"] is not None:
                st["max_exp_avg_sq"] = g["vmax"][off:off + k].view(p.shape)
            off += k

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for g, group in zip(self._groups, self.param_groups):
            fb, start, count = g["fb"], g["start"], g["count"]
            fb.check_alive()
            self._upload(g, self._pack(group))
            if fb.pending is not None and start == 0 and count == fb.numel and self.grad_scale == 1.0:
                # backward left the per-tile partials: sum them and update in ONE launch
                dp, fb.pending_dp = fb.pending_dp, None
                if dp is not None:      # data parallel over peer memory: reduce + all-reduce + update in ONE launch
                    pending, fb.pending = fb.pending, None
                    dp.p2p.reduce_adam_(pending, fb, g)
                    continue
                scratch, n, beta, scale, stats, gscale = fb.pending
                fb.pending = None
                call("pdt_adam_step_reduce", scratch.data_ptr(), n, float(beta), float(scale), ptr(stats),
                     fb.grad.data_ptr(), fb.data.data_ptr(), ptr(g["m"]), ptr(g["v"]), ptr(g["vmax"]),
                     ptr(g["hyper"]), ptr(gscale), stream())
                continue
            fb.materialize()
            call("pdt_adam_step_flat", fb.data.data_ptr() + 4 * start, fb.grad.data_ptr() + 4 * start,
                 ptr(g["m"]), ptr(g["v"]), ptr(g["vmax"]), count, float(self.grad_scale), ptr(g["hyper"]), stream())
        return loss

    def state_dict(self):
        """torch.optim.Adam-compatible: every parameter's state carries a 'step' tensor."""
        for gi, group in enumerate(self.param_groups):
            step = float(self.device_step(gi))
            for p in group["params"]:
                self.state[p]["step"] = torch.tensor(step)
        return super().state_dict()

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        for gi, (g, group) in enumerate(zip(self._groups, self.param_groups)):
            sts = [self.state[p] for p in group["params"]]
            if sts and all("exp_avg" in s for s in sts):
                cat = lambda key: torch.cat([s[key].reshape(-1).to(g["dev"], torch.float32) for s in sts])
                g["m"], g["v"] = cat("exp_avg"), cat("exp_avg_sq")
                if group["amsgrad"]:
                    g["vmax"] = cat("max_exp_avg_sq") if all("max_exp_avg_sq" in s for s in sts) \
                        else torch.zeros_like(g["m"])
                step = sts[0].get("step", 0)
                self._set_device_step(gi, int(float(step)))
                self._link_state(g, group)
            g["cached"] = None
