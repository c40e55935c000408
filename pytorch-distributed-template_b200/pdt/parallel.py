"""Data parallelism over one flat gradient buffer and ONE NCCL all-reduce per step.

Stands where the reference puts torch.nn.parallel.DistributedDataParallel (train.py:46-52):
same wrapping contract (state_dict keys gain the `module.` prefix, base/base_trainer.py:121,
151-154; parameters are broadcast from rank 0 at construction), but instead of 25 MB buckets
copied in and out of .grad, gradients already live in a flat buffer whose 4-float tail carries
the mean loss and the metric counts, so BaseTrainer.reduce_loss (base/base_trainer.py:165-174)
needs no collective of its own.  The all-reduce is ncclAllReduce called directly through the
C ABI on the current stream (capturable in a CUDA graph).
"""
import ctypes
import weakref

import torch
import torch.distributed as dist
import torch.nn as nn

from . import flat as _flat
from . import native
from .native import call, stream


class NcclComm:
    """ncclComm_t created from an id shipped through the existing torch process group."""

    def __init__(self, device):
        self.world_size = dist.get_world_size()
        self.rank = dist.get_rank()
        ident = torch.zeros(128, dtype=torch.uint8)
        if self.rank == 0:
            buf = ctypes.create_string_buffer(128)
            call("pdt_comm_unique_id", ctypes.cast(buf, ctypes.c_void_p))
            ident = torch.frombuffer(bytearray(buf.raw), dtype=torch.uint8).clone()
        box = [ident.tolist()]
        dist.broadcast_object_list(box, src=0)
        raw = bytes(box[0])
        self._id = ctypes.create_string_buffer(raw, 128)
        handle = ctypes.c_void_p()
        with torch.cuda.device(device):
            call("pdt_comm_init", ctypes.cast(self._id, ctypes.c_void_p), self.rank, self.world_size,
                 ctypes.byref(handle))
        self.handle = handle

    def allreduce_(self, tensor, count=None):
        count = tensor.numel() if count is None else count
        assert tensor.dtype == torch.float32 and tensor.is_contiguous()
        call("pdt_allreduce_flat", self.handle, tensor.data_ptr(), count, stream())

    def destroy(self):
        if self.handle:
            call("pdt_comm_destroy", self.handle)
            self.handle = None


class FlatDataParallel(nn.Module):
    def __init__(self, module, device_ids=None, output_device=None, broadcast_buffers=True):
        super().__init__()
        if not (dist.is_available() and dist.is_initialized()):
            raise RuntimeError("FlatDataParallel needs an initialised torch.distributed process group")
        self.module = module
        self.world_size = dist.get_world_size()
        self.broadcast_buffers = broadcast_buffers
        params = [p for p in module.parameters() if p.requires_grad]
        native.require_cuda(*params)
        self.native = bool(getattr(getattr(module, "flat", None), "native_backward", False))
        self.flat, self._start, self._count = _flat.ensure(params)
        self.device = self.flat.data.device
        self.comm = NcclComm(self.device)
        # rank 0's weights everywhere (what DDP's constructor does)
        dist.broadcast(self.flat.data, src=0)
        for b in module.buffers():
            dist.broadcast(b, src=0)
        if self.native:
            module._pdt_dp = weakref.ref(self)
        else:
            self._pending = 0
            self._n_params = len(params)
            for p in params:
                p.register_post_accumulate_grad_hook(self._on_grad)

    def forward(self, *args, **kwargs):
        if self.broadcast_buffers and self.training:
            for b in self.module.buffers():
                dist.broadcast(b, src=0)
        return self.module(*args, **kwargs)

    # native model: called from the model's backward once every gradient is in the flat buffer
    def allreduce_native(self):
        self.comm.allreduce_(self.flat.grad, self.flat.numel + _flat.TAIL)

    # any other module: count accumulated gradients, then scale by 1/W and all-reduce once
    def _on_grad(self, param):
        self._pending += 1
        if self._pending == self._n_params:
            self._pending = 0
            g = self.flat.grad
            call("pdt_flat_scale_cast", g.data_ptr(), g.data_ptr(), self.flat.numel, 1.0 / self.world_size, 0, stream())
            self.comm.allreduce_(g, self.flat.numel + _flat.TAIL)

    def reduced_tail(self):
        """[mean loss over ranks, top1, topk, n_valid] of the last step, after the all-reduce."""
        return self.flat.tail
