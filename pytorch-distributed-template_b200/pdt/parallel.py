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
import os
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


class P2PComm:
    """One-shot all-reduce over NVLink peer memory (pdt_p2p_*): every rank owns an exchange block allocated
    by the library, the CUDA IPC handles travel through the existing torch process group, and each call is
    ONE kernel.  Meant for small flat buffers (capacity_floats), where NCCL is latency-bound."""

    MAX_WORLD = 8

    def __init__(self, device, capacity_floats):
        self.world_size = dist.get_world_size()
        self.rank = dist.get_rank()
        self.device = device
        self.capacity = int(capacity_floats)
        if self.world_size > self.MAX_WORLD:
            raise native.PdtError("P2PComm supports at most %d ranks on one node" % self.MAX_WORLD)
        own = ctypes.c_void_p()
        handle = ctypes.create_string_buffer(64)
        with torch.cuda.device(device):
            call("pdt_p2p_alloc", self.capacity, ctypes.byref(own), ctypes.cast(handle, ctypes.c_void_p))
        self.own = own.value
        handles = [None] * self.world_size
        dist.all_gather_object(handles, bytes(handle.raw))
        self.bases = (ctypes.c_void_p * self.world_size)()
        self._opened = []
        with torch.cuda.device(device):
            for q, h in enumerate(handles):
                if q == self.rank:
                    self.bases[q] = self.own
                else:
                    peer = ctypes.c_void_p()
                    hb = ctypes.create_string_buffer(h, 64)
                    call("pdt_p2p_open", ctypes.cast(hb, ctypes.c_void_p), ctypes.byref(peer))
                    self.bases[q] = peer.value
                    self._opened.append(peer.value)
        dist.barrier()

    def allreduce_(self, tensor, count=None):
        count = tensor.numel() if count is None else count
        assert tensor.dtype == torch.float32 and tensor.is_contiguous()
        call("pdt_p2p_allreduce", ctypes.cast(self.bases, ctypes.c_void_p), self.rank, self.world_size,
             tensor.data_ptr(), count, self.capacity, stream())

    def reduce_adam_(self, pending, fb, state):
        """Backward's partial gradients -> flat gradient -> sum over ranks -> Adam, in ONE kernel."""
        scratch, n, beta, scale, stats, gscale = pending
        call("pdt_p2p_reduce_adam", ctypes.cast(self.bases, ctypes.c_void_p), self.rank, self.world_size,
             scratch.data_ptr(), n, float(beta), float(scale), native.ptr(stats), native.ptr(gscale),
             fb.grad.data_ptr(), fb.data.data_ptr(), native.ptr(state["m"]), native.ptr(state["v"]),
             native.ptr(state["vmax"]), native.ptr(state["hyper"]), self.capacity, stream())

    def check(self):
        """Raise if a wait inside the kernel timed out on this or any other rank (a rank missed a call or fell
        behind by more than the time-out).  From that call on nothing was summed and no weight was updated, so
        the replicas are intact up to the step before; torch DDP would have blocked instead (train.py:46-52).
        Reads 4 bytes from the device: synchronises with the work queued so far."""
        err = ctypes.c_int(0)
        with torch.cuda.device(self.device):
            call("pdt_p2p_error", self.own, ctypes.byref(err))
        if err.value:
            raise native.PdtError(self.TIMEOUT_MESSAGE)

    TIMEOUT_MESSAGE = ("peer-memory all-reduce: a rank did not arrive within the time-out (PDT_P2P_TIMEOUT_MS); "
                       "the gradient exchange and every update since were skipped")

    def error_async(self, pinned_int32):
        """Queue a copy of the error word into a pinned int32[1] tensor on the current stream (no synchronisation);
        the caller looks at it once the stream has passed that point."""
        call("pdt_p2p_error_async", self.own, pinned_int32.data_ptr(), stream())

    def destroy(self):
        with torch.cuda.device(self.device):
            for p in self._opened:
                call("pdt_p2p_close", p)
            self._opened = []
            if self.own:
                call("pdt_p2p_free", self.own)
                self.own = None


class _DeviceMemory:
    """Device memory that torch did not allocate, presented through __cuda_array_interface__ so that
    torch.as_tensor can view it (the view keeps this object alive; the memory itself is released by its owner)."""

    def __init__(self, ptr, numel):
        self.__cuda_array_interface__ = {"shape": (int(numel),), "typestr": "<f4", "data": (int(ptr), False), "version": 2}


class FlatDataParallel(nn.Module):
    P2P_MAX_FLOATS = 262144      # <= 1 MB goes over peer memory in one kernel; beyond that NCCL wins (sweep)
    # From here on the flat gradient buffer is taken from ncclMemAlloc and registered on the communicator
    # (ncclCommRegister): NCCL then reduces straight out of and into it (NVLS / peer mappings).  Measured on 8 B200s
    # (profiles/r02_allreduce_sweep_w8.jsonl): 4 MB 118 -> 126, 12 MB 225 -> 274, 40 MB 359 -> 504, 120 MB 565 -> 621
    # GB/s bus bandwidth; no difference at 2 or 4 ranks, none at 400 MB.
    REGISTER_MIN_FLOATS = 1 << 20

    def __init__(self, module, device_ids=None, output_device=None, broadcast_buffers=True, comm="auto",
                 register_buffer="auto"):
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
        n_reduce = self.flat.numel + _flat.TAIL
        use_p2p = comm == "p2p" or (comm == "auto" and n_reduce <= self.P2P_MAX_FLOATS
                                    and self.world_size <= P2PComm.MAX_WORLD
                                    and os.environ.get("PDT_P2P", "1") != "0")
        self.p2p = P2PComm(self.device, (n_reduce + 3) // 4 * 4) if use_p2p else None
        self._registered = None
        if register_buffer is True or (register_buffer == "auto" and self.p2p is None
                                       and n_reduce >= self.REGISTER_MIN_FLOATS):
            self._register_grad_buffer()
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

    def _register_grad_buffer(self):
        """Move the flat gradient buffer into memory from ncclMemAlloc, registered on the communicator; every
        parameter's .grad is re-pointed at it.  The memory lives until release_registered_buffer() or process exit."""
        import ctypes
        fb = self.flat
        numel = fb.grad.numel()
        mem, handle = ctypes.c_void_p(), ctypes.c_void_p()
        with torch.cuda.device(self.device):
            call("pdt_comm_mem_alloc", 4 * numel, ctypes.byref(mem))
            call("pdt_comm_register", self.comm.handle, mem, 4 * numel, ctypes.byref(handle))
            view = torch.as_tensor(_DeviceMemory(mem.value, numel), device=self.device)
        view.copy_(fb.grad)
        fb.grad = view
        fb.attach()
        self._registered = (mem, handle)

    def release_registered_buffer(self):
        if self._registered is not None:
            mem, handle = self._registered
            self._registered = None
            fb = self.flat
            plain = fb.grad.clone()
            fb.grad = plain
            fb.attach()
            torch.cuda.synchronize(self.device)
            with torch.cuda.device(self.device):
                call("pdt_comm_deregister", self.comm.handle, handle)
                call("pdt_comm_mem_free", mem)

    def forward(self, *args, **kwargs):
        self.flat.check_alive()
        if not self.native:
            if self._pending != 0:
                # the last backward produced a gradient for only some parameters, so the all-reduce never ran and the
                # optimizer stepped on unreduced local gradients; torch DDP raises in the same situation
                # (find_unused_parameters=False, the reference's setting: train.py:46-52)
                got, self._pending = self._pending, 0
                raise RuntimeError("FlatDataParallel: the previous backward produced gradients for %d of %d parameters; "
                                   "parameters that take no part in the loss are not supported (the gradient all-reduce "
                                   "did not run for that step)" % (got, self._n_params))
        if self.broadcast_buffers and self.training:
            for b in self.module.buffers():
                dist.broadcast(b, src=0)
        return self.module(*args, **kwargs)

    # native model: called from the model's backward once every gradient is in the flat buffer
    def _allreduce_flat(self):
        n = self.flat.numel + _flat.TAIL
        if self.p2p is not None:
            self.p2p.allreduce_(self.flat.grad, (n + 3) // 4 * 4)
        else:
            self.comm.allreduce_(self.flat.grad, n)

    def allreduce_native(self):
        self._allreduce_flat()

    # any other module: count accumulated gradients, then scale by 1/W and all-reduce once
    def _on_grad(self, param):
        self._pending += 1
        if self._pending == self._n_params:
            self._pending = 0
            g = self.flat.grad
            call("pdt_flat_scale_cast", g.data_ptr(), g.data_ptr(), self.flat.numel, 1.0 / self.world_size, 0, stream())
            self._allreduce_flat()

    def reduced_tail(self):
        """[mean loss over ranks, top1, topk, n_valid] of the last step, after the all-reduce."""
        return self.flat.tail
