"""Distributed helpers with the reference's five entry points (utils/dist.py:7-74).

`all_gather` is device-agnostic here: the reference hard-codes "cuda" (utils/dist.py:49-65) and so
cannot run under gloo; this version stages the pickled bytes on whatever device the backend
needs.  `reduce_tensor` is an addition named by the task; the reference has no counterpart.
"""
import pickle

import torch
import torch.distributed as dist


def _active():
    return dist.is_available() and dist.is_initialized()


def get_world_size():
    return dist.get_world_size() if _active() else 1


def get_rank():
    return dist.get_rank() if _active() else 0


def is_main_process():
    return get_rank() == 0


def synchronize():
    """Barrier across ranks; a no-op for single-process runs."""
    if _active() and dist.get_world_size() > 1:
        dist.barrier()


def _comm_device():
    return torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")


def all_gather(data):
    """Gather one picklable object per rank into a list ordered by rank."""
    world = get_world_size()
    if world == 1:
        return [data]
    dev = _comm_device()
    payload = torch.frombuffer(bytearray(pickle.dumps(data)), dtype=torch.uint8).to(dev)
    mine = torch.tensor([payload.numel()], dtype=torch.int64, device=dev)
    sizes = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(sizes, mine)
    sizes = [int(s.item()) for s in sizes]
    width = max(sizes)
    if payload.numel() < width:
        payload = torch.cat([payload, torch.zeros(width - payload.numel(), dtype=torch.uint8, device=dev)])
    slots = [torch.empty(width, dtype=torch.uint8, device=dev) for _ in range(world)]
    dist.all_gather(slots, payload)
    return [pickle.loads(slot[:size].cpu().numpy().tobytes()) for size, slot in zip(sizes, slots)]


def reduce_tensor(tensor, average=True):
    """Sum (or mean) of `tensor` over ranks, returned on every rank."""
    world = get_world_size()
    if world == 1:
        return tensor
    out = tensor.clone()
    dist.all_reduce(out)
    if average:
        out /= world
    return out
