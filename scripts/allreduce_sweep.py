"""BASELINE config 5: flat gradient buffer all-reduce (+4-float metric tail) sweep, 0.1M-100M fp32
parameters, this framework's direct-NCCL path (pdt_allreduce_flat, one in-place call over the flat
buffer) next to torch.distributed.all_reduce (what DDP issues per bucket) on the same communicator
fabric.  Run under torchrun with W = 2, 4 or 8; rank 0 prints one JSON line per size.

busBW = (S / t) * 2 (W - 1) / W   (nccl-tests convention), S = 4 * (P + 4) bytes.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "pytorch-distributed-template_b200"), ROOT]
import torch
import torch.distributed as dist


def timed(fn, reps, world, dev):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    g.replay()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()) * 1e3   # us


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", init_method="env://", device_id=dev)
    from pdt.parallel import NcclComm
    comm = NcclComm(dev)
    sizes = [100_000, 300_000, 1_000_000, 3_000_000, 10_000_000, 30_000_000, 100_000_000]
    if "--mnist" in sys.argv:
        sizes = [21_840] + sizes
    for p in sizes:
        count = p + 4
        g = torch.Generator(device=dev).manual_seed(rank)
        buf = (torch.rand(count, device=dev, generator=g) * 2 - 1)
        ref = buf.clone()
        # correctness of the direct path against torch's collective on the same data
        a, b = buf.clone(), buf.clone()
        comm.allreduce_(a)
        dist.all_reduce(b)
        torch.cuda.synchronize()
        assert torch.allclose(a, b, rtol=1e-6, atol=1e-6), "direct NCCL path disagrees with torch all_reduce"
        reps = 20 if p <= 3_000_000 else 5
        t_ours = timed(lambda: comm.allreduce_(buf), reps, world, dev)
        t_torch = timed(lambda: dist.all_reduce(ref), reps, world, dev)
        if rank == 0:
            s = 4 * count
            f = 2 * (world - 1) / world / 1e9
            print(json.dumps({"params": p, "bytes": s, "world": world,
                              "pdt_us": round(t_ours, 2), "pdt_bus_gbs": round(s / (t_ours * 1e-6) * f, 2),
                              "torch_us": round(t_torch, 2), "torch_bus_gbs": round(s / (t_torch * 1e-6) * f, 2),
                              "nvlink_peak_gbs_per_dir": 900}), flush=True)
    dist.barrier()
    comm.destroy()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
