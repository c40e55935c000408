"""BASELINE config 5: flat gradient buffer all-reduce (+4-float metric tail) sweep, 0.1M-100M fp32
parameters, this framework's direct-NCCL path (pdt_allreduce_flat, one in-place call over the flat
buffer) next to torch.distributed.all_reduce (what DDP issues per bucket) on the same communicator
fabric, and the same direct call on a REGISTERED buffer (ncclMemAlloc + ncclCommRegister: NCCL may then reduce through
NVLS / peer mappings straight out of the user's buffer).  Run under torchrun with W = 2, 4 or 8; rank 0 prints one JSON
line per size.

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
        t_reg = None
        try:
            import ctypes
            from pdt.native import call, stream
            rptr, handle = ctypes.c_void_p(), ctypes.c_void_p()
            call("pdt_comm_mem_alloc", 4 * count, ctypes.byref(rptr))
            call("pdt_comm_register", comm.handle, rptr, 4 * count, ctypes.byref(handle))
            call("pdt_memcpy_async", rptr, b.data_ptr(), 4 * count, stream())     # any finite data
            fn = lambda: call("pdt_allreduce_flat", comm.handle, rptr, count, torch.cuda.current_stream().cuda_stream)
            # correctness on the registered buffer
            src = (torch.rand(count, device=dev, generator=g) * 2 - 1)
            chk = src.clone()
            call("pdt_memcpy_async", rptr, src.data_ptr(), 4 * count, stream())
            fn()
            out = torch.empty_like(src)
            call("pdt_memcpy_async", out.data_ptr(), rptr, 4 * count, stream())
            dist.all_reduce(chk)
            torch.cuda.synchronize()
            assert torch.allclose(out, chk, rtol=1e-6, atol=1e-6), "registered-buffer all-reduce disagrees"
            call("pdt_memcpy_async", rptr, torch.zeros(count, device=dev).data_ptr(), 4 * count, stream())
            t_reg = timed(fn, reps, world, dev)
            call("pdt_comm_deregister", comm.handle, handle)
            call("pdt_comm_mem_free", rptr)
        except Exception as e:          # registration is optional in NCCL: report, do not fail the sweep
            if rank == 0:
                print(json.dumps({"params": p, "registered_error": repr(e)[:200]}), flush=True)
        if rank == 0:
            s = 4 * count
            f = 2 * (world - 1) / world / 1e9
            print(json.dumps({"params": p, "bytes": s, "world": world,
                              "pdt_us": round(t_ours, 2), "pdt_bus_gbs": round(s / (t_ours * 1e-6) * f, 2),
                              "torch_us": round(t_torch, 2), "torch_bus_gbs": round(s / (t_torch * 1e-6) * f, 2),
                              "registered_us": None if t_reg is None else round(t_reg, 2),
                              "registered_bus_gbs": None if t_reg is None else round(s / (t_reg * 1e-6) * f, 2),
                              "nvlink_peak_gbs_per_dir": 900}), flush=True)
    dist.barrier()
    comm.destroy()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
