"""CPU cost of the host-loop primitives (us per call), to see what bounds the end-to-end step."""
import os, sys, time, argparse
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "pytorch-distributed-template_b200"), ROOT]
import torch, bench
from pdt import native
dev = torch.device("cuda")
args = argparse.Namespace(batch=1024, optimizer="adam", steps=50, warmup=20)
tr, loader = bench.build_trainer(args, dev, 1, 8)
tr._train_epoch(1)
slot = [s for ring in tr._stager.slots.values() for s in ring if s.graph is not None][0]
x, t = loader.batches[0]
cs = tr._stager.copy_stream
main = torch.cuda.current_stream()
ev = torch.cuda.Event()
def bench_fn(name, fn, n=2000, sync_every=50):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for i in range(n):
        fn()
        if i % sync_every == sync_every - 1: torch.cuda.synchronize()
    torch.cuda.synchronize()
    print("%-28s %.2f us" % (name, 1e6 * (time.perf_counter() - t0) / n))
bench_fn("empty lambda", lambda: None)
bench_fn("event.record(main)", lambda: ev.record(main))
bench_fn("main.wait_event", lambda: main.wait_event(ev))
bench_fn("x.is_pinned()", lambda: x.is_pinned())
bench_fn("x.data_ptr()", lambda: x.data_ptr())
bench_fn("native memcpy 8KB", lambda: native.call("pdt_memcpy_async", slot.target.data_ptr(), t.data_ptr(), 8192, cs.cuda_stream))
bench_fn("native memcpy 3.2MB", lambda: native.call("pdt_memcpy_async", slot.data.data_ptr(), x.data_ptr(), x.numel() * 4, cs.cuda_stream), n=500)
bench_fn("torch copy_ 8KB", lambda: slot.target.copy_(t, non_blocking=True))
bench_fn("graph.replay()", lambda: slot.graph.replay(), n=500)
bench_fn("stage()", lambda: tr._stager.stage((x, t), 0), n=500)
bench_fn("native.stream()", lambda: native.stream())
