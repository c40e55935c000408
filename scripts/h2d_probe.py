"""Host->device transfer of one b1024 batch (3.06 MB) from pinned memory: one DMA copy, the copy split over
several streams, and a kernel that reads the pinned buffer in place (zero-copy over PCIe).  Explains when the
end-to-end step is transfer-bound rather than compute-bound on a given box."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "pytorch-distributed-template_b200"), ROOT]
import torch
from pdt import native

NB = 1024 * 784 * 4
host = torch.empty(NB // 4, dtype=torch.float32).pin_memory()
host.fill_(1)
dev = torch.empty(NB // 4, dtype=torch.float32, device="cuda")


def timed(label, fn, reps=200, nbytes=NB):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    s = torch.cuda.current_stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(s)
    for _ in range(reps):
        fn()
    e1.record(s)
    torch.cuda.synchronize()
    us = 1e3 * e0.elapsed_time(e1) / reps
    print("%-40s %7.1f us = %5.1f GB/s" % (label, us, nbytes / us / 1e3))


timed("one cudaMemcpyAsync", lambda: native.call("pdt_memcpy_async", dev.data_ptr(), host.data_ptr(), NB, native.stream()))
for k in (2, 4, 8):
    streams = [torch.cuda.Stream() for _ in range(k)]
    evs = [torch.cuda.Event() for _ in range(k)]
    start = torch.cuda.Event()
    chunk = NB // k

    def split():
        main = torch.cuda.current_stream()
        start.record(main)
        for i, st in enumerate(streams):
            st.wait_event(start)
            native.call("pdt_memcpy_async", dev.data_ptr() + i * chunk, host.data_ptr() + i * chunk, chunk, st.cuda_stream)
            evs[i].record(st)
            main.wait_event(evs[i])
    timed("copy split over %d streams" % k, split)
timed("kernel reads pinned memory (scale_cast)", lambda: native.call("pdt_flat_scale_cast", host.data_ptr(), dev.data_ptr(),
                                                                   NB // 4, 1.0, 0, native.stream()))
for mb in (0.25, 1, 16, 64):
    nb = int(mb * 2**20)
    h = torch.empty(nb, dtype=torch.uint8).pin_memory()
    d = torch.empty(nb, dtype=torch.uint8, device="cuda")
    timed("one cudaMemcpyAsync of %.2f MB" % mb, lambda: native.call("pdt_memcpy_async", d.data_ptr(), h.data_ptr(), nb, native.stream()),
          reps=50, nbytes=nb)
