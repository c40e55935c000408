"""Times conv2's forward on the CUDA cores and on the tensor cores (graph-timed, 40 launches each); with a library
built with -DPDT_TC_STAMPS (make PTXAS_V=-DPDT_TC_STAMPS) also prints the tensor-core kernel's phase timeline."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "pytorch-distributed-template_b200"), ROOT, os.path.join(ROOT, "tests")]
import torch
from pdt import ops
from pdt.native import call, ptr, stream
from model.model import MnistModel


def graph_time(fn, reps=40):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    g.replay()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, 1e3 * e0.elapsed_time(e1) / reps)
    return best


for n in (128, 1024, 4096):
    dev = torch.device("cuda")
    torch.manual_seed(0)
    m = MnistModel().to(dev)
    x = torch.randn(n, 1, 28, 28, device=dev)
    k2 = (torch.rand(n, 20, device=dev) < 0.5).to(torch.uint8)
    k1 = (torch.rand(n, 50, device=dev) < 0.5).to(torch.uint8)
    _, saved = ops.mnist_forward_raw(x, m.flat.data, k2, k1, True)
    p1, idx1, p2, idx2 = saved[:4]
    base = m.flat.data.data_ptr()
    for name in ("pdt_conv2_drop_pool_relu_fwd", "pdt_conv2_drop_pool_relu_fwd_tc"):
        us = graph_time(lambda: call(name, ptr(p1), base + 1040, base + 21040, ptr(k2), 2.0, ptr(p2), ptr(idx2), n,
                                     torch.cuda.current_stream().cuda_stream))
        print("n=%d %s %.2f us  (%.1f TFLOP/s of the 640 kFLOP/sample)" % (n, name, us, 640e3 * n / us / 1e6))

import ctypes
from pdt import native
lib = native.load()
buf = (ctypes.c_longlong * 64)()
n = 1024
call("pdt_conv2_drop_pool_relu_fwd_tc", ptr(p1), base + 1040, base + 21040, ptr(k2), 2.0, ptr(p2), ptr(idx2), n,
     torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
lib.pdt_debug_tc_stamps(buf)
st = list(buf)[:25]
names = {1: "weights staged", 2: "A0 built+sync", 24: "written"}
for p in range(4):
    names.update({4 + 5 * p: "p%d next built" % p, 7 + 5 * p: "p%d mma done" % p, 8 + 5 * p: "p%d epilogue" % p})
prev = st[0]
for i in sorted(names):
    print("%-16s +%6d cycles (total %6d)" % (names[i], st[i] - prev, st[i] - st[0]))
    prev = st[i]

# accuracy of both kernels against the fp64 evaluation (relative to the output's maximum)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_gpu_conv2_tc import _run
from util_gpu import max_err
for nn, train in ((128, True), (1024, False), (1024, True)):
    p2_tc, _, p2_core, _, ref = _run(nn, 4000 + nn, train)
    print("accuracy n=%d train=%s: tensor cores %.2e, CUDA cores %.2e" % (nn, train, max_err(p2_tc, ref), max_err(p2_core, ref)))
