"""Incremental device time of the step, stage by stage: graphs of growing prefixes of the kernel sequence
are replayed into an idle GPU; the differences are what each stage adds (kernel + graph edge)."""
import os, sys, struct
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "pytorch-distributed-template_b200"), ROOT]
import torch
from pdt import native, ops
from pdt.native import call, ptr, stream
from model.model import MnistModel
n = int(os.environ.get("B", 1024)); T = (n + 31) // 32
dev = torch.device("cuda")
g = torch.Generator().manual_seed(1)
x = torch.randn(n, 1, 28, 28, generator=g).to(dev); t = torch.randint(0, 10, (n,), generator=g).to(dev)
torch.manual_seed(0); m = MnistModel().to(dev); flat = m.flat.data; base = flat.data_ptr()
st = ops._DeviceState.get(dev); rng = st.rng_state()
k2 = torch.empty(n, 20, dtype=torch.uint8, device=dev); k1 = torch.empty(n, 50, dtype=torch.uint8, device=dev)
_, (p1, p2, hd, idx1, idx2) = ops._carve(dev, [T * 1440 * 128, T * 320 * 128, T * 50 * 128, T * 360 * 32, T * 320 * 32])
p1, p2, hd = p1.view(torch.float32), p2.view(torch.float32), hd.view(torch.float32)
logp = torch.empty(n, 10, device=dev); dlogp = torch.empty(n, 10, device=dev)
stats = torch.empty(8, dtype=torch.int32, device=dev)
dq2, dq1, scratch = st.bwd_workspace(n)
grad = torch.zeros(21848, device=dev); am, av, avm = (torch.zeros(21840, device=dev) for _ in range(3))
hyper = torch.zeros(64, dtype=torch.uint8, device=dev)
hyper[:44].copy_(torch.frombuffer(bytearray(struct.pack("dddddi", 1e-3, 0.9, 0.999, 1e-8, 0.0, 1)), dtype=torch.uint8))
nll_scr = st.nll_scratch(n)
side, (forked, joined) = st.side_stream()

hscr = st.head_nll_scratch(n)

def stages(upto, wg2=True):
    s = stream(); main = torch.cuda.current_stream()
    if upto >= 1:
        call("pdt_conv1_fwd_masks", ptr(x), base, base + 1000, ptr(p1), ptr(idx1), ptr(k2), ptr(k1), 0.5, ptr(rng), n, s)
    if upto >= 2:
        call("pdt_conv2_head_train", ptr(p1), base + 1040, base + 21040, ptr(k2), 2.0, ptr(p2), ptr(idx2), base + 4 * 5280,
             base + 4 * 21280, base + 4 * 21330, base + 4 * 21830, ptr(k1), 2.0, ptr(hd), ptr(logp), ptr(t), 3, -100,
             ptr(stats), ptr(hscr), None, None, 0, ptr(dq2), ptr(scratch), n, s)
    if upto >= 3:   # dgrad+wgrad1 (+ forked wgrad2 from stage 4 on)
        forked.record(main)
        call("pdt_conv2_bwd_data_conv1_bwd_weight", ptr(dq2), ptr(idx2), base + 1040, ptr(p1), ptr(x), ptr(idx1),
             ptr(scratch), n, s)
        if upto >= 4:
            side.wait_event(forked)
            call("pdt_conv2_bwd_weight", ptr(dq2), ptr(idx2), ptr(p1), ptr(scratch), n, side.cuda_stream)
            joined.record(side)
            main.wait_event(joined)
        if upto >= 5:
            call("pdt_adam_step_reduce", ptr(scratch), n, 0.0, 1.0, ptr(stats), ptr(grad), base, ptr(am), ptr(av), ptr(avm), ptr(hyper), None, s)

names = ["conv1+masks", "conv2+head train", "dgrad+wgrad1", "+wgrad2 (forked)", "reduce+adam"]
stages(5); torch.cuda.synchronize()
prev = 0.0
for k in range(1, 6):
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        stages(k)
    for _ in range(3):
        gr.replay()
    ts = []
    for _ in range(30):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); gr.replay(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort(); cur = ts[len(ts) // 2]
    print("%-18s prefix %.1f us   (+%.1f)" % (names[k - 1], cur, cur - prev))
    prev = cur
