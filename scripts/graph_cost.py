"""Is the step graph's replay bound by the host (cudaGraphLaunch cost) or by the GPU?"""
import os, sys, time, argparse
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "pytorch-distributed-template_b200"), ROOT]
import torch, bench
dev = torch.device("cuda")
args = argparse.Namespace(batch=int(os.environ.get("B", 1024)), optimizer="adam", steps=50, warmup=20)
tr, loader = bench.build_trainer(args, dev, 1, 8, cuda_graph=False)
x, t = loader.batches[0]
x, t = x.to(dev), t.to(dev)
tr.model.train()
tr._step(x, t)
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    tr._step(x, t)
for _ in range(5):
    g.replay()
torch.cuda.synchronize()
# (a) host cost of a launch into an idle GPU, (b) GPU duration of one replay, (c) back-to-back rate
n = 200
t_host = 0.0
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
for i in range(n):
    torch.cuda.synchronize()
    ev[i][0].record()
    t0 = time.perf_counter()
    g.replay()
    t_host += time.perf_counter() - t0
    ev[i][1].record()
torch.cuda.synchronize()
gpu = sorted(a.elapsed_time(b) * 1e3 for a, b in ev)
print("host cost of replay() into an idle GPU: %.1f us; GPU time of one isolated replay: median %.1f us" % (1e6 * t_host / n, gpu[n // 2]))
torch.cuda.synchronize(); t0 = time.perf_counter()
for i in range(1000):
    g.replay()
t_ret = time.perf_counter() - t0
torch.cuda.synchronize(); t_all = time.perf_counter() - t0
print("1000 back-to-back replays: host returned after %.1f us/replay, all done after %.1f us/replay" % (1e3 * t_ret, 1e3 * t_all))
