"""Where does the end-to-end step time go?  Times Trainer._train_epoch variants."""
import os, sys, time, argparse
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "pytorch-distributed-template_b200"), ROOT]
import torch, bench
args = argparse.Namespace(batch=int(os.environ.get("B", 1024)), optimizer="adam", steps=300, warmup=20)
dev = torch.device("cuda")
tr, loader = bench.build_trainer(args, dev, 1, 8)
base = list(loader.batches)
def run(n, label, cpu_only=False):
    loader.batches = [base[i % 8] for i in range(n)]; loader.n_batches = n; tr.len_epoch = n
    torch.cuda.synchronize(); t0 = time.perf_counter()
    tr._train_epoch(2)
    t_cpu = time.perf_counter() - t0
    torch.cuda.synchronize(); t1 = time.perf_counter() - t0
    print("%-28s cpu-return %.1f us/step   total %.1f us/step" % (label, 1e6 * t_cpu / n, 1e6 * t1 / n))
run(20, "warmup")
run(300, "default")
tr.log_step = 10 ** 9
run(300, "no log flush")
# device-resident batches: no H2D bytes, same loop
loader.batches = None
base = [(x.to(dev), t.to(dev)) for x, t in base]
run(300, "device-resident inputs")
# per-step device timeline and the effect of the Python GC
import gc
loader.batches = None
base = list(bench.build_trainer(args, dev, 1, 8)[1].batches)
def timeline(label):
    n = 400
    loader.batches = [base[i % 8] for i in range(n)]; loader.n_batches = n; tr.len_epoch = n
    evs = []
    orig = tr._run_staged
    def wrapped(slot):
        e = torch.cuda.Event(enable_timing=True); e.record(torch.cuda.current_stream()); evs.append(e)
        return orig(slot)
    tr._run_staged = wrapped
    torch.cuda.synchronize(); t0 = time.perf_counter()
    tr._train_epoch(2)
    torch.cuda.synchronize(); wall = time.perf_counter() - t0
    tr._run_staged = orig
    d = sorted(evs[i].elapsed_time(evs[i + 1]) * 1e3 for i in range(len(evs) - 1))
    print("%-12s wall %.1f us/step; step gaps us: median %.1f p90 %.1f p99 %.1f max %.1f sum/n %.1f" % (
        label, 1e6 * wall / n, d[len(d) // 2], d[int(0.9 * len(d))], d[int(0.99 * len(d))], d[-1], sum(d) / len(d)))
timeline("gc on")
gc.collect(); gc.disable()
timeline("gc off")
timeline("gc off")
