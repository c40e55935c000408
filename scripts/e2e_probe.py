"""Where does the end-to-end step time go?  Times Trainer._train_epoch variants."""
import os, sys, time, argparse
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "pytorch-distributed-template_b200"), ROOT]
import torch, bench
args = argparse.Namespace(batch=1024, optimizer="adam", steps=300, warmup=20)
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
