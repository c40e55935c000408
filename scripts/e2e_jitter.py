"""Is the end-to-end loop disturbed by the nvidia-smi sampler?  Times epochs with PDT_SMI_MS = 20 / 200 / off."""
import os, sys, time, argparse
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "pytorch-distributed-template_b200"), ROOT]
import torch, bench

class Null:
    def __enter__(self): return self
    def __exit__(self, *a): pass

dev = torch.device("cuda")
args = argparse.Namespace(batch=1024, optimizer="adam", steps=400, warmup=20)
tr, loader = bench.build_trainer(args, dev, 1, 8)
for mode in ("off", "20", "off", "200", "off", "20"):
    if mode == "off":
        s = Null()
    else:
        os.environ["PDT_SMI_MS"] = mode
        s = bench.ClockSampler(0)
        time.sleep(0.3)
    ms, log, runs = bench.measure_e2e(tr, loader, args, dev, 1, s)
    print("sampler %-4s epochs us/step %s" % (mode, [round(1e3 * r / args.steps, 1) for r in runs]), flush=True)
    if mode != "off":
        s.summary()
