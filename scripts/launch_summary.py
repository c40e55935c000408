"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list next to bench.py's own graph-timed kernels:
per-kernel count, mean duration and each step kernel's SHARE (ncu's launches are cold-cache and serialised, so only
the shares are comparable).  usage: launch_summary.py launches.csv bench_line.json > summary.txt"""
import collections
import csv
import json
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
h = rows[hi]
kn, mv, mu = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[hi + 1:]:
    if len(r) <= mv:
        continue
    name = r[kn].split("(")[0].replace("pdt::", "").replace("void ", "")
    v = float(r[mv].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}[r[mu]]
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += v
step = ["conv1_fwd_masks_kernel", "conv2_head_train_kernel", "conv2_bwd_data_wgrad1_kernel", "conv2_bwd_weight_kernel",
        "reduce_adam_kernel"]
in_step = lambda k: any(k.startswith(s) for s in step)
tot = sum(t / c for k, (c, t) in agg.items() if in_step(k))
print("%-52s %6s %10s %8s" % ("kernel (ncu launch list)", "count", "avg us", "share"))
for k, (c, t) in agg.items():
    print("%-52s %6d %10.2f %8s" % (k[:52], c, t / c, ("%.1f %%" % (100 * t / c / tot)) if in_step(k) else ""))
bench = json.load(open(sys.argv[2]))
bk = {k["kernel"]: k["us"] for k in bench["kernels"] if k["in_step"]}
print("\nbench.py (%s), the step's launches timed alone in a CUDA graph:" % sys.argv[2])
for k, v in bk.items():
    print("%-52s %17.2f %7.1f %%" % (k, v, 100 * v / sum(bk.values())))
