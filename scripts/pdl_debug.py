import os, sys, tempfile, logging
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "pytorch-distributed-template_b200"), ROOT, os.path.join(ROOT, "tests")]
import torch
from oracle import ref_step as R
from util_gpu import make_model
from test_gpu_trainer import make_trainer, ListLoader
from pdt import ops
import pathlib
for graph in (False, True):
    params = R.init_params(0)
    model = make_model(params)
    ops.seed_dropout(99)
    loader = ListLoader(); loader.batch_size = 64
    for i in range(6):
        x, t, _, _ = R.synthetic_batch(64, 40 + i)
        loader.append((x.pin_memory(), t.pin_memory()))
    tr = make_trainer(pathlib.Path(tempfile.mkdtemp()), model, loader, graph=graph)
    tr.sync_every_step = True
    losses = []
    orig = tr.train_metrics.update
    tr.train_metrics.update = lambda k, v, n=1: (losses.append(v), orig(k, v, n))[1]
    logging.disable(logging.CRITICAL)
    log = tr._train_epoch(1)
    print("PDL", os.environ.get("PDT_PDL"), "graph", graph, ["%.6f" % l for l in losses], "w", float(model.flat.data[:21840].double().sum()))
