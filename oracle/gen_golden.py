"""Generate tests/golden/*.npz by running the UNMODIFIED reference from /root/reference.

TEST INFRASTRUCTURE ONLY.  Run in the build container (where /root/reference is mounted):

    PYTHONDONTWRITEBYTECODE=1 python oracle/gen_golden.py

What runs is the reference's own code: model.model.MnistModel (model/model.py:6-22),
model.loss.nll_loss (model/loss.py:4-5), model.metric.accuracy / top_k_acc
(model/metric.py:4-20), trainer.Trainer._train_epoch (trainer/trainer.py:36-92),
BaseTrainer.reduce_loss (base/base_trainer.py:165-174) and torch.optim / DDP exactly as
train.py:41-52 wires them.  Nothing in /root/reference is edited.  Two things are done
from the outside:

  * torch.nn.functional.dropout / dropout2d are wrapped by a recorder that replays the
    RNG state on a tensor of ones to learn the keep mask the reference is about to use,
    then lets the original call proceed untouched (the reference's results are unchanged;
    we only learn the masks so the CUDA side can drop the same units);
  * utils.util.MetricTracker is replaced by a dict-based tracker of identical semantics
    because the original raises under pandas 3.

The fixtures record torch's version; they are small (a few hundred KB each).
"""
import os
import sys
import tempfile
import logging

import numpy as np
import torch
import torch.nn.functional as F

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")

sys.dont_write_bytecode = True
sys.path.insert(0, REF)

_orig_dropout, _orig_dropout2d = F.dropout, F.dropout2d
RECORDED = []


def _record(orig, kind):
    def wrapper(input, p=0.5, training=True, inplace=False):
        if training:
            state = torch.get_rng_state()
            probe = orig(torch.ones_like(input), p, True, False)
            torch.set_rng_state(state)
            keep = (probe != 0)
            if kind == "2d":
                keep = keep[:, :, 0, 0]
            RECORDED.append((kind, keep.to(torch.uint8).clone()))
        return orig(input, p, training, inplace)
    return wrapper


F.dropout = _record(_orig_dropout, "1d")
F.dropout2d = _record(_orig_dropout2d, "2d")


def pop_masks():
    k2 = [m for k, m in RECORDED if k == "2d"]
    k1 = [m for k, m in RECORDED if k == "1d"]
    RECORDED.clear()
    return k2, k1


def batch(n, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, 1, 28, 28, generator=g)
    t = torch.randint(0, 10, (n,), generator=g, dtype=torch.int64)
    return x, t


def state_np(model, prefix=""):
    return {prefix + k.replace("module.", ""): v.detach().cpu().numpy().copy()
            for k, v in model.state_dict().items()}


def run_steps(name, n, make_opt, steps, seed0):
    """trainer/trainer.py:48-58 driven by hand on the reference model/loss/optimizer."""
    from model.model import MnistModel
    from model.loss import nll_loss
    torch.manual_seed(0)
    model = MnistModel()
    opt = make_opt(model.parameters())
    model.train()
    rec = {}
    rec.update(state_np(model, "init/"))
    torch.manual_seed(777)
    for s in range(steps):
        x, t = batch(n, seed0 + s)
        opt.zero_grad()
        out = model(x)
        loss = nll_loss(out, t)
        loss.backward()
        opt.step()
        k2, k1 = pop_masks()
        rec[f"x/{s}"] = x.numpy()
        rec[f"t/{s}"] = t.numpy()
        rec[f"keep2d/{s}"] = k2[0].numpy()
        rec[f"keep1d/{s}"] = k1[0].numpy()
        rec[f"loss/{s}"] = np.float32(loss.item())
        rec[f"out/{s}"] = out.detach().numpy().copy()
        for k, p in model.named_parameters():
            rec[f"grad/{s}/{k}"] = p.grad.detach().numpy().copy()
    rec.update(state_np(model, "final/"))
    rec["steps"] = np.int64(steps)
    rec["torch_version"] = np.array(torch.__version__)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
    print("wrote", name)


def run_eval(name, n):
    """Eval-mode forward + the reference's metrics and loss, with ignore_index rows."""
    from model.model import MnistModel
    from model.loss import nll_loss
    from model.metric import accuracy, top_k_acc
    torch.manual_seed(0)
    model = MnistModel()
    model.eval()
    x, t = batch(n, 4242)
    with torch.no_grad():
        out = model(x)
    t_ign = t.clone()
    t_ign[::7] = -100
    rec = dict(state_np(model, "init/"))
    rec.update({
        "x": x.numpy(), "t": t.numpy(), "t_ign": t_ign.numpy(), "out": out.numpy(),
        "loss": np.float32(nll_loss(out, t).item()),
        "loss_ign": np.float32(nll_loss(out, t_ign).item()),
        "accuracy": np.float64(accuracy(out, t)),
        "top_k_acc": np.float64(top_k_acc(out, t)),
        "top_k_acc_k5": np.float64(top_k_acc(out, t, k=5)),
        "torch_version": np.array(torch.__version__),
    })
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
    print("wrote", name)


def _ddp_worker(rank, world, port, n, steps, ret):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                      RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, REF)
    F.dropout = _record(_orig_dropout, "1d")
    F.dropout2d = _record(_orig_dropout2d, "2d")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from model.model import MnistModel
    from model.loss import nll_loss
    from base.base_trainer import BaseTrainer
    torch.manual_seed(rank)            # DDP broadcasts rank 0's weights (train.py:46-52)
    if rank == 0:
        torch.manual_seed(0)
    model = MnistModel()
    opt = torch.optim.Adam(model.parameters(), lr=1e-3, weight_decay=0, amsgrad=True)
    ddp = torch.nn.parallel.DistributedDataParallel(model)
    rec = {}
    if rank == 0:
        rec.update(state_np(ddp, "init/"))
    ddp.train()
    torch.manual_seed(100 + rank)
    for s in range(steps):
        xg, tg = batch(n * world, 9000 + s)          # global batch; rank takes its rows
        x, t = xg[rank * n:(rank + 1) * n], tg[rank * n:(rank + 1) * n]
        opt.zero_grad()
        out = ddp(x)
        loss = nll_loss(out, t)
        red = BaseTrainer.reduce_loss(None, loss)     # base_trainer.py:165-174, unmodified
        loss.backward()
        opt.step()
        k2, k1 = pop_masks()
        rec[f"keep2d/{s}/r{rank}"] = k2[0].numpy()
        rec[f"keep1d/{s}/r{rank}"] = k1[0].numpy()
        rec[f"local_loss/{s}/r{rank}"] = np.float32(loss.item())
        if rank == 0:
            rec[f"x/{s}"] = xg.numpy()
            rec[f"t/{s}"] = tg.numpy()
            rec[f"reduced_loss/{s}"] = np.float32(red.item())
            for k, p in model.named_parameters():
                rec[f"grad/{s}/{k}"] = p.grad.detach().numpy().copy()
    if rank == 0:
        rec.update(state_np(ddp, "final/"))
    ret[rank] = rec
    dist.barrier()
    dist.destroy_process_group()


def run_ddp(name, n, steps, world=2):
    """The reference's data parallelism (train.py:46-52: torch DDP) on gloo/CPU, two ranks."""
    import torch.multiprocessing as mp
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_ddp_worker, args=(world, 29517, n, steps, ret), nprocs=world, join=True)
    rec = {}
    for r in range(world):
        rec.update(ret[r])
    rec["steps"] = np.int64(steps)
    rec["world"] = np.int64(world)
    rec["n_per_rank"] = np.int64(n)
    rec["torch_version"] = np.array(torch.__version__)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
    print("wrote", name)


class DictTracker:
    """Same semantics as utils/util.py:46-67 (running mean per key), without pandas."""

    def __init__(self, *keys, writer=None):
        self.writer = writer
        self.keys = keys
        self.reset()

    def reset(self):
        self.total = {k: 0.0 for k in self.keys}
        self.counts = {k: 0 for k in self.keys}

    def update(self, key, value, n=1):
        if self.writer is not None:
            self.writer.add_scalar(key, value)
        self.total[key] += value * n
        self.counts[key] += n

    def avg(self, key):
        return self.total[key] / self.counts[key]

    def result(self):
        return {k: (self.total[k] / self.counts[k] if self.counts[k] else 0) for k in self.keys}


def run_trainer(name, n, n_batches):
    """One epoch of the reference's Trainer (trainer/trainer.py:36-92) on CPU over a synthetic
    loader, template-default optimizer (config/config.json:38-45), validation included."""
    cwd = os.getcwd()
    os.chdir(REF)     # logger/logger.py:7 looks for its json relative to cwd
    try:
        import utils.util
        utils.util.MetricTracker = DictTracker
        import utils
        utils.MetricTracker = DictTracker
        import trainer.trainer as tt
        tt.MetricTracker = DictTracker
        from parse_config import ConfigParser
        from model.model import MnistModel
        from model.loss import nll_loss
        from model.metric import accuracy, top_k_acc
        from utils import read_json
        cfg = read_json(os.path.join(REF, "config", "config.json"))
        tmp = tempfile.mkdtemp()
        cfg["trainer"]["save_dir"] = tmp
        cfg["trainer"]["epochs"] = 1
        cfg["trainer"]["tensorboard"] = False
        cfg["trainer"]["save_period"] = 100
        config = ConfigParser(cfg, run_id="golden")
        logging.disable(logging.CRITICAL)

        class Loader(list):
            batch_size = n
            sampler = list(range(n * n_batches))

        train = Loader(batch(n, 5000 + i) for i in range(n_batches))
        valid = Loader(batch(n, 6000 + i) for i in range(2))
        valid.sampler = list(range(2 * n))
        torch.manual_seed(0)
        model = MnistModel()
        rec = dict(state_np(model, "init/"))
        opt = config.init_obj("optimizer", torch.optim, model.parameters())
        sched = config.init_obj("lr_scheduler", torch.optim.lr_scheduler, opt)
        tr = tt.Trainer(model, nll_loss, [accuracy, top_k_acc], opt, config=config,
                        device=torch.device("cpu"), data_loader=train,
                        valid_data_loader=valid, lr_scheduler=sched)
        torch.manual_seed(31337)
        log = tr._train_epoch(1)
        k2, k1 = pop_masks()
        for i in range(n_batches):
            rec[f"x/{i}"], rec[f"t/{i}"] = train[i][0].numpy(), train[i][1].numpy()
            rec[f"keep2d/{i}"], rec[f"keep1d/{i}"] = k2[i].numpy(), k1[i].numpy()
        for i in range(2):
            rec[f"vx/{i}"], rec[f"vt/{i}"] = valid[i][0].numpy(), valid[i][1].numpy()
        for k, v in log.items():
            rec["log/" + k] = np.float64(v)
        rec.update(state_np(model, "final/"))
        rec["n_batches"] = np.int64(n_batches)
        rec["torch_version"] = np.array(torch.__version__)
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **rec)
        print("wrote", name, {k: float(v) for k, v in log.items()})
    finally:
        logging.disable(logging.NOTSET)
        os.chdir(cwd)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(1)       # one thread: bit-reproducible reductions
    run_steps("ref_sgd_b37", 37, lambda p: torch.optim.SGD(p, lr=0.01), 3, 1234)
    run_steps("ref_sgdm_b32", 32, lambda p: torch.optim.SGD(p, lr=0.05, momentum=0.9, weight_decay=1e-4,
                                                         nesterov=True), 3, 2234)
    run_steps("ref_adam_b64", 64, lambda p: torch.optim.Adam(p, lr=1e-3, weight_decay=0, amsgrad=True), 3, 3234)
    run_eval("ref_eval_b50", 50)
    run_trainer("ref_trainer_b24", 24, 3)
    run_ddp("ref_ddp2_b16", 16, 2)
