"""The drop-in Trainer on the GPU against the reference's own Trainer epoch (golden fixture made
by oracle/gen_golden.py: trainer/trainer.py:36-92 on CPU, template-default Adam+amsgrad)."""
import logging
import os

import numpy as np
import pytest
import torch

from oracle import ref_step as R
from test_gpu_parity import check_adam_weights
from util_gpu import close, make_model

pytestmark = pytest.mark.gpu


class ListLoader(list):
    batch_size = 24
    sampler = None


def make_trainer(tmp_path, model, loader, valid=None, graph=True, **trainer_cfg):
    from parse_config import ConfigParser
    from utils import read_json
    import model.loss as L
    import model.metric as M
    from pdt import optim as O
    from trainer import Trainer
    pkg = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "pytorch-distributed-template_b200")
    cfg = read_json(os.path.join(pkg, "config", "config.json"))
    cfg["trainer"].update(save_dir=str(tmp_path), tensorboard=False, epochs=1, cuda_graph=graph, **trainer_cfg)
    config = ConfigParser(cfg, run_id="t")
    opt = config.init_obj("optimizer", O, model.parameters())
    sched = config.init_obj("lr_scheduler", torch.optim.lr_scheduler, opt)
    tr = Trainer(model, L.nll_loss, [M.accuracy, M.top_k_acc], opt, config=config, device=torch.device("cuda"),
                 data_loader=loader, valid_data_loader=valid, lr_scheduler=sched)
    return tr


def test_trainer_epoch_matches_reference_trainer(tmp_path, golden_dir):
    g = np.load(os.path.join(golden_dir, "ref_trainer_b24.npz"))
    nb = int(g["n_batches"])
    params = {k: torch.from_numpy(g["init/" + k]) for k, _ in R.PARAM_SHAPES}
    model = make_model(params)
    loader = ListLoader((torch.from_numpy(g[f"x/{i}"]), torch.from_numpy(g[f"t/{i}"])) for i in range(nb))
    valid = ListLoader((torch.from_numpy(g[f"vx/{i}"]), torch.from_numpy(g[f"vt/{i}"])) for i in range(2))
    # eager: injected masks are fresh tensors per step (graph replay is covered by the next test)
    tr = make_trainer(tmp_path, model, loader, valid, graph=False)
    masks = [(torch.from_numpy(g[f"keep2d/{i}"]).cuda(), torch.from_numpy(g[f"keep1d/{i}"]).cuda()) for i in range(nb)]
    model.inject_dropout_masks(masks)
    log = tr._train_epoch(1)
    assert abs(log["loss"] - float(g["log/loss"])) <= 1e-5 * abs(float(g["log/loss"]))
    assert log["val_accuracy"] == float(g["log/val_accuracy"])
    assert log["val_top_k_acc"] == float(g["log/val_top_k_acc"])
    assert log["val_loss"] == float(g["log/val_loss"]) == 0.0   # reference never updates it
    check_adam_weights(model.state_dict(), g, nb, R.Adam(lr=1e-3, weight_decay=0.0, amsgrad=True))


def test_graph_replay_equals_eager(tmp_path):
    """The CUDA-graph Trainer path gives bit-identical weights to the eager path when both draw the
    same dropout masks (same generator seed and offsets)."""
    from pdt import ops
    results = []
    for graph in (False, True):
        params = R.init_params(0)
        model = make_model(params)
        ops.seed_dropout(99)
        loader = ListLoader()
        loader.batch_size = 64
        for i in range(6):
            x, t, _, _ = R.synthetic_batch(64, 40 + i)
            loader.append((x.pin_memory(), t.pin_memory()))
        tr = make_trainer(tmp_path / str(graph), model, loader, graph=graph)
        log = tr._train_epoch(1)
        results.append((log["loss"], model.flat.data[:21840].clone()))
        if graph:
            assert sum(s.graph is not None for ring in tr._stager.slots.values() for s in ring) == 2
    assert results[0][0] == results[1][0]
    assert torch.equal(results[0][1], results[1][1])


def test_lr_schedule_reaches_replayed_graph(tmp_path):
    """StepLR(step_size=1) over three epochs: the CUDA-graph Trainer (every full batch after the first is a
    replay that never passes through optimizer.step()) must end with the same weights as the eager Trainer,
    and both must differ from a run whose learning rate never changes (reference trainer.py:90-91)."""
    from pdt import ops
    finals = {}
    for name, graph, gamma in (("eager", False, 0.1), ("graph", True, 0.1), ("graph_const_lr", True, 1.0)):
        model = make_model(R.init_params(0))
        ops.seed_dropout(321)
        loader = ListLoader()
        loader.batch_size = 64
        for i in range(5):
            x, t, _, _ = R.synthetic_batch(64, 60 + i)
            loader.append((x.pin_memory(), t.pin_memory()))
        tr = make_trainer(tmp_path / name, model, loader, graph=graph)
        tr.lr_scheduler = torch.optim.lr_scheduler.StepLR(tr.optimizer, step_size=1, gamma=gamma)
        for epoch in (1, 2, 3):
            tr._train_epoch(epoch)
        assert abs(tr.optimizer.param_groups[0]["lr"] - 1e-3 * gamma ** 3) < 1e-12
        assert tr.optimizer.device_step() == 15
        finals[name] = model.flat.data[:21840].clone()
    assert torch.equal(finals["eager"], finals["graph"])
    assert not torch.equal(finals["graph"], finals["graph_const_lr"])


def test_trainer_ragged_last_batch_and_checkpoint(tmp_path):
    params = R.init_params(0)
    model = make_model(params)
    loader = ListLoader()
    loader.batch_size = 48
    for i, n in enumerate([48, 48, 48, 17]):
        x, t, _, _ = R.synthetic_batch(n, 70 + i)
        loader.append((x, t))
    tr = make_trainer(tmp_path, model, loader, save_period=1)
    tr.train()
    ckpt = torch.load(str(tr.checkpoint_dir / "checkpoint-epoch1.pth"), weights_only=True)
    assert set(ckpt) >= {"arch", "epoch", "state_dict", "optimizer", "monitor_best", "config"}
    assert list(ckpt["state_dict"].keys()) == [k for k, _ in R.PARAM_SHAPES]
    assert ckpt["optimizer"]["state"][0]["step"].item() == 4
    for k, v in ckpt["state_dict"].items():
        assert torch.equal(v.cuda(), model.state_dict()[k])
    # resume into a fresh model/optimizer
    model2 = make_model(seed=3)
    from parse_config import ConfigParser
    tr2 = make_trainer(tmp_path / "b", model2, loader)
    tr2._resume_checkpoint(tr.checkpoint_dir / "checkpoint-epoch1.pth")
    assert tr2.start_epoch == 2
    assert torch.equal(model2.flat.data[:21840], model.flat.data[:21840])
    assert tr2.optimizer.device_step() == 4
    logging.disable(logging.NOTSET)


def test_validation_counts_equal_gather_path_and_oracle(tmp_path):
    """Validation metrics from per-rank integer counts (N1) == the reference's gather-everything path ==
    the oracle's metrics on the concatenated outputs; test.py's evaluate() gives the same plus the loss."""
    import importlib.util
    params = R.init_params(0)
    valid = ListLoader()
    valid.batch_size = 50
    xs, ts = [], []
    for i, n in enumerate([50, 50, 23]):
        x, t, _, _ = R.synthetic_batch(n, 900 + i)
        valid.append((x, t))
        xs.append(x)
        ts.append(t)
    train = ListLoader([valid[0]])
    train.batch_size = 50
    logs = []
    for counts in (True, False):
        model = make_model(params)
        tr = make_trainer(tmp_path / str(counts), model, train, valid, count_metrics=counts, val_loss=True)
        tr.optimizer.param_groups[0]["lr"] = 0.0          # keep the weights: both runs validate the same model
        logs.append(tr._train_epoch(1))
    ref_out = R.mnist_forward(params, torch.cat(xs))
    ref_t = torch.cat(ts)
    for log in logs:
        assert log["val_accuracy"] == R.accuracy(ref_out, ref_t)
        assert log["val_top_k_acc"] == R.top_k_acc(ref_out, ref_t)
    assert abs(logs[0]["val_loss"] - float(R.nll_loss(ref_out, ref_t))) < 1e-5 * abs(float(R.nll_loss(ref_out, ref_t)))
    assert logs[1]["val_loss"] == 0.0                      # the reference's gather path never sets it
    pkg = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "pytorch-distributed-template_b200")
    spec = importlib.util.spec_from_file_location("pdt_test_entry", os.path.join(pkg, "test.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    import model.metric as M
    log = mod.evaluate(make_model(params), valid, [M.accuracy, M.top_k_acc], torch.device("cuda"))
    assert log["accuracy"] == R.accuracy(ref_out, ref_t) and log["top_k_acc"] == R.top_k_acc(ref_out, ref_t)
    assert abs(log["loss"] - float(R.nll_loss(ref_out, ref_t))) < 1e-5 * abs(log["loss"])
