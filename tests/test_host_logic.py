"""CPU tests of the host-side mirror of the reference's plug-in surface (no GPU, no kernels)."""
import argparse
import collections
import ctypes
import json
import os
import re
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "pytorch-distributed-template_b200")


def base_config(tmp_path):
    from utils import read_json
    cfg = read_json(os.path.join(PKG, "config", "config.json"))
    cfg["trainer"]["save_dir"] = str(tmp_path)
    cfg["trainer"]["tensorboard"] = False
    return cfg


def test_c_abi_exports_every_declared_symbol():
    """include/pdt.h <-> lib/libpdt.so <-> pdt.native.SIGNATURES agree; the library loads without a GPU."""
    from pdt import native
    header = open(os.path.join(ROOT, "include", "pdt.h")).read()
    declared = set(re.findall(r"\b(pdt_[a-z0-9_]+)\s*\(", header))
    assert declared == set(native.SIGNATURES), declared ^ set(native.SIGNATURES)
    lib = native.load()
    for name in declared:
        assert getattr(lib, name) is not None
    assert lib.pdt_version() >= 100
    assert lib.pdt_num_tiles(33) == 2
    assert lib.pdt_mnist_scratch_bytes(1024) == 32 * 23140 * 4
    assert lib.pdt_nll_scratch_bytes(10) >= 32
    # argument validation happens before any CUDA call
    rc = lib.pdt_conv1_pool_relu_fwd(None, None, None, None, None, 0, None)
    assert rc == -1 and b"pdt_conv1_pool_relu_fwd" in lib.pdt_last_error()


def test_product_path_refuses_cpu():
    from model.model import MnistModel
    from model.metric import accuracy
    from pdt.native import PdtError
    m = MnistModel()
    with pytest.raises(PdtError):
        m(torch.zeros(2, 1, 28, 28))
    with pytest.raises(PdtError):
        accuracy(torch.zeros(2, 10), torch.zeros(2, dtype=torch.int64))
    with pytest.raises(NotImplementedError):
        MnistModel(num_classes=7)


def test_model_surface_matches_reference_names():
    from model.model import MnistModel
    from oracle import ref_step as R
    torch.manual_seed(0)
    m = MnistModel()
    sd = m.state_dict()
    assert list(sd.keys()) == [k for k, _ in R.PARAM_SHAPES]
    ref = R.init_params(0)
    for k in sd:
        assert torch.equal(sd[k], ref[k])      # same seed -> same initial weights as the reference
    assert "Trainable parameters: 21840" in str(m)


def test_config_parser_overrides_and_init_obj(tmp_path):
    from parse_config import ConfigParser
    cfg_file = tmp_path / "cfg.json"
    json.dump(base_config(tmp_path), open(cfg_file, "w"))
    parser = argparse.ArgumentParser()
    parser.add_argument("-c", "--config", default=None, type=str)
    parser.add_argument("-r", "--resume", default=None, type=str)
    parser.add_argument("-s", "--save_dir", default=None, type=str)
    CustomArgs = collections.namedtuple("CustomArgs", "flags type target")
    options = [CustomArgs(["--lr", "--learning_rate"], type=float, target="optimizer;args;lr"),
               CustomArgs(["--bs", "--batch_size"], type=int, target="train_loader;args;batch_size")]
    argv, sys.argv = sys.argv, ["x", "-c", str(cfg_file), "--lr", "0.5", "--bs", "256"]
    try:
        args, config = ConfigParser.from_args(parser, options)
    finally:
        sys.argv = argv
    assert config["optimizer"]["args"]["lr"] == 0.5
    assert config["train_loader"]["args"]["batch_size"] == 256
    assert config.save_dir.is_dir() and (config.save_dir / "config.json").is_file()
    assert config.save_dir.parts[-3:-1] == (config["name"], "train")
    w = [torch.nn.Parameter(torch.zeros(3))]
    opt = config.init_obj("optimizer", torch.optim, w)
    assert isinstance(opt, torch.optim.Adam) and opt.defaults["amsgrad"] and opt.defaults["lr"] == 0.5
    sched = config.init_obj("lr_scheduler", torch.optim.lr_scheduler, opt)
    assert sched.step_size == 50
    with pytest.raises(AssertionError):
        config.init_obj("optimizer", torch.optim, w, lr=1.0)
    ftn = config.init_ftn("optimizer", torch.optim)
    assert isinstance(ftn(w), torch.optim.Adam)
    assert config.get_logger("x", 1).level == 20
    with pytest.raises(AssertionError):
        config.get_logger("x", 5)


def test_metric_tracker_and_dist_getters():
    from utils import MetricTracker, inf_loop
    from utils.dist import all_gather, get_rank, get_world_size, is_main_process, reduce_tensor, synchronize

    class W:
        def __init__(self):
            self.seen = []

        def add_scalar(self, k, v):
            self.seen.append((k, v))
    w = W()
    mt = MetricTracker("loss", "acc", writer=w)
    mt.update("loss", 2.0)
    mt.update("loss", 4.0, n=3)
    assert mt.avg("loss") == 3.5 and mt.result() == {"loss": 3.5, "acc": 0}
    assert w.seen == [("loss", 2.0), ("loss", 4.0)]
    mt.reset()
    assert mt.result() == {"loss": 0, "acc": 0}
    assert (get_rank(), get_world_size(), is_main_process()) == (0, 1, True)
    synchronize()
    assert all_gather({"a": 1}) == [{"a": 1}]
    t = torch.ones(2)
    assert reduce_tensor(t) is t
    it = inf_loop([1, 2])
    assert [next(it) for _ in range(5)] == [1, 2, 1, 2, 1]


class TinyNet(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.fc = torch.nn.Linear(8, 3)

    def forward(self, x):
        return torch.log_softmax(self.fc(x), dim=1)


class Loader(list):
    batch_size = 4
    sampler = None


def test_trainer_host_loop_monitor_checkpoint_resume(tmp_path):
    """The generic (device-agnostic) Trainer path with a plain torch module on CPU: epoch loop,
    min-monitor bookkeeping, checkpoint format and resume (reference base_trainer.py:60-163)."""
    import torch.nn.functional as F
    from parse_config import ConfigParser
    from trainer import Trainer
    cfg = base_config(tmp_path)
    cfg["trainer"].update(epochs=3, save_period=1, monitor="min loss", early_stop=2, verbosity=0)
    cfg["optimizer"] = {"type": "SGD", "args": {"lr": 0.1}}
    config = ConfigParser(cfg, run_id="a")
    torch.manual_seed(0)
    net = TinyNet()
    opt = config.init_obj("optimizer", torch.optim, net.parameters())
    sched = torch.optim.lr_scheduler.StepLR(opt, step_size=1, gamma=0.5)
    g = torch.Generator().manual_seed(1)
    loader = Loader((torch.randn(4, 8, generator=g), torch.randint(0, 3, (4,), generator=g)) for _ in range(5))

    def acc(o, t):
        return float((o.argmax(1) == t).float().mean())
    tr = Trainer(net, F.nll_loss, [acc], opt, config=config, device="cpu", data_loader=loader,
                 valid_data_loader=loader, lr_scheduler=sched)
    assert tr.log_step == 2 and tr.len_epoch == 5
    tr.train()
    files = sorted(p.name for p in config.save_dir.glob("*.pth"))
    assert files == ["checkpoint-epoch1.pth", "checkpoint-epoch2.pth", "checkpoint-epoch3.pth", "model_best.pth"]
    ck = torch.load(config.save_dir / "checkpoint-epoch3.pth", weights_only=True)
    assert ck["epoch"] == 3 and ck["arch"] == "TinyNet" and list(ck["state_dict"]) == ["fc.weight", "fc.bias"]
    assert ck["config"]["optimizer"]["type"] == "SGD" and "lr_scheduler" in ck
    assert abs(opt.param_groups[0]["lr"] - 0.1 * 0.5 ** 3) < 1e-12
    # resume
    config2 = ConfigParser(base_config(tmp_path) | {"optimizer": cfg["optimizer"]}, resume=config.save_dir / "checkpoint-epoch3.pth", run_id="b")
    net2 = TinyNet()
    opt2 = torch.optim.SGD(net2.parameters(), lr=0.1)
    tr2 = Trainer(net2, F.nll_loss, [acc], opt2, config=config2, device="cpu", data_loader=loader,
                  lr_scheduler=torch.optim.lr_scheduler.StepLR(opt2, step_size=1, gamma=0.5))
    assert tr2.start_epoch == 4
    assert torch.equal(net2.fc.weight, net.fc.weight)
    assert tr2.lr_scheduler.last_epoch == 3


def test_early_stop_counts(tmp_path):
    import torch.nn.functional as F
    from parse_config import ConfigParser
    from trainer import Trainer
    cfg = base_config(tmp_path)
    cfg["trainer"].update(epochs=50, save_period=100, monitor="max val_never_there", early_stop=1, verbosity=0)
    config = ConfigParser(cfg, run_id="c")
    net = TinyNet()
    opt = torch.optim.SGD(net.parameters(), lr=0.0)
    loader = Loader([(torch.zeros(4, 8), torch.zeros(4, dtype=torch.int64))])
    tr = Trainer(net, F.nll_loss, [], opt, config=config, device="cpu", data_loader=loader)
    tr.train()                                  # unknown metric: monitoring switches itself off
    assert tr.mnt_mode == "off"
    cfg["trainer"]["monitor"] = "min loss"
    cfg["trainer"]["epochs"] = 10
    tr = Trainer(net, F.nll_loss, [], opt, config=ConfigParser(cfg, run_id="d"), device="cpu", data_loader=loader)
    seen = []
    orig = tr._train_epoch
    tr._train_epoch = lambda e: (seen.append(e), {"loss": float(e)})[1]   # gets worse every epoch
    tr.train()
    assert seen == [1, 2, 3]      # best at 1; 2 and 3 do not improve; count 2 > early_stop 1 -> stop


def test_synthetic_loader_rank_slices():
    from data_loader.data_loaders import SyntheticDataLoader
    a = SyntheticDataLoader(4, n_batches=2, seed=5, pin=False)
    assert len(a) == 2 and a.batch_size == 4 and a.n_samples == 8
    x, t = next(iter(a))
    assert tuple(x.shape) == (4, 1, 28, 28) and t.dtype == torch.int64 and int(t.max()) <= 9
    b = SyntheticDataLoader(4, n_batches=2, seed=5, pin=False)
    assert torch.equal(b.batches[1][0], a.batches[1][0])


def test_bench_reference_arm_runs_on_cpu():
    """`bench.py --impl reference` is the reference's CPU step timed by the oracle port: one JSON line with the
    contract's keys, no GPU involved (it also runs on the box without touching /root/reference)."""
    import json
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "3",
                          "--warmup", "3", "--batch", "64"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines
    line = json.loads(lines[0])
    assert line["impl"] == "reference" and line["unit"] == "samples/s" and line["value"] > 0
    assert line["steps"] == 3 and line["n_gpus"] == 1 and line["higher_is_better"] is True
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"]["value"] == line["value"] and line["e2e"]["h2d_bytes_per_step"] == 0


def test_shipped_configs_keep_the_reference_keys():
    """config/*.json are the user's interface to the template: same sections, types and default values as the
    reference's config/config.json, plus optional trainer keys of this build."""
    from utils import read_json
    pkg = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "pytorch-distributed-template_b200")
    for name in ("config.json", "debug.json"):
        cfg = read_json(os.path.join(pkg, "config", name))
        assert cfg["arch"] == {"type": "MnistModel", "args": {}}
        assert cfg["loss"] == "nll_loss" and cfg["metrics"] == ["accuracy", "top_k_acc"]
        assert cfg["optimizer"]["type"] == "Adam"
        assert dict(cfg["optimizer"]["args"]) == {"lr": 0.001, "weight_decay": 0, "amsgrad": True}
        assert cfg["lr_scheduler"]["type"] == "StepLR" and dict(cfg["lr_scheduler"]["args"]) == {"step_size": 50, "gamma": 0.1}
        for sec, training in (("train_loader", True), ("valid_loader", False), ("test_loader", False)):
            a = cfg[sec]["args"]
            assert cfg[sec]["type"] == "MnistDataLoader" and a["batch_size"] == 128 and a["training"] is training
            assert a["data_dir"] == "data/" and a["shuffle"] is training
        t = cfg["trainer"]
        assert t["monitor"] == "min val_loss" and t["early_stop"] == 10 and t["save_period"] == 1
        assert t["save_dir"] == "checkpoints/" and t["verbosity"] == 2 and t["tensorboard"] is True


def test_near_gpu_is_a_noop_without_topology():
    from utils import near_gpu
    before = os.sched_getaffinity(0)
    with near_gpu(0) as ctx:
        assert ctx.cpus is None or ctx.cpus <= before
    assert os.sched_getaffinity(0) == before


def test_fused_optimizers_accept_only_implementation_selecting_torch_kwargs():
    """torch.optim keywords that merely pick one of torch's implementations are accepted and ignored; anything that
    would change the arithmetic is refused (a reference config may carry either kind, config.json optimizer.args)."""
    from pdt.optim import _check_torch_kwargs
    _check_torch_kwargs("Adam", {"foreach": None, "fused": True, "capturable": False, "differentiable": False, "maximize": False})
    with pytest.raises(ValueError):
        _check_torch_kwargs("SGD", {"maximize": True})
    with pytest.raises(ValueError):
        _check_torch_kwargs("Adam", {"differentiable": True})
    with pytest.raises(TypeError):
        _check_torch_kwargs("Adam", {"nonsense": 1})


def test_bench_reference_arm_runs_as_gloo_processes_for_more_than_one_rank():
    """`bench.py --impl reference --gpus 2` launched under torchrun with two ranks, as the driver launches it: rank 0
    alone runs the reference's CPU job as two processes over gloo (the way train.py launches it) and prints one JSON
    line whose config equals the repo arm's for the same arguments; rank 1 exits without work.  (The spawned job must
    not pick up torchrun's "use the agent's store" environment: it did once, and hung.)"""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29578", os.path.join(root, "bench.py"),
                          "--impl", "reference", "--gpus", "2", "--steps", "2", "--warmup", "1", "--batch", "32"],
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["n_gpus"] == 2 and line["value"] > 0
    assert "2 processes over gloo" in line["cpu_baseline"]["sample"]
    sys.path.insert(0, root)
    import bench
    args = type("A", (), {"batch": 32, "optimizer": "adam"})()
    assert line["config"] == bench.workload_config(args, 2)
