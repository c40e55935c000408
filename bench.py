"""Benchmark of the data-parallel training step (BASELINE.json metric: train samples/sec,
MnistModel, per-GPU batch 1024, template-default Adam+amsgrad).

    python bench.py [--gpus N --steps K --warmup W]            # this framework
    python bench.py --impl reference [...]                      # reference CPU step (oracle port)
    torchrun --nproc-per-node N ... bench.py --gpus N ...       # N > 1, one rank per GPU

One "step" = trainer/trainer.py:48-58 (zero_grad, forward, nll_loss, reduce_loss, backward with
the gradient all-reduce, optimizer.step) on one synthetic batch.  Two measurements:
  value : inputs already resident in HBM (ring of device batches larger than L2), the step
          replayed as a CUDA graph, CUDA events on the launching stream, max over ranks.
  e2e   : the user-facing path, Trainer._train_epoch over a host (pinned) loader: every step
          copies its batch H2D and reads its loss back D2H inside the timed region.
"""
import argparse
import json
import os
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "pytorch-distributed-template_b200")
for _p in (PKG, ROOT):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402

MIN_REGION_MS = 250.0
E2E_MIN_STEPS = 2500          # end-to-end epochs are at least this long (~0.25 s at batch 1024)
FLOP_PER_SAMPLE = 2_595_000        # fwd 961,000 + bwd 1,634,000
FP32_NOMINAL_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12   # 74.4: SMs x lanes x 2 x max clock


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=500)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=1024, help="per-GPU batch")
    ap.add_argument("--optimizer", default="adam", choices=["adam", "sgd"])
    ap.add_argument("--ring", type=int, default=0, help="device-resident batches (0 = enough to exceed L2)")
    ap.add_argument("--no-extras", action="store_true", help="skip cpu_baseline / torch-eager / per-kernel legs")
    ap.add_argument("--cpu-seconds", type=float, default=10.0)
    ap.add_argument("--input", default="uint8", choices=["uint8", "fp32"],
                    help="host format of the end-to-end batches: raw uint8 pixels normalised inside conv1's load "
                         "(what MnistDataLoader(raw_bytes=True) delivers), or the loader's fp32")
    ap.add_argument("--no-parity", action="store_true", help="skip the 3-step parity probe against the oracle")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------
# clocks: sampled DURING the timed regions through NVML
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi polled every 20 ms in a SEPARATE process (a sampling thread in this process would
    steal the GIL from the host loop being timed).  Samples are kept only if their timestamp falls
    inside one of the windows entered with `with sampler:` (the timed regions, plus one sustained
    replay of the same step long enough to be sampled at all)."""
    FIELDS = ("timestamp,clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        import subprocess
        self.windows, self.err, self.proc = [], None, None
        self.path = os.path.join(tempfile.mkdtemp(prefix="pdt_clk_"), "clocks.csv")
        try:
            self.out = open(self.path, "w")
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.FIELDS,
                                          "--format=csv,noheader,nounits", "-lms", os.environ.get("PDT_SMI_MS", "20")],
                                         stdout=self.out, stderr=subprocess.DEVNULL)
        except Exception as e:
            self.err = repr(e)

    def __enter__(self):
        self._t0 = time.time()
        return self

    def __exit__(self, *a):
        self.windows.append((self._t0, time.time()))

    def summary(self):
        import datetime
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable: %s" % self.err]}
        time.sleep(0.05)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.out.close()
        clocks, reasons, mx, total = [], set(), None, 0
        for line in open(self.path):
            parts = [q.strip() for q in line.split(",")]
            if len(parts) < 7:
                continue
            total += 1
            try:
                ts = datetime.datetime.strptime(parts[0], "%Y/%m/%d %H:%M:%S.%f").timestamp()
                mhz, mx = float(parts[1]), float(parts[2])
            except ValueError:
                continue
            if not any(a - 0.02 <= ts <= b + 0.02 for a, b in self.windows):
                continue
            clocks.append(mhz)
            for name, val in zip(self.NAMES, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        clocks.sort()
        return {"sm_mhz": clocks[len(clocks) // 2] if clocks else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(clocks), "samples_total": total, "how": "nvidia-smi -lms 20 in a child process, "
                "samples inside the timed regions"}


# ---------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the reference's CPU step (oracle port of trainer.py:48-58)
# ---------------------------------------------------------------------------------------------
def host_threads():
    """All host threads this process may use (torchrun exports OMP_NUM_THREADS=1: undo that here)."""
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    torch.set_num_threads(n)
    return n


def cpu_reference_step_fn(batch, optimizer, world=1):
    """Returns step(i) running the oracle's restatement of the reference step on host cores: `world`
    ranks of `batch` samples each, gradients averaged as DDP does, one optimizer step.
    This is the one place outside tests/ that executes oracle/ (as the thing measured for the
    baseline arms only, never on the product path)."""
    from oracle import ref_step as R
    params = R.init_params(0)
    opt = R.Adam(lr=1e-3, amsgrad=True) if optimizer == "adam" else R.SGD(lr=0.01)
    ring = []
    for s in range(4):
        x, t, k2, k1 = R.synthetic_batch(batch * world, 1234 + s)
        ring.append([(x[r * batch:(r + 1) * batch], t[r * batch:(r + 1) * batch], k2[r * batch:(r + 1) * batch],
                      k1[r * batch:(r + 1) * batch]) for r in range(world)])
    state = {"p": params}

    def step(i):
        p, losses, _ = R.train_steps(state["p"], [ring[i % 4]], opt, world_size=world)
        state["p"] = p
        return float(losses[0])     # the reference's loss.item() (trainer.py:62)
    return step


def time_cpu(batch, optimizer, steps=None, warmup=2, seconds=10.0, world=1):
    step = cpu_reference_step_fn(batch, optimizer, world)
    for i in range(warmup):
        step(i)
    t0 = time.perf_counter()
    n = 0
    while True:
        step(n)
        n += 1
        el = time.perf_counter() - t0
        if (steps is not None and n >= steps) or (steps is None and el >= seconds):
            break
    return n, el


def _cpu_ddp_worker(rank, world, port, batch, optimizer, steps, warmup, threads, out_path):
    """One rank of the reference's CPU job (train.py:24-52 with backend gloo): the oracle's restatement of the step
    on this rank's slice, gradients averaged with one gloo all-reduce as DDP does, the blocking reduce_loss of
    base_trainer.py:165-174, optimizer step, loss.item()."""
    import torch.distributed as dist
    from oracle import ref_step as R
    torch.set_num_threads(threads)
    # under torchrun the parent's environment says "use the elastic agent's store": this job brings its own
    for key in ("TORCHELASTIC_USE_AGENT_STORE", "TORCHELASTIC_RUN_ID", "TORCHELASTIC_RESTART_COUNT", "TORCHELASTIC_MAX_RESTARTS"):
        os.environ.pop(key, None)
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    params = R.init_params(0)
    opt = R.Adam(lr=1e-3, amsgrad=True) if optimizer == "adam" else R.SGD(lr=0.01)
    ring = []
    for s in range(4):
        x, t, k2, k1 = R.synthetic_batch(batch * world, 1234 + s)
        sl = slice(rank * batch, (rank + 1) * batch)
        ring.append((x[sl], t[sl], k2[sl], k1[sl]))
    keys = [k for k, _ in R.PARAM_SHAPES]

    def step(i):
        nonlocal params
        loss, _, g = R.loss_and_grads(params, *ring[i % 4])
        red = loss.clone()
        dist.reduce(red, dst=0)
        flat = torch.cat([g[k].reshape(-1) for k in keys]) / world
        dist.all_reduce(flat)
        off = 0
        for k in keys:
            cnt = g[k].numel()
            g[k] = flat[off:off + cnt].view_as(g[k])
            off += cnt
        params = opt.step(dict(params), g)
        return float(red) / world
    for i in range(warmup):
        step(i)
    dist.barrier()
    t0 = time.perf_counter()
    for i in range(steps):
        step(i)
    dist.barrier()
    el = time.perf_counter() - t0
    if rank == 0:
        with open(out_path, "w") as f:
            json.dump({"steps": steps, "seconds": el}, f)
    dist.destroy_process_group()


def time_cpu_ddp(batch, optimizer, world, steps, warmup, threads_total):
    """The W-rank CPU job as W processes over gloo on this host, the cores divided between them."""
    import torch.multiprocessing as mp
    import socket
    sock = socket.socket()
    sock.bind(("127.0.0.1", 0))
    port = sock.getsockname()[1]
    sock.close()
    per = max(1, threads_total // world)
    out_path = os.path.join(tempfile.mkdtemp(prefix="pdt_cpu_ddp_"), "result.json")
    mp.spawn(_cpu_ddp_worker, args=(world, port, batch, optimizer, steps, warmup, per, out_path), nprocs=world, join=True)
    r = json.load(open(out_path))
    return r["steps"], r["seconds"], per


def run_reference(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    threads = host_threads()
    world = args.gpus
    budget = 100.0
    # bounded sample: K steps of the global batch, capped so the whole run stays within a few minutes
    probe_n, probe_t = time_cpu(args.batch, args.optimizer, steps=1, warmup=1, world=world)
    per = probe_t / probe_n
    k = max(1, min(args.steps, int(budget / per)))
    w = max(1, min(args.warmup, int(15.0 / per)))
    if world == 1:
        n, el = time_cpu(args.batch, args.optimizer, steps=k, warmup=w, world=1)
        how = "one process, %d threads" % threads
    else:
        n, el, per_rank = time_cpu_ddp(args.batch, args.optimizer, world, k, w, threads)
        how = "%d processes over gloo (as train.py launches them), %d threads each" % (world, per_rank)
    value = n * args.batch * world / el
    sample = "%d steps of global batch %d (%d rank(s) x %d, warmup %d) of the same synthetic workload, torch %s CPU, %s" % (
        n, args.batch * world, world, args.batch, w, torch.__version__, how)
    line = {
        "impl": "reference", "metric": "train samples/sec", "value": value, "unit": "samples/s",
        "n_gpus": args.gpus, "steps": n, "warmup": w, "ms_per_step": 1e3 * el / n, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, world),
        "cpu_baseline": {"value": value, "unit": "samples/s", "cores": threads, "kind": "port", "sample": sample,
                         "host_cpus": os.cpu_count()},
        "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def workload_config(args, world):
    return {"workload": "MnistModel train step (conv-pool-relu x2, fc x2, log_softmax+nll_loss, %s), "
                        "per-GPU batch %d, 1x28x28 fp32, 10 classes" % (
                            "Adam lr=1e-3 amsgrad" if args.optimizer == "adam" else "SGD lr=0.01", args.batch),
            "global_batch": args.batch * world, "per_gpu_batch": args.batch,
            "parallelism": "dp%d" % world, "optimizer": args.optimizer}


# ---------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------
def build_trainer(args, device, world, n_batches, cuda_graph=True, input_dtype="float32"):
    import logging
    from parse_config import ConfigParser
    from utils import read_json
    import data_loader.data_loaders as module_data
    import model.loss as module_loss
    import model.metric as module_metric
    import model.model as module_arch
    from pdt import optim as pdt_optim
    from pdt.parallel import FlatDataParallel
    from trainer import Trainer

    cfg = read_json(os.path.join(PKG, "config", "config.json"))
    cfg["trainer"].update(save_dir=tempfile.mkdtemp(prefix="pdt_bench_"), tensorboard=False, verbosity=0,
                          cuda_graph=cuda_graph, epochs=1, save_period=10 ** 9)
    if args.optimizer == "sgd":
        cfg["optimizer"] = {"type": "SGD", "args": {"lr": 0.01}}
    config = ConfigParser(cfg, run_id="bench_r%d" % int(os.environ.get("RANK", 0)))
    logging.disable(logging.CRITICAL)
    torch.manual_seed(0)
    model = config.init_obj("arch", module_arch).to(device)
    criterion = getattr(module_loss, config["loss"])
    metrics = [getattr(module_metric, m) for m in config["metrics"]]
    optimizer = config.init_obj("optimizer", pdt_optim, [p for p in model.parameters() if p.requires_grad])
    sched = config.init_obj("lr_scheduler", torch.optim.lr_scheduler, optimizer)
    if world > 1:
        model = FlatDataParallel(model, device_ids=[device.index], output_device=device.index)
    loader = module_data.SyntheticDataLoader(args.batch, n_batches=n_batches, seed=1234, dtype=input_dtype)
    trainer = Trainer(model, criterion, metrics, optimizer, config=config, device=device, data_loader=loader,
                      valid_data_loader=None, lr_scheduler=sched)
    return trainer, loader


class CountLaunches:
    """Counts the kernel-launching C-ABI calls made while active (our kernels only)."""
    NON_KERNEL = {"pdt_comm_unique_id", "pdt_comm_init", "pdt_comm_destroy", "pdt_allreduce_flat"}

    def __enter__(self):
        from pdt import native
        self.native, self.orig, self.n = native, native.call, 0

        def counting(name, *a):
            if name not in self.NON_KERNEL:
                self.n += 1
            return self.orig(name, *a)
        native.call = counting
        import pdt.ops, pdt.optim, pdt.parallel  # noqa: E401
        self.mods = [pdt.ops, pdt.optim, pdt.parallel]
        for m in self.mods:
            m.call = counting
        return self

    def __exit__(self, *a):
        self.native.call = self.orig
        for m in self.mods:
            m.call = self.orig


def barrier_sync(world):
    if world > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize()


def max_over_ranks(ms, world, device):
    if world == 1:
        return ms
    t = torch.tensor([ms], dtype=torch.float64, device=device)
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    return float(t.item())


def measure_value(trainer, args, device, world, sampler):
    """Inputs resident in HBM; one CUDA graph per ring slot; events on the launching stream."""
    from data_loader.data_loaders import SyntheticDataLoader
    per_batch = args.batch * 784 * 4 + args.batch * 8
    ring = args.ring or max(8, -(-200 * 2 ** 20 // per_batch))     # > 126 MB L2: no slot is still cached
    ring = min(ring, 640)
    src = SyntheticDataLoader(args.batch, n_batches=8, seed=1234, pin=False)
    dev_batches = []
    for s in range(ring):
        x, t = src.batches[s % 8]
        dev_batches.append((x.to(device).clone(), t.to(device).clone()))
    trainer.model.train()
    # one eager step creates the lazy workspaces; then the steps over ALL ring slots are captured, in order, into
    # ONE CUDA graph: a replay runs `ring` consecutive steps with no host involvement (one graph per slot made the
    # host re-upload a different graph every step: measured 88 us per step at batch 128 where the loop over two
    # graphs of the end-to-end path took 73)
    trainer._step(*dev_batches[0])
    torch.cuda.synchronize()
    note("  capturing %d steps into one graph" % ring)
    losses, launches = [], 0
    graph = torch.cuda.CUDAGraph()
    with CountLaunches() as cl:
        with torch.cuda.graph(graph):
            for x, t in dev_batches:
                losses.append(trainer._step(x, t))
    launches = cl.n // ring
    stream = torch.cuda.current_stream()
    for i in range(max(1, -(-args.warmup // ring))):
        graph.replay()

    def region(replays):
        barrier_sync(world)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for i in range(replays):
            graph.replay()
        e1.record(stream)
        barrier_sync(world)
        return max_over_ranks(e0.elapsed_time(e1), world, device)
    # --steps K is the unit; a region runs whole passes over the ring, at least K steps and at least MIN_REGION_MS
    # (the driver's --steps 20 alone would be a 2 ms sample), the same count on every rank; FIVE regions are timed
    # and the MEDIAN is reported per step
    base_replays = max(1, -(-args.steps // ring))
    note("  timing")
    probe_ms = region(base_replays)
    replays = base_replays * max(1, int(-(-MIN_REGION_MS // max(probe_ms, 1e-3))))
    count = replays * ring
    with sampler:
        regions = [region(replays) for _ in range(5)]
    ms = sorted(regions)[2] / count * args.steps         # per K steps, as the contract's ms_per_step * K
    timing = {"stat": "median of 5 timed regions", "steps_per_region": count, "region_ms": [round(r, 4) for r in regions],
              "min_region_ms": MIN_REGION_MS, "ring": ring}
    last_loss = float(losses[-1].item())
    return ms, launches, ring, last_loss, timing


def measure_e2e(trainer, loader, args, device, world, sampler):
    """Trainer._train_epoch over host batches: per step one H2D copy of the batch from pinned
    memory and one D2H read of the step's loss, both inside the timed region."""
    # warm-up epoch (first step eager, second captures the graph)
    base = list(loader.batches)
    loader.n_batches = max(args.warmup, 3)
    loader.batches = [base[i % len(base)] for i in range(loader.n_batches)]
    trainer.len_epoch = loader.n_batches
    trainer._train_epoch(1)
    note("  warm epoch done")
    # a fresh process finds the GPU at idle clocks and W steps last ~2 ms: one more untimed epoch of a FIXED 2,000
    # steps (the same count on every rank: each step carries a collective) so the timed epochs see the clocks
    # the step runs at when sustained
    loader.batches = [base[i % len(base)] for i in range(2000)]
    loader.n_batches = trainer.len_epoch = 2000
    trainer._train_epoch(1)
    torch.cuda.synchronize()
    # an epoch of --steps K steps, or of E2E_MIN_STEPS if K is smaller (the driver's --steps 20 would time 2 ms epochs that
    # are mostly the epoch's own start-up and final flush); the result is reported per K steps
    k_run = max(args.steps, E2E_MIN_STEPS)
    loader.batches = [base[i % len(base)] for i in range(k_run)]
    loader.n_batches = k_run
    trainer.len_epoch = k_run
    # The loop is driven by the host, so one descheduling of this process lands in the number.  Three epochs
    # of exactly K steps each are timed; the MEDIAN epoch is reported and all three are listed.
    runs, log = [], None
    stream = torch.cuda.current_stream()
    for rep in range(3):
        barrier_sync(world)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with sampler:
            t0 = time.perf_counter()
            e0.record(stream)
            log = trainer._train_epoch(2 + rep)
            e1.record(stream)
            barrier_sync(world)
            wall_ms = 1e3 * (time.perf_counter() - t0)
        ms = max(e0.elapsed_time(e1), wall_ms)       # the host side is part of end-to-end
        runs.append(max_over_ranks(ms, world, device) * args.steps / k_run)
    return sorted(runs)[1], log, runs


def measure_h2d_copy(args, device, bytes_per_pixel=4, reps=50):
    """Time of the step's input copy alone (pinned host -> device, one cudaMemcpyAsync): when it exceeds the
    device step time, the end-to-end loop is transfer-bound on this box (PCIe rate varies 4x between boxes)."""
    from pdt import native
    nb = args.batch * 784 * bytes_per_pixel
    host = torch.empty(nb, dtype=torch.uint8).pin_memory()
    dev = torch.empty(nb, dtype=torch.uint8, device=device)
    s = torch.cuda.current_stream()
    fn = lambda: native.call("pdt_memcpy_async", dev.data_ptr(), host.data_ptr(), nb, native.stream())
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(s)
    for _ in range(reps):
        fn()
    e1.record(s)
    torch.cuda.synchronize()
    return 1e3 * e0.elapsed_time(e1) / reps


def kernel_breakdown(args, device, reps=40):
    from pdt import native, ops
    from pdt.native import call, ptr, stream
    n = args.batch
    T = (n + 31) // 32
    g = torch.Generator().manual_seed(1)
    x = torch.randn(n, 1, 28, 28, generator=g).to(device)
    t = torch.randint(0, 10, (n,), generator=g).to(device)
    from model.model import MnistModel
    torch.manual_seed(0)
    m = MnistModel().to(device)
    flat = m.flat.data
    k2 = (torch.rand(n, 20, generator=g) < 0.5).to(torch.uint8).to(device)
    k1 = (torch.rand(n, 50, generator=g) < 0.5).to(torch.uint8).to(device)
    logp, saved = ops.mnist_forward_raw(x, flat, k2, k1, True)
    p1, idx1, p2, idx2, hd = saved[:5]
    st = ops._DeviceState.get(device)
    dq2, dq1, scratch = st.bwd_workspace(n)
    stats = torch.empty(8, dtype=torch.int32, device=device)
    dlogp = torch.empty(n, 10, device=device)
    grad = torch.zeros(21840 + 4, device=device)
    adam_m, adam_v, adam_vm = (torch.zeros(21840, device=device) for _ in range(3))
    import struct
    hyper = torch.zeros(64, dtype=torch.uint8, device=device)
    hyper[:44].copy_(torch.frombuffer(bytearray(struct.pack("dddddi", 1e-3, 0.9, 0.999, 1e-8, 0.0, 1)), dtype=torch.uint8))
    rng = st.rng_state()
    base = flat.data_ptr()
    s = stream()   # the lambdas below read `s` when called, so it can follow the capture stream
    fma = lambda per_sample: per_sample * n
    kernels = [
        ("dropout_masks", lambda: call("pdt_dropout_masks", ptr(k2), ptr(k1), n, 0.5, ptr(rng), s), 0, 70 * n),
        ("conv1_pool_relu_fwd", lambda: call("pdt_conv1_pool_relu_fwd", ptr(x), base, base + 1000, ptr(p1), ptr(idx1), n, s),
         fma(288_000), n * (3136 + 5760 + 360)),
        ("conv2_drop_pool_relu_fwd", lambda: call("pdt_conv2_drop_pool_relu_fwd", ptr(p1), base + 1040, base + 21040, ptr(k2), 2.0,
                                                 ptr(p2), ptr(idx2), n, s), fma(640_000), n * (5760 + 1280 + 320)),
        ("conv2_drop_pool_relu_fwd_tc", lambda: call("pdt_conv2_drop_pool_relu_fwd_tc", ptr(p1), base + 1040, base + 21040, ptr(k2),
                                                    2.0, ptr(p2), ptr(idx2), n, s), fma(640_000), n * (5760 + 1280 + 320)),
        ("head_fwd", lambda: call("pdt_head_fwd", ptr(p2), base + 4 * 5280, base + 4 * 21280, base + 4 * 21330, base + 4 * 21830,
                                  ptr(k1), 2.0, ptr(hd), ptr(logp), n, s), fma(33_000), n * (1280 + 200 + 40)),
        ("nll_metrics_fwd", lambda: call("pdt_nll_metrics_fwd", ptr(logp), ptr(t), n, 10, 3, -100, ptr(stats),
                                         ptr(st.nll_scratch(n)), s), 0, n * 48),
        ("nll_bwd", lambda: call("pdt_nll_bwd", ptr(t), ptr(stats), None, ptr(dlogp), n, 10, -100, s), 0, n * 48),
        ("head_bwd", lambda: call("pdt_head_bwd", ptr(dlogp), ptr(logp), ptr(hd), ptr(p2), base + 4 * 5280, base + 4 * 21330,
                                  2.0, 2.0, ptr(dq2), ptr(scratch), n, s), fma(66_000), n * (80 + 200 + 1280 + 1280) + T * 16560 * 4),
        ("conv2_bwd_data", lambda: call("pdt_conv2_bwd_data", ptr(dq2), ptr(idx2), base + 1040, ptr(p1), ptr(dq1), n, s),
         fma(640_000), n * (1280 + 160 + 5760 + 5760)),
        ("conv2_bwd_weight", lambda: call("pdt_conv2_bwd_weight", ptr(dq2), ptr(idx2), ptr(p1), ptr(scratch), n, s),
         fma(640_000), n * (1280 + 160 + 5760) + T * 5020 * 4),
        ("conv1_bwd_weight", lambda: call("pdt_conv1_bwd_weight", ptr(dq1), ptr(idx1), ptr(x), ptr(scratch), n, s),
         fma(288_000), n * (5760 + 360 + 3136) + T * 6 * 260 * 4),
        ("grad_reduce", lambda: call("pdt_grad_reduce", ptr(scratch), n, 0.0, 1.0, ptr(stats), ptr(grad), None, s),
         0, T * 23140 * 4 + 21840 * 4),
        ("adam_step_flat", lambda: call("pdt_adam_step_flat", base, ptr(grad), ptr(adam_m), ptr(adam_v), ptr(adam_vm),
                                        21840, 1.0, ptr(hyper), s), 0, 21840 * 36),
    ]
    # the launches as the trainer's step issues them (fused variants of the entries above)
    hscr = st.head_nll_scratch(n)
    kernels += [
        ("conv1_fwd_masks", lambda: call("pdt_conv1_fwd_masks", ptr(x), base, base + 1000, ptr(p1), ptr(idx1), ptr(k2), ptr(k1),
                                         0.5, ptr(rng), n, s), fma(288_000), n * (3136 + 5760 + 360 + 70)),
        ("head_fwd_nll", lambda: call("pdt_head_fwd_nll", ptr(p2), base + 4 * 5280, base + 4 * 21280, base + 4 * 21330,
                                      base + 4 * 21830, ptr(k1), 2.0, ptr(hd), ptr(logp), n, ptr(t), 3, -100, ptr(stats),
                                      ptr(hscr), None, None, 0, s), fma(33_000), n * (1280 + 200 + 40 + 8)),
        ("head_bwd_nll", lambda: call("pdt_head_bwd_nll", ptr(t), ptr(stats), None, -100, ptr(logp), ptr(hd), ptr(p2),
                                      base + 4 * 5280, base + 4 * 21330, 2.0, 2.0, ptr(dq2), ptr(scratch), n, s),
         fma(66_000), n * (8 + 40 + 200 + 1280 + 1280) + T * 16560 * 4),
        ("adam_step_reduce", lambda: call("pdt_adam_step_reduce", ptr(scratch), n, 0.0, 1.0, ptr(stats), ptr(grad), base,
                                          ptr(adam_m), ptr(adam_v), ptr(adam_vm), ptr(hyper), None, s),
         0, T * 23140 * 4 + 21840 * 40),
    ]
    kernels += [
        ("conv2_head_train", lambda: call("pdt_conv2_head_train", ptr(p1), base + 1040, base + 21040, ptr(k2), 2.0, ptr(p2),
                                          ptr(idx2), base + 4 * 5280, base + 4 * 21280, base + 4 * 21330, base + 4 * 21830,
                                          ptr(k1), 2.0, ptr(hd), ptr(logp), ptr(t), 3, -100, ptr(stats), ptr(hscr), None,
                                          None, 0, ptr(dq2), ptr(scratch), n, s),
         fma(640_000 + 33_000 + 66_000), n * (5760 + 1280 + 320 + 200 + 40 + 8 + 1280) + T * 16560 * 4),
        ("conv2_head_train_tc", lambda: call("pdt_conv2_head_train_tc", ptr(p1), base + 1040, base + 21040, ptr(k2), 2.0, ptr(p2),
                                             ptr(idx2), base + 4 * 5280, base + 4 * 21280, base + 4 * 21330, base + 4 * 21830,
                                             ptr(k1), 2.0, ptr(hd), ptr(logp), ptr(t), 3, -100, ptr(stats), ptr(hscr), None,
                                             None, 0, ptr(dq2), ptr(scratch), n, s),
         fma(640_000 + 33_000 + 66_000), n * (5760 + 1280 + 320 + 200 + 40 + 8 + 1280) + T * 16560 * 4),
        ("conv2_bwd_data_conv1_bwd_weight", lambda: call("pdt_conv2_bwd_data_conv1_bwd_weight", ptr(dq2), ptr(idx2), base + 1040,
                                                         ptr(p1), ptr(x), ptr(idx1), ptr(scratch), n, s),
         fma(640_000 + 288_000), n * (1280 + 160 + 5760 + 3136 + 360) + T * 6 * 260 * 4),
    ]
    if n <= ops.fused_rows(device):     # the launches of the Trainer's step at this batch size (pdt/ops.py)
        in_step = {"conv1_fwd_masks", "conv2_head_train_tc" if ops.TRAIN_TC else "conv2_head_train",
                   "conv2_bwd_data_conv1_bwd_weight", "conv2_bwd_weight", "adam_step_reduce"}
    else:
        in_step = {"conv1_fwd_masks", "conv2_drop_pool_relu_fwd_tc" if ops.CONV2_TC else "conv2_drop_pool_relu_fwd",
                   "head_fwd_nll", "head_bwd_nll", "conv2_bwd_data",
                   "conv2_bwd_weight", "conv1_bwd_weight", "adam_step_reduce"}
    if args.optimizer == "sgd":      # no fused reduce+update for SGD: the reduction and the update are two launches
        sgd_hyper = torch.zeros(64, dtype=torch.uint8, device=device)
        sgd_hyper[:36].copy_(torch.frombuffer(bytearray(struct.pack("ddddi", 0.01, 0.0, 0.0, 0.0, 0)), dtype=torch.uint8))
        kernels.append(("sgd_step_flat", lambda: call("pdt_sgd_step_flat", base, ptr(grad), None, 21840, 1.0, ptr(sgd_hyper), s),
                        0, 21840 * 12))
        in_step = (in_step - {"adam_step_reduce"}) | {"grad_reduce", "sgd_step_flat"}
    out = []
    for name, fn, flops, bytes_ in kernels:
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        # `reps` back-to-back launches recorded in one CUDA graph: no host launch cost in the number
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            s = stream()
            for _ in range(reps):
                fn()
        s = stream()
        g.replay()
        torch.cuda.synchronize()
        cur = torch.cuda.current_stream()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(cur)
        g.replay()
        e1.record(cur)
        torch.cuda.synchronize()
        us = 1e3 * e0.elapsed_time(e1) / reps
        out.append({"kernel": name, "us": round(us, 3), "flop": flops, "bytes": bytes_, "in_step": name in in_step})
    return out


NCU_NAMES = {"conv2_drop_pool_relu_fwd_tc": "conv2_fwd_tc_kernel", "conv1_pool_relu_fwd": "conv1_fwd_kernel", "conv2_drop_pool_relu_fwd": "conv2_fwd_kernel",
             "conv2_bwd_weight": "conv2_bwd_weight_kernel", "conv2_bwd_data": "conv2_bwd_data_kernel",
             "conv1_bwd_weight": "conv1_bwd_weight_kernel", "head_fwd": "head_fwd_kernel", "head_bwd": "head_bwd_kernel",
             "conv2_head_train": "conv2_head_train_kernel", "conv2_head_train_tc": "conv2_head_train_kernel_tc",
             "conv2_bwd_data_conv1_bwd_weight": "conv2_bwd_data_wgrad1_kernel", "conv1_fwd_masks": "conv1_fwd_masks_kernel"}


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel`, from the committed ncu --set full
    captures at N=1024 (profiles/r02_dram_traffic.json: round 1's captures of the unchanged kernels plus round 2's of
    the tensor-core training kernel); None if the captures lack the kernel."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "r02_dram_traffic.json")))["kernels"]
        return d[NCU_NAMES.get(kernel, kernel)]["traffic_bytes"]
    except Exception:
        return None


def measure_fma_peak(device, mode=0):
    """FP32 CUDA-core FMA peak of this GPU (MEASURED_PEAKS.json has no such entry).
    mode 0: scalar FFMA chains; 1: packed FFMA2 (fma.rn.f32x2); 2: FFMA2 with shared loads mixed in."""
    from pdt.native import call, stream
    out = torch.zeros(4, device=device)
    blocks, threads, iters = 148 * 8, 256, 4096
    call("pdt_fma_peak", out.data_ptr(), blocks, threads, 64, mode, stream())
    torch.cuda.synchronize()
    best = 0.0
    cur = torch.cuda.current_stream()
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(cur)
        call("pdt_fma_peak", out.data_ptr(), blocks, threads, iters, mode, stream())
        e1.record(cur)
        torch.cuda.synchronize()
        flops = blocks * threads * iters * 64 * 2
        best = max(best, flops / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    return best


def measure_allreduce(dp, device, world, reps=50):
    """The step's one collective alone: in-place fp32 sum over [21840 grads | 4 tail] floats, `reps`
    launches recorded in a CUDA graph, max over ranks; both transports: the one-shot NVLink peer-memory
    kernel the step uses at this size, and NCCL.  busBW = (S/t) * 2(W-1)/W (nccl-tests)."""
    buf = dp.flat.grad
    count = (dp.flat.numel + 4 + 3) // 4 * 4
    out = {"bytes": count * 4, "nvlink_peak_gbs": 900.0, "note": "87 KB message: latency-bound by construction",
           "used_by_step": "p2p" if dp.p2p is not None else "nccl"}
    for name, comm in (("p2p", dp.p2p), ("nccl", dp.comm)):
        if comm is None:
            continue
        for _ in range(3):
            comm.allreduce_(buf, count)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(reps):
                comm.allreduce_(buf, count)
        g.replay()
        barrier_sync(world)
        cur = torch.cuda.current_stream()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(cur)
        g.replay()
        e1.record(cur)
        barrier_sync(world)
        us = max_over_ranks(e0.elapsed_time(e1), world, device) * 1e3 / reps
        out[name + "_us"] = us
        out[name + "_bus_gbs"] = count * 4 / (us * 1e-6) * 2 * (world - 1) / world / 1e9
        buf.zero_()
    if dp.p2p is not None:
        dp.p2p.check()
    return out


def torch_eager_gpu(args, device, steps=100, warmup=20, world=1):
    """The reference's step as stock torch eager on this GPU (model.py layers via torch.nn, F.nll_loss,
    torch.optim; under torch DistributedDataParallel + NCCL when world > 1, as train.py:46-52 wraps it, with the
    blocking reduce_loss of base_trainer.py:165-174 and the per-step loss.item() of trainer.py:62) -- the 'torch
    DDP+NCCL step on the same box' the north_star names.  Not our code path.  Every rank must call it."""
    import torch.nn as nn
    import torch.nn.functional as F

    class Net(nn.Module):
        def __init__(self):
            super().__init__()
            self.conv1 = nn.Conv2d(1, 10, kernel_size=5)
            self.conv2 = nn.Conv2d(10, 20, kernel_size=5)
            self.conv2_drop = nn.Dropout2d()
            self.fc1 = nn.Linear(320, 50)
            self.fc2 = nn.Linear(50, 10)

        def forward(self, x):
            x = F.relu(F.max_pool2d(self.conv1(x), 2))
            x = F.relu(F.max_pool2d(self.conv2_drop(self.conv2(x)), 2))
            x = F.relu(self.fc1(x.view(-1, 320)))
            x = F.dropout(x, training=self.training)
            return F.log_softmax(self.fc2(x), dim=1)

    res = {}
    for mode, tf32 in (("tf32_default", True), ("ieee_fp32", False)):
        torch.backends.cudnn.allow_tf32 = tf32
        torch.manual_seed(0)
        net = Net().to(device)
        if world > 1:
            net = torch.nn.parallel.DistributedDataParallel(net, device_ids=[device.index], output_device=device.index)
        opt = torch.optim.Adam(net.parameters(), lr=1e-3, amsgrad=True) if args.optimizer == "adam" \
            else torch.optim.SGD(net.parameters(), lr=0.01)
        g = torch.Generator().manual_seed(1234)
        ring = [(torch.randn(args.batch, 1, 28, 28, generator=g).to(device),
                 torch.randint(0, 10, (args.batch,), generator=g).to(device)) for _ in range(8)]
        net.train()

        def step(i):
            x, t = ring[i % 8]
            opt.zero_grad()
            loss = F.nll_loss(net(x), t)
            red = loss
            if world > 1:
                with torch.no_grad():
                    red = loss.clone()
                    torch.distributed.reduce(red, dst=0)
            loss.backward()
            opt.step()
            return red.item()
        for i in range(warmup):
            step(i)
        barrier_sync(world)
        t0 = time.perf_counter()
        for i in range(steps):
            step(i)
        barrier_sync(world)
        el = time.perf_counter() - t0
        el = max_over_ranks(el, world, device)
        res[mode] = {"samples_per_s": world * args.batch * steps / el, "ms_per_step": 1e3 * el / steps}
        del net, opt
    torch.backends.cudnn.allow_tf32 = True
    if world > 1:
        res["how"] = "torch.nn.parallel.DistributedDataParallel + NCCL, %d ranks, whole-job samples/s" % world
    return res


_T0 = time.time()


def note(msg):
    """progress on stderr (rank 0): a run that stalls says where"""
    if int(os.environ.get("RANK", 0)) == 0:
        print("[bench %6.1f s] %s" % (time.time() - _T0, msg), file=sys.stderr, flush=True)


def run_ours(args):
    import faulthandler
    import signal
    faulthandler.register(signal.SIGTERM, chain=True)       # a run killed by a time-out says where every rank stood
    world = int(os.environ.get("WORLD_SIZE", 1))
    rank = int(os.environ.get("RANK", 0))
    local = int(os.environ.get("LOCAL_RANK", 0))
    assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU fallback)"
    # libraries (NCCL's version banner) write to fd 1: keep stdout for the ONE JSON line
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        torch.distributed.init_process_group("nccl", init_method="env://", device_id=device)
    assert world == args.gpus, "launch with torchrun --nproc-per-node %d" % args.gpus
    sampler = ClockSampler(local)

    # before anything is timed: three steps of this exact configuration (per-GPU batch, the Trainer's fused step
    # under data parallelism) against the oracle's W-rank restatement of the reference step and against this
    # framework on one GPU over the concatenated global batch.  The oracle is the checker here, nothing else.
    parity = None
    if args.optimizer == "adam" and not args.no_parity:
        from oracle.probe import dp_parity_probe
        # the oracle runs on rank 0's host cores while the other ranks spin in a collective: a thread per core
        # would oversubscribe the box (measured at 4 ranks: the probe did not finish in 240 s; with a few threads
        # it takes seconds)
        try:
            cores = len(os.sched_getaffinity(0))
        except AttributeError:
            cores = os.cpu_count() or 1
        torch.set_num_threads(max(1, min(8, cores // max(world, 1) - 1)))
        note("parity probe")
        parity = dp_parity_probe(rank, world, device, n=args.batch, steps=3, seed=9100 + args.batch, trainer_path=True)
        if rank == 0 and not parity["ok"]:
            print("parity probe FAILED: %s" % json.dumps(parity), file=sys.stderr)

    note("end-to-end loop")
    e2e_dtype = "uint8" if args.input == "uint8" else "float32"
    trainer, loader = build_trainer(args, device, world, 8, input_dtype=e2e_dtype)
    e2e_ms, log, e2e_runs = measure_e2e(trainer, loader, args, device, world, sampler)
    other = None
    if not args.no_extras:      # the other host format, for comparison (same loop, same step)
        other_dtype = "float32" if e2e_dtype == "uint8" else "uint8"
        note("end-to-end loop, %s batches" % other_dtype)
        trainer_o, loader_o = build_trainer(args, device, world, 8, input_dtype=other_dtype)
        o_ms, _, o_runs = measure_e2e(trainer_o, loader_o, args, device, world, sampler)
        px = 1 if other_dtype == "uint8" else 4
        other = {"input": other_dtype, "value": args.batch * world * args.steps / (o_ms * 1e-3), "unit": "samples/s",
                 "ms_per_step": o_ms / args.steps, "h2d_bytes_per_step": args.batch * (784 * px + 8),
                 "epochs_ms_per_step": [round(r / args.steps, 6) for r in o_runs]}
        del trainer_o, loader_o
    note("device-resident steps")
    trainer2, _ = build_trainer(args, device, world, 8, cuda_graph=False)
    ms, launches, ring, last_loss, timing = measure_value(trainer2, args, device, world, sampler)

    samples = args.batch * world * args.steps
    value = samples / (ms * 1e-3)
    e2e_value = samples / (e2e_ms * 1e-3)
    line = {
        "metric": "train samples/sec", "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, world),
        "timing": dict(timing, l2="device ring of %d batches (%.0f MB; the L2 holds 126 MB)" % (
            ring, ring * args.batch * 3144 / 2 ** 20), graph="the steps over all ring slots captured in one CUDA graph"),
        "e2e": {"value": e2e_value, "unit": "samples/s",
                "h2d_bytes_per_step": args.batch * (784 * (1 if e2e_dtype == "uint8" else 4) + 8),
                "input": "uint8 raw pixels, ToTensor + Normalize applied inside conv1's load (the format of "
                         "MnistDataLoader(raw_bytes=True)); the step computes on the same fp32 values"
                         if e2e_dtype == "uint8" else "fp32, normalised by the loader",
                "d2h_bytes_per_step": 4, "ms_per_step": e2e_ms / args.steps, "epoch_loss": log.get("loss"),
                "stat": "median of 3 epochs of %d steps" % max(args.steps, E2E_MIN_STEPS),
                "epochs_ms_per_step": [round(r / args.steps, 6) for r in e2e_runs]},
        "gpu_launches": launches * timing["steps_per_region"], "launches_per_step": launches, "last_loss": last_loss,
    }
    if parity is not None:
        line["parity"] = parity
    if other is not None:
        line["e2e_other_input"] = other
    h2d_us = measure_h2d_copy(args, device, 1 if e2e_dtype == "uint8" else 4)
    line["e2e"]["h2d_copy_us"] = round(h2d_us, 1)
    line["e2e"]["bound"] = "transfer (input copy slower than the device step on this box)" \
        if h2d_us > 1e3 * ms / args.steps else "compute"
    note("per-kernel timings, cpu baseline, torch arms")
    if rank == 0 and not args.no_extras:
        with sampler:
            fma_peak = measure_fma_peak(device)
            kernels = kernel_breakdown(args, device)
        top = max((k for k in kernels if k["in_step"]), key=lambda k: k["us"])
        total_us = sum(k["us"] for k in kernels if k["in_step"])     # the launches of the trainer's Adam step
        achieved = top["flop"] / (top["us"] * 1e-6) / 1e12
        line["roofline"] = {
            "bound": "fp32_fma", "kernel": top["kernel"], "achieved": achieved, "peak": fma_peak, "unit": "TFLOP/s",
            "frac": achieved / fma_peak, "traffic": ncu_traffic(top["kernel"]),
            "peak_source": "measured in this run by pdt_fma_peak (register-resident FFMA chains); nominal %.1f"
                           % FP32_NOMINAL_TFLOPS,
            "share_of_step": top["us"] / total_us,
            "step_frac_of_fp32_peak": (FLOP_PER_SAMPLE * args.batch / (ms / args.steps * 1e-3) / 1e12) / fma_peak,
        }
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        adam = [k for k in kernels if k["kernel"] == "adam_step_flat"][0]
        line["roofline_hbm"] = {"bound": "hbm", "kernel": "adam_step_flat",
                                "achieved": adam["bytes"] / (adam["us"] * 1e-6) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                                "frac": adam["bytes"] / (adam["us"] * 1e-6) / 1e9 / hbm_peak,
                                "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback",
                                "note": "786 KB per launch: latency-bound, not bandwidth-bound"}
        line["kernels"] = kernels
        if world == 1:
            threads = host_threads()
            n, el = time_cpu(args.batch, args.optimizer, seconds=args.cpu_seconds)
            line["cpu_baseline"] = {"value": n * args.batch / el, "unit": "samples/s", "cores": threads,
                                    "kind": "port", "host_cpus": os.cpu_count(),
                                    "sample": "%d steps of batch %d in %.1f s (oracle port of the reference step, torch CPU)"
                                              % (n, args.batch, el)}
            line["torch_eager_gpu"] = torch_eager_gpu(args, device)
    if world > 1:
        torch.distributed.barrier()      # rank 0 was busy above: the peer-memory kernels below must start together
        note("all-reduce alone, torch DDP arm")
        line["allreduce"] = measure_allreduce(trainer2.model, device, world)
        if not args.no_extras:      # every rank takes part (collectives); rank 0 reports
            line["torch_ddp_gpu"] = torch_eager_gpu(args, device, world=world)
    torch_arm = line.get("torch_ddp_gpu") or line.get("torch_eager_gpu")
    if torch_arm:       # the bar the north_star names: stock torch (DDP+NCCL when N > 1) on the same GPUs
        base = torch_arm["ieee_fp32"]["samples_per_s"]
        line["vs_torch_ddp"] = {"value": value / base, "e2e": e2e_value / base, "torch_samples_per_s": base,
                                "torch_mode": "ieee_fp32 (the TF32-default run is listed beside it)",
                                "what": "torch DistributedDataParallel + NCCL, eager" if world > 1 else "torch eager, one GPU"}
    line["clocks"] = sampler.summary()
    sys.stdout.flush()
    os.dup2(json_fd, 1)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
