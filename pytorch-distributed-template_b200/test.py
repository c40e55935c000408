"""Evaluation entry point (reference: test.py:14-128; same flags).

    python test.py -r saved/.../checkpoint-epoch1.pth
    torchrun --nproc_per_node=N test.py -r ...

The reference gathers every output tensor of every batch to every rank through pickle
(test.py:87-90) and computes the metrics on rank 0.  Here each rank reduces its shard to the
fused kernel's integer counts and loss sum on the device; one small all-reduce combines them.
"""
import argparse
import os

import numpy as np
import torch

import data_loader.data_loaders as module_data
import model.loss as module_loss
import model.metric as module_metric
import model.model as module_arch
from parse_config import ConfigParser
from pdt import ops
from pdt.parallel import FlatDataParallel
from utils.dist import get_rank, get_world_size, synchronize


def evaluate(model, loader, metric_fns, device):
    """Returns {'loss': mean nll over the samples seen by all ranks, metric_name: value, ...}."""
    model.eval()
    ks = sorted({getattr(m, "pdt_count", ("topk", 3))[1] for m in metric_fns}) or [3]
    totals = {k: torch.zeros(4, dtype=torch.int64, device=device) for k in ks}
    loss_sum = torch.zeros(1, dtype=torch.float64, device=device)
    with torch.no_grad():
        for data, target in loader:
            data, target = data.to(device, non_blocking=True), target.to(device, non_blocking=True)
            output = model(data)
            for k in ks:
                st = ops.stats_async(output, target, k)
                totals[k] += st[2:6]
            loss_sum += st[1:2].view(torch.float32).double()
    packed = torch.cat([totals[k].double() for k in ks] + [loss_sum])
    if get_world_size() > 1:
        torch.distributed.all_reduce(packed)
    host = packed.cpu().tolist()
    per_k = {k: host[4 * i:4 * i + 4] for i, k in enumerate(ks)}
    n_rows = per_k[ks[0]][3]
    log = {"loss": host[-1] / max(n_rows, 1.0)}
    for m in metric_fns:
        kind, k = getattr(m, "pdt_count")
        top1, topk, _, rows = per_k[k if kind == "topk" else ks[0]]
        log[m.__name__] = (top1 if kind == "top1" else topk) / rows
    return log


def main(args, config):
    logger = config.get_logger("test")
    world_size = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", args.local_rank))
    torch.cuda.set_device(local_rank)
    if world_size > 1:
        torch.distributed.init_process_group(backend="nccl", init_method="env://")
        synchronize()
    device = torch.device("cuda", local_rank)
    model = config.init_obj("arch", module_arch).to(device)
    loader = config.init_obj("test_loader", module_data)
    metric_fns = [getattr(module_metric, name) for name in config["metrics"]]
    getattr(module_loss, config["loss"])          # the configured loss must exist; nll is what the kernel sums

    if get_rank() == 0:
        logger.info(model)
        logger.info("Loading checkpoint: {} ...".format(config.resume))
    checkpoint = torch.load(str(config.resume), map_location="cpu", weights_only=False)
    model.load_state_dict(checkpoint["state_dict"])
    if world_size > 1:
        model = FlatDataParallel(model, device_ids=[local_rank], output_device=local_rank)
    log = evaluate(model, loader, metric_fns, device)
    if get_rank() == 0:
        logger.info(log)
    if world_size > 1:
        torch.distributed.destroy_process_group()
    return log


if __name__ == "__main__":
    parser = argparse.ArgumentParser(description="B200-native PyTorch distributed template: evaluation")
    parser.add_argument("-c", "--config", default=None, type=str, help="config file path (default: None)")
    parser.add_argument("-r", "--resume", default=None, type=str, help="path to the checkpoint to evaluate")
    parser.add_argument("-l", "--local_rank", "--local-rank", dest="local_rank", default=0, type=int)
    parser.add_argument("-s", "--save_dir", default=None, type=str, help="dir of save path")
    parser.add_argument("--seed", type=int, default=None, help="random seed")
    parser.add_argument("--deterministic", action="store_true")
    args, config = ConfigParser.from_args(parser, training=False)
    if args.seed is not None:
        torch.manual_seed(args.seed)
        np.random.seed(args.seed)
    main(args, config)
