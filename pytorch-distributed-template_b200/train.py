"""Training entry point (reference: train.py:16-108; same flags, same construction order).

    torchrun --nproc_per_node=N train.py -c config/config.json
    python train.py -c config/synthetic.json

The optimizer named in the config is looked up in pdt.optim first (fused flat SGD / Adam with
torch's constructor arguments) and in torch.optim otherwise; the wrapper that stands where the
reference puts DistributedDataParallel is pdt.parallel.FlatDataParallel.
"""
import argparse
import collections
import os

import numpy as np
import torch

import data_loader.data_loaders as module_data
import model.loss as module_loss
import model.metric as module_metric
import model.model as module_arch
from parse_config import ConfigParser
from pdt import ops, optim as pdt_optim
from pdt.parallel import FlatDataParallel
from trainer import Trainer
from utils.dist import get_rank, synchronize


class _OptimNamespace:
    """getattr(ns, 'Adam') -> fused class when there is one, torch.optim's otherwise."""

    def __getattr__(self, name):
        return getattr(pdt_optim, name, None) or getattr(torch.optim, name)


def main(args, config):
    logger = config.get_logger("train")
    world_size = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", args.local_rank))
    distributed = world_size > 1
    torch.cuda.set_device(local_rank)
    if args.seed is not None:
        ops.seed_dropout(args.seed)  # same stream on every rank, as the reference's torch.manual_seed
    if distributed:
        torch.distributed.init_process_group(backend="nccl", init_method="env://")
        synchronize()

    device = torch.device("cuda", local_rank)
    model = config.init_obj("arch", module_arch).to(device)
    criterion = getattr(module_loss, config["loss"])
    metrics = [getattr(module_metric, name) for name in config["metrics"]]

    trainable = [p for p in model.parameters() if p.requires_grad]
    optimizer = config.init_obj("optimizer", _OptimNamespace(), trainable)
    lr_scheduler = config.init_obj("lr_scheduler", torch.optim.lr_scheduler, optimizer)
    if distributed:
        model = FlatDataParallel(model, device_ids=[local_rank], output_device=local_rank)

    data_loader = config.init_obj("train_loader", module_data)
    valid_data_loader = None if args.no_validate else config.init_obj("valid_loader", module_data)
    if get_rank() == 0:
        logger.info(model)

    Trainer(model, criterion, metrics, optimizer, config=config, device=device, data_loader=data_loader,
            valid_data_loader=valid_data_loader, lr_scheduler=lr_scheduler).train()
    if distributed:
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    parser = argparse.ArgumentParser(description="B200-native PyTorch distributed template")
    parser.add_argument("-c", "--config", default=None, type=str, help="config file path (default: None)")
    parser.add_argument("-r", "--resume", default=None, type=str, help="path to latest checkpoint (default: None)")
    # torch >= 2.0 launchers pass --local-rank; older ones --local_rank; torchrun sets LOCAL_RANK
    parser.add_argument("-l", "--local_rank", "--local-rank", dest="local_rank", default=0, type=int,
                        help="local rank of gpu")
    parser.add_argument("-s", "--save_dir", default=None, type=str, help="dir of save path")
    parser.add_argument("--no-validate", action="store_true", help="skip validation during training")
    parser.add_argument("--seed", type=int, default=None, help="random seed")
    parser.add_argument("--deterministic", action="store_true", help="kept for CLI parity (kernels are deterministic)")
    CustomArgs = collections.namedtuple("CustomArgs", "flags type target")
    options = [
        CustomArgs(["--lr", "--learning_rate"], type=float, target="optimizer;args;lr"),
        CustomArgs(["--bs", "--batch_size"], type=int, target="train_loader;args;batch_size"),
    ]
    args, config = ConfigParser.from_args(parser, options, training=True)
    if args.seed is not None:
        torch.manual_seed(args.seed)
        np.random.seed(args.seed)
        torch.backends.cudnn.deterministic = args.deterministic
        torch.backends.cudnn.benchmark = False
    main(args, config)
