"""world_size-2 gloo tests (CPU) of the distributed host logic: utils.dist, the generic
BaseTrainer.reduce_loss path (reference base_trainer.py:165-174), early-stop agreement and the
synthetic loader's rank slicing."""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "pytorch-distributed-template_b200")


def _worker(rank, world, port, tmp, ret):
    for p in (PKG, ROOT):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from utils.dist import all_gather, get_rank, get_world_size, is_main_process, reduce_tensor, synchronize
    from data_loader.data_loaders import SyntheticDataLoader
    from base import BaseTrainer
    out = {}
    out["ids"] = (get_rank(), get_world_size(), is_main_process())
    synchronize()
    out["gather"] = all_gather({"rank": rank, "blob": "x" * (10 + 100 * rank)})   # ragged payloads
    out["reduce_mean"] = reduce_tensor(torch.tensor([float(rank + 1)])).item()
    out["reduce_sum"] = reduce_tensor(torch.tensor([float(rank + 1)]), average=False).item()
    # reduce_loss, generic path: rank 0 gets the mean, other ranks' value is unspecified
    loss = torch.tensor(2.0 + rank)
    out["reduce_loss"] = BaseTrainer.reduce_loss(type("T", (), {"_folded_tail": lambda s, l: None})(), loss).item()
    # early-stop agreement: rank 0 owns the counter, every rank takes the max (base_trainer.py:101-107)
    out["early"] = max(all_gather(5 if rank == 0 else 0))
    ld = SyntheticDataLoader(4, n_batches=1, seed=11, pin=False)
    out["slice"] = ld.batches[0][0].clone()
    ret[rank] = out
    dist.barrier()
    dist.destroy_process_group()


def test_world2_gloo(tmp_path):
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(2, 29561, str(tmp_path), ret), nprocs=2, join=True)
    r0, r1 = ret[0], ret[1]
    assert r0["ids"] == (0, 2, True) and r1["ids"] == (1, 2, False)
    for r in (r0, r1):
        assert [g["rank"] for g in r["gather"]] == [0, 1]
        assert len(r["gather"][1]["blob"]) == 110
        assert r["reduce_mean"] == 1.5 and r["reduce_sum"] == 3.0
        assert r["early"] == 5
    assert r0["reduce_loss"] == 2.5          # mean(2, 3) on rank 0
    # the two ranks serve the two halves of one global batch
    g = torch.Generator().manual_seed(11)
    x = torch.randn(8, 1, 28, 28, generator=g)
    assert torch.equal(r0["slice"], x[:4]) and torch.equal(r1["slice"], x[4:])
