"""`train_loader` / `valid_loader` plug-ins (reference: data_loader/data_loaders.py:8-27).

MnistDataLoader keeps the reference's signature (torchvision MNIST, ToTensor + Normalize,
DistributedSampler when world_size > 1).  SyntheticDataLoader is an addition: MNIST-shaped
N(0,1) batches from pinned host memory, for machines without the dataset or a network --
it is what bench.py and the tests drive the Trainer with (dtype="uint8": uniform random raw pixels).
"""
import torch

from base import BaseDataLoader
from utils.dist import get_rank, get_world_size
from utils.util import near_gpu


class MnistDataLoader(BaseDataLoader):
    """raw_bytes=True (an addition): batches stay uint8 [N,1,28,28] -- the pixels as the MNIST files hold them -- and
    MnistModel applies ToTensor + Normalize((0.1307,), (0.3081,)) while conv1 loads them on the GPU: the same
    values, a quarter of the host-to-device traffic, no per-pixel work in the loader's workers."""

    def __init__(self, data_dir, batch_size, shuffle=True, num_workers=1, training=True, raw_bytes=False):
        from torchvision import datasets, transforms
        if raw_bytes:
            trsfm = transforms.PILToTensor()
        else:
            trsfm = transforms.Compose([transforms.ToTensor(), transforms.Normalize((0.1307,), (0.3081,))])
        self.data_dir = data_dir
        self.dataset = datasets.MNIST(self.data_dir, train=training, download=True, transform=trsfm)
        sampler = None
        if get_world_size() > 1:
            sampler = torch.utils.data.distributed.DistributedSampler(self.dataset, shuffle=shuffle)
        super().__init__(self.dataset, batch_size, shuffle, num_workers, sampler=sampler,
                         pin_memory=torch.cuda.is_available())


class SyntheticDataLoader:
    """`n_batches` batches of x ~ N(0,1) [batch_size,1,28,28] fp32 and t ~ U{0..9} int64, drawn once
    from torch.Generator().manual_seed(seed + slot) and served from pinned memory.  With
    world_size W, rank r serves rows [r*B, (r+1)*B) of each W*B global batch."""

    def __init__(self, batch_size, n_batches=8, seed=1234, shape=(1, 28, 28), num_classes=10, pin=True,
                 training=True, dtype="float32"):
        self.batch_size = batch_size
        self.n_batches = n_batches
        self.n_samples = batch_size * n_batches
        world, rank = get_world_size(), get_rank()
        self.batches = []
        pin = pin and torch.cuda.is_available()
        for slot in range(n_batches):
            g = torch.Generator().manual_seed(seed + slot)
            if dtype == "uint8":     # raw pixels (MnistDataLoader(raw_bytes=True)'s format): normalised on the GPU
                x = torch.randint(0, 256, (world * batch_size,) + tuple(shape), generator=g, dtype=torch.uint8)
            else:
                x = torch.randn(world * batch_size, *shape, generator=g)
            t = torch.randint(0, num_classes, (world * batch_size,), generator=g, dtype=torch.int64)
            x = x[rank * batch_size:(rank + 1) * batch_size].contiguous()
            t = t[rank * batch_size:(rank + 1) * batch_size].contiguous()
            if pin:
                with near_gpu():        # pinned pages next to this rank's GPU (no-op on single-node hosts)
                    x, t = x.pin_memory(), t.pin_memory()
            self.batches.append((x, t))
        self.sampler = range(self.n_samples)

    def __len__(self):
        return self.n_batches

    def __iter__(self):
        return iter(self.batches)
