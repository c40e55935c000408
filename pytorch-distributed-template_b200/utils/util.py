"""Host-side utilities (reference: utils/util.py).  MetricTracker keeps the reference's
interface but is a plain dict accumulator: the pandas version (utils/util.py:46-67) raises under
pandas 3 and would force a DataFrame update into every training step."""
import json
from collections import OrderedDict
from itertools import repeat
from pathlib import Path

import torch


def ensure_dir(dirname):
    Path(dirname).mkdir(parents=True, exist_ok=True)


def read_json(fname):
    with Path(fname).open("rt") as fh:
        return json.load(fh, object_hook=OrderedDict)


def write_json(content, fname):
    with Path(fname).open("wt") as fh:
        json.dump(content, fh, indent=4, sort_keys=False)


def inf_loop(data_loader):
    """Endless iteration over a finite loader (iteration-based training)."""
    for loader in repeat(data_loader):
        yield from loader


def prepare_device(n_gpu_use):
    """Kept for interface parity; one process drives one GPU in this framework."""
    n_gpu = torch.cuda.device_count()
    n_gpu_use = max(0, min(n_gpu_use, n_gpu))
    device = torch.device("cuda:0" if n_gpu_use > 0 else "cpu")
    return device, list(range(n_gpu_use))


class MetricTracker:
    def __init__(self, *keys, writer=None):
        self.writer = writer
        self._keys = list(keys)
        self.reset()

    def reset(self):
        self._total = {k: 0.0 for k in self._keys}
        self._count = {k: 0 for k in self._keys}

    def update(self, key, value, n=1):
        if self.writer is not None:
            self.writer.add_scalar(key, value)
        self._total[key] += value * n
        self._count[key] += n

    def avg(self, key):
        return self._total[key] / self._count[key] if self._count[key] else 0

    def result(self):
        return {k: self.avg(k) for k in self._keys}


class near_gpu:
    """Context manager: run the enclosed host allocations on the CPUs next to a GPU, so that pinned staging
    buffers (first touched, hence placed, by the allocating thread) land on that GPU's NUMA node.  On an
    8-GPU box every step moves 8 x 3.2 MB host->device; buffers on the far socket halve that rate.  Topology
    comes from /sys (pci bus id -> local_cpulist); anything missing or odd makes this a no-op."""

    def __init__(self, device_index=None):
        self.cpus, self.saved = None, None
        try:
            import os
            import torch
            if device_index is None:
                device_index = torch.cuda.current_device()
            if os.environ.get("PDT_NUMA_BIND", "1") == "0" or not hasattr(os, "sched_setaffinity"):
                return
            props = torch.cuda.get_device_properties(device_index)
            bdf = "%04x:%02x:%02x.0" % (props.pci_domain_id, props.pci_bus_id, props.pci_device_id)
            path = "/sys/bus/pci/devices/%s/local_cpulist" % str(bdf).lower()
            cpus = set()
            for part in open(path).read().strip().split(","):
                if part:
                    a, _, b = part.partition("-")
                    cpus.update(range(int(a), int(b or a) + 1))
            allowed = os.sched_getaffinity(0)
            cpus &= allowed
            if cpus and cpus != allowed:
                self.cpus = cpus
        except Exception:
            self.cpus = None

    def __enter__(self):
        if self.cpus:
            try:
                import os
                self.saved = os.sched_getaffinity(0)
                os.sched_setaffinity(0, self.cpus)
            except Exception:
                self.saved = None
        return self

    def __exit__(self, *a):
        if self.saved is not None:
            try:
                import os
                os.sched_setaffinity(0, self.saved)
            except Exception:
                pass
        return False
