"""Helpers of the plug-in surface (reference: utils/__init__.py re-exports util.py and dist.py)."""
from .util import MetricTracker, ensure_dir, inf_loop, near_gpu, prepare_device, read_json, write_json  # noqa: F401
from .dist import all_gather, get_rank, get_world_size, is_main_process, reduce_tensor, synchronize  # noqa: F401
