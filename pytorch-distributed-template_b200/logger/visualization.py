"""TensorBoard proxy (reference: logger/visualization.py:5-73): add_* calls get the current step
and a '/train' or '/valid' tag suffix; when TensorBoard is off or missing they do nothing."""
import importlib
from datetime import datetime

_FORWARDED = {"add_scalar", "add_scalars", "add_image", "add_images", "add_audio", "add_text",
              "add_histogram", "add_pr_curve", "add_embedding"}
_UNTAGGED = {"add_histogram", "add_embedding"}


class TensorboardWriter:
    def __init__(self, log_dir, logger, enabled):
        self.writer = None
        self.selected_module = ""
        self.step = 0
        self.mode = ""
        self.timer = datetime.now()
        if enabled:
            for name in ("torch.utils.tensorboard", "tensorboardX"):
                try:
                    self.writer = importlib.import_module(name).SummaryWriter(str(log_dir))
                    self.selected_module = name
                    break
                except ImportError:
                    continue
            if self.writer is None:
                logger.warning("Warning: visualization (Tensorboard) is configured to use, but currently not "
                               "installed on this machine. Install tensorboard / tensorboardX or set "
                               "'tensorboard': false in the config file.")

    @property
    def enabled(self):
        return self.writer is not None

    def set_step(self, step, mode="train"):
        self.mode, self.step = mode, step
        now = datetime.now()
        if step != 0:
            elapsed = (now - self.timer).total_seconds()
            if elapsed > 0:
                self.add_scalar("steps_per_sec", 1.0 / elapsed)
        self.timer = now

    def __getattr__(self, name):
        if name in _FORWARDED:
            target = getattr(self.writer, name, None)

            def forward(tag, data, *args, **kwargs):
                if target is None:
                    return
                full = tag if name in _UNTAGGED else "{}/{}".format(tag, self.mode)
                target(full, data, self.step, *args, **kwargs)
            return forward
        raise AttributeError("type object '{}' has no attribute '{}'".format(self.selected_module, name))
