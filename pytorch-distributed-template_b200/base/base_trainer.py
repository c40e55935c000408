"""BaseTrainer: epoch loop, monitoring / early stop, checkpoint save and resume, reduce_loss
(reference: base/base_trainer.py:10-180; same constructor, attributes and method names)."""
from abc import abstractmethod
from math import inf

import torch
import torch.distributed as dist

from logger import TensorboardWriter
from utils.dist import all_gather, get_world_size, is_main_process, synchronize


class BaseTrainer:
    def __init__(self, model, criterion, metric_ftns, optimizer, config):
        self.config = config
        cfg = config["trainer"]
        self.logger = config.get_logger("trainer", cfg["verbosity"])
        self.model, self.criterion = model, criterion
        self.metric_ftns, self.optimizer = metric_ftns, optimizer
        self.epochs = cfg["epochs"]
        self.save_period = cfg["save_period"]
        self.monitor = cfg.get("monitor", "off")
        if self.monitor == "off":
            self.mnt_mode, self.mnt_best = "off", 0
        else:
            self.mnt_mode, self.mnt_metric = self.monitor.split()
            assert self.mnt_mode in ("min", "max")
            self.mnt_best = inf if self.mnt_mode == "min" else -inf
            self.early_stop = cfg.get("early_stop", inf)
            if self.early_stop <= 0:
                self.early_stop = inf
        self.start_epoch = 1
        self.checkpoint_dir = config.save_dir
        self.writer = TensorboardWriter(config.log_dir, self.logger, cfg["tensorboard"])
        if config.resume is not None:
            self._resume_checkpoint(config.resume)

    @abstractmethod
    def _train_epoch(self, epoch):
        raise NotImplementedError

    def _improved(self, log):
        """True / False, or None when the monitored metric is missing (monitoring is then disabled)."""
        try:
            value = log[self.mnt_metric]
        except KeyError:
            self.logger.warning("Warning: Metric '{}' is not found. Model performance monitoring is "
                                "disabled.".format(self.mnt_metric))
            self.mnt_mode = "off"
            return None
        return value <= self.mnt_best if self.mnt_mode == "min" else value >= self.mnt_best

    def train(self):
        not_improved_count = 0
        for epoch in range(self.start_epoch, self.epochs + 1):
            result = self._train_epoch(epoch)
            if is_main_process():
                log = {"epoch": epoch}
                log.update(result)
                for key, value in log.items():
                    self.logger.info("    {:15s}: {}".format(str(key), value))
                best = False
                if self.mnt_mode != "off":
                    improved = self._improved(log)
                    if improved:
                        self.mnt_best = log[self.mnt_metric]
                        not_improved_count = 0
                        best = True
                    elif improved is not None:
                        not_improved_count += 1
                if epoch % self.save_period == 0:
                    self._save_checkpoint(epoch, save_best=best)
            # every rank agrees on stopping: rank 0 owns the counter, all take the max
            synchronize()
            if max(all_gather(not_improved_count)) > getattr(self, "early_stop", inf):
                if is_main_process():
                    self.logger.info("Validation performance didn't improve for {} epochs. "
                                     "Training stops.".format(self.early_stop))
                break

    def _save_checkpoint(self, epoch, save_best=False):
        """Same keys as the reference (arch, epoch, state_dict, optimizer, monitor_best, config);
        state_dict keys are un-prefixed; `config` is stored as a plain dict so the file loads with
        torch.load(weights_only=True)."""
        state = {
            "arch": type(self.model).__name__,
            "epoch": epoch,
            "state_dict": {k.replace("module.", ""): v.detach().clone() for k, v in self.model.state_dict().items()},
            "optimizer": self.optimizer.state_dict(),
            "monitor_best": self.mnt_best,
            "config": _plain(self.config.config),
        }
        sched = getattr(self, "lr_scheduler", None)
        if sched is not None:
            state["lr_scheduler"] = sched.state_dict()
        filename = str(self.checkpoint_dir / "checkpoint-epoch{}.pth".format(epoch))
        torch.save(state, filename)
        self.logger.info("Saving checkpoint: {} ...".format(filename))
        if save_best:
            torch.save(state, str(self.checkpoint_dir / "model_best.pth"))
            self.logger.info("Saving current best: model_best.pth ...")

    def _resume_checkpoint(self, resume_path):
        resume_path = str(resume_path)
        if is_main_process():
            self.logger.info("Loading checkpoint: {} ...".format(resume_path))
        # weights_only=False: checkpoints written by the reference pickle a ConfigParser
        checkpoint = torch.load(resume_path, map_location="cpu", weights_only=False)
        self.start_epoch = checkpoint["epoch"] + 1
        self.mnt_best = checkpoint["monitor_best"]
        if checkpoint["config"]["arch"] != self.config["arch"]:
            self.logger.warning("Warning: Architecture configuration given in config file is different from that "
                                "of checkpoint. This may yield an exception while state_dict is being loaded.")
        state = checkpoint["state_dict"]
        if any(k.startswith("module.") for k in self.model.state_dict().keys()):
            state = {"module." + k: v for k, v in state.items()}
        self.model.load_state_dict(state)
        if checkpoint["config"]["optimizer"]["type"] != self.config["optimizer"]["type"]:
            self.logger.warning("Warning: Optimizer type given in config file is different from that of "
                                "checkpoint. Optimizer parameters not being resumed.")
        else:
            self.optimizer.load_state_dict(checkpoint["optimizer"])
        self._resumed_scheduler_state = checkpoint.get("lr_scheduler")
        self.logger.info("Checkpoint loaded. Resume training from epoch {}".format(self.start_epoch))

    def reduce_loss(self, loss):
        """Mean of the per-rank losses, defined on rank 0 (reference base_trainer.py:165-174).
        With the native model under FlatDataParallel the value rides in the tail of the gradient
        all-reduce: the returned 0-dim tensor is that tail slot and holds the mean once
        loss.backward() has run (every rank gets it)."""
        world_size = get_world_size()
        if world_size < 2:
            return loss
        tail = self._folded_tail(loss)
        if tail is not None:
            return tail[0]
        with torch.no_grad():
            all_loss = loss.clone()
            dist.reduce(all_loss, dst=0)
            if dist.get_rank() == 0:
                all_loss /= world_size
        return all_loss

    def _folded_tail(self, loss):
        dp = self.model
        if getattr(dp, "native", False) and getattr(loss, "_pdt_stats", None) is not None \
                and self.config["trainer"].get("fold_loss_reduce", True):
            return dp.reduced_tail()
        return None

    def _accumulate_predictions_from_multiple_gpus(self, predictions_per_gpu):
        gathered = all_gather(predictions_per_gpu)
        return gathered if is_main_process() else None


def _plain(obj):
    if isinstance(obj, dict):
        return {k: _plain(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return [_plain(v) for v in obj]
    return obj
