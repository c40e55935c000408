"""Trainer: the hot loop (reference: trainer/trainer.py:11-123; inner loop :45-72).

Per step the call sequence is the reference's -- zero_grad, model(data), criterion,
reduce_loss, backward, optimizer.step -- through the same plug-in objects.  What differs is
how it is driven on a B200:
  * `trainer.cuda_graph` (default on for CUDA runs): after one eager step the whole sequence
    is captured once into a CUDA graph on static input buffers and replayed; a step is then
    one H2D copy plus one graph launch instead of ~230 dispatcher calls.
  * the rank-0 `loss.item()` of every step (reference :62) is deferred: per-step losses stay
    on the device and are read in one transfer at each log step / epoch end, in order, so the
    logged numbers are the same but the stream is never drained inside the epoch.
"""
import numpy as np
import torch

from base import BaseTrainer
from utils import MetricTracker, inf_loop
from utils.dist import is_main_process, synchronize


class Trainer(BaseTrainer):
    def __init__(self, model, criterion, metric_ftns, optimizer, config, device,
                 data_loader, valid_data_loader=None, lr_scheduler=None, len_epoch=None):
        super().__init__(model, criterion, metric_ftns, optimizer, config)
        self.config = config
        self.device = torch.device(device)
        self.data_loader = data_loader
        if len_epoch is None:
            self.len_epoch = len(self.data_loader)  # epoch-based
        else:
            self.data_loader = inf_loop(data_loader)  # iteration-based
            self.len_epoch = len_epoch
        self.valid_data_loader = valid_data_loader
        self.do_validation = valid_data_loader is not None
        self.lr_scheduler = lr_scheduler
        if lr_scheduler is not None and getattr(self, "_resumed_scheduler_state", None):
            lr_scheduler.load_state_dict(self._resumed_scheduler_state)
        self.log_step = max(1, int(np.sqrt(data_loader.batch_size)))
        self.train_metrics = MetricTracker("loss", writer=self.writer)
        self.valid_metrics = MetricTracker("loss", *[m.__name__ for m in self.metric_ftns], writer=self.writer)
        cfg = config["trainer"]
        self.use_graph = bool(cfg.get("cuda_graph", True)) and self.device.type == "cuda"
        self.sync_every_step = bool(cfg.get("sync_every_step", False))
        self._graph = None
        self._eager_steps = 0

    # -- one optimisation step (reference trainer.py:48-58) ---------------------------------
    def _step(self, data, target):
        self.optimizer.zero_grad()
        output = self.model(data)
        loss = self.criterion(output, target)
        loss_reduced = self.reduce_loss(loss)
        loss.backward()
        self.optimizer.step()
        return loss_reduced

    def _graphed_step(self, data, target):
        g = self._graph
        if g is None or g["shape"] != (tuple(data.shape), tuple(target.shape)):
            if self._eager_steps == 0 or (g is not None):
                # first step ever (lazy workspaces get created), or an odd-sized batch
                self._eager_steps += 1
                data = data.to(self.device, non_blocking=True)
                target = target.to(self.device, non_blocking=True)
                return self._step(data, target).detach().clone()
            g = self._capture(data, target)
        g["data"].copy_(data, non_blocking=True)
        g["target"].copy_(target, non_blocking=True)
        g["graph"].replay()
        return g["loss"].detach().clone()

    def _capture(self, data, target):
        static_data = torch.empty(data.shape, dtype=data.dtype, device=self.device)
        static_target = torch.empty(target.shape, dtype=target.dtype, device=self.device)
        static_data.copy_(data)
        static_target.copy_(target)
        torch.cuda.synchronize(self.device)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):  # records, does not execute: the replay below IS this step
            loss = self._step(static_data, static_target)
        self._graph = dict(graph=graph, data=static_data, target=static_target, loss=loss,
                           shape=(tuple(data.shape), tuple(target.shape)))
        return self._graph

    # -- deferred loss read-out ---------------------------------------------------------------
    def _flush(self, pending, epoch, log_idx=None):
        if not pending:
            return
        values = torch.stack([v for _, v in pending]).float().cpu().tolist()
        for (batch_idx, _), value in zip(pending, values):
            self.writer.set_step((epoch - 1) * self.len_epoch + batch_idx)
            self.train_metrics.update("loss", value)
            if batch_idx == log_idx:
                self.logger.debug("Train Epoch: {} {} Loss: {:.6f}".format(epoch, self._progress(batch_idx + 1), value))
        pending.clear()

    def _train_epoch(self, epoch):
        self.model.train()
        self.train_metrics.reset()
        main = is_main_process()
        pending = []
        for batch_idx, (data, target) in enumerate(self.data_loader):
            if self.use_graph:
                loss_reduced = self._graphed_step(data, target)
            else:
                data, target = data.to(self.device), target.to(self.device)
                loss_reduced = self._step(data, target)
            if main:
                pending.append((batch_idx, loss_reduced.detach()))
                at_log = batch_idx % self.log_step == 0
                if at_log or self.sync_every_step:
                    self._flush(pending, epoch, batch_idx if at_log else None)
                    if at_log and self.writer.enabled:
                        from torchvision.utils import make_grid
                        self.writer.add_image("input", make_grid(data.cpu(), nrow=8, normalize=True))
            if batch_idx == self.len_epoch:
                break
        if main:
            self._flush(pending, epoch)
        log = self.train_metrics.result()

        if self.do_validation:
            self.valid_metrics.reset()
            outputs, targets = self._valid_epoch(epoch)
            synchronize()
            outputs = self._accumulate_predictions_from_multiple_gpus(outputs)
            targets = self._accumulate_predictions_from_multiple_gpus(targets)
            if main:
                out = torch.cat([o.to(self.device) for o in outputs], dim=0)
                tgt = torch.cat([t.to(self.device) for t in targets], dim=0)
                for met in self.metric_ftns:
                    self.valid_metrics.update(met.__name__, met(out, tgt))
                log.update(**{"val_" + k: v for k, v in self.valid_metrics.result().items()})

        if self.lr_scheduler is not None:
            self.lr_scheduler.step()
        return log

    def _valid_epoch(self, epoch):
        self.model.eval()
        outputs, targets = [], []
        with torch.no_grad():
            for data, target in self.valid_data_loader:
                data, target = data.to(self.device), target.to(self.device)
                outputs.append(self.model(data).clone())
                targets.append(target.clone())
        return torch.cat(outputs, dim=0), torch.cat(targets, dim=0)

    def _progress(self, batch_idx):
        if hasattr(self.data_loader, "n_samples"):
            current, total = batch_idx * self.data_loader.batch_size, self.data_loader.n_samples
        else:
            current, total = batch_idx, self.len_epoch
        return "[{}/{} ({:.0f}%)]".format(current, total, 100.0 * current / total)
