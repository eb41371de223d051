"""Monitors: fan out ``(tag, value, step)`` events on rank 0 to TensorBoard / W&B / CSV / Comet.

Reference: ``monitor/monitor.py:30-60`` (MonitorMaster), per-backend writers ``monitor/{tensorboard,wandb,
csv_monitor,comet}.py``.
"""
import csv
import os

from .. import comm as dist
from ..utils.logging import logger


class Monitor:

    def __init__(self, cfg):
        self.cfg = cfg
        self.enabled = getattr(cfg, "enabled", False)

    def write_events(self, event_list):
        raise NotImplementedError


class TensorBoardMonitor(Monitor):

    def __init__(self, cfg):
        super().__init__(cfg)
        self.summary_writer = None
        if self.enabled and dist.get_rank() == 0:
            try:
                from torch.utils.tensorboard import SummaryWriter
                base = cfg.output_path or os.path.join(os.path.expanduser("~"), "tensorboard")
                log_dir = os.path.join(base, cfg.job_name)
                os.makedirs(log_dir, exist_ok=True)
                self.summary_writer = SummaryWriter(log_dir=log_dir)
            except Exception as e:  # tensorboard not installed
                logger.warning(f"TensorBoard monitor disabled: {e}")
                self.enabled = False

    def write_events(self, event_list, flush=True):
        if self.enabled and self.summary_writer is not None:
            for name, value, step in event_list:
                self.summary_writer.add_scalar(name, value, step)
            if flush:
                self.summary_writer.flush()


class WandbMonitor(Monitor):

    def __init__(self, cfg):
        super().__init__(cfg)
        self._wandb = None
        if self.enabled and dist.get_rank() == 0:
            try:
                import wandb
                wandb.init(project=cfg.project, group=cfg.group, entity=cfg.team)
                self._wandb = wandb
            except Exception as e:
                logger.warning(f"W&B monitor disabled: {e}")
                self.enabled = False

    def write_events(self, event_list):
        if self.enabled and self._wandb is not None:
            for name, value, step in event_list:
                self._wandb.log({name: value}, step=step)


class EventsLogScheduler:
    """Rate limit per event name: an event is written when at least ``samples_log_interval`` samples have passed since
    it was last written (reference ``monitor/comet.py:74``)."""

    def __init__(self, samples_log_interval):
        self._interval = samples_log_interval
        self._last = {}

    def needs_logging(self, name, current_sample):
        last = self._last.get(name)
        if last is None or current_sample - last >= self._interval:
            self._last[name] = current_sample
            return True
        return False


class CometMonitor(Monitor):

    def __init__(self, cfg):
        super().__init__(cfg)
        self._exp = None
        if self.enabled and dist.get_rank() == 0:
            try:
                import comet_ml
                self._exp = comet_ml.start(api_key=cfg.api_key, project=cfg.project, workspace=cfg.workspace,
                                           experiment_key=cfg.experiment_key, mode=cfg.mode, online=cfg.online)
                if cfg.experiment_name:
                    self._exp.set_name(cfg.experiment_name)
            except Exception as e:
                logger.warning(f"Comet monitor disabled: {e}")
                self.enabled = False
        self.samples_log_interval = getattr(cfg, "samples_log_interval", 100)

    def write_events(self, event_list):
        if self.enabled and self._exp is not None:
            if getattr(self, "_sched", None) is None:
                self._sched = EventsLogScheduler(int(getattr(self.cfg, "samples_log_interval", 0) or 0))
            for name, value, step in event_list:
                if self._sched.needs_logging(name, step):
                    self._exp.log_metric(name, value, step=step)


class csvMonitor(Monitor):

    def __init__(self, cfg):
        super().__init__(cfg)
        self.filenames = {}
        self.log_dir = None
        if self.enabled and dist.get_rank() == 0:
            base = cfg.output_path or os.path.join(os.path.expanduser("~"), "csv_monitor")
            self.log_dir = os.path.join(base, cfg.job_name)
            os.makedirs(self.log_dir, exist_ok=True)

    def write_events(self, event_list):
        if not (self.enabled and self.log_dir):
            return
        for name, value, step in event_list:
            short = name.split("/")[-1]
            fname = os.path.join(self.log_dir, short.replace(" ", "_") + ".csv")
            new = fname not in self.filenames and not os.path.exists(fname)
            self.filenames[fname] = True
            with open(fname, "a+", newline="") as f:
                w = csv.writer(f)
                if new:
                    w.writerow(["step", short])
                w.writerow([step, value])


class MonitorMaster(Monitor):

    def __init__(self, monitor_config):
        self.cfg = monitor_config
        self.enabled = monitor_config.enabled
        self.tb_monitor = self.wandb_monitor = self.csv_monitor = self.comet_monitor = None
        if dist.get_rank() == 0:
            if monitor_config.tensorboard.enabled:
                self.tb_monitor = TensorBoardMonitor(monitor_config.tensorboard)
            if monitor_config.wandb.enabled:
                self.wandb_monitor = WandbMonitor(monitor_config.wandb)
            if monitor_config.csv_monitor.enabled:
                self.csv_monitor = csvMonitor(monitor_config.csv_monitor)
            if monitor_config.comet.enabled:
                self.comet_monitor = CometMonitor(monitor_config.comet)

    def write_events(self, event_list):
        if dist.get_rank() == 0:
            for m in (self.tb_monitor, self.wandb_monitor, self.csv_monitor, self.comet_monitor):
                if m is not None:
                    m.write_events(event_list)
