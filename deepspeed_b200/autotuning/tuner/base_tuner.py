"""Tuner loop (reference ``autotuning/tuner/base_tuner.py``): ask for a batch of experiments, run them, keep the best,
stop after ``early_stopping`` consecutive batches without improvement."""
import sys

from ...utils.logging import logger


class BaseTuner:

    def __init__(self, exps, resource_manager, metric):
        self.all_exps = exps
        self.rm = resource_manager
        self.best_iter = 0
        self.best_exp = None
        self.best_metric_val = None
        self.metric = metric if metric else "throughput"
        logger.info(f"total number of exps =  {len(self.all_exps)}")

    def has_next(self):
        return len(self.all_exps) > 0

    def next_batch(self, sample_size):
        raise NotImplementedError

    def update(self):
        """Called after a batch ran (model-based tuners refit here)."""

    def tune(self, sample_size=1, n_trials=1000, early_stopping=None):
        i = 0
        try:
            while i < n_trials and self.has_next():
                sampled = self.next_batch(sample_size)
                self.rm.schedule_experiments(sampled)
                self.rm.run()
                exp, metric_val = self.rm.parse_results(self.metric)
                if self.best_exp is None or (metric_val is not None and
                                             (self.best_metric_val is None or metric_val > self.best_metric_val)):
                    if metric_val is not None:
                        self.best_exp, self.best_metric_val, self.best_iter = exp, metric_val, i
                i += len(sampled)
                self.update()
                self.rm.clear()
                if early_stopping and i >= self.best_iter + early_stopping:
                    logger.info(f"Tuner early stopped at iteration {i}. Best iteration is {self.best_iter}.")
                    break
            return i
        except Exception:
            logger.info(f"Tuner Error: {sys.exc_info()[0]}")
            return i
