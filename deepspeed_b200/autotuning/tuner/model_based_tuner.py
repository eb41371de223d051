"""Model-based tuner (reference ``tuner/model_based_tuner.py`` with an XGBoost cost model): fit a cost model on the
experiments seen so far and evaluate the most promising unseen configs next, with epsilon-greedy exploration.
The cost model here is ridge regression on one-hot / numeric features (no external dependency)."""
import random

import numpy as np

from ..utils import _flatten
from .base_tuner import BaseTuner

INIT_NUM = 2


class RidgeCostModel:

    def __init__(self, l2=1e-3):
        self.l2 = l2
        self.w = None

    def fit(self, xs, ys):
        X = np.asarray(xs, dtype=np.float64)
        y = np.asarray(ys, dtype=np.float64)
        y = y / max(np.max(np.abs(y)), 1e-9)
        X = np.concatenate([X, np.ones((len(X), 1))], axis=1)
        self.w = np.linalg.solve(X.T @ X + self.l2 * np.eye(X.shape[1]), X.T @ y)

    def predict(self, xs):
        X = np.asarray(xs, dtype=np.float64)
        X = np.concatenate([X, np.ones((len(X), 1))], axis=1)
        return X @ self.w


class ModelBasedTuner(BaseTuner):

    def __init__(self, exps, resource_manager, metric, tuning_space=None):
        super().__init__(exps, resource_manager, metric)
        self.tuning_space = tuning_space
        self.cost_model = RidgeCostModel()
        self.evaluated_configs, self.evaluated_perf = [], []
        self.visited = set()
        self.trials = []
        self.random_exploration_ratio = 0.2
        self.num_all_configs = len(exps)
        flats = [_flatten(e["ds_config"]) for e in exps]
        self.keys = sorted({k for f in flats for k, v in f.items() if isinstance(v, (int, float, bool))})
        self.features = [[float(f.get(k, 0)) for k in self.keys] for f in flats]
        scale = np.maximum(np.max(np.abs(np.asarray(self.features)), axis=0), 1e-9) if self.features else 1
        self.features = (np.asarray(self.features) / scale).tolist() if self.features else []

    def has_next(self):
        return len(self.visited) < self.num_all_configs

    def next_batch(self, sample_size):
        batch = []
        while len(batch) < sample_size and self.has_next():
            unseen = [i for i in range(self.num_all_configs) if i not in self.visited]
            if len(self.evaluated_configs) < INIT_NUM or random.random() < self.random_exploration_ratio:
                idx = random.choice(unseen)
            else:
                preds = self.cost_model.predict([self.features[i] for i in unseen])
                idx = unseen[int(np.argmax(preds))]
            self.visited.add(idx)
            self.trials.append(idx)
            batch.append(self.all_exps[idx])
        return batch

    def update(self):
        for exp_id, (exp, err) in self.rm.finished_experiments.items():
            idx = next((i for i, e in enumerate(self.all_exps) if e["name"] == exp["name"]), None)
            if idx is None or idx in [c for c in self.evaluated_configs]:
                continue
            val = self.rm.metric_of(exp, self.metric) if not err else 0.0
            self.evaluated_configs.append(idx)
            self.evaluated_perf.append(val or 0.0)
        if len(self.evaluated_configs) >= INIT_NUM:
            self.cost_model.fit([self.features[i] for i in self.evaluated_configs], self.evaluated_perf)
