import random

from .base_tuner import BaseTuner


class RandomTuner(BaseTuner):
    """Explore the space in random order."""

    def __init__(self, exps, resource_manager, metric):
        super().__init__(exps, resource_manager, metric)

    def next_batch(self, sample_size=1):
        sample_size = min(sample_size, len(self.all_exps))
        batch = random.sample(self.all_exps, sample_size)
        self.all_exps = [e for e in self.all_exps if e not in batch]
        return batch


class GridSearchTuner(BaseTuner):
    """Explore the space in order."""

    def __init__(self, exps, resource_manager, metric):
        super().__init__(exps, resource_manager, metric)

    def next_batch(self, sample_size=1):
        sample_size = min(sample_size, len(self.all_exps))
        batch, self.all_exps = self.all_exps[:sample_size], self.all_exps[sample_size:]
        return batch
