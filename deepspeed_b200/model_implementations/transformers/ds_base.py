import torch.nn as nn


class DeepSpeedTransformerBase(nn.Module):
    """Placeholder base of a future layer hierarchy in the reference; kept for imports."""

    def __init__(self):
        super().__init__()
