from .indexed_dataset import MMapIndexedDataset, MMapIndexedDatasetBuilder  # noqa: F401
from .data_sampler import DeepSpeedDataSampler  # noqa: F401
