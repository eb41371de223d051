from . import flops_profiler
from ..config.sections import DeepSpeedFlopsProfilerConfig
