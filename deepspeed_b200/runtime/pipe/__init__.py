"""Pipeline parallelism: PipelineModule / LayerSpec / 1F1B schedule / NCCL p2p."""
from .module import PipelineModule, LayerSpec, TiedLayerSpec
from .topology import ProcessTopology, PipeDataParallelTopology, PipeModelDataParallelTopology, PipelineParallelGrid
