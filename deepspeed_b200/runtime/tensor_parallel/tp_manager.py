"""``TpTrainingManager``: shards a model for tensor-parallel training.

Reference: ``runtime/tensor_parallel/tp_manager.py:13``. Order of business: build the TP group for ``tp_size``, make
the lanes of a replica draw the same random numbers (dropout masks on replicated activations must agree: the RNG state
of the group's first rank is broadcast), then replace the linears / embeddings (``parallel/tp.py::AutoTP``)."""
import torch

from ... import comm as dist
from ...config.sections import TPConfig, TPTrainingConfig
from ...utils import groups


class TpTrainingManager:

    def __init__(self, model, tp_size, dtype, partition_config=None):
        from ...parallel.tp import apply_autotp
        self.module = model
        self.config = TPTrainingConfig(autotp_size=tp_size, dtype=dtype)
        if partition_config is not None:
            self.config.partition_config = partition_config
        self.model_config = getattr(model, "config", None)
        if not dist.is_initialized():
            dist.init_distributed()
        groups.initialize(tp_size=tp_size, force=True)
        self.tp_config = TPConfig(tp_size=tp_size)
        self.tp_config.tp_group = groups._get_model_parallel_group()
        self.config.tensor_parallel = self.tp_config
        self.sync_rng_state()
        apply_autotp(model, self.config, self.tp_config.tp_group)
        model.ds_autotp_parsed = True

    def sync_rng_state(self):
        group = self.tp_config.tp_group
        if group is None or dist.get_world_size(group=group) == 1:
            return
        src = groups.get_tensor_model_parallel_src_rank()
        cuda = torch.cuda.is_available()
        state = torch.cuda.get_rng_state() if cuda else torch.get_rng_state()
        buf = state.to(torch.device("cuda", torch.cuda.current_device())) if cuda else state.clone()
        dist.broadcast(buf, src, group=group)
        if cuda:
            torch.cuda.set_rng_state(buf.cpu())
        else:
            torch.set_rng_state(buf)
