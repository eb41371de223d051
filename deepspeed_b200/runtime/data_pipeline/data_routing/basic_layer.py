"""``RandomLayerTokenDrop``: wrap a transformer layer so that it only processes a random sorted subset of the tokens
(reference ``runtime/data_pipeline/data_routing/basic_layer.py:13``). Layer 0 draws the indices for all wrapped
layers of the step and parks them in the scheduler state."""
from torch import Tensor
from torch.nn import Module

from ....ops.random_ltd.dropping_utils import GatherTokens, ScatterTokens, bert_sample_tokens, gpt_sample_tokens


class RandomLayerTokenDrop(Module):

    def __init__(self, layer):
        super().__init__()
        self.random_ltd_layer = layer
        self.random_ltd_scheduler = None
        self.max_length = None
        self.reserved_length = -1
        self.curr_seq = -1
        self.batch_first = False

    def init_config(self, config, scheduler, random_ltd_layer_id):
        self.random_ltd_scheduler = scheduler
        self.random_ltd_layer_id = random_ltd_layer_id
        self.max_length = scheduler.state["max_value"]
        self.mask_name = config["model_mask_name"]
        self.micro_bs = config["micro_batch_size"]
        self.random_ltd_num_layer = scheduler.random_ltd_layer_num
        order = config["hidden_state_order"]
        if order not in ("batch_seq_dim", "seq_batch_dim"):
            raise NotImplementedError("hidden_state_order must be batch_seq_dim or seq_batch_dim")
        self.batch_first = order == "batch_seq_dim"
        self.model_type = config["model_type"]
        if self.model_type == "encoder":
            self.index_generator = bert_sample_tokens
        elif self.model_type == "decoder":
            self.index_generator = gpt_sample_tokens
        else:
            raise NotImplementedError("model_type must be encoder or decoder")

    def _shape(self, h):
        if self.batch_first:
            self.curr_micro_batch, self.curr_seq = h.size(0), h.size(1)
        else:
            self.curr_seq, self.curr_micro_batch = h.size(0), h.size(1)

    def forward(self, hidden_states, **kwargs) -> Tensor:
        sch = self.random_ltd_scheduler
        if sch is not None:
            self.reserved_length = sch.get_current_seq()
            self._shape(hidden_states)
        if not (self.training and sch is not None and self.reserved_length < self.curr_seq):
            return self.random_ltd_layer(hidden_states, **kwargs)
        mask = kwargs[self.mask_name] if self.mask_name is not None else None
        if self.random_ltd_layer_id == 0:
            sampled, part_mask = self.index_generator(self.reserved_length, self.curr_seq, self.curr_micro_batch,
                                                      self.random_ltd_num_layer, hidden_states.device, mask)
            sch.state["sample_idx"], sch.state["attention_mask"] = sampled, part_mask
        else:
            sampled, part_mask = sch.state["sample_idx"], sch.state["attention_mask"]
        mine = sampled[self.random_ltd_layer_id]
        hidden_states, part = GatherTokens.apply(hidden_states, mine, self.batch_first)
        if self.mask_name is not None:
            kwargs[self.mask_name] = part_mask[self.random_ltd_layer_id] if self.model_type == "encoder" else part_mask
        outputs = self.random_ltd_layer(part, **kwargs)
        if isinstance(outputs, tuple):
            merged = ScatterTokens.apply(hidden_states, outputs[0], mine, self.batch_first)
            return (merged, ) + tuple(outputs[1:])
        if isinstance(outputs, Tensor):
            return ScatterTokens.apply(hidden_states, outputs, mine, self.batch_first)
        raise NotImplementedError("the wrapped layer must return a tensor or a tuple whose first item is one")
