"""Random-LTD kept-sequence-length schedule + consumed layer-token accounting (reference
``runtime/data_pipeline/data_routing/scheduler.py:38``)."""
import math


class BaseScheduler:

    def __init__(self):
        self.state = {}

    def _fixed_root(self, global_steps, root_degree=None):
        sc = self.state["schedule_config"]
        root_degree = sc["root_degree"] if root_degree is None else root_degree
        lo, hi = self.state["min_value"], self.state["max_value"]
        nxt = math.floor((float(global_steps) / sc["require_steps"])**(1.0 / root_degree) * (hi - lo) + lo)
        nxt -= nxt % sc["seq_per_step"]
        return min(nxt, hi)

    def get_value(self, global_steps):
        if self.state["schedule_type"] == "fixed_linear":
            return self._fixed_root(global_steps, 1)
        raise RuntimeError("Unsupported random LTD schedule type")


class RandomLTDScheduler(BaseScheduler):

    def __init__(self, config):
        super().__init__()
        self.model_layer_num = config["total_layer_num"]
        self.random_ltd_layer_num = config["random_ltd_layer_num"]
        self.config_schedule = config["random_ltd_schedule"]
        self.global_batch_size = config["global_batch_size"]
        self.reset_to_init()
        if config.get("layer_token_lr_schedule", {}).get("enabled", False):
            raise NotImplementedError("layer_token_lr_schedule is not implemented (nor in the reference)")
        self.state["consumed_layer_tokens"] = 0

    def get_total_layer_tokens(self, train_iters):
        for step in range(train_iters):
            self.update_seq(step)
        return self.state["consumed_layer_tokens"]

    def reset_to_init(self):
        cs = self.config_schedule
        if cs is not None:
            self.state.update(min_value=cs["min_value"], max_value=cs["max_value"], current_value=cs["min_value"],
                              schedule_config=cs["schedule_config"], schedule_type=cs["schedule_type"])
        self.state["consumed_layer_tokens"] = 0
        self.state["current_steps"] = -1

    def get_current_seq(self):
        return self.state["current_value"]

    def set_current_seq(self, seq_length):
        self.state["current_value"] = seq_length

    def get_random_ltd_layer_num(self):
        return self.random_ltd_layer_num

    def get_state(self):
        return self.state

    def set_state(self, state):
        self.state = state

    def update_seq(self, global_steps):
        st = self.state
        if st["current_value"] < st["max_value"]:
            st["current_value"] = self.get_value(global_steps)
        if global_steps != st["current_steps"]:
            st["consumed_layer_tokens"] += self.global_batch_size * (
                st["current_value"] * self.random_ltd_layer_num + st["max_value"] *
                (self.model_layer_num - self.random_ltd_layer_num))
            st["current_steps"] = global_steps

    _SAVED = ("consumed_layer_tokens", "current_steps", "current_value", "min_value", "max_value")

    def state_dict(self):
        return {k: self.state[k] for k in self._SAVED}

    def load_state_dict(self, state_dict):
        for k in self._SAVED:
            self.state[k] = state_dict[k]
