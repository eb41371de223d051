"""Reference ``utils/mixed_precision_linkage.py``: attaches ``_hp_mapping`` (the fp32 fragment view) and the
``get_full_hp_param`` family of bound methods to every low-precision parameter. The unit bookkeeping of
``ZeroShardedOptimizer`` owns that mapping here; these functions attach the same attribute names for code that expects
them."""
import types

from . import tensor_fragment as tf


def link_hp_params(lp_param_list, flat_hp_partition=None, gradient_dict=None, offload_gradient_dict=None,
                   use_offload=False, param_group_index=0, partition_start=0, partition_size=0, dp_group=None):
    for lp in lp_param_list:
        if tf._zopt(lp) is None:
            lp._hp_mapping = None
            continue
        lp._hp_mapping = tf.tensor_fragment(lp)
        lp._dp_group = dp_group
        lp.get_full_hp_param = types.MethodType(tf.get_full_hp_param, lp)
        lp.get_full_hp_grad = types.MethodType(tf.get_full_hp_grad, lp)
        lp.set_full_hp_param = types.MethodType(tf.set_full_hp_param, lp)
        lp.set_full_hp_grad = types.MethodType(tf.set_full_hp_grad, lp)


def lazy_init_hp_params_optimizer_state(lp_param_list, flat_hp_partition=None, optimizer_state=None):
    """Optimizer state fragments are read straight from the sharded optimizer on access: nothing to initialise."""
    return None
