"""Memory estimators for model states under ZeRO-2 / ZeRO-3.

Reference: ``estimate_zero2_model_states_mem_needs*`` (``runtime/zero/stage_1_and_2.py:3081``) and
``estimate_zero3_model_states_mem_needs*`` (``runtime/zero/stage3.py:3744``). The numbers follow from the bytes each
parameter costs in mixed-precision Adam training:

    ``P16 = 2`` half-precision weight, ``G16 = 2`` half-precision gradient, and ``FP32_STATE = 16`` for the fp32
    gradient (4), fp32 master weight (4) and the two Adam moments (8) - ``G16 + FP32_STATE = 18`` partitionable bytes.

What is partitioned / offloaded decides on which side a term lands; host numbers carry ``additional_buffer_factor`` for
pinned staging buffers. ``zero_init`` means the model is built under ``zero.Init`` (never whole in host memory)."""
P16 = 2
G16 = 2
FP32_STATE = 16  # fp32 gradient + fp32 master + exp_avg + exp_avg_sq
FP32 = 4


def _unique_param_count(model):
    return sum({p.data_ptr(): p.numel() for p in model.parameters()}.values())


def model_to_params(model, with_largest_layer=False):
    """Total parameter count (shared tensors once); with ``with_largest_layer`` also the largest single module's own
    parameters (what ZeRO-3 must be able to hold gathered at once)."""
    total = _unique_param_count(model)
    if not with_largest_layer:
        return total
    largest = max((sum(p.numel() for p in m.parameters(recurse=False)) for m in model.modules()), default=0)
    return total, largest


def estimate_zero2_model_states_mem_needs(total_params, num_gpus_per_node=1, num_nodes=1, cpu_offload=True,
                                          additional_buffer_factor=1.5):
    """(host bytes per node, device bytes per GPU)."""
    gpus = num_nodes * num_gpus_per_node
    if cpu_offload:  # weights stay whole on the device, every optimizer-side byte lives on the host
        gpu = P16 * total_params
        cpu = total_params * max(FP32 * gpus, FP32_STATE) * additional_buffer_factor
    else:
        gpu = P16 * total_params + int((G16 + FP32_STATE) * total_params / gpus)
        cpu = total_params * FP32 * num_gpus_per_node * additional_buffer_factor  # fp32 copy per rank while loading
    return int(cpu), int(gpu)


def estimate_zero3_model_states_mem_needs(total_params, largest_layer_params, num_gpus_per_node=1, num_nodes=1,
                                          cpu_offload=True, cpu_offload_params=True, zero_init=True,
                                          additional_buffer_factor=1.5):
    """(host bytes per node, device bytes per GPU, bytes of the largest gathered layer)."""
    gpus = num_nodes * num_gpus_per_node
    node_share = 1 / num_nodes
    largest = FP32 * largest_layer_params
    load_whole = FP32 * num_gpus_per_node  # without zero.Init every rank of a node first holds the fp32 model
    if cpu_offload and cpu_offload_params:
        gpu = largest
        per_param = (P16 + FP32_STATE) * node_share
        cpu = total_params * (per_param if zero_init else max(load_whole, per_param)) * additional_buffer_factor
    elif cpu_offload:
        gpu = largest + int(P16 * total_params / gpus)
        per_param = FP32_STATE * node_share
        cpu = total_params * (per_param if zero_init else max(load_whole, per_param)) * additional_buffer_factor
    else:
        gpu = largest + int((P16 + FP32_STATE) * total_params / gpus)
        cpu = (largest_layer_params if zero_init else total_params) * load_whole * additional_buffer_factor
    return int(cpu), int(gpu), largest


def _gb(n):
    return f"{n / 2**30:7.2f}GB"


def _header(kind, total_params, num_gpus_per_node, num_nodes, largest=None):
    sw = f"SW: Model with {int(total_params / 1e6)}M total params"
    if largest is not None:
        sw += f", {int(largest / 1e6)}M largest layer params"
    print(f"Estimated memory needed for params, optim states and gradients for a:\n"
          f"HW: Setup with {num_nodes} node{'s' if num_nodes > 1 else ''}, {num_gpus_per_node} "
          f"GPU{'s' if num_gpus_per_node > 1 else ''} per node.\n{sw}.")
    print("  per CPU  |  per GPU |   Options")


def estimate_zero2_model_states_mem_needs_all_cold(total_params, num_gpus_per_node=1, num_nodes=1,
                                                   additional_buffer_factor=1.5):
    """Print the table for a hypothetical model of ``total_params`` parameters; returns the rows."""
    _header("zero2", total_params, num_gpus_per_node, num_nodes)
    rows = []
    for offload in (True, False):
        cpu, gpu = estimate_zero2_model_states_mem_needs(total_params, num_gpus_per_node, num_nodes, offload,
                                                         additional_buffer_factor)
        opts = f"offload_optimizer={'cpu ' if offload else 'none'}"
        print(f" {_gb(cpu)} | {_gb(gpu)} | {opts}")
        rows.append((cpu, gpu, opts))
    return rows


def estimate_zero2_model_states_mem_needs_all_live(model, num_gpus_per_node=1, num_nodes=1,
                                                   additional_buffer_factor=1.5):
    return estimate_zero2_model_states_mem_needs_all_cold(model_to_params(model), num_gpus_per_node, num_nodes,
                                                          additional_buffer_factor)


def estimate_zero3_model_states_mem_needs_all_cold(total_params, largest_layer_params, num_gpus_per_node=1,
                                                   num_nodes=1, additional_buffer_factor=1.5):
    _header("zero3", total_params, num_gpus_per_node, num_nodes, largest_layer_params)
    rows = []
    for offload in (True, False):
        for offload_params in (True, False):
            if offload_params and not offload:
                continue
            for zero_init in (True, False):
                cpu, gpu, _ = estimate_zero3_model_states_mem_needs(total_params, largest_layer_params,
                                                                    num_gpus_per_node, num_nodes, offload,
                                                                    offload_params, zero_init,
                                                                    additional_buffer_factor)
                opts = (f"offload_param={'cpu ' if offload_params else 'none'}, "
                        f"offload_optimizer={'cpu ' if offload else 'none'}, zero_init={int(zero_init)}")
                print(f" {_gb(cpu)} | {_gb(gpu)} | {opts}")
                rows.append((cpu, gpu, opts))
    return rows


def estimate_zero3_model_states_mem_needs_all_live(model, num_gpus_per_node=1, num_nodes=1,
                                                   additional_buffer_factor=1.5):
    total, largest = model_to_params(model, with_largest_layer=True)
    return estimate_zero3_model_states_mem_needs_all_cold(total, largest, num_gpus_per_node, num_nodes,
                                                          additional_buffer_factor)
