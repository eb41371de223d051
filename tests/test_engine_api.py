"""Engine API surface of the reference (runtime/engine.py): config accessors, coalesce_grad_reduction, caller-owned
gradient accumulation (managed_gradient_accumulation=false), no_sync, the plain data-parallel reduction helpers.
Reference tests: runtime/test_ds_initialize.py, test_no_sync_ctxt.py, v1/zero/test_zero_coalesce_grad_reduction.py."""
import pytest
import torch

from common import run_distributed
from simple_model import SimpleModel, random_batches

# every zero-argument method of the reference engine (deepspeed 0.19.6 runtime/engine.py)
REFERENCE_ACCESSORS = """active_timers aio_config amp_enabled amp_params autotp_size autotuning_enabled
autotuning_end_profile_step autotuning_metric autotuning_metric_path autotuning_model_info_path
autotuning_profile_model_info autotuning_start_profile_step bf16_master_weights_and_gradients bf16_optimizer_states
bfloat16_enabled checkpoint_serialization_enabled checkpoint_tag_validation_enabled checkpoint_tag_validation_fail
checkpoint_writer_enabled compile_autosp compile_autotp compile_zero_optimization_stage cpuadam_cores_perc
curriculum_enabled_legacy curriculum_params_legacy data_efficiency_enabled dataloader_drop_last dynamic_loss_scale
dynamic_loss_scale_args eigenvalue_enabled eigenvalue_gas_boundary_resolution eigenvalue_layer_name
eigenvalue_layer_num eigenvalue_max_iter eigenvalue_stability eigenvalue_tol eigenvalue_verbose elasticity_enabled
flops_profiler_detailed flops_profiler_enabled flops_profiler_module_depth flops_profiler_output_file
flops_profiler_profile_step flops_profiler_recompute_fwd_factor flops_profiler_top_modules fp16_auto_cast fp16_enabled
fp16_master_weights_and_gradients get_data_parallel_rank get_model_parallel_rank get_tensor_parallel_rank
get_pld_theta get_wall_clock_timers get_compile_time gradient_accumulation_steps gradient_clipping
gradient_predivide_factor graph_harvesting initial_dynamic_scale is_deepcompile_active is_deepcompile_enabled
is_elastic_model_parallel_supported is_first_weights_partition_group is_gradient_accumulation_boundary
is_sanity_checks_enabled load_universal_checkpoint log_level loss_scale managed_gradient_accumulation memory_breakdown
mics_shard_size optimizer_legacy_fusion optimizer_name optimizer_params pld_enabled pld_params postscale_gradients
quantize_training scheduler_name scheduler_params sparse_gradients_enabled steps_per_print super_offload
swap_tensor_config tensor_parallel_config tensorboard_enabled torch_autocast_dtype torch_autocast_enabled
torch_autocast_lower_precision_safe_modules train_batch_size train_micro_batch_size_per_gpu use_node_local_storage
uses_parallelization_pass_only wall_clock_breakdown zenflow_config zero_allgather_bucket_size
zero_allgather_partitions zero_allgather_sequential zero_allow_untested_optimizer zero_contiguous_gradients
zero_cpu_offload zero_elastic_checkpoint zero_force_ds_cpu_optimizer zero_gather_16bit_weights_on_model_save
zero_grad_hooks zero_hpz_partition_size zero_ignore_unused_parameters zero_legacy_stage1 zero_load_from_fp32_weights
zero_log_trace_cache_warnings zero_max_live_parameters zero_max_reuse_distance zero_model_persistence_threshold
zero_module_granularity_threshold zero_multi_rank_bucket_allreduce zero_nvme_offload_optimizer zero_offload_optimizer
zero_offload_param zero_optimization zero_optimization_partition_gradients zero_optimization_partition_weights
zero_optimization_stage zero_overlap_comm zero_param_persistence_threshold zero_partial_offload
zero_prefetch_bucket_size zero_quantized_gradients zero_quantized_nontrainable_weights zero_quantized_weights
zero_reduce_bucket_size zero_reduce_scatter zero_round_robin_gradients zero_save_muon_momentum_buffer_in_memory
zero_sub_group_size zero_use_cpu_optimizer zeropp_loco_param""".split()

REFERENCE_METHODS = """all_gather_scalar allreduce_and_copy allreduce_bucket allreduce_gradients allreduce_no_retain
backward buffered_allreduce_fallback clip_fp32_gradients coalesce_grad_reduction compile deepspeed_io destroy
dump_state empty_partition_cache eval forward get_global_grad_norm get_lr get_mom get_type load_checkpoint
load_moe_state_dict module_state_dict no_sync offload_states print_forward_breakdown register_compile_pass
reload_states save_16bit_model save_checkpoint save_fp16_model scale set_gradient_accumulation_boundary
set_train_batch_size set_train_micro_batch_size sparse_all_gather sparse_allreduce sparse_allreduce_bucket
sparse_allreduce_no_retain step train train_batch was_step_applied zero_grad is_map_style_dataset
is_iterable_style_dataset""".split()


def _engine(stage=1, extra=None, gas=1, seed=0):
    import deepspeed_b200 as ds
    torch.manual_seed(seed)
    model = SimpleModel(16, 3)
    cfg = {"train_micro_batch_size_per_gpu": 4, "gradient_accumulation_steps": gas,
           "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}},
           "zero_optimization": {"stage": stage, "unit_max_numel": 300, "stage3_param_persistence_threshold": 0}}
    cfg.update(extra or {})
    return ds.initialize(model=model, config=cfg)[0]


def _accessors():
    eng = _engine(1, {"fp16": {"enabled": False}, "torch_autocast": {"enabled": True, "dtype": "bfloat16"},
                      "eigenvalue": {"enabled": False, "max_iter": 7}})
    missing = [n for n in REFERENCE_ACCESSORS + REFERENCE_METHODS if not hasattr(eng, n)]
    failed = []
    for n in REFERENCE_ACCESSORS:
        if n in missing:
            continue
        try:
            getattr(eng, n)()
        except Exception as e:  # noqa: BLE001
            failed.append((n, repr(e)))
    facts = (eng.eigenvalue_max_iter(), eng.torch_autocast_enabled(), eng.torch_autocast_dtype(),
             eng.zero_use_cpu_optimizer(), eng.managed_gradient_accumulation(), eng.zero_partial_offload(),
             eng.communication_data_type, isinstance(eng.torch_autocast_lower_precision_safe_modules(), list))
    return missing, failed, facts


def test_reference_accessors_exist_and_answer():
    missing, failed, facts = run_distributed(_accessors, 1)[0]
    assert missing == [], missing
    assert failed == [], failed
    assert facts == (7, True, torch.bfloat16, False, True, 1.0, torch.float32, True)


def _train(stage, mode):
    """mode: 'gas' = managed accumulation over 3 micro-batches; 'coalesce' = the same inside coalesce_grad_reduction;
    'unmanaged' = managed_gradient_accumulation false, the caller calls step() after 3 backwards."""
    import deepspeed_b200 as ds
    extra = {"managed_gradient_accumulation": False} if mode == "unmanaged" else {}
    eng = _engine(stage, extra, gas=3)
    batches = random_batches(24, 4, 16, "cpu")
    rank = ds.comm.get_rank()
    losses = []
    for it in range(3):
        mb = [batches[(it * 3 + k) * 2 + rank] for k in range(3)]
        if mode == "gas":
            for x, y in mb:
                loss = eng(x, y)
                eng.backward(loss)
                eng.step()
                losses.append(float(loss.detach()))
        elif mode == "coalesce":
            with eng.coalesce_grad_reduction():
                for x, y in mb:
                    loss = eng(x, y)
                    eng.backward(loss)
                    losses.append(float(loss.detach()))
            eng.set_gradient_accumulation_boundary(True)
            eng.step()
            eng.set_gradient_accumulation_boundary(None)
        else:
            for x, y in mb:
                assert not eng.is_gradient_accumulation_boundary()
                loss = eng(x, y)
                eng.backward(loss)
                losses.append(float(loss.detach()))
            eng.step()
            assert eng.was_step_applied()
    from deepspeed_b200.utils import safe_get_full_fp32_param
    w = safe_get_full_fp32_param(next(iter(eng.module.parameters()))).clone()
    return losses, w, eng.global_steps, eng.global_samples


@pytest.mark.parametrize("stage", [1, 2, 3])
def test_coalesced_and_unmanaged_accumulation_match_managed(stage):
    ref = run_distributed(_train, 2, args=(stage, "gas"))[0]
    co = run_distributed(_train, 2, args=(stage, "coalesce"))[0]
    un = run_distributed(_train, 2, args=(stage, "unmanaged"))[0]
    for other in (co, un):
        assert other[0] == pytest.approx(ref[0], rel=1e-5, abs=1e-6)
        assert torch.allclose(other[1], ref[1], rtol=1e-5, atol=1e-6)
    assert ref[2] == 3 and un[2] == 3 and co[2] == 3
    assert un[3] == ref[3]  # samples counted from the backward calls actually made


def _no_sync(stage):
    import deepspeed_b200 as ds
    eng = _engine(stage, gas=2)
    if stage >= 2:
        try:
            with eng.no_sync():
                pass
        except AssertionError as e:
            return "incompatible" in str(e)
        return False
    batches = random_batches(8, 4, 16, "cpu")
    rank = ds.comm.get_rank()
    with eng.no_sync():
        loss = eng(*batches[rank])
        eng.backward(loss)
        try:
            eng.step()
            stepped_inside = True
        except AssertionError:
            stepped_inside = False
    return not stepped_inside


@pytest.mark.parametrize("stage", [1, 2])
def test_no_sync_context(stage):
    assert all(run_distributed(_no_sync, 2, args=(stage, )))


def _dp_helpers():
    import deepspeed_b200 as ds
    from deepspeed_b200.runtime.sparse_tensor import SparseTensor
    from deepspeed_b200.utils import groups
    eng = _engine(0)
    rank, world = ds.comm.get_rank(), ds.comm.get_world_size()
    g = groups.get_data_parallel_group()
    a, b = torch.full((3, ), float(rank + 1)), torch.full((2, 2), float(10 * (rank + 1)))
    eng.allreduce_no_retain([a, b], g, numel_per_bucket=2)
    mean = sum(range(1, world + 1)) / world
    ok = bool(torch.allclose(a, torch.full((3, ), mean)) and torch.allclose(b, torch.full((2, 2), 10 * mean)))
    for p in eng.module.parameters():
        p.grad = torch.full_like(p, float(rank))
    eng.buffered_allreduce_fallback()
    ok &= all(bool(torch.allclose(p.grad, torch.full_like(p, (world - 1) / 2))) for p in eng.module.parameters())
    dense = torch.zeros(6, 4)
    dense[rank] = 1.0 + rank
    dense[5] = 2.0
    sp = eng.sparse_allreduce(SparseTensor(dense), g)
    want = torch.zeros(6, 4)
    for r in range(world):
        want[r] += (1.0 + r) / world
        want[5] += 2.0 / world
    ok &= bool(torch.allclose(sp.to_dense(), want))
    vals = eng.all_gather_scalar(torch.tensor([rank]), g)
    ok &= [int(v) for v in vals] == list(range(world))
    return ok


def test_data_parallel_reduction_helpers():
    assert all(run_distributed(_dp_helpers, 2))


def _custom_pass():
    import deepspeed_b200 as ds
    calls = []

    def drop_prefetch(profile, plan):
        calls.append((len(profile.unit_ids), len(plan.launch_at)))
        plan.launch_at = {}
        return plan

    eng = _engine(3, {"compile": {"deepcompile": True}})
    eng.register_compile_pass("drop_prefetch", drop_prefetch)
    eng.compile(schedule=[(0, ["prefetch", "drop_prefetch"])])
    batches = random_batches(12, 4, 16, "cpu")
    for i in range(5):
        loss = eng(*batches[2 * i + ds.comm.get_rank()])
        eng.backward(loss)
        eng.step()
    t = eng.get_compile_time()
    return len(calls), eng._dc_driver.plan.launch_at, sorted(t), eng.is_deepcompile_active()


def test_custom_compile_pass_edits_the_unit_plan():
    n_calls, launch_at, keys, active = run_distributed(_custom_pass, 2)[0]
    assert n_calls == 1 and launch_at == {} and active
    assert keys == ["drop_prefetch", "plan", "profile"]


def _zero_compat(stage):
    import deepspeed_b200 as ds
    eng = _engine(stage, {"fp16": {"enabled": False}})
    z = eng.optimizer
    batches = random_batches(4, 4, 16, "cpu")
    loss = eng(*batches[ds.comm.get_rank()])
    eng.backward(loss)
    p = next(iter(eng.module.parameters()))
    ok = []
    ok.append(not z.has_overflow())
    g = z.get_fp32_grad_for_param(p)
    ok.append(g.shape == tuple(getattr(p, "ds_shape", p.shape)) and float(g.abs().sum()) > 0)
    z.update_fp32_grad_for_param_vectorized(lambda grad, param: grad * 0 + 2.0, [p])
    ok.append(bool((z.get_fp32_grad_for_param(p) == 2.0).all()))
    z.set_local_grad_for_param(torch.full_like(z.get_local_fp32_grad_for_param(p), float("inf")), p)
    ok.append(z.has_overflow() or z.get_local_fp32_grad_for_param(p).numel() == 0 and z.has_overflow())
    z.check_overflow()
    ok.append(z.overflow)
    z.set_local_grad_for_param(torch.zeros_like(z.get_local_fp32_grad_for_param(p)), p)
    eng.step()
    z.set_lr(0.5)
    ok.append(z.get_lr() == 0.5 and eng.get_lr()[0] == 0.5)
    z.override_loss_scale(128.0)
    ok.append(z.cur_scale == 128.0 and z.custom_loss_scaler)
    ok.append(z.get_hp_param_device(p).type == "cpu" and isinstance(z.get_param_id(p), tuple))
    ok.append("exp_avg" in next(iter(z.state.values())))
    lean = z.get_lean_optimizer_state()
    ok.append("fp32" in lean and "exp_avg" in lean and lean["fp32"][0].dim() == 1)
    w0 = z.get_local_fp32_param(p).clone()
    z.set_local_hp_param(w0 + 1.0, p)
    ok.append(bool(torch.equal(z.get_local_fp32_param(p), w0 + 1.0)))
    z.update_lp_params()
    from deepspeed_b200.utils import safe_get_full_fp32_param
    full = safe_get_full_fp32_param(p)
    z.refresh_fp32_params()
    ok.append(bool(torch.allclose(safe_get_full_fp32_param(p), full)))
    z.empty_partition_cache()
    z.finalize_gradient_accumulation_boundary()
    ok.append(z.is_gradient_accumulation_boundary)
    return ok


@pytest.mark.parametrize("stage", [2, 3])
def test_zero_optimizer_reference_methods(stage):
    for ok in run_distributed(_zero_compat, 2, args=(stage, )):
        assert all(ok), [i for i, v in enumerate(ok) if not v]


def test_config_dictionary_getters():
    """``get_<field>(param_dict)`` readers with the reference's defaults (runtime/config.py and friends)."""
    import torch
    from deepspeed_b200.config import config as C
    d = {"train_batch_size": 32, "gradient_clipping": 1.0,
         "optimizer": {"type": "AdamW", "params": {"lr": 1e-3, "max_grad_norm": 2.0}},
         "scheduler": {"type": "WarmupLR", "params": {"warmup_num_steps": 10}},
         "sparse_attention": {"mode": "bigbird", "block": 32}, "fp16": {"enabled": True},
         "checkpoint": {"tag_validation": "fail", "parallel_write": {"pipeline_stage": True}},
         "data_efficiency": {"data_sampling": {"enabled": True, "curriculum_learning": {"enabled": True, "x": 1}},
                             "data_routing": {"enabled": True, "random_ltd": {"enabled": True, "layers": 2}}},
         "progressive_layer_drop": {"enabled": True, "theta": 0.5}, "communication_data_type": "bf16",
         "tensorboard": {"enabled": True}, "amp": {"enabled": True, "opt_level": "O1"},
         "torch_autocast": {"enabled": True, "dtype": "bfloat16"}}
    assert (C.get_train_batch_size(d), C.get_gradient_clipping(d), C.get_optimizer_name(d)) == (32, 1.0, "adamw")
    assert C.get_optimizer_gradient_clipping(d) == 2.0 and C.get_scheduler_params(d) == {"warmup_num_steps": 10}
    sa = C.get_sparse_attention(d)
    assert sa["mode"] == "bigbird" and sa["block"] == 32 and sa["num_sliding_window_blocks"] == 3
    assert C.get_sparse_attention({}) is None and C.get_sparse_fixed_config({})["num_local_blocks"] == 4
    assert C.get_float16_config(d).enabled and not C.get_bfloat16_config(d).enabled
    ck = C.get_checkpoint_params(d)
    assert C.get_checkpoint_tag_validation_mode(ck) == "FAIL" and C.get_checkpoint_parallel_write_pipeline(ck) is True
    assert C.get_curriculum_learning_enabled(d) and C.get_curriculum_learning_params(d) == {"x": 1}
    assert C.get_random_ltd_enabled(d) and C.get_random_ltd_params(d) == {"layers": 2}
    assert C.get_pld_enabled(d) and C.get_pld_params(d) == {"theta": 0.5} and C.get_pld_params({}) is False
    assert C.get_communication_data_type(d) is torch.bfloat16 and C.get_communication_data_type({}) is None
    assert C.get_amp_enabled(d) and C.get_amp_params(d) == {"opt_level": "O1"}
    assert C.get_torch_autocast_enabled(d) and C.get_torch_autocast_dtype(d) is torch.bfloat16
    assert C.get_eigenvalue_config(d)[2] == 100 and C.get_pipeline_config(d)["partition"] == "best"
    assert C.get_monitor_config(d).tensorboard.enabled and C.get_hybrid_engine_config(d).max_out_tokens == 512
    assert C.get_gradient_allreduce_op({}) == "avg" and C.get_zero_force_ds_cpu_optimizer({}) is True
    import pytest
    with pytest.raises(ValueError):
        C.get_gradient_allreduce_op({"gradient_allreduce_op": "max"})
    with pytest.raises(ValueError):
        C.get_checkpoint_tag_validation_mode({"tag_validation": "maybe"})
    with pytest.raises(NotImplementedError):
        C.get_sparse_attention({"sparse_attention": {"mode": "hexagonal"}})
