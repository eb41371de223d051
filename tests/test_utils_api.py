"""``deepspeed.utils`` surface (reference utils/__init__.py): leaf modules, OnDevice, bwc helpers, static KV cache,
native pinned memory, debug names, tensor-fragment views."""
import pytest
import torch

from common import run_distributed
from simple_model import SimpleModel, random_batches


def test_public_names_import():
    import deepspeed_b200.utils as u
    names = """logger log_dist log_dist_once set_log_level_from_string get_caller_func OnDevice instrument_w_nvtx
    get_full_hp_param get_hp_fragment_mapping fragment_address get_full_hp_grad map_to_flat_opt_states
    safe_get_full_fp32_param safe_get_full_grad safe_get_full_optimizer_state set_full_hp_param set_full_hp_grad
    safe_set_full_fp32_param safe_set_full_optimizer_state safe_set_full_grad safe_get_local_fp32_param
    safe_get_local_grad safe_get_local_optimizer_state safe_set_local_fp32_param safe_set_local_grad
    safe_set_local_optimizer_state safe_update_full_grad_vectorized set_z3_leaf_modules unset_z3_leaf_modules
    get_z3_leaf_modules z3_leaf_module z3_leaf_parameter set_z3_leaf_module set_z3_leaf_modules_by_name
    set_z3_leaf_modules_by_suffix link_hp_params lazy_init_hp_params_optimizer_state RepeatingLoader get_numactl_cmd
    get_data_parallel_group get_tensor_model_parallel_rank""".split()
    missing = [n for n in names if not hasattr(u, n)]
    assert missing == []
    for mod in ("bwc", "torch", "types", "exceptions", "debug", "comms_logging", "pin_memory", "pin_memory_tracker",
                "static_cache", "config", "zero_to_fp32", "timer", "numa", "nvtx"):
        __import__(f"deepspeed_b200.utils.{mod}")
    with pytest.raises(ValueError):
        u.set_log_level_from_string("chatty")


def test_on_device_meta_and_dtype():
    from deepspeed_b200.utils import OnDevice
    with OnDevice(dtype=torch.bfloat16, device="meta"):
        m = torch.nn.Linear(8, 8)
        idx = torch.zeros(3, dtype=torch.long)
    assert m.weight.is_meta and m.weight.dtype == torch.bfloat16 and idx.dtype == torch.long
    assert torch.get_default_dtype() == torch.float32
    with OnDevice(dtype=torch.float16, device="cpu", enabled=False):
        assert torch.nn.Linear(2, 2).weight.dtype == torch.float32


def test_bwc_prefers_newest_names():
    from deepspeed_b200.utils import bwc

    class New:

        def get_tensor_model_parallel_rank(self):
            return 3

        def get_model_parallel_rank(self):
            return 9

        def get_pipe_parallel_world_size(self):
            return 4

    class Old:

        def get_model_parallel_rank(self):
            return 9

        def get_model_parallel_world_size(self):
            return 2

    assert bwc.bwc_tensor_model_parallel_rank(None) == 0 and bwc.bwc_tensor_model_parallel_world_size(None) == 1
    assert bwc.bwc_tensor_model_parallel_rank(New()) == 3 and bwc.bwc_tensor_model_parallel_rank(Old()) == 9
    assert bwc.bwc_tensor_model_parallel_world_size(Old()) == 2
    assert bwc.bwc_pipeline_parallel_world_size(New()) == 4 and bwc.bwc_pipeline_parallel_world_size(Old()) == 1
    assert bwc.bwc_pipeline_parallel_group(None) is None


def test_static_cache_writes_at_external_position():
    from deepspeed_b200.utils.static_cache import DeepSpeedStaticCache

    class Cfg:
        num_hidden_layers, num_attention_heads, num_key_value_heads, hidden_size = 2, 4, 2, 32

    cache = DeepSpeedStaticCache(Cfg(), batch_size=1, max_cache_len=16, device="cpu", dtype=torch.float32)
    assert len(cache) == 2 and cache.layers[0].keys.shape == (1, 2, 16, 8)
    pos = torch.zeros((), dtype=torch.long)
    cache.set_write_position(pos)
    k0 = torch.arange(1 * 2 * 5 * 8, dtype=torch.float32).view(1, 2, 5, 8)
    keys, vals = cache.update(k0, -k0, 0)
    pos.fill_(5)  # advanced in place by the caller, as before a graph replay
    k1 = torch.full((1, 2, 1, 8), 7.0)
    keys, vals = cache.update(k1, k1, 0)
    assert torch.equal(keys[:, :, :5], k0) and torch.equal(keys[:, :, 5:6], k1) and float(keys[:, :, 6:].abs().sum()) == 0
    assert torch.equal(vals[:, :, :5], -k0)
    assert int(cache.get_seq_length()) == 6 and cache.get_mask_sizes(1) == (16, 0) and cache.get_max_cache_shape() == 16
    base = cache.layers[1].keys.data_ptr()
    cache.reset()
    assert float(cache.layers[0].keys.abs().sum()) == 0 and cache.layers[1].keys.data_ptr() == base
    with pytest.raises(IndexError):
        cache.update(k1, k1, 2)


def test_native_pinned_memory_roundtrip(monkeypatch):
    from deepspeed_b200.utils import pin_memory as pm
    assert pm.get_active_native_pinned_memory() is None
    monkeypatch.setenv("DS_PIN_MEMORY_BACKEND", "native")
    mgr = pm.get_active_native_pinned_memory()
    src = torch.randn(3, 1000)
    t = mgr.pin(src)
    assert t.shape == src.shape and torch.equal(t, src) and mgr.is_pinned(t) and mgr.is_pinned(t[1])
    assert t.data_ptr() % 4096 == 0
    assert not mgr.is_pinned(src)
    flat = mgr.pin(src, make_copy=False, match_shape=False)
    assert flat.shape == (3000, )
    assert mgr.unpin(t) and not mgr.unpin(t)
    n_before = len(mgr._ranges)
    del flat
    import gc
    gc.collect()
    assert len(mgr._ranges) == n_before - 1  # freed by the finalizer


class _Block(torch.nn.Module):

    def __init__(self):
        super().__init__()
        self.a = torch.nn.Linear(16, 16)
        self.b = torch.nn.Linear(16, 16)

    def forward(self, x):
        return self.b(torch.relu(self.a(x)))


class _Net(torch.nn.Module):

    def __init__(self):
        super().__init__()
        self.blocks = torch.nn.ModuleList([_Block() for _ in range(3)])
        self.head = torch.nn.Linear(16, 16)

    def forward(self, x, y):
        for b in self.blocks:
            x = b(x)
        return torch.nn.functional.mse_loss(self.head(x), y)


def _leaf(kind):
    import deepspeed_b200 as ds
    from deepspeed_b200 import utils as U
    torch.manual_seed(0)
    net = _Net()
    zo = {"stage": 3, "unit_max_numel": 100, "stage3_param_persistence_threshold": 0}
    if kind == "class":
        assert len(U.set_z3_leaf_modules(net, [_Block])) == 3
        assert len(U.unset_z3_leaf_modules(net, ["_Block"])) == 3 and U.get_z3_leaf_modules(net) == []
        U.set_z3_leaf_modules(net, ["_Block"])
    elif kind == "name":
        mods, missing = U.set_z3_leaf_modules_by_name(net, ["blocks.1", "nope"], raise_if_not_found=False)
        assert len(mods) == 1 and missing == ["nope"]
    elif kind == "suffix":
        U.set_z3_leaf_modules_by_suffix(net, ["blocks.2"])
    elif kind == "config":
        zo["leaf_module"] = {"name_suffixes": "blocks.0"}
    eng, _, _, _ = ds.initialize(model=net, config={"train_micro_batch_size_per_gpu": 2,
                                                    "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}},
                                                    "zero_optimization": zo})
    units = [u for u in eng.optimizer.units if not u.persistent]
    per_block = []
    for b in net.blocks:
        per_block.append(len({p.ds_unit.uid for p in b.parameters()}))
    leaf_params = [U.z3_leaf_parameter(p) for p in net.blocks[1].parameters()]
    x = torch.randn(2, 16)
    loss = eng(x, x)
    eng.backward(loss)
    eng.step()
    return per_block, len(units), all(leaf_params), any(leaf_params)


@pytest.mark.parametrize("kind,expect", [("none", [2, 2, 2]), ("class", [1, 1, 1]), ("name", [2, 1, 2]),
                                         ("suffix", [2, 2, 1]), ("config", [1, 2, 2])])
def test_leaf_modules_become_single_units(kind, expect):
    per_block, n_units, all_leaf, any_leaf = run_distributed(_leaf, 2, args=(kind, ))[0]
    assert per_block == expect
    assert all_leaf == (expect[1] == 1) and any_leaf == (expect[1] == 1)


def _fragments():
    import deepspeed_b200 as ds
    from deepspeed_b200 import utils as U
    from deepspeed_b200.utils import debug
    from deepspeed_b200.utils.tensor_fragment import tensor_fragment
    torch.manual_seed(0)
    model = SimpleModel(16, 2)
    eng, _, _, _ = ds.initialize(model=model, config={"train_micro_batch_size_per_gpu": 2,
                                                      "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}},
                                                      "zero_optimization": {"stage": 2}})
    batches = random_batches(2, 2, 16, "cpu")
    loss = eng(*batches[ds.comm.get_rank()])
    eng.backward(loss)
    eng.step()
    p = next(iter(model.parameters()))
    frag = tensor_fragment(p)
    local = U.safe_get_local_fp32_param(p)
    ok = []
    ok.append(frag.get_hp_fragment().numel() == local.numel() == frag.get_hp_fragment_address().numel)
    ok.append("exp_avg" in frag.get_optim_state_keys())
    ok.append(frag.get_optim_state_fragment("exp_avg").numel() == local.numel())
    U.safe_set_local_optimizer_state(p, torch.full_like(local, 0.5), "exp_avg")
    ok.append(bool((U.safe_get_local_optimizer_state(p, "exp_avg") == 0.5).all()))
    full = U.safe_get_full_optimizer_state(p, "exp_avg")
    ok.append(bool((full == 0.5).all()) and full.shape == p.shape)
    U.link_hp_params(list(model.parameters()))
    ok.append(torch.equal(p.get_full_hp_param(), U.safe_get_full_fp32_param(p)))
    debug.debug_extract_module_and_param_names(model)
    ok.append(debug.debug_param2name(p) != "unknown" and "shape=" in debug.debug_param2name_id_shape(p))
    debug.debug_clear_module_and_param_names()
    return ok


def test_tensor_fragment_views_and_linkage():
    res = run_distributed(_fragments, 2)
    assert all(all(r) for r in res), res


def test_small_reference_helpers():
    import pytest
    import torch
    from deepspeed_b200.launcher.launcher_helper import env_mapping
    from deepspeed_b200.runtime.zero import utils as Z
    from deepspeed_b200.utils.numa import check_for_numactl_pkg, parse_range
    env = env_mapping({"PMI_RANK": "3", "OMPI_COMM_WORLD_RANK": "3", "MPI_LOCALRANKID": "1"},
                      ["PMI_RANK", "OMPI_COMM_WORLD_RANK"], ["MPI_LOCALRANKID", "OMPI_COMM_WORLD_LOCAL_RANK"])
    assert env["RANK"] == "3" and env["LOCAL_RANK"] == "1"
    with pytest.raises(EnvironmentError):
        env_mapping({"A": "1", "B": "2", "L": "0"}, ["A", "B"], ["L"])
    with pytest.raises(EnvironmentError):
        env_mapping({"A": "1"}, ["A"], ["L"])
    assert list(parse_range("4")) == [4] and list(parse_range("2-5")) == [2, 3, 4, 5]
    for bad in ("5-2", "1-2-3"):
        with pytest.raises(ValueError):
            parse_range(bad)
    assert isinstance(check_for_numactl_pkg(), bool)
    ts = [torch.zeros(3, dtype=torch.float16), torch.zeros(2), torch.zeros(1, dtype=torch.bfloat16), torch.zeros(4)]
    b = Z.split_half_float_double(ts)
    assert [x[0].dtype for x in b] == [torch.float16, torch.float32, torch.bfloat16] and len(b[1]) == 2
    assert [dt for dt, _ in Z.split_half_float_double_sparse(ts)] == [torch.float16, torch.float32, torch.bfloat16]
    assert Z.get_alignment_padding(ts, 8) == 6 and Z.get_alignment_padding(ts[:1] + ts[:1] + ts[1:2], 8) == 0
    assert Z.lcm(4, 6) == 12 and Z.isclose(1.0, 1.0 + 1e-12) and not Z.isclose(1.0, 1.1)
    mk = Z.zero_wrapper_for_fp_tensor_constructor(torch.ones, torch.bfloat16)
    assert mk(2).dtype == torch.bfloat16 and mk(2, dtype=torch.int32).dtype == torch.int32
    p = torch.nn.Parameter(torch.zeros(2))
    assert Z.get_allgather_dtype(p, torch.zeros(1, dtype=torch.float16)) == torch.float16
    p.autocast_dtype = torch.bfloat16
    assert Z.get_allgather_dtype(p, torch.zeros(1, dtype=torch.float16)) == torch.bfloat16
