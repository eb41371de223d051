"""Autotuner search logic against an analytic performance/memory model (no subprocesses), plus space helpers.
Mirrors reference tests/unit/autotuning/test_autotuning.py."""
import json
import os

from deepspeed_b200.autotuning import Autotuner
from deepspeed_b200.autotuning.scheduler import ResourceManager
from deepspeed_b200.autotuning.tuner import GridSearchTuner, ModelBasedTuner, RandomTuner
from deepspeed_b200.autotuning.utils import canonical_name, get_all_configs, replace_dict


def test_config_space_helpers():
    space = {"zero_optimization": {"stage": 2, "overlap_comm": [True, False], "reduce_bucket_size": [5e7, 5e8]}}
    cfgs = get_all_configs(space)
    assert len(cfgs) == 4 and all(c["zero_optimization"]["stage"] == 2 for c in cfgs)
    assert {c["zero_optimization"]["reduce_bucket_size"] for c in cfgs} == {5e7, 5e8}
    merged = replace_dict({"a": {"b": 1, "c": 2}, "d": 3}, {"a": {"b": 5}, "d": None})
    assert merged == {"a": {"b": 5, "c": 2}}
    name = canonical_name({"train_micro_batch_size_per_gpu": 8, "gradient_accumulation_steps": 2,
                           "zero_optimization": {"stage": 3, "reduce_bucket_size": 5e8}},
                          ["zero_optimization.reduce_bucket_size"])
    assert name.startswith("z3_mbs8_gas2") and "rbs" in name


N_PARAMS = 2e9
GPU_MEM = 40e9
ACT_PER_SAMPLE = 1.5e9


def _perf_model(exp):
    """8 GPUs; model states per the ZeRO memory model; OOM -> failure; throughput grows with micro-batch size,
    higher stages pay a communication tax, overlap_comm helps."""
    cfg = exp["ds_config"]
    if (cfg.get("autotuning", {}).get("model_info") or {}).get("profile"):
        path = cfg["autotuning"]["model_info_path"]
        with open(path, "w") as f:
            json.dump({"num_params": N_PARAMS, "activation_mem_per_gpu": ACT_PER_SAMPLE}, f)
        return {"throughput": 1.0, "latency": 1.0}
    z = cfg["zero_optimization"]
    stage, mbs = z["stage"], cfg["train_micro_batch_size_per_gpu"]
    g = 8
    mem = N_PARAMS * (2 / (g if stage >= 3 else 1) + 2 / (g if stage >= 2 else 1) + 12 / (g if stage >= 1 else 1))
    mem += mbs * ACT_PER_SAMPLE
    if mem > GPU_MEM:
        raise RuntimeError("CUDA out of memory")
    tax = {0: 1.0, 1: 0.98, 2: 0.95, 3: 0.80}[stage]
    if z.get("overlap_comm") is True:
        tax *= 1.04
    if z.get("reduce_bucket_size") == 5e8:
        tax *= 1.02
    return {"throughput": g * mbs / (0.05 + 0.01 * mbs) * tax, "latency": (0.05 + 0.01 * mbs) * 1e3}


def test_autotuner_finds_best_stage_and_micro_batch(tmp_path):
    user = {"train_micro_batch_size_per_gpu": 1, "bf16": {"enabled": True},
            "autotuning": {"enabled": True, "results_dir": str(tmp_path / "results"), "exps_dir": str(tmp_path / "exps"),
                           "tuner_type": "gridsearch", "tuner_num_trials": 100, "tuner_early_stopping": 100,
                           "max_train_micro_batch_size_per_gpu": 64, "num_tuning_micro_batch_sizes": 4,
                           "fast": False}}
    tuner = Autotuner(None, {"localhost": list(range(8))}, user_config=user, run_fn=_perf_model, gpu_memory=GPU_MEM)
    best = tuner.tune()
    assert tuner.model_info["num_params"] == N_PARAMS
    # stage 0 needs 32 GB of states -> tiny batches; stage 2 (4.0+0.5+3 GB) fits ~16 samples and beats stage 3's tax
    exp, val, _ = best["global"]
    z = exp["ds_config"]["zero_optimization"]
    assert z["stage"] == 2, best
    assert exp["ds_config"]["train_micro_batch_size_per_gpu"] >= 8
    assert z.get("overlap_comm") is True and z.get("reduce_bucket_size") == 5e8
    assert "z3" in best and best["z3"][1] < val
    opt = json.load(open(os.path.join(user["autotuning"]["results_dir"], "ds_config_optimal.json")))
    assert opt["zero_optimization"]["stage"] == 2 and "autotuning" not in opt
    tuner.print_tuning_results()


def test_tuners_agree_on_small_space(tmp_path):
    exps = []
    for mbs in (1, 2, 4, 8):
        for oc in (True, False):
            cfg = {"train_micro_batch_size_per_gpu": mbs, "zero_optimization": {"stage": 1, "overlap_comm": oc}}
            exps.append({"name": f"e_mbs{mbs}_oc{int(oc)}", "ds_config": cfg, "num_gpus": 1, "num_nodes": 1})
    for cls in (GridSearchTuner, RandomTuner, ModelBasedTuner):
        rm = ResourceManager(None, ["localhost"], 8, str(tmp_path / cls.__name__), str(tmp_path / "e"),
                             run_fn=_perf_model)
        t = cls(list(exps), rm, "throughput")
        t.tune(sample_size=2, n_trials=100, early_stopping=100)
        assert t.best_exp["ds_config"]["train_micro_batch_size_per_gpu"] == 8
        assert t.best_exp["ds_config"]["zero_optimization"]["overlap_comm"] is True


def test_autotuning_dict_helpers():
    from deepspeed_b200.autotuning import utils as U
    cfg = {"a": "$MBS", "b": ["run_$NAME", 3], "c": {"d": "$NVME_PATH/x", "e": "$GAS"}}
    U.find_replace(cfg, {"mbs": 4, "name": "exp", "gas": "2"})
    assert cfg == {"a": 4, "b": ["run_exp", "3"], "c": {"d": "$NVME_PATH/x", "e": "2"}}
    assert U.combine_dict({"x": 1, "n": {"y": 2}}, {"x": 3, "n": {"y": [2, 5]}, "z": 0}) == \
        {"x": [1, 3], "n": {"y": [2, 5]}, "z": 0}
    d = {"zero_optimization": {"stage": 3, "offload_optimizer": {"device": "cpu"}}, "k": {"stage": 1}}
    assert U.get_val_by_key(d, "device") == "cpu" and U.get_val_by_key(d, "nope") is None
    U.set_val_by_key(d, "stage", 2)
    assert d["zero_optimization"]["stage"] == 2 and d["k"]["stage"] == 2
    assert not U.validate_ds_config({"zero_optimization": {"stage": 3, "offload_param": {"device": "nvme"}}})
    assert U.validate_ds_config({"zero_optimization": {"stage": 3, "offload_param": {"device": "nvme"}},
                                 "optimizer": {"type": "Adam"}})
    assert not U.validate_ds_config({"zero_optimization": {"stage": 2, "cpu_offload": True, "cpu_offload_params": True}})
    cs = [{"a": 1, "t": {"x": 1}}, {"a": 1, "t": {"x": 2}}, {"a": 2, "t": {"x": 1}}]
    assert U.prune_configs(cs, ["x"]) == [{"a": 1, "t": {}}, {"a": 2, "t": {}}]
    assert sorted(U.get_tuning_keys({"mbs": [1, 2], "z": {"stage": [1, 2, 3], "bucket": [5]}, "gas": 1})) == \
        ["mbs", "stage"]
    assert U.was_interrupted is U.was_interruptted
