"""Checkpoint round trips: save/load (all stages), zero_to_fp32 (ours AND the reference's script reading our
files), universal conversion with DP reshape 2 -> 1 (reference tests/unit/checkpoint/*)."""
import os
import subprocess
import sys

import pytest
import torch

from common import run_distributed
from simple_model import SimpleModel, random_batches

HID = 16
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _cfg(stage, universal=False):
    c = {"train_micro_batch_size_per_gpu": 4, "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}},
         "zero_optimization": {"stage": stage, "unit_max_numel": 300, "stage3_param_persistence_threshold": 20,
                               "reduce_bucket_size": 400}, "gradient_clipping": 1.0}
    if universal:
        c["checkpoint"] = {"load_universal": True}
    return c


def _train_and_save(stage, ckpt_dir, steps=3):
    import deepspeed_b200 as ds
    from deepspeed_b200.utils.tensor_fragment import safe_get_full_fp32_param, safe_get_full_optimizer_state
    torch.manual_seed(0)
    model = SimpleModel(HID, 3)
    eng, _, _, _ = ds.initialize(model=model, config=_cfg(stage))
    rank, world = ds.comm.get_rank(), ds.comm.get_world_size()
    batches = random_batches(steps * world + world, 4, HID, "cpu")
    for i in range(steps):
        x, y = batches[i * world + rank]
        loss = eng(x, y)
        eng.backward(loss)
        eng.step()
    eng.save_checkpoint(ckpt_dir, tag="t1", client_state={"hello": 7})
    full = {n: safe_get_full_fp32_param(p).clone() for n, p in model.named_parameters()}
    m = {n: safe_get_full_optimizer_state(p, "exp_avg").clone() for n, p in model.named_parameters()}
    # one more step: the reference trajectory the reloaded engine must reproduce
    x, y = batches[steps * world + rank]
    loss = eng(x, y)
    eng.backward(loss)
    eng.step()
    after = {n: safe_get_full_fp32_param(p).clone() for n, p in model.named_parameters()}
    return full, m, after


def _load_and_step(stage, ckpt_dir, steps=3, universal=False, tag=None):
    import deepspeed_b200 as ds
    from deepspeed_b200.utils.tensor_fragment import safe_get_full_fp32_param, safe_get_full_optimizer_state
    torch.manual_seed(123)  # different init: everything must come from the checkpoint
    model = SimpleModel(HID, 3)
    eng, _, _, _ = ds.initialize(model=model, config=_cfg(stage, universal))
    path, client = eng.load_checkpoint(ckpt_dir, tag=tag)
    assert path is not None
    full = {n: safe_get_full_fp32_param(p).clone() for n, p in model.named_parameters()}
    m = {n: safe_get_full_optimizer_state(p, "exp_avg").clone() for n, p in model.named_parameters()}
    rank, world = ds.comm.get_rank(), ds.comm.get_world_size()
    return full, m, client, eng.global_steps


@pytest.mark.parametrize("stage", [1, 2, 3])
def test_save_load_roundtrip_and_zero_to_fp32(stage, tmp_path):
    d = str(tmp_path / "ck")
    saved = run_distributed(_train_and_save, 2, args=(stage, d))
    full, m, after = saved[0]
    assert open(os.path.join(d, "latest")).read() == "t1"
    files = sorted(os.listdir(os.path.join(d, "t1")))
    assert any(f.endswith("_optim_states.pt") for f in files) and "zero_to_fp32.py" in files
    # ---- reload on 2 ranks ----
    loaded = run_distributed(_load_and_step, 2, args=(stage, d))
    lfull, lm, client, gsteps = loaded[0]
    assert client.get("hello") == 7 and gsteps == 3
    for n in full:
        torch.testing.assert_close(lfull[n], full[n], rtol=0, atol=0)
        torch.testing.assert_close(lm[n], m[n], rtol=0, atol=1e-7)
    # ---- offline consolidation with OUR script (copied into the checkpoint) ----
    out = str(tmp_path / "fp32.pt")
    subprocess.check_call([sys.executable, os.path.join(d, "t1", "zero_to_fp32.py"), d, out], cwd=str(tmp_path),
                          env=dict(os.environ, PYTHONPATH=""))
    sd = torch.load(out, weights_only=False)
    for n in full:
        torch.testing.assert_close(sd[n], full[n], rtol=0, atol=0)
    # ---- and with the REFERENCE's zero_to_fp32 reading our files (on-disk format compatibility) ----
    ref = os.path.join(REPO, "baseline", "_ref", "deepspeed", "utils", "zero_to_fp32.py")
    if os.path.exists(ref):
        out2 = str(tmp_path / "ref_out")
        subprocess.check_call([sys.executable, "-m", "deepspeed.utils.zero_to_fp32", d, out2], cwd=str(tmp_path),
                              env=dict(os.environ, PYTHONPATH=os.path.join(REPO, "baseline", "_ref")))
        got = {}
        for f in os.listdir(out2):
            if f.endswith(".bin"):
                got.update(torch.load(os.path.join(out2, f), weights_only=False))
        for n in full:
            torch.testing.assert_close(got[n].float(), full[n], rtol=0, atol=0)


@pytest.mark.parametrize("stage", [2, 3])
def test_universal_checkpoint_dp_reshape(stage, tmp_path):
    d = str(tmp_path / "ck")
    saved = run_distributed(_train_and_save, 2, args=(stage, d))
    full, m, _ = saved[0]
    from deepspeed_b200.checkpoint import ds_to_universal
    args = ds_to_universal.parse_arguments(["--input_folder", os.path.join(d, "t1"), "--output_folder",
                                            os.path.join(d, "t1_universal"), "--num_extract_workers", "1",
                                            "--num_merge_workers", "1"])
    ds_to_universal.main(args)
    assert open(os.path.join(d, "latest_universal")).read() == "t1_universal"
    # load on ONE rank (dp 2 -> 1)
    loaded = run_distributed(_load_and_step, 1, args=(stage, d), kwargs=dict(universal=True))
    lfull, lm, _, gsteps = loaded[0]
    assert gsteps == 3
    for n in full:
        torch.testing.assert_close(lfull[n], full[n], rtol=0, atol=0)
        torch.testing.assert_close(lm[n], m[n], rtol=0, atol=1e-7)


def _frozen_model(seed):
    torch.manual_seed(seed)
    model = SimpleModel(HID, 3)
    # frozen: one large (partitioned) weight and one small (persistent, replicated) bias -- LoRA-style fine-tuning
    model.linears[1].weight.requires_grad_(False)
    model.linears[0].bias.requires_grad_(False)
    return model


def _frozen_values(eng, model):
    import deepspeed_b200 as ds
    out = {}
    for n, p in model.named_parameters():
        with ds.zero.GatheredParameters([p]):
            out[n] = p.detach().float().cpu().clone()
    return out


def _train_and_save_frozen(ckpt_dir):
    import deepspeed_b200 as ds
    model = _frozen_model(0)
    eng, _, _, _ = ds.initialize(model=model, model_parameters=[p for p in model.parameters() if p.requires_grad],
                                 config=_cfg(3))
    rank, world = ds.comm.get_rank(), ds.comm.get_world_size()
    batches = random_batches(3 * world, 4, HID, "cpu")
    for i in range(3):
        x, y = batches[i * world + rank]
        loss = eng(x, y)
        eng.backward(loss)
        eng.step()
    eng.save_checkpoint(ckpt_dir, tag="t1")
    return _frozen_values(eng, model)


def _load_frozen(ckpt_dir):
    import deepspeed_b200 as ds
    model = _frozen_model(321)  # different init, frozen weights included: everything must come from the checkpoint
    eng, _, _, _ = ds.initialize(model=model, model_parameters=[p for p in model.parameters() if p.requires_grad],
                                 config=_cfg(3))
    path, _ = eng.load_checkpoint(ckpt_dir, tag="t1")
    assert path is not None
    vals = _frozen_values(eng, model)
    x, y = random_batches(1, 4, HID, "cpu")[0]
    loss = eng(x, y)  # the restored shards must also drive a forward/backward
    eng.backward(loss)
    eng.step()
    return vals


def test_zero3_frozen_params_roundtrip(tmp_path):
    """ZeRO-3 with frozen parameters: fragments are saved in the reference layout (this rank's ceil(n/world) slice),
    both zero_to_fp32 scripts rebuild them, and load restores them (reference engine.py:5292-5306,
    zero_to_fp32.py _zero3_merge_frozen_params)."""
    d = str(tmp_path / "ck")
    want = run_distributed(_train_and_save_frozen, 2, args=(d, ))[0]
    frozen = ["linears.1.weight", "linears.0.bias"]
    ms = torch.load(os.path.join(d, "t1", "zero_pp_rank_0_mp_rank_00_model_states.pt"), weights_only=False)
    assert set(ms["frozen_param_shapes"]) == set(frozen)
    assert ms["frozen_param_fragments"]["linears.1.weight"].numel() == HID * HID // 2
    for names in ms["param_shapes"]:
        assert not set(names) & set(frozen)
    got = run_distributed(_load_frozen, 2, args=(d, ))[0]
    for n in want:
        torch.testing.assert_close(got[n], want[n], rtol=0, atol=0)
    out = str(tmp_path / "fp32.pt")
    subprocess.check_call([sys.executable, os.path.join(d, "t1", "zero_to_fp32.py"), d, out], cwd=str(tmp_path),
                          env=dict(os.environ, PYTHONPATH=""))
    sd = torch.load(out, weights_only=False)
    for n in frozen:
        torch.testing.assert_close(sd[n], want[n], rtol=0, atol=0)
    ref = os.path.join(REPO, "baseline", "_ref", "deepspeed", "utils", "zero_to_fp32.py")
    if os.path.exists(ref):
        out2 = str(tmp_path / "ref_out")
        subprocess.check_call([sys.executable, "-m", "deepspeed.utils.zero_to_fp32", d, out2], cwd=str(tmp_path),
                              env=dict(os.environ, PYTHONPATH=os.path.join(REPO, "baseline", "_ref")))
        ref_sd = {}
        for f in os.listdir(out2):
            if f.endswith(".bin"):
                ref_sd.update(torch.load(os.path.join(out2, f), weights_only=False))
        for n in want:
            torch.testing.assert_close(ref_sd[n].float(), want[n], rtol=0, atol=1e-6)


def test_zero_to_fp32_lazy_gathered_tensors_and_file_helpers(tmp_path):
    import torch
    from deepspeed_b200.checkpoint import zero_to_fp32 as Z
    flats = [torch.arange(0, 10).float(), torch.arange(100, 110).float()]  # two ranks' flat buffers
    g = Z.GatheredTensor(flats, offset=2, per_rank_numel=3, shape=(5, ))
    assert g.numel() == 5 and g.contiguous().tolist() == [2.0, 3.0, 4.0, 102.0, 103.0]
    assert Z.zero3_partitioned_param_info(5, 2) == (3, 1) and Z.zero3_partitioned_param_info(6, 2) == (3, 0)
    lazy = Z._merge_stage3(flats, [{"a": torch.Size([5]), "b": torch.Size([2, 2])}], 2, lazy=True)
    assert isinstance(lazy["a"], Z.GatheredTensor)
    lazy["alias"] = lazy["a"]
    eager = Z.to_torch_tensor(lazy)
    assert eager["a"].tolist() == [0.0, 1.0, 2.0, 100.0, 101.0] and eager["b"].tolist() == [[3.0, 4.0], [103.0, 104.0]]
    assert eager["alias"] is eager["a"]
    assert Z.to_torch_tensor(lazy, return_empty_tensor=True)["b"].shape == (2, 2)
    (tmp_path / "mp_rank_00_model_states.pt").write_bytes(b"x")
    (tmp_path / "layer_0_expert_1_mp_rank_00_model_states.pt").write_bytes(b"x")
    assert [os.path.basename(f) for f in Z.get_model_state_files(str(tmp_path))] == ["mp_rank_00_model_states.pt"]
    assert Z.get_model_state_file(str(tmp_path), 2).endswith("mp_rank_00_model_states.pt")
    import pytest
    with pytest.raises(FileNotFoundError):
        Z.get_optim_files(str(tmp_path))
    with pytest.raises(FileNotFoundError):
        Z.get_model_state_file(str(tmp_path), 3)
