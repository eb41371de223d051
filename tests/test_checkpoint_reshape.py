"""Checkpoint directory readers and topology contraction (checkpoint/reshape.py). Reference: checkpoint/reshape_*.py,
zero_checkpoint.py, deepspeed_checkpoint.py; tests/unit/checkpoint/test_reshape_checkpoint.py."""
import os

import pytest
import torch

from deepspeed_b200.checkpoint import (DeepSpeedCheckpoint, ZeROCheckpoint, get_model_3d_descriptor,
                                       get_model_ckpt_name_for_rank, get_zero_ckpt_name_for_rank, merge_state,
                                       model_3d_desc, reshape_meg_2d_parallel)
from deepspeed_b200.checkpoint import constants as C


def test_2d_reshape_maps():
    m = reshape_meg_2d_parallel(old_pp_degree=2, old_tp_degree=4, new_pp_degree=1, new_tp_degree=2)
    assert m.get_data(0, 0) == [0, 1, 4, 5] and m.get_data(0, 1) == [2, 3, 6, 7]
    m = reshape_meg_2d_parallel(4, 2, 2, 2)
    assert m.get_data(0, 0) == [0, 2] and m.get_data(1, 1) == [5, 7]
    same = reshape_meg_2d_parallel(2, 2, 2, 2)
    assert [same.get_data(i, j) for i in range(2) for j in range(2)] == [[0], [1], [2], [3]]
    with pytest.raises(AssertionError):
        reshape_meg_2d_parallel(1, 2, 2, 2)


def test_3d_descriptor_reshape():
    src = model_3d_desc(pp_degree=2, tp_degree=2, dp_degree=4)
    assert src.world_size() == 16
    ok, errs = src.can_reshape(model_3d_desc(4, 2, 4))
    assert not ok and "PP" in errs[0]
    maps = src.reshape(model_3d_desc(pp_degree=1, tp_degree=2, dp_degree=2))
    assert len(maps) == 2
    # target (pp 0, tp 0): source 2-D ranks 0 and 2 (both pipeline stages), each at dp 0..3 -> 8 files, 4 per target dp
    assert maps[0].get_data(0, 0) == [0, 4, 8, 12] and maps[1].get_data(0, 0) == [2, 6, 10, 14]
    assert src.is_valid(1, 1, 3)[0] and not src.is_valid(2, 0, 0)[0]


def test_merge_state_concatenates_tensors_and_checks_structure():
    a = {"w": torch.arange(3.0), "n": 5, "l": [torch.zeros(1), {"x": torch.ones(2)}]}
    b = {"w": torch.arange(3.0, 5.0), "n": 7, "l": [torch.ones(1), {"x": torch.zeros(1)}], "extra": 1}
    m = merge_state(a, b)
    assert m["w"].tolist() == [0, 1, 2, 3, 4] and m["n"] == 5 and m["extra"] == 1
    assert m["l"][0].tolist() == [0, 1] and m["l"][1]["x"].tolist() == [1, 1, 0]
    with pytest.raises(ValueError):
        merge_state({"l": [1]}, {"l": [1, 2]})
    with pytest.raises(ValueError):
        merge_state({"a": 1}, {"a": "x"})


def _write_ckpt(d, tp, dp, numel=8, pad=2):
    """A stage-1/2 style checkpoint folder: one model-state file per tp rank, one optimizer file per (dp, tp)."""
    for t in range(tp):
        torch.save({"module": {"w": torch.full((2, ), float(t))}, C.ITERATION_KEY: 123,
                    C.CHECKPOINT_PARALLEL_DIMS: {C.CHECKPOINT_PP_DEGREE: 1, C.CHECKPOINT_TP_DEGREE: tp}},
                   get_model_ckpt_name_for_rank(d, f"{t:02d}"))
        for r in range(dp):
            last = r == dp - 1
            flat = torch.arange(numel, dtype=torch.float32) + 100 * r + 1000 * t
            state = {0: {"exp_avg": flat.clone(), "step": 3}}
            sd = {C.OPTIMIZER_STATE_DICT: {C.SINGLE_PARTITION_OF_FP32_GROUPS: [flat], C.PARTITION_COUNT: [dp],
                                           C.GROUP_PADDINGS: [pad if last else 0], C.ZERO_STAGE: 2,
                                           C.BASE_OPTIMIZER_STATE: {C.GROUP_STATE_KEY: state}}}
            torch.save(sd, get_zero_ckpt_name_for_rank(d, r, t))


def test_zero_checkpoint_reads_and_contracts_dp(tmp_path):
    d = str(tmp_path)
    _write_ckpt(d, tp=2, dp=4)
    desc = get_model_3d_descriptor(d)
    assert (desc.pp_degree, desc.tp_degree, desc.dp_degree) == (1, 2, 4)
    z = ZeROCheckpoint(d)
    assert z.num_files == 8 and z.get_src_dp_degree() == 4 and z.get_src_world_size() == 8
    assert os.path.basename(z.get_files_for_rank(0, 1, 2)[0]) == "zero_pp_rank_2_mp_rank_01_optim_states.pt"
    one = z.get_state_for_rank(0, 1, 3)  # last dp rank: padding stripped from the optimizer state
    opt = one[C.OPTIMIZER_STATE_DICT]
    assert opt[C.BASE_OPTIMIZER_STATE][C.GROUP_STATE_KEY][0]["exp_avg"].numel() == 6 and opt[C.GROUP_PADDINGS] == [0]
    z.reshape(model_3d_desc(pp_degree=1, tp_degree=2, dp_degree=2))
    files = [os.path.basename(f) for f in z.get_files_for_rank(0, 0, 1)]
    assert files == ["zero_pp_rank_2_mp_rank_00_optim_states.pt", "zero_pp_rank_3_mp_rank_00_optim_states.pt"]
    merged = z.get_state_for_rank(0, 0, 1)[C.OPTIMIZER_STATE_DICT]
    assert merged[C.PARTITION_COUNT] == [2]
    flat = merged[C.SINGLE_PARTITION_OF_FP32_GROUPS][0]
    assert flat.numel() == 16 and flat[0] == 200 and flat[8] == 300
    assert merged[C.BASE_OPTIMIZER_STATE][C.GROUP_STATE_KEY][0]["exp_avg"].numel() == 8 + 6
    with pytest.raises(AssertionError):
        z.reshape(model_3d_desc(1, 2, 8))


def test_deepspeed_checkpoint_topology_view(tmp_path):
    d = str(tmp_path)
    _write_ckpt(d, tp=2, dp=2)
    ck = DeepSpeedCheckpoint(d, tp_degree=1, dp_degree=1)
    assert (ck.original_tp_degree, ck.original_pp_degree, ck.original_dp_degree) == (2, 1, 2)
    assert ck.is_change_tp_degree() and ck.is_change_dp_degree() and not ck.is_change_pp_degree()
    assert ck.get_iteration() == 123 and ck.world_size == 1 and ck.original_world_size == 4
    assert len(ck.get_2d_parallel_files(tp_index=0, pp_index=0)) == 2
    assert ck.get_2d_parallel_state(0, 0)["module"]["w"].tolist() == [0, 0, 1, 1]
    assert len(ck.get_zero_files(0, 0, 0)) == 4
    st = ck.get_zero_checkpoint_state(0, 0, 0)
    assert st[C.OPTIMIZER_STATE_DICT][C.SINGLE_PARTITION_OF_FP32_GROUPS][0].numel() == 32
