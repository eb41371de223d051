"""AutoTP: tensor-parallel Llama (tp=2) reproduces the single-rank loss and weight update
(reference tests/unit/model_parallelism/test_autotp_training.py)."""
import torch

from common import run_distributed


def _run(tp):
    import deepspeed_b200 as ds
    from deepspeed_b200.models import LlamaConfig, LlamaForCausalLM
    from deepspeed_b200.parallel import LinearAllreduce, LinearLayer
    torch.manual_seed(0)
    cfg = LlamaConfig.tiny(dtype=torch.float32, gradient_checkpointing=False)
    model = LlamaForCausalLM(cfg)
    full_o = model.model.layers[0].self_attn.o_proj.weight.detach().clone()
    conf = {"train_micro_batch_size_per_gpu": 2, "optimizer": {"type": "SGD", "params": {"lr": 0.1}},
            "zero_optimization": {"stage": 1}}
    if tp > 1:
        conf["tensor_parallel"] = {"autotp_size": tp}
    eng, _, _, _ = ds.initialize(model=model, config=conf)
    if tp > 1:
        assert isinstance(model.model.layers[0].self_attn.qkv_proj, LinearLayer)
        assert isinstance(model.model.layers[0].mlp.down_proj, LinearAllreduce)
    torch.manual_seed(3)
    ids = torch.randint(0, 512, (2, 24))
    losses = []
    for _ in range(2):
        loss = eng(ids, labels=ids)
        eng.backward(loss)
        eng.step()
        losses.append(float(loss.detach()))
    from deepspeed_b200.utils.tensor_fragment import safe_get_full_fp32_param
    w = safe_get_full_fp32_param(model.model.layers[0].self_attn.o_proj.weight).clone()
    return losses, w, ds.comm.get_rank()


def test_autotp_matches_single_rank():
    single = run_distributed(_run, 1, args=(1, ))[0]
    tp = run_distributed(_run, 2, args=(2, ))
    for a, b in zip(single[0], tp[0][0]):
        assert abs(a - b) < 2e-4, (single[0], tp[0][0])
    # row-parallel o_proj: rank r holds columns [r*half, (r+1)*half)
    full = single[1]
    half = full.shape[1] // 2
    for losses, w, rank in tp:
        torch.testing.assert_close(w, full[:, rank * half:(rank + 1) * half], rtol=1e-4, atol=1e-5)
