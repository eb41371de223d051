"""Fused NVLink collectives (csrc/symm.cu) vs the NCCL oracle on >=2 GPUs; ZeRO-3 fused vs NCCL path parity."""
import pytest
import torch

from common import run_distributed

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


needs2 = pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 GPUs")


def _collectives(world):
    import torch.distributed as dist
    from deepspeed_b200.comm import symm
    rank = dist.get_rank()
    dev = torch.device("cuda", torch.cuda.current_device())
    assert symm.is_supported(None)
    ctx = symm.SymmContext.get(dist.group.WORLD)
    out = {}
    # ---- reduce-scatter (bf16 in, fp32 / bf16 out, accumulate on second call) ----
    shard = 256 * 1024 + 128
    n = shard * world
    buf = ctx.alloc(n * 2)
    for trial, (odt, use_mc) in enumerate([(torch.float32, False), (torch.bfloat16, False), (torch.float32, True)]):
        torch.manual_seed(100 + rank + 10 * trial)
        x = torch.randn(n, device=dev).to(torch.bfloat16)
        ctx.wait_rs_done(buf.last_rs_epoch)
        buf.view(n, torch.bfloat16).copy_(x)
        acc = torch.zeros(shard, dtype=odt, device=dev)
        ctx.reduce_scatter(buf, torch.bfloat16, shard, acc, accumulate=False, scale=0.5, use_multicast=use_mc)
        ctx.reduce_scatter(buf, torch.bfloat16, shard, acc, accumulate=True, scale=0.5, use_multicast=use_mc)
        ref = torch.empty(shard, dtype=torch.float32, device=dev)
        dist.reduce_scatter_tensor(ref, x.float())
        torch.cuda.synchronize()
        err = (acc.float() - ref).abs().max().item()
        out[f"rs_{trial}"] = (err, ref.abs().max().item(), bool(buf.mc_ptr))
    # ---- all-gather ----
    sb = ctx.alloc(shard * 2, want_multicast=False)
    torch.manual_seed(7 + rank)
    mine = torch.randn(shard, device=dev).to(torch.bfloat16)
    sb.view(shard, torch.bfloat16).copy_(mine)
    full = torch.empty(n, dtype=torch.bfloat16, device=dev)
    ctx.all_gather(sb, 0, shard * 2, full)
    ref = torch.empty(n, dtype=torch.bfloat16, device=dev)
    dist.all_gather_into_tensor(ref, mine)
    torch.cuda.synchronize()
    out["ag"] = bool(torch.equal(full, ref))
    # ---- all-to-all (plain and interleaved layouts) ----
    rows, rowb = 96, 512
    ib = ctx.alloc(world * rows * rowb, want_multicast=False)
    src = torch.randint(0, 255, (world * rows * rowb, ), dtype=torch.uint8, device=dev)
    ib.tensor[:src.numel()].copy_(src)
    o = torch.empty_like(src)
    ctx.all_to_all(ib, rows, rowb, o, interleave=False)
    ref = torch.empty_like(src)
    dist.all_to_all_single(ref, src)
    torch.cuda.synchronize()
    out["a2a"] = bool(torch.equal(o, ref))
    o2 = torch.empty_like(src)
    ctx.wait_a2a_done(ib.last_a2a_epoch)
    ctx.all_to_all(ib, rows, rowb, o2, interleave=True)
    torch.cuda.synchronize()
    ref2 = ref.view(world, rows, rowb).permute(1, 0, 2).contiguous().view(-1)
    out["a2a_interleave"] = bool(torch.equal(o2, ref2))
    dist.barrier()
    return out


@needs2
@pytest.mark.parametrize("world", [2])
def test_symm_collectives_match_nccl(world):
    outs = run_distributed(_collectives, min(world, _ngpu()), args=(min(world, _ngpu()), ))
    for o in outs:
        assert o["ag"] and o["a2a"] and o["a2a_interleave"], o
        for k in ("rs_0", "rs_1", "rs_2"):
            err, mag, _ = o[k]
            assert err <= 2e-2 * max(mag, 1.0), (k, o[k])


def _train(fused, steps):
    import deepspeed_b200 as ds
    from deepspeed_b200.models import LlamaConfig, LlamaForCausalLM
    dev = torch.device("cuda", torch.cuda.current_device())
    torch.manual_seed(0)
    cfg = LlamaConfig(vocab_size=2048, hidden_size=256, intermediate_size=512, num_hidden_layers=3,
                      num_attention_heads=4, num_key_value_heads=2, max_position_embeddings=256)
    model = LlamaForCausalLM(cfg, device=dev)
    eng, _, _, _ = ds.initialize(model=model, config={
        "train_micro_batch_size_per_gpu": 2, "bf16": {"enabled": True}, "gradient_clipping": 1.0,
        "optimizer": {"type": "AdamW", "params": {"lr": 1e-3}},
        "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 1000, "fused_collectives": fused}})
    assert (eng.optimizer.symm is not None) == fused
    torch.manual_seed(1 + ds.comm.get_rank())
    losses = []
    for _ in range(steps):
        ids = torch.randint(0, 2048, (2, 128), device=dev)
        loss = eng(ids, labels=ids)
        eng.backward(loss)
        eng.step()
        losses.append(float(loss.detach()))
    return losses


@needs2
def test_zero3_fused_collectives_match_nccl_path():
    a = run_distributed(_train, 2, args=(True, 6))
    b = run_distributed(_train, 2, args=(False, 6))
    for la, lb in zip(a[0], b[0]):
        assert abs(la - lb) < 5e-2, (a[0], b[0])
    assert a[0][-1] < a[0][0]
