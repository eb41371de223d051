#!/usr/bin/env python
"""The BASELINE.json configurations other than the headline one (which is ``bench.py``):

    zeropp    Llama-3 8B, ZeRO-3 + ZeRO++ (qwZ block-scaled FP8 all-gather, hpZ secondary partition, qgZ)
    ulysses   Llama-3 8B, Ulysses sequence parallel over all GPUs, long sequences (--seq total tokens)
    moe       Mixtral 8x7B-shaped MoE, expert parallel over all GPUs (AutoEP all-to-all dispatch/combine)
    gpt2      GPT-2 small, ZeRO-1 (runs on CPU/gloo too)

    torchrun --nproc-per-node N benchmarks/bench_configs.py --config zeropp [--layers L] [--seq S] [--steps K]

Timing: CUDA events, W warm-up steps, K timed steps, max over ranks, inputs copied from pinned host memory every
step. Prints ONE JSON line on rank 0. ``--layers`` below the model's depth marks the line ``"reduced": true``."""
import argparse
import json
import os
import sys
import time

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", required=True, choices=["zeropp", "ulysses", "moe", "gpt2"])
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--layers", type=int, default=0, help="0 = the model's real depth")
    ap.add_argument("--seq", type=int, default=0, help="tokens per sequence (ulysses: total, split over ranks)")
    ap.add_argument("--mbs", type=int, default=1)
    ap.add_argument("--no-ckpt", action="store_true")
    ap.add_argument("--local_rank", type=int, default=-1)
    return ap.parse_args()


def base_config(args, extra_zero=None, stage=3, clip=1.0):
    z = {"stage": stage, "overlap_comm": True, "stage3_param_persistence_threshold": 1e5}
    z.update(extra_zero or {})
    return {"train_micro_batch_size_per_gpu": args.mbs, "gradient_accumulation_steps": 1, "steps_per_print": 10**9,
            "bf16": {"enabled": True}, "gradient_clipping": clip,
            "optimizer": {"type": "AdamW", "params": {"lr": 1e-5, "betas": [0.9, 0.95], "weight_decay": 0.1}},
            "zero_optimization": z}


def build(args, world, dev):
    import deepspeed_b200 as ds
    from deepspeed_b200.models import LlamaConfig, LlamaForCausalLM
    ckpt = not args.no_ckpt
    meta = {}
    if args.config in ("zeropp", "ulysses"):
        layers = args.layers or 32
        cfg = LlamaConfig(vocab_size=128256, hidden_size=4096, intermediate_size=14336, num_hidden_layers=layers,
                          num_attention_heads=32, num_key_value_heads=8, max_position_embeddings=1 << 20,
                          rope_theta=500000.0, gradient_checkpointing=ckpt, dtype=torch.bfloat16)
        seq = args.seq or (4096 if args.config == "zeropp" else 32768 * world)
        torch.manual_seed(1234)
        model = LlamaForCausalLM(cfg, device=dev)
        if args.config == "zeropp":
            dcfg = base_config(args, {"zero_quantized_weights": True, "zero_quantized_gradients": True,
                                      "zero_hpz_partition_size": min(world, 8) if world > 1 else 1})
            meta = {"model": f"Llama-3-8B ({layers} layers)", "parallelism": f"zero3++ dp{world} (qwZ fp8, hpZ, qgZ)",
                    "seq_len": seq, "tokens_per_step": world * args.mbs * seq}
            local_seq = seq
        else:
            dcfg = base_config(args)
            dcfg["sequence_parallel_size"] = world
            meta = {"model": f"Llama-3-8B ({layers} layers)", "parallelism": f"ulysses sp{world} + zero3",
                    "seq_len": seq, "tokens_per_step": args.mbs * seq}
            local_seq = seq // world
        engine, _, _, _ = ds.initialize(model=model, config=dcfg)
        if args.config == "ulysses":
            from deepspeed_b200.sequence import enable_ulysses
            enable_ulysses(model)
        full_layers, vocab = 32, 128256
    elif args.config == "moe":
        from deepspeed_b200.models.mixtral import MixtralConfig, MixtralForCausalLM
        layers = args.layers or 32
        cfg = MixtralConfig(num_hidden_layers=layers, ep_size=world, gradient_checkpointing=ckpt, dtype=torch.bfloat16)
        seq = args.seq or 4096
        torch.manual_seed(1234)
        model = MixtralForCausalLM(cfg, device=dev)
        engine, _, _, _ = ds.initialize(model=model, config=base_config(args, stage=2))
        meta = {"model": f"Mixtral-8x7B shape ({layers} layers, 8 experts top-2)",
                "parallelism": f"ep{world} x zero2", "seq_len": seq, "tokens_per_step": world * args.mbs * seq}
        local_seq, full_layers, vocab = seq, 32, 32000
    else:
        from deepspeed_b200.models.gpt2 import GPT2Config, GPT2LMHeadModel
        layers = args.layers or 12
        cfg = GPT2Config(n_layer=layers)
        seq = args.seq or 1024
        torch.manual_seed(1234)
        model = GPT2LMHeadModel(cfg).to(dev)
        dcfg = base_config(args, stage=1)
        if dev.type == "cpu":
            dcfg.pop("bf16")
        engine, _, _, _ = ds.initialize(model=model, config=dcfg)
        meta = {"model": f"GPT-2 small ({layers} layers)", "parallelism": f"zero1 dp{world}", "seq_len": seq,
                "tokens_per_step": world * args.mbs * seq}
        local_seq, full_layers, vocab = seq, 12, cfg.vocab_size
    meta["reduced"] = layers < full_layers
    meta["params"] = sum(getattr(p, "ds_numel", p.numel()) for p in model.parameters())
    return engine, meta, local_seq, vocab


def main():
    args = parse()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    os.environ.setdefault("PYTORCH_CUDA_ALLOC_CONF", "expandable_segments:True")
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29512")
    os.environ.setdefault("RANK", "0"), os.environ.setdefault("WORLD_SIZE", "1"), os.environ.setdefault("LOCAL_RANK", "0")
    cuda = torch.cuda.is_available()
    if cuda:
        torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank) if cuda else torch.device("cpu")
    if not dist.is_initialized():
        dist.init_process_group("nccl" if cuda else "gloo", **({"device_id": dev} if cuda else {}))
    engine, meta, local_seq, vocab = build(args, world, dev)
    K, W, B = args.steps, args.warmup, args.mbs
    g = torch.Generator().manual_seed(1000 + (0 if args.config == "ulysses" else rank))
    host = [torch.randint(0, vocab, (B, local_seq if args.config != "ulysses" else local_seq * world), generator=g)
            for _ in range(W + K)]
    if args.config == "ulysses":  # every SP rank feeds its contiguous slice of the same sequence
        host = [h[:, rank * local_seq:(rank + 1) * local_seq].contiguous() for h in host]
    if cuda:
        host = [h.pin_memory() for h in host]

    def step(ids):
        if args.config == "ulysses":
            pos = torch.arange(rank * local_seq, (rank + 1) * local_seq, device=dev).unsqueeze(0).expand(B, -1)
            loss = engine(ids, labels=ids, position_ids=pos)
        else:
            loss = engine(ids, labels=ids)
        engine.backward(loss)
        engine.step()
        return loss

    for i in range(W):
        loss = step(host[i].to(dev, non_blocking=True))
    if cuda:
        torch.cuda.synchronize()
    dist.barrier()
    if cuda:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    t0 = time.perf_counter()
    for i in range(K):
        loss = step(host[W + i].to(dev, non_blocking=True))
        last = float(loss.detach())
    if cuda:
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    else:
        ms = (time.perf_counter() - t0) * 1e3
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t) / K
    if rank == 0:
        out = {"config": args.config, "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms,
               "tokens_per_s": meta["tokens_per_step"] / (ms / 1e3), "dtype": "bf16" if cuda else "fp32",
               "data": "synthetic random tokens, random-init weights", "final_loss": last,
               "activation_checkpointing": not args.no_ckpt,
               "peak_mem_gib": round(torch.cuda.max_memory_allocated() / 2**30, 2) if cuda else None, **meta}
        print(json.dumps(out))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
