#!/usr/bin/env python
"""Headline benchmark: Llama-3-8B, ZeRO-3, bf16, tokens/sec per box at N GPUs (BASELINE.json metric).

Contract (driver): ``python bench.py --gpus N --steps K --warmup W`` (N>1 via torch.distributed.run). Prints ONE
JSON line on rank 0. ``--impl reference`` runs the unmodified reference (``baseline/_ref``) on a Hugging Face
Llama of the same architecture through ``deepspeed.initialize`` for the same metric/config.

Two timed regions, K steps each, both bracketed by barrier + cuda synchronize, max over ranks:
  * ``value``: device-timed (CUDA events), inputs already resident on the device;
  * ``e2e``  : through the public API with, every step, a pinned-host -> device copy of that step's tokens and a
               device -> host read of the loss.
Synthetic data (random tokens), random-init weights. Inputs of consecutive steps are different buffers and every
step streams >126 MB (weights+optimizer state are >100 GB), so no L2 flush is needed ("inputs larger than L2").
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--seq", type=int, default=4096)
    ap.add_argument("--mbs", type=int, default=1)
    ap.add_argument("--layers", type=int, default=32, help="debug only: fewer layers => INVALID headline number")
    ap.add_argument("--model", default="llama3-8b")
    ap.add_argument("--zero", type=int, default=3)
    ap.add_argument("--act-ckpt", default="auto",
                    help="activation checkpointing: auto (only as many layers as HBM requires; same rule for both "
                    "arms), on, off, or an integer number of checkpointed layers (ours only)")
    ap.add_argument("--no-ckpt", action="store_true", help="same as --act-ckpt off")
    ap.add_argument("--local_rank", type=int, default=-1)
    return ap.parse_args()


class ClockSampler:
    """Samples SM clocks / throttle reasons of THIS rank's GPU while the timed region runs.

    NVML is queried in-process (pynvml, one device handle) from a background thread. Spawning ``nvidia-smi -lms``
    instead attaches every GPU of the box (persistence mode is off on the bench boxes) and stalled one or two timed
    steps by 0.2-1 s, seconds after the spawn (measured: a 950 ms and a 795 ms step inside a 10-step region at 2 GPUs);
    it is kept only as the fallback when pynvml is missing."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, gpu_index=0, period_s=None):
        if period_s is None:
            period_s = float(os.environ.get("DSB200_BENCH_SAMPLE_PERIOD", "0.5"))
        self.rows, self.proc, self.gpu_index, self.period = [], None, gpu_index, period_s
        self.t_region = 0.0
        self.nvml, self.handle, self._stop = None, None, False
        self.t = None

    def mark_region_start(self):
        """Samples before this moment (warm-up; start-up of the sampler) are discarded."""
        self.t_region = time.time()

    def _physical_index(self):
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:
            try:
                return int(vis.split(",")[self.gpu_index])
            except Exception:
                return self.gpu_index
        return self.gpu_index

    def start(self):
        if os.environ.get("DSB200_BENCH_NO_SAMPLER", "0") == "1":
            return
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(self._physical_index())
            self.max_sm = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self._physical_index())], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _poll(self):
        n = self.nvml
        bits = [n.nvmlClocksEventReasonHwSlowdown, n.nvmlClocksEventReasonHwThermalSlowdown,
                n.nvmlClocksEventReasonSwThermalSlowdown, n.nvmlClocksEventReasonSwPowerCap]
        while not self._stop:
            try:
                sm = float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM))
                mask = int(n.nvmlDeviceGetCurrentClocksEventReasons(self.handle))
                self.rows.append((time.time(), [self.gpu_index, sm, self.max_sm, 0.0, mask] +
                                  ["Active" if mask & b else "Not Active" for b in bits]))
            except Exception:
                pass
            time.sleep(self.period)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def stop(self):
        if self.nvml is not None:
            self._stop = True
            self.t.join(timeout=2)
        elif self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        else:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["clock sampler unavailable"]}
        sm, mx, reasons = [], [], set()
        for ts, r in self.rows:
            if ts < self.t_region:
                continue
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                for n, v in zip(self.NAMES, r[5:9]):
                    if str(v).lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm),
                "source": "nvml (in-process)" if self.nvml is not None else "nvidia-smi -lms 200"}


def model_dims(args):
    return dict(vocab_size=128256, hidden_size=4096, intermediate_size=14336, num_hidden_layers=args.layers,
                num_attention_heads=32, num_key_value_heads=8, max_position_embeddings=8192, rms_norm_eps=1e-5,
                rope_theta=500000.0)


def ckpt_layers(args, world, impl, total_mem_bytes):
    """How many transformer layers recompute their activations in backward.

    Recomputation is pure overhead (one extra forward = 1/4 of the GEMM work), so it is used only where the model
    states leave no room: per rank they take ~17 B/param / world (fp32 master + 2 moments + bf16 weights + grads, plus
    gather buffers). One layer's saved activations are ~150 KB/token with the fused kernels (ours) and ~3x that with
    the stock HF layer (reference), which can only checkpoint all layers or none."""
    mode = "off" if args.no_ckpt else args.act_ckpt
    if mode == "on":
        return args.layers
    if mode == "off":
        return 0
    if mode != "auto":
        return max(0, min(args.layers, int(mode)))
    tokens = args.mbs * args.seq
    params = 8.03e9 * args.layers / 32
    states = 17.0 * params / world + 6e9           # + gathered units, logits chunk, allocator slack
    per_layer = tokens * (150e3 if impl == "ours" else 450e3)
    budget = 0.90 * total_mem_bytes - states
    if impl == "reference":
        return 0 if budget > per_layer * args.layers else args.layers
    keep = int(max(0, budget) // per_layer)
    return max(0, args.layers - keep)


def ds_config(args, world):
    cfg = _ds_config(args, world)
    if os.environ.get("DSB200_BENCH_DEEPCOMPILE", "0") == "1" and args.impl == "ours":
        # extra (non-headline) configuration: profile-guided unit plan of the ZeRO-3 executor (needs warmup >= 4)
        cfg["compile"] = {"deepcompile": True}
    return cfg


def _ds_config(args, world):
    return {
        "train_micro_batch_size_per_gpu": args.mbs,
        "gradient_accumulation_steps": 1,
        "steps_per_print": 10**9,
        "bf16": {"enabled": True},
        "gradient_clipping": 1.0,
        "optimizer": {"type": "AdamW", "params": {"lr": 1e-5, "betas": [0.9, 0.95], "eps": 1e-8, "weight_decay": 0.1}},
        "zero_optimization": {"stage": args.zero, "overlap_comm": True, "contiguous_gradients": True,
                              "reduce_bucket_size": 5e8, "stage3_prefetch_bucket_size": 5e8,
                              "stage3_param_persistence_threshold": 1e5, "stage3_max_live_parameters": 1e9,
                              "stage3_max_reuse_distance": 1e9},
    }


def build_ours(args, torch, world):
    sys.path.insert(0, REPO)
    import deepspeed_b200 as ds
    from deepspeed_b200.models import LlamaConfig, LlamaForCausalLM
    dev = torch.device("cuda", torch.cuda.current_device())
    args.n_ckpt = ckpt_layers(args, world, "ours", torch.cuda.get_device_properties(dev).total_memory)
    cfg = LlamaConfig(**model_dims(args), gradient_checkpointing=args.n_ckpt, dtype=torch.bfloat16)
    torch.manual_seed(1234)
    model = LlamaForCausalLM(cfg, device=dev)
    engine, _, _, _ = ds.initialize(model=model, config=ds_config(args, world))
    if os.environ.get("DSB200_BENCH_DEEPCOMPILE", "0") == "1":
        engine.compile()
        args._dc_engine = engine

    def step(ids):
        loss = engine(ids, labels=ids)
        engine.backward(loss)
        engine.step()
        return loss

    def launches():
        from deepspeed_b200.ops import kernels
        return kernels.launch_count()

    return step, launches, cfg.num_params()


def build_reference(args, torch, world):
    ref = os.path.join(REPO, "baseline", "_ref")
    if not os.path.isdir(os.path.join(ref, "deepspeed")):
        return None, "baseline/_ref missing: reference not installed"
    sys.path.insert(0, ref)
    # the reference's op_builder cannot parse an "a"-suffixed arch ("10.0a") that the image may export
    os.environ["TORCH_CUDA_ARCH_LIST"] = "10.0"
    os.environ.setdefault("DS_SKIP_CUDA_CHECK", "1")
    try:
        import deepspeed
        from transformers import LlamaConfig, LlamaForCausalLM
    except Exception as e:  # noqa: BLE001
        return None, f"reference import failed: {type(e).__name__}: {e}"
    dims = model_dims(args)
    hf_cfg = LlamaConfig(**dims, tie_word_embeddings=False, attention_bias=False, use_cache=False,
                         attn_implementation="sdpa")
    dev = torch.device("cuda", torch.cuda.current_device())

    def make_model():
        torch.manual_seed(1234)
        old = torch.get_default_dtype()
        torch.set_default_dtype(torch.bfloat16)
        try:
            with torch.device(dev):
                m = LlamaForCausalLM(hf_cfg)
        finally:
            torch.set_default_dtype(old)
        if args.n_ckpt:
            m.gradient_checkpointing_enable(gradient_checkpointing_kwargs={"use_reentrant": False})
        m.train()
        return m

    args.n_ckpt = ckpt_layers(args, world, "reference", torch.cuda.get_device_properties(dev).total_memory)
    cfg = ds_config(args, world)
    model = make_model()
    try:
        engine, _, _, _ = deepspeed.initialize(model=model, model_parameters=model.parameters(), config=cfg)
    except Exception as e:  # FusedAdam JIT build can fail offline -> the reference's own torch_adam switch
        import traceback
        print(f"[bench] reference initialize failed with its fused optimizer ({type(e).__name__}: {e}); "
              f"retrying with optimizer.params.torch_adam=true", file=sys.stderr)
        traceback.print_exc()
        del model
        torch.cuda.empty_cache()
        cfg["optimizer"]["params"]["torch_adam"] = True
        model = make_model()
        engine, _, _, _ = deepspeed.initialize(model=model, model_parameters=model.parameters(), config=cfg)

    def step(ids):
        out = engine(input_ids=ids, labels=ids)
        loss = out.loss
        engine.backward(loss)
        engine.step()
        return loss

    return (step, lambda: 0, sum(p.ds_numel if hasattr(p, "ds_numel") else p.numel() for p in model.parameters())), None


def main():
    args = parse()
    world = int(os.environ.get("WORLD_SIZE", 1))
    rank = int(os.environ.get("RANK", 0))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference" and not os.path.isdir(os.path.join(REPO, "baseline", "_ref", "deepspeed")):
        if world == 1:  # a fresh checkout lacks the git-ignored install; redo it when the reference source is here
            try:
                sys.path.insert(0, REPO)
                import __graft_entry__
                __graft_entry__.ensure_reference(verbose=False)
            except Exception:  # noqa: BLE001
                pass
    if args.impl == "reference" and not os.path.isdir(os.path.join(REPO, "baseline", "_ref", "deepspeed")):
        print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref not present (reference not installed)"}))
        return 0
    # both arms: let the caching allocator grow segments in place (no fragmentation stalls close to the HBM limit)
    os.environ.setdefault("PYTORCH_CUDA_ALLOC_CONF", "expandable_segments:True")
    import torch
    if not torch.cuda.is_available():
        print(json.dumps({"impl": args.impl, "unavailable": "no CUDA device"}))
        return 0
    torch.cuda.set_device(local_rank)
    import torch.distributed as dist
    if world > 1 or "RANK" in os.environ:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
    if not dist.is_initialized():
        os.environ.setdefault("RANK", "0"), os.environ.setdefault("WORLD_SIZE", "1")
        os.environ.setdefault("LOCAL_RANK", "0"), os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    assert world == args.gpus or world == 1, f"--gpus {args.gpus} but WORLD_SIZE={world}"

    if args.impl == "ours":
        step, launches, nparams = build_ours(args, torch, world)
    else:
        built, err = build_reference(args, torch, world)
        if built is None:
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": err}))
            return 0
        step, launches, nparams = built

    dev = torch.device("cuda", local_rank)
    K, W, B, S = args.steps, args.warmup, args.mbs, args.seq
    g = torch.Generator().manual_seed(1000 + rank)
    n_buf = W + 2 * K
    host = [torch.randint(0, 128256, (B, S), generator=g, dtype=torch.int64).pin_memory() for _ in range(n_buf)]
    dev_inputs = [h.to(dev) for h in host[W:W + K]]

    def sync_all():
        dist.barrier()
        torch.cuda.synchronize()

    # the clock sampler is started BEFORE the warm-up so that spawning nvidia-smi (NVML start-up, device enumeration)
    # cannot stall a timed step; only samples taken inside the timed regions are reported
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for i in range(W):
        loss = step(host[i].to(dev, non_blocking=True))
    float(loss)
    sync_all()
    sampler.mark_region_start()
    # ---- region 1: device-timed ----
    l0 = launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    e0.record()
    for i in range(K):
        loss = step(dev_inputs[i])
        marks[i].record()
    e1.record()
    sync_all()
    dev_ms = e0.elapsed_time(e1)
    step_ms = [round((e0 if i == 0 else marks[i - 1]).elapsed_time(marks[i]), 2) for i in range(K)]
    n_launch = launches() - l0
    last_loss = float(loss)
    # ---- region 2: end to end (pinned H2D each step + loss D2H each step) ----
    sync_all()
    t0 = time.perf_counter()
    for i in range(K):
        ids = host[W + K + i].to(dev, non_blocking=True)
        loss = step(ids)
        last_loss = loss.item()
    sync_all()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.stop() if rank == 0 else None

    t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms = float(t[0]), float(t[1])
    tokens = world * B * S * K
    value = tokens / (dev_ms / 1e3)
    e2e_value = tokens / (e2e_ms / 1e3)
    peak_mem = torch.cuda.max_memory_allocated() / 2**30
    if rank == 0:
        baseline = None
        try:
            with open(os.path.join(REPO, "BASELINE.json")) as f:
                pub = json.load(f).get("published", {})
            baseline = pub.get("tokens_per_sec") if isinstance(pub, dict) else None
        except Exception:
            pass
        out = {
            "metric": "tokens/sec/box Llama-3-8B ZeRO-3 (max over ranks)",
            "impl": args.impl,
            "value": value,
            "unit": "tokens/s",
            "n_gpus": world,
            "steps": K,
            "warmup": W,
            "ms_per_step": dev_ms / K,
            "higher_is_better": True,
            "scaling": "weak",
            "vs_baseline": (value / baseline) if baseline else None,
            "dtype": "bf16",
            "data": "synthetic random tokens, random-init weights",
            "config": {"model": f"Llama-3-8B ({args.layers} layers, h4096, 32q/8kv heads, ffn14336, vocab128256)",
                       "params": nparams, "global_batch": world * B, "seq_len": S,
                       "parallelism": f"zero{args.zero}-dp{world}", "activation_checkpointing": f"{args.n_ckpt}/{args.layers} layers",
                       "optimizer": "AdamW (fused)", "grad_clip": 1.0,
                       "l2": "inputs larger than L2 (per-step working set >100 GB), no flush needed"},
            "e2e": {"value": e2e_value, "unit": "tokens/s", "ms_per_step": e2e_ms / K,
                    "h2d_bytes_per_step": B * S * 8, "d2h_bytes_per_step": 4},
            "gpu_launches": n_launch,
            "clocks": clocks,
            "step_ms": step_ms,
            "final_loss": last_loss,
            "peak_mem_gib": round(peak_mem, 2),
        }
        dc = getattr(getattr(args, "_dc_engine", None), "_dc_driver", None)
        if dc is not None:
            out["config"]["deepcompile"] = ("no plan yet (warmup too short)" if dc.plan is None else {
                "resident_units": len(dc.plan.resident), "kept_forward_to_backward": len(dc.plan.keep),
                "predicted_exposed_gather_ms": round(dc.plan.predicted["plan_exposed_us"] / 1e3, 2),
                "comm_model": f"{dc.profile.comm_latency_us:.0f} us + {dc.profile.comm_bytes_per_us / 1e3:.0f} GB/s"})
        if args.layers != 32:
            out["INVALID"] = "reduced layer count (debug run)"
        print(json.dumps(out))
    dist.barrier()
    return 0


if __name__ == "__main__":
    sys.exit(main())
