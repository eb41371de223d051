#!/bin/bash
N=$1
if [ "$N" = "1" ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551"; fi
$L bench.py --impl reference --gpus $N --steps 8 --warmup 3 2>gpurun_out/final_ref_n$N.err | tail -1 > gpurun_out/final_ref_n$N.json
$L bench.py --gpus $N --steps 8 --warmup 3 2>gpurun_out/final_ours_n$N.err | tail -1 > gpurun_out/final_ours_n$N.json
python - <<PY
import json
for a in ("ref","ours"):
    try:
        d=json.load(open(f"gpurun_out/final_{a}_n$N.json"))
        print(a, d.get("n_gpus"), round(d["ms_per_step"],1), round(d["value"]), "e2e", round(d["e2e"]["value"]), d.get("step_ms"), d["clocks"].get("sm_mhz"))
    except Exception as e:
        print(a, "ERR", e, open(f"gpurun_out/final_{a}_n$N.json").read()[:300])
PY
