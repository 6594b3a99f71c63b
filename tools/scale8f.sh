#!/bin/bash
# 8 GPUs, probe batch: copy-engine peer gather with 2 / 4 / 8 landing buffers per tensor (the coupling between ranks loosens with more).
cd "$(dirname "$0")/.."
o=gpurun_out/scale8f; mkdir -p $o
for nb in 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29880+nb)) bench.py --gpus 8 --workload probes --probes-gather p2p --probes-gather-buffers $nb --steps 64 --warmup 8 > $o/probes_p2p_nb$nb.json 2> $o/probes_p2p_nb$nb.err
  python - <<PY
import json
try:
    d=json.loads(open("$o/probes_p2p_nb$nb.json").read().strip().splitlines()[-1]); print("probes N=8 p2p, $nb buffers:", round(d["value"],3), "Gpix/s", round(d["ms_per_step"],4), "ms  per rank", d["per_rank_ms_per_step"], "timeouts", d["gather_wait_timeouts"])
except Exception as e: print("nb$nb failed", e); print(open("$o/probes_p2p_nb$nb.err").read()[-600:])
PY
done
