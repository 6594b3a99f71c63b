#!/bin/bash
cd "$(dirname "$0")/.."
o=gpurun_out/scale8e; mkdir -p $o
for cfg in "ctas8 --nccl-max-ctas 8" "ctas16 --nccl-max-ctas 16" "nogather --probes-gather none"; do
  set -- $cfg; name=$1; shift
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29871 bench.py --gpus 8 --workload probes --steps 48 --warmup 8 "$@" > $o/probes_$name.json 2> $o/probes_$name.err
  python - <<PY
import json
try:
    d=json.loads(open("$o/probes_$name.json").read().strip().splitlines()[-1]); print("probes N=8 $name:", round(d["value"],3), "Gpix/s", round(d["ms_per_step"],4), "ms  -> x", round(3.8118/d["ms_per_step"],3), "of the same build's N=1 (3.8118 ms)")
except Exception as e: print("$name failed", e); print(open("$o/probes_$name.err").read()[-500:])
PY
done
