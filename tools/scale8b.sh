#!/bin/bash
# 8-GPU A/B of the NVML clock sampler's effect on the replica workload: off / rank 0 every 20 ms (default) / every rank every 5 ms (round 1)
cd "$(dirname "$0")/.."
o=gpurun_out/scale8b; mkdir -p $o
for mode in off rank0 all-ranks-5ms; do
  BENCH_CLOCKS=$mode timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29811 bench.py --gpus 8 --steps 64 --warmup 4 --no-sub --no-e2e > $o/bench8_$mode.json 2> $o/bench8_$mode.err
  python - <<PY
import json
d=json.loads(open("$o/bench8_$mode.json").read().strip().splitlines()[-1])
print("$mode", "ms/step", round(d["ms_per_step"],4), "Gpix/s", round(d["value"],2), "clocks", d["clocks"], d["roofline"]["per_task_source"][60:130])
PY
done
