#!/bin/bash
# 8 GPUs, probe batch: the gather fused into the blur / accumulate kernels (peer stores) against the NCCL all-gather and no gather,
# same build, same box; the 1-GPU point of the same build first.
cd "$(dirname "$0")/.."
o=gpurun_out/scale8g; mkdir -p $o
python bench.py --workload probes --steps 32 --warmup 8 > $o/probes_n1.json 2> $o/probes_n1.err
for mode in fused allgather none; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29891 bench.py --gpus 8 --workload probes --probes-gather $mode --steps 64 --warmup 8 > $o/probes_$mode.json 2> $o/probes_$mode.err
done
python - <<PY
import json
def load(n):
    try: return json.loads(open("$o/probes_%s.json" % n).read().strip().splitlines()[-1])
    except Exception as e:
        print(n, "failed:", e); print(open("$o/probes_%s.err" % n).read()[-800:]); return None
n1 = load("n1")
print("N=1:", round(n1["ms_per_step"], 4), "ms", round(n1["value"], 3), "Gpix/s")
for m in ("fused", "allgather", "none"):
    d = load(m)
    if d: print("N=8 %-9s: %.4f ms  %.3f Gpix/s  = %.3fx of N=1   per rank %s  timeouts %s" % (m, d["ms_per_step"], d["value"], d["value"] / n1["value"], d["per_rank_ms_per_step"], d["gather_wait_timeouts"]))
PY
