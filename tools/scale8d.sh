#!/bin/bash
# 8-GPU: the default bench line (replicas of ONE sequence + tiles / probes sub-records), then the probe batch with the copy-engine gather.
cd "$(dirname "$0")/.."
o=gpurun_out/scale8d; mkdir -p $o
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29861 bench.py --gpus 8 --steps 48 --warmup 4 > $o/bench_8.json 2> $o/bench_8.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open("$o/bench_8.json").read().strip().splitlines()[-1])
print("chain N=8:", {k:round(d[k],4) for k in ("value","ms_per_step")}, "e2e", round(d["e2e"]["value"],3), "host enqueue", round(d["host_enqueue_ms_per_frame"],4))
for k in ("tiles","probes"): print(" ", k, {x:(round(d[k][x],4) if isinstance(d[k][x],float) else d[k][x]) for x in ("value","ms_per_step","speedup_vs_n1","per_rank_ms_per_step")}, d[k]["n1_same_run"])
PY
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29862 bench.py --gpus 8 --workload probes --probes-gather p2p --steps 48 --warmup 8 > $o/probes_p2p_8.json 2> $o/probes_p2p_8.err
python - <<PY
import json
d=json.loads(open("$o/probes_p2p_8.json").read().strip().splitlines()[-1])
print("probes p2p N=8:", round(d["value"],3), "Gpix/s", round(d["ms_per_step"],4), "ms", d.get("per_rank_ms_per_step"), "timeouts", d["gather_wait_timeouts"])
PY
