#!/bin/bash
# What costs the replica workload 10 % when >= 4 processes run at once?  (a) 8 independent single-GPU processes, no NCCL;
# (b) torchrun x8 with NCCL_NVLS_ENABLE=0; (c) torchrun x8 with NCCL_P2P_DISABLE=1.
cd "$(dirname "$0")/.."
o=gpurun_out/scale8c; mkdir -p $o
common="--steps 48 --warmup 4 --no-sub --no-e2e --no-cpu-baseline"
pids=""
for i in 0 1 2 3 4 5 6 7; do
  CUDA_VISIBLE_DEVICES=$i BENCH_CLOCKS=off python bench.py $common > $o/indep_$i.json 2> $o/indep_$i.err &
  pids="$pids $!"
done
for p in $pids; do wait $p; done
python - <<PY
import json
v=[json.loads(open("$o/indep_%d.json"%i).read().strip().splitlines()[-1])["ms_per_step"] for i in range(8)]
print("(a) 8 independent processes, no NCCL: ms/step per GPU", [round(x,4) for x in v])
PY
for cfg in "NCCL_NVLS_ENABLE=0" "NCCL_P2P_DISABLE=1"; do
  env $cfg timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29821 bench.py --gpus 8 $common > $o/bench8_$cfg.json 2> $o/bench8_$cfg.err
  python - <<PY
import json
d=json.loads(open("$o/bench8_$cfg.json").read().strip().splitlines()[-1])
print("$cfg", "ms/step", round(d["ms_per_step"],4), "Gpix/s", round(d["value"],2))
PY
done
