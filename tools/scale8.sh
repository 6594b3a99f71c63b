#!/bin/bash
# One 8-GPU call: PCIe concurrency at 1/2/4/8 GPUs, the default bench line (with its probes / tiles sub-records) at 8 and 4 GPUs.
cd "$(dirname "$0")/.."
o=gpurun_out/scale8; mkdir -p $o
nvidia-smi topo -m > $o/topo.txt 2>&1; lscpu | grep -iE "numa|^CPU\(s\)|Model name|Socket" >> $o/topo.txt
for n in 1 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) tools/pcie_bw_multi.py 2>/dev/null | tail -1 | tee $o/pcie_$n.json
done
for n in 8 4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700+n)) bench.py --gpus $n --steps 32 --warmup 4 > $o/bench_$n.json 2> $o/bench_$n.err
  echo "bench $n rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("$o/bench_$n.json").read().strip().splitlines()[-1])
    print($n, {k:d[k] for k in ("value","ms_per_step")}, "e2e", d["e2e"]["value"])
    for k in ("probes","tiles"): print("  ", k, {x:d[k][x] for x in ("value","ms_per_step","speedup_vs_n1")}, d[k]["n1_same_run"])
except Exception as e: print("parse failed", e)
PY
done
