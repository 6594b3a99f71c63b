#!/bin/bash
# Scaling runs on one box: tools/scale_run.sh "1 2 4 8" -> gpurun_out/scale_{probes,chain}_N.json
mkdir -p gpurun_out
for n in $1; do
  if [ "$n" = "1" ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n))"; fi
  $L bench.py --gpus $n --workload probes --steps 16 --warmup 4 > gpurun_out/scale_probes_$n.json 2> gpurun_out/scale_probes_$n.err
  $L bench.py --gpus $n --steps 16 --warmup 4 --no-cpu-baseline > gpurun_out/scale_chain_$n.json 2> gpurun_out/scale_chain_$n.err
  tail -c 400 gpurun_out/scale_probes_$n.err
done
grep -h -o '"value": [0-9.]*, "unit": "Gpix/s", [^}]*"ms_per_step": [0-9.]*' gpurun_out/scale_*.json
