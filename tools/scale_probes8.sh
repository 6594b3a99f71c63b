#!/bin/bash
# 8-GPU probe-batch runs on one box: N=1, then N=8 with the NCCL all_gather and with the copy-engine peer gather.
#   tools/scale_probes8.sh  -> gpurun_out/probes8_<tag>.json
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
python bench.py --gpus 1 --workload probes --steps 64 --warmup 4 > gpurun_out/probes8_n1.json 2> gpurun_out/probes8_n1.err
$T --master-port 29611 bench.py --gpus 8 --workload probes --steps 64 --warmup 4 > gpurun_out/probes8_allgather.json 2> gpurun_out/probes8_allgather.err
$T --master-port 29612 bench.py --gpus 8 --workload probes --steps 64 --warmup 4 --probes-gather p2p > gpurun_out/probes8_p2p.json 2> gpurun_out/probes8_p2p.err
for t in n1 allgather p2p; do echo "$t: $(grep -h -o '"value": [0-9.]*, "unit": "Gpix/s"' gpurun_out/probes8_$t.json) $(grep -h -o '"ms_per_step": [0-9.]*' gpurun_out/probes8_$t.json) $(grep -h -o '"gather_wait_timeouts": [0-9]*' gpurun_out/probes8_$t.json) $(grep -v "OMP_NUM_THREADS\|\*\*\*\*" gpurun_out/probes8_$t.err | tail -c 300 | tr '\n' ' ')"; done
