#!/bin/bash
# Row-tiled 8K frame (BASELINE configs[4]) on N GPUs of one box: tools/scale_tiles.sh N -> gpurun_out/scale_tiles_N.json
mkdir -p gpurun_out
n=$1
if [ "$n" = "1" ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700 + n))"; fi
timeout 600 $L bench.py --gpus $n --workload tiles --steps 16 --warmup 4 $2 > gpurun_out/scale_tiles_$n.json 2> gpurun_out/scale_tiles_$n.err
tail -c 600 gpurun_out/scale_tiles_$n.err
cat gpurun_out/scale_tiles_$n.json
