#!/bin/bash
cd "$(dirname "$0")/.."
o=gpurun_out/scale4c; mkdir -p $o; rm -f /tmp/go_flag2
echo "--- alone (GPU 0)"; CUDA_VISIBLE_DEVICES=0 tools/ubench/bin/launch_gap 8 | tail -2
echo "--- 4 copies started together"
for i in 0 1 2 3; do CUDA_VISIBLE_DEVICES=$i tools/ubench/bin/launch_gap 8 /tmp/go_flag2 > $o/gap_$i.txt & done
sleep 4; touch /tmp/go_flag2; wait
for i in 0 1 2 3; do echo "GPU $i: $(tail -2 $o/gap_$i.txt | head -1)"; done
C="--steps 64 --warmup 4 --no-sub --no-e2e --no-cpu-baseline"
BENCH_STAGGER_S=0.6 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29851 bench.py --gpus 4 $C > $o/stagger.json 2> $o/stagger.err
python -c "
import json; d=json.loads(open('$o/stagger.json').read().strip().splitlines()[-1]); print('N=4, timed regions staggered by 0.6 s per rank: ms/step (max over ranks)', round(d['ms_per_step'],4))"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29852 bench.py --gpus 4 $C > $o/sync.json 2> $o/sync.err
python -c "
import json; d=json.loads(open('$o/sync.json').read().strip().splitlines()[-1]); print('N=4, common start: ms/step', round(d['ms_per_step'],4), 'host enqueue', round(d['host_enqueue_ms_per_frame'],4))"
