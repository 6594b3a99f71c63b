#!/bin/bash
# Does a kernel boundary cost more when several GPUs of the box submit at the same time?  launch_gap alone, then 4 copies started
# together (one per GPU); then the chain at N=4 with one lane / no frame graph.
cd "$(dirname "$0")/.."
o=gpurun_out/scale4b; mkdir -p $o; rm -f /tmp/go_flag
echo "--- alone (GPU 0)"; CUDA_VISIBLE_DEVICES=0 tools/ubench/bin/launch_gap 12 | tail -1
echo "--- 4 copies started together"
for i in 0 1 2 3; do CUDA_VISIBLE_DEVICES=$i tools/ubench/bin/launch_gap 12 /tmp/go_flag > $o/gap_$i.txt & done
sleep 4; touch /tmp/go_flag; wait
for i in 0 1 2 3; do echo "GPU $i: $(tail -1 $o/gap_$i.txt)"; done
C="--steps 48 --warmup 4 --no-sub --no-e2e --no-cpu-baseline"
for opts in "graph.concurrent_branches=0" "graph.frame_graph=0" "graph.concurrent_branches=0,graph.frame_graph=0"; do
  BENCH_APP_OPTIONS=$opts timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29841 bench.py --gpus 4 $C > $o/b.json 2> $o/b.err
  python -c "
import json; d=json.loads(open('$o/b.json').read().strip().splitlines()[-1]); print('N=4 $opts: ms/step', round(d['ms_per_step'],4))"
done
BENCH_APP_OPTIONS="graph.concurrent_branches=0" python bench.py $C 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N=1 single lane: ms/step', round(d['ms_per_step'],4))"
