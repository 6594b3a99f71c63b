#!/bin/bash
# 4-GPU matrix: which part of an initialised NCCL communicator costs the replica workload its 10 %, and does the probe batch
# do better with the copy-engine peer gather when no NCCL communicator exists at all?
cd "$(dirname "$0")/.."
o=gpurun_out/scale4; mkdir -p $o
run() {  # name, env..., -- bench args
  name=$1; shift; envs=""; while [ "$1" != "--" ]; do envs="$envs $1"; shift; done; shift
  env $envs timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29831 bench.py --gpus 4 "$@" > $o/$name.json 2> $o/$name.err
  python - <<PY
import json
try:
    d=json.loads(open("$o/$name.json").read().strip().splitlines()[-1]); print("$name", "ms/step", round(d["ms_per_step"],4), d["unit"], round(d["value"],2))
except Exception as e: print("$name failed", e); print(open("$o/$name.err").read()[-600:])
PY
}
C="--steps 48 --warmup 4 --no-sub --no-e2e --no-cpu-baseline"
run chain_gloo -- $C
run chain_nccl BENCH_CONTROL=nccl -- $C
run chain_nccl_nocumem BENCH_CONTROL=nccl NCCL_CUMEM_ENABLE=0 -- $C
run chain_nccl_nowatchdog BENCH_CONTROL=nccl TORCH_NCCL_ENABLE_MONITORING=0 TORCH_NCCL_ASYNC_ERROR_HANDLING=0 -- $C
run chain_nccl_1chan BENCH_CONTROL=nccl NCCL_MAX_NCHANNELS=1 NCCL_MIN_NCHANNELS=1 -- $C
P="--workload probes --steps 48 --warmup 8"
run probes_allgather -- $P
run probes_p2p_no_nccl -- $P --probes-gather p2p
run probes_nogather -- $P --probes-gather none
