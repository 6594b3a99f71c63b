"""Pinned-host <-> device copy bandwidth with N GPUs copying AT THE SAME TIME (the ceiling of bench.py's e2e number at N GPUs;
VERDICT r01 next-round item 9).  One process per GPU:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/pcie_bw_multi.py [MB]
Every rank copies MB (default 166 = one 4K G-buffer) H2D, 0.4 * MB D2H (one TAA target), and both at once, 10 times after a
barrier; rank 0 prints one JSON line with per-rank and aggregate GB/s (aggregate = total bytes / slowest rank's time)."""
import json
import os
import sys

import torch
import torch.distributed as dist

mb = int(sys.argv[1]) if len(sys.argv) > 1 else 166
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", rank))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n_in, n_out = mb * 1000 * 1000, int(mb * 0.4) * 1000 * 1000
h_in, d_in = torch.empty(n_in, dtype=torch.uint8).pin_memory(), torch.empty(n_in, dtype=torch.uint8, device="cuda")
h_out, d_out = torch.empty(n_out, dtype=torch.uint8).pin_memory(), torch.empty(n_out, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def up():
    with torch.cuda.stream(s1):
        d_in.copy_(h_in, non_blocking=True)


def down():
    with torch.cuda.stream(s2):
        h_out.copy_(d_out, non_blocking=True)


def timed(fns, reps=10):
    for f in fns:
        f()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    cur = torch.cuda.current_stream()
    a.record()
    s1.wait_event(a); s2.wait_event(a)
    for _ in range(reps):
        for f in fns:
            f()
    cur.wait_stream(s1); cur.wait_stream(s2)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / reps
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        g = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(g, t)
        return [float(x.item()) for x in g]
    return [ms]


res = {}
for name, fns, nbytes in (("h2d", [up], n_in), ("d2h", [down], n_out), ("both", [up, down], n_in + n_out)):
    ms = timed(fns)
    res[name] = {"per_rank_GBps": [round(nbytes / t / 1e6, 2) for t in ms], "aggregate_GBps": round(world * nbytes / max(ms) / 1e6, 2)}
if rank == 0:
    numa = ""
    try:
        numa = open("/sys/devices/system/node/online").read().strip()
    except OSError:
        pass
    print(json.dumps({"n_gpus": world, "h2d_MB": mb, "d2h_MB": int(mb * 0.4), "numa_nodes_online": numa, "host_cores": len(os.sched_getaffinity(0)), **res}))
if world > 1:
    dist.destroy_process_group()
