"""Per-task device times of the CUDA chain (CUDA events on the launch stream). Usage: python tools/time_passes.py W H frames"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "tests"), os.path.join(ROOT, "vk-renderer_b200")]
import torch  # noqa: E402

import chain  # noqa: E402
import vkr  # noqa: E402

w, h, frames = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
scene = vkr.synth_scene(0, 64)
cc = chain.Chain(w, h, chain.CudaBackend())
N = w * h
ALG = {"DownsampleGbuffer+DownsampleDepth": 10.67, "SSSR_trace": 10.33, "SSSR_filter": 16, "SSSR_blur": 14, "GTAO_main": 13, "GTAO_filter": 3.5,
       "GTAO_accumulate": 5.5, "DeferedShading": 23, "TAA": 32}   # algorithmic bytes per full-res pixel, SURVEY §8d
tot = {}
warm = 3
for f in range(frames + warm):
    P = cc.begin_frame(scene)
    evs = []
    for name, thunk, _ in cc.steps(P):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); thunk(); b.record()
        evs.append((name, a, b))
    torch.cuda.synchronize()
    if f >= warm:
        for name, a, b in evs:
            tot[name] = tot.get(name, 0.0) + a.elapsed_time(b)
    cc.end_frame()
rows = []
s = 0.0
for name, t in tot.items():
    ms = t / frames
    s += ms
    gbs = ALG[name] * N / (ms * 1e-3) / 1e9
    rows.append(dict(task=name, ms=round(ms, 4), alg_GBps=round(gbs, 1), frac_of_6544=round(gbs / 6544.3, 3)))
    print(f"{name:36s} {ms:8.4f} ms  {gbs:8.1f} GB/s  {gbs/6544.3:6.3f}")
print(f"{'chain':36s} {s:8.4f} ms  {128*N/(s*1e-3)/1e9:8.1f} GB/s  {128*N/(s*1e-3)/1e9/6544.3:6.3f}   {N/(s*1e-3)/1e9:.2f} Gpix/s")
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(dict(w=w, h=h, frames=frames, rows=rows, chain_ms=s), open(os.path.join(ROOT, "gpurun_out", f"time_passes_{w}x{h}.json"), "w"), indent=1)
