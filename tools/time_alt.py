"""Device times of the earlier passes (SSAO, mirror SSR) on a synthetic frame. Usage: python tools/time_alt.py W H iters"""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "tests"), os.path.join(ROOT, "vk-renderer_b200")]
import torch  # noqa: E402

import chain  # noqa: E402
import vkr  # noqa: E402

w, h, iters = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
scene = vkr.synth_scene(0, 64)
cc = chain.Chain(w, h, chain.CudaBackend())
imgs, P = cc.render(scene)
t = P["trace"]
p = vkr.SsaoParams()
C.memmove(p.projection, vkr.perspective(t.fovy, t.aspect, t.znear, t.zfar), 64)
p.fovy, p.aspect, p.znear, p.zfar = t.fovy, t.aspect, t.znear, t.zfar
vkr.set_ssao_samples(p, vkr.sphere_samples(3))
sp = vkr.SsrParams()
C.memmove(sp.normal_mat, t.normal_mat, 64)
sp.fovy, sp.aspect, sp.znear, sp.zfar = t.fovy, t.aspect, t.znear, t.zfar
ao, refl = vkr.Image(w, h, vkr.R8_UNORM), vkr.Image(w, h, vkr.R8G8B8A8_UNORM)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
calls = {"SSAO": (lambda: vkr.ssao(p, imgs["depth"], ao), 5.0),                       # D24 read + R8 write
         "SSR": (lambda: vkr.ssr_legacy(sp, imgs["depth"], imgs["normal"], imgs["color"], imgs["material"], refl), 20.0)}
rows = []
for name, (fn, alg) in calls.items():
    for _ in range(3):
        fn()
    tot = 0.0
    for _ in range(iters):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    ms = tot / iters
    gbs = alg * w * h / (ms * 1e-3) / 1e9
    rows.append(dict(task=name, ms=round(ms, 4), alg_GBps=round(gbs, 1)))
    print(f"{name:8s} {ms:8.4f} ms  {gbs:8.1f} GB/s  {w*h/(ms*1e-3)/1e9:.2f} Gpix/s")
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(dict(w=w, h=h, iters=iters, rows=rows), open(os.path.join(ROOT, "gpurun_out", f"time_alt_{w}x{h}.json"), "w"), indent=1)
