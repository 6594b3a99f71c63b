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
import numpy as np  # noqa: E402
import vkr_probes  # noqa: E402
F16 = vkr.R16G16B16A16_SFLOAT
st = vkr.ScreenTraceParams()
C.memmove(st.normal_mat, t.normal_mat, 64)
st.random_offset, st.angle_offset, st.fovy, st.aspect, st.znear, st.zfar = 0.37, 0.3, t.fovy, t.aspect, t.znear, t.zfar
st_raw, st_f, st_a = vkr.Image(w, h, F16), vkr.Image(w, h, F16), vkr.Image(w, h, F16)
cam = np.array([t.fovy, t.aspect, t.znear, t.zfar], dtype=np.float32)
gp = vkr.GtaoParams()
C.memmove(gp.normal_mat, t.normal_mat, 64)
gp.fovy, gp.aspect, gp.znear, gp.zfar, gp.angle_offset = t.fovy, t.aspect, t.znear, t.zfar, 0.21
w2, h2 = w // 2, h // 2
gv_raw = vkr.Image(w2, h2, F16)
gv_arr = vkr.ImageArray(w2 // 4, h2 // 4, vkr.R32_SFLOAT, 1, layers=16)
r16 = [vkr.Image(w2, h2, vkr.R16_SFLOAT) for _ in range(3)]
grid = vkr_probes.OctahedralProbeGrid(4, 256)
grid.min, grid.max = (-3.0, 1.0, -5.0), (3.0, 1.0, 5.0)
cubes = vkr_probes.CubeMapSet(16, 128).render_synthetic(scene, grid.positions())
pr, ptp = vkr_probes.ProbeRenderer(128), vkr_probes.ProbeTracePass()
pparams = vkr_probes.ProbeTraceParams(vkr.mat_inverse(P["shade"].camera), t.fovy, t.aspect, t.znear, t.zfar)
probe_out = vkr.Image(w, h, vkr.R8G8B8A8_UNORM)
pr.render_octahedral(cubes, grid.color_array, grid.depth_array, 0, 16)
pr.probe_downsample(grid.depth_array, 0, 16)
NP = 16 * 256 * 256
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
calls = {"SSAO": (lambda: vkr.ssao(p, imgs["depth"], ao), 5.0),                       # D24 read + R8 write
         "SSR": (lambda: vkr.ssr_legacy(sp, imgs["depth"], imgs["normal"], imgs["color"], imgs["material"], refl), 20.0),
         "ScreenTrace": (lambda: vkr.screen_trace_main(st, imgs["depth"], imgs["normal"], imgs["color"], imgs["material"], st_raw), 24.0),
         "ScreenTraceFilter": (lambda: vkr.screen_trace_filter(t.znear, t.zfar, st_raw, imgs["depth"], st_f), 20.0),
         "ScreenTraceAccumulate": (lambda: vkr.screen_trace_accumulate(cam, imgs["depth"], imgs["prev_depth"], st_f, st_a), 32.0),
         "GTAO(frag, half-res)": (lambda: vkr.gtao_main_graphics(gp, imgs["depth"], imgs["normal"], gv_raw), (4 + 4 * 4 + 8) / 4),
         "GTAO_reproject(half-res)": (lambda: vkr.gtao_reproject(cam, imgs["depth"], imgs["prev_depth"], r16[0], r16[1], r16[2]), (8 + 6) / 4),
         "Cubemap2Octahedral x16": (lambda: pr.render_octahedral(cubes, grid.color_array, grid.depth_array, 0, 16), None),
         "DownsampleProbe x16": (lambda: pr.probe_downsample(grid.depth_array, 0, 16), None),
         "TraceProbe": (lambda: ptp.run(grid, imgs["depth"], imgs["normal"], probe_out, pparams), 12.0)}
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
    if alg is None:                    # probe conversion: 16 probes of 256^2, 6 B/texel out + cube reads / 1.33 * 2 B pyramid
        alg_bytes = NP * (6 + 6) if "Cubemap" in name else NP * 2 * (1 + 1 / 3)
    else:
        alg_bytes = alg * w * h
    gbs = alg_bytes / (ms * 1e-3) / 1e9
    rows.append(dict(task=name, ms=round(ms, 4), alg_GBps=round(gbs, 1)))
    print(f"{name:28s} {ms:8.4f} ms  {gbs:8.1f} GB/s")
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(dict(w=w, h=h, iters=iters, rows=rows), open(os.path.join(ROOT, "gpurun_out", f"time_alt_{w}x{h}.json"), "w"), indent=1)
