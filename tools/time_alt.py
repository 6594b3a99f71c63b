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
# SSSR tile path + capture path
ntiles = ((w2 + 7) // 8) * ((h2 + 7) // 8)
t_r, t_g = torch.zeros(ntiles, dtype=torch.int32, device="cuda"), torch.zeros(ntiles, dtype=torch.int32, device="cuda")
i_r, i_g = torch.zeros(3, dtype=torch.int32, device="cuda"), torch.zeros(3, dtype=torch.int32, device="cuda")
cls = vkr.ClassifyParams(w2, h2, 1.0, 0.5)
hal_dev = torch.from_numpy(vkr.halton23_seq(128)).cuda()
tile_rays = vkr.Image(w2, h2, vkr.R16G16B16A16_UNORM)
planes = vkr.Image(max(1, w // 16), max(1, h // 16), vkr.R32G32B32A32_SFLOAT)
rg = vkr.RegressionParams()
C.memmove(rg.camera_to_world, vkr.mat_transpose(t.normal_mat), 64)
rg.fovy, rg.aspect, rg.znear, rg.zfar = t.fovy, t.aspect, t.znear, t.zfar
rb_img = vkr.Image(w, h, vkr.R8G8B8A8_SRGB)
rb_tight = torch.zeros(w * h * 4, dtype=torch.uint8, device="cuda")


def classify():
    vkr.sssr_clear_indirect(i_r, i_g)
    vkr.sssr_classification(cls, imgs["material"], t_r, t_g, i_r, i_g)


classify()
torch.cuda.synchronize()
print("tiles: reflective", int(i_r[0]), "glossy", int(i_g[0]), "of", ntiles)
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
         "SSSR_Clear+Classification": (classify, 1.0 + 2 * 4 / 256),          # material at half-res corners: every full-res texel once
         "SSSR_trace(mirror tiles)": (lambda: vkr.sssr_trace_indirect(t, 0, imgs["depth"], imgs["dn"], imgs["material"], hal_dev, t_r, i_r, tile_rays), None),
         "SSSR_trace(glossy tiles)": (lambda: vkr.sssr_trace_indirect(t, 1, imgs["depth"], imgs["dn"], imgs["material"], hal_dev, t_g, i_g, tile_rays), None),
         "SSSR_Tile_Regression": (lambda: vkr.sssr_tile_regression(rg, imgs["depth"], planes), 1.0 + 16.0 / 256),
         "CopyImage(TAA->sRGB8)": (lambda: vkr.blit_image(imgs["taa"], rb_img), 12.0),
         "ImageRead(pack RGBA8)": (lambda: vkr.readback_pack(rb_img, rb_tight, 0, vkr.READBACK_OPAQUE_ALPHA), 8.0),
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
    if alg is None and name.startswith("SSSR_trace"):   # share of the per-pixel trace's 10.33 B/px that this list's tiles cover
        alg_bytes = 10.33 * w * h * (int(i_r[0]) if "mirror" in name else int(i_g[0])) / ntiles
    elif alg is None:                  # probe conversion: 16 probes of 256^2, 6 B/texel out + cube reads / 1.33 * 2 B pyramid
        alg_bytes = NP * (6 + 6) if "Cubemap" in name else NP * 2 * (1 + 1 / 3)
    else:
        alg_bytes = alg * w * h
    gbs = alg_bytes / (ms * 1e-3) / 1e9
    rows.append(dict(task=name, ms=round(ms, 4), alg_GBps=round(gbs, 1)))
    print(f"{name:28s} {ms:8.4f} ms  {gbs:8.1f} GB/s")
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(dict(w=w, h=h, iters=iters, rows=rows), open(os.path.join(ROOT, "gpurun_out", f"time_alt_{w}x{h}.json"), "w"), indent=1)
