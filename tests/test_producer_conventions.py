"""The G-buffer producer against an INDEPENDENT restatement of the reference's conventions.

vk-renderer_b200/csrc/synth.cu ray-casts an analytic scene instead of rasterising gbuf/opaque_taa.*; what the image-space
chain depends on is not the geometry but three conventions, restated in oracle/passes_gbuf.cpp from the reference:
  * the 4-sample jitter table and its scaling                       reference src/main.cpp:93-108
  * jitter enters as clip += clip.w * jitter (NDC shift)            reference src/shaders/gbuf/opaque_taa.vert:40
  * velocity = 0.5 * (prev_ndc - current UNJITTERED ndc)            reference src/shaders/gbuf/opaque_taa.frag:45, .vert:42-43
For a sample of geometry pixels the surface point is rebuilt from the producer's own depth texel (the oracle's
reconstruct_view_vec, gbuffer_encode.glsl:58-69, at the pixel's un-jittered uv), pushed through the restated vertex /
fragment arithmetic with the frame's mvp / prev_mvp, and must land (a) on the pixel centre once jittered, (b) at the stored
depth (GLM_FORCE_DEPTH_ZERO_TO_ONE projection, main.cpp:294), (c) with the stored motion vector.
The CPU test runs the producer's host twin, the GPU test the CUDA kernel.
"""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as orc
import vkr

D24 = np.float32(16777215.0)


def f32(*v):
    return np.array(v, dtype=np.float32)


def check_producer(gb, dp, w, h, frame, stride=7):
    depth = gb.depth.download(0) & 0xFFFFFF
    vel = gb.velocity_vectors.download().astype(np.float32)
    mvp, prev_mvp = np.array(list(dp.mvp), np.float32), np.array(list(dp.prev_mvp), np.float32)
    jit = np.array(list(dp.jitter), np.float32)
    cam = np.array(list(dp.fovy_aspect_znear_zfar), np.float32)
    inv_view = np.linalg.inv(np.array(list(dp.camera), np.float64).reshape(4, 4).T)     # column-major -> math
    worst = dict(ndc=0.0, depth=0.0, vel=0.0)
    n = 0
    pos4, vel2, view3 = f32(0, 0, 0, 0), f32(0, 0), f32(0, 0, 0)
    for py in range(3, h, stride):
        for px in range(2, w, stride):
            k = int(depth[py, px])
            if k == 0xFFFFFF:
                continue
            ndc = f32(2 * (px + 0.5) / w - 1, 2 * (py + 0.5) / h - 1)
            uv = ((ndc - jit[:2]) + 1) / 2                                              # the rasteriser shifted the surface by +jitter
            orc.lib.orc_probe_reconstruct(orc.fptr(uv.astype(np.float32)), float(np.float32(k) / D24), orc.fptr(cam), orc.fptr(view3))
            world = (inv_view @ np.array([*view3.astype(np.float64), 1.0]))[:3].astype(np.float32)
            orc.lib.orc_gbuf_point(orc.fptr(mvp), orc.fptr(prev_mvp), orc.fptr(jit), orc.fptr(world), orc.fptr(pos4), orc.fptr(vel2))
            p = pos4.astype(np.float64)
            worst["ndc"] = max(worst["ndc"], abs(p[0] / p[3] - ndc[0]) * w / 2, abs(p[1] / p[3] - ndc[1]) * h / 2)   # in pixels
            worst["depth"] = max(worst["depth"], abs(p[2] / p[3] - k / float(D24)))
            tol = 2.0 ** -10 * np.abs(vel2).max() + 3e-5
            worst["vel"] = max(worst["vel"], float(np.abs(vel[py, px] - vel2).max() / tol))
            n += 1
    assert n > 500, n
    print(f"frame {frame}: {n} pixels, worst {worst}")
    assert worst["ndc"] < 0.02, worst          # lands on the pixel centre (pixels)
    assert worst["depth"] < 4e-6, worst        # clip z/w is the stored depth
    assert worst["vel"] <= 1.0, worst          # fp16 texel of the restated velocity
    return n


def test_jitter_table_matches_the_restated_next_taa_offset():
    out = f32(0, 0, 0, 0)
    for w, h in ((1920, 1080), (3840, 2160), (328, 200)):
        for frame in range(9):
            dp = vkr.synth_camera(0, frame, w, h, True)
            orc.lib.orc_taa_offset(frame, w, h, orc.fptr(out))
            assert list(dp.jitter) == list(out), (w, h, frame, list(dp.jitter), list(out))
        assert list(vkr.synth_camera(0, 3, w, h, False).jitter) == [0.0] * 4


def test_host_twin_honours_the_restated_conventions():
    w, h = 320, 184
    scene = vkr.synth_scene(2, 48)
    for frame in (0, 1, 6):
        dp = vkr.synth_camera(2, frame, w, h, True)
        gb = vkr.Gbuffer(w, h, device="cpu")
        vkr.synth_gbuffer(scene, dp, gb)
        check_producer(gb, dp, w, h, frame)


@pytest.mark.gpu
def test_cuda_producer_honours_the_restated_conventions(cuda):
    import torch
    w, h = 640, 360
    scene = vkr.synth_scene(5, 64)
    for frame in (0, 3, 10):
        dp = vkr.synth_camera(5, frame, w, h, True)
        gb = vkr.Gbuffer(w, h)
        vkr.synth_gbuffer(scene, dp, gb)
        torch.cuda.synchronize()
        check_producer(gb, dp, w, h, frame, stride=11)
