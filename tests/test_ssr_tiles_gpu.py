"""The SSSR tile path (classification, indirect trace, tile regression; reference src/advanced_ssr.cpp:216-302,:440-538) through
the C ABI and through the C++ host against the CPU oracle.

All three kernels keep the oracle's unfused arithmetic and the shaders' tree-sum association, so the bar is: tile lists equal as
sets (their order comes out of atomicAdd on both the reference's GPU and this one), planes bit-identical, rays identical except
within the bar of the per-pixel trace (tests/parity.py "rays": the ray direction goes through device sin / cos / sqrt).
"""
import ctypes as C

import numpy as np
import pytest

import chain
import oracle_lib as orc
import parity
import vkr
from test_ssr_tiles import classify

pytestmark = pytest.mark.gpu


def _frame(w, h, objects=64, seed=1):
    scene = vkr.synth_scene(seed, objects)
    oc = chain.Chain(w, h, chain.OracleBackend(), luts=(None, None))
    P = oc.begin_frame(scene)
    oc.steps(P)[0][1]()                                  # Hi-Z on the CPU: the tile passes read mips >= 1
    return oc, P


def _gpu_classify(material_dev, w2, h2, max_rough=1.0, glossy=0.5):
    import torch
    n = ((w2 + 7) // 8) * ((h2 + 7) // 8)
    rt = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    gt = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    ri = torch.full((3,), 77, dtype=torch.int32, device="cuda")
    gi = torch.full((3,), 77, dtype=torch.int32, device="cuda")
    vkr.sssr_clear_indirect(ri, gi)
    p = vkr.ClassifyParams(w2, h2, max_rough, glossy)
    vkr.sssr_classification(p, material_dev, rt, gt, ri, gi)
    torch.cuda.synchronize()
    assert ri.tolist()[1:] == [1, 1] and gi.tolist()[1:] == [1, 1]
    return rt, gt, ri, gi


@pytest.mark.parametrize("w,h,glossy", [(640, 360, 0.5), (328, 200, 0.3), (1280, 720, 0.5), (40, 32, 0.5)])
def test_classification_lists_match_oracle(cuda, w, h, glossy):
    oc, _ = _frame(w, h)
    w2, h2 = w // 2, h // 2
    r_ref, g_ref = classify(oc.gb.material, w2, h2, glossy=glossy)
    rt, gt, ri, gi = _gpu_classify(oc.gb.material.clone_to("cuda"), w2, h2, glossy=glossy)
    nr, ng = int(ri[0]), int(gi[0])
    assert nr == len(r_ref) and ng == len(g_ref), (nr, len(r_ref), ng, len(g_ref))
    assert sorted(rt[:nr].tolist()) == sorted(r_ref.tolist())
    assert sorted(gt[:ng].tolist()) == sorted(g_ref.tolist())
    if w >= 328:
        assert nr > 0 and ng > 0, "the scene should have both mirror-like and glossy tiles"


@pytest.mark.parametrize("w,h", [(640, 360), (328, 200), (1280, 720)])
def test_trace_indirect_matches_oracle(cuda, w, h):
    import torch
    oc, P = _frame(w, h)
    w2, h2 = w // 2, h // 2
    r_ref, g_ref = classify(oc.gb.material, w2, h2)
    hal = orc.halton23_seq(128)
    sentinel = np.full((h2, w2, 4), 4242, dtype=np.uint16)
    oc.rays.upload(sentinel)
    for rtype, tiles in ((0, r_ref), (1, g_ref)):
        orc.ok(orc.lib.orc_sssr_trace_indirect(C.byref(P["trace"]), rtype, oc.gb.depth.ptr, oc.gb.downsampled_normals.ptr, oc.gb.material.ptr,
                                               orc.fptr(hal), tiles.ctypes.data, len(tiles), oc.rays.ptr))
    ref = oc.rays.download()

    dev = {k: getattr(oc.gb, k).clone_to("cuda") for k in ("depth", "downsampled_normals", "material")}
    rays = vkr.Image(w2, h2, vkr.R16G16B16A16_UNORM).upload(sentinel)
    hal_dev = torch.from_numpy(hal).cuda()
    rt, gt, ri, gi = _gpu_classify(dev["material"], w2, h2)
    vkr.sssr_trace_indirect(P["trace"], 0, dev["depth"], dev["downsampled_normals"], dev["material"], hal_dev, rt, ri, rays)
    vkr.sssr_trace_indirect(P["trace"], 1, dev["depth"], dev["downsampled_normals"], dev["material"], hal_dev, gt, gi, rays)
    torch.cuda.synchronize()
    got = rays.download()
    # every pixel of the image belongs to exactly one listed tile; partial tiles write only their inside part
    assert not (got == 4242).all(axis=-1).any() and not (ref == 4242).all(axis=-1).any()
    # the bar of the per-pixel trace (tests/parity.py "rays"): the VNDF direction goes through device sin/cos/sqrt, so the last
    # place of a ray differs on a few pixels; a march whose branch flips on it may end elsewhere on <= 1e-4 of the pixels
    m = parity.compare("rays", got, ref)
    d = np.abs(got.astype(np.int64) - ref.astype(np.int64)).max(axis=-1)
    print(w, h, m, "differing px", (d > 0).mean(), "max code diff", d.max())
    assert m["valid_frac"] > 0.01, "the scene should produce hits"
    assert parity.TOLERANCES["rays"](m), m
    assert (d > 0).mean() <= 2e-3, (d > 0).mean()


@pytest.mark.parametrize("w,h", [(640, 360), (328, 200), (1280, 720), (3840, 2160)])
def test_tile_regression_matches_oracle(cuda, w, h):
    import torch
    if w >= 3840:                                         # full size: device producer, Hi-Z on the device, oracle on the same bytes
        scene = vkr.synth_scene(1, 64)
        dp = vkr.synth_camera(0, 0, w, h, True)
        gbd = vkr.Gbuffer(w, h)
        vkr.synth_gbuffer(scene, dp, gbd)
        vkr.hiz_build(gbd.depth, gbd.normal, gbd.velocity_vectors, gbd.downsampled_normals, gbd.downsampled_velocity_vectors)
        torch.cuda.synchronize()
        depth_dev, depth_cpu = gbd.depth, gbd.depth.clone_to("cpu")
        P = chain.frame_params(dp, 0, 0)
    else:
        oc, P = _frame(w, h)
        depth_cpu, depth_dev = oc.gb.depth, oc.gb.depth.clone_to("cuda")
    c2w = vkr.mat_transpose(P["trace"].normal_mat)
    cam = np.array([P["trace"].fovy, P["trace"].aspect, P["trace"].znear, P["trace"].zfar], dtype=np.float32)
    pw, ph = max(1, w // 16), max(1, h // 16)
    ref = vkr.Image(pw, ph, vkr.R32G32B32A32_SFLOAT, device="cpu")
    orc.ok(orc.lib.orc_sssr_tile_regression(c2w, orc.fptr(cam), depth_cpu.ptr, ref.ptr))
    p = vkr.RegressionParams()
    C.memmove(p.camera_to_world, c2w, 64)
    p.fovy, p.aspect, p.znear, p.zfar = [float(x) for x in cam]
    out = vkr.Image(pw, ph, vkr.R32G32B32A32_SFLOAT)
    vkr.sssr_tile_regression(p, depth_dev, out)
    torch.cuda.synchronize()
    g, r = out.download().view(np.uint32), ref.download().view(np.uint32)
    same = (g == r) | (np.isnan(g.view(np.float32)) & np.isnan(r.view(np.float32)))
    print(w, h, "bit-identical", same.mean())
    assert same.mean() >= 1.0 - 1e-5, (~same).sum()


def test_host_runs_the_tile_tasks_in_reference_order(cuda):
    """AdvancedSSR::run with the block the reference has commented out (src/advanced_ssr.cpp:547-550): task order, the lists and
    the planes the host's images and buffers hold afterwards."""
    import vkr_host
    w, h = 640, 360
    scene = vkr.synth_scene(1, 64)
    app = vkr_host.App(w, h)
    app.set_option("alt.ssr_tiles", 1.0)
    oc = chain.Chain(w, h, chain.OracleBackend(), luts=(None, None))
    P = oc.begin_frame(scene)
    dp = vkr.synth_camera(0, 0, w, h, True)
    app.upload_gbuffer(oc.gb)
    app.render_frame(dp, chain.angle_jitter(0))
    tasks = app.last_tasks()
    i = tasks.index("SSSR_Clear")
    assert tasks[i:i + 5] == ["SSSR_Clear", "SSSR_Classification", "SSSR_Tile_Regression", "SSSR_trace", "SSSR_trace"], tasks
    assert tasks[i + 5] == "SSSR_filter"
    oc.steps(P)[0][1]()
    w2, h2 = w // 2, h // 2
    r_ref, g_ref = classify(oc.gb.material, w2, h2)
    ri, gi = app.readback_buffer("reflective_indirect", "uint32"), app.readback_buffer("glossy_indirect", "uint32")
    assert ri.tolist() == [len(r_ref), 1, 1] and gi.tolist() == [len(g_ref), 1, 1]
    assert sorted(app.readback_buffer("reflective_tiles")[:ri[0]].tolist()) == sorted(r_ref.tolist())
    assert sorted(app.readback_buffer("glossy_tiles")[:gi[0]].tolist()) == sorted(g_ref.tolist())
    c2w = vkr.mat_transpose(P["trace"].normal_mat)
    cam = np.array([P["trace"].fovy, P["trace"].aspect, P["trace"].znear, P["trace"].zfar], dtype=np.float32)
    ref = vkr.Image(w // 16, h // 16, vkr.R32G32B32A32_SFLOAT, device="cpu")
    orc.ok(orc.lib.orc_sssr_tile_regression(c2w, orc.fptr(cam), oc.gb.depth.ptr, ref.ptr))
    got = app.readback("tile_planes")
    g, r = got.download().view(np.uint32), ref.download().view(np.uint32)
    assert ((g == r).mean()) >= 1.0 - 1e-5
    # a second frame clears and refills the lists (counts do not accumulate)
    app.upload_gbuffer(oc.gb)
    app.render_frame(dp, chain.angle_jitter(1))
    assert app.readback_buffer("reflective_indirect", "uint32").tolist() == [len(r_ref), 1, 1]
    app.close()
