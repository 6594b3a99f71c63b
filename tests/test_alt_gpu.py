"""The renderer's earlier AO / reflection passes (reference src/ssao.cpp, src/ssr.cpp) through the C ABI vs the CPU oracle.

Both passes are evaluated with the oracle's unfused arithmetic (SSAO's output is a count of 16 strict comparisons, the SSR
march picks texels), so the bar is bit-exact bytes, allowing a 1e-4 fraction of pixels for libm-vs-device differences in the
last place of a division feeding a comparison.
"""
import ctypes as C

import numpy as np
import pytest

import chain
import oracle_lib as orc
import vkr
import vkr_host

pytestmark = pytest.mark.gpu


def _oracle_frame(w, h, scene_seed=0, objects=64, frames=1):
    scene = vkr.synth_scene(scene_seed, objects)
    oc = chain.Chain(w, h, chain.OracleBackend(), lut_size=64)
    for _ in range(frames):
        imgs, P = oc.render(scene)
    return imgs, P


def _ssao_params(w, h, P, seed=3, reference_packing=False):
    p = vkr.SsaoParams()
    fovy, aspect, zn, zf = P["trace"].fovy, P["trace"].aspect, P["trace"].znear, P["trace"].zfar
    C.memmove(p.projection, vkr.perspective(fovy, aspect, zn, zf), 64)
    p.fovy, p.aspect, p.znear, p.zfar = fovy, aspect, zn, zf
    vkr.set_ssao_samples(p, vkr.sphere_samples(seed), reference_packing)
    return p


def _ssr_params(P):
    sp = vkr.SsrParams()
    C.memmove(sp.normal_mat, P["trace"].normal_mat, 64)
    sp.fovy, sp.aspect, sp.znear, sp.zfar = P["trace"].fovy, P["trace"].aspect, P["trace"].znear, P["trace"].zfar
    return sp


@pytest.mark.parametrize("w,h,packing", [(320, 180, False), (250, 130, True), (1280, 720, False)])
def test_ssao_matches_oracle(cuda, w, h, packing):
    import torch
    imgs, P = _oracle_frame(w, h)
    p = _ssao_params(w, h, P, reference_packing=packing)
    ref = vkr.Image(w, h, vkr.R8_UNORM, device="cpu")
    orc.ok(orc.lib.orc_ssao(C.byref(p), imgs["depth"].ptr, ref.ptr, 0, 0))
    out = vkr.Image(w, h, vkr.R8_UNORM)
    vkr.ssao(p, imgs["depth"].clone_to("cuda"), out)
    torch.cuda.synchronize()
    g, r = out.download(), ref.download()
    assert len(np.unique(r)) > 8, "the scene should produce a spread of occlusion values"
    bad = (g != r)
    assert bad.mean() <= 1e-4, (bad.sum(), np.abs(g.astype(int) - r.astype(int)).max())
    assert np.abs(g.astype(int) - r.astype(int)).max() <= 16         # a flipped comparison moves the count by one sample


@pytest.mark.parametrize("w,h,frame_name", [(320, 180, "color"), (254, 126, "taa"), (1280, 720, "color")])
def test_ssr_legacy_matches_oracle(cuda, w, h, frame_name):
    import torch
    imgs, P = _oracle_frame(w, h, frames=2)
    sp = _ssr_params(P)
    ref = vkr.Image(w, h, vkr.R8G8B8A8_UNORM, device="cpu")
    orc.ok(orc.lib.orc_ssr_legacy(C.byref(sp), imgs["depth"].ptr, imgs["normal"].ptr, imgs[frame_name].ptr, imgs["material"].ptr, ref.ptr, 0, 0))
    out = vkr.Image(w, h, vkr.R8G8B8A8_UNORM)
    dev = {k: imgs[k].clone_to("cuda") for k in ("depth", "normal", "material", frame_name)}
    vkr.ssr_legacy(sp, dev["depth"], dev["normal"], dev[frame_name], dev["material"], out)
    torch.cuda.synchronize()
    g, r = out.download().astype(np.int32), ref.download().astype(np.int32)
    hit = (r[..., :3].max(axis=-1) > 0)
    assert hit.mean() > 0.02, "the scene should have reflecting pixels"
    d = np.abs(g - r).max(axis=-1)
    # a march whose branch flips on the last place of an IEEE-vs-device division may end elsewhere: <= 1e-4 of the pixels
    assert (d > 1).mean() <= 1e-4, ((d > 1).sum(), d.max())
    assert (d > 0).mean() <= 1e-3, (d > 0).sum()


def test_alt_passes_row_windows_and_layers(cuda):
    """Row windows write the bytes of the whole pass; two array layers equal two single launches."""
    import torch
    w, h = 256, 160
    imgs, P = _oracle_frame(w, h)
    p, sp = _ssao_params(w, h, P), _ssr_params(P)
    dev = {k: imgs[k].clone_to("cuda") for k in ("depth", "normal", "material", "color")}
    whole_ao, whole_ssr = vkr.Image(w, h, vkr.R8_UNORM), vkr.Image(w, h, vkr.R8G8B8A8_UNORM)
    vkr.ssao(p, dev["depth"], whole_ao)
    vkr.ssr_legacy(sp, dev["depth"], dev["normal"], dev["color"], dev["material"], whole_ssr)
    part_ao, part_ssr = vkr.Image(w, h, vkr.R8_UNORM), vkr.Image(w, h, vkr.R8G8B8A8_UNORM)
    mid = (h // 2) // vkr.ROW_ALIGN * vkr.ROW_ALIGN
    for rows in ((0, mid), (mid, h)):
        vkr.ssao(p, dev["depth"], part_ao, ext=vkr.make_ext([], rows=rows))
        vkr.ssr_legacy(sp, dev["depth"], dev["normal"], dev["color"], dev["material"], part_ssr, ext=vkr.make_ext([], rows=rows))
    torch.cuda.synchronize()
    assert torch.equal(whole_ao.buf, part_ao.buf) and torch.equal(whole_ssr.buf, part_ssr.buf)

    # layers: the same view twice with different cameras' parameter blocks
    imgs2, P2 = _oracle_frame(w, h, scene_seed=5)
    arrs = {}
    for k, fmt, mips in (("depth", vkr.D24_UNORM_S8_UINT, imgs["depth"].mips), ("normal", vkr.R16G16_UNORM, 1), ("material", vkr.R8G8B8A8_SRGB, 1),
                         ("color", vkr.R8G8B8A8_SRGB, 1)):
        a = vkr.ImageArray(w, h, fmt, mips, layers=2)
        a.layer(0).buf.copy_(imgs[k].buf)
        a.layer(1).buf.copy_(imgs2[k].buf)
        arrs[k] = a
    out_ao, out_ssr = vkr.ImageArray(w, h, vkr.R8_UNORM, 1, layers=2), vkr.ImageArray(w, h, vkr.R8G8B8A8_UNORM, 1, layers=2)
    scratch = torch.zeros(2 * vkr.PARAMS_SCRATCH_STRIDE, dtype=torch.uint8, device="cuda")
    p2, sp2 = _ssao_params(w, h, P2), _ssr_params(P2)
    vkr.ssao(vkr.params_array(vkr.SsaoParams, [p, p2]), arrs["depth"], out_ao, ext=vkr.make_ext([arrs["depth"], out_ao], layers=2, scratch=scratch))
    torch.cuda.synchronize()
    vkr.ssr_legacy(vkr.params_array(vkr.SsrParams, [sp, sp2]), arrs["depth"], arrs["normal"], arrs["color"], arrs["material"], out_ssr,
                   ext=vkr.make_ext([arrs["depth"], arrs["normal"], arrs["color"], arrs["material"], out_ssr], layers=2, scratch=scratch))
    torch.cuda.synchronize()
    dev2 = {k: imgs2[k].clone_to("cuda") for k in ("depth", "normal", "material", "color")}
    single_ao, single_ssr = vkr.Image(w, h, vkr.R8_UNORM), vkr.Image(w, h, vkr.R8G8B8A8_UNORM)
    vkr.ssao(p2, dev2["depth"], single_ao)
    vkr.ssr_legacy(sp2, dev2["depth"], dev2["normal"], dev2["color"], dev2["material"], single_ssr)
    torch.cuda.synchronize()
    assert torch.equal(out_ao.layer(0).buf, whole_ao.buf) and torch.equal(out_ao.layer(1).buf, single_ao.buf)
    assert torch.equal(out_ssr.layer(0).buf, whole_ssr.buf) and torch.equal(out_ssr.layer(1).buf, single_ssr.buf)


def test_alt_passes_validation(cuda):
    w, h = 64, 32
    depth = vkr.Image(w, h, vkr.D24_UNORM_S8_UINT, vkr.full_mip_count(w, h))
    bad = vkr.Image(w, h, vkr.R8G8B8A8_UNORM)
    with pytest.raises(vkr.VkrError):
        vkr.ssao(vkr.SsaoParams(), depth, bad)                     # SSAO target must be R8_UNORM
    nrm, mat, frame = vkr.Image(w, h, vkr.R16G16_UNORM), vkr.Image(w, h, vkr.R8G8B8A8_SRGB), vkr.Image(w, h, vkr.R8G8B8A8_SRGB)
    with pytest.raises(vkr.VkrError):
        vkr.ssr_legacy(vkr.SsrParams(), depth, nrm, frame, mat, vkr.Image(w, h, vkr.R8_UNORM))


def test_host_app_runs_the_alt_passes(cuda):
    """SSAOPass / add_ssr_pass scheduled behind TAA by the C++ host equal the oracle fed with the blocks the host built."""
    w, h = 320, 180
    scene = vkr.synth_scene(0, 32)
    app = vkr_host.App(w, h, device=0)
    app.set_option("alt.ssao", 1)
    app.set_option("alt.ssr", 1)
    oc = chain.Chain(w, h, chain.OracleBackend(), lut_size=1024)
    for f in range(2):
        dp = vkr.synth_camera(0, f, w, h, True)
        P = oc.begin_frame(scene)
        app.upload_gbuffer(oc.gb)
        app.render_frame(dp, chain.angle_jitter(f))
        oc.passes(P)
        ref = oc.images()
        assert app.last_tasks()[-2:] == ["SSAO", "SSR"]
        a, r = app.alt_params()
        ref_ao, ref_ssr = vkr.Image(w, h, vkr.R8_UNORM, device="cpu"), vkr.Image(w, h, vkr.R8G8B8A8_UNORM, device="cpu")
        color_dev = app.readback("color")                 # the host's own lit frame is the SSR input (1 LSB from the oracle's)
        orc.ok(orc.lib.orc_ssao(C.byref(a), ref["depth"].ptr, ref_ao.ptr, 0, 0))
        orc.ok(orc.lib.orc_ssr_legacy(C.byref(r), ref["depth"].ptr, ref["normal"].ptr, color_dev.ptr, ref["material"].ptr, ref_ssr.ptr, 0, 0))
        g_ao, g_ssr = app.readback("ssao").download(), app.readback("ssr_legacy").download().astype(np.int32)
        assert (g_ao != ref_ao.download()).mean() <= 1e-4
        d = np.abs(g_ssr - ref_ssr.download().astype(np.int32)).max(axis=-1)
        assert (d > 1).mean() <= 1e-4 and (d > 0).mean() <= 1e-3
        oc.end_frame()
    app.close()


# ---- ScreenSpaceTrace (reference src/screen_trace.cpp) -------------------------------------------------------------------
def _st_params(P, angle=0.3, rnd=0.37):
    p = vkr.ScreenTraceParams()
    C.memmove(p.normal_mat, P["trace"].normal_mat, 64)
    p.random_offset, p.angle_offset = rnd, angle
    p.fovy, p.aspect, p.znear, p.zfar = P["trace"].fovy, P["trace"].aspect, P["trace"].znear, P["trace"].zfar
    return p


def _st_compare(got, ref):
    """fp16 RGBA: a march that breaks / hits one sample earlier on a last-place difference changes the pixel (and its 3x3
    tile neighbours' gather) visibly, so the bar is a fraction of pixels: |diff| <= 2e-3 on >= 99.9 %."""
    g, r = got.astype(np.float32), ref.astype(np.float32)
    assert (np.isnan(g) == np.isnan(r)).mean() >= 0.9999
    d = np.abs(np.nan_to_num(g) - np.nan_to_num(r)).max(axis=-1)
    return float((d > 2e-3).mean()), float(d.max())


@pytest.mark.parametrize("w,h,color_name", [(320, 180, "color"), (250, 132, "taa"), (1280, 720, "color")])
def test_screen_trace_chain_matches_oracle(cuda, w, h, color_name):
    import torch
    imgs, P = _oracle_frame(w, h, frames=2)
    p = _st_params(P)
    F = vkr.R16G16B16A16_SFLOAT
    ref = {k: vkr.Image(w, h, F, device="cpu") for k in ("raw", "filtered", "acc")}
    got = {k: vkr.Image(w, h, F) for k in ("raw", "filtered", "acc")}
    dev = {k: imgs[k].clone_to("cuda") for k in ("depth", "prev_depth", "normal", "material", color_name)}
    cam = np.array([p.fovy, p.aspect, p.znear, p.zfar], dtype=np.float32)
    for it in range(2):                                    # second iteration accumulates over the first (static camera -> blends)
        orc.ok(orc.lib.orc_screen_trace_main(C.byref(p), imgs["depth"].ptr, imgs["normal"].ptr, imgs[color_name].ptr, imgs["material"].ptr, ref["raw"].ptr, 0, 0))
        orc.ok(orc.lib.orc_screen_trace_filter(p.znear, p.zfar, ref["raw"].ptr, imgs["depth"].ptr, ref["filtered"].ptr, 0, 0))
        prev = imgs["depth"] if it else imgs["prev_depth"]
        orc.ok(orc.lib.orc_screen_trace_accumulate(orc.fptr(cam), imgs["depth"].ptr, prev.ptr, ref["filtered"].ptr, ref["acc"].ptr, 0, 0))
        vkr.screen_trace_main(p, dev["depth"], dev["normal"], dev[color_name], dev["material"], got["raw"])
        vkr.screen_trace_filter(p.znear, p.zfar, got["raw"], dev["depth"], got["filtered"])
        vkr.screen_trace_accumulate(cam, dev["depth"], dev["depth"] if it else dev["prev_depth"], got["filtered"], got["acc"])
        torch.cuda.synchronize()
        r = ref["raw"].download().astype(np.float32)
        assert (r[..., 3] > 0.05).mean() > 0.5 and (r[..., :3].max(axis=-1) > 0).mean() > 0.01, "AO and gathered colour should both be populated"
        for k in ("raw", "filtered", "acc"):
            frac, dmax = _st_compare(got[k].download(), ref[k].download())
            assert frac <= 1e-3, (it, k, frac, dmax)
    # the filter and the accumulation alone, on identical inputs, are exact arithmetic: bit-identical fp16
    got["raw"].buf.copy_(ref["raw"].buf)
    vkr.screen_trace_filter(p.znear, p.zfar, got["raw"], dev["depth"], got["filtered"])
    torch.cuda.synchronize()
    a, b = got["filtered"].download().view(np.uint16), ref["filtered"].download().view(np.uint16)
    assert (a != b).mean() <= 1e-5, (a != b).sum()


def test_screen_trace_row_windows(cuda):
    import torch
    w, h = 256, 160
    imgs, P = _oracle_frame(w, h)
    p = _st_params(P)
    F = vkr.R16G16B16A16_SFLOAT
    dev = {k: imgs[k].clone_to("cuda") for k in ("depth", "prev_depth", "normal", "material", "color")}
    cam = np.array([p.fovy, p.aspect, p.znear, p.zfar], dtype=np.float32)
    whole = {k: vkr.Image(w, h, F) for k in ("raw", "filtered", "acc")}
    part = {k: vkr.Image(w, h, F) for k in ("raw", "filtered", "acc")}
    vkr.screen_trace_main(p, dev["depth"], dev["normal"], dev["color"], dev["material"], whole["raw"])
    vkr.screen_trace_filter(p.znear, p.zfar, whole["raw"], dev["depth"], whole["filtered"])
    vkr.screen_trace_accumulate(cam, dev["depth"], dev["depth"], whole["filtered"], whole["acc"])
    mid = (h // 2) // vkr.ROW_ALIGN * vkr.ROW_ALIGN
    for rows in ((0, mid), (mid, h)):
        vkr.screen_trace_main(p, dev["depth"], dev["normal"], dev["color"], dev["material"], part["raw"], ext=vkr.make_ext([], rows=rows))
    for rows in ((0, mid), (mid, h)):
        vkr.screen_trace_filter(p.znear, p.zfar, part["raw"], dev["depth"], part["filtered"], ext=vkr.make_ext([], rows=rows))
    for rows in ((0, mid), (mid, h)):
        vkr.screen_trace_accumulate(cam, dev["depth"], dev["depth"], part["filtered"], part["acc"], ext=vkr.make_ext([], rows=rows))
    torch.cuda.synchronize()
    for k in whole:
        assert torch.equal(whole[k].buf, part[k].buf), k


def test_host_app_runs_screen_trace(cuda):
    w, h = 320, 176
    scene = vkr.synth_scene(0, 32)
    app = vkr_host.App(w, h, device=0)
    app.set_option("alt.screen_trace", 1)
    oc = chain.Chain(w, h, chain.OracleBackend(), lut_size=1024)
    F = vkr.R16G16B16A16_SFLOAT
    ref = {k: vkr.Image(w, h, F, device="cpu") for k in ("raw", "filtered", "acc")}
    for f in range(2):
        dp = vkr.synth_camera(0, f, w, h, True)
        P = oc.begin_frame(scene)
        app.upload_gbuffer(oc.gb)
        app.render_frame(dp, chain.angle_jitter(f))
        oc.passes(P)
        im = oc.images()
        assert app.last_tasks()[-3:] == ["ScreenTrace", "ScreenTraceFilter", "ScreenTraceAccumulate"]
        p = app.screen_trace_params()
        color_dev = app.readback("color")
        cam = np.array([p.fovy, p.aspect, p.znear, p.zfar], dtype=np.float32)
        orc.ok(orc.lib.orc_screen_trace_main(C.byref(p), im["depth"].ptr, im["normal"].ptr, color_dev.ptr, im["material"].ptr, ref["raw"].ptr, 0, 0))
        orc.ok(orc.lib.orc_screen_trace_filter(p.znear, p.zfar, ref["raw"].ptr, im["depth"].ptr, ref["filtered"].ptr, 0, 0))
        orc.ok(orc.lib.orc_screen_trace_accumulate(orc.fptr(cam), im["depth"].ptr, im["prev_depth"].ptr, ref["filtered"].ptr, ref["acc"].ptr, 0, 0))
        for k, name in (("raw", "st_raw"), ("filtered", "st_filtered"), ("acc", "st_accumulated")):
            frac, dmax = _st_compare(app.readback(name).download(), ref[k].download())
            assert frac <= 1e-3, (f, k, frac, dmax)
        oc.end_frame()
    app.close()


# ---- GTAO variants (reference src/gtao.cpp:241-284,:349-413,:445-526) ----------------------------------------------------
def _gv_params(P, angle=0.21):
    p = vkr.GtaoParams()
    C.memmove(p.normal_mat, P["trace"].normal_mat, 64)
    p.fovy, p.aspect, p.znear, p.zfar, p.angle_offset = P["trace"].fovy, P["trace"].aspect, P["trace"].znear, P["trace"].zfar, angle
    return p


def _ao_compare(got, ref):
    """fp16 AO: |diff| <= 2e-3 except where a thickness break flipped on a last-place difference; returns that fraction."""
    g, r = got.astype(np.float32), ref.astype(np.float32)
    assert (np.isnan(g) == np.isnan(r)).mean() >= 0.9999
    d = np.abs(np.nan_to_num(g) - np.nan_to_num(r))
    return float((d > 2e-3).mean()), float(d.max())


@pytest.mark.parametrize("w,h", [(320, 180), (1280, 720)])
def test_gtao_variants_match_oracle(cuda, w, h):
    import torch
    imgs, P = _oracle_frame(w, h, frames=2)
    p = _gv_params(P)
    w2, h2 = w // 2, h // 2
    dev = {k: imgs[k].clone_to("cuda") for k in ("depth", "prev_depth", "normal")}
    # graphics main
    ref_raw, got_raw = vkr.Image(w2, h2, vkr.R16G16B16A16_SFLOAT, device="cpu"), vkr.Image(w2, h2, vkr.R16G16B16A16_SFLOAT)
    orc.ok(orc.lib.orc_gtao_main_graphics(C.byref(p), imgs["depth"].ptr, imgs["normal"].ptr, ref_raw.ptr, 1, 0, 0))
    vkr.gtao_main_graphics(p, dev["depth"], dev["normal"], got_raw)
    torch.cuda.synchronize()
    r = ref_raw.download().astype(np.float32)
    # (the variant's projected-normal angle uses normalize(end - pos) instead of the slice tangent: values above 1 and a few
    #  NaN from acos of 1 + 1 ulp are what the formula gives)
    assert 0.2 < np.nanmean(r[..., 0]) < 1.5 and np.nanstd(r[..., 0]) > 0.02 and np.isnan(r[..., 0]).mean() < 0.05
    frac, dmax = _ao_compare(got_raw.download(), ref_raw.download())
    assert frac <= 1e-3, (frac, dmax)
    # deinterleave: pure data movement -> bit-exact
    ref_arr = vkr.ImageArray(w2 // 4, h2 // 4, vkr.R32_SFLOAT, 1, layers=16, device="cpu")
    got_arr = vkr.ImageArray(w2 // 4, h2 // 4, vkr.R32_SFLOAT, 1, layers=16)
    orc.ok(orc.lib.orc_gtao_deinterleave(imgs["depth"].ptr, 1, ref_arr.ptr, ref_arr.layer_stride, 2))
    vkr.gtao_deinterleave(dev["depth"], got_arr, 2)
    torch.cuda.synchronize()
    assert torch.equal(got_arr.buf.cpu(), ref_arr.buf)
    assert (ref_arr.layer(5).download() > 0).any()
    # deinterleaved main, two layers
    for layer in (0, 6):
        orc.ok(orc.lib.orc_gtao_main_deinterleaved(C.byref(p), ref_arr.ptr, ref_arr.layer_stride, imgs["normal"].ptr, ref_raw.ptr, 2, layer, 0, 0))
        vkr.gtao_main_deinterleaved(p, got_arr, dev["normal"], got_raw, 2, layer)
    torch.cuda.synchronize()
    frac, dmax = _ao_compare(got_raw.download(), ref_raw.download())
    assert frac <= 1e-3, (frac, dmax)
    # reprojection: exact arithmetic -> bit-exact fp16
    F = vkr.R16_SFLOAT
    rng = np.random.default_rng(1)
    ao_np, prev_np = rng.random((h2, w2)).astype(np.float16), rng.random((h2, w2)).astype(np.float16)
    ao, prev, ref_out = vkr.Image(w2, h2, F, device="cpu"), vkr.Image(w2, h2, F, device="cpu"), vkr.Image(w2, h2, F, device="cpu")
    ao.upload(ao_np); prev.upload(prev_np)
    cam = np.array([p.fovy, p.aspect, p.znear, p.zfar], dtype=np.float32)
    got_out = vkr.Image(w2, h2, F)
    for pd_name in ("prev_depth", "depth"):              # a moving and a static previous frame
        orc.ok(orc.lib.orc_gtao_reproject(orc.fptr(cam), imgs["depth"].ptr, imgs[pd_name].ptr, 1, ao.ptr, prev.ptr, ref_out.ptr, 0, 0))
        vkr.gtao_reproject(cam, dev["depth"], dev[pd_name], ao.clone_to("cuda"), prev.clone_to("cuda"), got_out)
        torch.cuda.synchronize()
        assert np.array_equal(got_out.download().view(np.uint16), ref_out.download().view(np.uint16)), pd_name
    assert (ref_out.download() != ao_np).mean() > 0.5    # static previous frame: most texels blended


def test_host_app_runs_gtao_variants(cuda):
    w, h = 320, 192
    scene = vkr.synth_scene(0, 32)
    app = vkr_host.App(w, h, device=0)
    app.set_option("alt.gtao_variants", 1)
    oc = chain.Chain(w, h, chain.OracleBackend(), lut_size=1024)
    w2, h2 = w // 2, h // 2
    ref_raw = vkr.Image(w2, h2, vkr.R16G16B16A16_SFLOAT, device="cpu")
    ref_filt, ref_prev, ref_out = [vkr.Image(w2, h2, vkr.R16_SFLOAT, device="cpu") for _ in range(3)]
    ref_arr = vkr.ImageArray(w2 // 4, h2 // 4, vkr.R32_SFLOAT, 1, layers=16, device="cpu")
    for f in range(2):
        dp = vkr.synth_camera(0, f, w, h, True)
        P = oc.begin_frame(scene)
        app.upload_gbuffer(oc.gb)
        app.render_frame(dp, chain.angle_jitter(f))
        oc.passes(P)
        im = oc.images()
        assert app.last_tasks()[-5:] == ["GTAO", "GTAO_filter", "GTAO_reproject", "GTAO_deinterleave", "GTAO_deinterleaved"]
        # the variant object draws two angles per frame from the table (graphics main, then deinterleaved main)
        jit = np.float32(chain.angle_jitter(f))
        ang = [float(np.float32(np.float32(chain.ANGLE_TABLE[(2 * f + i) % 12]) / np.float32(360.0)) + jit) for i in range(2)]
        assert app.get_value("gv.angle_offset") == ang[1]
        p = _gv_params(P, ang[0])
        cam = np.array([p.fovy, p.aspect, p.znear, p.zfar], dtype=np.float32)
        orc.ok(orc.lib.orc_gtao_main_graphics(C.byref(p), im["depth"].ptr, im["normal"].ptr, ref_raw.ptr, 1, 0, 0))
        orc.ok(orc.lib.orc_gtao_filter(C.byref(p), im["depth"].ptr, ref_raw.ptr, ref_filt.ptr, 1, 0, 0))
        orc.ok(orc.lib.orc_gtao_reproject(orc.fptr(cam), im["depth"].ptr, im["prev_depth"].ptr, 1, ref_filt.ptr, ref_prev.ptr, ref_out.ptr, 0, 0))
        for name, ref in (("gv_filtered", ref_filt), ("gv_output", ref_out)):
            frac, dmax = _ao_compare(app.readback(name).download(), ref.download())
            assert frac <= 2e-3, (f, name, frac, dmax)
        p2 = _gv_params(P, ang[1])
        orc.ok(orc.lib.orc_gtao_deinterleave(im["depth"].ptr, 1, ref_arr.ptr, ref_arr.layer_stride, 2))
        orc.ok(orc.lib.orc_gtao_main_deinterleaved(C.byref(p2), ref_arr.ptr, ref_arr.layer_stride, im["normal"].ptr, ref_raw.ptr, 2, 0, 0, 0))
        frac, dmax = _ao_compare(app.readback("gv_raw").download(), ref_raw.download())    # graphics raw with layer 0's sparse overwrite
        assert frac <= 2e-3, (f, "gv_raw", frac, dmax)
        ref_prev, ref_out = ref_out, ref_prev                                   # main.cpp:417 remap(output, prev_frame)
        oc.end_frame()
    app.close()
