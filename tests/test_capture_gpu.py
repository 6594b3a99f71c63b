"""Capture path on the device and through the C++ host (blit_image, ReadBackSystem, the PNG / CSV callbacks) vs the oracle."""
import numpy as np
import pytest

import chain
import oracle_lib as orc
import vkr

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("sw,sh,dw,dh,sfmt,dfmt", [
    (640, 360, 640, 360, vkr.R16G16B16A16_SFLOAT, vkr.R8G8B8A8_SRGB),     # the reference's use: TAA target -> read-back image
    (640, 360, 320, 180, vkr.R16G16B16A16_SFLOAT, vkr.R8G8B8A8_SRGB),
    (250, 130, 333, 77, vkr.R8G8B8A8_SRGB, vkr.R8G8B8A8_UNORM),
    (96, 64, 96, 64, vkr.R8G8B8A8_UNORM, vkr.R8G8B8A8_SRGB),
])
def test_blit_matches_oracle(cuda, sw, sh, dw, dh, sfmt, dfmt):
    import torch
    rng = np.random.default_rng(sw + dw)
    if sfmt == vkr.R16G16B16A16_SFLOAT:
        data = (rng.random((sh, sw, 4)) * 1.2 - 0.1).astype(np.float16)
    else:
        data = rng.integers(0, 256, size=(sh, sw, 4), dtype=np.uint8)
    src = vkr.Image(sw, sh, sfmt, device="cpu").upload(data)
    ref = vkr.Image(dw, dh, dfmt, device="cpu")
    orc.ok(orc.lib.orc_blit_image(src.ptr, ref.ptr))
    out = vkr.Image(dw, dh, dfmt)
    vkr.blit_image(src.clone_to("cuda"), out)
    torch.cuda.synchronize()
    g, r = out.download().astype(np.int32), ref.download().astype(np.int32)
    d = np.abs(g - r)
    print(sw, sh, dw, dh, "differing", (d > 0).mean(), "max", d.max())
    assert d.max() <= 1 and (d > 0).mean() <= 1e-3          # device powf vs libm pow in the sRGB OETF: a rounding tie at most


@pytest.mark.parametrize("w,h,fmt,mips", [(37, 11, vkr.D24_UNORM_S8_UINT, 3), (640, 360, vkr.R8G8B8A8_SRGB, 1), (250, 130, vkr.R16G16B16A16_SFLOAT, 1),
                                          (3840, 2160, vkr.D24_UNORM_S8_UINT, 12)])
def test_pack_matches_oracle(cuda, w, h, fmt, mips):
    import torch
    rng = np.random.default_rng(w)
    img = vkr.Image(w, h, fmt, mips, device="cpu")
    img.buf.copy_(torch.from_numpy(rng.integers(0, 256, size=img.buf.numel(), dtype=np.uint8)))
    dev = img.clone_to("cuda")
    ts = vkr.lib.vkr_format_texel_size(fmt)
    for mip in sorted({0, mips - 1, min(1, mips - 1)}):
        mw, mh = img.mip_size(mip)
        for mode in ((0, 1, 2) if ts == 4 else (0,)):
            ref = np.zeros(mw * mh * ts, dtype=np.uint8)
            orc.ok(orc.lib.orc_readback_pack(img.ptr, mip, mode, ref.ctypes.data))
            out = torch.zeros(mw * mh * ts, dtype=torch.uint8, device="cuda")
            vkr.readback_pack(dev, out, mip, mode)
            torch.cuda.synchronize()
            assert (out.cpu().numpy() == ref).all(), (mip, mode)


def test_host_capture_buttons(cuda, tmp_path):
    """Final frame -> PNG (blit + read-back, alpha 255), depth -> PNG and CSV of the 24-bit codes, through the C++ host."""
    import vkr_host
    w, h = 320, 180
    scene = vkr.synth_scene(2, 32)
    app = vkr_host.App(w, h)
    oc = chain.Chain(w, h, chain.OracleBackend(), lut_size=1024)
    for f in range(2):
        dp = vkr.synth_camera(0, f, w, h, True)
        P = oc.begin_frame(scene)
        app.upload_gbuffer(oc.gb)
        app.render_frame(dp, chain.angle_jitter(f))
        oc.passes(P)
        if f == 0:
            oc.end_frame()
    final = tmp_path / "final.png"
    app.capture(app.FINAL_FRAME, final)
    assert app.last_tasks() == ["CopyImage", "ImageRead"]
    got = vkr_host.read_png_rgba8(final)
    ref = vkr.Image(w, h, vkr.R8G8B8A8_SRGB, device="cpu")
    orc.ok(orc.lib.orc_blit_image(oc.taa_target.ptr, ref.ptr))
    r = ref.download().astype(np.int32)
    assert (got[..., 3] == 255).all()
    d = np.abs(got[..., :3].astype(np.int32) - r[..., :3])
    print("final frame: differing", (d > 0).mean(), "max", d.max())
    assert (d <= 1).mean() >= 0.999                          # the TAA image itself is within 4e-3 of the oracle's (tests/parity.py)

    depth_codes = oc.gb.depth.download(0) & 0xFFFFFF
    dpng = tmp_path / "depth.png"
    app.capture(app.DEPTH_PNG, dpng)
    assert app.last_tasks() == ["ImageRead"]
    px = vkr_host.read_png_rgba8(dpng)
    words = px.view(np.uint32)[..., 0]
    assert (words == depth_codes).all()                      # depth is bit-exact, and the stencil byte is masked off

    dcsv = tmp_path / "depth.csv"
    app.capture(app.DEPTH_CSV, dcsv)
    lines = open(dcsv).read().splitlines()
    assert lines[0] == "y, " + ",".join(str(x) for x in range(w)) and len(lines) == h + 1
    for y in (0, h // 2, h - 1):
        cells = lines[1 + y].split(",")
        assert int(cells[0]) == y and [int(c, 16) for c in cells[1:]] == depth_codes[y].tolist()
    app.close()
