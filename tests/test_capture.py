"""Capture path (reference src/util_passes.cpp:152-180, src/image_readback.cpp, src/main.cpp:118-176): analytic
anchors of the oracle's blit / pack and the host's PNG writer.  CPU only."""
import numpy as np

import oracle_lib as orc
import vkr


def test_png_writer_roundtrip(tmp_path):
    import vkr_host
    rng = np.random.default_rng(0)
    for w, h in ((1, 1), (7, 5), (300, 200), (130, 127)):     # 300x200 crosses the 65535-byte stored-block limit several times
        img = rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)
        path = tmp_path / f"t_{w}x{h}.png"
        vkr_host.write_png_rgba8(path, img)
        back = vkr_host.read_png_rgba8(path)                  # checks signature, IHDR fields, chunk CRCs, zlib Adler-32
        assert back.shape == img.shape and (back == img).all()
    with __import__("pytest").raises(vkr_host.HostError):
        vkr_host.write_png_rgba8(tmp_path / "no_such_dir" / "x.png", img)


def test_blit_same_extent_converts_the_format_only():
    """1:1 blit: the sample point is the texel centre, so the result is the texel itself pushed through the format conversion —
    fp16 linear -> sRGB8 with the OETF, alpha linear."""
    w, h = 40, 24
    rng = np.random.default_rng(1)
    src_v = rng.random((h, w, 4)).astype(np.float16)
    src = vkr.Image(w, h, vkr.R16G16B16A16_SFLOAT, device="cpu").upload(src_v)
    dst = vkr.Image(w, h, vkr.R8G8B8A8_SRGB, device="cpu")
    orc.ok(orc.lib.orc_blit_image(src.ptr, dst.ptr))
    got = dst.download().astype(np.int32)
    lin = src_v.astype(np.float64)
    enc = np.where(lin < 0.0031308, 12.92 * lin, 1.055 * np.power(np.maximum(lin, 1e-12), 1 / 2.4) - 0.055)
    expect = np.floor(enc * 255 + 0.5)
    expect[..., 3] = np.floor(lin[..., 3] * 255 + 0.5)
    assert np.abs(got - expect).max() <= 1 and (got != expect).mean() < 0.01
    # sRGB8 -> sRGB8 at the same extent is the identity (decode then encode of a representable code)
    codes = rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)
    a = vkr.Image(w, h, vkr.R8G8B8A8_SRGB, device="cpu").upload(codes)
    b = vkr.Image(w, h, vkr.R8G8B8A8_SRGB, device="cpu")
    orc.ok(orc.lib.orc_blit_image(a.ptr, b.ptr))
    assert (b.download() == codes).all()


def test_blit_halving_averages_2x2_blocks():
    """2:1 linear blit: destination centres fall on source texel corners -> the mean of the 2x2 block (UNORM, no transfer curve)."""
    w, h = 32, 16
    rng = np.random.default_rng(2)
    codes = rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)
    a = vkr.Image(w, h, vkr.R8G8B8A8_UNORM, device="cpu").upload(codes)
    b = vkr.Image(w // 2, h // 2, vkr.R8G8B8A8_UNORM, device="cpu")
    orc.ok(orc.lib.orc_blit_image(a.ptr, b.ptr))
    mean = codes.reshape(h // 2, 2, w // 2, 2, 4).astype(np.float64).mean(axis=(1, 3))
    assert np.abs(b.download().astype(np.float64) - mean).max() <= 0.5 + 1e-3


def test_pack_is_tight_and_applies_the_callback_fixups():
    w, h = 37, 11                                              # pitch align16(37*4) = 160 > 148: rows really are re-packed
    rng = np.random.default_rng(3)
    words = rng.integers(0, 2**32, size=(h, w), dtype=np.uint32)
    d = vkr.Image(w, h, vkr.D24_UNORM_S8_UINT, 3, device="cpu").upload(words)
    assert d.desc.row_pitch[0] == 160
    out = np.zeros((h, w), dtype=np.uint32)
    orc.ok(orc.lib.orc_readback_pack(d.ptr, 0, 0, out.ctypes.data))
    assert (out == words).all()
    orc.ok(orc.lib.orc_readback_pack(d.ptr, 0, 1, out.ctypes.data))
    assert (out == (words & 0xFFFFFF)).all()                   # main.cpp:156
    orc.ok(orc.lib.orc_readback_pack(d.ptr, 0, 2, out.ctypes.data))
    assert (out == (words | 0xFF000000)).all()                 # main.cpp:166
    assert orc.lib.orc_readback_pack(d.ptr, 3, 0, out.ctypes.data) != 0
