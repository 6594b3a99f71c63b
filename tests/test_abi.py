"""The C-ABI libraries load on a machine without a GPU and export every symbol include/*.h declares; compute entries
refuse to run without a device instead of falling back (no compute calls here)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "vk-renderer_b200")
DECL = re.compile(r"^\s*(?:const\s+)?[A-Za-z_][\w\s\*]*?\b(vkr_\w+)\s*\(", re.M)


def declared(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(DECL.findall(text)))


@pytest.mark.parametrize("header,libname", [("vkr_passes.h", "libvkr_b200.so"), ("vkr_synth.h", "libvkr_b200.so"), ("vkr_host.h", "libvkr_host.so"),
                                            ("vkr_peer.h", "libvkr_b200.so"), ("vkr_probe.h", "libvkr_b200.so")])
def test_every_declared_symbol_is_exported(header, libname):
    if not os.path.exists(os.path.join(ROOT, "include", header)):
        pytest.skip(f"{header} not present")
    names = declared(header)
    assert len(names) >= 3, names
    lib = C.CDLL(os.path.join(PKG, libname))
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_headers_compile_as_plain_c(tmp_path):
    """The boundary is C: no C++ or torch types in the signatures."""
    import subprocess
    src = tmp_path / "abi.c"
    hdrs = [h for h in sorted(os.listdir(os.path.join(ROOT, "include"))) if h.endswith(".h")]
    src.write_text("".join(f'#include "{h}"\n' for h in hdrs) + "int main(void){return (int)sizeof(VkrImage) - 176;}\n")
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-c", str(src), "-o", str(tmp_path / "abi.o")], check=True)


def test_version_and_layout_helpers_work_without_a_gpu():
    import vkr
    assert vkr.lib.vkr_full_mip_count(3840, 2160) == 12 and vkr.lib.vkr_full_mip_count(256, 256) == 9    # scene_renderer.cpp:13
    img = vkr.VkrImage()
    vkr.lib.vkr_image_layout.restype = C.c_uint64
    n = vkr.lib.vkr_image_layout(250, 130, vkr.D24_UNORM_S8_UINT, 8, C.byref(img))
    assert n > 0 and img.row_pitch[0] == 1008 and all(img.mip_offset[i] % 256 == 0 for i in range(8))
    assert vkr.lib.vkr_format_texel_size(vkr.R16G16B16A16_SFLOAT) == 8


def test_compute_entry_fails_loudly_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    import vkr
    d = vkr.Image(64, 64, vkr.D24_UNORM_S8_UINT, 7, device="cpu")
    n = vkr.Image(64, 64, vkr.R16G16_UNORM, device="cpu")
    v = vkr.Image(64, 64, vkr.R16G16_SFLOAT, device="cpu")
    on = vkr.Image(32, 32, vkr.R16G16_UNORM, device="cpu")
    ov = vkr.Image(32, 32, vkr.R16G16_SFLOAT, device="cpu")
    assert vkr.lib.vkr_hiz_build(d.ptr, n.ptr, v.ptr, on.ptr, ov.ptr, None) == -3      # VKR_ERR_CUDA: no CPU fallback


def test_pixel_uv_division_is_exact(tmp_path):
    """common.cuh pixel_uv / texel_uv: q = k*r, e = fma(-q, W, k), q + e*r with r = RN(1/W) equals the IEEE quotient for
    every extent W <= 16384 and every pixel p < W, for k = p + 0.5 and k = p (host check with hardware FMA)."""
    import subprocess
    src = tmp_path / "div.c"
    src.write_text("""
#include <math.h>
#include <stdio.h>
int main(void) {
  long bad = 0;
  for (int w = 1; w <= 16384; w++) {
    float d = (float)w, r = 1.0f / d;
    for (int p = 0; p < w; p++) {
      float k = (float)p + 0.5f, q = k * r, e = fmaf(-q, d, k);
      if (fmaf(e, r, q) != k / d) bad++;
      k = (float)p; q = k * r; e = fmaf(-q, d, k);
      if (fmaf(e, r, q) != k / d) bad++;
    }
  }
  printf("%ld", bad);
  return bad != 0;
}
""")
    exe = tmp_path / "div"
    subprocess.run(["gcc", "-O2", "-mfma", "-ffp-contract=off", "-o", str(exe), str(src), "-lm"], check=True)
    r = subprocess.run([str(exe)], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.strip() == "0", r.stdout


def test_new_entries_validate_arguments_before_touching_the_device():
    """The tile-path and capture entries reject bad arguments with VKR_ERR_INVALID_ARGUMENT / VKR_ERR_UNSUPPORTED and a message,
    on a machine without a GPU too (validation comes before any CUDA call)."""
    import vkr
    L = vkr.lib
    L.vkr_last_error.restype = C.c_char_p
    mat = vkr.Image(64, 64, vkr.R8G8B8A8_SRGB, device="cpu")
    wrong = vkr.Image(64, 64, vkr.R8G8B8A8_UNORM, device="cpu")
    buf = (C.c_int32 * 64)()
    ind = (C.c_uint32 * 3)()
    p = vkr.ClassifyParams(32, 32, 1.0, 0.5)
    assert L.vkr_sssr_clear_indirect(None, ind, None) == -1
    assert L.vkr_sssr_classification(C.byref(p), wrong.ptr, buf, buf, ind, ind, None) == -1      # material must be RGBA8_SRGB
    assert b"sssr_classification" in L.vkr_last_error()
    assert L.vkr_sssr_classification(C.byref(vkr.ClassifyParams(0, 32, 1.0, 0.5)), mat.ptr, buf, buf, ind, ind, None) == -1
    tp = vkr.TraceParams()
    d = vkr.Image(64, 64, vkr.D24_UNORM_S8_UINT, 7, device="cpu")
    dn = vkr.Image(32, 32, vkr.R16G16_UNORM, device="cpu")
    rays = vkr.Image(32, 32, vkr.R16G16B16A16_UNORM, device="cpu")
    assert L.vkr_sssr_trace_indirect(C.byref(tp), 2, d.ptr, dn.ptr, mat.ptr, buf, buf, ind, 16, rays.ptr, None) == -1   # reflection_type > 1
    assert L.vkr_sssr_trace_indirect(C.byref(tp), 0, d.ptr, dn.ptr, mat.ptr, buf, buf, ind, 0, rays.ptr, None) == 0     # empty list capacity: no-op
    planes = vkr.Image(4, 4, vkr.R32G32B32A32_SFLOAT, device="cpu")
    assert L.vkr_sssr_tile_regression(None, d.ptr, planes.ptr, None) == -1
    assert L.vkr_sssr_tile_regression(C.byref(vkr.RegressionParams()), d.ptr, rays.ptr, None) == -1                    # planes must be RGBA32F
    assert L.vkr_blit_image(d.ptr, mat.ptr, None) == -4                                                                  # depth is not a blit source
    assert L.vkr_blit_image(None, mat.ptr, None) == -1
    out = (C.c_uint8 * (64 * 64 * 8))()
    assert L.vkr_readback_pack(mat.ptr, 1, 0, out, None) == -1                                                           # mip out of range
    assert L.vkr_readback_pack(mat.ptr, 0, 3, out, None) == -1                                                           # unknown mode
    assert L.vkr_readback_pack(rays.ptr, 0, 1, out, None) == -4                                                          # fix-up modes need 4-byte texels
    e = vkr.LaunchExt()
    e.first_param = 3                                                                                                    # needs layers > 1 + PARAMS_PREPARED
    raw = vkr.Image(32, 32, vkr.R16G16B16A16_SFLOAT, device="cpu")
    filt = vkr.Image(32, 32, vkr.R16_SFLOAT, device="cpu")
    assert L.vkr_gtao_filter_ex(C.byref(vkr.GtaoParams()), d.ptr, raw.ptr, filt.ptr, 1, C.byref(e), None) == -1
    assert b"first_param" in L.vkr_last_error()


def test_misaligned_or_short_pitched_images_are_refused():
    """the kernels read rows as 128-bit vectors, so an image whose base / mip offset / pitch is not
    16-byte aligned, or whose pitch does not hold a row, must come back as VKR_ERR_INVALID_ARGUMENT naming the field --
    from every pass entry (they share image_ok) and from vkr_image_validate -- instead of faulting on the device."""
    import torch
    import vkr
    L = vkr.lib
    L.vkr_last_error.restype = C.c_char_p
    L.vkr_image_validate.restype = C.c_int
    L.vkr_image_validate.argtypes = [C.POINTER(vkr.VkrImage)]
    INVALID = -1
    w, h = 30, 8                                        # 30 texels x 4 bytes = 120: a tight pitch is not a multiple of 16

    def img(fmt=vkr.R8G8B8A8_SRGB, ww=w, hh=h):
        return vkr.Image(ww, hh, fmt, device="cpu")

    good = img()
    assert L.vkr_image_validate(good.ptr) == 0
    # (a) base offset by 4 bytes
    buf = torch.zeros(good.buf.numel() + 16, dtype=torch.uint8)
    off = vkr.Image.over(buf[4:], w, h, vkr.R8G8B8A8_SRGB)
    if off.desc.base % 16 == 0:                          # the allocation itself happened to sit at 16k + 12
        off = vkr.Image.over(buf[8:], w, h, vkr.R8G8B8A8_SRGB)
    assert L.vkr_image_validate(off.ptr) == INVALID and b"base" in L.vkr_last_error()
    # (b) tight pitch w*4 with w % 4 != 0
    tight = img()
    tight.desc.row_pitch[0] = w * 4
    assert L.vkr_image_validate(tight.ptr) == INVALID and b"row_pitch[0] = 120" in L.vkr_last_error()
    # (c) pitch aligned but shorter than a row; (d) mip offset not aligned
    short = img()
    short.desc.row_pitch[0] = 112
    assert L.vkr_image_validate(short.ptr) == INVALID and b"smaller than a row" in L.vkr_last_error()
    depth = vkr.Image(32, 32, vkr.D24_UNORM_S8_UINT, 3, device="cpu")
    depth.desc.mip_offset[1] += 4
    assert L.vkr_image_validate(depth.ptr) == INVALID and b"mip_offset[1]" in L.vkr_last_error()
    # the pass entries apply the same rule before any CUDA call: TAA with a tight-pitched colour image
    hist, out = img(vkr.R16G16B16A16_SFLOAT), img(vkr.R16G16B16A16_SFLOAT)
    pd, d, vel = img(vkr.D24_UNORM_S8_UINT), img(vkr.D24_UNORM_S8_UINT), img(vkr.R16G16_SFLOAT)
    rc = L.vkr_taa_resolve(C.byref(vkr.TaaParams()), hist.ptr, pd.ptr, d.ptr, vel.ptr, tight.ptr, out.ptr, None)
    assert rc == INVALID and b"image layout: row_pitch[0] = 120" in L.vkr_last_error(), (rc, L.vkr_last_error())


def test_cpu_arm_never_loads_the_product_library(tmp_path):
    """bench.py --impl reference lays out images and produces its input G-buffers through oracle/libvkr_hostonly.so (the host
    helpers of csrc/api.cu, synth.cu, synth_scene.cu compiled by g++ with VKR_HOST_ONLY: no device code, no CUDA runtime call).
    The process must map that and liboracle.so only, the twin must write the bytes the product's host twin writes, and a
    compute entry must raise instead of silently doing nothing."""
    import subprocess
    import sys
    subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "libvkr_hostonly.so"], check=True, capture_output=True)
    code = r"""
import json, os, sys
sys.argv = ['bench.py', '--impl', 'reference', '--steps', '1', '--warmup', '0', '--width', '256', '--height', '144']
sys.path.insert(0, %r)
import runpy
runpy.run_path(os.path.join(%r, 'bench.py'), run_name='__main__')
maps = open('/proc/self/maps').read()
assert 'libvkr_b200' not in maps and 'libvkr_host.so' not in maps, 'product library mapped in the CPU arm'
assert 'libvkr_hostonly.so' in maps and 'liboracle.so' in maps
import vkr, numpy as np
try:
    vkr.lib.vkr_taa_resolve(None)
    raise SystemExit('compute entry did not raise')
except RuntimeError:
    pass
gb = vkr.Gbuffer(96, 64, device='cpu')
vkr.synth_gbuffer(vkr.synth_scene(1, 16), vkr.synth_camera(1, 2, 96, 64, True), gb)
np.save(%r, np.concatenate([i.buf.numpy() for i in gb.input_images()]))
""" % (ROOT, ROOT, str(tmp_path / "ho.npy"))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env={**os.environ, "OMP_NUM_THREADS": "2"})
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1]
    import json
    assert json.loads(line)["native_libraries"] == ["oracle/liboracle.so", "oracle/libvkr_hostonly.so"]
    import numpy as np
    import vkr
    gb = vkr.Gbuffer(96, 64, device="cpu")
    vkr.synth_gbuffer(vkr.synth_scene(1, 16), vkr.synth_camera(1, 2, 96, 64, True), gb)
    assert np.array_equal(np.load(tmp_path / "ho.npy"), np.concatenate([i.buf.numpy() for i in gb.input_images()]))
