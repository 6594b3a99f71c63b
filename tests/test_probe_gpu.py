"""probe_renderer compute path through the C ABI (include/vkr_probe.h) vs the CPU oracle on identical bytes.

cube2oct and the probe pyramid are integer-valued outputs of exact arithmetic: bit-exact (a 1e-4 fraction allowed for cube2oct,
whose face pick / unorm rounding can sit on a last-place tie).  The probe trace ends in a discrete hit / miss per pixel: a
mismatching fraction <= 1e-3 of pixels, the rest within 1 LSB.
"""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as orc
import vkr
import vkr_probes
from test_oracle_anchors import _probe_setup

pytestmark = pytest.mark.gpu


def _to_dev(arr):
    d = vkr.ImageArray(arr.width, arr.height, arr.format, arr.mips, layers=arr.layers)
    d.buf.copy_(arr.buf)
    return d


@pytest.mark.parametrize("grid_sz,size,cube,objects", [(2, 128, 64, 0), (3, 256, 128, 48)])
def test_probe_grid_and_trace_match_oracle(cuda, grid_sz, size, cube, objects):
    import torch
    w, h = 320, 180
    scene, grid, cubes, gb, dp, params = _probe_setup(w, h, grid_sz=grid_sz, size=size, cube=cube, objects=objects)
    n = grid_sz * grid_sz
    for i in range(n):
        orc.ok(orc.lib.orc_cube2oct(cubes.color.layer(6 * i).ptr, cubes.color.layer_stride, cubes.distance.layer(6 * i).ptr, cubes.distance.layer_stride,
                                    grid.color_array.layer(i).ptr, grid.depth_array.layer(i).ptr))
        orc.ok(orc.lib.orc_probe_downsample(grid.depth_array.layer(i).ptr))
    dgrid = vkr_probes.OctahedralProbeGrid(grid_sz, size)
    dgrid.min, dgrid.max = grid.min, grid.max
    dcubes = vkr_probes.CubeMapSet(n, cube)
    dcubes.color.buf.copy_(cubes.color.buf)
    dcubes.distance.buf.copy_(cubes.distance.buf)
    pr = vkr_probes.ProbeRenderer(cube)
    pr.render_octahedral(dcubes, dgrid.color_array, dgrid.depth_array, 0, n)          # all probes, one launch
    pr.probe_downsample(dgrid.depth_array, 0, n)
    torch.cuda.synchronize()
    for i in range(n):
        gc, rc = dgrid.color_array.layer(i).download(), grid.color_array.layer(i).download()
        assert (gc != rc).any(axis=-1).mean() <= 1e-4, i
        assert np.abs(gc.astype(int) - rc.astype(int)).max() <= 1
        for lod in range(grid.depth_array.mips):
            gd, rd = dgrid.depth_array.layer(i).download(lod), grid.depth_array.layer(i).download(lod)
            assert (gd != rd).mean() <= 1e-4 and np.abs(gd.astype(int) - rd.astype(int)).max() <= 1, (i, lod)
    # per-layer calls (the reference's one-task-per-probe order) give the same bytes as the batched launch
    d2 = vkr_probes.OctahedralProbeGrid(grid_sz, size)
    for i in range(n):
        one = vkr_probes.CubeMapSet(1, cube)
        one.color.buf.copy_(cubes.color.buf[6 * i * cubes.color.layer_stride:(6 * i + 6) * cubes.color.layer_stride])
        one.distance.buf.copy_(cubes.distance.buf[6 * i * cubes.distance.layer_stride:(6 * i + 6) * cubes.distance.layer_stride])
        pr.render_octahedral(one, d2.color_array, d2.depth_array, i)
        pr.probe_downsample(d2.depth_array, i)
    torch.cuda.synchronize()
    assert torch.equal(d2.color_array.buf, dgrid.color_array.buf) and torch.equal(d2.depth_array.buf, dgrid.depth_array.buf)

    # trace on the ORACLE's probe bytes so that both sides march the same pyramid
    dgrid.color_array.buf.copy_(grid.color_array.buf)
    dgrid.depth_array.buf.copy_(grid.depth_array.buf)
    ref = vkr.Image(w, h, vkr.R8G8B8A8_UNORM, device="cpu")
    k = vkr_probes.trace_constants(grid, params)
    orc.ok(orc.lib.orc_trace_probe(C.byref(k), gb.depth.ptr, gb.normal.ptr, grid.color_array.ptr, grid.color_array.layer_stride, grid.depth_array.ptr,
                                   grid.depth_array.layer_stride, ref.ptr, 0, 0))
    out = vkr.Image(w, h, vkr.R8G8B8A8_UNORM)
    ddepth, dnormal = gb.depth.clone_to("cuda"), gb.normal.clone_to("cuda")
    vkr_probes.ProbeTracePass().run(dgrid, ddepth, dnormal, out, params)
    torch.cuda.synchronize()
    g, r = out.download().astype(np.int32), ref.download().astype(np.int32)
    assert (r[..., 3] > 0).mean() > 0.02, "some pixels should find a reflection"
    d = np.abs(g - r).max(axis=-1)
    assert (d > 1).mean() <= 1e-3, ((d > 1).sum(), d.max())
    # row windows write the bytes of the whole pass
    part = vkr.Image(w, h, vkr.R8G8B8A8_UNORM)
    mid = (h // 2) // vkr.ROW_ALIGN * vkr.ROW_ALIGN
    for rows in ((0, mid), (mid, h)):
        vkr_probes.ProbeTracePass().run(dgrid, ddepth, dnormal, part, params, ext=vkr.make_ext([], rows=rows))
    torch.cuda.synchronize()
    assert torch.equal(part.buf, out.buf)


def test_probe_validation(cuda):
    grid = vkr_probes.OctahedralProbeGrid(2, 64)
    cubes = vkr_probes.CubeMapSet(1, 32)
    with pytest.raises(vkr.VkrError):
        vkr.cube2oct(cubes.color, cubes.distance, grid.depth_array, grid.depth_array)          # colour target must be RGBA8_UNORM
    with pytest.raises(vkr.VkrError):
        vkr.probe_downsample(grid.color_array)                                                  # pyramid must be R16_UNORM
    k = vkr.ProbeTraceParams()
    k.grid_size = 1
    d, nrm, out = vkr.Image(64, 32, vkr.D24_UNORM_S8_UINT), vkr.Image(64, 32, vkr.R16G16_UNORM), vkr.Image(64, 32, vkr.R8G8B8A8_UNORM)
    with pytest.raises(vkr.VkrError):
        vkr.trace_probe(k, d, nrm, grid.color_array, grid.depth_array, out)                     # grid_size < 2 ("Ooops", probe_renderer.cpp:266)


def test_host_app_probe_grid_and_trace(cuda):
    """The C++ ProbeRenderer / ProbeTracePass classes behind the host app: the cube faces of the device producer equal the
    Python producer's, the octahedral probes and their pyramids equal the oracle's on those faces, and the traced reflection
    equals the oracle's on the app's own probe bytes."""
    import chain
    import vkr_host
    w, h = 320, 180
    scene = vkr.synth_scene(0, 32)
    app = vkr_host.App(w, h, device=0)
    pmin, pmax = (-3.0, 1.0, -5.0), (3.0, 1.0, 5.0)
    app.render_probe_grid(scene, pmin, pmax)
    tasks = app.last_tasks()
    assert tasks == ["CubemapProbe", "Cubemap2Octahedral", "DownsampleProbe"] * 16
    color, depth = app.readback_array("probe_color"), app.readback_array("probe_depth")
    cube_c, cube_d = app.readback_array("probe_cube_color"), app.readback_array("probe_cube_distance")
    grid = vkr_probes.OctahedralProbeGrid(4, 256, device="cpu")
    grid.min, grid.max = pmin, pmax
    # the cube map in the app holds the LAST probe of the grid
    last = vkr_probes.CubeMapSet(1, 128, device="cpu").render_synthetic(scene, [grid.positions()[15]])
    assert np.array_equal(cube_c.buf.numpy(), last.color.buf.numpy())
    dd = np.abs(cube_d.buf.numpy().view(np.int16).astype(np.int32) - last.distance.buf.numpy().view(np.int16).astype(np.int32))
    assert dd.max() <= 1                                                       # fp16 codes: at most one ulp between the two producers
    ref_c, ref_d = vkr.Image(256, 256, vkr.R8G8B8A8_UNORM, device="cpu"), vkr.Image(256, 256, vkr.R16_UNORM, depth.mips, device="cpu")
    orc.ok(orc.lib.orc_cube2oct(cube_c.ptr, cube_c.layer_stride, cube_d.ptr, cube_d.layer_stride, ref_c.ptr, ref_d.ptr))
    orc.ok(orc.lib.orc_probe_downsample(ref_d.ptr))
    assert (color.layer(15).download() != ref_c.download()).any(axis=-1).mean() <= 1e-4
    for lod in range(depth.mips):
        assert (depth.layer(15).download(lod) != ref_d.download(lod)).mean() <= 1e-4, lod
    assert all((depth.layer(i).download(3) > 0).any() for i in range(16))

    app.set_option("alt.trace_probe", 1)
    oc = chain.Chain(w, h, chain.OracleBackend(), lut_size=1024)
    dp = vkr.synth_camera(0, 0, w, h, True)
    P = oc.begin_frame(scene)
    app.upload_gbuffer(oc.gb)
    app.render_frame(dp, chain.angle_jitter(0))
    assert app.last_tasks()[-1] == "TraceProbe"
    grid.color_array.buf.copy_(color.buf)
    grid.depth_array.buf.copy_(depth.buf)
    params = vkr_probes.ProbeTraceParams(vkr.mat_inverse(dp.camera), *list(dp.fovy_aspect_znear_zfar))
    k = vkr_probes.trace_constants(grid, params)
    ref = vkr.Image(w, h, vkr.R8G8B8A8_UNORM, device="cpu")
    orc.ok(orc.lib.orc_trace_probe(C.byref(k), oc.gb.depth.ptr, oc.gb.normal.ptr, grid.color_array.ptr, grid.color_array.layer_stride, grid.depth_array.ptr,
                                   grid.depth_array.layer_stride, ref.ptr, 0, 0))
    g, r = app.readback("probe_reflection").download().astype(np.int32), ref.download().astype(np.int32)
    assert (r[..., 3] > 0).mean() > 0.02
    d = np.abs(g - r).max(axis=-1)
    assert (d > 1).mean() <= 1e-3, ((d > 1).sum(), d.max())
    app.close()
