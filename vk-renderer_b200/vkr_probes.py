"""Host-side mirror of the reference's probe classes (src/probe_renderer.hpp:9-81) over the C ABI of include/vkr_probe.h.

  OctahedralProbeGrid   color_array RGBA8_UNORM + depth_array R16_UNORM with a full mip chain, grid_size^2 layers (:19-29)
  ProbeRenderer         render_octahedral / probe_downsample / render_probe_grid (:31-53); the cube faces, which the
                        reference rasterises (render_cubemap), come from a caller-supplied producer — here the synthetic
                        scene's analytic ray-caster seen through six 90-degree cameras
  ProbeTracePass        run(grid, gbuffer_depth, gbuffer_norm, out_image, params) (:63-79)

No CPU fallback: every call goes to libvkr_b200.so and fails without a CUDA device.
"""
import ctypes as C
import math

import numpy as np
import torch

import vkr

PROBE_SIZE = 256    # src/probe_renderer.hpp:6-7
CUBE_SIZE = 128

# Cube face f sees direction s' * X + t' * Y - Z for face coordinates s', t' in [-1, 1] (Vulkan face table, +x -x +y -y +z -z)
FACE_AXES = [
    ((0, 0, -1), (0, -1, 0), (-1, 0, 0)),
    ((0, 0, 1), (0, -1, 0), (1, 0, 0)),
    ((1, 0, 0), (0, 0, 1), (0, -1, 0)),
    ((1, 0, 0), (0, 0, -1), (0, 1, 0)),
    ((1, 0, 0), (0, -1, 0), (0, 0, -1)),
    ((-1, 0, 0), (0, -1, 0), (0, 0, 1)),
]
ZNEAR, ZFAR = 0.05, 80.0


def face_camera(face, pos):
    """DrawTAAParams of a 90-degree square camera at `pos` whose pixel (s, t) looks along cube face `face`'s texel (s, t)."""
    X, Y, Z = FACE_AXES[face]
    inv_view = np.eye(4, dtype=np.float64)
    inv_view[:3, 0], inv_view[:3, 1], inv_view[:3, 2], inv_view[:3, 3] = X, Y, Z, pos
    view = np.linalg.inv(inv_view)
    dp = vkr.DrawTAAParams()
    col = (C.c_float * 16)(*view.T.reshape(16).astype(np.float32))           # column-major
    for field in (dp.camera, dp.prev_camera, dp.mvp, dp.prev_mvp):
        C.memmove(field, col, 64)
    dp.fovy_aspect_znear_zfar[:] = [math.pi / 2, 1.0, ZNEAR, ZFAR]
    return dp


class CubeMapSet:
    """Cube colour (RGBA8_SRGB) + radial distance (R16_SFLOAT) for `count` probes: 6 * count faces each."""

    def __init__(self, count=1, size=CUBE_SIZE, device="cuda"):
        self.count, self.size, self.device = count, size, device
        self.color = vkr.ImageArray(size, size, vkr.R8G8B8A8_SRGB, 1, layers=6 * count, device=device)
        self.distance = vkr.ImageArray(size, size, vkr.R16_SFLOAT, 1, layers=6 * count, device=device)

    def render_synthetic(self, scene, positions):
        """Stand-in for ProbeRenderer::render_cubemap (reference :78-168): the synthetic scene ray-cast through each face."""
        n = self.size
        gb = vkr.Gbuffer(n, n, device=self.device)
        c = (2.0 * (torch.arange(n, dtype=torch.float32, device=self.device) + 0.5) / n - 1.0)
        ray_len = torch.sqrt(1.0 + c[None, :] ** 2 + c[:, None] ** 2)        # |(s', t', -1)|, tan(45 deg) = 1
        for i, pos in enumerate(positions):
            for f in range(6):
                vkr.synth_gbuffer(scene, face_camera(f, pos), gb)
                if self.device == "cuda":
                    torch.cuda.synchronize()
                d24 = torch.from_numpy(gb.depth.download(0).astype(np.int64)).to(self.device)
                d = d24.to(torch.float64) / 16777215.0
                z = ZNEAR * ZFAR / (d * (ZFAR - ZNEAR) - ZFAR)               # gbuffer_encode.glsl:53-56 (negative in front)
                dist = (-z).to(torch.float32) * ray_len
                self.distance.layer(6 * i + f).upload(dist.to(torch.float16).cpu().numpy())
                self.color.layer(6 * i + f).buf.copy_(gb.albedo.buf)
        return self


class OctahedralProbeGrid:
    def __init__(self, grid_sz=4, size=PROBE_SIZE, device="cuda"):
        self.grid_size, self.size = grid_sz, size
        self.min = (0.0, 0.0, 0.0)
        self.max = (0.0, 0.0, 0.0)
        mips = int(math.floor(math.log2(size))) + 1                          # probe_renderer.cpp:306
        self.color_array = vkr.ImageArray(size, size, vkr.R8G8B8A8_UNORM, 1, layers=grid_sz * grid_sz, device=device)
        self.depth_array = vkr.ImageArray(size, size, vkr.R16_UNORM, mips, layers=grid_sz * grid_sz, device=device)

    def positions(self):
        """min + step * (x, 0, y), layer = y * grid_size + x (probe_renderer.cpp:271-276)."""
        g = self.grid_size
        step = [(b - a) / float(g - 1) for a, b in zip(self.min, self.max)]
        return [(self.min[0] + step[0] * x, self.min[1], self.min[2] + step[2] * y) for y in range(g) for x in range(g)]


class ProbeRenderer:
    def __init__(self, cubemap_res=CUBE_SIZE, device="cuda"):
        self.cubemap_res, self.device = cubemap_res, device

    def render_octahedral(self, cubes, probe_color, probe_depth, array_layer=0, probe_count=1, stream=None):
        vkr.cube2oct(cubes.color, cubes.distance, probe_color, probe_depth, array_layer, probe_count, stream)

    def probe_downsample(self, probe_depth, array_layer=0, layer_count=1, stream=None):
        vkr.probe_downsample(probe_depth, array_layer, layer_count, stream)

    def render_probe_grid(self, scene, pmin, pmax, grid, stream=None):
        """reference :253-283; all probes of the grid go through ONE cube2oct launch and one launch per pyramid level."""
        if grid.grid_size < 2:
            raise RuntimeError("Ooops")                                       # :266-268
        grid.min = tuple(min(a, b) for a, b in zip(pmin, pmax))
        grid.max = tuple(max(a, b) for a, b in zip(pmin, pmax))
        n = grid.grid_size * grid.grid_size
        cubes = CubeMapSet(n, self.cubemap_res, self.device).render_synthetic(scene, grid.positions())
        self.render_octahedral(cubes, grid.color_array, grid.depth_array, 0, n, stream)
        self.probe_downsample(grid.depth_array, 0, n, stream)
        return cubes


class ProbeTraceParams:
    """reference src/probe_renderer.hpp:55-61"""

    def __init__(self, inv_view, fovy, aspect, znear, zfar):
        self.inv_view, self.fovy, self.aspect, self.znear, self.zfar = inv_view, fovy, aspect, znear, zfar


def trace_constants(grid, params):
    """The Constants block ProbeTracePass::run builds (:340-360)."""
    k = vkr.ProbeTraceParams()
    C.memmove(k.inverse_view, params.inv_view, 64)
    k.probe_min[:] = [*grid.min, 1.0]
    k.probe_max[:] = [*grid.max, 1.0]
    k.grid_size = grid.grid_size
    k.fovy, k.aspect, k.znear, k.zfar = params.fovy, params.aspect, params.znear, params.zfar
    return k


class ProbeTracePass:
    def run(self, probe, gbuffer_depth, gbuffer_norm, out_image, params, ext=None, stream=None):
        vkr.trace_probe(trace_constants(probe, params), gbuffer_depth, gbuffer_norm, probe.color_array, probe.depth_array, out_image, ext, stream)
