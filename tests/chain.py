"""Frame-chain driver used by the parity tests, smoke() and bench.py's CPU legs.

Runs the reference's per-frame pass order (src/main.cpp:345-391) and ping-pong remaps
(:416-420) on either backend:
  * "oracle": oracle/liboracle.so on CPU images (the checker / CPU baseline),
  * "cuda":   libvkr_b200.so through the C ABI on device images.
Per-frame pass parameters follow SURVEY §8d: angle table[f%12]/360 + u_f - 0.5 (u_f from a
seeded generator instead of libc rand(), src/gtao.cpp:109-111), frame_random = f % 16
(src/advanced_ssr.cpp:153-171), flags = 7, histories zero, prev_depth cleared to 1.0.
"""
import ctypes as C

import numpy as np
import torch

import vkr

ANGLE_TABLE = [60., 300., 180., 240., 120., 0., 300., 60., 180., 120., 240., 0.]  # src/gtao.cpp:109


def angle_jitter(frame, seed=0):
    """Stand-in for rand()/RAND_MAX - 0.5 of src/gtao.cpp:111, reproducible."""
    u = np.random.default_rng(1000003 * seed + frame).random()
    return float(np.float32(np.float32(u) - np.float32(0.5)))


def angle_offset(frame, seed=0):
    return float(np.float32(np.float32(ANGLE_TABLE[frame % 12]) / np.float32(360.0)) + np.float32(angle_jitter(frame, seed)))


class OracleBackend:
    device = "cpu"

    def __init__(self, rows=None):
        import oracle_lib
        self.o = oracle_lib
        self.rows = rows  # (y0, y1) fractions for bounded-sample timing, None = full frame

    def _r(self, h):
        if self.rows is None:
            return 0, 0
        return int(self.rows[0] * h), max(int(self.rows[1] * h), int(self.rows[0] * h) + 1)

    def halton(self):
        return self.o.halton23_seq(128)

    def hiz(self, depth, normal, vel, on, ov):
        self.o.ok(self.o.lib.orc_downsample_gbuffer(depth.ptr, normal.ptr, vel.ptr, on.ptr, ov.ptr, 0, 0))
        self.o.ok(self.o.lib.orc_depth_mips(depth.ptr, 1))

    def pdf_lut(self, out):
        self.o.ok(self.o.lib.orc_pdf_preintegrate(out.ptr, 0, 0))

    def brdf_lut(self, halton, out):
        self.o.ok(self.o.lib.orc_brdf_preintegrate(self.o.fptr(halton), out.ptr, 0, 0))

    def trace(self, p, depth, dn, mat, halton, pdf, rays, occ):
        self.o.ok(self.o.lib.orc_sssr_trace(C.byref(p), depth.ptr, dn.ptr, mat.ptr, self.o.fptr(halton), pdf.ptr, rays.ptr, occ.ptr, *self._r(rays.height)))

    def filter(self, p, rays, depth, albedo, normal, mat, out):
        self.o.ok(self.o.lib.orc_sssr_filter(C.byref(p), rays.ptr, depth.ptr, albedo.ptr, normal.ptr, mat.ptr, out.ptr, *self._r(out.height)))

    def blur(self, p, depth, normal, refl, mat, hist, dv, pdepth, out):
        self.o.ok(self.o.lib.orc_sssr_blur(C.byref(p), depth.ptr, normal.ptr, refl.ptr, mat.ptr, hist.ptr, dv.ptr, pdepth.ptr, out.ptr, *self._r(out.height)))

    def gtao_main(self, p, depth, normal, mat, pdf, raw):
        self.o.ok(self.o.lib.orc_gtao_main(C.byref(p), depth.ptr, normal.ptr, mat.ptr, pdf.ptr, raw.ptr, 1, *self._r(raw.height)))

    def gtao_filter(self, p, depth, raw, filtered):
        self.o.ok(self.o.lib.orc_gtao_filter(C.byref(p), depth.ptr, raw.ptr, filtered.ptr, 1, *self._r(filtered.height)))

    def gtao_accumulate(self, p, depth, pdepth, filtered, acc, dv, hist):
        self.o.ok(self.o.lib.orc_gtao_accumulate(C.byref(p), depth.ptr, pdepth.ptr, filtered.ptr, acc.ptr, dv.ptr, hist.ptr, 1, *self._r(acc.height)))

    def shading(self, p, albedo, normal, mat, depth, occ, brdf, refl, out):
        self.o.ok(self.o.lib.orc_deferred_shading(C.byref(p), albedo.ptr, normal.ptr, mat.ptr, depth.ptr, occ.ptr, brdf.ptr, refl.ptr, out.ptr, *self._r(out.height)))

    def taa(self, p, hist, pdepth, depth, vel, color, out):
        self.o.ok(self.o.lib.orc_taa_resolve(C.byref(p), hist.ptr, pdepth.ptr, depth.ptr, vel.ptr, color.ptr, out.ptr, *self._r(out.height)))


class CudaBackend:
    device = "cuda"

    def __init__(self, fused_hiz=True):
        self.fused_hiz = fused_hiz
        self._halton_dev = None

    def halton(self):
        h = vkr.halton23_seq(128)
        self._halton_dev = torch.from_numpy(h).cuda()
        return h

    def _hd(self, halton):
        if self._halton_dev is None:
            self._halton_dev = torch.from_numpy(np.ascontiguousarray(halton)).cuda()
        return self._halton_dev

    def hiz(self, depth, normal, vel, on, ov):
        if self.fused_hiz:
            vkr.hiz_build(depth, normal, vel, on, ov)
        else:
            vkr.downsample_gbuffer(depth, normal, vel, on, ov)
            vkr.depth_mips(depth, 1)

    def pdf_lut(self, out):
        vkr.pdf_preintegrate(out)

    def brdf_lut(self, halton, out):
        vkr.brdf_preintegrate(self._hd(halton), out)

    def trace(self, p, depth, dn, mat, halton, pdf, rays, occ):
        vkr.sssr_trace(p, depth, dn, mat, self._hd(halton), pdf, rays, occ)

    def filter(self, *a):
        vkr.sssr_filter(*a)

    def blur(self, *a):
        vkr.sssr_blur(*a)

    def gtao_main(self, *a):
        vkr.gtao_main(*a, depth_lod=1)

    def gtao_filter(self, *a):
        vkr.gtao_filter(*a, depth_lod=1)

    def gtao_accumulate(self, *a):
        vkr.gtao_accumulate(*a, depth_lod=1)

    def shading(self, *a):
        vkr.deferred_shading(*a)

    def taa(self, *a):
        vkr.taa_resolve(*a)


def frame_params(dp, frame, counter, seed=0):
    """Pass parameter blocks for one frame from DrawTAAParams (main.cpp:366-373, pass .cpp files)."""
    inv_cam, inv_prev = vkr.mat_inverse(dp.camera), vkr.mat_inverse(dp.prev_camera)
    normal_mat = vkr.mat_transpose(inv_cam)                      # transpose(inverse(view)), main.cpp:368
    fovy, aspect, zn, zf = list(dp.fovy_aspect_znear_zfar)
    ident = (C.c_float * 16)(*np.eye(4, dtype=np.float32).reshape(16))

    def cp(dst, src):
        C.memmove(dst, src, 64)

    tr = vkr.TraceParams(); cp(tr.normal_mat, normal_mat)
    tr.frame_random, tr.fovy, tr.aspect, tr.znear, tr.zfar, tr.max_roughness, tr.render_flags = counter, fovy, aspect, zn, zf, 1.0, 7
    bl = vkr.BlurParams(); cp(bl.inverse_camera, inv_cam); cp(bl.prev_inverse_camera, inv_prev)
    bl.fovy_aspect_znear_zfar[:] = [fovy, aspect, zn, zf]; bl.max_roughness, bl.accumulate, bl.disable_blur = 1.0, 1, 0
    gt = vkr.GtaoParams(); cp(gt.normal_mat, normal_mat)
    gt.fovy, gt.aspect, gt.znear, gt.zfar = fovy, aspect, zn, zf
    gt.angle_offset, gt.weight_ratio, gt.use_mis, gt.two_directions, gt.reflections_only = angle_offset(frame, seed), 1.0, 1, 0, 0
    ac = vkr.AccumParams(); cp(ac.inverse_camera, inv_cam); cp(ac.prev_inverse_camera, inv_prev); cp(ac.mvp, dp.mvp)
    ac.fovy_aspect_znear_zfar[:] = [fovy, aspect, zn, zf]; ac.clear_history = 0
    sh = vkr.ShadingParams(); cp(sh.inverse_camera, inv_cam); cp(sh.camera, dp.camera); cp(sh.shadow_mvp, ident)
    sh.fovy, sh.aspect, sh.znear, sh.zfar = fovy, aspect, zn, zf
    sh.min_max_roughness[:] = [0.0, 1.0]; sh.show_ao = 0
    ta = vkr.TaaParams(); cp(ta.inverse_camera, inv_cam); cp(ta.prev_inverse_camera, inv_prev)
    ta.fovy_aspect_znear_zfar[:] = [fovy, aspect, zn, zf]
    return dict(trace=tr, blur=bl, gtao=gt, accum=ac, shade=sh, taa=ta)


class Chain:
    """All images of Gbuffer + GTAO + AdvancedSSR + TAA + colour target for one view (main.cpp:263-291)."""

    def __init__(self, w, h, backend, lut_size=1024, luts=None):
        self.w, self.h, self.b = w, h, backend
        dev = backend.device
        w2, h2 = w // 2, h // 2
        I = lambda ww, hh, f, m=1: vkr.Image(ww, hh, f, m, device=dev)  # noqa: E731
        self.gb = vkr.Gbuffer(w, h, device=dev)
        # prev_depth cleared to 1.0 (main.cpp:306): every mip = 0xFFFFFF
        for lod in range(self.gb.prev_depth.mips):
            mw, mh = self.gb.prev_depth.mip_size(lod)
            self.gb.prev_depth.upload(np.full((mh, mw), 0xFFFFFF, dtype=np.uint32), lod)
        self.raw = I(w2, h2, vkr.R16G16B16A16_SFLOAT)                       # gtao.cpp:27-33
        self.filtered = I(w2, h2, vkr.R16_SFLOAT)
        self.accumulated_ao = I(w2, h2, vkr.R16G16_SFLOAT)
        self.accumulated_history = I(w2, h2, vkr.R16G16_SFLOAT)
        self.rays = I(w2, h2, vkr.R16G16B16A16_UNORM)                       # advanced_ssr.cpp:62-73
        self.reflections = I(w2, h2, vkr.R8G8B8A8_UNORM)
        self.blurred = I(w2, h2, vkr.R8G8B8A8_UNORM)
        self.blurred_history = I(w2, h2, vkr.R8G8B8A8_UNORM)
        self.color_out = I(w, h, vkr.R8G8B8A8_SRGB)                         # main.cpp:289-291
        self.taa_history = I(w, h, vkr.R16G16B16A16_SFLOAT)                 # taa.cpp:6-10
        self.taa_target = I(w, h, vkr.R16G16B16A16_SFLOAT)
        self.halton = backend.halton()
        if luts is None:
            self.pdf = I(lut_size, lut_size, vkr.R32_SFLOAT)                # advanced_ssr.cpp:88-92
            self.brdf = I(lut_size, lut_size, vkr.R16G16_SFLOAT)
            backend.pdf_lut(self.pdf)                                       # main.cpp:269-270
            backend.brdf_lut(self.halton, self.brdf)
        else:
            self.pdf, self.brdf = luts
        self.counter = 0
        self.frame_index = 0

    def steps(self, P, skip_lighting=False):
        """main.cpp:347-391 on the current G-buffer, as (task name, thunk, names of the images it writes)."""
        b, gb = self.b, self.gb
        st = [
            ("DownsampleGbuffer+DownsampleDepth", lambda: b.hiz(gb.depth, gb.normal, gb.velocity_vectors, gb.downsampled_normals,
                                                                gb.downsampled_velocity_vectors), ["depth", "dn", "dv"]),
            ("SSSR_trace", lambda: b.trace(P["trace"], gb.depth, gb.downsampled_normals, gb.material, self.halton, self.pdf, self.rays, self.raw),
             ["rays", "raw"]),
            ("SSSR_filter", lambda: b.filter(P["trace"], self.rays, gb.depth, gb.albedo, gb.normal, gb.material, self.reflections), ["reflections"]),
            ("SSSR_blur", lambda: b.blur(P["blur"], gb.depth, gb.normal, self.reflections, gb.material, self.blurred_history,
                                         gb.downsampled_velocity_vectors, gb.prev_depth, self.blurred), ["blurred"]),
            ("GTAO_main", lambda: b.gtao_main(P["gtao"], gb.depth, gb.normal, gb.material, self.pdf, self.raw), ["raw"]),
            ("GTAO_filter", lambda: b.gtao_filter(P["gtao"], gb.depth, self.raw, self.filtered), ["filtered"]),
            ("GTAO_accumulate", lambda: b.gtao_accumulate(P["accum"], gb.depth, gb.prev_depth, self.filtered, self.accumulated_ao,
                                                          gb.downsampled_velocity_vectors, self.accumulated_history), ["accumulated_ao"]),
        ]
        if not skip_lighting:
            st.append(("DeferedShading", lambda: b.shading(P["shade"], gb.albedo, gb.normal, gb.material, gb.depth, self.accumulated_ao, self.brdf,
                                                           self.blurred, self.color_out), ["color"]))
        st.append(("TAA", lambda: b.taa(P["taa"], self.taa_history, gb.prev_depth, gb.depth, gb.velocity_vectors, self.color_out, self.taa_target),
                   ["taa"]))
        return st

    def passes(self, P, skip_lighting=False):
        for _, thunk, _ in self.steps(P, skip_lighting):
            thunk()

    def images(self):
        """Every image slot by name (current assignment of the ping-pong pairs)."""
        gb = self.gb
        return dict(albedo=gb.albedo, normal=gb.normal, material=gb.material, velocity=gb.velocity_vectors, depth=gb.depth,
                    prev_depth=gb.prev_depth, dn=gb.downsampled_normals, dv=gb.downsampled_velocity_vectors, rays=self.rays, raw=self.raw,
                    reflections=self.reflections, blurred=self.blurred, blurred_history=self.blurred_history, filtered=self.filtered,
                    accumulated_ao=self.accumulated_ao, accumulated_history=self.accumulated_history, color=self.color_out,
                    taa=self.taa_target, taa_history=self.taa_history, pdf=self.pdf, brdf=self.brdf)

    def begin_frame(self, scene, view_seed=0, jitter=True, seed=0, gbuffer_from=None):
        f = self.frame_index
        dp = vkr.synth_camera(view_seed, f, self.w, self.h, jitter)
        if gbuffer_from is None:
            vkr.synth_gbuffer(scene, dp, self.gb)
        else:                                                # copy inputs produced elsewhere (identical bytes on both backends)
            for dst, src in zip(self.gb.input_images(), gbuffer_from.input_images()):
                dst.buf.copy_(src.buf)
        P = frame_params(dp, f, self.counter, seed)
        self.counter = (self.counter + 1) % 16               # advanced_ssr.cpp:166-171
        return P

    def end_frame(self):
        self.remap()
        self.frame_index += 1

    def remap(self):
        """main.cpp:416-420 — swap slots after submit."""
        gb = self.gb
        gb.depth, gb.prev_depth = gb.prev_depth, gb.depth
        self.taa_history, self.taa_target = self.taa_target, self.taa_history
        self.blurred, self.blurred_history = self.blurred_history, self.blurred
        self.accumulated_history, self.accumulated_ao = self.accumulated_ao, self.accumulated_history

    def render(self, scene, view_seed=0, jitter=True, seed=0, gbuffer_from=None):
        P = self.begin_frame(scene, view_seed, jitter, seed, gbuffer_from)
        self.passes(P)
        out = {k: v for k, v in self.images().items()}
        self.end_frame()
        return out, P
