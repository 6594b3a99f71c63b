"""Batched independent views (BASELINE.json configs[3]: 512 probe views of 256x256, GTAO + SSR) and their sharding.

A view batch keeps every image of the per-view chain as an ARRAY image (vkr.ImageArray: one allocation, one layer per
view) and runs each pass ONCE for the whole batch through the `_ex` entries with layers on grid.z — a single 256x256
view has 64..256 CTAs and cannot fill 148 SMs, 512 of them in one launch can.  Views are independent units
(SURVEY §8e): a rank owns a contiguous block of views, keeps their temporal state, and no collective touches the data
path; only the final outputs are gathered (NCCL all_gather on GPUs, gloo in the CPU tests).

The pass order is the reference's (src/main.cpp:347-390) without lighting and TAA, which probe views do not use
(reference src/probe_renderer.cpp renders cube faces and traces them; config 4 asks for GTAO + SSR on those views).
"""
import ctypes as C

import numpy as np
import torch

import vkr

ANGLE_TABLE = [60., 300., 180., 240., 120., 0., 300., 60., 180., 120., 240., 0.]  # src/gtao.cpp:109


def shard_range(n_units, world, rank):
    """Contiguous block [begin, end) of units owned by `rank`; the first n_units % world ranks get one extra."""
    base, extra = divmod(n_units, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def shard_sizes(n_units, world):
    return [shard_range(n_units, world, r)[1] - shard_range(n_units, world, r)[0] for r in range(world)]


def gather_blocks(local, n_units, group=None, out=None):
    """All-gather per-rank blocks of shape (units_r, ...) into (n_units, ...) on every rank (ragged blocks are padded
    to the largest one for the collective and trimmed afterwards).  Works on CUDA tensors over NCCL and CPU tensors
    over gloo.  Equal blocks go through ONE all_gather_into_tensor straight into `out` (preallocated (n_units, ...))."""
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return local
    sizes = shard_sizes(n_units, world)
    m = max(sizes)
    if min(sizes) == m and local.is_contiguous():
        if out is None:
            out = local.new_empty((n_units,) + tuple(local.shape[1:]))
        dist.all_gather_into_tensor(out, local, group=group)
        return out
    pad = local
    if local.shape[0] < m:
        pad = torch.cat([local, local.new_zeros((m - local.shape[0],) + tuple(local.shape[1:]))], dim=0)
    out = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(out, pad.contiguous(), group=group)
    return torch.cat([o[:s] for o, s in zip(out, sizes)], dim=0)


def _angle_offset(frame, jitter):
    return float(np.float32(np.float32(ANGLE_TABLE[frame % 12]) / np.float32(360.0)) + np.float32(jitter))


def view_params(dp, frame, counter, jitter=0.0):
    """Per-view parameter blocks of the GTAO + SSR passes from DrawTAAParams (src/main.cpp:366-373 and the pass .cpp files)."""
    inv_cam, inv_prev = vkr.mat_inverse(dp.camera), vkr.mat_inverse(dp.prev_camera)
    normal_mat = vkr.mat_transpose(inv_cam)
    fovy, aspect, zn, zf = list(dp.fovy_aspect_znear_zfar)
    tr = vkr.TraceParams(); C.memmove(tr.normal_mat, normal_mat, 64)
    tr.frame_random, tr.fovy, tr.aspect, tr.znear, tr.zfar, tr.max_roughness, tr.render_flags = counter, fovy, aspect, zn, zf, 1.0, 7
    bl = vkr.BlurParams(); C.memmove(bl.inverse_camera, inv_cam, 64); C.memmove(bl.prev_inverse_camera, inv_prev, 64)
    bl.fovy_aspect_znear_zfar[:] = [fovy, aspect, zn, zf]; bl.max_roughness, bl.accumulate, bl.disable_blur = 1.0, 1, 0
    gt = vkr.GtaoParams(); C.memmove(gt.normal_mat, normal_mat, 64)
    gt.fovy, gt.aspect, gt.znear, gt.zfar = fovy, aspect, zn, zf
    gt.angle_offset, gt.weight_ratio, gt.use_mis, gt.two_directions, gt.reflections_only = _angle_offset(frame, jitter), 1.0, 1, 0, 0
    ac = vkr.AccumParams(); C.memmove(ac.inverse_camera, inv_cam, 64); C.memmove(ac.prev_inverse_camera, inv_prev, 64); C.memmove(ac.mvp, dp.mvp, 64)
    ac.fovy_aspect_znear_zfar[:] = [fovy, aspect, zn, zf]; ac.clear_history = 0
    ident = (C.c_float * 16)(*np.eye(4, dtype=np.float32).reshape(16))
    sh = vkr.ShadingParams(); C.memmove(sh.inverse_camera, inv_cam, 64); C.memmove(sh.camera, dp.camera, 64); C.memmove(sh.shadow_mvp, ident, 64)
    sh.fovy, sh.aspect, sh.znear, sh.zfar = fovy, aspect, zn, zf
    sh.min_max_roughness[:] = [0.0, 1.0]; sh.show_ao = 0                      # defered_shading.hpp:30-31
    ta = vkr.TaaParams(); C.memmove(ta.inverse_camera, inv_cam, 64); C.memmove(ta.prev_inverse_camera, inv_prev, 64)
    ta.fovy_aspect_znear_zfar[:] = [fovy, aspect, zn, zf]
    return dict(trace=tr, blur=bl, gtao=gt, accum=ac, shade=sh, taa=ta)


class ViewBatch:
    """`views` independent w x h views on one device: images as arrays, passes batched over grid.z."""

    def __init__(self, w, h, view_ids, device="cuda", luts=None, lut_size=1024, stream=None, param_sets=1, fused=None):
        """fused: a vkr_gather.FusedGather whose units are ALL views of the job (this batch = its rank's shard): the two gathered
        results (accumulated AO, blurred reflections) and their histories then live in its peer-shared full arrays, the blur and
        the accumulate pass store every texel into every peer's copy as well, and passes() brackets them with the epoch kernels."""
        self.w, self.h, self.view_ids, self.V = w, h, list(view_ids), len(view_ids)
        self.device, self.stream = torch.device(device), stream
        V, w2, h2 = self.V, w // 2, h // 2
        mips = vkr.full_mip_count(w, h)
        A = lambda ww, hh, f, m=1: vkr.ImageArray(ww, hh, f, m, V, device=device)  # noqa: E731
        self.albedo, self.normal, self.material = A(w, h, vkr.R8G8B8A8_SRGB), A(w, h, vkr.R16G16_UNORM), A(w, h, vkr.R8G8B8A8_SRGB)
        self.velocity = A(w, h, vkr.R16G16_SFLOAT)
        self.depth, self.prev_depth = A(w, h, vkr.D24_UNORM_S8_UINT, mips), A(w, h, vkr.D24_UNORM_S8_UINT, mips)
        self.prev_depth.buf.view(torch.int32).fill_(0x00FFFFFF)           # clear_depth(prev_depth, 1.0), src/main.cpp:306
        self.dn, self.dv = A(w2, h2, vkr.R16G16_UNORM), A(w2, h2, vkr.R16G16_SFLOAT)
        self.raw, self.filtered = A(w2, h2, vkr.R16G16B16A16_SFLOAT), A(w2, h2, vkr.R16_SFLOAT)
        self.rays = A(w2, h2, vkr.R16G16B16A16_UNORM)
        self.reflections = A(w2, h2, vkr.R8G8B8A8_UNORM)
        self.fused, self._sync_stream = fused, None
        if fused is None:
            self.accumulated_ao, self.accumulated_history = A(w2, h2, vkr.R16G16_SFLOAT), A(w2, h2, vkr.R16G16_SFLOAT)
            self.blurred, self.blurred_history = A(w2, h2, vkr.R8G8B8A8_UNORM), A(w2, h2, vkr.R8G8B8A8_UNORM)
        else:                                                 # epoch 1 (the first frame) writes slot 1, its history is slot 0
            assert fused.end_unit - fused.first == V
            self.accumulated_ao, self.accumulated_history = fused.own("accumulated_ao", 1), fused.own("accumulated_ao", 0)
            self.blurred, self.blurred_history = fused.own("blurred", 1), fused.own("blurred", 0)
        self.halton = torch.from_numpy(vkr.halton23_seq(128)).to(device)
        if luts is None:
            self.pdf = vkr.Image(lut_size, lut_size, vkr.R32_SFLOAT, device=device)
            vkr.pdf_preintegrate(self.pdf, stream=stream)
        else:
            self.pdf = luts
        # one slice of device scratch per (parameter set, pass): a set prepared once can be replayed without host work
        self.scratch = torch.zeros(max(V, 1) * vkr.PARAMS_SCRATCH_STRIDE * 7 * param_sets, dtype=torch.uint8, device=device)
        self.counter, self.frame = 0, 0
        self.scenes = [vkr.synth_scene(v, 64) for v in self.view_ids]

    def _scratch(self, i):
        n = self.V * vkr.PARAMS_SCRATCH_STRIDE
        return self.scratch[i * n:(i + 1) * n]

    def synth_inputs(self, frame=None):
        """Produce every view's G-buffer for `frame` on the device (the path's INPUT; not part of the timed chain)."""
        f = self.frame if frame is None else frame
        for i, v in enumerate(self.view_ids):
            dp = vkr.synth_camera(v, f, self.w, self.h, True)
            a = (self.albedo.layer(i).ptr, self.normal.layer(i).ptr, self.material.layer(i).ptr, self.velocity.layer(i).ptr, self.depth.layer(i).ptr)
            vkr.check(vkr.lib.vkr_synth_gbuffer(C.byref(self.scenes[i]), C.byref(dp), *a, vkr.stream_ptr(self.stream)))

    def params(self, jitters=None):
        ps = []
        for i, v in enumerate(self.view_ids):
            dp = vkr.synth_camera(v, self.frame, self.w, self.h, True)
            ps.append(view_params(dp, self.frame, self.counter, 0.0 if jitters is None else jitters[i]))
        return {k: vkr.params_array(type(ps[0][k]), [p[k] for p in ps]) for k in ps[0]}

    def passes(self, P, flags=0, param_set=0, aux_stream=None, group_streams=None):
        """Hi-Z, SSSR trace/filter/blur, GTAO main/filter/accumulate for every view: 8 launches in total.
        flags: vkr.EXT_PREPARE_ONLY uploads the per-view parameter blocks of `param_set` without launching,
        vkr.EXT_PARAMS_PREPARED launches on blocks prepared earlier (pure kernel launches: CUDA-graph capturable).
        aux_stream: issue the SSR filter + blur on a second stream beside the three GTAO passes (both branches only need the
        trace's outputs; the same split RenderGraph::submit derives for the chain), joined back before returning.  With a few
        dozen views per launch every kernel is only a handful of waves, and the two branches fill each other's tails.
        group_streams: [(main, aux), ...] — with EXT_PARAMS_PREPARED the views are split into len(group_streams) + 1 contiguous
        layer ranges, each running the whole chain on its own pair of streams (the first range on self.stream / aux_stream):
        views are independent, so the ranges fill each other's launch gaps and tails.  Everything joins self.stream at the end."""
        V, st = self.V, self.stream
        prepared = bool(flags & vkr.EXT_PARAMS_PREPARED) and V > 1
        lanes = [(st, aux_stream)] + (list(group_streams) if (group_streams and prepared) else [])
        G = len(lanes)
        if G > 1 and V < 2 * G:
            lanes, G = lanes[:1], 1
        peers = self.fused.deltas if (self.fused is not None and not (flags & vkr.EXT_PREPARE_ONLY)) else None
        ev_begin = None
        if peers is not None:
            # "every peer is done with the buffers this frame overwrites": only the frame's LAST kernels (blur, accumulate) need it,
            # so the wait runs on a side branch forked here and joined just before them — off the frame's critical path
            if self._sync_stream is None:
                self._sync_stream = torch.cuda.Stream(device=self.device)
            ev_fork = torch.cuda.Event()
            ev_fork.record(st)
            self._sync_stream.wait_event(ev_fork)
            self.fused.begin(self._sync_stream)
            ev_begin = torch.cuda.Event()
            ev_begin.record(self._sync_stream)
        if G > 1:
            ev0 = torch.cuda.Event()
            ev0.record(st)
        for g, (sm, sa) in enumerate(lanes):
            first, count = (V * g) // G, (V * (g + 1)) // G - (V * g) // G
            if g > 0:
                sm.wait_event(ev0)

            def R(a):                                          # the range's view of an array image (plain images are shared)
                return vkr.LayerRange(a, first, count) if (G > 1 and isinstance(a, (vkr.ImageArray, vkr.LayerRange))) else a

            def ext(imgs, slot, peers=None):
                return vkr.make_ext(imgs, layers=count, scratch=self._scratch(param_set * 7 + slot), flags=flags if V > 1 else 0,
                                    first_param=first if G > 1 else 0, peers=peers)

            depth, normal, velocity, dn, dv, material, albedo = (R(x) for x in (self.depth, self.normal, self.velocity, self.dn, self.dv, self.material, self.albedo))
            rays, raw, reflections, blurred, blurred_history = (R(x) for x in (self.rays, self.raw, self.reflections, self.blurred, self.blurred_history))
            prev_depth, filtered, accumulated_ao, accumulated_history = (R(x) for x in (self.prev_depth, self.filtered, self.accumulated_ao, self.accumulated_history))
            vkr.hiz_build_ex(depth, normal, velocity, dn, dv, ext([depth, normal, velocity, dn, dv], 0), sm)
            vkr.sssr_trace_ex(P["trace"], depth, dn, material, self.halton, self.pdf, rays, raw, ext([depth, dn, material, self.pdf, rays, raw], 1), sm)
            fork = sa is not None and not (flags & vkr.EXT_PREPARE_ONLY)
            s2 = sa if fork else sm
            if fork:
                ev = torch.cuda.Event()
                ev.record(sm)
                s2.wait_event(ev)
            vkr.sssr_filter_ex(P["trace"], rays, depth, albedo, normal, material, reflections, ext([rays, depth, albedo, normal, material, reflections], 2), s2)
            if ev_begin is not None:
                s2.wait_event(ev_begin)
            vkr.sssr_blur_ex(P["blur"], depth, normal, reflections, material, blurred_history, dv, prev_depth, blurred,
                             ext([depth, normal, reflections, material, blurred_history, dv, prev_depth, blurred], 3, peers), s2)
            vkr.gtao_main_ex(P["gtao"], depth, normal, material, self.pdf, raw, ext([depth, normal, material, self.pdf, raw], 4), 1, sm)
            vkr.gtao_filter_ex(P["gtao"], depth, raw, filtered, ext([depth, raw, filtered], 5), 1, sm)
            if ev_begin is not None and s2 is not sm:
                sm.wait_event(ev_begin)
            vkr.gtao_accumulate_ex(P["accum"], depth, prev_depth, filtered, accumulated_ao, dv, accumulated_history,
                                   ext([depth, prev_depth, filtered, accumulated_ao, dv, accumulated_history], 6, peers), 1, sm)
            if fork:
                ev2 = torch.cuda.Event()
                ev2.record(s2)
                sm.wait_event(ev2)
            if g > 0:
                evg = torch.cuda.Event()
                evg.record(sm)
                st.wait_event(evg)
        if peers is not None:
            self.fused.end(st)                                 # this frame's results are complete on every rank

    def end_frame(self):
        """The remaps of src/main.cpp:416-420 that concern these passes."""
        self.depth, self.prev_depth = self.prev_depth, self.depth
        self.blurred, self.blurred_history = self.blurred_history, self.blurred
        self.accumulated_ao, self.accumulated_history = self.accumulated_history, self.accumulated_ao
        self.counter = (self.counter + 1) % 16
        self.frame += 1

    def outputs(self):
        """(V, bytes) uint8 tensors of the two per-view results (accumulated AO, blurred reflections) for the gather."""
        def rows(a):
            return a.buf.view(a.layers, a.layer_stride)[:, :a.nbytes]
        return rows(self.accumulated_ao), rows(self.blurred)

    KERNELS_PER_FRAME = 9   # hiz tile + tail, trace, filter, blur, gtao main/filter/accumulate
