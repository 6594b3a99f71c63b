"""Array-layer (batched views) and row-window launches through the `_ex` C-ABI entries.

  * a batch of independent views in ONE launch per pass equals the per-view CPU oracle (same tolerances as the
    single-view parity tests) — BASELINE configs[3] shape;
  * a pass computed in two row windows writes exactly the bytes of the whole-frame pass — BASELINE configs[4] shape.
"""
import numpy as np
import pytest

import chain
import parity
import vkr
import vkr_views

pytestmark = pytest.mark.gpu


# views 64..127 at 256x256: rank 1's shard of BASELINE.json configs[3] (512 probe views over 8 GPUs) in one batch
@pytest.mark.parametrize("w,h,views,frames", [(128, 64, [3, 11, 12, 40, 7], 3), (256, 256, [0, 5], 2), (256, 256, list(range(64, 128)), 2)])
def test_batched_views_match_oracle(cuda, w, h, views, frames):
    import torch
    ob = chain.OracleBackend()
    pdf_ref = vkr.Image(256, 256, vkr.R32_SFLOAT, device="cpu")
    ob.pdf_lut(pdf_ref)
    brdf_ref = vkr.Image(16, 16, vkr.R16G16_SFLOAT, device="cpu")      # lighting is not part of the view batch
    oracles = [chain.Chain(w, h, chain.OracleBackend(), luts=(pdf_ref, brdf_ref)) for _ in views]
    vb = vkr_views.ViewBatch(w, h, views, luts=pdf_ref.clone_to("cuda"))
    failures = []
    for f in range(frames):
        refs = [oc.render(vkr.synth_scene(v, 64), view_seed=v)[0] for oc, v in zip(oracles, views)]
        vb.synth_inputs()
        vb.passes(vb.params(jitters=[chain.angle_jitter(f)] * len(views)))
        torch.cuda.synchronize()
        got = dict(depth=vb.depth, dn=vb.dn, dv=vb.dv, rays=vb.rays, raw=vb.raw, reflections=vb.reflections, blurred=vb.blurred,
                   filtered=vb.filtered, accumulated_ao=vb.accumulated_ao)
        for i, v in enumerate(views):
            for name, arr in got.items():
                m = parity.compare_image(name, arr.layer(i), refs[i][name])
                if not parity.TOLERANCES[name](m):
                    failures.append((f, v, name, m))
        vb.end_frame()
    assert not failures, failures


def _split(hh):
    mid = (hh // 2) // vkr.ROW_ALIGN * vkr.ROW_ALIGN
    return [(0, mid), (mid, hh)]


def test_row_windows_equal_whole_pass(cuda):
    import torch
    w, h = 256, 160
    scene = vkr.synth_scene(2, 64)
    cc = chain.Chain(w, h, chain.CudaBackend(), lut_size=64)
    cc.render(scene)                                   # frame 0 fills the histories
    P = cc.begin_frame(scene)
    gb = cc.gb
    hd = cc.b._hd(cc.halton)
    E = lambda imgs, rows: vkr.make_ext(imgs, rows=rows)  # noqa: E731
    # (name, whole-pass thunk, windowed call taking rows, outputs, output height)
    passes = [
        ("SSSR_trace", lambda r: vkr.sssr_trace_ex(P["trace"], gb.depth, gb.downsampled_normals, gb.material, hd, cc.pdf, cc.rays, cc.raw, E([], r)), [cc.rays, cc.raw]),
        ("SSSR_filter", lambda r: vkr.sssr_filter_ex(P["trace"], cc.rays, gb.depth, gb.albedo, gb.normal, gb.material, cc.reflections, E([], r)), [cc.reflections]),
        ("SSSR_blur", lambda r: vkr.sssr_blur_ex(P["blur"], gb.depth, gb.normal, cc.reflections, gb.material, cc.blurred_history, gb.downsampled_velocity_vectors,
                                                 gb.prev_depth, cc.blurred, E([], r)), [cc.blurred]),
        ("GTAO_main", lambda r: vkr.gtao_main_ex(P["gtao"], gb.depth, gb.normal, gb.material, cc.pdf, cc.raw, E([], r)), [cc.raw]),
        ("GTAO_filter", lambda r: vkr.gtao_filter_ex(P["gtao"], gb.depth, cc.raw, cc.filtered, E([], r)), [cc.filtered]),
        ("GTAO_accumulate", lambda r: vkr.gtao_accumulate_ex(P["accum"], gb.depth, gb.prev_depth, cc.filtered, cc.accumulated_ao, gb.downsampled_velocity_vectors,
                                                             cc.accumulated_history, E([], r)), [cc.accumulated_ao]),
        ("DeferedShading", lambda r: vkr.deferred_shading_ex(P["shade"], gb.albedo, gb.normal, gb.material, gb.depth, cc.accumulated_ao, cc.brdf, cc.blurred,
                                                             cc.color_out, E([], r)), [cc.color_out]),
        ("TAA", lambda r: vkr.taa_resolve_ex(P["taa"], cc.taa_history, gb.prev_depth, gb.depth, gb.velocity_vectors, cc.color_out, cc.taa_target, E([], r)),
         [cc.taa_target]),
    ]
    cc.b.hiz(gb.depth, gb.normal, gb.velocity_vectors, gb.downsampled_normals, gb.downsampled_velocity_vectors)
    for name, call, outs in passes:
        before = [o.buf.clone() for o in outs]
        call(None)                                      # rows = None -> 0,0 = the whole image
        torch.cuda.synchronize()
        whole = [o.buf.clone() for o in outs]
        for o, b in zip(outs, before):
            o.buf.copy_(b)
        for rows in _split(outs[0].height):
            call(rows)
        torch.cuda.synchronize()
        for o, wbuf in zip(outs, whole):
            assert torch.equal(o.buf, wbuf), f"{name}: windowed result differs from the whole pass"


def test_row_window_validation(cuda):
    cc = chain.Chain(128, 64, chain.CudaBackend(), lut_size=16)
    P = chain.frame_params(vkr.synth_camera(0, 0, 128, 64, True), 0, 0)
    with pytest.raises(vkr.VkrError, match="aligned"):
        vkr.gtao_filter_ex(P["gtao"], cc.gb.depth, cc.raw, cc.filtered, vkr.make_ext([], rows=(8, 32)))
    with pytest.raises(vkr.VkrError, match="row windows are not supported"):
        vkr.hiz_build_ex(cc.gb.depth, cc.gb.normal, cc.gb.velocity_vectors, cc.gb.downsampled_normals, cc.gb.downsampled_velocity_vectors,
                         vkr.make_ext([], rows=(0, 16)))


def test_view_ranges_on_separate_streams_write_the_same_bytes(cuda):
    """ViewBatch.passes with prepared parameter blocks split into layer ranges on their own stream pairs (what bench.py's probe
    workload captures into its frame graph) writes exactly the bytes of the single-launch form, two frames in a row."""
    import torch
    w = h = 128
    views = list(range(7))                               # odd count: ranges of 2, 2, 3 views
    pdf = vkr.Image(64, 64, vkr.R32_SFLOAT)
    vkr.pdf_preintegrate(pdf)
    st = torch.cuda.Stream()
    a = vkr_views.ViewBatch(w, h, views, luts=pdf, stream=st)
    b = vkr_views.ViewBatch(w, h, views, luts=pdf, stream=st)
    gs = [(torch.cuda.Stream(), torch.cuda.Stream()), (torch.cuda.Stream(), None)]
    with torch.cuda.stream(st):
        for f in range(2):
            for vb in (a, b):
                vb.synth_inputs()
            P = a.params()
            a.passes(P)
            b.passes(P, flags=vkr.EXT_PREPARE_ONLY)
            b.passes(P, flags=vkr.EXT_PARAMS_PREPARED, aux_stream=gs[1][0], group_streams=gs[:1] + [(gs[0][1], None)])
            st.synchronize()
            for name in ("depth", "dn", "dv", "rays", "raw", "reflections", "blurred", "filtered", "accumulated_ao"):
                assert torch.equal(getattr(a, name).buf, getattr(b, name).buf), (f, name)
            a.end_frame(); b.end_frame()


# the last case: 4 landing buffers per tensor (epochs > buffers, so every buffer is reused)
@pytest.mark.parametrize("world,n_units,epochs,buffers", [(2, 16, 5, 2), (3, 10, 4, 2), (3, 16, 11, 4)])
def test_peer_gather_equals_all_gather(cuda, world, n_units, epochs, buffers):
    """vkr_gather.PeerGather (copy-engine pushes into every rank's cudaIpc slab + stream-side flags) against the blocks an
    all_gather would deliver, several epochs in a row (buffer reuse), ragged shards, strided source rows.  One process per rank;
    on a single-GPU box the ranks share cuda:0 (the IPC mapping, pushes and flags are exercised all the same)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    port = 29300 + (os.getpid() + world) % 500
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(root, "tests", "gather_worker.py"), root, str(n_units), str(epochs), str(buffers)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("ok") == world


@pytest.mark.parametrize("world,n_views,frames", [(2, 5, 4), (3, 7, 5)])
def test_fused_gather_equals_local_batch(cuda, world, n_views, frames):
    """The gather fused into the producing kernels (VkrLaunchExt.peer_delta + the epoch kernels of include/vkr_peer.h,
    vkr_gather.FusedGather): on every rank the full arrays equal, bit for bit and for every view, what a local batch of ALL views
    computes — ragged shards, several frames (ping-pong slots reused), the last two through a captured CUDA graph."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    port = 29400 + (os.getpid() + 7 * world) % 500
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(root, "tests", "fused_worker.py"), root, str(n_views), str(frames)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("ok") == world

