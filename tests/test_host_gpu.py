"""The C++ host (rendergraph shim + pass classes + frame loop, libvkr_host.so) against the oracle chain.

This is the drop-in boundary exercised end to end: host-memory G-buffers go in through
vkr_app_upload_gbuffer, the reference's pass classes enqueue their tasks through RenderGraph::add_task,
and the TAA target comes back to the host.
"""
import numpy as np
import pytest

import chain
import parity
import vkr

pytestmark = pytest.mark.gpu

REFERENCE_TASKS = ["DownsampleGbuffer", "DownsampleDepth", "SSSR_trace", "SSSR_filter", "SSSR_blur", "GTAO_main", "GTAO_filter",
                   "GTAO_accumulate", "DeferedShading", "TAA"]   # src/main.cpp:347-391 (one DownsampleDepth task stands for the L-2 of the reference)


@pytest.mark.parametrize("w,h,fused", [(640, 360, True), (256, 256, False)])
def test_host_app_matches_oracle_chain(cuda, w, h, fused):
    import vkr_host
    frames = 4
    scene = vkr.synth_scene(2, 64)
    app = vkr_host.App(w, h)
    app.set_option("fused_hiz", 1.0 if fused else 0.0)
    oc = chain.Chain(w, h, chain.OracleBackend())       # full 1024^2 LUTs, as the app builds
    failures = []
    for f in range(frames):
        dp = vkr.synth_camera(0, f, w, h, True)
        P = oc.begin_frame(scene)                       # fills oc.gb on the CPU with the host twin of the producer
        app.upload_gbuffer(oc.gb)
        app.render_frame(dp, chain.angle_jitter(f))
        oc.passes(P)
        got_tasks = app.last_tasks()
        if fused:
            assert got_tasks == ["DownsampleGbuffer+DownsampleDepth"] + REFERENCE_TASKS[2:]
        else:
            assert got_tasks == REFERENCE_TASKS
        ref = oc.images()
        for name in ("depth", "dn", "dv", "rays", "raw", "filtered", "accumulated_ao", "reflections", "blurred", "color", "taa"):
            got = app.readback(name)
            m = parity.compare_image(name, got, ref[name])
            ok = parity.TOLERANCES[name](m)
            print(f, name, m)
            if not ok:
                failures.append((f, name, m))
        oc.end_frame()
    app.close()
    assert not failures, failures


def test_host_rejects_bad_sizes(cuda):
    import vkr_host
    with pytest.raises(vkr_host.HostError):
        vkr_host.App(8, 8)


def test_device_producer_equals_host_twin(cuda):
    """The CUDA ray-caster and its host twin write identical G-buffers (unfused fp32 on both sides)."""
    import torch
    w, h = 640, 360
    scene = vkr.synth_scene(5, 64)
    for frame in (0, 7):
        dp = vkr.synth_camera(3, frame, w, h, True)
        a, b = vkr.Gbuffer(w, h, device="cpu"), vkr.Gbuffer(w, h, device="cuda")
        vkr.synth_gbuffer(scene, dp, a)
        vkr.synth_gbuffer(scene, dp, b)
        torch.cuda.synchronize()
        for x, y, name in zip(a.input_images(), b.input_images(), ("albedo", "normal", "material", "velocity", "depth")):
            ga, gb_ = x.download(), y.download()
            if ga.dtype == np.float16:
                ga, gb_ = ga.view(np.uint16), gb_.view(np.uint16)
            mism = float((ga != gb_).mean())
            assert mism <= 1e-5, (name, mism)      # a handful of silhouette pixels may flip on libm-free but FMA-free code paths
