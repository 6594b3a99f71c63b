"""vkr_sssr_trace_tiles_static — the allocation-free entry point (include/vkr_passes.h; no counterpart in the reference)
— against the entries it bundles, through the C ABI.

The static entry keeps the tile lists, the indirect commands and the sRGB table in a file-scope device array of the library
(csrc/ssr_tiles.cu) that every call shares and that keeps the last call's contents.  So the calls are issued IN A ROW on one
stream without a synchronisation between them, at different extents (the second, smaller one leaves stale list entries of the
first beyond its counters; the third returns to the first extent with another frame), and each result must equal what
vkr_sssr_clear_indirect + vkr_sssr_classification + 2 x vkr_sssr_trace_indirect write with caller-allocated buffers: within the
"rays" tolerance of tests/parity.py — and, the kernels being the same and every pixel belonging to exactly one tile, bit for bit.

This file sorts last on purpose: it must pass alone in a fresh process and after the other tests.
"""
import numpy as np
import pytest

import chain
import parity
import vkr

pytestmark = pytest.mark.gpu

SENTINEL = 4242


def _inputs(w, h, seed, frame):
    """Device G-buffer of one synthetic frame with its Hi-Z pyramid + the frame's trace parameters."""
    import torch
    gb = vkr.Gbuffer(w, h)
    dp = vkr.synth_camera(seed, frame, w, h, True)
    vkr.synth_gbuffer(vkr.synth_scene(seed, 64), dp, gb)
    vkr.hiz_build(gb.depth, gb.normal, gb.velocity_vectors, gb.downsampled_normals, gb.downsampled_velocity_vectors)
    torch.cuda.synchronize()
    return gb, chain.frame_params(dp, frame, frame % 16)["trace"]


def _rays(w2, h2):
    return vkr.Image(w2, h2, vkr.R16G16B16A16_UNORM).upload(np.full((h2, w2, 4), SENTINEL, dtype=np.uint16))


def _separate_entries(P, glossy, gb, hal_dev, rays):
    """The existing path: caller-allocated lists and indirect commands."""
    import torch
    w2, h2 = rays.width, rays.height
    n = ((w2 + 7) // 8) * ((h2 + 7) // 8)
    rt = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    gt = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    ri = torch.full((3,), 77, dtype=torch.int32, device="cuda")
    gi = torch.full((3,), 77, dtype=torch.int32, device="cuda")
    vkr.sssr_clear_indirect(ri, gi)
    vkr.sssr_classification(vkr.ClassifyParams(w2, h2, P.max_roughness, glossy), gb.material, rt, gt, ri, gi)
    vkr.sssr_trace_indirect(P, 0, gb.depth, gb.downsampled_normals, gb.material, hal_dev, rt, ri, rays)
    vkr.sssr_trace_indirect(P, 1, gb.depth, gb.downsampled_normals, gb.material, hal_dev, gt, gi, rays)
    torch.cuda.synchronize()
    return int(ri[0]), int(gi[0])


def test_static_scratch_entry_equals_separate_entries_on_calls_in_a_row(cuda):
    import torch
    hal_dev = torch.from_numpy(vkr.halton23_seq(128)).cuda()
    cases = [(1280, 720, 1, 0, 0.5), (328, 200, 2, 3, 0.3), (1280, 720, 1, 5, 0.5)]      # w, h, scene seed, frame, glossy threshold
    frames = [_inputs(w, h, seed, f) for w, h, seed, f, _ in cases]

    # set-up call (symbol address + sRGB table) on a throw-away image, then the calls under test back to back on the current stream
    vkr.sssr_trace_tiles_static(frames[1][1], 0.5, frames[1][0].depth, frames[1][0].downsampled_normals, frames[1][0].material, hal_dev, _rays(164, 100))
    torch.cuda.synchronize()
    got = [_rays(w // 2, h // 2) for w, h, *_ in cases]
    free_before = torch.cuda.mem_get_info()[0]
    for (w, h, _, _, glossy), (gb, P), rays in zip(cases, frames, got):
        vkr.sssr_trace_tiles_static(P, glossy, gb.depth, gb.downsampled_normals, gb.material, hal_dev, rays)
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    assert free_before - free_after < (1 << 20), f"device memory shrank by {free_before - free_after} bytes across the allocation-free calls"

    for i, ((w, h, _, _, glossy), (gb, P), rays) in enumerate(zip(cases, frames, got)):
        want = _rays(w // 2, h // 2)
        nr, ng = _separate_entries(P, glossy, gb, hal_dev, want)
        g, r = rays.download(), want.download()
        assert not (r == SENTINEL).all(axis=-1).any() and not (g == SENTINEL).all(axis=-1).any(), "every pixel belongs to a listed tile"
        m = parity.compare("rays", g, r)
        mism = int((g != r).any(axis=-1).sum())
        print(f"call {i}: {w // 2}x{h // 2} rays, {nr} mirror + {ng} glossy tiles, {m}, pixels differing in any bit: {mism}")
        assert nr > 0 and ng > 0 and m["valid_frac"] > 0.01, "the frame should have both tile classes and some hits"
        assert parity.TOLERANCES["rays"](m), (i, m)
        assert mism == 0, (i, mism)                    # same kernels, same inputs, one tile per pixel: identical


def test_static_scratch_entry_rejects_extents_beyond_its_maximum(cuda):
    gb, P = _inputs(64, 64, 0, 0)
    import torch
    hal_dev = torch.from_numpy(vkr.halton23_seq(128)).cuda()
    too_big = vkr.Image(2056, 2048, vkr.R16G16B16A16_UNORM)              # 257 x 256 = 65 792 tiles > 65 536
    with pytest.raises(vkr.VkrError, match=str(vkr.SSSR_STATIC_MAX_TILES)):
        vkr.sssr_trace_tiles_static(P, 0.5, gb.depth, gb.downsampled_normals, gb.material, hal_dev, too_big)
