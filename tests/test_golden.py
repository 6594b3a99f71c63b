"""Golden fixtures (tests/golden/chain_96x64.npz, written by tools/make_golden.py from the CPU oracle).

Regression pins, not upstream outputs: the reference ships no vectors and cannot run here.
  * CPU: the oracle still reproduces them (guards the checker against silent edits);
  * GPU: the CUDA chain, fed the fixture's input bytes through the C ABI, lands within the stated tolerances.
"""
import os

import numpy as np
import pytest

import chain
import parity
import vkr

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "chain_96x64.npz")
INPUTS = ("albedo", "normal", "material", "velocity", "depth")
OUTPUTS = ("dn", "dv", "rays", "raw", "reflections", "blurred", "filtered", "accumulated_ao", "color", "taa")


def _load():
    g = np.load(GOLD)
    w, h, frames, view, objects, lut = [int(v) for v in g["meta"]]
    return g, w, h, frames, view, objects, lut


def _check(g, f, images, failures):
    for k in OUTPUTS:
        m = parity.compare(k, images[k].download(), g[f"f{f}_{k}"])
        if not parity.TOLERANCES[k](m):
            failures.append((f, k, m))
    for lod in range(1, images["depth"].mips):
        if not np.array_equal(images["depth"].download(lod), g[f"f{f}_depth_mip{lod}"]):
            failures.append((f, f"depth mip {lod}", "bits differ"))


def _host_gbuffer(g, f, w, h):
    gb = vkr.Gbuffer(w, h, device="cpu")
    for img, k in zip(gb.input_images(), INPUTS):
        img.upload(g[f"f{f}_in_{k}"])
    return gb


def test_oracle_reproduces_golden():
    g, w, h, frames, view, objects, lut = _load()
    scene = vkr.synth_scene(view, objects)
    oc = chain.Chain(w, h, chain.OracleBackend(), lut_size=lut)
    np.testing.assert_array_equal(oc.pdf.download().view(np.uint32), g["pdf"].view(np.uint32))
    failures = []
    for f in range(frames):
        out, _ = oc.render(scene, view_seed=view)
        for img, k in zip([out[k] for k in INPUTS], INPUTS):      # the host twin of the synthetic producer is pinned too
            np.testing.assert_array_equal(img.download(0).view(np.uint8), g[f"f{f}_in_{k}"].view(np.uint8), err_msg=k)
        _check(g, f, out, failures)
    assert not failures, failures


@pytest.mark.gpu
def test_cuda_chain_matches_golden(cuda):
    import torch
    g, w, h, frames, view, objects, lut = _load()
    scene = vkr.synth_scene(view, objects)
    cc = chain.Chain(w, h, chain.CudaBackend(), lut_size=lut)
    torch.cuda.synchronize()
    np.testing.assert_array_equal(cc.pdf.download().view(np.uint32), g["pdf"].view(np.uint32))
    failures = []
    for f in range(frames):
        out, _ = cc.render(scene, view_seed=view, gbuffer_from=_host_gbuffer(g, f, w, h))
        torch.cuda.synchronize()
        _check(g, f, out, failures)
    assert not failures, failures
