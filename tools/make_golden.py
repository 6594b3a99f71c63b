"""Writes tests/golden/chain_96x64.npz: inputs and every output image of 3 frames of the full chain, produced by the CPU
oracle (oracle/liboracle.so) on the synthetic G-buffer's host twin.

The reference has no golden vectors and cannot be run here, so these are REGRESSION pins of the oracle
(itself pinned by tests/test_oracle_anchors.py), not upstream outputs.  Usage: python tools/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "tests"), os.path.join(ROOT, "vk-renderer_b200")]
import chain  # noqa: E402
import vkr  # noqa: E402

W, H, FRAMES, VIEW, OBJECTS, LUT = 96, 64, 3, 2, 12, 64
INPUTS = ("albedo", "normal", "material", "velocity", "depth")
OUTPUTS = ("dn", "dv", "rays", "raw", "reflections", "blurred", "filtered", "accumulated_ao", "color", "taa")


def main():
    scene = vkr.synth_scene(VIEW, OBJECTS)
    oc = chain.Chain(W, H, chain.OracleBackend(), lut_size=LUT)
    out = {"meta": np.array([W, H, FRAMES, VIEW, OBJECTS, LUT], dtype=np.int32)}
    out["pdf"], out["brdf"] = oc.pdf.download(), oc.brdf.download()
    for f in range(FRAMES):
        o, _ = oc.render(scene, view_seed=VIEW)
        for k in INPUTS:
            out[f"f{f}_in_{k}"] = o[k].download(0)
        for k in OUTPUTS:
            out[f"f{f}_{k}"] = o[k].download()
        for lod in range(1, o["depth"].mips):
            out[f"f{f}_depth_mip{lod}"] = o["depth"].download(lod)
    path = os.path.join(ROOT, "tests", "golden", "chain_96x64.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
