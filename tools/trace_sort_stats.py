"""Would permuting rays inside a CTA raise the SSR march's lane utilisation?   (needs a -DVKR_TRACE_STATS build of ssr_trace.cu:
   tools/variant.sh stats ssr_trace.cu "-DVKR_TRACE_STATS";  VKR_B200_LIB=.../libvkr_b200_stats.so python tools/trace_sort_stats.py W H)

The stats build stores every ray's step count in raw.z.  For CTA tiles of 16x8 (the kernel's) and 16x16 rays this prints
  mean steps / mean over warps of the warp's maximum  (= lane utilisation of the march)
for the kernel's layout (a warp = 16x2 pixels), for the UPPER BOUND of any permutation inside the tile (sorted by the true count),
and for predictors available BEFORE the march: the pixel's Halton sample (U1, U2 from the bit-defined hash), its roughness, both."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "tests"), os.path.join(ROOT, "vk-renderer_b200")]
import numpy as np
import torch

import chain
import vkr

w, h = int(sys.argv[1]), int(sys.argv[2])
scene = vkr.synth_scene(0, 64)
cc = chain.Chain(w, h, chain.CudaBackend(), lut_size=64)
hal = vkr.halton23_seq(128)
f32 = np.float32


def halton_index(w2, h2, frame_random):
    """trace.comp:61-63 with the bit-defined hash of ssr_trace.cu rand_hash."""
    su = ((np.arange(w2, dtype=f32) + f32(0.5)) / f32(w2))[None, :].repeat(h2, 0)
    sv = ((np.arange(h2, dtype=f32) + f32(0.5)) / f32(h2))[:, None].repeat(w2, 1)
    d = (su * f32(12.9898)).astype(f32) + (sv * f32(78.233)).astype(f32)
    s = np.sin(d.astype(np.float64)).astype(f32)
    p = (s * f32(43758.5453)).astype(f32)
    fr = (p - np.floor(p)).astype(f32)
    return ((fr * f32(128.0)).astype(np.uint32) + np.uint32(frame_random)) & np.uint32(127)


def util(it, key, th, tw=16):
    """Tiles of tw x th rays; inside a tile rays are ordered by `key` (None = the kernel's row-major order) and cut into warps of 32."""
    h2, w2 = it.shape
    hh, ww = h2 // th * th, w2 // tw * tw
    t = it[:hh, :ww].reshape(hh // th, th, ww // tw, tw).transpose(0, 2, 1, 3).reshape(-1, th * tw)
    if key is not None:
        k = key[:hh, :ww].reshape(hh // th, th, ww // tw, tw).transpose(0, 2, 1, 3).reshape(-1, th * tw)
        order = np.argsort(k, axis=1, kind="stable")
        t = np.take_along_axis(t, order, axis=1)
    wmax = t.reshape(t.shape[0], -1, 32).max(axis=2)
    return float(t.mean() / wmax.mean()), float(wmax.mean())


def compaction(it, th, points, tw=16):
    """Warp-steps per ray if the tile's surviving rays are re-packed into full warps (in thread order) after the step counts in
    `points` (CTA-wide, through shared memory): segment k runs from points[k-1] to points[k]; a warp's cost in a segment is the
    largest remaining count of its lanes, clipped to the segment."""
    h2, w2 = it.shape
    hh, ww = h2 // th * th, w2 // tw * tw
    t = it[:hh, :ww].reshape(hh // th, th, ww // tw, tw).transpose(0, 2, 1, 3).reshape(-1, th * tw)
    total = 0.0
    edges = [0] + list(points) + [10 ** 6]
    for a, b in zip(edges[:-1], edges[1:]):
        live = t > a                                                   # rays that still march after step a
        rem = np.clip(t - a, 0, b - a) * live
        if a == 0:
            packed = rem                                               # the kernel's layout
        else:                                                          # survivors first, in thread order (a stable partition)
            order = np.argsort(~live, axis=1, kind="stable")
            packed = np.take_along_axis(rem, order, axis=1)
        total += packed.reshape(packed.shape[0], -1, 32).max(axis=2).sum()
    return total / t.size


for f in range(3):
    P = cc.begin_frame(scene)
    st = cc.steps(P)
    st[0][1](); st[1][1](); torch.cuda.synchronize()
    raw = cc.raw.download().astype(np.float32)
    it = raw[..., 2]
    h2, w2 = it.shape
    idx = halton_index(w2, h2, P["trace"].frame_random)
    u1, u2 = hal[idx, 0], hal[idx, 1]
    mat = cc.gb.material.download()
    rough = mat[::2, ::2, 1][:h2, :w2].astype(np.float32)          # nearest full-res texel: a proxy for the bilinear sample
    print(f"frame {f}: mean steps {it.mean():.2f}")
    for th in (8, 16, 32):
        base = util(it, None, th)
        rows = [("kernel layout", base), ("sorted by true count (bound)", util(it, it, th)), ("sorted by U1", util(it, u1, th)), ("sorted by U2", util(it, u2, th)),
                ("sorted by Halton index", util(it, idx.astype(np.float32), th)), ("sorted by roughness", util(it, rough, th)),
                ("sorted by roughness, then U1", util(it, rough * 4 + u1, th))]
        if f > 0:
            rows.append(("sorted by last frame's count", util(it, prev[:h2, :w2], th)))
        print(f"  tile 16x{th}: " + "; ".join(f"{n}: {u:.3f} (warp max {m:.1f})" for n, (u, m) in rows))
    for th in (8, 16):
        base = compaction(it, th, [])
        res = [(pts, compaction(it, th, pts)) for pts in ([15], [16], [20], [24], [16, 24], [16, 32], [16, 24, 40], [20, 40])]
        print(f"  tile 16x{th}: warp-steps per ray, no compaction {base / 32 * 32:.2f} (x 1/32 lanes = {base:.3f}); " +
              "; ".join(f"after {pts}: {v / base:.3f} of it" for pts, v in res))
    prev = it.copy()
    for _, t, _ in st[2:]:
        t()
    cc.end_frame()
