"""Iteration statistics of the SSR march (needs a build with EXTRA=-DVKR_TRACE_STATS): python tools/trace_stats.py W H"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "tests"), os.path.join(ROOT, "vk-renderer_b200")]
import numpy as np, torch, chain, vkr
w, h = int(sys.argv[1]), int(sys.argv[2])
scene = vkr.synth_scene(0, 64)
cc = chain.Chain(w, h, chain.CudaBackend(), lut_size=64)
for f in range(3):
    P = cc.begin_frame(scene)
    st = cc.steps(P)
    st[0][1](); st[1][1](); torch.cuda.synchronize()
    it = cc.raw.download()[..., 2].astype(np.float32)
    h2, w2 = it.shape
    hh, ww = h2 // 2 * 2, w2 // 16 * 16
    blk = it[:hh, :ww].reshape(hh // 2, 2, ww // 16, 16)
    wmax = blk.max(axis=(1, 3))
    b8 = it[:h2 // 8 * 8, :w2 // 4 * 4].reshape(h2 // 8, 8, w2 // 4, 4).max(axis=(1, 3))
    b84 = it[:h2 // 4 * 4, :w2 // 8 * 8].reshape(h2 // 4, 4, w2 // 8, 8).max(axis=(1, 3))
    print(f"frame {f}: mean iters {it.mean():.2f}  mean of warp(16x2) max {wmax.mean():.2f}  (8x4: {b84.mean():.2f}, 4x8: {b8.mean():.2f})  utilisation {it.mean()/wmax.mean():.3f}  hist",
          np.histogram(it, bins=[0, 1, 5, 10, 16, 20, 30, 40, 60, 79, 81])[0] / it.size)
    for _, t, _ in st[2:]:
        t()
    cc.end_frame()
