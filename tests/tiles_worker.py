"""Worker of tests/test_tiles_gpu.py (one process per rank): a row-tiled frame sequence must reproduce, bit for bit on the
rank's own rows, the whole-frame chain computed locally.  argv: root width height frames [oracle]
With "oracle" the rank's rows are ALSO held to the CPU oracle's free-running chain of the whole frame, with the tolerances of
tests/parity.py (BASELINE.json configs[4]'s shape against the checker, not only against the repo's own single-GPU result)."""
import os
import sys

root, w, h, frames = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
vs_oracle = len(sys.argv) > 5 and sys.argv[5] == "oracle"
sys.path[:0] = [os.path.join(root, "tests"), os.path.join(root, "vk-renderer_b200")]
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import chain  # noqa: E402
import parity  # noqa: E402
import vkr  # noqa: E402
import vkr_tiles  # noqa: E402

dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
dev = rank % torch.cuda.device_count()
torch.cuda.set_device(dev)


def allgather_bytes(b):
    out = [None] * world
    dist.all_gather_object(out, b)
    return out


scene = vkr.synth_scene(4, 64)
tf = vkr_tiles.RowTiledFrame(w, h, rank, world, allgather_bytes, device=f"cuda:{dev}", motion_rows=8, lut_size=64)
ref = chain.Chain(w, h, chain.CudaBackend(), lut_size=64)
orc_chain = chain.Chain(w, h, chain.OracleBackend(), lut_size=64) if vs_oracle else None
bad = []
for f in range(frames):
    dp = vkr.synth_camera(4, f, w, h, True)
    vkr.synth_gbuffer(scene, dp, tf.gb)
    out = tf.render_frame(dp, chain.angle_jitter(f))
    ro, _ = ref.render(scene, view_seed=4)
    oo = orc_chain.render(scene, view_seed=4)[0] if vs_oracle else None
    torch.cuda.synchronize()
    y0, y1 = tf.own_rows(True)
    a, b = tf.own_rows(False)
    # after render_frame the ping-pong pairs are swapped: the frame's outputs sit in the *_history / prev slots
    pairs = [("taa", out, ro["taa"], y0, y1), ("color", tf.color_out, ro["color"], y0, y1), ("accumulated_ao", tf.accumulated_history, ro["accumulated_ao"], a, b),
             ("blurred", tf.blurred_history, ro["blurred"], a, b), ("reflections", tf.reflections, ro["reflections"], a, b), ("rays", tf.rays, ro["rays"], a, b),
             ("raw", tf.raw, ro["raw"], a, b)]
    for name, got, want, r0, r1 in pairs:
        g, r = got.download()[r0:r1], want.download()[r0:r1]
        if not np.array_equal(g.view(np.uint8), r.view(np.uint8)):
            bad.append((f, name, int((g.view(np.uint8) != r.view(np.uint8)).sum())))
        if vs_oracle and r1 > r0:
            m = parity.compare(name, g, oo[name].download()[r0:r1])
            if not parity.TOLERANCES[name](m):
                bad.append((f, name, "vs oracle", m))
            elif rank == 0 and f == frames - 1:
                print("rank 0 rows", (r0, r1), name, {k: (round(v, 6) if isinstance(v, float) else v) for k, v in m.items()})
tf.finish()
fails = tf.wait_failures()
tf.close(dist.barrier)
dist.destroy_process_group()
assert not bad and fails == 0, (rank, bad, fails)
print("rank", rank, "rows", tf.own_rows(True), "ok")
