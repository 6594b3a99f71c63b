"""Worker of tests/test_views_gpu.py::test_fused_gather_equals_local_batch (one process per rank; on a single-GPU box the ranks share
cuda:0 — the IPC mapping, the peer stores and the epoch kernels are exercised all the same).
Every rank renders ITS shard of the views with the gather fused into the blur / accumulate kernels (vkr_gather.FusedGather) and, as
the reference, ALL views with a plain local ViewBatch.  After every frame the full arrays gathered on this rank must equal the
reference's results bit for bit, for every view.  The last frames replay a captured CUDA graph (the epoch comes from device memory).
argv: root n_views frames"""
import os
import sys

root, n_views, frames = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
sys.path[:0] = [os.path.join(root, "tests"), os.path.join(root, "vk-renderer_b200")]
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import vkr  # noqa: E402
import vkr_gather  # noqa: E402
import vkr_views  # noqa: E402

dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
dev = rank % torch.cuda.device_count()
torch.cuda.set_device(dev)
w = h = 64


def allgather_bytes(b):
    out = [None] * world
    dist.all_gather_object(out, b)
    return out


st = torch.cuda.Stream()
pdf = vkr.Image(64, 64, vkr.R32_SFLOAT)
vkr.pdf_preintegrate(pdf)
torch.cuda.synchronize()
fg = vkr_gather.FusedGather(n_views, {"accumulated_ao": (w // 2, h // 2, vkr.R16G16_SFLOAT), "blurred": (w // 2, h // 2, vkr.R8G8B8A8_UNORM)},
                            rank, world, allgather_bytes, f"cuda:{dev}")
b, e = vkr_views.shard_range(n_views, world, rank)
mine = vkr_views.ViewBatch(w, h, list(range(b, e)), luts=pdf, stream=st, fused=fg)
ref = vkr_views.ViewBatch(w, h, list(range(n_views)), luts=pdf, stream=st)
bad, graph = [], None
with torch.cuda.stream(st):
    for f in range(frames):
        for vb in (mine, ref):
            vb.synth_inputs()
        P, Pr = mine.params(), ref.params()
        if f < frames - 2 or e - b < 2:
            mine.passes(P)
        else:                                             # prepared parameter blocks + graph capture, then replay (as bench.py does)
            mine.passes(P, flags=vkr.EXT_PREPARE_ONLY)
            st.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=st):
                mine.passes(P, flags=vkr.EXT_PARAMS_PREPARED)
            g.replay()
        ref.passes(Pr)
        got = fg.gathered(fg.epoch, st)
        st.synchronize()
        want_ao, want_bl = ref.outputs()
        for name, want in (("accumulated_ao", want_ao), ("blurred", want_bl)):
            if not torch.equal(got[name], want):
                rows = (got[name] != want).any(dim=1).nonzero().flatten().tolist()
                bad.append((f, name, rows[:8]))
        dist.barrier()                                    # everybody has consumed epoch f + 1 before anybody finishes f + 2
        mine.end_frame(); ref.end_frame()
fails = fg.wait_failures()
mine.fused = None
fg.close(dist.barrier)
dist.destroy_process_group()
assert not bad and fails == 0, (rank, bad, fails)
print("rank", rank, "views", (b, e), "epochs", frames, "ok")
