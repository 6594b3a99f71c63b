"""Worker of tests/test_views_gpu.py::test_peer_gather (one process per rank): PeerGather must deliver, on every rank and for
every epoch, exactly what an all_gather of the ranks' blocks delivers.  argv: root n_units epochs"""
import os
import sys

root, n_units, epochs = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
sys.path[:0] = [os.path.join(root, "tests"), os.path.join(root, "vk-renderer_b200")]
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import vkr_gather  # noqa: E402
import vkr_views  # noqa: E402

dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
dev = rank % torch.cuda.device_count()
torch.cuda.set_device(dev)


def allgather_bytes(b):
    out = [None] * world
    dist.all_gather_object(out, b)
    return out


row_bytes = [4096, 1000]                       # the second tensor's rows are strided in the source (pitch 1024)
nbuf = int(sys.argv[4]) if len(sys.argv) > 4 else 2
pg = vkr_gather.PeerGather(n_units, row_bytes, rank, world, allgather_bytes, f"cuda:{dev}", buffers=nbuf)
b, e = vkr_views.shard_range(n_units, world, rank)
st = torch.cuda.Stream()
bad = []


def block(epoch, r, t):
    """Deterministic content of rank r's block of tensor t at `epoch` (any rank can rebuild any block)."""
    b0, e0 = vkr_views.shard_range(n_units, world, r)
    g = torch.Generator().manual_seed(1000 * epoch + 10 * r + t)
    return torch.randint(0, 256, (e0 - b0, row_bytes[t]), dtype=torch.uint8, generator=g)


expect, keep = {}, []
with torch.cuda.stream(st):
    for ep in range(1, epochs + 1):
        if ep >= 2:                            # consume epoch ep - 1 on `st`, THEN push ep behind it (the rule of vkr_gather)
            outs = pg.gathered(ep - 1, st)
            got = [o.clone() for o in outs]
            st.synchronize()
            for t in range(2):
                if not torch.equal(got[t].cpu(), expect[ep - 1][t]):
                    bad.append((ep - 1, t))
        src0 = block(ep, rank, 0).cuda()
        wide = torch.zeros((e - b, 1024), dtype=torch.uint8, device="cuda")
        wide[:, :1000] = block(ep, rank, 1).cuda()
        keep.append((src0, wide))              # sources stay untouched until their copies have run (pg.sources_free otherwise)
        got_epoch = pg.push([src0, wide[:, :1000]], st)
        assert got_epoch == ep
        expect[ep] = [torch.cat([block(ep, r, t) for r in range(world)]) for t in range(2)]
    outs = pg.finish()
    pg.xfer.synchronize()
    for t in range(2):
        if not torch.equal(outs[t].cpu(), expect[epochs][t]):
            bad.append((epochs, t))
fails = pg.wait_failures()
pg.close(dist.barrier)
dist.destroy_process_group()
assert not bad and fails == 0, (rank, bad, fails)
print("rank", rank, "units", (b, e), "ok")
