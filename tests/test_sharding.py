"""Host-side multi-GPU logic on CPU: unit sharding and the final gather over gloo with world_size 2 (and 3, ragged)."""
import os
import subprocess
import sys
import textwrap

import pytest

import vkr_views

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_ranges_partition_the_units():
    for n in (0, 1, 5, 512, 513):
        for world in (1, 2, 3, 4, 8):
            spans = [vkr_views.shard_range(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [e - b for b, e in spans]
            assert max(sizes) - min(sizes) <= 1 and sizes == vkr_views.shard_sizes(n, world)


WORKER = textwrap.dedent("""
    import os, sys
    sys.path[:0] = [os.path.join(sys.argv[1], "vk-renderer_b200")]
    import torch, torch.distributed as dist
    import vkr_views
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    n = int(sys.argv[2])
    full = (torch.arange(n * 6, dtype=torch.int32).reshape(n, 6) * 7 + 3).to(torch.uint8)
    b, e = vkr_views.shard_range(n, world, rank)
    got = vkr_views.gather_blocks(full[b:e].clone(), n)
    assert got.shape == full.shape and torch.equal(got, full), (rank, got.shape)
    dist.barrier()
    dist.destroy_process_group()
    print("rank", rank, "ok")
""")


@pytest.mark.parametrize("world,n", [(2, 512), (2, 5), (3, 7)])
def test_gather_blocks_over_gloo(tmp_path, world, n):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    port = 29500 + (os.getpid() + world * 13 + n) % 2000
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), str(script), ROOT, str(n)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.count("ok") == world
