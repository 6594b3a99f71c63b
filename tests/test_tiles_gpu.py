"""Row-tiled frames with peer-memory halo exchange (vkr_tiles.RowTiledFrame, include/vkr_peer.h): every rank's own rows
equal the whole-frame chain bit for bit.  Runs one process per rank; on a single-GPU box the ranks share cuda:0 (the IPC
mapping, pushes and flags are exercised all the same), on a multi-GPU box they spread over the devices (NVLink P2P)."""
import os
import subprocess
import sys

import pytest

import vkr_tiles

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_tile_rows_partition():
    import vkr
    for h2 in (2160, 1080, 360, 96, 40):
        for world in (1, 2, 3, 4, 8):
            spans = [vkr_tiles.tile_rows(h2, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == h2
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert all(a % vkr.ROW_ALIGN == 0 or a == h2 for a, _ in spans)      # a rank may be empty when strips < world
    assert vkr_tiles.halo_ranges(0, 64, 11, 128) == (None, (53, 64))
    assert vkr_tiles.halo_ranges(64, 128, 11, 128) == ((64, 75), None)


@pytest.mark.gpu
# the last case also compares every rank's rows with the CPU oracle's whole frame at 1920x1080 (tests/parity.py tolerances)
@pytest.mark.parametrize("world,w,h,frames,mode", [(2, 512, 384, 3, ""), (3, 384, 320, 2, ""), (2, 1920, 1080, 2, "oracle")])
def test_row_tiled_frames_are_bit_identical(cuda, world, w, h, frames, mode):
    port = 29900 + (os.getpid() + world + w) % 500
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "tiles_worker.py"), ROOT, str(w), str(h), str(frames)] + ([mode] if mode else [])
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("ok") == world
