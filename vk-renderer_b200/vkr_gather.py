"""The final gather of sharded per-unit results as peer stores over NVLink (BASELINE.json configs[3]: the batched probe views).

Every rank owns, inside a cudaIpc slab (include/vkr_peer.h), two gather buffers per result tensor — (n_units, row_bytes) each,
used alternately by even and odd epochs — and maps every other rank's slab.  Per epoch (= frame) a rank copies its block of rows
into the same rows of every peer's buffer with the COPY ENGINES (vkr_peer_push_2d: cudaMemcpyAsync on peer-mapped memory, no SM
involved) and then raises its flag in that peer's slab; the gathered epoch is complete on a rank once its world - 1 flags have
reached the epoch.  Nothing synchronises the host and nothing runs on the SMs except one-thread flag kernels, so the gather costs
the passes that run beside it no issue slots — an NCCL all_gather of the same 64 MB per frame takes about 7 % of them
(DESIGN.md §5).  Reuse of a buffer needs no further handshake: a rank pushes epoch e only after it has seen every peer's flag
for e - 1, and a peer raises that flag behind everything its `after_stream` held when it called push(e - 1).  So the one rule for
a consumer is: read epoch e - 2 (gathered(e - 2, stream)) on a stream that push(e - 1) is then given as `after_stream` — consume
the previous epoch, then push the next one.  The SOURCE blocks of an epoch may be overwritten once sources_free(stream) has made
the writer wait for that epoch's copies.
"""
import ctypes as C

import torch

import vkr
from vkr_tiles import _P, PeerSlab, PeerWaitTimeout, peer_wait_failures_total
from vkr_views import shard_range


class PeerGather:
    def __init__(self, n_units, row_bytes, rank, world, allgather_bytes, device, timeout_ms=2000, buffers=2):
        """row_bytes: bytes per unit of each result tensor, e.g. [65536, 65536].  allgather_bytes(b) -> list over ranks (host side,
        any transport: it only carries the 64-byte IPC handles once)."""
        self.n_units, self.row_bytes, self.rank, self.world = n_units, list(row_bytes), rank, world
        self.device, self.timeout = torch.device(device), timeout_ms
        # `buffers` landing buffers per tensor on every rank.  Epoch e lands in buffer e % buffers, whose last tenant was epoch
        # e - buffers; a rank may overwrite it once the owner has pushed epoch e - buffers + 1 (the consumer rule below).  With 2
        # buffers every rank is held within ONE epoch of the slowest rank, and since the compute stream in turn waits for its own
        # copies (sources_free), jitter on any rank stalls all of them every frame; more buffers loosen that coupling at
        # n_units * sum(row_bytes) bytes of slab each.
        assert buffers >= 2
        self.nbuf = buffers
        self._fail_base = peer_wait_failures_total()
        self.first, self.end = shard_range(n_units, world, rank)
        sizes = [n_units * rb for rb in self.row_bytes]
        self.slab = PeerSlab(self.nbuf * sum((s + 255) & ~255 for s in sizes), self.device)
        assert world * 4 <= PeerSlab.FLAG_BYTES
        self.bufs = [[self.slab.carve(s).view(n_units, rb) for s, rb in zip(sizes, self.row_bytes)] for _ in range(self.nbuf)]
        handles = allgather_bytes(self.slab.handle_bytes())
        for r in range(world):
            if r != rank:
                self.slab.open_peer(r, handles[r])
        self.epoch = 0
        self.xfer = torch.cuda.Stream(device=self.device)
        # one stream per peer: copies to different peers run on different copy engines and NVLink paths at the same time
        # (on ONE stream the 2 x (world - 1) copies serialise: 72 GB/s out of a rank at 8 GPUs, slower than the frame)
        self.pstreams = {r: torch.cuda.Stream(device=self.device) for r in range(world) if r != rank}
        self._go = torch.cuda.Event()
        self._pdone = {r: torch.cuda.Event() for r in self.pstreams}
        self._ready = torch.cuda.Event()
        self._sent = [torch.cuda.Event(), torch.cuda.Event()]   # behind the copies of the last even / odd epoch

    def _flag(self, base, src_rank):
        return base + 4 * src_rank

    def _wait_epoch(self, epoch, st):
        for r in range(self.world):
            if r != self.rank:
                vkr.check(_P.vkr_peer_wait(C.c_void_p(self._flag(self.slab.ptr, r)), epoch, self.timeout, st))

    def push(self, blocks, after_stream):
        """blocks: this rank's (units_r, >= row_bytes) uint8 views, one per result tensor (rows may be strided).  Queues, behind
        everything `after_stream` holds now, the copy of the blocks into every rank's buffers of the new epoch.  Returns the epoch."""
        self.epoch += 1
        e, st = self.epoch, C.c_void_p(self.xfer.cuda_stream)
        self._ready.record(after_stream)
        self.xfer.wait_event(self._ready)
        if e > self.nbuf - 1:
            self._wait_epoch(e - self.nbuf + 1, st)   # every peer has finished with epoch e - buffers, the last tenant of this buffer
        units = self.end - self.first
        self._go.record(self.xfer)
        for r in range(self.world):
            if r == self.rank:
                s, sp, base = self.xfer, st, self.slab.ptr
            else:
                s, base = self.pstreams[r], self.slab.peers[r]
                s.wait_event(self._go)
                sp = C.c_void_p(s.cuda_stream)
            for t, (blk, rb) in enumerate(zip(blocks, self.row_bytes)):
                assert blk.shape[0] == units and blk.shape[1] >= rb and blk.stride(1) == 1
                dst = base + (self.bufs[e % self.nbuf][t].data_ptr() - self.slab.ptr) + self.first * rb
                vkr.check(_P.vkr_peer_push_2d(C.c_void_p(blk.data_ptr()), blk.stride(0), C.c_void_p(dst), rb, rb, units, sp))
            if r != self.rank:
                self._pdone[r].record(s)
                vkr.check(_P.vkr_peer_signal(C.c_void_p(self._flag(base, self.rank)), e, sp))
        for r in self.pstreams:
            self.xfer.wait_event(self._pdone[r])
        self._sent[e & 1].record(self.xfer)
        return e

    def sources_free(self, stream, epoch=None):
        """Make `stream` wait until the copies of `epoch` (default: the one before the last pushed — with results that alternate
        between two images, the image the next frame writes) have read their source blocks."""
        e = self.epoch - 1 if epoch is None else epoch
        if e >= 1 and e > self.epoch - 2:
            stream.wait_event(self._sent[e & 1])

    def gathered(self, epoch, stream=None):
        """Holds `stream` (default: the transfer stream) until `epoch` is complete on this rank and returns its (n_units, row_bytes)
        tensors.  Consume them on a stream that is (or is ordered before) the `after_stream` of push(epoch + 1): that push's
        flag is what lets the peers overwrite this buffer with epoch + buffers."""
        st = self.xfer if stream is None else stream
        if stream is not None:                       # own rows arrive through the transfer stream
            ev = torch.cuda.Event()
            ev.record(self.xfer)
            st.wait_event(ev)
        self._wait_epoch(epoch, C.c_void_p(st.cuda_stream))
        return self.bufs[epoch % self.nbuf]

    def finish(self):
        """Wait (stream-side) for the last epoch to be complete everywhere it lands here; returns its tensors."""
        return self.gathered(self.epoch) if self.epoch else self.bufs[0]

    def wait_failures(self):
        """Timed-out peer waits in this process since this gather was created (synchronises the device)."""
        return peer_wait_failures_total() - self._fail_base

    def close(self, barrier=None):
        """Joins, releases the slab and RAISES PeerWaitTimeout if any wait gave up (the epochs read after it were stale)."""
        for s in list(self.pstreams.values()) + [self.xfer]:
            s.synchronize()
        fails = self.wait_failures()
        self.bufs = None
        self.slab.close(barrier)
        if fails:
            raise PeerWaitTimeout(f"{fails} peer wait(s) timed out on rank {self.rank}: gathered blocks were stale")


class FusedGather:
    """The gather fused into the kernels that produce the results (VkrLaunchExt.peer_delta, include/vkr_passes.h): every rank holds
    the FULL result arrays — `names` x 2 (ping-pong) array images of n_units layers — in a cudaIpc-shared slab; a rank's passes
    write its own layers [first, end) and store every texel into the same address of every peer's slab as well (NVLink stores from
    the producing kernel: no collective, no copy engine, no SMs taken from the passes).  Ordering is the epoch protocol of
    include/vkr_peer.h: begin(stream) before the frame's writers, end(stream) after them; both read the epoch from device memory,
    so a frame containing them replays as a CUDA graph.

    Consumer rule: read the gathered arrays of epoch k (gathered(k, stream)) on a stream ordered before this rank's end() of k + 1.
    The arrays of epoch k live in slot k & 1."""

    def __init__(self, n_units, specs, rank, world, allgather_bytes, device, timeout_ms=4000):
        """specs: {name: (width, height, format)} of the per-unit result images."""
        assert world <= 8, "the epoch kernels handle at most 8 ranks"
        self.n_units, self.rank, self.world, self.timeout = n_units, rank, world, timeout_ms
        self.device = torch.device(device)
        self.first, self.end_unit = shard_range(n_units, world, rank)
        sizes = {n: vkr.ImageArray.array_bytes(w, h, f, 1, n_units) for n, (w, h, f) in specs.items()}
        self.slab = PeerSlab(2 * sum(sizes.values()) + 256, self.device)
        assert world * 4 + 4 <= PeerSlab.FLAG_BYTES
        self.slab.tensor.zero_()                       # flags, epoch word and the arrays (histories start at zero) ...
        torch.cuda.synchronize(self.device)            # ... before anybody can learn this slab's handle below
        # flags[r] at slab + 4 r (written by rank r), this rank's epoch word behind them
        self.ctrl = self.slab.ptr + 4 * 8
        self.full = {n: [vkr.ImageArray.over(self.slab.carve(sizes[n]), *specs[n], 1, n_units) for _ in range(2)] for n in specs}
        handles = allgather_bytes(self.slab.handle_bytes())
        for r in range(world):
            if r != rank:
                self.slab.open_peer(r, handles[r])
        self.peer_ranks = [r for r in range(world) if r != rank]
        self.deltas = [self.slab.peers[r] - self.slab.ptr for r in self.peer_ranks]           # same layout on every rank
        self._remote = (C.c_void_p * 8)(*[(self.slab.peers[r] + 4 * rank) if r != rank and r < world else None for r in range(8)])
        self._fail_base = peer_wait_failures_total()
        self.epoch = 0                                                                        # host-side count of end() calls

    def own(self, name, slot):
        """This rank's layers of the full array `name` in ping-pong slot 0 / 1, as the `_ex` entries take them."""
        return vkr.LayerRange(self.full[name][slot], self.first, self.end_unit - self.first)

    def begin(self, stream):
        vkr.check(_P.vkr_peer_epoch_begin(C.c_void_p(self.ctrl), C.c_void_p(self.slab.ptr), self.world, self.rank, self.timeout, C.c_void_p(stream.cuda_stream)))

    def end(self, stream):
        """Call once per frame (also counts the epoch on the host; a CUDA-graph REPLAY of a frame that contains end() must be
        followed by note_replayed())."""
        vkr.check(_P.vkr_peer_epoch_end(C.c_void_p(self.ctrl), self._remote, self.world, C.c_void_p(stream.cuda_stream)))
        self.epoch += 1

    def note_replayed(self):
        self.epoch += 1

    def gathered(self, epoch, stream):
        """Holds `stream` until every rank has finished writing `epoch` here; returns {name: (n_units, layer bytes) uint8 view}."""
        st = C.c_void_p(stream.cuda_stream)
        for r in self.peer_ranks:
            vkr.check(_P.vkr_peer_wait(C.c_void_p(self.slab.ptr + 4 * r), epoch, self.timeout, st))
        out = {}
        for n, pair in self.full.items():
            a = pair[epoch & 1]
            out[n] = a.buf.view(a.layers, a.layer_stride)[:, :a.nbytes]
        return out

    def wait_failures(self):
        return peer_wait_failures_total() - self._fail_base

    def close(self, barrier=None):
        torch.cuda.synchronize(self.device)
        fails = self.wait_failures()
        self.full = None
        self.slab.close(barrier)
        if fails:
            raise PeerWaitTimeout(f"{fails} peer wait(s) timed out on rank {self.rank}: gathered arrays were stale")
