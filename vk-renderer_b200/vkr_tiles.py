"""One very large frame split into row tiles across the GPUs of a box (BASELINE.json configs[4]: 7680x4320 at 2/4/8 B200).

One process per GPU.  Every rank holds the full-size image set; the full-res INPUTS (albedo, normal, material,
velocity, depth mip 0) and the Hi-Z pyramid are replicated — SSR rays and the resolve's hit-colour fetch are unbounded
in screen space, and rebuilding the pyramid costs 1/12 of the chain — while every other pass computes only
the rank's own rows through the row-window `_ex` entries.  Intermediates that a neighbour's next pass reads beyond its
window travel as HALO ROWS written straight into the neighbour's copy of the image (cudaIpc-mapped peer memory over
NVLink: include/vkr_peer.h), ordered by stream-side flags; the host never synchronises inside a frame and no
collective is involved (NCCL is only used by the caller for the final gather of the TAA target).

Row ownership is in half-res rows, aligned to vkr.ROW_ALIGN (= the tallest tile of any pass); a full-res pass owns the
doubled range.  The result on a rank's own rows is bit-identical to the single-GPU frame provided the halos cover the
footprints: 1 row of rays (5-tap resolve), 11 of reflections (blur radius), 2 of blurred and raw (lighting pick, 4x4 AO
filter), `motion_rows` of accumulated AO and 2*motion_rows+2 of the TAA target (reprojection reach).
"""
import ctypes as C

import torch

import vkr
import vkr_views

_P = vkr.lib


class IpcHandle(C.Structure):
    _fields_ = [("bytes", C.c_ubyte * 64)]


for _n, _a in [("vkr_peer_alloc", [C.c_uint64, C.POINTER(C.c_void_p)]), ("vkr_peer_free", [C.c_void_p]),
               ("vkr_peer_export", [C.c_void_p, C.POINTER(IpcHandle)]), ("vkr_peer_open", [C.POINTER(IpcHandle), C.POINTER(C.c_void_p)]),
               ("vkr_peer_close", [C.c_void_p]),
               ("vkr_peer_push_rows", [vkr.IP, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]),
               ("vkr_peer_push_2d", [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p]),
               ("vkr_peer_signal", [C.c_void_p, C.c_uint32, C.c_void_p]), ("vkr_peer_wait", [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]),
               ("vkr_peer_epoch_begin", [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]),
               ("vkr_peer_epoch_end", [C.c_void_p, C.POINTER(C.c_void_p), C.c_uint32, C.c_void_p]),
               ("vkr_peer_wait_failed", [])]:
    getattr(_P, _n).restype = C.c_int
    getattr(_P, _n).argtypes = _a


class PeerWaitTimeout(RuntimeError):
    """A stream-side wait for a peer's flag gave up (include/vkr_peer.h): the pass behind it ran on stale halo / gather rows."""


class MotionOutOfHalo(RuntimeError):
    """A frame moved further per frame than the halos of the row-tiled frame were sized for (motion_rows)."""


_wait_failures_total = 0


def peer_wait_failures_total():
    """The device-side counter of timed-out waits is read-and-cleared by vkr_peer_wait_failed(); every user of the peer layer in
    this process (RowTiledFrame, PeerGather) goes through this running total instead, so that they do not steal each other's
    counts.  Synchronises the device."""
    global _wait_failures_total
    n = int(_P.vkr_peer_wait_failed())
    if n > 0:
        _wait_failures_total += n
    return _wait_failures_total


class _CudaMem:
    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class PeerSlab:
    """A device allocation shared with the neighbouring ranks.  Layout: FLAG_BYTES of 32-bit flags, then images."""
    FLAG_BYTES = 1024

    def __init__(self, nbytes, device):
        self.nbytes = nbytes + self.FLAG_BYTES
        p = C.c_void_p()
        with torch.cuda.device(device):
            vkr.check(_P.vkr_peer_alloc(self.nbytes, C.byref(p)))
        self.ptr = p.value
        self.tensor = torch.as_tensor(_CudaMem(self.ptr, self.nbytes), device=device)
        self.peers = {}       # rank -> base address of that rank's slab in this process
        self.cursor = self.FLAG_BYTES

    def carve(self, nbytes):
        off = self.cursor
        self.cursor += (nbytes + 255) & ~255
        assert self.cursor <= self.nbytes, "slab too small"
        return self.tensor[off:off + nbytes]

    def handle_bytes(self):
        h = IpcHandle()
        vkr.check(_P.vkr_peer_export(C.c_void_p(self.ptr), C.byref(h)))
        return bytes(h.bytes)

    def open_peer(self, rank, handle_bytes):
        h = IpcHandle()
        C.memmove(h.bytes, handle_bytes, 64)
        p = C.c_void_p()
        vkr.check(_P.vkr_peer_open(C.byref(h), C.byref(p)))
        self.peers[rank] = p.value

    def peer_address(self, rank, local_address):
        return self.peers[rank] + (local_address - self.ptr)

    def close(self, barrier=None):
        """Unmap the neighbours' slabs, then free our own.  `barrier` (a callable that synchronises all ranks) is called between
        the two steps so that no slab is freed while another process still has it mapped."""
        for p in self.peers.values():
            _P.vkr_peer_close(C.c_void_p(p))
        self.peers = {}
        if barrier is not None:
            barrier()
        self.tensor = None
        _P.vkr_peer_free(C.c_void_p(self.ptr))


def tile_rows(h2, world, rank):
    """Half-res rows [a, b) owned by `rank`: contiguous blocks of ROW_ALIGN-row strips, as even as possible."""
    strips = (h2 + vkr.ROW_ALIGN - 1) // vkr.ROW_ALIGN
    s0, s1 = vkr_views.shard_range(strips, world, rank)
    return min(s0 * vkr.ROW_ALIGN, h2), min(s1 * vkr.ROW_ALIGN, h2)


def balance_bounds(bounds, times_ms, min_strips=3):
    """New strip boundaries that equalise the per-rank time, taking the measured time of each current tile as uniformly spread
    over its strips (SSR march length and blur radius vary down the frame, so equal rows are not equal work)."""
    world = len(times_ms)
    total = bounds[-1]
    dens = []
    for r in range(world):
        n = max(bounds[r + 1] - bounds[r], 1)
        dens += [times_ms[r] / n] * (bounds[r + 1] - bounds[r])
    cum = [0.0]
    for d in dens:
        cum.append(cum[-1] + d)
    out = [0]
    for r in range(1, world):
        target = cum[-1] * r / world
        s = min(range(total + 1), key=lambda i: abs(cum[i] - target))
        s = max(s, out[-1] + min_strips)
        out.append(s)
    out.append(total)
    for r in range(world - 1, 0, -1):                       # keep min_strips at the tail as well
        out[r] = min(out[r], out[r + 1] - min_strips)
    return out


def even_bounds(h2, world):
    strips = (h2 + vkr.ROW_ALIGN - 1) // vkr.ROW_ALIGN
    return [vkr_views.shard_range(strips, world, r)[0] for r in range(world)] + [strips]


def halo_ranges(a, b, halo, height):
    """Rows of an image with own rows [a, b) that go to the upper / lower neighbour: the `halo` rows at each end."""
    up = (a, min(a + halo, b)) if a > 0 else None
    down = (max(b - halo, a), b) if b < height else None
    return up, down


class RowTiledFrame:
    """The full chain of one w x h frame sequence on rank `rank` of `world`, own rows only (see the module docstring).

    exchange_handles(fn): fn(list_of_bytes_per_rank) must return every rank's slab handle — the caller supplies the
    transport (torch.distributed.all_gather_object in bench.py / the tests), keeping this class free of any collective.
    """
    EXCHANGES = ("rays", "reflections", "blurred", "raw", "accumulated_ao", "taa")

    def __init__(self, w, h, rank, world, allgather_bytes, device="cuda", motion_rows=8, lut_size=1024, stream=None, wait_timeout_ms=4000, bounds=None):
        self.w, self.h, self.rank, self.world, self.stream = w, h, rank, world, stream
        self.device = torch.device(device)
        self.timeout = wait_timeout_ms
        w2, h2 = w // 2, h // 2
        if bounds is None:
            self.a, self.b = tile_rows(h2, world, rank)          # own half-res rows: an even split of the 16-row strips
        else:                                                    # caller-chosen strip boundaries (cost-balanced tiles, see balance_bounds)
            assert len(bounds) == world + 1 and bounds[0] == 0 and all(x <= y for x, y in zip(bounds, bounds[1:]))
            self.a, self.b = min(bounds[rank] * vkr.ROW_ALIGN, h2), min(bounds[rank + 1] * vkr.ROW_ALIGN, h2)
        if self.a >= self.b:
            raise vkr.VkrError(f"rank {rank} of {world} owns no rows of a {w}x{h} frame: use fewer ranks")
        self.halo = dict(rays=1, reflections=11, blurred=2, raw=2, accumulated_ao=motion_rows, taa=2 * motion_rows + 2)
        self.motion_rows = motion_rows
        self._fail_base = peer_wait_failures_total()
        self._vmax = torch.zeros((), dtype=torch.float16, device=device)   # running max |velocity.y| (uv units) of the frames rendered
        self.gb = vkr.Gbuffer(w, h, device=device)
        self.gb.prev_depth.buf.view(torch.int32).fill_(0x00FFFFFF)   # clear_depth(prev_depth, 1.0), src/main.cpp:306
        shared = [("rays", w2, h2, vkr.R16G16B16A16_UNORM), ("raw", w2, h2, vkr.R16G16B16A16_SFLOAT), ("reflections", w2, h2, vkr.R8G8B8A8_UNORM),
                  ("blurred", w2, h2, vkr.R8G8B8A8_UNORM), ("blurred_history", w2, h2, vkr.R8G8B8A8_UNORM),
                  ("accumulated_ao", w2, h2, vkr.R16G16_SFLOAT), ("accumulated_history", w2, h2, vkr.R16G16_SFLOAT),
                  ("taa_target", w, h, vkr.R16G16B16A16_SFLOAT), ("taa_history", w, h, vkr.R16G16B16A16_SFLOAT)]
        self.slab = PeerSlab(sum(vkr.Image.layout_bytes(ww, hh, f) for _, ww, hh, f in shared), device)
        for name, ww, hh, f in shared:
            setattr(self, name, vkr.Image.over(self.slab.carve(vkr.Image.layout_bytes(ww, hh, f)), ww, hh, f))
        self.filtered = vkr.Image(w2, h2, vkr.R16_SFLOAT, device=device)
        self.color_out = vkr.Image(w, h, vkr.R8G8B8A8_SRGB, device=device)
        self.halton = torch.from_numpy(vkr.halton23_seq(128)).to(device)
        self.pdf = vkr.Image(lut_size, lut_size, vkr.R32_SFLOAT, device=device)
        self.brdf = vkr.Image(lut_size, lut_size, vkr.R16G16_SFLOAT, device=device)
        vkr.pdf_preintegrate(self.pdf, stream=stream)
        vkr.brdf_preintegrate(self.halton, self.brdf, stream=stream)
        handles = allgather_bytes(self.slab.handle_bytes())
        self.up = rank - 1 if rank > 0 else None
        self.down = rank + 1 if rank < world - 1 else None
        for r in (self.up, self.down):
            if r is not None:
                self.slab.open_peer(r, handles[r])
        self.epoch, self.counter, self.frame = 0, 0, 0
        self.xfer = torch.cuda.Stream(device=self.device)        # halo pushes + flag stores run here, beside the compute stream

    # ---- peer plumbing -------------------------------------------------------------------------------------
    def _flag(self, base, exchange, from_upper):
        """Address of the flag (in the slab at `base`) that the upper / lower neighbour raises for `exchange`."""
        return base + 4 * (2 * self.EXCHANGES.index(exchange) + (0 if from_upper else 1))

    def _compute_stream(self):
        return self.stream if self.stream is not None else torch.cuda.current_stream(self.device)

    exchange_enabled = True      # False: render own rows without touching the neighbours (timing a tile on its own)

    def push(self, name, img, full_res=False):
        """After everything queued on the compute stream so far: copy this rank's boundary rows of `img` into both neighbours'
        copies and raise their flags — on the transfer stream, so the compute stream goes on with work that does not need them."""
        if (self.up is None and self.down is None) or not self.exchange_enabled:
            return
        ev = torch.cuda.Event()
        ev.record(self._compute_stream())
        self.xfer.wait_event(ev)
        st = C.c_void_p(self.xfer.cuda_stream)
        k = 2 if full_res else 1
        a, b, hgt = self.a * k, self.b * k, img.height
        up, down = halo_ranges(a, b, self.halo[name], hgt)
        for nb, rows, i_am_upper in ((self.up, up, False), (self.down, down, True)):
            if nb is None or rows is None:
                continue
            peer_img = self.slab.peer_address(nb, img.desc.base)
            vkr.check(_P.vkr_peer_push_rows(img.ptr, 0, rows[0], rows[1], C.c_void_p(peer_img), st))
            vkr.check(_P.vkr_peer_signal(C.c_void_p(self._flag(self.slab.peers[nb], name, from_upper=i_am_upper)), self.epoch, st))

    def wait(self, name, epoch=None):
        """Hold the compute stream until both neighbours' rows of exchange `name` (of this frame, or of `epoch`) have landed."""
        if not self.exchange_enabled:
            return
        st = vkr.stream_ptr(self.stream)
        for nb, from_upper in ((self.up, True), (self.down, False)):
            if nb is not None:
                vkr.check(_P.vkr_peer_wait(C.c_void_p(self._flag(self.slab.ptr, name, from_upper)), self.epoch if epoch is None else epoch, self.timeout, st))

    def exchange(self, name, img, full_res=False):
        """push + wait back to back (no overlap)."""
        self.push(name, img, full_res)
        self.wait(name)

    # ---- one frame -------------------------------------------------------------------------------------------
    def render_frame(self, dp, gtao_angle_jitter=0.0, swap_depth=True):
        """src/main.cpp:345-391 + the remaps :416-420 on the CURRENT contents of self.gb (inputs), own rows only.

        The passes are issued in an order that puts one independent pass between every halo push and the pass that reads it
        (GTAO main behind the ray exchange, the SSR filter behind the raw-AO exchange, ...), and the TAA halo of frame f is only
        waited for behind frame f+1's Hi-Z: five of the six exchanges are hidden behind compute.  Every pass still sees exactly
        the inputs of the reference order — passes that are swapped do not depend on each other (main.cpp:375-388: ssr.run only
        feeds gtao through `raw`, which stays in order)."""
        self.epoch += 1
        P = vkr_views.view_params(dp, self.frame, self.counter, gtao_angle_jitter)
        gb, st = self.gb, self.stream
        # own rows equal the single-GPU frame only while the reprojection reach stays inside the exchanged halos: |velocity.y| <=
        # motion_rows half-res rows.  Folded into a running maximum on the compute stream (one 2-byte-per-texel reduction, no
        # host sync); close() / check_motion() compare it with the bound and raise MotionOutOfHalo.
        v = gb.velocity_vectors
        with torch.cuda.stream(self._compute_stream()):
            vy = v.buf[:v.desc.row_pitch[0] * v.height].view(torch.float16).view(v.height, -1)[:, 1:2 * v.width:2]
            self._vmax = torch.maximum(self._vmax, vy.abs().amax())
        half = vkr.make_ext([], rows=(self.a, self.b))
        full = vkr.make_ext([], rows=self.own_rows(full_res=True))
        self._compute_stream().wait_stream(self.xfer)            # last frame's pushes read images this frame overwrites
        vkr.hiz_build(gb.depth, gb.normal, gb.velocity_vectors, gb.downsampled_normals, gb.downsampled_velocity_vectors, stream=st)
        if self.frame > 0:
            self.wait("taa", self.epoch - 1)                     # neighbours are done with frame f-1: their halos may be overwritten
        vkr.sssr_trace_ex(P["trace"], gb.depth, gb.downsampled_normals, gb.material, self.halton, self.pdf, self.rays, self.raw, half, st)
        self.push("rays", self.rays)
        vkr.gtao_main_ex(P["gtao"], gb.depth, gb.normal, gb.material, self.pdf, self.raw, half, 1, st)
        self.push("raw", self.raw)
        self.wait("rays")
        vkr.sssr_filter_ex(P["trace"], self.rays, gb.depth, gb.albedo, gb.normal, gb.material, self.reflections, half, st)
        self.push("reflections", self.reflections)
        self.wait("raw")
        vkr.gtao_filter_ex(P["gtao"], gb.depth, self.raw, self.filtered, half, 1, st)
        vkr.gtao_accumulate_ex(P["accum"], gb.depth, gb.prev_depth, self.filtered, self.accumulated_ao, gb.downsampled_velocity_vectors,
                               self.accumulated_history, half, 1, st)
        self.push("accumulated_ao", self.accumulated_ao)
        self.wait("reflections")
        vkr.sssr_blur_ex(P["blur"], gb.depth, gb.normal, self.reflections, gb.material, self.blurred_history, gb.downsampled_velocity_vectors,
                         gb.prev_depth, self.blurred, half, st)
        self.push("blurred", self.blurred)
        self.wait("accumulated_ao")
        self.wait("blurred")
        vkr.deferred_shading_ex(P["shade"], gb.albedo, gb.normal, gb.material, gb.depth, self.accumulated_ao, self.brdf, self.blurred, self.color_out, full, st)
        vkr.taa_resolve_ex(P["taa"], self.taa_history, gb.prev_depth, gb.depth, gb.velocity_vectors, self.color_out, self.taa_target, full, st)
        self.push("taa", self.taa_target, full_res=True)
        out = self.taa_target
        if swap_depth:
            gb.depth, gb.prev_depth = gb.prev_depth, gb.depth
        self.taa_history, self.taa_target = self.taa_target, self.taa_history
        self.blurred, self.blurred_history = self.blurred_history, self.blurred
        self.accumulated_history, self.accumulated_ao = self.accumulated_ao, self.accumulated_history
        self.counter = (self.counter + 1) % 16
        self.frame += 1
        return out

    def own_rows(self, full_res=True):
        """Own rows of a full-res (default) or half-res image; the last rank also takes an odd last full-res row."""
        if not full_res:
            return self.a, self.b
        return 2 * self.a, (self.h if self.b >= self.h // 2 else 2 * self.b)

    def wait_failures(self):
        """Timed-out peer waits in this process since this frame object was created (synchronises the device)."""
        return peer_wait_failures_total() - self._fail_base

    def motion_rows_seen(self):
        """Largest |velocity.y| of the frames rendered so far, in half-res rows (synchronises)."""
        return float(self._vmax.item()) * (self.h // 2)

    def check_motion(self):
        seen = self.motion_rows_seen()
        if seen > self.motion_rows:
            raise MotionOutOfHalo(f"a frame moved {seen:.2f} half-res rows; the halos were sized for motion_rows={self.motion_rows}: "
                                  f"tile seams read history rows that were never exchanged")

    def finish(self):
        """Wait (on the compute stream) for the neighbours' last TAA halo: after this every exchanged row of the last frame is in place."""
        if self.frame > 0:
            self.wait("taa", self.epoch)

    def close(self, barrier=None):
        """Joins, releases the slab, and RAISES if the sequence was not valid: a peer wait that timed out (PeerWaitTimeout) or
        motion beyond the halos (MotionOutOfHalo) would otherwise leave a plausible-looking wrong frame behind."""
        torch.cuda.synchronize()
        fails = self.wait_failures()
        if barrier is not None:
            barrier()                                            # nobody is still pushing into our slab
        self.slab.close(barrier)
        if fails:
            raise PeerWaitTimeout(f"{fails} peer wait(s) timed out on rank {self.rank}: halo rows were stale")
        self.check_motion()

    KERNELS_PER_FRAME = 10   # hiz tile + tail, trace, filter, blur, gtao main/filter/accumulate, shading, taa (+ signal/wait kernels)
