"""ctypes binding of libvkr_host.so — the headless C++ host (rendergraph shim + the reference's
pass classes + frame loop, include/vkr_host.h).  This is the public API a user of the repo calls:
upload (or synthesise) a G-buffer, render_frame(), read the TAA target back.
"""
import ctypes as C
import os

import numpy as np
import torch

import vkr

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvkr_host.so")
if not os.path.exists(LIB_PATH):
    raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
lib = C.CDLL(LIB_PATH)

lib.vkr_host_last_error.restype = C.c_char_p
lib.vkr_app_create.restype = C.c_void_p
lib.vkr_app_create.argtypes = [C.c_int, C.c_uint32, C.c_uint32, C.c_void_p]
lib.vkr_app_destroy.restype = None
lib.vkr_app_destroy.argtypes = [C.c_void_p]
lib.vkr_app_upload_gbuffer.restype = C.c_int
lib.vkr_app_upload_gbuffer.argtypes = [C.c_void_p] + [C.c_void_p] * 5
lib.vkr_app_synth_gbuffer.restype = C.c_int
lib.vkr_app_synth_gbuffer.argtypes = [C.c_void_p, C.POINTER(vkr.SynthScene), C.POINTER(vkr.DrawTAAParams)]
lib.vkr_app_render_frame.restype = C.c_int
lib.vkr_app_render_frame.argtypes = [C.c_void_p, C.POINTER(vkr.DrawTAAParams), C.c_float]
lib.vkr_app_get_image.restype = C.c_int
lib.vkr_app_get_image.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(vkr.VkrImage)]
lib.vkr_app_readback.restype = C.c_int
lib.vkr_app_readback.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_uint64]
lib.vkr_app_readback_buffer.restype = C.c_int
lib.vkr_app_readback_buffer.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]
lib.vkr_app_capture.restype = C.c_int
lib.vkr_app_capture.argtypes = [C.c_void_p, C.c_uint32, C.c_char_p]
lib.vkr_write_png_rgba8.restype = C.c_int
lib.vkr_write_png_rgba8.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, C.c_void_p]
lib.vkr_app_sync.restype = C.c_int
lib.vkr_app_sync.argtypes = [C.c_void_p]
lib.vkr_app_stream.restype = C.c_void_p
lib.vkr_app_stream.argtypes = [C.c_void_p]
lib.vkr_app_last_tasks.restype = C.c_char_p
lib.vkr_app_last_tasks.argtypes = [C.c_void_p]
lib.vkr_app_last_lanes.restype = C.c_char_p
lib.vkr_app_last_lanes.argtypes = [C.c_void_p]
lib.vkr_app_frame_graph_stats.restype = C.c_int
lib.vkr_app_frame_graph_stats.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
lib.vkr_app_set_option.restype = C.c_int
lib.vkr_app_set_option.argtypes = [C.c_void_p, C.c_char_p, C.c_float]


lib.vkr_app_bind_gbuffer.restype = C.c_int
lib.vkr_app_bind_gbuffer.argtypes = [C.c_void_p] + [C.POINTER(vkr.VkrImage)] * 5
lib.vkr_app_task_timing.restype = C.c_int
lib.vkr_app_render_probe_grid.restype = C.c_int
lib.vkr_app_render_probe_grid.argtypes = [C.c_void_p, C.POINTER(vkr.SynthScene), C.POINTER(C.c_float), C.POINTER(C.c_float)]
lib.vkr_app_get_image_array.restype = C.c_int
lib.vkr_app_get_image_array.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(vkr.VkrImage), C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)]
lib.vkr_app_get_value.restype = C.c_int
lib.vkr_app_get_value.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_float)]
lib.vkr_app_screen_trace_params.restype = C.c_int
lib.vkr_app_screen_trace_params.argtypes = [C.c_void_p, C.POINTER(vkr.ScreenTraceParams)]
lib.vkr_app_alt_params.restype = C.c_int
lib.vkr_app_alt_params.argtypes = [C.c_void_p, C.POINTER(vkr.SsaoParams), C.POINTER(vkr.SsrParams)]
lib.vkr_app_task_timing.argtypes = [C.c_void_p, C.c_int]
lib.vkr_app_task_times.restype = C.c_int
lib.vkr_app_task_times.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.POINTER(C.c_double), C.POINTER(C.c_uint64), C.c_uint32]


class HostError(RuntimeError):
    pass


def _check(rc):
    if rc != 0:
        raise HostError(f"vkr_host status {rc}: {lib.vkr_host_last_error().decode()}")


class App:
    """Headless frame loop (reference src/main.cpp:255-421) on one device."""

    def __init__(self, width, height, device=0, stream=None):
        self.width, self.height, self.device = width, height, device
        sp = None if stream is None else C.c_void_p(stream.cuda_stream)
        self.h = lib.vkr_app_create(device, width, height, sp)
        if not self.h:
            raise HostError(lib.vkr_host_last_error().decode())

    def close(self):
        if self.h:
            lib.vkr_app_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload_gbuffer(self, gb):
        """gb: vkr.Gbuffer on the CPU (pinned for asynchronous copies)."""
        ims = gb.input_images()
        _check(lib.vkr_app_upload_gbuffer(self.h, *[C.c_void_p(i.buf.data_ptr()) for i in ims]))

    def bind_gbuffer(self, gb):
        """gb: vkr.Gbuffer on this device; its five input images become the frame's G-buffer (no copy)."""
        _check(lib.vkr_app_bind_gbuffer(self.h, *[i.ptr for i in gb.input_images()]))

    def task_timing(self, enable=True):
        _check(lib.vkr_app_task_timing(self.h, 1 if enable else 0))

    def task_times(self):
        """{task name: (total ms, launches)} since task_timing(True)."""
        names = C.create_string_buffer(4096)
        ms, cnt = (C.c_double * 64)(), (C.c_uint64 * 64)()
        n = lib.vkr_app_task_times(self.h, names, 4096, ms, cnt, 64)
        if n < 0:
            _check(n)
        keys = names.value.decode().split("\n")[:n]
        return {k: (ms[i], int(cnt[i])) for i, k in enumerate(keys)}

    def synth_gbuffer(self, scene, params):
        _check(lib.vkr_app_synth_gbuffer(self.h, C.byref(scene), C.byref(params)))

    def render_frame(self, params, angle_jitter=0.0):
        _check(lib.vkr_app_render_frame(self.h, C.byref(params), angle_jitter))

    def sync(self):
        _check(lib.vkr_app_sync(self.h))

    def set_option(self, name, value):
        _check(lib.vkr_app_set_option(self.h, name.encode(), float(value)))

    def alt_params(self):
        """Parameter blocks the optional SSAO / SSR tasks of the last frame received (options alt.ssao / alt.ssr)."""
        a, r = vkr.SsaoParams(), vkr.SsrParams()
        _check(lib.vkr_app_alt_params(self.h, C.byref(a), C.byref(r)))
        return a, r

    def render_probe_grid(self, scene, pmin, pmax):
        _check(lib.vkr_app_render_probe_grid(self.h, C.byref(scene), (C.c_float * 3)(*pmin), (C.c_float * 3)(*pmax)))

    def readback_array(self, name):
        """A named array image (all layers, all mips) as a CPU vkr.ImageArray."""
        import torch
        d, stride, layers = vkr.VkrImage(), C.c_uint64(), C.c_uint32()
        _check(lib.vkr_app_get_image_array(self.h, name.encode(), C.byref(d), C.byref(stride), C.byref(layers)))
        self.sync()
        out = vkr.ImageArray(d.width, d.height, d.format, d.mip_count, layers=layers.value, device="cpu")
        assert out.layer_stride == stride.value
        nbytes = stride.value * layers.value

        class _Mem:
            __cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (d.base, False), "version": 2}
        out.buf.copy_(torch.as_tensor(_Mem(), device=f"cuda:{self.device}").cpu())
        return out

    def get_value(self, name):
        v = C.c_float()
        _check(lib.vkr_app_get_value(self.h, name.encode(), C.byref(v)))
        return v.value

    def screen_trace_params(self):
        p = vkr.ScreenTraceParams()
        _check(lib.vkr_app_screen_trace_params(self.h, C.byref(p)))
        return p

    def last_tasks(self):
        return lib.vkr_app_last_tasks(self.h).decode().split()

    def last_lanes(self):
        """Stream lane (0 or 1) of each task of the last frame, in the order of last_tasks()."""
        return [int(c) for c in lib.vkr_app_last_lanes(self.h).decode()]

    def frame_graph_stats(self):
        """dict(launches, instantiations, eager_submits, capture_failures) of the rendergraph's submits (include/vkr_host.h)."""
        out = (C.c_uint64 * 4)()
        _check(lib.vkr_app_frame_graph_stats(self.h, out))
        return dict(launches=int(out[0]), instantiations=int(out[1]), eager_submits=int(out[2]), capture_failures=int(out[3]))

    def image_desc(self, name):
        d = vkr.VkrImage()
        _check(lib.vkr_app_get_image(self.h, name.encode(), C.byref(d)))
        return d

    FINAL_FRAME, DEPTH_PNG, DEPTH_CSV = 0, 1, 2

    def capture(self, kind, path):
        """The reference's capture buttons (main.cpp:352-357,:393-412): blit + read-back + PNG / CSV file at `path`."""
        _check(lib.vkr_app_capture(self.h, kind, str(path).encode()))

    def readback_buffer(self, name, dtype="int32"):
        """A named rendergraph buffer of the SSSR tile path as a numpy array."""
        import numpy as np
        size = C.c_uint64()
        _check(lib.vkr_app_readback_buffer(self.h, name.encode(), None, 0, C.byref(size)))
        out = np.empty(size.value // np.dtype(dtype).itemsize, dtype=dtype)
        _check(lib.vkr_app_readback_buffer(self.h, name.encode(), out.ctypes.data_as(C.c_void_p), out.nbytes, None))
        self.sync()
        return out

    def readback(self, name, host_image=None, sync=True):
        """Copy a named image (all mips) into a CPU vkr.Image (allocated if not given)."""
        d = self.image_desc(name)
        if host_image is None:
            host_image = vkr.Image(d.width, d.height, d.format, d.mip_count, device="cpu")
        _check(lib.vkr_app_readback(self.h, name.encode(), C.c_void_p(host_image.buf.data_ptr()), host_image.nbytes))
        if sync:
            self.sync()
        return host_image


def write_png_rgba8(path, rgba):
    """Host-only PNG writer of the capture path; rgba = uint8 array (h, w, 4)."""
    import numpy as np
    a = np.ascontiguousarray(rgba, dtype=np.uint8)
    _check(lib.vkr_write_png_rgba8(str(path).encode(), a.shape[1], a.shape[0], a.ctypes.data_as(C.c_void_p)))


def read_png_rgba8(path):
    """Minimal decoder for the files write_png_rgba8 produces (8-bit RGBA, filter 0 scanlines); checks every chunk CRC."""
    import struct
    import zlib

    import numpy as np
    b = open(path, "rb").read()
    assert b[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w, h = 8, b"", 0, 0
    while pos < len(b):
        n, typ = struct.unpack(">I4s", b[pos:pos + 8])
        data = b[pos + 8:pos + 8 + n]
        (crc,) = struct.unpack(">I", b[pos + 8 + n:pos + 12 + n])
        assert zlib.crc32(typ + data) == crc, typ
        if typ == b"IHDR":
            w, h, depth, ctype, comp, flt, inter = struct.unpack(">IIBBBBB", data)
            assert (depth, ctype, comp, flt, inter) == (8, 6, 0, 0, 0)
        elif typ == b"IDAT":
            idat += data
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), dtype=np.uint8).reshape(h, w * 4 + 1)
    assert (raw[:, 0] == 0).all()
    return raw[:, 1:].reshape(h, w, 4).copy()
