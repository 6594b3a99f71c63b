"""ctypes binding of libvkr_b200.so (the C ABI in include/vkr_passes.h) plus image helpers.

PyTorch is used only for device memory and streams.  There is no CPU fallback: importing
this module without the built library raises, and every compute entry needs a CUDA device.
"""
import ctypes as C
import os

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvkr_b200.so")
if not os.path.exists(LIB_PATH):
    raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                      "(there is no CPU fallback for the vkr passes)")
lib = C.CDLL(LIB_PATH)

# VkFormat values (include/vkr_passes.h VkrFormat)
R8G8B8A8_UNORM, R8G8B8A8_SRGB = 37, 43
R16_UNORM, R16_SFLOAT, R16G16_UNORM, R16G16_SFLOAT = 70, 76, 77, 83
R16G16B16A16_UNORM, R16G16B16A16_SFLOAT = 91, 97
R32_SFLOAT, R32G32B32A32_SFLOAT, D24_UNORM_S8_UINT = 100, 109, 129

_NP = {  # format -> (numpy dtype, channels)
    R8G8B8A8_UNORM: (np.uint8, 4), R8G8B8A8_SRGB: (np.uint8, 4),
    R16_UNORM: (np.uint16, 1), R16_SFLOAT: (np.float16, 1),
    R16G16_UNORM: (np.uint16, 2), R16G16_SFLOAT: (np.float16, 2),
    R16G16B16A16_UNORM: (np.uint16, 4), R16G16B16A16_SFLOAT: (np.float16, 4),
    R32_SFLOAT: (np.float32, 1), R32G32B32A32_SFLOAT: (np.float32, 4),
    D24_UNORM_S8_UINT: (np.uint32, 1),
}


class VkrImage(C.Structure):
    _fields_ = [("base", C.c_void_p), ("width", C.c_uint32), ("height", C.c_uint32), ("format", C.c_uint32),
                ("mip_count", C.c_uint32), ("mip_offset", C.c_uint64 * 16), ("row_pitch", C.c_uint32 * 16)]


def _mat(name):
    return (name, C.c_float * 16)


class TraceParams(C.Structure):
    _fields_ = [_mat("normal_mat"), ("frame_random", C.c_uint32), ("fovy", C.c_float), ("aspect", C.c_float),
                ("znear", C.c_float), ("zfar", C.c_float), ("max_roughness", C.c_float), ("render_flags", C.c_uint32)]


class BlurParams(C.Structure):
    _fields_ = [_mat("inverse_camera"), _mat("prev_inverse_camera"), ("fovy_aspect_znear_zfar", C.c_float * 4),
                ("max_roughness", C.c_float), ("accumulate", C.c_uint32), ("disable_blur", C.c_uint32)]


class GtaoParams(C.Structure):
    _fields_ = [_mat("normal_mat"), ("fovy", C.c_float), ("aspect", C.c_float), ("znear", C.c_float), ("zfar", C.c_float),
                ("angle_offset", C.c_float), ("weight_ratio", C.c_float), ("use_mis", C.c_uint32),
                ("two_directions", C.c_uint32), ("reflections_only", C.c_uint32)]


class AccumParams(C.Structure):
    _fields_ = [_mat("inverse_camera"), _mat("prev_inverse_camera"), _mat("mvp"),
                ("fovy_aspect_znear_zfar", C.c_float * 4), ("clear_history", C.c_uint32)]


class ShadingParams(C.Structure):
    _fields_ = [_mat("inverse_camera"), _mat("camera"), _mat("shadow_mvp"), ("fovy", C.c_float), ("aspect", C.c_float),
                ("znear", C.c_float), ("zfar", C.c_float), ("min_max_roughness", C.c_float * 2), ("show_ao", C.c_uint32)]


class TaaParams(C.Structure):
    _fields_ = [_mat("inverse_camera"), _mat("prev_inverse_camera"), ("fovy_aspect_znear_zfar", C.c_float * 4)]


def set_mat(field, m):
    """m: 4x4 numpy array in math (row, col) convention -> column-major float[16] (glm::mat4)."""
    flat = np.asarray(m, dtype=np.float32).T.reshape(16)
    for i in range(16):
        field[i] = float(flat[i])


IP = C.POINTER(VkrImage)
lib.vkr_version.restype = C.c_char_p
lib.vkr_last_error.restype = C.c_char_p
lib.vkr_format_texel_size.restype = C.c_uint32
lib.vkr_format_texel_size.argtypes = [C.c_uint32]
lib.vkr_full_mip_count.restype = C.c_uint32
lib.vkr_full_mip_count.argtypes = [C.c_uint32, C.c_uint32]
lib.vkr_image_layout.restype = C.c_uint64
lib.vkr_image_layout.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, IP]
lib.vkr_halton23_seq.restype = None
lib.vkr_halton23_seq.argtypes = [C.c_uint32, C.POINTER(C.c_float)]


class VkrError(RuntimeError):
    pass


def check(rc):
    """Non-zero return -> exception (the reference's error convention, src/gpu/common.hpp:9)."""
    if rc != 0:
        raise VkrError(f"vkr status {rc}: {lib.vkr_last_error().decode()}")


def full_mip_count(w, h):
    return int(lib.vkr_full_mip_count(w, h))


def halton23_seq(count=128):
    out = (C.c_float * (4 * count))()
    lib.vkr_halton23_seq(count, out)
    return np.frombuffer(out, dtype=np.float32).reshape(count, 4).copy()


class Image:
    """A mip chain in one allocation (torch uint8 buffer on `device`), laid out by vkr_image_layout."""

    def __init__(self, width, height, fmt, mips=1, device="cuda", pinned=False):
        self.desc = VkrImage()
        nbytes = lib.vkr_image_layout(width, height, fmt, mips, C.byref(self.desc))
        if nbytes == 0:
            raise VkrError("vkr_image_layout rejected the arguments")
        self.nbytes = int(nbytes)
        self.device = torch.device(device)
        if self.device.type == "cpu" and pinned:
            self.buf = torch.zeros(self.nbytes, dtype=torch.uint8).pin_memory()
        else:
            self.buf = torch.zeros(self.nbytes, dtype=torch.uint8, device=self.device)
        self.desc.base = self.buf.data_ptr()
        self.width, self.height, self.format, self.mips = width, height, fmt, mips

    @property
    def ptr(self):
        return C.byref(self.desc)

    def mip_size(self, lod):
        return max(1, self.width >> lod), max(1, self.height >> lod)

    def _np_view(self, host_u8, lod):
        dt, ch = _NP[self.format]
        w, h = self.mip_size(lod)
        pitch, off = self.desc.row_pitch[lod], self.desc.mip_offset[lod]
        rows = host_u8[off:off + pitch * h].reshape(h, pitch)
        v = rows[:, : w * ch * np.dtype(dt).itemsize].view(dt)
        return v.reshape(h, w, ch) if ch > 1 else v.reshape(h, w)

    def upload(self, arr, lod=0):
        """Copy a (h, w[, c]) numpy array into mip `lod`."""
        host = self.buf.cpu().numpy() if self.device.type != "cpu" else self.buf.numpy()
        dt, _ = _NP[self.format]
        self._np_view(host, lod)[...] = np.asarray(arr).astype(dt, copy=False)
        if self.device.type != "cpu":
            self.buf.copy_(torch.from_numpy(host))
        return self

    def download(self, lod=0):
        host = self.buf.cpu().numpy() if self.device.type != "cpu" else self.buf.numpy()
        return self._np_view(host, lod).copy()

    def copy_from(self, other, non_blocking=False):
        assert other.nbytes == self.nbytes
        self.buf.copy_(other.buf, non_blocking=non_blocking)
        return self

    def clone_to(self, device, pinned=False):
        o = Image(self.width, self.height, self.format, self.mips, device=device, pinned=pinned)
        o.buf.copy_(self.buf)
        return o

    def fill_bytes(self, value=0):
        self.buf.fill_(value)
        return self


def _sig(name, *argtypes):
    f = getattr(lib, name)
    f.restype = C.c_int
    f.argtypes = list(argtypes)
    return f


VP = C.c_void_p
FP = C.c_void_p  # device float pointer
_sig("vkr_downsample_gbuffer", IP, IP, IP, IP, IP, VP)
_sig("vkr_depth_mips", IP, C.c_uint32, VP)
_sig("vkr_hiz_build", IP, IP, IP, IP, IP, VP)


def stream_ptr(stream=None):
    if stream is None:
        stream = torch.cuda.current_stream()
    return C.c_void_p(stream.cuda_stream)


def downsample_gbuffer(depth, normal, velocity, out_normal, out_velocity, stream=None):
    check(lib.vkr_downsample_gbuffer(depth.ptr, normal.ptr, velocity.ptr, out_normal.ptr, out_velocity.ptr, stream_ptr(stream)))


def depth_mips(depth, src_mip=1, stream=None):
    check(lib.vkr_depth_mips(depth.ptr, src_mip, stream_ptr(stream)))


def hiz_build(depth, normal, velocity, out_normal, out_velocity, stream=None):
    check(lib.vkr_hiz_build(depth.ptr, normal.ptr, velocity.ptr, out_normal.ptr, out_velocity.ptr, stream_ptr(stream)))
