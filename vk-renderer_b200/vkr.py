"""ctypes binding of libvkr_b200.so (the C ABI in include/vkr_passes.h) plus image helpers.

PyTorch is used only for device memory and streams.  There is no CPU fallback: importing
this module without the built library raises, and every compute entry needs a CUDA device.
"""
import ctypes as C
import os

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
# VKR_HOST_ONLY=1 (set only by bench.py's CPU arm): bind oracle/libvkr_hostonly.so -- the layout / Halton / producer-twin helpers
# compiled without any device code -- so that a process that measures the CPU side never loads the product library.  Every other
# entry then raises on call.  Without that variable a missing product library is an ImportError: there is no fallback.
HOST_ONLY = os.environ.get("VKR_HOST_ONLY") == "1"
if HOST_ONLY:
    LIB_PATH = os.path.join(os.path.dirname(_HERE), "oracle", "libvkr_hostonly.so")
else:
    LIB_PATH = os.environ.get("VKR_B200_LIB") or os.path.join(_HERE, "libvkr_b200.so")   # override: kernel-variant experiments (tools/variant.sh)
if not os.path.exists(LIB_PATH):
    raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                      "(there is no CPU fallback for the vkr passes)")


class _HostOnlyLib:
    """The host-only library: symbols it lacks (every compute entry) become stubs that raise when called."""

    def __init__(self, cdll):
        object.__setattr__(self, "_cdll", cdll)

    def __getattr__(self, name):
        try:
            f = getattr(self._cdll, name)
        except AttributeError:
            class _Missing:
                restype = argtypes = None

                def __call__(self, *a, **k):
                    raise RuntimeError(f"{name} is not part of oracle/libvkr_hostonly.so (VKR_HOST_ONLY=1: no device code in this process)")
            f = _Missing()
        object.__setattr__(self, name, f)
        return f


lib = _HostOnlyLib(C.CDLL(LIB_PATH)) if HOST_ONLY else C.CDLL(LIB_PATH)

# VkFormat values (include/vkr_passes.h VkrFormat)
R8_UNORM, R8G8B8A8_UNORM, R8G8B8A8_SRGB = 9, 37, 43
R16_UNORM, R16_SFLOAT, R16G16_UNORM, R16G16_SFLOAT = 70, 76, 77, 83
R16G16B16A16_UNORM, R16G16B16A16_SFLOAT = 91, 97
R32_SFLOAT, R32G32B32A32_SFLOAT, D24_UNORM_S8_UINT = 100, 109, 129

_NP = {  # format -> (numpy dtype, channels)
    R8_UNORM: (np.uint8, 1), R8G8B8A8_UNORM: (np.uint8, 4), R8G8B8A8_SRGB: (np.uint8, 4),
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


class ClassifyParams(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("max_roughness", C.c_float), ("glossy_value", C.c_float)]


class RegressionParams(C.Structure):
    _fields_ = [_mat("camera_to_world"), ("fovy", C.c_float), ("aspect", C.c_float), ("znear", C.c_float), ("zfar", C.c_float)]


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


class SsaoParams(C.Structure):
    """SSAOParams as the shader declares it (std140: samples are vec3 at a 16-byte stride)."""
    _fields_ = [_mat("projection"), ("fovy", C.c_float), ("aspect", C.c_float), ("znear", C.c_float), ("zfar", C.c_float),
                ("samples", (C.c_float * 4) * 16)]


class ScreenTraceParams(C.Structure):
    _fields_ = [_mat("normal_mat"), ("random_offset", C.c_float), ("angle_offset", C.c_float), ("fovy", C.c_float), ("aspect", C.c_float),
                ("znear", C.c_float), ("zfar", C.c_float)]


class ProbeTraceParams(C.Structure):
    _fields_ = [_mat("inverse_view"), ("probe_min", C.c_float * 4), ("probe_max", C.c_float * 4), ("grid_size", C.c_uint32), ("fovy", C.c_float),
                ("aspect", C.c_float), ("znear", C.c_float), ("zfar", C.c_float)]


class SsrParams(C.Structure):
    _fields_ = [_mat("normal_mat"), ("fovy", C.c_float), ("aspect", C.c_float), ("znear", C.c_float), ("zfar", C.c_float)]


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

    @classmethod
    def over(cls, buf, width, height, fmt, mips=1):
        """An Image laid out over the start of an existing uint8 tensor (e.g. a slice of a peer-exported slab)."""
        im = cls.__new__(cls)
        im.desc = VkrImage()
        im.nbytes = int(lib.vkr_image_layout(width, height, fmt, mips, C.byref(im.desc)))
        if im.nbytes == 0 or buf.numel() < im.nbytes:
            raise VkrError("Image.over: bad layout or buffer too small")
        im.device, im.buf = buf.device, buf[:im.nbytes]
        im.desc.base = im.buf.data_ptr()
        im.width, im.height, im.format, im.mips = width, height, fmt, mips
        return im

    @staticmethod
    def layout_bytes(width, height, fmt, mips=1):
        d = VkrImage()
        return (int(lib.vkr_image_layout(width, height, fmt, mips, C.byref(d))) + 255) & ~255

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


# ---- synthetic G-buffer producer (include/vkr_synth.h) -----------------------------------
class SynthObject(C.Structure):
    _fields_ = [("center", C.c_float * 3), ("half_extent", C.c_float * 3), ("kind", C.c_uint32),
                ("albedo_srgb8", C.c_uint32), ("material_srgb8", C.c_uint32)]


class SynthScene(C.Structure):
    _fields_ = [("object_count", C.c_uint32), ("objects", SynthObject * 64), ("ground_albedo", C.c_uint32 * 2),
                ("ground_material", C.c_uint32), ("wall_albedo", C.c_uint32), ("wall_material", C.c_uint32),
                ("sky_albedo", C.c_uint32), ("wall_z", C.c_float), ("wall_height", C.c_float), ("wall_half_width", C.c_float)]


class DrawTAAParams(C.Structure):
    """reference src/scene_renderer.hpp:26-33"""
    _fields_ = [_mat("mvp"), _mat("prev_mvp"), _mat("camera"), _mat("prev_camera"), ("jitter", C.c_float * 4),
                ("fovy_aspect_znear_zfar", C.c_float * 4)]


lib.vkr_synth_scene.restype = None
lib.vkr_synth_scene.argtypes = [C.c_uint32, C.c_uint32, C.POINTER(SynthScene)]
lib.vkr_synth_camera.restype = None
lib.vkr_synth_camera.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.POINTER(DrawTAAParams)]
lib.vkr_mat4_inverse.restype = None
lib.vkr_mat4_inverse.argtypes = [C.POINTER(C.c_float), C.POINTER(C.c_float)]
lib.vkr_mat4_transpose.restype = None
lib.vkr_mat4_transpose.argtypes = [C.POINTER(C.c_float), C.POINTER(C.c_float)]
_sig("vkr_synth_gbuffer", C.POINTER(SynthScene), C.POINTER(DrawTAAParams), IP, IP, IP, IP, IP, VP)
_sig("vkr_synth_gbuffer_host", C.POINTER(SynthScene), C.POINTER(DrawTAAParams), IP, IP, IP, IP, IP, C.c_int)


def synth_scene(seed=0, objects=64):
    s = SynthScene()
    lib.vkr_synth_scene(seed, objects, C.byref(s))
    return s


def synth_camera(view_seed, frame, width, height, jitter=True):
    p = DrawTAAParams()
    lib.vkr_synth_camera(view_seed, frame, width, height, 1 if jitter else 0, C.byref(p))
    return p


def mat_inverse(m16):
    out = (C.c_float * 16)()
    lib.vkr_mat4_inverse(m16, out)
    return out


def mat_transpose(m16):
    out = (C.c_float * 16)()
    lib.vkr_mat4_transpose(m16, out)
    return out


class Gbuffer:
    """Image set of `struct Gbuffer` (reference src/scene_renderer.hpp:11-24, ctor src/scene_renderer.cpp:8-44)."""

    def __init__(self, w, h, device="cuda", pinned=False):
        self.w, self.h = w, h
        mips = full_mip_count(w, h)
        mk = lambda f, ww=w, hh=h, m=1: Image(ww, hh, f, m, device=device, pinned=pinned)  # noqa: E731
        self.albedo = mk(R8G8B8A8_SRGB)
        self.normal = mk(R16G16_UNORM)
        self.material = mk(R8G8B8A8_SRGB)
        self.velocity_vectors = mk(R16G16_SFLOAT)
        self.depth = mk(D24_UNORM_S8_UINT, m=mips)
        self.prev_depth = mk(D24_UNORM_S8_UINT, m=mips)
        self.downsampled_normals = mk(R16G16_UNORM, w // 2, h // 2)
        self.downsampled_velocity_vectors = mk(R16G16_SFLOAT, w // 2, h // 2)

    def input_images(self):
        return [self.albedo, self.normal, self.material, self.velocity_vectors, self.depth]


def synth_gbuffer(scene, params, gb, stream=None, host_threads=0):
    """Fill gb's input images for one frame: CUDA ray-caster for device images, host twin for CPU images."""
    a = (gb.albedo.ptr, gb.normal.ptr, gb.material.ptr, gb.velocity_vectors.ptr, gb.depth.ptr)
    if gb.depth.device.type == "cpu":
        check(lib.vkr_synth_gbuffer_host(C.byref(scene), C.byref(params), *a, host_threads))
    else:
        check(lib.vkr_synth_gbuffer(C.byref(scene), C.byref(params), *a, stream_ptr(stream)))


# ---- remaining pass entries ---------------------------------------------------------------
PT = C.POINTER
_sig("vkr_pdf_preintegrate", IP, VP)
_sig("vkr_brdf_preintegrate", FP, IP, VP)
_sig("vkr_sssr_trace", PT(TraceParams), IP, IP, IP, FP, IP, IP, IP, VP)
_sig("vkr_sssr_filter", PT(TraceParams), IP, IP, IP, IP, IP, IP, VP)
_sig("vkr_sssr_blur", PT(BlurParams), IP, IP, IP, IP, IP, IP, IP, IP, VP)
_sig("vkr_gtao_main", PT(GtaoParams), IP, IP, IP, IP, IP, C.c_uint32, VP)
_sig("vkr_gtao_filter", PT(GtaoParams), IP, IP, IP, C.c_uint32, VP)
_sig("vkr_gtao_accumulate", PT(AccumParams), IP, IP, IP, IP, IP, IP, C.c_uint32, VP)
_sig("vkr_deferred_shading", PT(ShadingParams), IP, IP, IP, IP, IP, IP, IP, IP, VP)
_sig("vkr_taa_resolve", PT(TaaParams), IP, IP, IP, IP, IP, IP, VP)


_sig("vkr_sssr_clear_indirect", VP, VP, VP)
_sig("vkr_sssr_classification", PT(ClassifyParams), IP, VP, VP, VP, VP, VP)
_sig("vkr_sssr_trace_indirect", PT(TraceParams), C.c_uint32, IP, IP, IP, FP, VP, VP, C.c_uint32, IP, VP)
_sig("vkr_sssr_tile_regression", PT(RegressionParams), IP, IP, VP)


def _dp(t):
    return C.c_void_p(t.data_ptr())


def sssr_clear_indirect(reflective_indirect, glossy_indirect, stream=None):
    """SSSR_Clear: both int32[3] device tensors <- {0, 1, 1}."""
    check(lib.vkr_sssr_clear_indirect(_dp(reflective_indirect), _dp(glossy_indirect), stream_ptr(stream)))


def sssr_classification(p, material, reflective_tiles, glossy_tiles, reflective_indirect, glossy_indirect, stream=None):
    check(lib.vkr_sssr_classification(C.byref(p), material.ptr, _dp(reflective_tiles), _dp(glossy_tiles), _dp(reflective_indirect),
                                      _dp(glossy_indirect), stream_ptr(stream)))


def sssr_trace_indirect(p, reflection_type, depth, dnormal, material, halton_dev, tiles, indirect, rays, stream=None):
    check(lib.vkr_sssr_trace_indirect(C.byref(p), reflection_type, depth.ptr, dnormal.ptr, material.ptr, _dp(halton_dev), _dp(tiles), _dp(indirect),
                                      tiles.numel(), rays.ptr, stream_ptr(stream)))


def sssr_tile_regression(p, depth, out_planes, stream=None):
    check(lib.vkr_sssr_tile_regression(C.byref(p), depth.ptr, out_planes.ptr, stream_ptr(stream)))


SSSR_STATIC_MAX_TILES = 65536
_sig("vkr_sssr_trace_tiles_static", PT(TraceParams), C.c_float, IP, IP, IP, FP, IP, VP)


def sssr_trace_tiles_static(p, glossy_value, depth, dnormal, material, halton_dev, rays, stream=None):
    """Clear + classification + both indirect traces with the library's static device scratch: allocates nothing (include/vkr_passes.h)."""
    check(lib.vkr_sssr_trace_tiles_static(C.byref(p), float(glossy_value), depth.ptr, dnormal.ptr, material.ptr, _dp(halton_dev), rays.ptr, stream_ptr(stream)))


_sig("vkr_blit_image", IP, IP, VP)
_sig("vkr_readback_pack", IP, C.c_uint32, C.c_uint32, VP, VP)
READBACK_RAW, READBACK_DEPTH24, READBACK_OPAQUE_ALPHA = 0, 1, 2


def blit_image(src, dst, stream=None):
    check(lib.vkr_blit_image(src.ptr, dst.ptr, stream_ptr(stream)))


def readback_pack(image, out_tensor, mip=0, mode=READBACK_RAW, stream=None):
    check(lib.vkr_readback_pack(image.ptr, mip, mode, _dp(out_tensor), stream_ptr(stream)))


def pdf_preintegrate(out, stream=None):
    check(lib.vkr_pdf_preintegrate(out.ptr, stream_ptr(stream)))


def brdf_preintegrate(halton_dev, out, stream=None):
    check(lib.vkr_brdf_preintegrate(C.c_void_p(halton_dev.data_ptr()), out.ptr, stream_ptr(stream)))


def sssr_trace(p, depth, dnormal, material, halton_dev, pdf, rays, occlusion, stream=None):
    check(lib.vkr_sssr_trace(C.byref(p), depth.ptr, dnormal.ptr, material.ptr, C.c_void_p(halton_dev.data_ptr()), pdf.ptr, rays.ptr,
                             occlusion.ptr, stream_ptr(stream)))


def sssr_filter(p, rays, depth, albedo, normal, material, out, stream=None):
    check(lib.vkr_sssr_filter(C.byref(p), rays.ptr, depth.ptr, albedo.ptr, normal.ptr, material.ptr, out.ptr, stream_ptr(stream)))


def sssr_blur(p, depth, normal, refl, material, history, dvel, prev_depth, out, stream=None):
    check(lib.vkr_sssr_blur(C.byref(p), depth.ptr, normal.ptr, refl.ptr, material.ptr, history.ptr, dvel.ptr, prev_depth.ptr, out.ptr,
                            stream_ptr(stream)))


def gtao_main(p, depth, normal, material, pdf, raw, depth_lod=1, stream=None):
    check(lib.vkr_gtao_main(C.byref(p), depth.ptr, normal.ptr, material.ptr, pdf.ptr, raw.ptr, depth_lod, stream_ptr(stream)))


def gtao_filter(p, depth, raw, filtered, depth_lod=1, stream=None):
    check(lib.vkr_gtao_filter(C.byref(p), depth.ptr, raw.ptr, filtered.ptr, depth_lod, stream_ptr(stream)))


def gtao_accumulate(p, depth, prev_depth, filtered, accumulated, dvel, history, depth_lod=1, stream=None):
    check(lib.vkr_gtao_accumulate(C.byref(p), depth.ptr, prev_depth.ptr, filtered.ptr, accumulated.ptr, dvel.ptr, history.ptr, depth_lod,
                                  stream_ptr(stream)))


def deferred_shading(p, albedo, normal, material, depth, occlusion, brdf, reflections, out, stream=None):
    check(lib.vkr_deferred_shading(C.byref(p), albedo.ptr, normal.ptr, material.ptr, depth.ptr, occlusion.ptr, brdf.ptr, reflections.ptr,
                                   out.ptr, stream_ptr(stream)))


def taa_resolve(p, history, prev_depth, depth, velocity, color, out, stream=None):
    check(lib.vkr_taa_resolve(C.byref(p), history.ptr, prev_depth.ptr, depth.ptr, velocity.ptr, color.ptr, out.ptr, stream_ptr(stream)))


# ---- extended launches: array layers and row windows (include/vkr_passes.h VkrLaunchExt) ---------------
class LaunchExt(C.Structure):
    _fields_ = [("layers", C.c_uint32), ("row_begin", C.c_uint32), ("row_end", C.c_uint32), ("flags", C.c_uint32), ("layer_stride", C.c_uint64 * 10),
                ("params_scratch", C.c_void_p), ("first_param", C.c_uint32), ("reserved", C.c_uint32),
                ("peer_count", C.c_uint32), ("reserved2", C.c_uint32), ("peer_delta", C.c_int64 * 7)]


PARAMS_SCRATCH_STRIDE = 256
ROW_ALIGN = 16
EP = PT(LaunchExt)
_sig("vkr_hiz_build_ex", IP, IP, IP, IP, IP, EP, VP)
_sig("vkr_sssr_trace_ex", PT(TraceParams), IP, IP, IP, FP, IP, IP, IP, EP, VP)
_sig("vkr_sssr_filter_ex", PT(TraceParams), IP, IP, IP, IP, IP, IP, EP, VP)
_sig("vkr_sssr_blur_ex", PT(BlurParams), IP, IP, IP, IP, IP, IP, IP, IP, EP, VP)
_sig("vkr_gtao_main_ex", PT(GtaoParams), IP, IP, IP, IP, IP, C.c_uint32, EP, VP)
_sig("vkr_gtao_filter_ex", PT(GtaoParams), IP, IP, IP, C.c_uint32, EP, VP)
_sig("vkr_gtao_accumulate_ex", PT(AccumParams), IP, IP, IP, IP, IP, IP, C.c_uint32, EP, VP)
_sig("vkr_deferred_shading_ex", PT(ShadingParams), IP, IP, IP, IP, IP, IP, IP, IP, EP, VP)
_sig("vkr_taa_resolve_ex", PT(TaaParams), IP, IP, IP, IP, IP, IP, EP, VP)
_sig("vkr_screen_trace_main_ex", PT(ScreenTraceParams), IP, IP, IP, IP, IP, EP, VP)
_sig("vkr_screen_trace_filter_ex", C.c_float, C.c_float, IP, IP, IP, EP, VP)
_sig("vkr_screen_trace_accumulate_ex", PT(C.c_float), IP, IP, IP, IP, EP, VP)
_sig("vkr_gtao_main_graphics", PT(GtaoParams), IP, IP, IP, C.c_uint32, EP, VP)
_sig("vkr_gtao_deinterleave", IP, C.c_uint32, IP, C.c_uint64, C.c_int, VP)
_sig("vkr_gtao_main_deinterleaved", PT(GtaoParams), IP, C.c_uint64, IP, IP, C.c_int, C.c_uint32, EP, VP)
_sig("vkr_gtao_reproject", PT(C.c_float), IP, IP, C.c_uint32, IP, IP, IP, EP, VP)
_sig("vkr_cube2oct", IP, C.c_uint64, IP, C.c_uint64, IP, C.c_uint64, IP, C.c_uint64, C.c_uint32, C.c_uint32, VP)
_sig("vkr_probe_downsample", IP, C.c_uint64, C.c_uint32, C.c_uint32, VP)
_sig("vkr_trace_probe", PT(ProbeTraceParams), IP, IP, IP, C.c_uint64, IP, C.c_uint64, IP, EP, VP)
_sig("vkr_ssao", PT(SsaoParams), IP, IP, VP)
_sig("vkr_ssao_ex", PT(SsaoParams), IP, IP, EP, VP)
_sig("vkr_ssr_legacy", PT(SsrParams), IP, IP, IP, IP, IP, VP)
_sig("vkr_ssr_legacy_ex", PT(SsrParams), IP, IP, IP, IP, IP, EP, VP)


class ImageArray:
    """`layers` images of one extent/format in ONE allocation, layer i at base + i * layer_stride (an array image:
    gpu::ImageInfo.array_layers, reference src/gpu/resource_info.hpp).  `desc` describes layer 0."""

    def __init__(self, width, height, fmt, mips=1, layers=1, device="cuda"):
        self.desc = VkrImage()
        nbytes = int(lib.vkr_image_layout(width, height, fmt, mips, C.byref(self.desc)))
        if nbytes == 0:
            raise VkrError("vkr_image_layout rejected the arguments")
        self.layer_stride = (nbytes + 255) & ~255
        self.layers, self.width, self.height, self.format, self.mips = layers, width, height, fmt, mips
        self.device = torch.device(device)
        self.buf = torch.zeros(self.layer_stride * layers, dtype=torch.uint8, device=self.device)
        self.desc.base = self.buf.data_ptr()
        self.nbytes = nbytes

    @classmethod
    def over(cls, buf, width, height, fmt, mips=1, layers=1):
        """An ImageArray laid out over the start of an existing uint8 tensor (e.g. a slice of a peer-exported slab)."""
        a = cls.__new__(cls)
        a.desc = VkrImage()
        nbytes = int(lib.vkr_image_layout(width, height, fmt, mips, C.byref(a.desc)))
        a.layer_stride = (nbytes + 255) & ~255
        if nbytes == 0 or buf.numel() < a.layer_stride * layers:
            raise VkrError("ImageArray.over: bad layout or buffer too small")
        a.layers, a.width, a.height, a.format, a.mips = layers, width, height, fmt, mips
        a.device, a.buf, a.nbytes = buf.device, buf[:a.layer_stride * layers], nbytes
        a.desc.base = a.buf.data_ptr()
        return a

    @staticmethod
    def array_bytes(width, height, fmt, mips=1, layers=1):
        d = VkrImage()
        return ((int(lib.vkr_image_layout(width, height, fmt, mips, C.byref(d))) + 255) & ~255) * layers

    @property
    def ptr(self):
        return C.byref(self.desc)

    def layer(self, i):
        """An Image aliasing layer i (shares storage)."""
        im = Image.__new__(Image)
        im.desc = VkrImage.from_buffer_copy(self.desc)
        im.nbytes, im.device = self.nbytes, self.device
        im.buf = self.buf[i * self.layer_stride: i * self.layer_stride + self.nbytes]
        im.desc.base = im.buf.data_ptr()
        im.width, im.height, im.format, im.mips = self.width, self.height, self.format, self.mips
        return im


class LayerRange:
    """Layers [first, first + count) of an ImageArray as the `_ex` entries want them: descriptor of the first layer of the range
    plus the array's layer stride (shares storage)."""

    def __init__(self, arr, first, count):
        assert 0 <= first and first + count <= arr.layers
        self.desc = VkrImage.from_buffer_copy(arr.desc)
        self.desc.base = arr.desc.base + first * arr.layer_stride
        self.layer_stride, self.layers = arr.layer_stride, count
        self.nbytes = arr.nbytes
        self.buf = arr.buf[first * arr.layer_stride:(first + count) * arr.layer_stride]     # a range of a range works too
        self.device, self.width, self.height, self.format, self.mips = arr.device, arr.width, arr.height, arr.format, arr.mips
        self._keep = arr

    def layer(self, i):
        return ImageArray.layer(self, i)

    @property
    def ptr(self):
        return C.byref(self.desc)


EXT_PREPARE_ONLY, EXT_PARAMS_PREPARED = 1, 2


def make_ext(images, layers=1, rows=None, scratch=None, flags=0, first_param=0, peers=None):
    """VkrLaunchExt for a call whose VkrImage arguments are `images` (ImageArray or Image, in argument order).
    peers: byte offsets from the pass's output to the same array on other GPUs (fused gather; blur and accumulate only)."""
    e = LaunchExt()
    if peers:
        e.peer_count = len(peers)
        for i, d in enumerate(peers):
            e.peer_delta[i] = int(d)
    e.layers = layers
    e.first_param = first_param
    e.flags = flags
    if rows is not None:
        e.row_begin, e.row_end = rows
    for i, im in enumerate(images):
        e.layer_stride[i] = getattr(im, "layer_stride", 0) if layers > 1 else 0
    e.params_scratch = scratch.data_ptr() if scratch is not None else None
    return e


def params_array(ptype, items):
    arr = (ptype * len(items))()
    for i, it in enumerate(items):
        C.memmove(C.byref(arr, i * C.sizeof(ptype)), C.byref(it), C.sizeof(ptype))
    return arr


def _p(p):
    """A single params struct or a ctypes array of them -> pointer to the first."""
    return p if isinstance(p, C.Array) else C.byref(p)


def hiz_build_ex(depth, normal, velocity, out_normal, out_velocity, ext, stream=None):
    check(lib.vkr_hiz_build_ex(depth.ptr, normal.ptr, velocity.ptr, out_normal.ptr, out_velocity.ptr, C.byref(ext), stream_ptr(stream)))


def sssr_trace_ex(p, depth, dnormal, material, halton_dev, pdf, rays, occlusion, ext, stream=None):
    check(lib.vkr_sssr_trace_ex(_p(p), depth.ptr, dnormal.ptr, material.ptr, C.c_void_p(halton_dev.data_ptr()), pdf.ptr, rays.ptr, occlusion.ptr,
                                C.byref(ext), stream_ptr(stream)))


def sssr_filter_ex(p, rays, depth, albedo, normal, material, out, ext, stream=None):
    check(lib.vkr_sssr_filter_ex(_p(p), rays.ptr, depth.ptr, albedo.ptr, normal.ptr, material.ptr, out.ptr, C.byref(ext), stream_ptr(stream)))


def sssr_blur_ex(p, depth, normal, refl, material, history, dvel, prev_depth, out, ext, stream=None):
    check(lib.vkr_sssr_blur_ex(_p(p), depth.ptr, normal.ptr, refl.ptr, material.ptr, history.ptr, dvel.ptr, prev_depth.ptr, out.ptr, C.byref(ext),
                               stream_ptr(stream)))


def gtao_main_ex(p, depth, normal, material, pdf, raw, ext, depth_lod=1, stream=None):
    check(lib.vkr_gtao_main_ex(_p(p), depth.ptr, normal.ptr, material.ptr, pdf.ptr, raw.ptr, depth_lod, C.byref(ext), stream_ptr(stream)))


def gtao_filter_ex(p, depth, raw, filtered, ext, depth_lod=1, stream=None):
    check(lib.vkr_gtao_filter_ex(_p(p), depth.ptr, raw.ptr, filtered.ptr, depth_lod, C.byref(ext), stream_ptr(stream)))


def gtao_accumulate_ex(p, depth, prev_depth, filtered, accumulated, dvel, history, ext, depth_lod=1, stream=None):
    check(lib.vkr_gtao_accumulate_ex(_p(p), depth.ptr, prev_depth.ptr, filtered.ptr, accumulated.ptr, dvel.ptr, history.ptr, depth_lod, C.byref(ext),
                                     stream_ptr(stream)))


def deferred_shading_ex(p, albedo, normal, material, depth, occlusion, brdf, reflections, out, ext, stream=None):
    check(lib.vkr_deferred_shading_ex(_p(p), albedo.ptr, normal.ptr, material.ptr, depth.ptr, occlusion.ptr, brdf.ptr, reflections.ptr, out.ptr,
                                      C.byref(ext), stream_ptr(stream)))


def taa_resolve_ex(p, history, prev_depth, depth, velocity, color, out, ext, stream=None):
    check(lib.vkr_taa_resolve_ex(_p(p), history.ptr, prev_depth.ptr, depth.ptr, velocity.ptr, color.ptr, out.ptr, C.byref(ext), stream_ptr(stream)))


# ---- earlier passes kept beside the live ones (reference src/ssao.cpp, src/ssr.cpp) ------------------------------
def ssao(p, depth, out, ext=None, stream=None):
    check(lib.vkr_ssao_ex(_p(p), depth.ptr, out.ptr, C.byref(ext) if ext is not None else None, stream_ptr(stream)))


def ssr_legacy(p, depth, normal, frame, material, out, ext=None, stream=None):
    check(lib.vkr_ssr_legacy_ex(_p(p), depth.ptr, normal.ptr, frame.ptr, material.ptr, out.ptr, C.byref(ext) if ext is not None else None,
                                stream_ptr(stream)))


def _e(ext):
    return C.byref(ext) if ext is not None else None


def screen_trace_main(p, depth, normal, color, material, raw, ext=None, stream=None):
    check(lib.vkr_screen_trace_main_ex(_p(p), depth.ptr, normal.ptr, color.ptr, material.ptr, raw.ptr, _e(ext), stream_ptr(stream)))


def screen_trace_filter(znear, zfar, raw, depth, filtered, ext=None, stream=None):
    check(lib.vkr_screen_trace_filter_ex(znear, zfar, raw.ptr, depth.ptr, filtered.ptr, _e(ext), stream_ptr(stream)))


def screen_trace_accumulate(fovy_aspect_znear_zfar, depth, prev_depth, filtered, accumulated, ext=None, stream=None):
    f4 = (C.c_float * 4)(*fovy_aspect_znear_zfar)
    check(lib.vkr_screen_trace_accumulate_ex(f4, depth.ptr, prev_depth.ptr, filtered.ptr, accumulated.ptr, _e(ext), stream_ptr(stream)))


def gtao_main_graphics(p, depth, normal, raw, depth_lod=1, ext=None, stream=None):
    check(lib.vkr_gtao_main_graphics(C.byref(p), depth.ptr, normal.ptr, raw.ptr, depth_lod, _e(ext), stream_ptr(stream)))


def gtao_deinterleave(depth, out_array, pattern_n=2, depth_lod=1, stream=None):
    """out_array: ImageArray of R32_SFLOAT with 4**pattern_n layers."""
    check(lib.vkr_gtao_deinterleave(depth.ptr, depth_lod, out_array.ptr, out_array.layer_stride, pattern_n, stream_ptr(stream)))


def gtao_main_deinterleaved(p, depth_array, normal, raw, pattern_n=2, layer=0, ext=None, stream=None):
    check(lib.vkr_gtao_main_deinterleaved(C.byref(p), depth_array.ptr, depth_array.layer_stride, normal.ptr, raw.ptr, pattern_n, layer, _e(ext),
                                          stream_ptr(stream)))


def gtao_reproject(fovy_aspect_znear_zfar, depth, prev_depth, gtao, prev_gtao, out, depth_lod=1, ext=None, stream=None):
    f4 = (C.c_float * 4)(*fovy_aspect_znear_zfar)
    check(lib.vkr_gtao_reproject(f4, depth.ptr, prev_depth.ptr, depth_lod, gtao.ptr, prev_gtao.ptr, out.ptr, _e(ext), stream_ptr(stream)))


# ---- probe_renderer compute path (include/vkr_probe.h) -------------------------------------------------------------
def cube2oct(cube_color, cube_distance, oct_color, oct_depth, first_layer=0, probe_count=1, stream=None):
    """cube_*: ImageArray with 6 * probe_count faces; oct_*: ImageArray of probe layers."""
    check(lib.vkr_cube2oct(cube_color.ptr, cube_color.layer_stride, cube_distance.ptr, cube_distance.layer_stride, oct_color.ptr, oct_color.layer_stride,
                           oct_depth.ptr, oct_depth.layer_stride, first_layer, probe_count, stream_ptr(stream)))


def probe_downsample(probe_depth, first_layer=0, layer_count=1, stream=None):
    check(lib.vkr_probe_downsample(probe_depth.ptr, probe_depth.layer_stride, first_layer, layer_count, stream_ptr(stream)))


def trace_probe(p, depth, normal, probe_color, probe_depth, out, ext=None, stream=None):
    check(lib.vkr_trace_probe(C.byref(p), depth.ptr, normal.ptr, probe_color.ptr, probe_color.layer_stride, probe_depth.ptr, probe_depth.layer_stride,
                              out.ptr, _e(ext), stream_ptr(stream)))


def sphere_samples(seed=0, count=16):
    """Unit vectors drawn like SSAOPass's constructor (reference src/ssao.cpp:34-52): rejection-sample the unit ball, then
    normalise.  The reference draws from libc rand() with no seed; here a seeded numpy generator."""
    rng = np.random.RandomState(seed)
    out = np.zeros((count, 4), dtype=np.float32)
    i = 0
    while i < count:
        v = (2.0 * rng.rand(3) - 1.0).astype(np.float32)
        l2 = np.float32(v[0] * v[0] + v[1] * v[1] + v[2] * v[2])
        if l2 < 1.0 and l2 > 0.0:
            out[i, :3] = v / np.sqrt(l2)
            i += 1
    return out


def set_ssao_samples(p, samples, reference_packing=False):
    """samples: (16, >=3) floats.  reference_packing=True reproduces the bytes the reference's C++ block holds (glm::vec3
    at a 12-byte stride, src/ssao.cpp:15-22) as its shader reads them (16-byte stride): 48 packed floats land in
    samples[0..11][0..3]; entries 12..15 lie beyond the block in the reference (undefined there, zero here)."""
    a = np.zeros((16, 4), dtype=np.float32)
    s3 = np.asarray(samples, dtype=np.float32)[:16, :3]
    if reference_packing:
        a.reshape(-1)[:48] = s3.reshape(-1)
    else:
        a[:, :3] = s3
    C.memmove(C.addressof(p.samples), a.ctypes.data, a.nbytes)
    return p


def perspective(fovy, aspect, znear, zfar):
    """glm::perspectiveRH_ZO as column-major float[16] — the projection the synthetic camera uses (csrc/synth_scene.cu)."""
    t = np.tan(float(fovy) / 2.0)
    m = np.zeros(16, dtype=np.float64)
    m[0], m[5], m[10], m[11], m[14] = 1.0 / (float(aspect) * t), 1.0 / t, zfar / (znear - zfar), -1.0, -(zfar * znear) / (zfar - znear)
    return (C.c_float * 16)(*m.astype(np.float32))
