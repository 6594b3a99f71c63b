"""ctypes binding of oracle/liboracle.so — the CPU restatement used as the parity checker.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  OrcImage has the same POD layout as VkrImage, so
vkr.Image(device="cpu") buffers are passed straight in.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(_ROOT, "oracle")
LIB_PATH = os.path.join(ORACLE_DIR, "liboracle.so")


def build():
    subprocess.run(["make", "-C", ORACLE_DIR], check=True, capture_output=True)


if not os.path.exists(LIB_PATH):
    build()
lib = C.CDLL(LIB_PATH)

import sys  # noqa: E402
sys.path.insert(0, os.path.join(_ROOT, "vk-renderer_b200"))
import vkr  # noqa: E402  (struct layouts + Image helper only; no product compute is called here)

IP = C.POINTER(vkr.VkrImage)
I = C.c_int
U = C.c_uint32
FP = C.POINTER(C.c_float)


def _sig(name, *argtypes, restype=C.c_int):
    f = getattr(lib, name)
    f.restype = restype
    f.argtypes = list(argtypes)
    return f


_sig("orc_set_threads", I, restype=None)
_sig("orc_get_max_threads", restype=C.c_int)
_sig("orc_halton23_seq", U, FP, restype=None)
_sig("orc_downsample_gbuffer", IP, IP, IP, IP, IP, I, I)
_sig("orc_depth_mips", IP, U)
_sig("orc_sssr_trace", C.POINTER(vkr.TraceParams), IP, IP, IP, FP, IP, IP, IP, I, I)
_sig("orc_sssr_filter", C.POINTER(vkr.TraceParams), IP, IP, IP, IP, IP, IP, I, I)
_sig("orc_sssr_blur", C.POINTER(vkr.BlurParams), IP, IP, IP, IP, IP, IP, IP, IP, I, I)
_sig("orc_pdf_preintegrate", IP, I, I)
_sig("orc_brdf_preintegrate", FP, IP, I, I)
_sig("orc_gtao_main", C.POINTER(vkr.GtaoParams), IP, IP, IP, IP, IP, U, I, I)
_sig("orc_gtao_filter", C.POINTER(vkr.GtaoParams), IP, IP, IP, U, I, I)
_sig("orc_gtao_accumulate", C.POINTER(vkr.AccumParams), IP, IP, IP, IP, IP, IP, U, I, I)
_sig("orc_deferred_shading", C.POINTER(vkr.ShadingParams), IP, IP, IP, IP, IP, IP, IP, IP, I, I)
_sig("orc_taa_resolve", C.POINTER(vkr.TaaParams), IP, IP, IP, IP, IP, IP, I, I)
_sig("orc_ssao", C.POINTER(vkr.SsaoParams), IP, IP, I, I)
_sig("orc_ssr_legacy", C.POINTER(vkr.SsrParams), IP, IP, IP, IP, IP, I, I)
_sig("orc_screen_trace_main", C.POINTER(vkr.ScreenTraceParams), IP, IP, IP, IP, IP, I, I)
_sig("orc_screen_trace_filter", C.c_float, C.c_float, IP, IP, IP, I, I)
_sig("orc_screen_trace_accumulate", FP, IP, IP, IP, IP, I, I)
_sig("orc_gtao_main_graphics", C.POINTER(vkr.GtaoParams), IP, IP, IP, U, I, I)
_sig("orc_gtao_deinterleave", IP, U, IP, C.c_uint64, I)
_sig("orc_gtao_main_deinterleaved", C.POINTER(vkr.GtaoParams), IP, C.c_uint64, IP, IP, I, U, I, I)
_sig("orc_gtao_reproject", FP, IP, IP, U, IP, IP, IP, I, I)
_sig("orc_cube2oct", IP, C.c_uint64, IP, C.c_uint64, IP, IP)
_sig("orc_probe_downsample", IP)
_sig("orc_trace_probe", C.POINTER(vkr.ProbeTraceParams), IP, IP, IP, C.c_uint64, IP, C.c_uint64, IP, I, I)
_sig("orc_sssr_classification", I, I, C.c_float, C.c_float, IP, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p)
_sig("orc_sssr_trace_indirect", C.POINTER(vkr.TraceParams), U, IP, IP, IP, FP, C.c_void_p, U, IP)
_sig("orc_sssr_tile_regression", FP, FP, IP, IP)
_sig("orc_blit_image", IP, IP)
_sig("orc_readback_pack", IP, U, U, C.c_void_p)
_sig("orc_probe_encode_normal", FP, FP, restype=None)
_sig("orc_probe_decode_normal", FP, FP, restype=None)
_sig("orc_probe_reconstruct", FP, C.c_float, FP, FP, restype=None)
_sig("orc_probe_project", FP, FP, FP, restype=None)
_sig("orc_probe_half_roundtrip", C.c_float, restype=C.c_float)
_sig("orc_probe_srgb_decode", U, restype=C.c_float)
_sig("orc_probe_srgb_encode", C.c_float, restype=C.c_float)
_sig("orc_taa_offset", U, U, U, FP, restype=None)
_sig("orc_gbuf_point", FP, FP, FP, FP, FP, FP, restype=None)


def fptr(a):
    assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(FP)


def halton23_seq(count=128):
    out = np.zeros((count, 4), dtype=np.float32)
    lib.orc_halton23_seq(count, fptr(out))
    return out


def ok(rc):
    assert rc == 0, f"oracle returned {rc}"


def cpu_image(w, h, fmt, mips=1):
    return vkr.Image(w, h, fmt, mips, device="cpu")
