"""ctypes binding of the C ABI in include/mnenc.h (imageexperiments_b200/libmnenc.so).

There is no fallback: a missing or unloadable library raises ImportError-like RuntimeError, and every GPU
entry point returns an error code (raised as MnencError) when no sm_100 device is usable.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MNENC_LIB") or os.path.join(HERE, "libmnenc.so")  # MNENC_LIB: A/B runs against another build

_vp, _sz = C.c_void_p, C.c_size_t

#: every symbol include/mnenc.h declares (checked by tests/test_abi.py)
EXPORTS = (
    "mnenc_last_error", "mnenc_create", "mnenc_destroy", "mnenc_quant_tables", "mnenc_set_quant", "mnenc_encode_rgb8",
    "mnenc_encode_rgb8_device", "mnenc_encode_batch_rgb8", "mnenc_free", "mnenc_get_symbols", "mnenc_pack_symbols", "mnenc_last_timing", "mnenc_last_work",
    "mnenc_debug_plan_sequence", "mnenc_debug_stitch_runs", "mnenc_debug_rle_chunk_size", "mnenc_debug_plan_stripes",
    "mnenc_decode_rgb8", "mnenc_decode_rgb8_device", "mnenc_decode_batch_rgb8", "mnenc_stream_info", "mnenc_debug_parse",
    "mnenc_stripe_begin", "mnenc_stripe_rle", "mnenc_stripe_histogram", "mnenc_stripe_plan", "mnenc_stripe_pack", "mnenc_stripe_finish",
    "mnenc_comm_unique_id", "mnenc_comm_init_rank", "mnenc_comm_destroy", "mnenc_encode_striped_rgb8", "mnenc_last_stripe_timing",
    "mnenc_create_multi", "mnenc_create_auto", "mnenc_last_plan", "mnenc_debug_plan_sharded", "mnenc_debug_fail_next_backend", "mnenc_build_dictionary",
    "mnenc_scatter_streams_static",
)


class MnencError(RuntimeError):
    def __init__(self, code: int, text: str):
        super().__init__(f"mnenc error {code}: {text}")
        self.code = code


_lib = None


def load() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(make -C imageexperiments_b200/csrc). There is no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    L.mnenc_last_error.restype = C.c_char_p
    L.mnenc_create.argtypes = [_vp, C.c_int, C.c_int, C.c_int, _vp, C.c_int, _vp, _vp, _vp, _vp]
    L.mnenc_destroy.argtypes = [_vp]
    L.mnenc_destroy.restype = None
    L.mnenc_quant_tables.argtypes = [C.c_int, C.c_int, C.c_double, _vp]
    L.mnenc_set_quant.argtypes = [_vp, _vp]
    L.mnenc_encode_rgb8.argtypes = [_vp, _vp, _sz, _sz, _sz, _vp, _vp]
    L.mnenc_encode_rgb8_device.argtypes = [_vp, _vp, _sz, _sz, _sz, _vp, _vp]
    L.mnenc_encode_batch_rgb8.argtypes = [_vp, _sz, _vp, C.c_int, _vp, _vp, _vp, _vp, _vp]
    L.mnenc_free.argtypes = [_vp]
    L.mnenc_free.restype = None
    L.mnenc_get_symbols.argtypes = [_vp, _vp, _vp]
    L.mnenc_pack_symbols.argtypes = [_vp, _sz, _sz, _vp, _vp, _vp, _vp]
    L.mnenc_last_timing.argtypes = [_vp, _vp]
    L.mnenc_last_work.argtypes = [_vp, _vp, _vp, _vp]
    L.mnenc_last_plan.argtypes = [_vp, _vp, _vp, _vp]
    L.mnenc_debug_plan_sequence.argtypes = [_vp, _vp, _vp, _sz, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp, _vp]
    L.mnenc_debug_stitch_runs.argtypes = [_vp, C.c_int, C.c_int, C.c_int, _vp, _vp]
    L.mnenc_debug_rle_chunk_size.argtypes = [C.c_uint32, C.c_uint32, C.c_uint64]
    L.mnenc_debug_rle_chunk_size.restype = C.c_uint32
    L.mnenc_debug_plan_stripes.argtypes = [_vp, C.c_int, C.c_int, _vp, _vp, _vp, _vp]
    L.mnenc_stripe_begin.argtypes = [_vp, _vp, C.c_int, _sz, _sz, _sz, C.c_int, C.c_int, C.c_uint32, C.c_uint32, _vp]
    L.mnenc_stripe_rle.argtypes = [_vp, _vp, _vp]
    L.mnenc_stripe_histogram.argtypes = [_vp, _vp, _vp, _vp, _vp, _vp, _vp]
    L.mnenc_stripe_plan.argtypes = [_vp, _vp]
    L.mnenc_stripe_pack.argtypes = [_vp, _vp, _vp, _vp, _vp]
    L.mnenc_stripe_finish.argtypes = [_vp, _vp, _vp]
    L.mnenc_comm_unique_id.argtypes = [_vp]
    L.mnenc_comm_init_rank.argtypes = [_vp, C.c_int, C.c_int, _vp]
    L.mnenc_comm_destroy.argtypes = [_vp]
    L.mnenc_encode_striped_rgb8.argtypes = [_vp, _vp, C.c_int, _sz, _sz, _sz, C.c_uint32, C.c_uint32, _vp, _vp]
    L.mnenc_last_stripe_timing.argtypes = [_vp, _vp]
    L.mnenc_create_multi.argtypes = [_vp, _vp, C.c_int, C.c_int, C.c_int, _vp, C.c_int, _vp, _vp, _vp, _vp]
    L.mnenc_debug_plan_sharded.argtypes = [_vp, _vp, _vp, _vp, C.c_int, C.c_int, _vp, _vp, _vp, _vp]
    L.mnenc_debug_fail_next_backend.argtypes = [_vp]
    L.mnenc_build_dictionary.argtypes = [C.c_int, C.c_int, _vp, _vp, _vp, _vp, _vp, _vp]
    L.mnenc_create_auto.argtypes = [_vp, C.c_int, C.c_int, _vp, C.c_int, _vp, _vp, _vp, _vp]
    L.mnenc_decode_rgb8.argtypes = [_vp, _vp, _sz, _vp, _vp, _vp]
    L.mnenc_decode_rgb8_device.argtypes = [_vp, _vp, _sz, _vp, _sz, _sz, _vp, _vp]
    L.mnenc_decode_batch_rgb8.argtypes = [_vp, _sz, _vp, _vp, _vp, _vp, _vp]
    L.mnenc_stream_info.argtypes = [_vp, _sz, _vp, _vp, _vp, _vp, _vp]
    L.mnenc_debug_parse.argtypes = [_vp, _sz, _vp, _sz, _vp, _vp, _sz, C.c_int]
    L.mnenc_scatter_streams_static.argtypes = [_vp, _vp, C.c_int, _sz, _vp, _sz, _vp, _vp]
    _lib = L
    return L


def check(rc: int) -> int:
    if rc < 0:
        raise MnencError(rc, load().mnenc_last_error().decode(errors="replace"))
    return rc
