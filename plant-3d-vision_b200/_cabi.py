"""ctypes binding of ``libbp_b200.so`` (C ABI declared in ``include/bp_b200.h``).

This is the only place the package touches native code.  A missing library is a hard error:
there is deliberately no CPU / numpy / torch fallback for the hot path.
"""
import ctypes
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
# BP_B200_LIB: another build of the SAME sources (compile-time A/B variants, tools/build_variants.py) - never a fallback
LIB_PATH = os.environ.get("BP_B200_LIB") or os.path.join(_PKG, "libbp_b200.so")

MODE_CARVING, MODE_AVERAGING = 0, 1
MASK_U8, MASK_I32, MASK_F32, MASK_U16 = 0, 1, 2, 3
FLAG_INVERT, FLAG_BOOL_INVERT, FLAG_DEVICE_MASKS, FLAG_PACKED_MASKS = 1, 2, 4, 8
TAIL_F64, TAIL_EXP_CLIP = 1, 2

# every symbol include/bp_b200.h declares; tests check the library exports all of them
SYMBOLS = [
    "bp_create", "bp_destroy", "bp_clear", "bp_process_view", "bp_process_views",
    "bp_get_values", "bp_get_values_tail", "bp_get_values_device", "bp_device_state", "bp_sync", "bp_stream", "bp_timing",
    "bp_host_alloc", "bp_host_free", "bp_last_error", "bp_version", "bp_selftest_division",
    "bp_pack_views", "bp_packed_size", "bp_create_striped", "bp_ingest_bytes", "bp_view_rect", "bp_pack_views_cams", "bp_set_readback_target", "bp_spec_stats",
    "bp_work_stats", "bp_box_bounds_host", "bp_spec_snapshot", "bp_tail_static",
]

# every symbol include/vol2pcd_b200.h declares (volume -> point-cloud handoff)
V2P_SYMBOLS = [
    "bp_vol2pcd", "bp_vol2pcd_info", "bp_vol2pcd_fetch", "bp_vol2pcd_device", "bp_vol2pcd_signed_distance",
    "bp_vol2pcd_free", "bp_multiclass_select", "bp_device_free", "bp_score_points",
]
V2P_U8, V2P_I8, V2P_I32, V2P_F32, V2P_F64 = 0, 1, 2, 3, 4
# every symbol include/masks_b200.h declares (mask producer)
MASKS_SYMBOLS = ["bp_masks_filter"]
MASKS_LINEAR, MASKS_EXCESS_GREEN = 0, 1
V2P_BIN_GT, V2P_BIN_EQ, V2P_BIN_GE = 0, 1, 2
V2P_FLAG_DEVICE_VOLUME = 1

_lib = None


class BackprojectionError(RuntimeError):
    """A libbp_b200 call failed (message from bp_last_error)."""


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
                "g.build()'` (nvcc, sm_100a). plant3dvision_b200 has no CPU fallback.")
        L = ctypes.CDLL(LIB_PATH)
        c_f = ctypes.POINTER(ctypes.c_float)
        c_i = ctypes.POINTER(ctypes.c_int32)
        vp = ctypes.c_void_p
        L.bp_create.argtypes = [ctypes.POINTER(vp), c_i, ctypes.c_int32, ctypes.c_int32, c_f,
                                ctypes.c_float, ctypes.c_int, ctypes.c_float, ctypes.c_int]
        L.bp_create_striped.argtypes = [ctypes.POINTER(vp), c_i, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32, c_f,
                                        ctypes.c_float, ctypes.c_int, ctypes.c_float, ctypes.c_int]
        L.bp_destroy.argtypes = [vp]
        L.bp_clear.argtypes = [vp]
        L.bp_process_view.argtypes = [vp, c_f, c_f, c_f, vp, ctypes.c_int32, ctypes.c_int32,
                                      ctypes.c_int, ctypes.c_uint32, c_f]
        L.bp_process_views.argtypes = [vp, ctypes.c_int32, c_f, ctypes.POINTER(vp), c_i,
                                       ctypes.c_int, ctypes.c_uint32, c_f]
        L.bp_pack_views.argtypes = [vp, ctypes.c_int32, ctypes.POINTER(vp), c_i, ctypes.c_int, ctypes.c_uint32, vp,
                                    ctypes.c_int64]
        L.bp_pack_views_cams.argtypes = [vp, ctypes.c_int32, c_f, ctypes.POINTER(vp), c_i, ctypes.c_int, ctypes.c_uint32,
                                         vp, ctypes.c_int64]
        L.bp_packed_size.argtypes = [ctypes.c_int32, ctypes.c_int32, ctypes.POINTER(ctypes.c_int64)]
        L.bp_get_values.argtypes = [vp, vp]
        L.bp_get_values_device.argtypes = [vp, vp]
        L.bp_get_values_tail.argtypes = [vp, vp, ctypes.c_uint32, ctypes.POINTER(ctypes.c_double),
                                         ctypes.POINTER(ctypes.c_double)]
        L.bp_device_state.argtypes = [vp, ctypes.POINTER(vp), ctypes.POINTER(ctypes.c_int64), c_i]
        L.bp_sync.argtypes = [vp]
        L.bp_stream.argtypes = [vp, ctypes.POINTER(vp)]
        L.bp_timing.argtypes = [vp, c_f, c_f, ctypes.POINTER(ctypes.c_int64)]
        L.bp_selftest_division.argtypes = [ctypes.c_int, ctypes.c_uint64, ctypes.c_uint64,
                                           ctypes.POINTER(ctypes.c_uint64)]
        L.bp_set_readback_target.argtypes = [vp, vp]
        L.bp_spec_snapshot.argtypes = [vp]
        L.bp_tail_static.argtypes = [ctypes.c_int, ctypes.c_int, vp, ctypes.c_int64, ctypes.c_int32, ctypes.c_int32, c_i,
                                     ctypes.c_uint32, vp, vp, vp]
        L.bp_spec_stats.argtypes = [vp, ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(ctypes.c_int64)]
        L.bp_work_stats.argtypes = [vp, ctypes.POINTER(ctypes.c_int64), ctypes.POINTER(ctypes.c_int64),
                                    ctypes.POINTER(ctypes.c_int64)]
        L.bp_ingest_bytes.argtypes = [vp, ctypes.POINTER(ctypes.c_int64)]
        L.bp_view_rect.argtypes = [c_i, ctypes.c_int32, ctypes.c_int32, c_f, ctypes.c_float, c_f, ctypes.c_int32,
                                   ctypes.c_int32, c_i]
        L.bp_box_bounds_host.argtypes = [c_i, c_f, ctypes.c_float, c_f, ctypes.c_int32, ctypes.c_int32,
                                         ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double), c_i]
        L.bp_host_alloc.argtypes = [ctypes.POINTER(vp), ctypes.c_size_t]
        L.bp_host_free.argtypes = [vp]
        c_d = ctypes.POINTER(ctypes.c_double)
        c_l = ctypes.POINTER(ctypes.c_int64)
        L.bp_vol2pcd.argtypes = [ctypes.POINTER(vp), ctypes.c_int, vp, ctypes.c_int, ctypes.c_uint32, c_i,
                                 ctypes.c_int64, ctypes.c_int, ctypes.c_double, c_d, ctypes.c_double,
                                 ctypes.c_double, ctypes.c_double, ctypes.c_double, c_d, ctypes.c_int]
        L.bp_vol2pcd_info.argtypes = [vp, c_l, c_f, c_f, c_l, c_l]
        L.bp_vol2pcd_fetch.argtypes = [vp, vp, vp]
        L.bp_vol2pcd_device.argtypes = [vp, ctypes.POINTER(vp), ctypes.POINTER(vp)]
        L.bp_vol2pcd_signed_distance.argtypes = [vp, vp, c_d]
        L.bp_vol2pcd_free.argtypes = [vp]
        L.bp_multiclass_select.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.POINTER(vp), ctypes.c_int,
                                           ctypes.c_uint32, c_i, ctypes.c_int, ctypes.c_double, ctypes.c_double,
                                           ctypes.c_double, ctypes.c_int, ctypes.POINTER(vp)]
        L.bp_device_free.argtypes = [ctypes.c_int, vp]
        L.bp_score_points.argtypes = [ctypes.c_int, ctypes.c_int64, c_d, ctypes.c_int32, c_d, ctypes.POINTER(vp), c_i, c_i,
                                      ctypes.c_int32, c_d, c_i, c_f]
        L.bp_masks_filter.argtypes = [ctypes.c_int, ctypes.c_int32, ctypes.POINTER(vp), ctypes.c_int32, ctypes.c_int32,
                                      ctypes.c_int32, ctypes.c_int, c_d, ctypes.c_double, ctypes.c_int32,
                                      ctypes.POINTER(ctypes.c_uint8), c_i, ctypes.POINTER(vp), ctypes.POINTER(vp), c_f]
        for name in V2P_SYMBOLS + MASKS_SYMBOLS:
            getattr(L, name).restype = ctypes.c_int
        L.bp_last_error.restype = ctypes.c_char_p
        L.bp_version.restype = ctypes.c_char_p
        for name in SYMBOLS:
            if name not in ("bp_last_error", "bp_version"):
                getattr(L, name).restype = ctypes.c_int
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        msg = lib().bp_last_error().decode("utf-8", "replace")
        if rc == -1:
            raise ValueError(f"libbp_b200: {msg}")
        if rc == -3:
            raise MemoryError(f"libbp_b200: {msg}")
        raise BackprojectionError(f"libbp_b200 (code {rc}): {msg}")
