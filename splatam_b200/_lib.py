"""ctypes binding of libsplat_b200.so (the C ABI declared in include/splat_b200.h).

There is no fallback: if the shared library is missing or does not export a symbol the header
declares, importing this module raises.  Build it with `python __graft_entry__.py` or
`make -C splatam_b200/csrc`.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# SPLAT_B200_LIB: developer override, used by the A/B scripts to load an alternative build of the same ABI
LIB_PATH = os.environ.get("SPLAT_B200_LIB") or os.path.join(_HERE, "libsplat_b200.so")


class SplatCamera(C.Structure):
    """Mirror of `struct SplatCamera` (include/splat_b200.h)."""
    _fields_ = [
        ("image_height", C.c_int32),
        ("image_width", C.c_int32),
        ("tanfovx", C.c_float),
        ("tanfovy", C.c_float),
        ("bg", C.c_void_p),
        ("scale_modifier", C.c_float),
        ("viewmatrix", C.c_void_p),
        ("projmatrix", C.c_void_p),
        ("sh_degree", C.c_int32),
        ("campos", C.c_void_p),
        ("prefiltered", C.c_int32),
        ("near_plane", C.c_float),
        ("depth_sil", C.c_int32),
    ]


class SplatAdamTensor(C.Structure):
    """Mirror of `struct SplatAdamTensor` (include/splat_b200.h)."""
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
                ("n", C.c_int64), ("lr", C.c_float), ("step", C.c_int32)]


class SplatRowTensor(C.Structure):
    """Mirror of `struct SplatRowTensor` (include/splat_b200.h)."""
    _fields_ = [("src", C.c_void_p), ("dst", C.c_void_p), ("width", C.c_int32)]


_P = C.c_void_p
_SIGNATURES = {
    "splat_geom_state_bytes": (C.c_size_t, [C.c_int32]),
    "splat_image_state_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "splat_binning_state_bytes": (C.c_size_t, [C.c_int32, C.c_int32, C.c_int64]),
    "splat_forward_geometry": (C.c_int, [C.POINTER(SplatCamera), C.c_int32, _P, _P, _P, _P, _P, _P, C.c_int32,
                                         _P, _P, _P, _P, _P, _P]),
    "splat_forward_render": (C.c_int, [C.POINTER(SplatCamera), C.c_int32, _P, _P, _P, _P, _P, C.c_int64, _P, _P, _P]),
    "splat_backward": (C.c_int, [C.POINTER(SplatCamera), C.c_int32, C.c_int64, _P, _P, _P, _P, _P, C.c_int32,
                                 _P, _P, _P, _P, _P, _P, _P, C.c_int32, C.c_int32,
                                 _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, C.c_size_t, _P]),
    "splat_backward_scratch_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "splat_mark_visible": (C.c_int, [C.c_int32, _P, _P, _P, C.c_float, _P, _P]),
    "splat_state_ranges": (_P, [_P, C.c_int32, C.c_int32]),
    "splat_state_n_contrib": (_P, [_P, C.c_int32, C.c_int32]),
    "splat_state_final_T": (_P, [_P, C.c_int32, C.c_int32]),
    "splat_state_point_list": (_P, [_P, C.c_int32, C.c_int32, C.c_int64]),
    "splat_state_gaussian_records": (_P, [_P, C.c_int32]),
    "splat_state_instance_records": (_P, [_P, C.c_int32, C.c_int32, C.c_int64]),
    "splat_rendervar_forward": (C.c_int, [C.c_int32, C.c_int32, _P, _P, _P, _P, _P, _P, C.c_int32, C.c_int32,
                                          _P, _P, _P, _P, _P]),
    "splat_rendervar_scratch_bytes": (C.c_size_t, [C.c_int32]),
    "splat_rendervar_backward": (C.c_int, [C.c_int32, C.c_int32, _P, _P, _P, _P, _P, _P, C.c_int32, C.c_int32,
                                           _P, _P, _P, _P, C.c_int32, C.c_int32, _P, _P, _P, _P, _P, _P,
                                           _P, C.c_size_t, _P]),
    "splat_slam_loss_scratch_bytes": (C.c_size_t, [C.c_int32, C.c_int32, C.c_int32]),
    "splat_slam_loss": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_float, C.c_float, C.c_float,
                                  _P, _P, _P, _P, _P, _P, _P, _P, C.c_size_t, _P]),
    "splat_slam_track_loss_noalloc": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_float, C.c_float, C.c_float,
                                                _P, _P, _P, _P, _P, _P, _P, _P]),
    "splat_adam_step": (C.c_int, [C.POINTER(SplatAdamTensor), C.c_int32, C.c_double, C.c_double, C.c_double, _P, _P, _P]),
    "splat_forward_guard": (C.c_int, [_P, C.c_int32, C.c_int32, C.c_int64, _P, _P]),
    "splat_scale_inplace": (C.c_int, [_P, C.c_int64, _P, _P]),
    "splat_seen_update": (C.c_int, [C.c_int32, _P, _P, _P, _P]),
    "splat_best_pose_update": (C.c_int, [_P, _P, _P, _P, C.c_int32, C.c_int32, _P, _P, _P]),
    "splat_prune_flags": (C.c_int, [C.c_int32, C.c_int32, _P, _P, C.c_float, C.c_float, _P, _P]),
    "splat_compact_scratch_bytes": (C.c_size_t, [C.c_int64]),
    "splat_compact_rows": (C.c_int, [C.c_int64, _P, C.POINTER(SplatRowTensor), C.c_int32, _P, C.c_size_t, _P, _P]),
    "splat_spawn_scratch_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "splat_new_gaussian_mask": (C.c_int, [C.c_int32, C.c_int32, _P, _P, C.c_float, _P, _P, C.c_size_t, _P, _P]),
    "splat_spawn_gaussians": (C.c_int, [C.c_int32, C.c_int32, _P, _P, _P, _P, C.c_float, C.c_float, C.c_float, C.c_float,
                                        _P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_int64, _P, _P, _P, _P, _P, _P, _P]),
    "splat_peer_allreduce_sum": (C.c_int, [_P, _P, C.c_uint64, C.c_int32, C.c_int32, C.c_uint32, C.c_int64, _P]),
    "splat_peer_adam_shard_floats": (C.c_size_t, [C.c_int64, C.c_int32]),
    "splat_peer_set_tuning": (None, [C.c_int32, C.c_int32]),
    "splat_peer_reduce_adam": (C.c_int, [_P, _P, _P, C.c_uint64, C.c_uint64, C.c_int32, C.c_int32, C.c_uint32, C.c_int64,
                                         _P, _P, C.c_int32, C.c_int32, C.c_double, C.c_double, C.c_double, _P, _P, _P]),
    "splat_profile_enable": (C.c_int, [C.c_int]),
    "splat_profile_read": (C.c_int, [_P, _P, C.c_int32]),
    "splat_profile_read_gaps": (C.c_int, [_P, C.c_int32]),
    "splat_profile_kernel_name": (C.c_char_p, [C.c_int]),
    "splat_debug_count_blended": (C.c_int, [C.POINTER(SplatCamera), C.c_int32, _P, _P, _P, _P, C.c_int64, _P, _P, _P]),
    "splat_debug_fp32_peak": (C.c_int, [_P, _P, _P]),
    "splat_set_gather": (C.c_int, [C.c_int]),
    "splat_debug_exp_mismatches": (C.c_int, [C.c_uint32, C.c_uint64, _P, _P]),
    "splat_abi_version": (C.c_int, []),
    "splat_strerror": (C.c_char_p, [C.c_int]),
    "splat_last_error": (C.c_char_p, []),
    "splat_kernel_launch_count": (C.c_int, []),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "libsplat_b200.so not found at %s -- the CUDA library is mandatory (no CPU fallback). "
            "Build it: python -c 'import __graft_entry__ as g; g.build()'" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing: intended
        fn.restype = res
        fn.argtypes = args
    if lib.splat_abi_version() != 7:
        raise ImportError("libsplat_b200.so ABI version mismatch")
    return lib


lib = _load()


class SplatError(RuntimeError):
    pass


def check(rc):
    if rc != 0:
        raise SplatError("libsplat_b200: %s: %s" % (lib.splat_strerror(rc).decode(), lib.splat_last_error().decode()))
