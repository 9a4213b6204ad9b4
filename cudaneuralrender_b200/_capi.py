"""ctypes binding of libnsdf.so (include/nsdf.h).  The library is the product; this file only
declares its prototypes.  There is no fallback: if the shared object is missing, loading raises."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# NSDF_LIB=dev selects the development build (make -C csrc dev: A/B schedules, knobs from the environment) for tools/ sweeps
LIB_PATH = os.path.join(_HERE, "libnsdf_dev.so" if os.environ.get("NSDF_LIB") == "dev" else "libnsdf.so")
if os.environ.get("NSDF_LIB_PATH"):      # development: an experimental build of the library
    LIB_PATH = os.environ["NSDF_LIB_PATH"]

NSDF_OK = 0
ACT_RELU, ACT_TANH_LINEAR = 0, 1
PREC_FP32, PREC_TC_FP16, PREC_TC_BF16, PREC_TC_FP16X3, PREC_FP32_WARP, PREC_TC_FP16_ACC16 = 0, 1, 2, 3, 4, 5
COLOR_FACING, COLOR_MATCAP = 0, 1
RENDER_SIMPLE, RENDER_NO_STRAGGLER, RENDER_PHASE_THROUGHPUT, RENDER_PHASE_LATENCY, RENDER_TWO_KERNELS, RENDER_QUEUE, RENDER_NO_QUIET_SMS, RENDER_TC_FARFIELD, RENDER_TWO_PASS, RENDER_HYBRID = 1, 2, 4, 8, 16, 32, 64, 128, 256, 512
RENDER_STAGED_STORES = 1024
IPC_HANDLE_BYTES = 64


class Camera(C.Structure):
    _fields_ = [("inv_view", C.c_float * 12), ("normal", C.c_float * 16), ("frame", C.c_int)]


class SceneOp(C.Structure):
    """nsdf_scene_op (include/nsdf.h): one step of a scene program."""
    _fields_ = [("prim", C.c_int), ("combine", C.c_int), ("a", C.c_float * 8)]


class Tuning(C.Structure):
    """nsdf_tuning (include/nsdf.h)."""
    _fields_ = [(n, C.c_int) for n in ("express_period", "express_steps", "express_warps", "express_old_steps", "express_old_mask",
                                       "express_from_bulk", "express_rays", "express_solo_steps", "tile_priority_steps",
                                       "tile_priority_min_jobs", "tail_lanes", "tail_push_max", "bulk4", "bulk4_evict_steps",
                                       "bulk4_period", "bulk4_drain", "bulk4_refill_min", "bulk4_unroll", "max_ctas")]


class RenderOpts(C.Structure):
    _fields_ = [("color_type", C.c_int), ("scene", C.c_int), ("max_steps", C.c_int), ("tile_rank", C.c_int),
                ("tile_world", C.c_int), ("evict_steps", C.c_int), ("flags", C.c_int), ("program", C.POINTER(SceneOp)),
                ("n_ops", C.c_int), ("math", C.c_int), ("tuning", C.POINTER(Tuning))]


SCENE_TANH, SCENE_MANYSPHERE, SCENE_PROGRAM = 0, 1, 2
MATH_REFERENCE, MATH_DETERMINISTIC = 0, 1   # nsdf_render_opts.math
PRIM_NSDF, PRIM_TANH, PRIM_SPHERE, PRIM_BOX, PRIM_PLANE, PRIM_CYLINDER = range(6)
MOD_DISPLACE, MOD_ROUND, MOD_ONION, MOD_MANY_SPHERE, MOD_MANY_CYLINDER_CUT = range(16, 21)
(COMB_SET, COMB_UNION, COMB_INTERSECT, COMB_SUBTRACT, COMB_SUBTRACT_FROM, COMB_SMOOTH_UNION, COMB_SMOOTH_SUBTRACT,
 COMB_SMOOTH_SUBTRACT_FROM) = range(8)


def scene_ops(ops):
    """[(prim, combine, [a0..]) ...] -> (SceneOp * n) array."""
    arr = (SceneOp * len(ops))()
    for i, (prim, comb, a) in enumerate(ops):
        arr[i].prim, arr[i].combine = int(prim), int(comb)
        for j, v in enumerate(a):
            arr[i].a[j] = float(v)
    return arr


class RenderStats(C.Structure):
    _fields_ = [("queries", C.c_ulonglong), ("hit_pixels", C.c_ulonglong), ("rays", C.c_ulonglong),
                ("evicted", C.c_ulonglong), ("unfinished", C.c_ulonglong), ("pixels", C.c_ulonglong)]


# name -> (restype, argtypes); every symbol include/nsdf.h declares
_vp, _i, _i64, _f, _sz = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t
_fp, _ip, _u32p = C.POINTER(C.c_float), C.POINTER(C.c_int), C.POINTER(C.c_uint32)
PROTOTYPES = {
    "nsdf_version": (C.c_char_p, []),
    "nsdf_error_string": (C.c_char_p, [_i]),
    "nsdf_device_sm_count": (_i, []),
    "nsdf_dev_malloc": (_i, [C.POINTER(_vp), _sz]),
    "nsdf_dev_free": (_i, [_vp]),
    "nsdf_dev_memset": (_i, [_vp, _i, _sz]),
    "nsdf_memcpy_h2d": (_i, [_vp, _vp, _sz]),
    "nsdf_memcpy_d2h": (_i, [_vp, _vp, _sz]),
    "nsdf_device_synchronize": (_i, []),
    "nsdf_set_device": (_i, [_i]),
    "nsdf_device_count": (_i, []),
    "nsdf_model_create": (_i, [C.POINTER(_fp), C.POINTER(_fp), _ip, _ip, _ip, _i, C.POINTER(_vp)]),
    "nsdf_model_create_packed": (_i, [_fp, _ip, _i, C.POINTER(_vp)]),
    "nsdf_model_load_h5": (_i, [C.c_char_p, C.POINTER(_vp)]),
    "nsdf_h5_read_keras": (_i, [C.c_char_p, _ip, _ip, _i, _fp, _sz, C.POINTER(_sz)]),
    "nsdf_h5_write_keras": (_i, [C.c_char_p, _fp, _ip, _i]),
    "nsdf_h5_write_keras_ex": (_i, [C.c_char_p, _fp, _ip, _i, _i]),
    "nsdf_model_destroy": (None, [_vp]),
    "nsdf_model_num_layers": (_i, [_vp]),
    "nsdf_model_dims": (_i, [_vp, _ip]),
    "nsdf_model_is_fused": (_i, [_vp]),
    "nsdf_model_params": (_i, [_vp, _fp, _sz, C.POINTER(_sz)]),
    "nsdf_mlp_forward": (_i, [_vp, _vp, _vp, _i64, _i, _vp]),
    "nsdf_mlp_forward_host": (_i, [_vp, _vp, _vp, _i64, _i]),
    "nsdf_mlp_forward_lattice": (_i, [_vp, _vp, _i, _f, _f, _i64, _i64, _i, _vp]),
    "nsdf_dense_layer_forward": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i64, _i, _vp]),
    "nsdf_render_opts_default": (None, [C.POINTER(RenderOpts)]),
    "nsdf_tuning_default": (None, [C.POINTER(Tuning)]),
    "nsdf_build_has_dev_schedules": (_i, []),
    "nsdf_scene_eval": (_i, [C.POINTER(SceneOp), _i, _vp, _vp, _i, _i, _vp, _i64, _vp]),
    "nsdf_camera_from_view": (None, [_f, _f, _f, _f, _f, _i, C.POINTER(Camera)]),
    "nsdf_render": (_i, [_vp, C.POINTER(Camera), _vp, _i, _i, C.POINTER(RenderOpts), _vp, _i, _i, _vp, _vp]),
    "nsdf_render_noalloc": (_i, [_vp, C.POINTER(Camera), _vp, _i, _i, C.POINTER(RenderOpts), _vp, _i, _i, _vp, _vp]),
    "nsdf_render_views": (_i, [_vp, C.POINTER(Camera), _i, _vp, _i, _i, C.POINTER(RenderOpts), _vp, _i, _i, _vp, _vp]),
    "nsdf_render_views_host": (_i, [_vp, C.POINTER(Camera), _i, _vp, _i, _i, C.POINTER(RenderOpts), _vp, _i, _i,
                                    C.POINTER(RenderStats)]),
    "nsdf_render_host": (_i, [_vp, C.POINTER(Camera), _vp, _i, _i, C.POINTER(RenderOpts), _vp, _i, _i,
                              C.POINTER(RenderStats)]),
    "nsdf_multi_create": (_i, [_vp, _ip, _i, C.POINTER(_vp)]),
    "nsdf_multi_destroy": (None, [_vp]),
    "nsdf_multi_device_count": (_i, [_vp]),
    "nsdf_multi_render": (_i, [_vp, C.POINTER(Camera), _i, _vp, _i, _i, C.POINTER(RenderOpts), _vp, _i, _i, C.POINTER(RenderStats)]),
    "nsdf_ipc_export": (_i, [_vp, C.c_char_p]),
    "nsdf_ipc_open": (_i, [C.c_char_p, C.POINTER(_vp)]),
    "nsdf_ipc_close": (_i, [_vp]),
    "nsdf_png_load": (_i, [C.c_char_p, _vp, _sz, _ip, _ip]),
    "nsdf_png_save": (_i, [C.c_char_p, _vp, _i, _i, _i]),
    "nsdf_ppm_save": (_i, [C.c_char_p, _vp, _i, _i]),
    "nsdf_zlib_inflate": (_i, [_vp, _sz, _vp, _sz, C.POINTER(_sz)]),
}

_lib = None


class NsdfError(RuntimeError):
    def __init__(self, status: int, where: str):
        self.status = status
        msg = lib().nsdf_error_string(status)
        super().__init__(f"{where}: {msg.decode() if msg else status} (status {status})")


def lib():
    """Loads libnsdf.so (built by __graft_entry__.build() / csrc/Makefile).  Raises if absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              f"or `make -C cudaneuralrender_b200/csrc`; there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(status: int, where: str) -> None:
    if status != NSDF_OK:
        raise NsdfError(status, where)
