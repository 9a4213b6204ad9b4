"""ctypes front-end of the CPU oracle (oracle/liboracle.so) -- TEST INFRASTRUCTURE ONLY.

Importers allowed: tests/, __graft_entry__.smoke(), bench.py (cpu_baseline and
--impl reference legs).  The product package never imports this.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Optional, Tuple

import numpy as np

from . import h5mini

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")
_lib = None

SCENE_TANH = 0
SCENE_MANYSPHERE = 1


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "nsdf_oracle.cpp")
    if force or not os.path.exists(_LIB_PATH) or (
            os.path.exists(src) and os.path.getmtime(src) > os.path.getmtime(_LIB_PATH)):
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        fp = C.POINTER(C.c_float)
        ip = C.POINTER(C.c_int)
        L.oracle_hardware_threads.restype = C.c_int
        L.oracle_tanhf.restype = C.c_float
        L.oracle_tanhf.argtypes = [C.c_float]
        L.oracle_tanhf_array.argtypes = [fp, fp, C.c_int64]
        L.oracle_scene_sdf.restype = C.c_float
        L.oracle_scene_sdf.argtypes = [C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int]
        L.oracle_set_scene_program.argtypes = [C.c_void_p, C.c_int]
        L.oracle_scene_program_eval.argtypes = [fp, fp, C.c_int, fp, C.c_int64]
        L.oracle_sinf.restype = C.c_float
        L.oracle_sinf.argtypes = [C.c_float]
        L.oracle_mlp_forward.restype = C.c_int
        L.oracle_mlp_forward.argtypes = [fp, ip, C.c_int, fp, C.c_int64, fp, C.c_int]
        L.oracle_view_matrices.argtypes = [C.c_float] * 5 + [fp, fp]
        L.oracle_init_rays.restype = C.c_int
        L.oracle_init_rays.argtypes = [fp, C.c_int, C.c_int, fp, C.POINTER(C.c_uint8)]
        L.oracle_render.restype = C.c_int
        L.oracle_render.argtypes = [fp, ip, C.c_int, fp, fp, C.c_int, C.POINTER(C.c_uint32), C.c_int, C.c_int,
                                    C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_uint32),
                                    C.POINTER(C.c_int32), C.POINTER(C.c_uint8), C.POINTER(C.c_int64), C.c_int]
        L.oracle_render_rows.restype = C.c_int
        L.oracle_render_rows.argtypes = [fp, ip, C.c_int, fp, fp, C.c_int, C.POINTER(C.c_uint32), C.c_int, C.c_int,
                                         C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                         C.POINTER(C.c_uint32), C.POINTER(C.c_int32), C.POINTER(C.c_uint8),
                                         C.POINTER(C.c_int64), C.c_int]
        _lib = L
    return _lib


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def hardware_threads() -> int:
    return int(lib().oracle_hardware_threads())


class Model:
    """Flat packed weights: per layer W[out][in] then b[out]; dims = [in0, out0, ...]."""

    def __init__(self, params: np.ndarray, dims: np.ndarray):
        self.params = np.ascontiguousarray(params, dtype=np.float32)
        self.dims = np.ascontiguousarray(dims, dtype=np.int32)
        self.n_layers = len(self.dims) - 1

    @classmethod
    def from_h5(cls, path: str) -> "Model":
        return cls(*h5mini.pack_model(h5mini.load_keras_mlp(path)))

    @classmethod
    def random(cls, dims, seed=0, scale=0.3) -> "Model":
        rng = np.random.default_rng(seed)
        chunks = []
        for i, o in zip(dims[:-1], dims[1:]):
            chunks.append((rng.standard_normal((o, i)) * scale).astype(np.float32).ravel())
            chunks.append((rng.standard_normal(o) * 0.1).astype(np.float32))
        return cls(np.concatenate(chunks), np.asarray(dims, dtype=np.int32))

    def layers(self):
        """[(W[out][in], b[out])]"""
        out, p = [], 0
        for i, o in zip(self.dims[:-1], self.dims[1:]):
            W = self.params[p:p + i * o].reshape(o, i); p += i * o
            b = self.params[p:p + o]; p += o
            out.append((W, b))
        return out


def mlp_forward(model: Model, x: np.ndarray, nthreads: int = 0) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, int(model.dims[0]))
    n = x.shape[0]
    y = np.empty((n, int(model.dims[-1])), dtype=np.float32)
    rc = lib().oracle_mlp_forward(_fp(model.params), model.dims.ctypes.data_as(C.POINTER(C.c_int)),
                                  model.n_layers, _fp(x), n, _fp(y), nthreads)
    if rc:
        raise RuntimeError(f"oracle_mlp_forward failed: {rc}")
    return y


def tanhf(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32)
    y = np.empty_like(x)
    lib().oracle_tanhf_array(_fp(x), _fp(y), x.size)
    return y


def scene_sdf(scene: int, p, nsdf: float, frame: int = 0) -> float:
    return float(lib().oracle_scene_sdf(scene, float(p[0]), float(p[1]), float(p[2]), float(nsdf), frame))


SCENE_PROGRAM = 2
# op codes of include/nsdf.h (nsdf_scene_op)
PRIM_NSDF, PRIM_TANH, PRIM_SPHERE, PRIM_BOX, PRIM_PLANE, PRIM_CYLINDER = range(6)
MOD_DISPLACE, MOD_ROUND, MOD_ONION, MOD_MANY_SPHERE, MOD_MANY_CYLINDER_CUT = range(16, 21)
(COMB_SET, COMB_UNION, COMB_INTERSECT, COMB_SUBTRACT, COMB_SUBTRACT_FROM, COMB_SMOOTH_UNION, COMB_SMOOTH_SUBTRACT,
 COMB_SMOOTH_SUBTRACT_FROM) = range(8)
SCENE_OP_DTYPE = np.dtype([("prim", np.int32), ("combine", np.int32), ("a", np.float32, (8,))])


def scene_ops(ops) -> np.ndarray:
    """[(prim, combine, [a0..]) ...] -> packed nsdf_scene_op records (40 bytes each)."""
    out = np.zeros(len(ops), dtype=SCENE_OP_DTYPE)
    for i, (prim, comb, a) in enumerate(ops):
        out[i]["prim"], out[i]["combine"] = prim, comb
        out[i]["a"][:len(a)] = np.asarray(a, dtype=np.float32)
    return out


def set_scene_program(ops) -> np.ndarray:
    """Program evaluated by scene == SCENE_PROGRAM in scene_sdf / render (process-global in the oracle library)."""
    packed = ops if isinstance(ops, np.ndarray) else scene_ops(ops)
    lib().oracle_set_scene_program(packed.ctypes.data, len(packed))
    return packed


def scene_program_eval(xyz: np.ndarray, nsdf: np.ndarray, frame: int = 0) -> np.ndarray:
    xyz = np.ascontiguousarray(xyz, dtype=np.float32)
    nsdf = np.ascontiguousarray(nsdf, dtype=np.float32)
    out = np.empty(nsdf.size, dtype=np.float32)
    lib().oracle_scene_program_eval(_fp(xyz), _fp(nsdf), frame, _fp(out), nsdf.size)
    return out


def sinf(x: float) -> float:
    return float(lib().oracle_sinf(float(x)))


def view_matrices(rx: float = 0.0, ry: float = 0.0, translation=(0.0, 0.0, -2.0)) -> Tuple[np.ndarray, np.ndarray]:
    """(inv_view 3x4, normal 4x4) for viewRotation=(rx,ry), viewTranslation (default zoom -2,
    src/main.cpp:608-626)."""
    iv = np.empty(12, dtype=np.float32)
    nm = np.empty(16, dtype=np.float32)
    lib().oracle_view_matrices(rx, ry, translation[0], translation[1], translation[2], _fp(iv), _fp(nm))
    return iv, nm


def init_rays(inv_view: np.ndarray, W: int, H: int):
    rays = np.empty((H * W, 10), dtype=np.float32)
    hit = np.empty(H * W, dtype=np.uint8)
    iv = np.ascontiguousarray(inv_view, dtype=np.float32)
    lib().oracle_init_rays(_fp(iv), W, H, _fp(rays), hit.ctypes.data_as(C.POINTER(C.c_uint8)))
    return rays, hit


def render(model: Model, inv_view, normal, W: int, H: int, matcap: Optional[np.ndarray] = None,
           color_type: Optional[int] = None, scene: int = SCENE_TANH, frame: int = 0, max_steps: int = 6000,
           nthreads: int = 0, want_stats: bool = True, rows=None):
    """Returns dict(image uint32[H,W], steps int32[H,W], hit uint8[H,W], queries int).
    rows=(y0, y1, stride) renders only those rows (the others stay 0)."""
    iv = np.ascontiguousarray(inv_view, dtype=np.float32)
    nm = np.ascontiguousarray(normal, dtype=np.float32)
    if color_type is None:
        color_type = 0 if matcap is None else 1
    if matcap is not None:
        matcap = np.ascontiguousarray(matcap, dtype=np.uint32)
        mh, mw = matcap.shape
        mptr = matcap.ctypes.data_as(C.POINTER(C.c_uint32))
    else:
        mh = mw = 0
        mptr = None
    out = np.zeros((H, W), dtype=np.uint32)
    steps = np.zeros((H, W), dtype=np.int32) if want_stats else None
    hit = np.zeros((H, W), dtype=np.uint8) if want_stats else None
    q = C.c_int64(0)
    y0, y1, ys = rows if rows is not None else (0, H, 1)
    rc = lib().oracle_render_rows(_fp(model.params), model.dims.ctypes.data_as(C.POINTER(C.c_int)), model.n_layers,
                             _fp(iv), _fp(nm), frame, mptr, mw, mh, color_type, scene, W, H, max_steps, y0, y1, ys,
                             out.ctypes.data_as(C.POINTER(C.c_uint32)),
                             steps.ctypes.data_as(C.POINTER(C.c_int32)) if want_stats else None,
                             hit.ctypes.data_as(C.POINTER(C.c_uint8)) if want_stats else None,
                             C.byref(q), nthreads)
    if rc:
        raise RuntimeError(f"oracle_render failed: {rc}")
    return {"image": out, "steps": steps, "hit": hit, "queries": int(q.value)}


def load_matcap_png(path: str) -> np.ndarray:
    """Matcap texels packed like Image::loadPNG (src/neuralUtils/image.cu:36-65):
    decode to RGBA8, texel = a<<24 | b<<16 | g<<8 | r.  Uses PIL (test side only)."""
    from PIL import Image
    im = np.asarray(Image.open(path).convert("RGBA"), dtype=np.uint32)
    return (im[..., 3] << 24) | (im[..., 2] << 16) | (im[..., 1] << 8) | im[..., 0]
