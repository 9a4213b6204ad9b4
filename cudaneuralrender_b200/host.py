"""Python mirror of the reference's host interface for the neural-SDF hot path, over the C ABI.

Names, argument meaning and error behaviour follow the reference (paths relative to its root):
  Shape / Matrix / Image     src/neuralUtils/{shape,matrix,image}.hh
  DenseLayer / NeuralNetwork src/layers/denseLayer.hh, src/neuralNetwork.hh
  render_kernel / copyViewMatrices / copyStaticSettings   src/volumeRender_kernel.cu:608-706
  updateViewMatrices         src/main.cpp:207-222

torch is used for device memory and streams only; all arithmetic happens in libnsdf.so.
The C++ mirror of the same interface is in include/*.hh (used by apps/).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import numpy as np
import torch

from . import _capi as capi
from ._capi import (ACT_RELU, ACT_TANH_LINEAR, COLOR_FACING, COLOR_MATCAP, PREC_FP32, PREC_FP32_WARP, PREC_TC_BF16,
                    PREC_TC_FP16, PREC_TC_FP16_ACC16, PREC_TC_FP16X3, SCENE_MANYSPHERE, SCENE_PROGRAM, SCENE_TANH, MATH_REFERENCE,
                    MATH_DETERMINISTIC, Camera, RenderOpts, RenderStats)

ReLU, Tanh = ACT_RELU, ACT_TANH_LINEAR  # enum Activation, src/layers/denseLayer.hh:6-9


def _stream_ptr() -> int:
    return int(torch.cuda.current_stream().cuda_stream)


class Shape:
    """src/neuralUtils/shape.hh:3-7"""

    def __init__(self, x: int = 1, y: int = 1):
        self.x, self.y = int(x), int(y)

    def __iter__(self):
        return iter((self.x, self.y))

    def __eq__(self, o):
        return isinstance(o, Shape) and (self.x, self.y) == (o.x, o.y)

    def __repr__(self):
        return f"Shape({self.x}, {self.y})"


class _Buffer:
    """Host + device buffer pair with explicit copies (Matrix / Image, matrix.cu:12-72).
    Element (x, y) lives at flat index y*shape.x + x.  Copying the object aliases storage."""
    _np_dtype = np.float32
    _torch_dtype = torch.float32

    def __init__(self, x_dim=1, y_dim=1, hostOnly: bool = False):
        if isinstance(x_dim, Shape):
            self.shape = Shape(x_dim.x, x_dim.y)
            if isinstance(y_dim, bool):   # Matrix(Shape, hostOnly)
                hostOnly = y_dim
        else:
            self.shape = Shape(x_dim, y_dim)
        self.hostOnly = hostOnly
        self.hostData: Optional[np.ndarray] = None
        self.deviceData: Optional[torch.Tensor] = None

    def size(self) -> int:
        return self.shape.x * self.shape.y

    def allocateMemory(self):
        if self.hostData is None:
            self.hostData = np.zeros(self.size(), dtype=self._np_dtype)
        if not self.hostOnly and self.deviceData is None:
            self.deviceData = torch.zeros(self.size(), dtype=self._torch_dtype, device="cuda")

    def maybeAllocateMemory(self, shape: Shape):
        # matrix.cu:40-45: allocates only when nothing is allocated yet (never re-allocates)
        if self.hostData is None and self.deviceData is None:
            self.shape = Shape(shape.x, shape.y)
            self.allocateMemory()

    def copyHostToDevice(self):
        if self.hostData is None or self.deviceData is None:
            print("[ERROR] Unable to copy host to device, data not allocated...")  # matrix.cu:52-54
            return
        xfer = np.int32 if self._torch_dtype == torch.int32 else self._np_dtype
        self.deviceData.copy_(torch.from_numpy(self.hostData.view(xfer)))

    def copyDeviceToHost(self):
        if self.hostData is None or self.deviceData is None:
            print("[ERROR] Unable to copy device to host, data not allocated...")
            return
        n = self.size()
        self.hostData[:n] = self.deviceData[:n].cpu().numpy().view(self._np_dtype)

    def __getitem__(self, i):
        return self.hostData[i]

    def __setitem__(self, i, v):
        self.hostData[i] = v


class Matrix(_Buffer):
    """src/neuralUtils/matrix.hh:9-37 (float)."""


class Image(_Buffer):
    """src/neuralUtils/image.hh:12-43 (uint RGBA8 packed a<<24|b<<16|g<<8|r)."""
    _np_dtype = np.uint32
    _torch_dtype = torch.int32

    def loadPNG(self, filename: str) -> bool:
        L = capi.lib()
        w, h = C.c_int(0), C.c_int(0)
        rc = L.nsdf_png_load(filename.encode(), None, 0, C.byref(w), C.byref(h))
        if rc:
            print("Error reading png: " + str(filename))
            return False
        self.maybeAllocateMemory(Shape(w.value, h.value))
        if self.size() < w.value * h.value:
            return False
        rc = L.nsdf_png_load(filename.encode(), self.hostData.ctypes.data, self.hostData.size, C.byref(w), C.byref(h))
        if rc:
            return False
        if not self.hostOnly:
            self.copyHostToDevice()
        return True

    def savePNG(self, filename: str, doFlip: bool = True, doMirror: bool = True) -> bool:
        if self.hostData is None:
            print("[ERROR] no data to save...")
            return False
        rc = capi.lib().nsdf_png_save(filename.encode(), self.hostData.ctypes.data, self.shape.x, self.shape.y,
                                      1 if doFlip else 0)
        if rc:
            print("[ERROR] Unable to save png: " + filename)
        return rc == 0

    def savePPM(self, filename: str) -> bool:
        return capi.lib().nsdf_ppm_save(filename.encode(), self.hostData.ctypes.data, self.shape.x, self.shape.y) == 0

    def as_array(self) -> np.ndarray:
        return self.hostData.reshape(self.shape.y, self.shape.x)


class DenseLayer:
    """src/layers/denseLayer.hh:11-41.  weights is [in][out] (the Keras kernel); W is stored
    [out][in] (denseLayer.cu:217-227).  forward = one nsdf_dense_layer_forward launch."""

    def __init__(self, name: str, weights, biases, activation: int, hostOnly: bool = False):
        w = np.asarray(weights, dtype=np.float32)
        assert w.ndim == 2
        self.name = name
        self.activation = int(activation)
        self.hostOnly = hostOnly
        self.W = Matrix(w.shape[0], w.shape[1], hostOnly)   # shape.x = in, shape.y = out
        self.b = Matrix(w.shape[1], 1, hostOnly)
        self.W.allocateMemory()
        self.b.allocateMemory()
        self.W.hostData[:] = np.ascontiguousarray(w.T).ravel()
        self.b.hostData[:] = np.asarray(biases, dtype=np.float32).ravel()
        if not hostOnly:
            self.W.copyHostToDevice()
            self.b.copyHostToDevice()
        self.Z = Matrix(1, 1, hostOnly)
        self.Z.hostData = None

    def getName(self): return self.name
    def getType(self): return 0  # eDense
    def getXDim(self): return self.W.shape.x
    def getYDim(self): return self.W.shape.y
    def getWeightsMatrix(self): return self.W
    def getBiasVector(self): return self.b
    def getNumWeightParams(self): return self.W.size()
    def getNumBiasParams(self): return self.b.size()

    def forward(self, A: Matrix, maxBatchSize: int = -1) -> Matrix:
        batch = A.shape.y
        cap = max(batch, maxBatchSize)
        if self.Z.deviceData is None:  # allocated once with maxBatch columns (denseLayer.cu:236-240)
            self.Z.shape = Shape(self.W.shape.y, cap)
            self.Z.deviceData = torch.empty(self.W.shape.y * cap, dtype=torch.float32, device="cuda")
            self.Z.hostData = np.zeros(self.W.shape.y * cap, dtype=np.float32)
        if self.Z.deviceData.numel() < self.W.shape.y * batch:
            raise RuntimeError("DenseLayer::forward: batch larger than the Z buffer allocated on first use")
        capi.check(capi.lib().nsdf_dense_layer_forward(self.W.deviceData.data_ptr(), A.deviceData.data_ptr(),
                                                       self.Z.deviceData.data_ptr(), self.b.deviceData.data_ptr(),
                                                       self.W.shape.y, 1, self.W.shape.x, batch, self.activation,
                                                       _stream_ptr()), "nsdf_dense_layer_forward")
        self.Z.shape = Shape(self.W.shape.y, batch)
        return self.Z


class NeuralNetwork:
    """src/neuralNetwork.hh:9-30.  forward() runs the fused all-layer kernel through
    nsdf_mlp_forward (precision selects the arithmetic rung); forward_layered() walks the layers
    one launch each like the reference (src/neuralNetwork.cpp:54-63)."""

    def __init__(self, path: Optional[str] = None):
        self.layers: List[DenseLayer] = []
        self._handle = C.c_void_p(None)
        self._Y = Matrix(1, 1)
        self.precision = PREC_FP32
        if path is not None and not self.load(path):
            raise SystemExit(1)   # ctor-from-path exits (neuralNetwork.cpp:36-42)

    # -- model handle ---------------------------------------------------------------------
    def _invalidate(self):
        if self._handle:
            capi.lib().nsdf_model_destroy(self._handle)
            self._handle = C.c_void_p(None)

    def handle(self) -> C.c_void_p:
        if not self._handle:
            n = len(self.layers)
            if n == 0:
                raise RuntimeError("NeuralNetwork has no layers")
            fpp = C.POINTER(C.c_float)
            Ws = (fpp * n)(*[l.W.hostData.ctypes.data_as(fpp) for l in self.layers])
            bs = (fpp * n)(*[l.b.hostData.ctypes.data_as(fpp) for l in self.layers])
            ins = (C.c_int * n)(*[l.getXDim() for l in self.layers])
            outs = (C.c_int * n)(*[l.getYDim() for l in self.layers])
            acts = (C.c_int * n)(*[l.activation for l in self.layers])
            h = C.c_void_p(None)
            capi.check(capi.lib().nsdf_model_create(Ws, bs, ins, outs, acts, n, C.byref(h)), "nsdf_model_create")
            self._handle = h
        return self._handle

    def __del__(self):
        try:
            self._invalidate()
        except Exception:
            pass

    # -- reference interface ----------------------------------------------------------------
    def addLayer(self, layer: DenseLayer):
        self.layers.append(layer)
        self._invalidate()

    def getLayers(self): return self.layers
    def getNumWeightParams(self): return sum(l.getNumWeightParams() for l in self.layers)
    def getNumBiasParams(self): return sum(l.getNumBiasParams() for l in self.layers)

    def load(self, path: str, hostOnly: bool = False) -> bool:
        """NeuralNetwork::load (src/neuralNetwork.cpp:85-151) through the library's own HDF5 reader."""
        L = capi.lib()
        h = C.c_void_p(None)
        rc = L.nsdf_model_load_h5(path.encode(), C.byref(h))
        if rc == 10003:
            import os
            if not os.path.exists(path):
                raise FileNotFoundError(path)   # the reference lets HighFive's exception escape
            return False
        capi.check(rc, "nsdf_model_load_h5")
        n = L.nsdf_model_num_layers(h)
        dims = (C.c_int * (n + 1))()
        L.nsdf_model_dims(h, dims)
        cnt = C.c_size_t(0)
        L.nsdf_model_params(h, None, 0, C.byref(cnt))
        params = np.empty(cnt.value, dtype=np.float32)
        L.nsdf_model_params(h, params.ctypes.data_as(C.POINTER(C.c_float)), params.size, C.byref(cnt))
        p = 0
        names = ["dense"] + [f"dense_{i}" for i in range(1, n)]
        for l in range(n):
            i, o = dims[l], dims[l + 1]
            W = params[p:p + i * o].reshape(o, i); p += i * o
            b = params[p:p + o]; p += o
            self.layers.append(DenseLayer(names[l], W.T, b, Tanh if l == n - 1 else ReLU, hostOnly))
        self._invalidate()
        self._handle = h
        return True

    def save(self, path: str) -> None:
        """Not in the reference: writes the layers back as a Keras save_weights-style file (nsdf_h5_write_keras)."""
        write_keras_h5(path, [(l.W.hostData.reshape(l.getYDim(), l.getXDim()).T, l.b.hostData) for l in self.layers])

    def forward(self, X: Matrix, maxBatchSize: int = -1, precision: Optional[int] = None) -> Matrix:
        batch = X.shape.y
        out_dim = self.layers[-1].getYDim()
        need = out_dim * max(batch, maxBatchSize, 1)
        if self._Y.deviceData is None or self._Y.deviceData.numel() < need:
            self._Y.deviceData = torch.empty(need, dtype=torch.float32, device="cuda")
            self._Y.hostData = np.zeros(need, dtype=np.float32)
        capi.check(capi.lib().nsdf_mlp_forward(self.handle(), X.deviceData.data_ptr(), self._Y.deviceData.data_ptr(),
                                               batch, self.precision if precision is None else precision,
                                               _stream_ptr()), "nsdf_mlp_forward")
        self._Y.shape = Shape(out_dim, batch)
        return self._Y

    def forward_layered(self, X: Matrix, maxBatchSize: int = -1) -> Matrix:
        Z = X
        for layer in self.layers:
            Z = layer.forward(Z, maxBatchSize)
        return Z

    # -- tensor-level conveniences (bench / tests) -------------------------------------------------
    def forward_tensor(self, x: torch.Tensor, precision: int = PREC_FP32, out: Optional[torch.Tensor] = None):
        assert x.is_cuda and x.dtype == torch.float32 and x.is_contiguous()
        n = x.numel() // self.layers[0].getXDim()
        if out is None:
            out = torch.empty(n * self.layers[-1].getYDim(), dtype=torch.float32, device=x.device)
        capi.check(capi.lib().nsdf_mlp_forward(self.handle(), x.data_ptr(), out.data_ptr(), n, precision,
                                               _stream_ptr()), "nsdf_mlp_forward")
        return out

    def forward_host(self, x: np.ndarray, precision: int = PREC_FP32) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=np.float32)
        n = x.size // self.layers[0].getXDim()
        y = np.empty(n * self.layers[-1].getYDim(), dtype=np.float32)
        capi.check(capi.lib().nsdf_mlp_forward_host(self.handle(), x.ctypes.data, y.ctypes.data, n, precision),
                   "nsdf_mlp_forward_host")
        return y

    def forward_lattice(self, res: int, first: int = 0, count: Optional[int] = None, lo: float = -1.0, hi: float = 1.0,
                        precision: int = PREC_FP32, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        if count is None:
            count = res ** 3 - first
        if out is None:
            out = torch.empty(count, dtype=torch.float32, device="cuda")
        capi.check(capi.lib().nsdf_mlp_forward_lattice(self.handle(), out.data_ptr(), res, lo, hi, first, count,
                                                       precision, _stream_ptr()), "nsdf_mlp_forward_lattice")
        return out


# ---- weight files without a device ------------------------------------------------
def read_keras_h5(path: str):
    """[(kernel[in][out], bias[out])] in network order, through the library's own HDF5 reader (nsdf_h5_read_keras)."""
    L = capi.lib()
    n, cnt = C.c_int(0), C.c_size_t(0)
    capi.check(L.nsdf_h5_read_keras(path.encode(), C.byref(n), None, 0, None, 0, C.byref(cnt)), "nsdf_h5_read_keras")
    dims = (C.c_int * (n.value + 1))()
    params = np.empty(cnt.value, dtype=np.float32)
    capi.check(L.nsdf_h5_read_keras(path.encode(), C.byref(n), dims, n.value + 1, params.ctypes.data_as(C.POINTER(C.c_float)),
                                    params.size, C.byref(cnt)), "nsdf_h5_read_keras")
    out, p = [], 0
    for l in range(n.value):
        i, o = dims[l], dims[l + 1]
        W = params[p:p + i * o].reshape(o, i); p += i * o
        out.append((np.ascontiguousarray(W.T), params[p:p + o].copy())); p += o
    return out


def write_keras_h5(path: str, layers, storage: int = 0) -> None:
    """layers: [(kernel[in][out], bias[out])] -> Keras save_weights-style file (groups dense, dense_1, ...; nsdf_h5_write_keras).
    storage: 0 contiguous datasets, 1 chunked, 2 chunked + shuffle + deflate filters (NSDF_H5_*)."""
    dims, chunks = [], []
    for k, (kernel, bias) in enumerate(layers):
        kernel = np.asarray(kernel, dtype=np.float32); bias = np.asarray(bias, dtype=np.float32).ravel()
        assert kernel.ndim == 2 and bias.size == kernel.shape[1]
        if k == 0:
            dims.append(kernel.shape[0])
        assert kernel.shape[0] == dims[-1]
        dims.append(kernel.shape[1])
        chunks += [np.ascontiguousarray(kernel.T).ravel(), bias]
    params = np.concatenate(chunks).astype(np.float32)
    capi.check(capi.lib().nsdf_h5_write_keras_ex(path.encode(), params.ctypes.data_as(C.POINTER(C.c_float)),
                                                 (C.c_int * len(dims))(*dims), len(layers), int(storage)), "nsdf_h5_write_keras")


# ---- renderer: the three extern "C" entry points of src/volumeRender_kernel.cu ------------------
class _RenderState:
    """What the reference keeps in __constant__ symbols and compile-time constants.  Starts as nsdf_render_opts_default does:
    the reference's HEAD build (sceneSDF = manySphere, reference math)."""

    def __init__(self):
        self.cam = Camera()
        self.color_type = COLOR_FACING
        self.num_inputs = 3
        self.scene = SCENE_MANYSPHERE
        self.program = None     # (SceneOp * n) for SCENE_PROGRAM
        self.max_steps = 6000
        self.flags = 0
        self.evict_steps = 0
        self.math = MATH_REFERENCE
        self.tuning = None      # capi.Tuning or None (library defaults)


_state = _RenderState()


def copyStaticSettings(colorType: int, numInputs: int = 3):
    """src/volumeRender_kernel.cu:702-706"""
    _state.color_type = int(colorType)
    _state.num_inputs = int(numInputs)


def copyViewMatrices(invViewMatrix: Sequence[float], normalMatrix: Sequence[float], frameNumber: int = 0,
                     numInputs: int = 3):
    """src/volumeRender_kernel.cu:694-700 (12 + 16 floats, row-major)."""
    iv = np.asarray(invViewMatrix, dtype=np.float32).ravel()
    nm = np.asarray(normalMatrix, dtype=np.float32).ravel()
    assert iv.size == 12 and nm.size == 16
    for i in range(12):
        _state.cam.inv_view[i] = float(iv[i])
    for i in range(16):
        _state.cam.normal[i] = float(nm[i])
    _state.cam.frame = int(frameNumber)
    _state.num_inputs = int(numInputs)


def updateViewMatrices(viewRotation=(0.0, 0.0, 0.0), viewTranslation=(0.0, 0.0, -2.0), frameNumber: int = 0):
    """src/main.cpp:207-222 followed by copyViewMatrices."""
    capi.lib().nsdf_camera_from_view(float(viewRotation[0]), float(viewRotation[1]), float(viewTranslation[0]),
                                     float(viewTranslation[1]), float(viewTranslation[2]), int(frameNumber),
                                     C.byref(_state.cam))
    return (np.array(_state.cam.inv_view[:], dtype=np.float32), np.array(_state.cam.normal[:], dtype=np.float32))


def setSceneOptions(scene: int = SCENE_TANH, max_steps: int = 6000, flags: int = 0, evict_steps: int = 0,
                    math: int = MATH_DETERMINISTIC):
    """Compile-time settings of the reference (sceneSDF variant, MAX_STEPS) as run-time options.  This is the switch the
    oracle-facing tests use, so `math` defaults to the deterministic forms the CPU oracle restates; pass
    math=MATH_REFERENCE (the library's own default) to get the reference binary's frames."""
    _state.scene, _state.max_steps, _state.flags, _state.evict_steps = int(scene), int(max_steps), int(flags), int(evict_steps)
    _state.math = int(math)
    if _state.scene != SCENE_PROGRAM:
        _state.program = None


def setSceneProgram(ops, max_steps: int = 6000, flags: int = 0, evict_steps: int = 0, math: int = MATH_DETERMINISTIC):
    """Run-time form of editing sceneSDF (src/volumeRender_kernel.cu:217-230): ops = [(prim, combine, [params]) ...] over the
    reference's analytic SDF library (:67-196), see include/nsdf.h nsdf_scene_op."""
    setSceneOptions(SCENE_PROGRAM, max_steps, flags, evict_steps, math)
    _state.program = ops if not isinstance(ops, (list, tuple)) else capi.scene_ops(ops)


def _opts(tile_rank: int = 0, tile_world: int = 1) -> RenderOpts:
    o = RenderOpts(_state.color_type, _state.scene, _state.max_steps, tile_rank, tile_world, _state.evict_steps, _state.flags)
    if _state.program is not None:
        o.program = C.cast(_state.program, C.POINTER(capi.SceneOp))
        o.n_ops = len(_state.program)
    o.math = _state.math
    if _state.tuning is not None:
        o.tuning = C.pointer(_state.tuning)
    return o


def setTuning(**knobs):
    """nsdf_tuning for the following renders: setTuning() restores the library defaults, setTuning(express_steps=96, ...)
    changes the named knobs of the defaults."""
    if not knobs:
        _state.tuning = None
        return
    t = capi.Tuning()
    capi.lib().nsdf_tuning_default(C.byref(t))
    for k, v in knobs.items():
        if not hasattr(t, k):
            raise KeyError(k)
        setattr(t, k, int(v))
    _state.tuning = t


def dev_schedules_available() -> bool:
    return bool(capi.lib().nsdf_build_has_dev_schedules())


def scene_eval(ops, xyz: torch.Tensor, nsdf: torch.Tensor, frame: int = 0, math: int = MATH_DETERMINISTIC) -> torch.Tensor:
    """nsdf_scene_eval: the program's sceneSDF at n device points with given network outputs."""
    prog = ops if not isinstance(ops, (list, tuple)) else capi.scene_ops(ops)
    out = torch.empty_like(nsdf)
    capi.check(capi.lib().nsdf_scene_eval(C.cast(prog, C.POINTER(capi.SceneOp)), len(prog), xyz.data_ptr(), nsdf.data_ptr(),
                                          frame, math, out.data_ptr(), nsdf.numel(), _stream_ptr()), "nsdf_scene_eval")
    return out


def render_kernel(d_output: torch.Tensor, imageW: int, imageH: int, nn: NeuralNetwork, matcap: Optional[Image] = None,
                  stats: Optional[torch.Tensor] = None, tile_rank: int = 0, tile_world: int = 1,
                  d_output_ptr: Optional[int] = None):
    """render_kernel (src/volumeRender_kernel.cu:608-692).  gridSize/blockSize of the reference have no
    meaning for the persistent tracer and are dropped.  d_output: W*H int32/uint32 device tensor
    (or d_output_ptr: a raw, possibly peer-mapped, device address).  stats: 6 x int64 device tensor
    accumulated into (nsdf_render_stats)."""
    opts = _opts(tile_rank, tile_world)
    mat_ptr, mw, mh = None, 0, 0
    if matcap is not None and matcap.deviceData is not None:
        mat_ptr, mw, mh = matcap.deviceData.data_ptr(), matcap.shape.x, matcap.shape.y
    out_ptr = d_output_ptr if d_output_ptr is not None else d_output.data_ptr()
    capi.check(capi.lib().nsdf_render(nn.handle(), C.byref(_state.cam), mat_ptr, mw, mh, C.byref(opts), out_ptr,
                                      imageW, imageH, stats.data_ptr() if stats is not None else None, _stream_ptr()),
               "nsdf_render")


def render_kernel_noalloc(d_output: torch.Tensor, imageW: int, imageH: int, nn: NeuralNetwork, matcap: Optional[Image] = None,
                          stats: Optional[torch.Tensor] = None):
    """nsdf_render_noalloc: render_kernel for callers that must not allocate device memory during the call (no counterpart in
    the reference; the tracer's scratch is a fixed device array inside the library, frames of at most 2048 64x64 tiles).
    Same frame bytes as render_kernel; calls that overlap on the GPU race on the shared scratch."""
    opts = _opts(0, 1)
    mat_ptr, mw, mh = None, 0, 0
    if matcap is not None and matcap.deviceData is not None:
        mat_ptr, mw, mh = matcap.deviceData.data_ptr(), matcap.shape.x, matcap.shape.y
    capi.check(capi.lib().nsdf_render_noalloc(nn.handle(), C.byref(_state.cam), mat_ptr, mw, mh, C.byref(opts), d_output.data_ptr(),
                                              imageW, imageH, stats.data_ptr() if stats is not None else None, _stream_ptr()),
               "nsdf_render_noalloc")


def view_cameras(views, frames=None):
    """[(rx_deg, ry_deg) ...] (+ optional frame numbers) -> (Camera * n) array: updateViewMatrices (src/main.cpp:207-222) per view,
    default viewTranslation (0, 0, -2)."""
    n = len(views)
    arr = (Camera * n)()
    for i, (rx, ry) in enumerate(views):
        capi.lib().nsdf_camera_from_view(float(rx), float(ry), 0.0, 0.0, -2.0, int(frames[i]) if frames is not None else 0,
                                         C.byref(arr[i]))
    return arr


def render_views(d_output: torch.Tensor, imageW: int, imageH: int, nn: NeuralNetwork, cams, matcap: Optional[Image] = None,
                 stats: Optional[torch.Tensor] = None, tile_rank: int = 0, tile_world: int = 1):
    """nsdf_render_views: len(cams) frames in ONE persistent launch (the `--spin` loop of src/main.cpp:470-478 as a batch);
    view v lands in d_output[v*W*H : (v+1)*W*H].  cams: (Camera * n) from view_cameras()."""
    opts = _opts(tile_rank, tile_world)
    mat_ptr, mw, mh = None, 0, 0
    if matcap is not None and matcap.deviceData is not None:
        mat_ptr, mw, mh = matcap.deviceData.data_ptr(), matcap.shape.x, matcap.shape.y
    capi.check(capi.lib().nsdf_render_views(nn.handle(), cams, len(cams), mat_ptr, mw, mh, C.byref(opts), d_output.data_ptr(),
                                            imageW, imageH, stats.data_ptr() if stats is not None else None, _stream_ptr()),
               "nsdf_render_views")


def render_host(nn: NeuralNetwork, imageW: int, imageH: int, matcap: Optional[np.ndarray] = None):
    """Host-buffer call (nsdf_render_host): uploads the matcap, renders, returns (uint32[H,W], stats dict)."""
    opts = _opts()
    out = np.empty((imageH, imageW), dtype=np.uint32)
    st = RenderStats()
    mp, mw, mh = None, 0, 0
    if matcap is not None:
        matcap = np.ascontiguousarray(matcap, dtype=np.uint32)
        mp, mh, mw = matcap.ctypes.data, matcap.shape[0], matcap.shape[1]
    capi.check(capi.lib().nsdf_render_host(nn.handle(), C.byref(_state.cam), mp, mw, mh, C.byref(opts),
                                           out.ctypes.data, imageW, imageH, C.byref(st)), "nsdf_render_host")
    return out, {k: int(getattr(st, k)) for k, _ in RenderStats._fields_}
