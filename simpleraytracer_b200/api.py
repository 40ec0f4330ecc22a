"""ctypes binding of libModelRunner.so -- the C ABI in include/model_runner.h and
include/srt_render.h.  Names, argument meaning and error behaviour are the C
API's (reference model_runner/model_runner.h:115-258); this module adds nothing
but struct declarations and thin RAII wrappers, and it never computes anything
itself: without the CUDA library it raises.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

_LIB_PATH = Path(__file__).resolve().parent / "lib" / "libModelRunner.so"

ML_OK, ML_FAIL = 0, 1
MLX_PRECISION_FP32, MLX_PRECISION_BF16, MLX_PRECISION_BF16X3 = 0, 1, 2
ML_FLOAT32, ML_FLOAT16 = 0, 1
SRT_SHADE_ALBEDO, SRT_SHADE_LAMBERT, SRT_SHADE_SHADOW = 0, 1, 2
SRT_ACCEL_BRUTE, SRT_ACCEL_BVH = 0, 1
SRT_NOALLOC_MAX_TRIANGLES, SRT_NOALLOC_MAX_TILES = 32768, 16384


class MLError(RuntimeError):
    pass


class ml_model_params(C.Structure):
    _fields_ = [("model_path", C.c_char_p), ("input_node", C.c_char_p), ("output_node", C.c_char_p)]


class ml_image_info(C.Structure):
    _fields_ = [("dtype", C.c_int), ("width", C.c_size_t), ("height", C.c_size_t), ("channels", C.c_size_t)]


class srt_camera(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("forward", C.c_float * 3), ("right", C.c_float * 3),
                ("up", C.c_float * 3), ("half_w", C.c_float), ("half_h", C.c_float)]


class srt_light(C.Structure):
    _fields_ = [("position", C.c_float * 3), ("intensity", C.c_float * 3)]


class srt_render_params(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("y0", C.c_uint32), ("y1", C.c_uint32),
                ("buf_y0", C.c_uint32), ("row_block_stride", C.c_uint32), ("row_block_phase", C.c_uint32),
                ("spp", C.c_uint32), ("jitter", C.c_uint32), ("seed", C.c_uint32),
                ("shade", C.c_uint32), ("accel", C.c_uint32), ("collect_stats", C.c_uint32),
                ("t_min", C.c_float), ("shadow_eps", C.c_float), ("background", C.c_float * 3)]


class srt_render_stats(C.Structure):
    _fields_ = [("primary_rays", C.c_uint64), ("shadow_rays", C.c_uint64), ("box_tests", C.c_uint64),
                ("tri_tests", C.c_uint64), ("kernel_ms", C.c_float), ("grid_build_ms", C.c_float),
                ("light_grid_build_ms", C.c_float)]


class srt_scene_info(C.Structure):
    _fields_ = [("n_triangles", C.c_uint64), ("n_bvh_nodes", C.c_uint64), ("device_bytes", C.c_uint64),
                ("bvh_build_ms", C.c_float), ("upload_ms", C.c_float),
                ("bounds_min", C.c_float * 3), ("bounds_max", C.c_float * 3)]


# name -> (restype, argtypes): every symbol include/*.h declares.
_PROTOTYPES = {
    "mlCreateContext": (C.c_void_p, []),
    "mlGetContextError": (C.c_void_p, [C.c_void_p, C.c_char_p, C.c_size_t]),
    "mlReleaseContext": (None, [C.c_void_p]),
    "mlCreateImage": (C.c_void_p, [C.c_void_p, C.POINTER(ml_image_info)]),
    "mlGetImageInfo": (C.c_int, [C.c_void_p, C.POINTER(ml_image_info)]),
    "mlMapImage": (C.c_void_p, [C.c_void_p, C.POINTER(C.c_size_t)]),
    "mlUnmapImage": (C.c_int, [C.c_void_p, C.c_void_p]),
    "mlReleaseImage": (None, [C.c_void_p]),
    "mlCreateModel": (C.c_void_p, [C.c_void_p, C.POINTER(ml_model_params)]),
    "mlGetModelError": (C.c_void_p, [C.c_void_p, C.c_char_p, C.c_size_t]),
    "mlGetModelInfo": (C.c_int, [C.c_void_p, C.POINTER(ml_image_info), C.POINTER(ml_image_info)]),
    "mlSetModelInputInfo": (C.c_int, [C.c_void_p, C.POINTER(ml_image_info)]),
    "mlInfer": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "mlReleaseModel": (None, [C.c_void_p]),
    "srtCreateScene": (C.c_void_p, [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int]),
    "srtGetSceneInfo": (C.c_int, [C.c_void_p, C.POINTER(srt_scene_info)]),
    "srtGetSceneError": (C.c_void_p, [C.c_void_p, C.c_char_p, C.c_size_t]),
    "srtReleaseScene": (None, [C.c_void_p]),
    "srtGetSceneBvh": (C.c_size_t, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]),
    "srtGenCornellBox": (C.c_size_t, [C.c_void_p, C.c_void_p, C.POINTER(srt_camera), C.POINTER(srt_light),
                                      C.c_uint32, C.c_uint32]),
    "srtGenMesh": (C.c_int, [C.c_uint32, C.c_size_t, C.c_void_p, C.c_void_p, C.POINTER(srt_camera),
                             C.POINTER(srt_light), C.c_uint32, C.c_uint32]),
    "srtRender": (C.c_int, [C.c_void_p, C.POINTER(srt_camera), C.POINTER(srt_light),
                            C.POINTER(srt_render_params), C.c_void_p, C.c_void_p, C.c_void_p,
                            C.POINTER(srt_render_stats)]),
    "srtRenderDevice": (C.c_int, [C.c_void_p, C.POINTER(srt_camera), C.POINTER(srt_light),
                                  C.POINTER(srt_render_params), C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_void_p]),
    "srtRenderDeviceNoAlloc": (C.c_int, [C.c_void_p, C.POINTER(srt_camera), C.POINTER(srt_light),
                                         C.POINTER(srt_render_params), C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_void_p]),
    "srtGetRenderStats": (C.c_int, [C.c_void_p, C.POINTER(srt_render_stats)]),
    "srtRenderToImage": (C.c_int, [C.c_void_p, C.POINTER(srt_camera), C.POINTER(srt_light),
                                   C.POINTER(srt_render_params), C.c_void_p, C.POINTER(srt_render_stats)]),
    "srtMeasureFp32Peak": (C.c_double, [C.c_void_p, C.c_int]),
    "mlxInspectModel": (C.c_int, [C.c_char_p, C.c_char_p, C.c_char_p, C.c_char_p, C.c_size_t]),
    "mlxImageDevicePointer": (C.c_void_p, [C.c_void_p, C.c_int, C.c_void_p]),
    "mlxModelLaunchCount": (C.c_size_t, [C.c_void_p]),
    "mlxModelLastBatchMs": (C.c_float, [C.c_void_p]),
    "mlxSetModelPrecision": (C.c_int, [C.c_void_p, C.c_int]),
    "mlxProfileModel": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t,
                                  C.c_void_p]),
    "mlxInferBatch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]),
    "mlxModelTensorCoreLayers": (C.c_size_t, [C.c_void_p]),
    "mlxCreateContextOnDevice": (C.c_void_p, [C.c_int]),
    "mlxIpcExportImage": (C.c_int, [C.c_void_p, C.c_void_p]),
    "mlxIpcOpen": (C.c_void_p, [C.c_void_p, C.c_void_p]),
    "mlxIpcClose": (None, [C.c_void_p, C.c_void_p]),
    "mlxDeviceAlloc": (C.c_void_p, [C.c_void_p, C.c_size_t]),
    "mlxDeviceFree": (None, [C.c_void_p, C.c_void_p]),
    "mlxIpcExportPointer": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "srtChunkLayout": (C.c_int, [C.POINTER(srt_render_params), C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "srtRenderDeviceSignal": (C.c_int, [C.c_void_p, C.POINTER(srt_camera), C.POINTER(srt_light),
                                        C.POINTER(srt_render_params), C.c_void_p, C.c_void_p, C.c_void_p,
                                        C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32]),
    "srtStreamWaitFlags": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]),
    "srtStreamWriteFlag": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32]),
    "srtFlagWaitTimeouts": (C.c_uint32, [C.c_void_p]),
    "srtGatherFrameHost": (C.c_int, [C.c_void_p, C.POINTER(srt_render_params), C.c_void_p, C.c_void_p, C.c_uint32,
                                     C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint32]),
    "srtCreateSceneGroup": (C.c_void_p, [C.c_void_p, C.c_uint32, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int,
                                         C.c_char_p, C.c_size_t]),
    "srtRenderGroup": (C.c_int, [C.c_void_p, C.POINTER(srt_camera), C.POINTER(srt_light),
                                 C.POINTER(srt_render_params), C.c_void_p, C.POINTER(srt_render_stats)]),
    "srtSceneGroupFrame": (C.c_void_p, [C.c_void_p]),
    "srtSceneGroupSize": (C.c_uint32, [C.c_void_p]),
    "srtGetSceneGroupError": (C.c_void_p, [C.c_void_p, C.c_char_p, C.c_size_t]),
    "srtReleaseSceneGroup": (None, [C.c_void_p]),
}

_lib = None


def declared_symbols():
    return sorted(_PROTOTYPES)


def library_path() -> Path:
    return _LIB_PATH


def load_library():
    """Loads libModelRunner.so; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not _LIB_PATH.exists():
        raise MLError(f"{_LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                      "(there is no CPU fallback)")
    lib = C.CDLL(str(_LIB_PATH), mode=getattr(os, "RTLD_LOCAL", 0) | getattr(os, "RTLD_NOW", 2))
    for name, (restype, argtypes) in _PROTOTYPES.items():
        fn = getattr(lib, name)
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def inspect_model(path, input_node=None, output_node=None):
    """mlxInspectModel: host-only parse; returns (ok, text)."""
    lib = load_library()
    buf = C.create_string_buffer(1 << 16)
    st = lib.mlxInspectModel(os.fsencode(path) if path is not None else None, input_node.encode() if input_node else None,
                             output_node.encode() if output_node else None, buf, len(buf))
    return st == ML_OK, buf.value.decode(errors="replace")


def _err(fn, handle) -> str:
    buf = C.create_string_buffer(1024)
    fn(handle, buf, len(buf))
    return buf.value.decode(errors="replace")


class Context:
    def __init__(self, device: int | None = None):
        """device=None: mlCreateContext() (MLB200_DEVICE, else the library runtime's current device);
        an int: mlxCreateContextOnDevice(device).  One process per GPU passes LOCAL_RANK: the
        library links its own CUDA runtime and does not see torch.cuda.set_device()."""
        self.lib = load_library()
        self.handle = self.lib.mlCreateContext() if device is None else self.lib.mlxCreateContextOnDevice(device)
        if not self.handle:
            raise MLError("mlCreateContext failed (no usable CUDA device?)")

    def error(self) -> str:
        return _err(self.lib.mlGetContextError, self.handle)

    def create_image(self, width, height, channels, dtype=ML_FLOAT32) -> "Image":
        info = ml_image_info(dtype, width, height, channels)
        h = self.lib.mlCreateImage(self.handle, C.byref(info))
        if not h:
            raise MLError(self.error())
        return Image(self.lib, h)

    def create_model(self, path, input_node=None, output_node=None) -> "Model":
        params = ml_model_params(os.fsencode(path),
                                 input_node.encode() if input_node else None,
                                 output_node.encode() if output_node else None)
        h = self.lib.mlCreateModel(self.handle, C.byref(params))
        if not h:
            raise MLError(self.error())
        return Model(self.lib, h)

    def create_scene(self, vertices: np.ndarray, albedo: np.ndarray | None = None, build_bvh=True) -> "Scene":
        v = np.ascontiguousarray(vertices, dtype=np.float32).reshape(-1, 9)
        a = None if albedo is None else np.ascontiguousarray(albedo, dtype=np.float32).reshape(-1, 3)
        if a is not None and len(a) != len(v):
            raise ValueError("albedo and vertices disagree on the triangle count")
        h = self.lib.srtCreateScene(self.handle, len(v), v.ctypes.data if len(v) else None,
                                    a.ctypes.data if a is not None and len(a) else None, int(build_bvh))
        if not h:
            raise MLError(self.error())
        return Scene(self.lib, h)

    def ipc_open(self, handle: bytes) -> int:
        """mlxIpcOpen: maps another process's exported image buffer; returns a device pointer."""
        buf = C.create_string_buffer(handle, 64)
        p = self.lib.mlxIpcOpen(self.handle, buf)
        if not p:
            raise MLError(self.error())
        return int(p)

    def ipc_close(self, ptr: int):
        self.lib.mlxIpcClose(self.handle, ptr)

    def device_alloc(self, nbytes: int) -> int:
        """mlxDeviceAlloc: zero-filled raw device memory (e.g. the flag array of a multi-GPU frame)."""
        p = self.lib.mlxDeviceAlloc(self.handle, nbytes)
        if not p:
            raise MLError(self.error())
        return int(p)

    def device_free(self, ptr: int):
        self.lib.mlxDeviceFree(self.handle, ptr)

    def ipc_export_pointer(self, ptr: int) -> bytes:
        buf = C.create_string_buffer(64)
        if self.lib.mlxIpcExportPointer(self.handle, ptr, buf) != ML_OK:
            raise MLError(self.error())
        return buf.raw

    def stream_wait_flags(self, d_flags: int, count: int, value: int, stride=1, stream=0, timeout_ms=0):
        """srtStreamWaitFlags: device-side wait (spin kernel) until `count` flags have reached `value`."""
        if self.lib.srtStreamWaitFlags(self.handle, stream or None, d_flags, count, stride, value, timeout_ms) != ML_OK:
            raise MLError(self.error())

    def stream_write_flag(self, d_flag: int, value: int, stream=0):
        if self.lib.srtStreamWriteFlag(self.handle, stream or None, d_flag, value) != ML_OK:
            raise MLError(self.error())

    def flag_wait_timeouts(self) -> int:
        return int(self.lib.srtFlagWaitTimeouts(self.handle))

    def gather_frame_host(self, params, d_rgb: int, d_flags: int, n_ranks: int, n_chunks: int, value: int,
                          rgb_host: np.ndarray, timeout_ms=0):
        """srtGatherFrameHost: rank 0's half of a host-buffer multi-GPU frame."""
        if self.lib.srtGatherFrameHost(self.handle, C.byref(params), d_rgb, d_flags, n_ranks, n_chunks, value,
                                       rgb_host.ctypes.data, timeout_ms) != ML_OK:
            raise MLError(self.error())

    def measure_fp32_peak(self, repeats=5) -> float:
        return float(self.lib.srtMeasureFp32Peak(self.handle, repeats))

    def release(self):
        if self.handle:
            self.lib.mlReleaseContext(self.handle)
            self.handle = None

    __del__ = release


class Image:
    def __init__(self, lib, handle):
        self.lib, self.handle = lib, handle

    def info(self) -> ml_image_info:
        info = ml_image_info()
        if self.lib.mlGetImageInfo(self.handle, C.byref(info)) != ML_OK:
            raise MLError("mlGetImageInfo failed")
        return info

    def map(self) -> np.ndarray:
        """mlMapImage: a numpy view of the image's (pinned) host memory, HWC."""
        size = C.c_size_t(0)
        ptr = self.lib.mlMapImage(self.handle, C.byref(size))
        if not ptr:
            raise MLError("mlMapImage failed")
        info = self.info()
        dt = np.float32 if info.dtype == ML_FLOAT32 else np.float16
        buf = (C.c_char * size.value).from_address(ptr)
        arr = np.frombuffer(buf, dtype=dt).reshape(info.height, info.width, info.channels)
        self._mapped = ptr
        return arr

    def unmap(self, ptr=None):
        if self.lib.mlUnmapImage(self.handle, ptr if ptr is not None else self._mapped) != ML_OK:
            raise MLError("mlUnmapImage failed")

    def device_pointer(self, for_write=False, stream=0) -> int:
        p = self.lib.mlxImageDevicePointer(self.handle, int(for_write), stream or None)
        if not p:
            raise MLError("mlxImageDevicePointer failed")
        return int(p)

    def ipc_export(self) -> bytes:
        """mlxIpcExportImage: 64-byte handle other ranks pass to Context.ipc_open()."""
        buf = C.create_string_buffer(64)
        if self.lib.mlxIpcExportImage(self.handle, buf) != ML_OK:
            raise MLError("mlxIpcExportImage failed")
        return buf.raw

    def write(self, data: np.ndarray):
        view = self.map()
        np.copyto(view, np.asarray(data).reshape(view.shape))
        self.unmap()

    def read(self) -> np.ndarray:
        view = self.map()
        out = view.copy()
        self.unmap()
        return out

    def release(self):
        if self.handle:
            self.lib.mlReleaseImage(self.handle)
            self.handle = None

    __del__ = release


class Model:
    def __init__(self, lib, handle):
        self.lib, self.handle = lib, handle

    def error(self) -> str:
        return _err(self.lib.mlGetModelError, self.handle)

    def info(self):
        i, o = ml_image_info(), ml_image_info()
        if self.lib.mlGetModelInfo(self.handle, C.byref(i), C.byref(o)) != ML_OK:
            raise MLError(self.error())
        return i, o

    def set_input_info(self, info: ml_image_info):
        if self.lib.mlSetModelInputInfo(self.handle, C.byref(info)) != ML_OK:
            raise MLError(self.error())

    def launch_count(self) -> int:
        return int(self.lib.mlxModelLaunchCount(self.handle))

    def set_precision(self, precision: int):
        """mlxSetModelPrecision: MLX_PRECISION_FP32 (0), MLX_PRECISION_BF16 (1, tcgen05 layers) or
        MLX_PRECISION_BF16X3 (2, tcgen05 layers with split operands: fp32-grade results)."""
        if self.lib.mlxSetModelPrecision(self.handle, int(precision)) != ML_OK:
            raise MLError(self.error())

    def tensor_core_layers(self) -> int:
        return int(self.lib.mlxModelTensorCoreLayers(self.handle))

    def infer(self, inp: Image, out: Image):
        if self.lib.mlInfer(self.handle, inp.handle, out.handle) != ML_OK:
            raise MLError(self.error())

    def profile(self, inp: Image, out: Image, max_ops=256):
        """mlxProfileModel: [(ms, algorithmic flops, algorithmic bytes)] per fused op, device-timed."""
        ms = (C.c_float * max_ops)(); fl = (C.c_double * max_ops)(); by = (C.c_double * max_ops)()
        n = C.c_size_t(0)
        if self.lib.mlxProfileModel(self.handle, inp.handle, out.handle, ms, fl, by, max_ops, C.byref(n)) != ML_OK:
            raise MLError(self.error())
        return [(float(ms[k]), float(fl[k]), float(by[k])) for k in range(min(n.value, max_ops))]

    def infer_batch(self, inputs, outputs):
        """mlxInferBatch: pipelined batch of independent inferences."""
        n = len(inputs)
        ia = (C.c_void_p * n)(*[i.handle for i in inputs])
        oa = (C.c_void_p * n)(*[o.handle for o in outputs])
        if self.lib.mlxInferBatch(self.handle, ia, oa, n) != ML_OK:
            raise MLError(self.error())

    def last_batch_ms(self) -> float:
        """mlxModelLastBatchMs: device span of the last infer_batch (CUDA events inside the library)."""
        return float(self.lib.mlxModelLastBatchMs(self.handle))

    def release(self):
        if self.handle:
            self.lib.mlReleaseModel(self.handle)
            self.handle = None

    __del__ = release


def make_params(width, height, spp=1, jitter=0, seed=0, shade=SRT_SHADE_ALBEDO, accel=SRT_ACCEL_BRUTE,
                y0=0, y1=None, buf_y0=0, t_min=1e-4, shadow_eps=1e-3, background=(0.0, 0.0, 0.0),
                collect_stats=0, row_block_stride=0, row_block_phase=0) -> srt_render_params:
    p = srt_render_params()
    p.width, p.height = width, height
    p.y0, p.y1 = y0, height if y1 is None else y1
    p.buf_y0 = buf_y0
    p.row_block_stride, p.row_block_phase = row_block_stride, row_block_phase
    p.spp, p.jitter, p.seed = spp, jitter, seed
    p.shade, p.accel, p.collect_stats = shade, accel, collect_stats
    p.t_min, p.shadow_eps = t_min, shadow_eps
    p.background[:] = background
    return p


class Scene:
    def __init__(self, lib, handle):
        self.lib, self.handle = lib, handle

    def error(self) -> str:
        return _err(self.lib.srtGetSceneError, self.handle)

    def info(self) -> srt_scene_info:
        info = srt_scene_info()
        if self.lib.srtGetSceneInfo(self.handle, C.byref(info)) != ML_OK:
            raise MLError("srtGetSceneInfo failed")
        return info

    def bvh(self):
        n = self.lib.srtGetSceneBvh(self.handle, None, 0, None, 0)
        nodes = np.zeros((n, 16), dtype=np.float32)
        order = np.zeros(int(self.info().n_triangles), dtype=np.int32)
        self.lib.srtGetSceneBvh(self.handle, nodes.ctypes.data, n, order.ctypes.data, len(order))
        return nodes, order

    def render(self, camera, light, params, want_rgb=True, want_aov=True, buffers=None):
        """srtRender: host buffers; returns (rgb, hit_id, hit_t, stats) for rows [buf_y0, y1).
        `buffers` = (rgb, hit_id, hit_t) numpy arrays to render into (e.g. views of page-locked memory,
        which srtRender fills straight from the device) instead of fresh pageable arrays."""
        rows = params.height - params.buf_y0
        if buffers is not None:
            rgb, hid, ht = buffers
            want_rgb, want_aov = rgb is not None, hid is not None and ht is not None
        else:
            rgb = np.zeros((rows, params.width, 3), dtype=np.float32) if want_rgb else None
            hid = np.full((rows, params.width), -2, dtype=np.int32) if want_aov else None
            ht = np.zeros((rows, params.width), dtype=np.float32) if want_aov else None
        stats = srt_render_stats()
        st = self.lib.srtRender(self.handle, C.byref(camera), C.byref(light), C.byref(params),
                                rgb.ctypes.data if want_rgb else None,
                                hid.ctypes.data if want_aov else None,
                                ht.ctypes.data if want_aov else None, C.byref(stats))
        if st != ML_OK:
            raise MLError(self.error())
        return rgb, hid, ht, stats

    def render_device(self, camera, light, params, d_rgb=0, d_hit_id=0, d_hit_t=0, stream=0):
        st = self.lib.srtRenderDevice(self.handle, C.byref(camera), C.byref(light), C.byref(params),
                                      d_rgb or None, d_hit_id or None, d_hit_t or None, stream or None)
        if st != ML_OK:
            raise MLError(self.error())

    def render_device_signal(self, camera, light, params, d_rgb, d_flags, value, chunk_stride=1, n_chunks=1,
                             stream=0):
        """srtRenderDeviceSignal: srtRenderDevice whose last block(s) publish completion flags."""
        st = self.lib.srtRenderDeviceSignal(self.handle, C.byref(camera), C.byref(light), C.byref(params),
                                            d_rgb or None, None, None, stream or None, d_flags, chunk_stride,
                                            n_chunks, value)
        if st != ML_OK:
            raise MLError(self.error())

    def render_device_no_alloc(self, camera, light, params, d_rgb=0, d_hit_id=0, d_hit_t=0, stream=0):
        """srtRenderDeviceNoAlloc: srtRenderDevice with its scratch in the library's static device array."""
        st = self.lib.srtRenderDeviceNoAlloc(self.handle, C.byref(camera), C.byref(light), C.byref(params),
                                             d_rgb or None, d_hit_id or None, d_hit_t or None, stream or None)
        if st != ML_OK:
            raise MLError(self.error())

    def stats(self) -> srt_render_stats:
        s = srt_render_stats()
        if self.lib.srtGetRenderStats(self.handle, C.byref(s)) != ML_OK:
            raise MLError(self.error())
        return s

    def render_to_image(self, camera, light, params, image: Image) -> srt_render_stats:
        s = srt_render_stats()
        st = self.lib.srtRenderToImage(self.handle, C.byref(camera), C.byref(light), C.byref(params),
                                       image.handle, C.byref(s))
        if st != ML_OK:
            raise MLError(self.error())
        return s

    def release(self):
        if self.handle:
            self.lib.srtReleaseScene(self.handle)
            self.handle = None

    __del__ = release


def chunk_layout(params, requested_chunks):
    rows, n = C.c_uint32(0), C.c_uint32(0)
    if load_library().srtChunkLayout(C.byref(params), requested_chunks, C.byref(rows), C.byref(n)) != ML_OK:
        raise MLError("srtChunkLayout failed")
    return rows.value, n.value


class SceneGroup:
    """srtCreateSceneGroup / srtRenderGroup: one process, N GPUs, one frame (peer stores + device flags)."""

    def __init__(self, devices, vertices: np.ndarray, albedo: np.ndarray | None = None, build_bvh=True):
        self.lib = load_library()
        v = np.ascontiguousarray(vertices, dtype=np.float32).reshape(-1, 9)
        a = None if albedo is None else np.ascontiguousarray(albedo, dtype=np.float32).reshape(-1, 3)
        dev = (C.c_int * len(devices))(*devices)
        err = C.create_string_buffer(1024)
        self.handle = self.lib.srtCreateSceneGroup(dev, len(devices), len(v), v.ctypes.data if len(v) else None,
                                                   a.ctypes.data if a is not None and len(a) else None,
                                                   int(build_bvh), err, len(err))
        if not self.handle:
            raise MLError(err.value.decode(errors="replace"))

    def error(self) -> str:
        return _err(self.lib.srtGetSceneGroupError, self.handle)

    def size(self) -> int:
        return int(self.lib.srtSceneGroupSize(self.handle))

    def frame_pointer(self) -> int:
        return int(self.lib.srtSceneGroupFrame(self.handle) or 0)

    def render(self, camera, light, params, rgb_host: np.ndarray | None = None, want_host=True):
        """Returns (rgb or None, stats).  want_host=False leaves the frame on GPU 0."""
        if want_host and rgb_host is None:
            rgb_host = np.zeros((params.height, params.width, 3), dtype=np.float32)
        stats = srt_render_stats()
        st = self.lib.srtRenderGroup(self.handle, C.byref(camera), C.byref(light), C.byref(params),
                                     rgb_host.ctypes.data if want_host else None, C.byref(stats))
        if st != ML_OK:
            raise MLError(self.error())
        return (rgb_host if want_host else None), stats

    def release(self):
        if self.handle:
            self.lib.srtReleaseSceneGroup(self.handle)
            self.handle = None

    __del__ = release


def gen_cornell_box(width, height):
    lib = load_library()
    n = lib.srtGenCornellBox(None, None, None, None, width, height)
    v = np.zeros((n, 9), dtype=np.float32)
    a = np.zeros((n, 3), dtype=np.float32)
    cam, light = srt_camera(), srt_light()
    lib.srtGenCornellBox(v.ctypes.data, a.ctypes.data, C.byref(cam), C.byref(light), width, height)
    return v, a, cam, light


def gen_mesh(seed, n_triangles, width, height):
    lib = load_library()
    v = np.zeros((n_triangles, 9), dtype=np.float32)
    a = np.zeros((n_triangles, 3), dtype=np.float32)
    cam, light = srt_camera(), srt_light()
    if lib.srtGenMesh(seed, n_triangles, v.ctypes.data, a.ctypes.data, C.byref(cam), C.byref(light),
                      width, height) != ML_OK:
        raise MLError("srtGenMesh failed")
    return v, a, cam, light
