"""ctypes binding of oracle/liboracle.so (the CPU oracle).  Test infrastructure:
imported only by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs."""
import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
_SO = ROOT / "oracle" / "liboracle.so"


class orc_camera(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("forward", C.c_float * 3), ("right", C.c_float * 3),
                ("up", C.c_float * 3), ("half_w", C.c_float), ("half_h", C.c_float)]


class orc_light(C.Structure):
    _fields_ = [("position", C.c_float * 3), ("intensity", C.c_float * 3)]


class orc_params(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("x0", C.c_uint32), ("x1", C.c_uint32),
                ("y0", C.c_uint32), ("y1", C.c_uint32), ("spp", C.c_uint32), ("jitter", C.c_uint32),
                ("seed", C.c_uint32), ("shade", C.c_uint32), ("t_min", C.c_float),
                ("shadow_eps", C.c_float), ("background", C.c_float * 3)]


_lib = None
_scenes = None


def scenes_lib():
    """oracle/libsrtscenes.so: the seeded scene generators compiled alone (no CUDA, no product library)."""
    global _scenes
    if _scenes is None:
        so = ROOT / "oracle" / "libsrtscenes.so"
        if not so.exists():
            subprocess.check_call(["make", "-C", str(ROOT / "oracle"), "libsrtscenes.so"])
        _scenes = C.CDLL(str(so))
        _scenes.orc_gen_mesh.restype = C.c_int
        _scenes.orc_gen_mesh.argtypes = [C.c_uint32, C.c_size_t, C.c_void_p, C.c_void_p, C.POINTER(orc_camera),
                                         C.POINTER(orc_light), C.c_uint32, C.c_uint32]
        _scenes.orc_gen_cornell_box.restype = C.c_size_t
        _scenes.orc_gen_cornell_box.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(orc_camera), C.POINTER(orc_light),
                                                C.c_uint32, C.c_uint32]
    return _scenes


def gen_mesh(seed, n_triangles, width, height):
    v = np.zeros((n_triangles, 9), np.float32)
    a = np.zeros((n_triangles, 3), np.float32)
    cam, light = orc_camera(), orc_light()
    if scenes_lib().orc_gen_mesh(seed, n_triangles, v.ctypes.data, a.ctypes.data, C.byref(cam), C.byref(light),
                                 width, height) != 0:
        raise RuntimeError("orc_gen_mesh failed")
    return v, a, cam, light


def gen_cornell_box(width, height):
    n = scenes_lib().orc_gen_cornell_box(None, None, None, None, width, height)
    v = np.zeros((n, 9), np.float32)
    a = np.zeros((n, 3), np.float32)
    cam, light = orc_camera(), orc_light()
    scenes_lib().orc_gen_cornell_box(v.ctypes.data, a.ctypes.data, C.byref(cam), C.byref(light), width, height)
    return v, a, cam, light


def last_setup_ms() -> float:
    """Host time the last render() spent building the oracle's BVH."""
    return float(lib().orc_last_setup_ms())


def lib():
    global _lib
    if _lib is None:
        if not _SO.exists():
            subprocess.check_call(["make", "-C", str(ROOT / "oracle"), "liboracle.so"])
        _lib = C.CDLL(str(_SO))
        _lib.orc_render.restype = C.c_int
        _lib.orc_render.argtypes = [C.c_uint64, C.c_void_p, C.c_void_p, C.POINTER(orc_camera),
                                    C.POINTER(orc_light), C.POINTER(orc_params), C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        _lib.orc_last_setup_ms.restype = C.c_double
        _lib.orc_last_setup_ms.argtypes = []
        _lib.orc_intersect.restype = C.c_int
        _lib.orc_intersect.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_float)]
        _lib.orc_conv2d.restype = C.c_int
        _lib.orc_conv2d.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int,
                                    C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float,
                                    C.c_void_p, C.c_int]
        _lib.orc_conv_geometry.restype = None
        _lib.orc_conv_geometry.argtypes = [C.c_int] * 5 + [C.POINTER(C.c_int), C.POINTER(C.c_int)]
        _lib.orc_pool2d.restype = None
        _lib.orc_pool2d.argtypes = [C.c_void_p] + [C.c_int] * 7 + [C.c_void_p]
        _lib.orc_resize.restype = None
        _lib.orc_resize.argtypes = [C.c_void_p] + [C.c_int] * 8 + [C.c_void_p]
        _lib.orc_depth_to_space.restype = None
        _lib.orc_depth_to_space.argtypes = [C.c_void_p] + [C.c_int] * 4 + [C.c_void_p]
        _lib.orc_pad2d.restype = None
        _lib.orc_pad2d.argtypes = [C.c_void_p] + [C.c_int] * 8 + [C.c_void_p]
        _lib.orc_add_bias_act.restype = None
        _lib.orc_add_bias_act.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_int,
                                          C.c_float, C.c_void_p]
    return _lib


def _copy_struct(dst, src):
    """Field-wise copy between the product's srt_* and the oracle's orc_* structs."""
    for name, _ in dst._fields_:
        if hasattr(src, name):
            v = getattr(src, name)
            if hasattr(v, "__len__"):
                getattr(dst, name)[:] = list(v)
            else:
                setattr(dst, name, v)
    return dst


def render(vertices, albedo, camera, light, params, crop=None, threads=0, use_bvh=False):
    """Oracle render of `crop` = (x0, x1, y0, y1) (default: rows [y0,y1) of params, all columns).
    `camera`/`light`/`params` may be the product's ctypes structs; they are copied field by field."""
    L = lib()
    cam = _copy_struct(orc_camera(), camera)
    lgt = _copy_struct(orc_light(), light)
    p = _copy_struct(orc_params(), params)
    if crop is None:
        crop = (0, params.width, params.y0, params.y1)
    p.x0, p.x1, p.y0, p.y1 = crop
    h, w = p.y1 - p.y0, p.x1 - p.x0
    v = np.ascontiguousarray(vertices, dtype=np.float32).reshape(-1, 9)
    a = None if albedo is None else np.ascontiguousarray(albedo, dtype=np.float32).reshape(-1, 3)
    rgb = np.zeros((h, w, 3), np.float32)
    hid = np.zeros((h, w), np.int32)
    ht = np.zeros((h, w), np.float32)
    counts = np.zeros(2, np.uint64)
    used = L.orc_render(len(v), v.ctypes.data, a.ctypes.data if a is not None else None, C.byref(cam),
                        C.byref(lgt), C.byref(p), rgb.ctypes.data, hid.ctypes.data, ht.ctypes.data,
                        counts.ctypes.data, threads, int(use_bvh))
    return rgb, hid, ht, counts, used


def intersect(o, d, tri):
    o = np.asarray(o, np.float32); d = np.asarray(d, np.float32); tri = np.asarray(tri, np.float32).reshape(9)
    t = C.c_float(0)
    inside = lib().orc_intersect(o.ctypes.data, d.ctypes.data, tri.ctypes.data, C.byref(t))
    return bool(inside), t.value


ACTS = {None: 0, "Relu": 1, "LeakyRelu": 2, "Relu6": 3, "Sigmoid": 4, "Tanh": 5}


def conv2d(x, w, bias=None, stride=1, dilation=1, padding="SAME", act=None, alpha=0.2, threads=0):
    """x: HWC float32, w: HWIO.  Returns (OH, OW, Cout)."""
    L = lib()
    x = np.ascontiguousarray(x, np.float32); w = np.ascontiguousarray(w, np.float32)
    H, W, Cin = x.shape
    KH, KW, _, Cout = w.shape
    oh, ow, pad = C.c_int(), C.c_int(), C.c_int()
    same = int(padding == "SAME")
    L.orc_conv_geometry(H, KH, stride, dilation, same, C.byref(oh), C.byref(pad))
    L.orc_conv_geometry(W, KW, stride, dilation, same, C.byref(ow), C.byref(pad))
    out = np.zeros((oh.value, ow.value, Cout), np.float32)
    b = None if bias is None else np.ascontiguousarray(bias, np.float32)
    L.orc_conv2d(x.ctypes.data, H, W, Cin, w.ctypes.data, KH, KW, Cout, b.ctypes.data if b is not None else None,
                 stride, dilation, same, ACTS[act], alpha, out.ctypes.data, threads)
    return out


def run_denoiser(x, layers, residual=True, alpha=0.1, threads=0):
    """Oracle forward pass of graphdef_writer.make_denoiser()."""
    cur = np.ascontiguousarray(x, np.float32)
    for w, b, act in layers:
        cur = conv2d(cur, w, b, act=act, alpha=alpha, threads=threads)
    if residual:
        cur = (np.asarray(x, np.float32) + cur).astype(np.float32)
    return cur


def _geometry(n, k, stride, same):
    o, p = C.c_int(), C.c_int()
    lib().orc_conv_geometry(n, k, stride, 1, int(same), C.byref(o), C.byref(p))
    return o.value


def pool2d(x, k, stride, padding="SAME", kind="max"):
    x = np.ascontiguousarray(x, np.float32)
    H, W, Ch = x.shape
    same = padding == "SAME"
    out = np.zeros((_geometry(H, k, stride, same), _geometry(W, k, stride, same), Ch), np.float32)
    lib().orc_pool2d(x.ctypes.data, H, W, Ch, k, stride, int(same), int(kind == "max"), out.ctypes.data)
    return out


def resize(x, out_h, out_w, bilinear=False, align_corners=False, half_pixel_centers=False):
    x = np.ascontiguousarray(x, np.float32)
    H, W, Ch = x.shape
    out = np.zeros((out_h, out_w, Ch), np.float32)
    lib().orc_resize(x.ctypes.data, H, W, Ch, out_h, out_w, int(bilinear), int(align_corners), int(half_pixel_centers),
                     out.ctypes.data)
    return out


def depth_to_space(x, block):
    x = np.ascontiguousarray(x, np.float32)
    H, W, Ch = x.shape
    out = np.zeros((H * block, W * block, Ch // (block * block)), np.float32)
    lib().orc_depth_to_space(x.ctypes.data, H, W, Ch, block, out.ctypes.data)
    return out


def pad2d(x, top, bottom, left, right, mode="CONSTANT"):
    x = np.ascontiguousarray(x, np.float32)
    H, W, Ch = x.shape
    out = np.zeros((H + top + bottom, W + left + right, Ch), np.float32)
    lib().orc_pad2d(x.ctypes.data, H, W, Ch, top, bottom, left, right, {"CONSTANT": 0, "REFLECT": 1, "SYMMETRIC": 2}[mode],
                    out.ctypes.data)
    return out


def batch_norm(x, scale, offset, mean, variance, epsilon):
    """FusedBatchNorm, inference form, in fp32 like TensorFlow's CPU kernel: (x - mean) * rsqrt(var + eps) * scale + offset."""
    x = np.asarray(x, np.float32)
    inv = (np.float32(1.0) / np.sqrt(np.asarray(variance, np.float32) + np.float32(epsilon))).astype(np.float32)
    return ((x - np.asarray(mean, np.float32)) * (inv * np.asarray(scale, np.float32)) + np.asarray(offset, np.float32)).astype(np.float32)


def run_program(x, program):
    """Oracle interpreter of the op lists graphdef_writer.make_unet() returns: [(op, out, inputs, params)]."""
    env = {"input": np.ascontiguousarray(x, np.float32)}
    out = None
    for op, name, ins, kw in program:
        a = [env[i] for i in ins]
        if op == "conv":
            r = conv2d(a[0], kw["w"], kw.get("b"), stride=kw.get("stride", 1), padding=kw.get("padding", "SAME"),
                       act=kw.get("act"), alpha=kw.get("alpha", 0.2))
        elif op == "bn":
            r = batch_norm(a[0], kw["scale"], kw["offset"], kw["mean"], kw["variance"], kw["epsilon"])
        elif op == "act":
            r = lib_act(a[0], kw["act"], kw.get("alpha", 0.2))
        elif op == "pool":
            r = pool2d(a[0], kw["k"], kw["stride"], kw.get("padding", "SAME"), kw["kind"])
        elif op == "resize":
            h, w = (a[0].shape[0] * kw["factor"], a[0].shape[1] * kw["factor"]) if "factor" in kw else kw["size"]
            r = resize(a[0], h, w, kw.get("bilinear", False), kw.get("align_corners", False), kw.get("half_pixel_centers", False))
        elif op == "d2s":
            r = depth_to_space(a[0], kw["block"])
        elif op == "pad":
            r = pad2d(a[0], *kw["pads"], mode=kw.get("mode", "CONSTANT"))
        elif op == "concat":
            r = np.concatenate(a, axis=2)
        elif op == "add":
            r = (a[0] + a[1]).astype(np.float32)
        elif op == "sub":
            r = (a[0] - a[1]).astype(np.float32)
        elif op == "mul":
            r = (a[0] * a[1]).astype(np.float32)
        elif op == "mulc":
            r = (a[0] * np.asarray(kw["c"], np.float32)).astype(np.float32)
        elif op == "subc":
            r = (a[0] - np.asarray(kw["c"], np.float32)).astype(np.float32)
        elif op == "divc":
            r = (a[0] / np.asarray(kw["c"], np.float32)).astype(np.float32)
        else:
            raise ValueError(op)
        env[name] = r
        out = r
    return out


def lib_act(x, act, alpha=0.2):
    x = np.ascontiguousarray(x, np.float32)
    out = np.zeros_like(x)
    lib().orc_add_bias_act(x.ctypes.data, None, None, x.size, 1, ACTS[act], alpha, out.ctypes.data)
    return out
