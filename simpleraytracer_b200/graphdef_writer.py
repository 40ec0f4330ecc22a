"""Writes frozen TensorFlow GraphDef files (binary protobuf) without TensorFlow
or protoc: a minimal wire-format encoder for the message subset the model
runner reads (GraphDef.node, NodeDef{name,op,input,attr}, AttrValue, TensorProto,
TensorShapeProto).  Used to create synthetic, seeded, random-init models for the
tests, the bench and `test_app -m`: the reference ships no model file
(model_runner/README.md:158 only names color_only_denoiser.pb).

tests/test_graphdef.py cross-checks these bytes against the protobuf runtime.
"""
from __future__ import annotations

import struct

import numpy as np

DT_FLOAT, DT_INT32, DT_HALF = 1, 3, 19


def _varint(v: int) -> bytes:
    if v < 0:
        v += 1 << 64
    out = bytearray()
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _tag(field: int, wire: int) -> bytes:
    return _varint((field << 3) | wire)


def _ld(field: int, payload: bytes) -> bytes:
    return _tag(field, 2) + _varint(len(payload)) + payload


def _vi(field: int, value: int) -> bytes:
    return _tag(field, 0) + _varint(value)


def shape_proto(dims) -> bytes:
    out = b""
    for d in dims:
        out += _ld(2, _vi(1, int(d)))  # Dim{size}
    return out


def tensor_proto(array: np.ndarray, dtype: int) -> bytes:
    np_dtype = {DT_FLOAT: "<f4", DT_INT32: "<i4", DT_HALF: "<f2"}[dtype]
    a = np.ascontiguousarray(array, dtype=np_dtype)
    return _vi(1, dtype) + _ld(2, shape_proto(a.shape)) + _ld(4, a.tobytes())


def attr_type(dtype: int) -> bytes:
    return _vi(6, dtype)


def attr_shape(dims) -> bytes:
    return _ld(7, shape_proto(dims))


def attr_tensor(array, dtype) -> bytes:
    return _ld(8, tensor_proto(array, dtype))


def attr_s(text: str) -> bytes:
    return _ld(2, text.encode())


def attr_f(value: float) -> bytes:
    return _tag(4, 5) + struct.pack("<f", value)


def attr_b(value: bool) -> bytes:
    return _vi(5, 1 if value else 0)


def attr_i(value: int) -> bytes:
    return _vi(3, int(value))


def attr_list_i(values) -> bytes:
    packed = b"".join(_varint(int(v)) for v in values)
    return _ld(1, _ld(3, packed))


def attr_list_shape(shapes) -> bytes:
    return _ld(1, b"".join(_ld(7, shape_proto(s)) for s in shapes))


def node(name: str, op: str, inputs=(), attrs: dict | None = None) -> bytes:
    out = _ld(1, name.encode()) + _ld(2, op.encode())
    for i in inputs:
        out += _ld(3, i.encode())
    for key in sorted(attrs or {}):
        entry = _ld(1, key.encode()) + _ld(2, attrs[key])
        out += _ld(5, entry)
    return out


class GraphBuilder:
    """Accumulates NodeDefs; `finish()` returns the serialized GraphDef."""

    def __init__(self, dtype=DT_FLOAT, output_shapes=True):
        self.nodes: list[bytes] = []
        self.dtype = dtype
        self.output_shapes = output_shapes
        self._n = 0

    def _name(self, base):
        self._n += 1
        return f"{base}_{self._n}"

    def _shapes(self, attrs, shape):
        if self.output_shapes and shape is not None:
            attrs["_output_shapes"] = attr_list_shape([shape])
        return attrs

    def placeholder(self, name="input", shape=(-1, -1, -1, 3)):
        attrs = {"dtype": attr_type(self.dtype), "shape": attr_shape(shape)}
        self.nodes.append(node(name, "Placeholder", (), self._shapes(attrs, shape)))
        return name

    def const(self, array, name=None, dtype=None):
        name = name or self._name("const")
        dtype = dtype or self.dtype
        self.nodes.append(node(name, "Const", (), {"dtype": attr_type(dtype), "value": attr_tensor(array, dtype)}))
        return name

    def conv2d(self, x, weights_hwio, stride=1, padding="SAME", dilation=1, name=None, out_shape=None):
        name = name or self._name("conv")
        w = self.const(weights_hwio, name + "/kernel")
        attrs = {"T": attr_type(self.dtype), "strides": attr_list_i([1, stride, stride, 1]),
                 "padding": attr_s(padding), "data_format": attr_s("NHWC"),
                 "dilations": attr_list_i([1, dilation, dilation, 1])}
        self.nodes.append(node(name, "Conv2D", (x, w), self._shapes(attrs, out_shape)))
        return name

    def bias_add(self, x, bias, name=None, out_shape=None):
        name = name or self._name("bias")
        b = self.const(bias, name + "/bias")
        attrs = {"T": attr_type(self.dtype), "data_format": attr_s("NHWC")}
        self.nodes.append(node(name, "BiasAdd", (x, b), self._shapes(attrs, out_shape)))
        return name

    def unary(self, op, x, name=None, out_shape=None, alpha=None):
        name = name or self._name(op.lower())
        attrs = {"T": attr_type(self.dtype)}
        if alpha is not None:
            attrs["alpha"] = attr_f(alpha)
        self.nodes.append(node(name, op, (x,), self._shapes(attrs, out_shape)))
        return name

    def add(self, a, b, name=None, out_shape=None, op="AddV2"):
        name = name or self._name("add")
        self.nodes.append(node(name, op, (a, b), self._shapes({"T": attr_type(self.dtype)}, out_shape)))
        return name

    def concat(self, xs, name=None, out_shape=None):
        name = name or self._name("concat")
        axis = self.const(np.array(3, dtype=np.int32), name + "/axis", dtype=DT_INT32)
        attrs = {"T": attr_type(self.dtype), "N": _vi(3, len(xs)), "Tidx": attr_type(DT_INT32)}
        self.nodes.append(node(name, "ConcatV2", tuple(xs) + (axis,), self._shapes(attrs, out_shape)))
        return name

    # ---- the ops around the convolutions of an encoder / decoder graph ----------------------------------
    def pool(self, x, kind="max", k=2, stride=2, padding="SAME", name=None, out_shape=None):
        name = name or self._name(kind + "pool")
        attrs = {"T": attr_type(self.dtype), "ksize": attr_list_i([1, k, k, 1]), "strides": attr_list_i([1, stride, stride, 1]),
                 "padding": attr_s(padding), "data_format": attr_s("NHWC")}
        self.nodes.append(node(name, "MaxPool" if kind == "max" else "AvgPool", (x,), self._shapes(attrs, out_shape)))
        return name

    def resize(self, x, size=None, factor=None, bilinear=False, align_corners=False, half_pixel_centers=False,
               name=None, out_shape=None):
        """size=(h, w): a constant `size` input.  factor=f: the subgraph Keras UpSampling2D exports,
        size = StridedSlice(Shape(x), [1], [3], [1]) * [f, f]."""
        name = name or self._name("resize")
        if factor is not None:
            shp = name + "/Shape"
            self.nodes.append(node(shp, "Shape", (x,), {"T": attr_type(self.dtype), "out_type": attr_type(DT_INT32)}))
            b = self.const(np.array([1], np.int32), name + "/begin", DT_INT32)
            e = self.const(np.array([3], np.int32), name + "/end", DT_INT32)
            st = self.const(np.array([1], np.int32), name + "/strides", DT_INT32)
            sl = name + "/strided_slice"
            masks = {m: attr_i(0) for m in ("begin_mask", "end_mask", "ellipsis_mask", "new_axis_mask", "shrink_axis_mask")}
            self.nodes.append(node(sl, "StridedSlice", (shp, b, e, st),
                                   {"T": attr_type(DT_INT32), "Index": attr_type(DT_INT32), **masks}))
            k = self.const(np.array([factor, factor], np.int32), name + "/factor", DT_INT32)
            sz = name + "/mul"
            self.nodes.append(node(sz, "Mul", (sl, k), {"T": attr_type(DT_INT32)}))
        else:
            sz = self.const(np.array(size, np.int32), name + "/size", DT_INT32)
        attrs = {"T": attr_type(self.dtype), "align_corners": attr_b(align_corners),
                 "half_pixel_centers": attr_b(half_pixel_centers)}
        self.nodes.append(node(name, "ResizeBilinear" if bilinear else "ResizeNearestNeighbor", (x, sz),
                               self._shapes(attrs, out_shape)))
        return name

    def depth_to_space(self, x, block, name=None, out_shape=None):
        name = name or self._name("d2s")
        attrs = {"T": attr_type(self.dtype), "block_size": attr_i(block), "data_format": attr_s("NHWC")}
        self.nodes.append(node(name, "DepthToSpace", (x,), self._shapes(attrs, out_shape)))
        return name

    def pad(self, x, top, bottom, left, right, mode="CONSTANT", name=None, out_shape=None):
        name = name or self._name("pad")
        p = self.const(np.array([[0, 0], [top, bottom], [left, right], [0, 0]], np.int32), name + "/paddings", DT_INT32)
        attrs = {"T": attr_type(self.dtype), "Tpaddings": attr_type(DT_INT32)}
        if mode == "CONSTANT":
            self.nodes.append(node(name, "Pad", (x, p), self._shapes(attrs, out_shape)))
        else:
            attrs["mode"] = attr_s(mode)
            self.nodes.append(node(name, "MirrorPad", (x, p), self._shapes(attrs, out_shape)))
        return name

    def batch_norm(self, x, scale, offset, mean, variance, epsilon=1e-3, name=None, out_shape=None, op="FusedBatchNormV3"):
        name = name or self._name("bn")
        ins = (x, self.const(scale, name + "/gamma"), self.const(offset, name + "/beta"),
               self.const(mean, name + "/moving_mean"), self.const(variance, name + "/moving_variance"))
        attrs = {"T": attr_type(self.dtype), "U": attr_type(DT_FLOAT), "epsilon": attr_f(epsilon),
                 "is_training": attr_b(False), "data_format": attr_s("NHWC")}
        self.nodes.append(node(name, op, ins, self._shapes(attrs, out_shape)))
        return name

    def with_const(self, op, x, value, name=None, out_shape=None, const_first=False):
        """Mul / Sub / RealDiv / AddV2 of a tensor with a scalar or per-channel constant."""
        name = name or self._name(op.lower())
        c = self.const(np.asarray(value, np.float32), name + "/y")
        ins = (c, x) if const_first else (x, c)
        self.nodes.append(node(name, op, ins, self._shapes({"T": attr_type(self.dtype)}, out_shape)))
        return name

    def binary(self, op, a, b, name=None, out_shape=None):
        name = name or self._name(op.lower())
        self.nodes.append(node(name, op, (a, b), self._shapes({"T": attr_type(self.dtype)}, out_shape)))
        return name

    def squeeze(self, x, name=None):
        name = name or self._name("squeeze")
        self.nodes.append(node(name, "Squeeze", (x,), {"T": attr_type(self.dtype), "squeeze_dims": attr_list_i([0])}))
        return name

    def reshape(self, x, dims, name=None):
        name = name or self._name("reshape")
        sh = self.const(np.array(dims, np.int32), name + "/shape", DT_INT32)
        self.nodes.append(node(name, "Reshape", (x, sh), {"T": attr_type(self.dtype), "Tshape": attr_type(DT_INT32)}))
        return name

    def finish(self) -> bytes:
        return b"".join(_ld(1, n) for n in self.nodes)


def make_denoiser(path=None, *, channels=3, width=32, depth=4, seed=0, dtype=DT_FLOAT, residual=True,
                  act="Relu", kernel=3):
    """A small image-to-image CNN in the spirit of the reference README's
    "color_only_denoiser": conv(k) C->width + act, (depth-2) x [conv width->width + act],
    conv width->C, optional global residual add.  Random-init (He) weights from `seed`.
    Returns (graphdef_bytes, layers) where layers lists (weights_hwio, bias, act) for oracles.
    """
    rng = np.random.default_rng(seed)
    g = GraphBuilder(dtype=dtype)
    x = g.placeholder("input", (-1, -1, -1, channels))
    layers = []
    cur, cin = x, channels
    for i in range(depth):
        cout = channels if i == depth - 1 else width
        w = (rng.standard_normal((kernel, kernel, cin, cout)) * np.sqrt(2.0 / (kernel * kernel * cin))).astype(np.float32)
        b = (rng.standard_normal(cout) * 0.05).astype(np.float32)
        last = i == depth - 1
        cur = g.conv2d(cur, w, name=f"conv{i}", out_shape=(-1, -1, -1, cout))
        cur = g.bias_add(cur, b, name=f"conv{i}/BiasAdd", out_shape=(-1, -1, -1, cout))
        if not last:
            cur = g.unary(act, cur, name=f"conv{i}/{act}", out_shape=(-1, -1, -1, cout),
                          alpha=0.1 if act == "LeakyRelu" else None)
        layers.append((w, b, None if last else act))
        cin = cout
    if residual:
        cur = g.add(x, cur, name="residual", out_shape=(-1, -1, -1, channels))
    cur = g.unary("Identity", cur, name="output", out_shape=(-1, -1, -1, channels))
    data = g.finish()
    if path is not None:
        with open(path, "wb") as f:
            f.write(data)
    return data, layers


def make_unet(path=None, *, channels=3, width=16, levels=2, seed=0, up="nearest", down="maxpool", batch_norm=True,
              act="Relu"):
    """An encoder / decoder image-to-image net with skip concatenations, of the kind the reference README's denoiser
    is (model_runner/README.md:158): per level  conv3x3 [+ FusedBatchNorm] + act -> skip, downsample (MaxPool / AvgPool /
    stride-2 conv);  bottleneck conv;  per level  upsample (ResizeNearestNeighbor or ResizeBilinear by Keras' Shape *
    factor pattern, or conv + DepthToSpace) -> concat(skip) -> conv3x3 + act;  1x1 head; the input is normalised first
    ((x - 0.5) * 2) and the result added back to the input (global residual).  H and W must be multiples of 2**levels.
    Returns (graphdef_bytes, program): the program is the op list tests/oracle_lib.run_program() interprets."""
    rng = np.random.default_rng(seed)
    g = GraphBuilder()
    x = g.placeholder("input", (-1, -1, -1, channels))
    prog = []

    def he(k, cin, cout):
        return (rng.standard_normal((k, k, cin, cout)) * np.sqrt(2.0 / (k * k * cin))).astype(np.float32)

    def conv_block(src, cin, cout, name, k=3, stride=1, use_act=True, bn=False):
        w, b = he(k, cin, cout), (rng.standard_normal(cout) * 0.05).astype(np.float32)
        cur = g.conv2d(src, w, stride=stride, name=name, out_shape=(-1, -1, -1, cout))
        cur = g.bias_add(cur, b, name=name + "/BiasAdd", out_shape=(-1, -1, -1, cout))
        prog.append(("conv", cur, [src], dict(w=w, b=b, stride=stride)))
        if bn:
            p = dict(scale=rng.uniform(0.5, 1.5, cout).astype(np.float32), offset=(rng.standard_normal(cout) * 0.1).astype(np.float32),
                     mean=(rng.standard_normal(cout) * 0.1).astype(np.float32), variance=rng.uniform(0.5, 1.5, cout).astype(np.float32),
                     epsilon=1e-3)
            nxt = g.batch_norm(cur, p["scale"], p["offset"], p["mean"], p["variance"], p["epsilon"], name=name + "/bn",
                               out_shape=(-1, -1, -1, cout))
            prog.append(("bn", nxt, [cur], p))
            cur = nxt
        if use_act:
            nxt = g.unary(act, cur, name=name + "/" + act, out_shape=(-1, -1, -1, cout), alpha=0.1 if act == "LeakyRelu" else None)
            prog.append(("act", nxt, [cur], dict(act=act, alpha=0.1)))
            cur = nxt
        return cur

    # input normalisation: Sub and Mul with constants
    cur = g.with_const("Sub", x, 0.5, name="norm/sub", out_shape=(-1, -1, -1, channels))
    prog.append(("subc", cur, [x], dict(c=0.5)))
    nxt = g.with_const("Mul", cur, 2.0, name="norm/mul", out_shape=(-1, -1, -1, channels))
    prog.append(("mulc", nxt, [cur], dict(c=2.0)))
    cur, cin = nxt, channels
    skips = []
    for lv in range(levels):
        cout = width << lv
        cur = conv_block(cur, cin, cout, f"enc{lv}", bn=batch_norm)
        skips.append((cur, cout))
        if down == "conv":
            cur = conv_block(cur, cout, cout, f"down{lv}", stride=2)
        else:
            nxt = g.pool(cur, "max" if down == "maxpool" else "avg", 2, 2, "SAME", name=f"down{lv}", out_shape=(-1, -1, -1, cout))
            prog.append(("pool", nxt, [cur], dict(k=2, stride=2, kind="max" if down == "maxpool" else "avg")))
            cur = nxt
        cin = cout
    cur = conv_block(cur, cin, width << levels, "mid")
    cin = width << levels
    for lv in reversed(range(levels)):
        skip, sc = skips[lv]
        if up == "d2s":
            cur = conv_block(cur, cin, sc * 4, f"up{lv}/expand", k=1, use_act=False)
            nxt = g.depth_to_space(cur, 2, name=f"up{lv}", out_shape=(-1, -1, -1, sc))
            prog.append(("d2s", nxt, [cur], dict(block=2)))
            cur, cin = nxt, sc
        else:
            bil = up == "bilinear"
            nxt = g.resize(cur, factor=2, bilinear=bil, half_pixel_centers=bil, name=f"up{lv}", out_shape=(-1, -1, -1, cin))
            prog.append(("resize", nxt, [cur], dict(factor=2, bilinear=bil, half_pixel_centers=bil)))
            cur = nxt
        nxt = g.concat([cur, skip], name=f"cat{lv}", out_shape=(-1, -1, -1, cin + sc))
        prog.append(("concat", nxt, [cur, skip], {}))
        cur = conv_block(nxt, cin + sc, sc, f"dec{lv}")
        cin = sc
    cur = conv_block(cur, cin, channels, "head", k=1, use_act=False)
    nxt = g.add(x, cur, name="residual", out_shape=(-1, -1, -1, channels))
    prog.append(("add", nxt, [x, cur], {}))
    out = g.unary("Identity", nxt, name="output", out_shape=(-1, -1, -1, channels))
    prog.append(("act", out, [nxt], dict(act=None)))
    data = g.finish()
    if path is not None:
        with open(path, "wb") as f:
            f.write(data)
    return data, prog
