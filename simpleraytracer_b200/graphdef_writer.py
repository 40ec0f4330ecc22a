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
