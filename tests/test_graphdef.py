"""CPU: the GraphDef writer and the C++ wire-format reader, cross-checked
against the protobuf runtime (an independent encoder/decoder)."""
from pathlib import Path

import numpy as np
import pytest

import tf_protos
from simpleraytracer_b200 import api
from simpleraytracer_b200 import graphdef_writer as gw

P = tf_protos.build()
ROOT = Path(__file__).resolve().parent.parent


def test_writer_bytes_parse_with_protobuf_runtime():
    data, layers = gw.make_denoiser(channels=3, width=8, depth=3, seed=5)
    g = P["GraphDef"]()
    g.ParseFromString(data)
    ops = [n.op for n in g.node]
    assert ops[0] == "Placeholder" and ops[-1] == "Identity"
    assert ops.count("Conv2D") == 3 and ops.count("BiasAdd") == 3 and ops.count("Relu") == 2
    conv0 = next(n for n in g.node if n.name == "conv0")
    assert list(conv0.input) == ["input", "conv0/kernel"]
    assert conv0.attr["padding"].s == b"SAME"
    assert list(conv0.attr["strides"].list.i) == [1, 1, 1, 1]
    assert conv0.attr["T"].type == gw.DT_FLOAT
    shp = conv0.attr["_output_shapes"].list.shape[0]
    assert [d.size for d in shp.dim] == [-1, -1, -1, 8]
    k = next(n for n in g.node if n.name == "conv0/kernel").attr["value"].tensor
    assert [d.size for d in k.tensor_shape.dim] == [3, 3, 3, 8]
    assert np.array_equal(np.frombuffer(k.tensor_content, "<f4").reshape(3, 3, 3, 8), layers[0][0])
    # re-serialising with the runtime yields a graph the runtime parses identically
    g2 = P["GraphDef"](); g2.ParseFromString(g.SerializeToString())
    assert g2 == g


def test_reader_inspects_writer_graph(tmp_path, monkeypatch):
    monkeypatch.setenv("MLB200_PRECISION", "fp32")           # the plan of the plain CUDA-core path
    path = tmp_path / "m.pb"
    gw.make_denoiser(path, channels=3, width=16, depth=4, seed=1, act="LeakyRelu")
    ok, text = api.inspect_model(path)
    assert ok, text
    lines = text.splitlines()
    assert lines[0] == "input_node=input output_node=output"
    assert lines[1] == "input dtype=0 w=0 h=0 c=3"
    assert lines[2] == "output dtype=0 w=0 h=0 c=3"
    # Conv+BiasAdd+LeakyRelu fused: 4 launches; the residual add is folded into the last conv (Identity is free)
    assert lines[3] == "ops=4 launches=4"
    assert lines[4].startswith("Conv name=conv0 ") and "bias=yes act=leaky_relu alpha=0.1" in lines[4]
    assert "act=none" in lines[7] and "residual=t0" in lines[7] and len(lines) == 8


def runtime_graph(float_val=False, with_shapes=True):
    """A graph built by the protobuf runtime with encodings our writer never
    produces: float_val instead of tensor_content, device strings, versions,
    control inputs, ':0' suffixes."""
    rng = np.random.default_rng(3)
    g = P["GraphDef"]()
    g.versions.producer = 27

    def add(name, op, inputs=()):
        n = g.node.add(); n.name, n.op = name, op
        n.input.extend(inputs)
        n.device = "/device:CPU:0"
        return n

    x = add("x", "Placeholder")
    x.attr["dtype"].type = 1
    for d in (-1, 24, 32, 3):
        x.attr["shape"].shape.dim.add().size = d
    if with_shapes:
        s = x.attr["_output_shapes"].list.shape.add()
        for d in (-1, 24, 32, 3):
            s.dim.add().size = d
    w = rng.standard_normal((3, 3, 3, 4)).astype(np.float32)
    k = add("k", "Const"); k.attr["dtype"].type = 1
    t = k.attr["value"].tensor; t.dtype = 1
    for d in w.shape:
        t.tensor_shape.dim.add().size = d
    if float_val:
        t.float_val.extend(w.ravel().tolist())
    else:
        t.tensor_content = w.tobytes()
    c = add("c", "Conv2D", ["x:0", "k", "^x"])
    c.attr["T"].type = 1; c.attr["padding"].s = b"VALID"
    c.attr["strides"].list.i.extend([1, 2, 2, 1])
    b = add("b", "Const"); b.attr["dtype"].type = 1
    tb = b.attr["value"].tensor; tb.dtype = 1; tb.tensor_shape.dim.add().size = 4
    tb.float_val.append(0.5)                       # one value, broadcast to 4
    ba = add("ba", "BiasAdd", ["c", "b"]); ba.attr["T"].type = 1
    r = add("out", "Relu6", ["ba"]); r.attr["T"].type = 1
    return g.SerializeToString(), w


@pytest.mark.parametrize("float_val", [False, True])
def test_reader_handles_runtime_encodings(tmp_path, float_val):
    data, w = runtime_graph(float_val=float_val)
    path = tmp_path / "r.pb"; path.write_bytes(data)
    ok, text = api.inspect_model(path)
    assert ok, text
    lines = text.splitlines()
    assert lines[0] == "input_node=x output_node=out"
    assert lines[1] == "input dtype=0 w=32 h=24 c=3"
    # static H,W,C known -> output shape known without mlSetModelInputInfo: VALID 3x3 stride 2
    assert lines[2] == "output dtype=0 w=15 h=11 c=4"
    assert lines[3] == "ops=1 launches=1"
    assert "k=3x3 stride=2 dilation=1 padding=VALID" in lines[4] and "bias=yes act=relu6" in lines[4]
    wsum = float(lines[4].split("wsum=")[1].split()[0])
    assert abs(wsum - float(w.astype(np.float64).sum())) < 1e-4


def test_reader_node_selection_and_errors(tmp_path):
    data, _ = runtime_graph()
    path = tmp_path / "r.pb"; path.write_bytes(data)
    ok, text = api.inspect_model(path, None, "ba")        # explicit output node
    assert ok and "output_node=ba" in text and "act=none" in text
    ok, text = api.inspect_model(path, "x", "k")          # constant output
    assert not ok and "Unable to detect" not in text and "constant" in text
    ok, text = api.inspect_model(tmp_path / "missing.pb")
    assert not ok and text.startswith("Error reading graph definition: ")
    bad = tmp_path / "bad.pb"; bad.write_bytes(b"\x0a\xff\xff\xff")
    ok, text = api.inspect_model(bad)
    assert not ok and text.startswith("Error reading graph definition: ")
    # unsupported op is refused loudly, never skipped
    g = P["GraphDef"](); g.ParseFromString(data)
    g.node[-1].op = "Softmax"
    path2 = tmp_path / "u.pb"; path2.write_bytes(g.SerializeToString())
    ok, text = api.inspect_model(path2)
    assert not ok and "Unimplemented" in text and "Softmax" in text
    # node without dtype/T: reference message (model.cpp:32)
    g = P["GraphDef"](); g.ParseFromString(data)
    del g.node[-1].attr["T"]
    path3 = tmp_path / "d.pb"; path3.write_bytes(g.SerializeToString())
    ok, text = api.inspect_model(path3)
    assert not ok and text == "Unable to detect node dtype: out"
    ok, text = api.inspect_model(None)
    assert not ok and text == "Bad model_path model parameter value"


def test_half_graph_dtype(tmp_path):
    path = tmp_path / "h.pb"
    gw.make_denoiser(path, channels=3, width=8, depth=2, seed=2, dtype=gw.DT_HALF)
    ok, text = api.inspect_model(path)
    assert ok and "input dtype=1 " in text and "output dtype=1 " in text


def test_bf16_plan_is_decided_on_the_host(tmp_path, monkeypatch):
    """MLB200_PRECISION=bf16: which layers go to the tensor cores (and which tensors are stored as
    bf16) is host logic -- visible through mlxInspectModel without a GPU."""
    path = tmp_path / "m.pb"
    gw.make_denoiser(path, channels=3, width=64, depth=4, seed=1)
    monkeypatch.setenv("MLB200_PRECISION", "fp32")
    ok, text = api.inspect_model(path)
    assert ok and "tcgen05" not in text                      # fp32 CUDA-core kernels only
    monkeypatch.setenv("MLB200_PRECISION", "bf16")
    ok, text = api.inspect_model(path)
    assert ok, text
    convs = [l for l in text.splitlines() if l.startswith("Conv ")]
    assert len(convs) == 4 and all("engine=tcgen05_bf16" in l for l in convs)
    assert all("out=bf16" in l for l in convs[:3]) and "out=bf16" not in convs[3]   # last layer writes fp32
    assert "launches=6" in text                              # + the pad-to-16-channels pass + the residual add (its own pass after a tensor-core layer)
    # a 5x5 net has no tensor-core layer; a 128->128 layer does not fit in shared memory
    gw.make_denoiser(path, channels=3, width=32, depth=3, seed=1, kernel=5)
    ok, text = api.inspect_model(path)
    assert ok and "tcgen05" not in text
    gw.make_denoiser(path, channels=3, width=128, depth=3, seed=1)
    ok, text = api.inspect_model(path)
    convs = [l for l in text.splitlines() if l.startswith("Conv ")]
    assert ["engine=tcgen05_bf16" in l for l in convs] == [True, False, True]
    assert "out=bf16" in convs[1]                            # the fp32 layer feeds a tensor-core layer in bf16


def test_bf16x3_plan_is_the_default(tmp_path, monkeypatch):
    """Default precision: fp32-grade tensor-core layers (split bf16 operands).  Tensors between two such
    layers are written as [hi | lo] halves by the producer; the first layer's input is split by a pass
    of its own; the last layer writes fp32 and its residual add is a separate pass."""
    monkeypatch.delenv("MLB200_PRECISION", raising=False)
    path = tmp_path / "m.pb"
    gw.make_denoiser(path, channels=3, width=64, depth=4, seed=1)
    ok, text = api.inspect_model(path)
    assert ok, text
    convs = [l for l in text.splitlines() if l.startswith("Conv ")]
    assert len(convs) == 4 and all("engine=tcgen05_bf16x3" in l for l in convs)
    assert all("out=bf16_hi_lo" in l for l in convs[:3]) and "out=bf16_hi_lo" not in convs[3]
    assert "launches=6" in text                              # 4 convs + the input split + the residual add
    monkeypatch.setenv("MLB200_PRECISION", "bf16x3")
    assert api.inspect_model(path)[1] == text
    # 5x5: no tensor-core layer; 128->128: the hi and lo filters do not fit in shared memory
    gw.make_denoiser(path, channels=3, width=32, depth=3, seed=1, kernel=5)
    ok, text = api.inspect_model(path)
    assert ok and "tcgen05" not in text
    gw.make_denoiser(path, channels=3, width=128, depth=3, seed=1)
    ok, text = api.inspect_model(path)
    convs = [l for l in text.splitlines() if l.startswith("Conv ")]
    assert ["engine=tcgen05_bf16x3" in l for l in convs] == [True, False, True]
    assert "out=bf16_hi_lo" not in text                      # fp32 either side of the CUDA-core layer


def test_add_is_folded_into_the_conv_only_when_the_order_allows(tmp_path):
    """Add(conv, r) -> conv epilogue (host logic, visible through mlxInspectModel).  Not folded: an
    activation on the sum, a conv result with a second reader, a residual that is computed later."""
    rng = np.random.default_rng(0)
    k = lambda ci, co: (rng.standard_normal((3, 3, ci, co)) * 0.1).astype(np.float32)

    def build(fn):
        g = gw.GraphBuilder()
        x = g.placeholder("input", (-1, -1, -1, 4))
        out = fn(g, x)
        g.unary("Identity", out, name="output", out_shape=(-1, -1, -1, 4))
        path = tmp_path / "g.pb"
        path.write_bytes(g.finish())
        ok, text = api.inspect_model(path)
        assert ok, text
        return [l for l in text.splitlines()[4:]]

    # relu(conv(x) + x): the activation belongs to the sum -> the Add stays (with the Relu folded into it)
    ops = build(lambda g, x: g.unary("Relu", g.add(g.conv2d(x, k(4, 4), name="c", out_shape=(-1, -1, -1, 4)), x,
                                                   out_shape=(-1, -1, -1, 4)), out_shape=(-1, -1, -1, 4)))
    assert [o.split()[0] for o in ops] == ["Conv", "Add"] and "act=relu" in ops[1] and "residual" not in ops[0]

    # t = conv(x); u = conv(t); t + u: t has two readers, u is folded with t as its residual
    def two(g, x):
        t = g.conv2d(x, k(4, 4), name="t", out_shape=(-1, -1, -1, 4))
        u = g.conv2d(t, k(4, 4), name="u", out_shape=(-1, -1, -1, 4))
        return g.add(t, u, out_shape=(-1, -1, -1, 4))
    ops = build(two)
    assert [o.split()[0] for o in ops] == ["Conv", "Conv"] and "residual" not in ops[0] and "residual=t" in ops[1]

    # a = conv(x); b = conv(x); a + b: only the LATER conv can take the other as a residual
    def par(g, x):
        a = g.conv2d(x, k(4, 4), name="a", out_shape=(-1, -1, -1, 4))
        b = g.conv2d(x, k(4, 4), name="b", out_shape=(-1, -1, -1, 4))
        return g.add(a, b, out_shape=(-1, -1, -1, 4))
    ops = build(par)
    assert [o.split()[0] for o in ops] == ["Conv", "Conv"]
    assert "residual" not in ops[0] and "name=b " in ops[1] and "residual=t" in ops[1]


def _one_op_graph(tmp_path, build, shape=(1, 64, 48, 3)):
    g = gw.GraphBuilder()
    x = g.placeholder("input", shape=shape)
    build(g, x)
    path = tmp_path / "one.pb"
    path.write_bytes(g.finish())
    return api.inspect_model(path)


def test_untrusted_sizes_cannot_wrap_the_plan(tmp_path):
    """A graph is untrusted input (the reference hands it to TensorFlow, model.cpp:87).  Sizes that would wrap the plan's
    shape arithmetic or exceed what the kernels can index (2^31 elements, 2^20 rows or columns per tensor) are refused
    when the graph is read, with a message; ordinary values next to them still pass.  Host only: no device involved."""
    ok, text = _one_op_graph(tmp_path, lambda g, x: g.resize(x, size=(128, 96), bilinear=True, name="out"))
    assert ok and "output dtype=0 w=96 h=128 c=3" in text
    for size in ((2**31 - 1, 2**31 - 1), (65536, 65536), (2**20, 4096)):
        ok, text = _one_op_graph(tmp_path, lambda g, x: g.resize(x, size=size, bilinear=True, name="out"))
        assert not ok and ("beyond" in text or "exceeds the largest supported tensor" in text), text
    ok, text = _one_op_graph(tmp_path, lambda g, x: g.resize(x, factor=4, name="out"))
    assert ok and "output dtype=0 w=192 h=256 c=3" in text
    for factor in (65536, 2**24, 2**31 - 1):
        ok, text = _one_op_graph(tmp_path, lambda g, x: g.resize(x, factor=factor, name="out"))
        assert not ok and ("beyond" in text or "exceeds the largest supported tensor" in text), text
    w = np.zeros((3, 3, 3, 4), np.float32)
    for kw in (dict(stride=2**31 - 1), dict(dilation=2**31 - 1), dict(stride=65), dict(dilation=65)):
        ok, text = _one_op_graph(tmp_path, lambda g, x: g.conv2d(x, w, name="out", **kw))
        assert not ok and "Unimplemented" in text and ("strides" in text or "dilations" in text), text
    ok, text = _one_op_graph(tmp_path, lambda g, x: g.conv2d(x, w, stride=64, name="out"))
    assert ok and "output dtype=0 w=1 h=1 c=4" in text
    ok, text = _one_op_graph(tmp_path, lambda g, x: g.pool(x, "max", 2, 2**31 - 1, name="out"))
    assert not ok and "strides" in text
    for pad in (2**31 - 1, 2**20, -1):
        ok, text = _one_op_graph(tmp_path, lambda g, x: g.pad(x, pad, pad, pad, pad, name="out"))
        assert not ok and "padding amount" in text
    # a frame the kernels cannot index is refused at the input already
    ok, text = _one_op_graph(tmp_path, lambda g, x: g.unary("Relu", x, name="out"), shape=(1, 2**16, 2**16, 3))
    assert not ok and "exceeds the largest supported tensor" in text


def test_reader_survives_damaged_files(tmp_path):
    """Truncated, bit-flipped, spliced and cut copies of a valid U-Net graph: the reader either builds a plan or returns
    a message -- it never crashes, hangs or reports success without a plan.  Runs in a child process so that a crash is a
    test failure rather than the end of the test run.  Host only."""
    import subprocess
    import sys
    import textwrap
    base = tmp_path / "u.pb"
    gw.make_unet(base, channels=3, width=8, levels=2, seed=1, up="bilinear", down="max", batch_norm=True, act="Relu")
    ok, text = api.inspect_model(base)
    assert ok, text
    script = textwrap.dedent(f"""
        import sys
        import numpy as np
        sys.path.insert(0, {str(ROOT)!r})
        from simpleraytracer_b200 import api
        data = open({str(base)!r}, "rb").read()
        rng = np.random.default_rng(7)
        n_ok = n_err = 0
        for it in range(600):
            b = bytearray(data)
            mode = it % 4
            if mode == 0:
                b = b[:rng.integers(0, len(b))]
            elif mode == 1:
                for _ in range(rng.integers(1, 4)):
                    b[rng.integers(0, len(b))] = rng.integers(0, 256)
            elif mode == 2:
                i = rng.integers(0, len(b)); b[i:i] = bytes(rng.integers(0, 256, rng.integers(1, 9)).tolist())
            else:
                i = rng.integers(0, len(b)); del b[i:min(len(b), i + rng.integers(1, 64))]
            open({str(tmp_path / "f.pb")!r}, "wb").write(bytes(b))
            ok, text = api.inspect_model({str(tmp_path / "f.pb")!r})
            assert text, "neither a plan nor a message"
            n_ok += ok; n_err += not ok
        print("ok", n_ok, "err", n_err)
    """)
    r = subprocess.run([sys.executable, "-c", script], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.returncode, r.stdout[-500:], r.stderr[-1500:])
    n_ok, n_err = int(r.stdout.split()[1]), int(r.stdout.split()[3])
    assert n_ok + n_err == 600 and n_err > 100          # most damage is detected; some flips only touch weights
