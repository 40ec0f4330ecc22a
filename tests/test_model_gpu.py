"""GPU tests of the model_runner path: API conformance (every row of SURVEY.md
section 8b) and inference parity against the conv oracle.  Tolerance from
BASELINE.json north_star: 1e-3 relative for fp32."""
import ctypes as C
import subprocess
from pathlib import Path

import numpy as np
import pytest

import oracle_lib as O
from simpleraytracer_b200 import api
from simpleraytracer_b200 import graphdef_writer as gw

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent
GOLD = ROOT / "tests" / "golden"
RTOL = 1e-3


def close(got, ref, what=""):
    """Two metrics, both asserted: max error over the output RANGE, and -- north_star's '1e-3 relative' read
    element-wise -- relative error over the elements that carry signal (|ref| > 5 % of the range; below that
    a relative error measures cancellation noise, not the kernel)."""
    scale = max(float(np.abs(ref).max()), 1e-6)
    err = float(np.abs(got - ref).max()) / scale
    assert err <= RTOL, f"{what}: max error {err:.3e} of the output range"
    m = np.abs(ref) > 5e-2 * scale
    if m.any():
        rel = float((np.abs(got - ref)[m] / np.abs(ref)[m]).max())
        assert rel <= RTOL, f"{what}: element-relative error {rel:.3e}"


def prepared_model(ctx, path, w, h):
    model = ctx.create_model(path)
    i, o = model.info()
    i.width, i.height = w, h
    model.set_input_info(i)
    i, o = model.info()
    return model, i, o


def test_golden_network(ctx, tmp_path):
    g = np.load(GOLD / "convnet_16.npz")
    path = tmp_path / "net.pb"; path.write_bytes(g["graphdef"].tobytes())
    model, i, o = prepared_model(ctx, path, 16, 16)
    assert (o.width, o.height, o.channels) == (16, 16, 3)
    inp = ctx.create_image(16, 16, 3); out = ctx.create_image(16, 16, 3)
    inp.write(g["x"])
    model.infer(inp, out)
    close(out.read(), g["y"], "golden convnet")
    for x in (inp, out, model):
        x.release()


@pytest.mark.parametrize("cfg", [
    dict(width=32, depth=4, act="Relu", kernel=3, w=97, h=61),
    dict(width=64, depth=3, act="LeakyRelu", kernel=3, w=64, h=40),
    dict(width=16, depth=3, act="Relu", kernel=5, w=50, h=33),
    dict(width=8, depth=2, act="Tanh", kernel=1, w=19, h=23),
])
def test_denoiser_matches_oracle(ctx, tmp_path, cfg):
    path = tmp_path / "d.pb"
    _, layers = gw.make_denoiser(path, channels=3, width=cfg["width"], depth=cfg["depth"], seed=4,
                                 act=cfg["act"], kernel=cfg["kernel"])
    w, h = cfg["w"], cfg["h"]
    model, i, o = prepared_model(ctx, path, w, h)
    x = np.random.default_rng(2).random((h, w, 3), dtype=np.float32)
    inp = ctx.create_image(w, h, 3); out = ctx.create_image(o.width, o.height, o.channels)
    inp.write(x)
    model.infer(inp, out)                             # default precision: bf16x3 tensor-core layers where supported
    ref = O.run_denoiser(x, layers, alpha=0.1)
    close(out.read(), ref, str(cfg))
    model.set_precision(api.MLX_PRECISION_FP32)       # the plain CUDA-core path
    model.infer(inp, out)
    close(out.read(), ref, str(cfg) + " fp32")
    assert model.launch_count() == cfg["depth"]       # conv+bias+act per layer; the residual add rides in the last one
    for obj in (inp, out, model):
        obj.release()


def single_conv_graph(path, cin, cout, k, stride, padding, dilation, act, seed):
    rng = np.random.default_rng(seed)
    w = (rng.standard_normal((k, k, cin, cout)) * 0.3).astype(np.float32)
    b = rng.standard_normal(cout).astype(np.float32)
    g = gw.GraphBuilder()
    x = g.placeholder("in", (-1, -1, -1, cin))
    y = g.conv2d(x, w, stride=stride, padding=padding, dilation=dilation, name="c")
    y = g.bias_add(y, b, name="b")
    if act:
        y = g.unary(act, y, name="a", alpha=0.3 if act == "LeakyRelu" else None)
    Path(path).write_bytes(g.finish())
    return w, b


@pytest.mark.parametrize("case", [
    (5, 7, 3, 1, "VALID", 1, "Relu6", 41, 29),
    (4, 6, 3, 2, "SAME", 1, "Sigmoid", 40, 31),
    (4, 6, 3, 2, "VALID", 1, None, 41, 30),
    (3, 4, 3, 1, "SAME", 2, "LeakyRelu", 30, 30),
    (8, 70, 3, 1, "SAME", 1, "Relu", 33, 17),     # Cout > 64: two channel tiles
    (70, 5, 3, 1, "SAME", 1, None, 20, 18),       # Cin not a multiple of the chunk
    (3, 3, 7, 1, "SAME", 1, None, 24, 24),        # kernel size without a tiled path
])
def test_single_conv_variants(ctx, tmp_path, case):
    cin, cout, k, stride, padding, dil, act, w, h = case
    path = tmp_path / "c.pb"
    wt, b = single_conv_graph(path, cin, cout, k, stride, padding, dil, act, seed=sum(map(hash, map(str, case))) & 0xFFFF)
    model, i, o = prepared_model(ctx, path, w, h)
    x = np.random.default_rng(1).standard_normal((h, w, cin)).astype(np.float32)
    ref = O.conv2d(x, wt, b, stride=stride, dilation=dil, padding=padding, act=act, alpha=0.3)
    assert (o.height, o.width, o.channels) == ref.shape
    inp = ctx.create_image(w, h, cin); out = ctx.create_image(o.width, o.height, o.channels)
    inp.write(x)
    model.infer(inp, out)                             # default precision (bf16x3 where the shape is supported)
    close(out.read(), ref, str(case))
    model.set_precision(api.MLX_PRECISION_FP32)
    model.infer(inp, out)
    close(out.read(), ref, str(case) + " fp32")
    assert model.launch_count() == 1
    for obj in (inp, out, model):
        obj.release()


def test_concat_and_unfused_ops(ctx, tmp_path):
    rng = np.random.default_rng(8)
    w1 = (rng.standard_normal((3, 3, 3, 4)) * 0.3).astype(np.float32)
    w2 = (rng.standard_normal((1, 1, 3, 5)) * 0.3).astype(np.float32)
    g = gw.GraphBuilder()
    x = g.placeholder("in", (-1, -1, -1, 3))
    a = g.conv2d(x, w1, name="a")
    ar = g.unary("Relu", a, name="ar")
    b = g.conv2d(x, w2, name="b")
    cat = g.concat([ar, b, a], name="cat")          # `a` has two consumers: Relu must NOT be fused into it
    g.unary("Tanh", cat, name="out")
    path = tmp_path / "cat.pb"; path.write_bytes(g.finish())
    model, i, o = prepared_model(ctx, path, 21, 13)
    assert o.channels == 13
    xin = rng.standard_normal((13, 21, 3)).astype(np.float32)
    ra = O.conv2d(xin, w1); rb = O.conv2d(xin, w2)
    ref = np.tanh(np.concatenate([np.maximum(ra, 0), rb, ra], axis=2))
    inp = ctx.create_image(21, 13, 3); out = ctx.create_image(21, 13, 13)
    inp.write(xin)
    model.infer(inp, out)
    close(out.read(), ref, "concat")
    for obj in (inp, out, model):
        obj.release()


def test_residual_adds_fused_and_unfused(ctx, tmp_path):
    """Add(conv, r) rides in the conv's epilogue when the order allows (test_graphdef checks the plan);
    the numbers must not care, nor about inferring in place (input image == output image)."""
    rng = np.random.default_rng(13)
    k = lambda ci, co: (rng.standard_normal((3, 3, ci, co)) * 0.2).astype(np.float32)
    w1, w2 = k(4, 4), k(4, 4)
    cases = {
        "relu(conv(x)+x)": (lambda g, x: g.unary("Relu", g.add(g.conv2d(x, w1, name="c"), x)),
                            lambda x: np.maximum(O.conv2d(x, w1) + x, 0)),
        "t+conv(t)": (lambda g, x: (lambda t: g.add(t, g.conv2d(t, w2, name="u")))(g.conv2d(x, w1, name="t")),
                      lambda x: (lambda t: t + O.conv2d(t, w2))(O.conv2d(x, w1))),
        "conv_a(x)+conv_b(x)": (lambda g, x: g.add(g.conv2d(x, w1, name="a"), g.conv2d(x, w2, name="b")),
                                lambda x: O.conv2d(x, w1) + O.conv2d(x, w2)),
        "x+conv(x)": (lambda g, x: g.add(x, g.conv2d(x, w1, name="c")),
                      lambda x: x + O.conv2d(x, w1)),
    }
    W, H = 37, 19
    xin = rng.standard_normal((H, W, 4)).astype(np.float32)
    for name, (build, ref_fn) in cases.items():
        g = gw.GraphBuilder()
        x = g.placeholder("input", (-1, -1, -1, 4))
        g.unary("Identity", build(g, x), name="output", out_shape=(-1, -1, -1, 4))
        path = tmp_path / "r.pb"; path.write_bytes(g.finish())
        model, i, o = prepared_model(ctx, path, W, H)
        inp = ctx.create_image(W, H, 4); out = ctx.create_image(W, H, 4)
        inp.write(xin)
        model.infer(inp, out)
        ref = ref_fn(xin)
        close(out.read(), ref, name)
        model.infer(inp, inp)                      # in place
        close(inp.read(), ref, name + " in place")
        for obj in (inp, out, model):
            obj.release()


def test_float16_images(ctx, tmp_path):
    path = tmp_path / "h.pb"
    _, layers = gw.make_denoiser(path, channels=3, width=8, depth=3, seed=3, dtype=gw.DT_HALF)
    model = ctx.create_model(path)
    i, o = model.info()
    assert i.dtype == api.ML_FLOAT16
    i.width, i.height = 20, 12
    model.set_input_info(i)
    x = np.random.default_rng(5).random((12, 20, 3), dtype=np.float32).astype(np.float16)
    inp = ctx.create_image(20, 12, 3, api.ML_FLOAT16); out = ctx.create_image(20, 12, 3, api.ML_FLOAT16)
    assert inp.map().nbytes == 20 * 12 * 3 * 2
    inp.unmap()
    inp.write(x)
    model.infer(inp, out)
    layers16 = [(w.astype(np.float16).astype(np.float32), b.astype(np.float16).astype(np.float32), a) for w, b, a in layers]
    ref = O.run_denoiser(x.astype(np.float32), layers16)
    got = out.read().astype(np.float32)
    assert float(np.abs(got - ref).max()) <= 2e-3 * max(1.0, float(np.abs(ref).max()))   # fp16 output rounding
    f32 = ctx.create_image(20, 12, 3, api.ML_FLOAT32)
    with pytest.raises(api.MLError, match="input size does not match"):
        model.infer(f32, out)
    for obj in (inp, out, f32, model):
        obj.release()


def test_shape_negotiation_contract(ctx, tmp_path):
    """reference model.cpp:161-235"""
    path = tmp_path / "m.pb"
    gw.make_denoiser(path, channels=3, width=8, depth=2, seed=1)
    model = ctx.create_model(path)
    i, o = model.info()
    assert (i.dtype, i.width, i.height, i.channels) == (api.ML_FLOAT32, 0, 0, 3)
    assert (o.width, o.height, o.channels) == (0, 0, 3)
    lib = ctx.lib
    assert lib.mlSetModelInputInfo(model.handle, None) == api.ML_FAIL
    assert model.error() == "Bad info parameter"
    bad = api.ml_image_info(api.ML_FLOAT16, 8, 8, 3)
    assert lib.mlSetModelInputInfo(model.handle, C.byref(bad)) == api.ML_FAIL
    assert model.error() == "Overriding data type 0 with 1"
    bad = api.ml_image_info(api.ML_FLOAT32, 8, 8, 4)
    assert lib.mlSetModelInputInfo(model.handle, C.byref(bad)) == api.ML_FAIL
    assert model.error() == "Overriding channels dimension 3 with 4"
    bad = api.ml_image_info(api.ML_FLOAT32, 8, 0, 3)
    assert lib.mlSetModelInputInfo(model.handle, C.byref(bad)) == api.ML_FAIL
    assert model.error() == "Input image height dimension is not specified"
    # inference before the shape is set
    a = ctx.create_image(8, 8, 3); b = ctx.create_image(8, 8, 3)
    assert lib.mlInfer(model.handle, a.handle, b.handle) == api.ML_FAIL
    assert model.error() == "Output image width dimension 8 does not match 0"
    good = api.ml_image_info(api.ML_FLOAT32, 8, 8, 3)
    model.set_input_info(good)
    assert model.error() == ""                       # cleared on success, never throws (quirk Q2)
    model.set_input_info(good)                       # unchanged -> no-op OK
    i, o = model.info()
    assert (o.width, o.height, o.channels) == (8, 8, 3)
    changed = api.ml_image_info(api.ML_FLOAT32, 16, 8, 3)
    assert lib.mlSetModelInputInfo(model.handle, C.byref(changed)) == api.ML_FAIL
    assert model.error() == "Overriding width dimension 8 with 16"
    model.infer(a, b)
    assert lib.mlGetModelInfo(model.handle, None, None) == api.ML_OK
    for obj in (a, b, model):
        obj.release()


def test_infer_contract_and_messages(ctx, tmp_path):
    """reference model.cpp:237-347"""
    path = tmp_path / "m.pb"
    gw.make_denoiser(path, channels=3, width=8, depth=2, seed=1)
    model, i, o = prepared_model(ctx, path, 12, 10)
    lib = ctx.lib
    inp = ctx.create_image(12, 10, 3); out = ctx.create_image(12, 10, 3)
    assert lib.mlInfer(model.handle, None, out.handle) == api.ML_FAIL
    assert model.error() == "Bad input image handle"
    assert lib.mlInfer(model.handle, inp.handle, None) == api.ML_FAIL
    assert model.error() == "Bad output image handle"
    wrong = ctx.create_image(12, 11, 3)
    assert lib.mlInfer(model.handle, inp.handle, wrong.handle) == api.ML_FAIL
    assert model.error() == "Output image height dimension 11 does not match 10"
    small = ctx.create_image(6, 10, 3)
    assert lib.mlInfer(model.handle, small.handle, out.handle) == api.ML_FAIL
    assert model.error() == "Internal error: input size does not match: 720 vs 1440"
    half_out = ctx.create_image(12, 10, 3, api.ML_FLOAT16)      # dims match, dtype does not (not checked by dims)
    assert lib.mlInfer(model.handle, inp.handle, half_out.handle) == api.ML_FAIL
    assert model.error() == "Internal error: output size does not match: 720 vs 1440"
    model.infer(inp, out)
    assert model.error() == ""
    for obj in (inp, out, wrong, small, half_out, model):
        obj.release()


def test_image_contract(ctx):
    """reference image.cpp:22-75 and context.cpp:20-33"""
    lib = ctx.lib
    assert lib.mlCreateImage(ctx.handle, None) is None
    assert ctx.error() == "Bad image information argument"
    for field, name in (("width", "width"), ("height", "height"), ("channels", "channels")):
        info = api.ml_image_info(api.ML_FLOAT32, 4, 4, 3)
        setattr(info, field, 0)
        assert lib.mlCreateImage(ctx.handle, C.byref(info)) is None
        assert ctx.error() == f"Unspecified image {name} dimension"
    info = api.ml_image_info(7, 4, 4, 3)
    assert lib.mlCreateImage(ctx.handle, C.byref(info)) is None
    assert ctx.error() == "Unsupported image data type: 7"
    img = ctx.create_image(5, 4, 3)
    assert ctx.error() == ""
    size = C.c_size_t(0)
    p1 = lib.mlMapImage(img.handle, C.byref(size))
    assert size.value == 5 * 4 * 3 * 4
    p2 = lib.mlMapImage(img.handle, None)
    assert p1 == p2                                   # stable pointer
    assert not np.frombuffer((C.c_char * size.value).from_address(p1), np.uint8).any()   # zero filled
    assert lib.mlUnmapImage(img.handle, p1 + 4) == api.ML_FAIL
    assert lib.mlUnmapImage(img.handle, p1) == api.ML_OK
    got = api.ml_image_info()
    assert lib.mlGetImageInfo(img.handle, C.byref(got)) == api.ML_OK
    assert (got.dtype, got.width, got.height, got.channels) == (0, 5, 4, 3)
    assert lib.mlGetImageInfo(img.handle, None) == api.ML_FAIL
    img.release()


def test_model_creation_errors(ctx, tmp_path):
    lib = ctx.lib
    assert lib.mlCreateModel(ctx.handle, None) is None
    assert ctx.error() == "Bad parameters argument"
    params = api.ml_model_params(None, None, None)
    assert lib.mlCreateModel(ctx.handle, C.byref(params)) is None
    assert ctx.error() == "Bad model_path model parameter value"
    with pytest.raises(api.MLError, match="Error reading graph definition: .*nope.pb"):
        ctx.create_model(tmp_path / "nope.pb")


def test_children_outlive_the_context(tmp_path):
    """reference context.cpp:102-105: the context holds no references to its children."""
    c = api.Context()
    path = tmp_path / "m.pb"
    _, layers = gw.make_denoiser(path, channels=3, width=8, depth=2, seed=1)
    model, i, o = prepared_model(c, path, 8, 8)
    inp = c.create_image(8, 8, 3); out = c.create_image(8, 8, 3)
    c.release()
    x = np.random.default_rng(0).random((8, 8, 3), dtype=np.float32)
    inp.write(x)
    model.infer(inp, out)
    close(out.read(), O.run_denoiser(x, layers), "after context release")
    for obj in (inp, out, model):
        obj.release()


def test_stale_pointer_pattern_still_works(ctx, tmp_path):
    """A reference-style caller may map once and keep writing/reading through the
    same pointers around every mlInfer (Map is a no-op there, image.cpp:57-65)."""
    path = tmp_path / "m.pb"
    _, layers = gw.make_denoiser(path, channels=3, width=8, depth=2, seed=1)
    model, i, o = prepared_model(ctx, path, 10, 6)
    inp = ctx.create_image(10, 6, 3); out = ctx.create_image(10, 6, 3)
    vin, vout = inp.map(), out.map()
    inp.unmap(); out.unmap()
    rng = np.random.default_rng(0)
    for _ in range(3):
        x = rng.random((6, 10, 3), dtype=np.float32)
        vin[...] = x                                   # no map/unmap around the write
        model.infer(inp, out)
        close(vout.copy(), O.run_denoiser(x, layers), "stale pointer")
    for obj in (inp, out, model):
        obj.release()


def test_render_feeds_inference_on_device(ctx, tmp_path):
    """Config 5 hand-off: the rendered frame is the model input without a host copy."""
    W, H = 64, 48
    v, a, cam, light = api.gen_cornell_box(W, H)
    scene = ctx.create_scene(v, a)
    path = tmp_path / "m.pb"
    _, layers = gw.make_denoiser(path, channels=3, width=16, depth=3, seed=7)
    model, i, o = prepared_model(ctx, path, W, H)
    frame = ctx.create_image(W, H, 3); out = ctx.create_image(W, H, 3)
    p = api.make_params(W, H, shade=api.SRT_SHADE_SHADOW, accel=api.SRT_ACCEL_BVH, spp=2, jitter=1)
    scene.render_to_image(cam, light, p, frame)
    model.infer(frame, out)
    ref_frame = O.render(v, a, cam, light, p)[0]
    close(out.read(), O.run_denoiser(ref_frame, layers), "render -> infer")
    for obj in (frame, out, model, scene):
        obj.release()


def test_config1_reference_test_app_end_to_end(tmp_path):
    """BASELINE config 1: the reference's own test_app.cpp (compiled unmodified) on one
    synthetic 256x256 RGB fp32 image, raw binary in / raw binary out."""
    exe = ROOT / "oracle" / "_ref" / "test_app"
    if not exe.exists():
        pytest.skip("oracle/_ref/test_app not built")
    model = tmp_path / "denoiser.pb"
    _, layers = gw.make_denoiser(model, channels=3, width=32, depth=4, seed=0)
    x = np.random.default_rng(0).random((256, 256, 3), dtype=np.float32)
    (tmp_path / "in.bin").write_bytes(x.tobytes())
    r = subprocess.run([str(exe), "-m", str(model), "-w", "256", "-h", "256", "-i", str(tmp_path / "in.bin"),
                        "-o", str(tmp_path / "out.bin")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "Input (init): 0 x 0 x 3" in r.stderr and "Output: 256 x 256 x 3" in r.stderr
    y = np.frombuffer((tmp_path / "out.bin").read_bytes(), np.float32).reshape(256, 256, 3)
    close(y, O.run_denoiser(x, layers), "test_app")
    # stdin/stdout path and a CLI error (exit status 255)
    r = subprocess.run([str(exe), "-m", str(model), "-w", "256", "-h", "256"], input=x.tobytes(),
                       capture_output=True, timeout=300)
    assert r.returncode == 0 and np.array_equal(np.frombuffer(r.stdout, np.float32).reshape(256, 256, 3), y)
    r = subprocess.run([str(exe), "-m", str(model), "-w", "128", "-h", "256", "-i", str(tmp_path / "in.bin")],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 255 and "Bad input size" in r.stderr


# --------------------------------------------------------------------------- bf16 tensor-core path
BF16_VS_EMULATION = 4e-3   # of the output range, against an oracle that rounds to bf16 at the same points
BF16_VS_FP32 = 3e-2        # stated bound of the bf16 path against the pure fp32 oracle (north_star)


def _bf16(x):
    import torch
    return torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)).bfloat16().float().numpy()


def bf16_emulated_denoiser(x, layers, tc_layers, alpha=0.1):
    """The conv oracle with operands rounded to bf16 exactly where the CUDA path stores bf16:
    filters of tensor-core layers, and every activation that feeds a tensor-core layer."""
    cur = _bf16(x) if 0 in tc_layers else x      # a thin first layer reads a 16-channel bf16 copy of the image
    for k, (w, b, act) in enumerate(layers):
        wk = _bf16(w) if k in tc_layers else w
        cur = O.conv2d(cur, wk, b, act=act, alpha=alpha)
        if k + 1 in tc_layers:
            cur = _bf16(cur)
    return x + cur


@pytest.mark.parametrize("cfg", [
    dict(width=64, depth=4, act="Relu", w=300, h=37, tc={0, 1, 2, 3}),     # SW128 rows, ragged last tile
    dict(width=32, depth=3, act="LeakyRelu", w=128, h=9, tc={0, 1, 2}),    # SW64 rows
    dict(width=16, depth=3, act="Relu", w=131, h=5, tc={0, 1, 2}),         # SW32 rows
    dict(width=128, depth=3, act="Relu", w=70, h=20, tc={0, 2}),           # 128->128 too big for smem: fp32; 2 K-chunks
    dict(width=64, depth=5, act="Relu", w=1000, h=150, tc={0, 1, 2, 3, 4}),  # many tiles per CTA: ring + TMEM phases wrap
    dict(width=64, depth=3, act="Tanh", w=257, h=11, tc={0, 1, 2}),        # slow-path activation
    dict(width=64, depth=3, act="Relu", w=140, h=1, tc={0, 1, 2}),         # one image row: no row above or below
    dict(width=48, depth=3, act="Relu", w=129, h=2, tc={0}),               # Cin = 48 is not a tensor-core shape: only layer 0
    dict(width=64, depth=3, act="LeakyRelu", w=100, h=700, tc={0, 1, 2}),       # one strip, many short row segments per CTA
])
@pytest.mark.parametrize("kernel", ["rows", "tile", "rows-3-boxes"])
def test_bf16_tensor_core_denoiser(ctx, tmp_path, cfg, kernel, monkeypatch):
    # both tensor-core kernels: the row-stacked one (default where it fits) and the tile-per-stage one
    # (Cin > 64 or Cout > 80, or MLB200_TC_ROWS=0); the library reads the switches when the plan is built
    monkeypatch.setenv("MLB200_TC_ROWS", "0" if kernel == "tile" else "1")
    monkeypatch.setenv("MLB200_TC_HALO", "0" if kernel == "rows-3-boxes" else "1")
    path = tmp_path / "d.pb"
    _, layers = gw.make_denoiser(path, channels=3, width=cfg["width"], depth=cfg["depth"], seed=11, act=cfg["act"])
    w, h = cfg["w"], cfg["h"]
    model, i, o = prepared_model(ctx, path, w, h)
    x = np.random.default_rng(5).random((h, w, 3), dtype=np.float32)
    inp = ctx.create_image(w, h, 3); out = ctx.create_image(w, h, 3)
    inp.write(x)
    model.set_precision(api.MLX_PRECISION_FP32)
    model.infer(inp, out)
    fp32 = out.read().copy()
    assert model.tensor_core_layers() == 0
    model.set_precision(api.MLX_PRECISION_BF16)
    assert model.tensor_core_layers() == len(cfg["tc"])
    model.infer(inp, out)
    got = out.read().copy()
    ref = O.run_denoiser(x, layers, alpha=0.1)
    emu = bf16_emulated_denoiser(x, layers, cfg["tc"])
    scale = float(np.abs(ref).max())
    err_emu = float(np.abs(got - emu).max()) / scale
    err_fp32 = float(np.abs(got - ref).max()) / scale
    assert err_emu <= BF16_VS_EMULATION, f"{cfg}: {err_emu:.3e} vs the bf16-emulating oracle"
    assert err_fp32 <= BF16_VS_FP32, f"{cfg}: {err_fp32:.3e} vs the fp32 oracle"
    close(fp32, ref, "fp32 path")
    # switching back restores the fp32 path bit for bit
    model.set_precision(api.MLX_PRECISION_FP32)
    model.infer(inp, out)
    assert np.array_equal(out.read(), fp32)
    for obj in (inp, out, model):
        obj.release()


BF16X3_VS_FP32 = 1e-4   # stated bound of the split-operand tensor-core path, fraction of the output range


@pytest.mark.parametrize("cfg", [
    dict(width=64, depth=4, act="Relu", w=300, h=37, tc=4),        # 64->64: 3 K chunks over [hi|lo], SW128
    dict(width=32, depth=3, act="LeakyRelu", w=128, h=9, tc=3),    # SW64
    dict(width=16, depth=3, act="Relu", w=131, h=5, tc=3),         # SW32
    dict(width=64, depth=5, act="Relu", w=1000, h=150, tc=5),      # many tiles per CTA: ring + TMEM phases wrap
    dict(width=64, depth=3, act="Tanh", w=257, h=11, tc=3),        # slow-path activation
    dict(width=64, depth=3, act="Relu", w=140, h=1, tc=3),         # one image row
    dict(width=128, depth=3, act="Relu", w=70, h=20, tc=2),        # 128->128 does not fit: fp32; 128->3 = 6 K chunks
    dict(width=48, depth=3, act="Relu", w=129, h=2, tc=1),         # Cin = 48 is not a tensor-core shape: only layer 0
])
@pytest.mark.parametrize("form", ["wide", "narrow"])
def test_bf16x3_tensor_core_denoiser_is_fp32_grade(ctx, tmp_path, cfg, form, monkeypatch):
    """MLX_PRECISION_BF16X3: split-operand tcgen05 layers against the fp32 conv oracle (tolerance 1e-4 of
    the range, ten times inside the 1e-3 bar north_star sets for fp32 inference).  Both MMA forms: hi and
    lo filter rows stacked along N (default where 2N <= 256) and three N-wide products per K chunk."""
    monkeypatch.setenv("MLB200_X3_WIDE", "1" if form == "wide" else "0")
    path = tmp_path / "d.pb"
    _, layers = gw.make_denoiser(path, channels=3, width=cfg["width"], depth=cfg["depth"], seed=11, act=cfg["act"])
    w, h = cfg["w"], cfg["h"]
    model, i, o = prepared_model(ctx, path, w, h)
    x = np.random.default_rng(5).random((h, w, 3), dtype=np.float32)
    inp = ctx.create_image(w, h, 3); out = ctx.create_image(w, h, 3)
    inp.write(x)
    assert model.tensor_core_layers() == cfg["tc"]          # bf16x3 is the default precision
    model.set_precision(api.MLX_PRECISION_FP32)
    assert model.tensor_core_layers() == 0
    model.infer(inp, out)
    fp32 = out.read().copy()
    model.set_precision(api.MLX_PRECISION_BF16X3)
    assert model.tensor_core_layers() == cfg["tc"]
    model.infer(inp, out)
    got = out.read().copy()
    ref = O.run_denoiser(x, layers, alpha=0.1)
    scale = float(np.abs(ref).max())
    err = float(np.abs(got - ref).max()) / scale
    assert err <= BF16X3_VS_FP32, f"{cfg}: {err:.3e} vs the fp32 oracle"
    # in place (input image == output image), and back to fp32 bit for bit
    inp2 = ctx.create_image(w, h, 3)
    inp2.write(x)
    model.infer(inp2, inp2)
    assert np.array_equal(inp2.read(), got)
    model.set_precision(api.MLX_PRECISION_BF16)
    model.infer(inp, out)
    bf16 = out.read().copy()
    inp2.write(x)
    model.infer(inp2, inp2)                        # in place through the bf16 layers as well
    assert np.array_equal(inp2.read(), bf16)
    model.set_precision(api.MLX_PRECISION_FP32)
    model.infer(inp, out)
    assert np.array_equal(out.read(), fp32)
    for obj in (inp, inp2, out, model):
        obj.release()


@pytest.mark.parametrize("shape", [(1, 1), (127, 3), (128, 4), (129, 5), (255, 31), (256, 32), (257, 33), (3, 200),
                                   (385, 7), (640, 360), (37, 1080), (1025, 2)])
def test_bf16_row_stacked_kernel_tiles_any_frame(ctx, tmp_path, shape):
    """The row-stacked kernel cuts the (strip, row) tile sequence into 2 x 148 contiguous ranges that may
    start and end anywhere, cross strip boundaries, or be empty: frames around the 128-pixel tile width,
    shorter than the pipes, one pixel wide, one row high."""
    w, h = shape
    path = tmp_path / "d.pb"
    _, layers = gw.make_denoiser(path, channels=3, width=64, depth=3, seed=21, act="Relu")
    model, i, o = prepared_model(ctx, path, w, h)
    x = np.random.default_rng(w * 131 + h).random((h, w, 3), dtype=np.float32)
    inp = ctx.create_image(w, h, 3); out = ctx.create_image(w, h, 3)
    inp.write(x)
    model.set_precision(api.MLX_PRECISION_BF16)
    assert model.tensor_core_layers() == 3
    model.infer(inp, out)
    emu = bf16_emulated_denoiser(x, layers, {0, 1, 2})
    scale = max(float(np.abs(emu).max()), 1e-6)
    err = float(np.abs(out.read() - emu).max()) / scale
    assert err <= BF16_VS_EMULATION, f"{shape}: {err:.3e}"
    for obj in (inp, out, model):
        obj.release()


@pytest.mark.parametrize("precision", [api.MLX_PRECISION_FP32, api.MLX_PRECISION_BF16, api.MLX_PRECISION_BF16X3])
def test_infer_batch_equals_single_infers(ctx, tmp_path, precision):
    """mlxInferBatch: pipelined batch == the same mlInfer calls one by one, bit for bit;
    host-written inputs are uploaded inside the call, mapped outputs are read back inside the call."""
    path = tmp_path / "d.pb"
    gw.make_denoiser(path, channels=3, width=32, depth=3, seed=2)
    w, h, n = 200, 64, 5
    model, i, o = prepared_model(ctx, path, w, h)
    model.set_precision(precision)
    rng = np.random.default_rng(8)
    xs = [rng.random((h, w, 3), dtype=np.float32) for _ in range(n)]
    ins = [ctx.create_image(w, h, 3) for _ in range(n)]
    outs = [ctx.create_image(w, h, 3) for _ in range(n)]
    single = []
    for x, a, b in zip(xs, ins, outs):
        a.write(x)
        model.infer(a, b)
        single.append(b.read().copy())
        b.write(np.zeros_like(x))            # host pointer handed out + stale content
    model.infer_batch(ins, outs)
    for k in range(n):
        assert np.array_equal(outs[k].read(), single[k]), f"image {k}"
    # per-image checks keep mlInfer's messages
    bad = ctx.create_image(w + 1, h, 3)
    with pytest.raises(api.MLError, match=r"Image 1: Output image width dimension 201 does not match 200"):
        model.infer_batch(ins[:2], [outs[0], bad])
    model.infer_batch([], [])
    for obj in ins + outs + [bad, model]:
        obj.release()


def test_render_cli_pipes_into_reference_test_app(ctx, tmp_path):
    """srt_render writes the headerless HWC fp32 wire format of test_app.cpp:137-192, so a
    rendered frame can be piped straight into the reference's (unmodified) client."""
    cli = ROOT / "simpleraytracer_b200" / "bin" / "srt_render"
    exe = ROOT / "oracle" / "_ref" / "test_app"
    if not cli.exists() or not exe.exists():
        pytest.skip("srt_render / oracle/_ref/test_app not built")
    W, H, n, spp = 160, 96, 3000, 8
    r = subprocess.run([str(cli), "-s", "mesh", "-n", str(n), "-seed", "3", "-w", str(W), "-h", str(H),
                        "-spp", str(spp)], capture_output=True, timeout=300)
    assert r.returncode == 0, r.stderr
    frame = np.frombuffer(r.stdout, np.float32).reshape(H, W, 3)
    # the same frame through the library API
    v, a, cam, light = api.gen_mesh(3, n, W, H)
    scene = ctx.create_scene(v, a)
    p = api.make_params(W, H, spp=spp, jitter=1, seed=3, shade=api.SRT_SHADE_SHADOW, accel=api.SRT_ACCEL_BVH,
                        background=(0.5, 0.7, 0.9))
    assert np.array_equal(frame, scene.render(cam, light, p)[0])
    scene.release()
    model = tmp_path / "denoiser.pb"
    _, layers = gw.make_denoiser(model, channels=3, width=16, depth=3, seed=5)
    out = subprocess.run([str(exe), "-m", str(model), "-w", str(W), "-h", str(H)], input=r.stdout,
                         capture_output=True, timeout=300)
    assert out.returncode == 0, out.stderr
    y = np.frombuffer(out.stdout, np.float32).reshape(H, W, 3)
    close(y, O.run_denoiser(frame, layers), "srt_render | test_app")
    # CLI errors: status 255 like test_app
    assert subprocess.run([str(cli), "-s", "mesh"], capture_output=True).returncode == 255


def test_unsupported_op_fails_model_creation_loudly(ctx, tmp_path):
    """A graph with an op outside the supported set is refused by mlCreateModel itself: NULL handle and the
    reference-style 'Error creating graph: Unimplemented: ...' message on the CONTEXT (context.cpp:36-49) --
    never a silently skipped node."""
    g = gw.GraphBuilder()
    x = g.placeholder("input", (-1, -1, -1, 3))
    w = np.zeros((3, 3, 3, 4), np.float32)
    y = g.conv2d(x, w, name="c0", out_shape=(-1, -1, -1, 4))
    y = g.unary("Softmax", y, name="sm", out_shape=(-1, -1, -1, 4))
    path = tmp_path / "softmax.pb"
    path.write_bytes(g.finish())
    with pytest.raises(api.MLError, match=r"Error creating graph: Unimplemented: .*Softmax"):
        ctx.create_model(path)
    params = api.ml_model_params(str(path).encode(), None, None)
    assert not ctx.lib.mlCreateModel(ctx.handle, C.byref(params))
    assert "Unimplemented" in ctx.error() and "sm" in ctx.error()


def test_infer_batch_with_chained_images_runs_in_order(ctx, tmp_path):
    """An image that is item 0's output and item 1's input: the batch must behave like the two mlInfer calls in
    order (the pipelined path would upload it while it is still being produced)."""
    path = tmp_path / "d.pb"
    gw.make_denoiser(path, channels=3, width=16, depth=2, seed=4)
    w, h = 96, 40
    model, i, o = prepared_model(ctx, path, w, h)
    x = np.random.default_rng(2).random((h, w, 3), dtype=np.float32)
    a, b, c = (ctx.create_image(w, h, 3) for _ in range(3))
    a.write(x)
    model.infer(a, b); model.infer(b, c)
    want = c.read()
    b2, c2 = ctx.create_image(w, h, 3), ctx.create_image(w, h, 3)
    b2.read(); c2.read()                      # host pointers handed out: read-back inside the call
    model.infer_batch([a, b2], [b2, c2])
    assert np.array_equal(c2.read(), want)
    for obj in (a, b, c, b2, c2, model):
        obj.release()
