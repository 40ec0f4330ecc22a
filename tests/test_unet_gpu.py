"""GPU tests of the ops around the convolutions of an encoder / decoder graph ("resize / concat if a U-Net
is wanted"; the reference runs whatever its frozen GraphDef holds, model.cpp:87-139, and its README's example is a
denoiser, README.md:158): MaxPool / AvgPool, ResizeNearestNeighbor / ResizeBilinear (constant size and the
Shape -> StridedSlice -> Mul pattern Keras UpSampling2D exports), DepthToSpace, Pad / MirrorPad, FusedBatchNorm,
Mul / Sub / RealDiv with constants, element-wise Sub / Mul, Squeeze / Reshape.  Each against the CPU oracle
(tests/test_conv_oracle.py pins that oracle against torch) through the C ABI, then whole U-Nets, then a U-Net through
the reference's unmodified test_app."""
import subprocess
from pathlib import Path

import numpy as np
import pytest

import oracle_lib as O
from simpleraytracer_b200 import api
from simpleraytracer_b200 import graphdef_writer as gw
from test_model_gpu import close, prepared_model

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def run_graph(ctx, tmp_path, g, x, precision=api.MLX_PRECISION_FP32):
    path = tmp_path / "g.pb"
    path.write_bytes(g.finish())
    h, w, c = x.shape
    model, i, o = prepared_model(ctx, path, w, h)
    model.set_precision(precision)
    inp = ctx.create_image(w, h, c)
    out = ctx.create_image(o.width, o.height, o.channels)
    inp.write(x)
    model.infer(inp, out)
    y = out.read()
    for obj in (inp, out, model):
        obj.release()
    return y


SINGLE_OPS = [
    ("maxpool_same", lambda g, x: g.pool(x, "max", 2, 2, "SAME"), lambda x: O.pool2d(x, 2, 2, "SAME", "max")),
    ("maxpool_3x3s2_valid", lambda g, x: g.pool(x, "max", 3, 2, "VALID"), lambda x: O.pool2d(x, 3, 2, "VALID", "max")),
    ("avgpool_same", lambda g, x: g.pool(x, "avg", 2, 2, "SAME"), lambda x: O.pool2d(x, 2, 2, "SAME", "avg")),
    ("avgpool_3x3s1_same", lambda g, x: g.pool(x, "avg", 3, 1, "SAME"), lambda x: O.pool2d(x, 3, 1, "SAME", "avg")),
    ("nearest_x2_keras", lambda g, x: g.resize(x, factor=2), lambda x: O.resize(x, x.shape[0] * 2, x.shape[1] * 2)),
    ("nearest_const_size", lambda g, x: g.resize(x, size=(31, 50)), lambda x: O.resize(x, 31, 50)),
    ("nearest_half_pixel", lambda g, x: g.resize(x, size=(40, 33), half_pixel_centers=True),
     lambda x: O.resize(x, 40, 33, half_pixel_centers=True)),
    ("nearest_align_corners", lambda g, x: g.resize(x, size=(40, 33), align_corners=True),
     lambda x: O.resize(x, 40, 33, align_corners=True)),
    ("bilinear_x2_half_pixel", lambda g, x: g.resize(x, factor=2, bilinear=True, half_pixel_centers=True),
     lambda x: O.resize(x, x.shape[0] * 2, x.shape[1] * 2, True, False, True)),
    ("bilinear_legacy", lambda g, x: g.resize(x, size=(37, 45), bilinear=True), lambda x: O.resize(x, 37, 45, True)),
    ("bilinear_align_corners_down", lambda g, x: g.resize(x, size=(9, 11), bilinear=True, align_corners=True),
     lambda x: O.resize(x, 9, 11, True, True)),
    ("pad_constant", lambda g, x: g.pad(x, 2, 1, 3, 0), lambda x: O.pad2d(x, 2, 1, 3, 0)),
    ("mirror_reflect", lambda g, x: g.pad(x, 2, 3, 1, 2, "REFLECT"), lambda x: O.pad2d(x, 2, 3, 1, 2, "REFLECT")),
    ("mirror_symmetric", lambda g, x: g.pad(x, 2, 3, 1, 2, "SYMMETRIC"), lambda x: O.pad2d(x, 2, 3, 1, 2, "SYMMETRIC")),
    ("mul_scalar", lambda g, x: g.with_const("Mul", x, 1.7), lambda x: (x * np.float32(1.7)).astype(np.float32)),
    ("sub_scalar", lambda g, x: g.with_const("Sub", x, 0.25), lambda x: (x - np.float32(0.25)).astype(np.float32)),
    ("scalar_minus_x", lambda g, x: g.with_const("Sub", x, 1.0, const_first=True), lambda x: (np.float32(1.0) - x).astype(np.float32)),
    ("sub_mul_self", lambda g, x: g.binary("Mul", g.binary("Sub", x, g.with_const("Mul", x, 0.5)), x),
     lambda x: ((x - x * np.float32(0.5)) * x).astype(np.float32)),
    ("squeeze_reshape", lambda g, x: g.unary("Relu", g.reshape(g.squeeze(x), [1, -1, -1, 4])), lambda x: np.maximum(x, 0)),
]
EXACT = {"maxpool", "nearest", "pad", "mirror", "squeeze"}     # pure data movement or max: bit-identical


@pytest.mark.parametrize("name,build,ref", SINGLE_OPS, ids=[s[0] for s in SINGLE_OPS])
def test_single_ops_against_the_oracle(ctx, tmp_path, name, build, ref):
    x = np.random.default_rng(len(name)).standard_normal((19, 26, 4)).astype(np.float32)
    g = gw.GraphBuilder()
    inp = g.placeholder("input", (-1, -1, -1, 4))
    g.unary("Identity", build(g, inp), name="output")
    got, want = run_graph(ctx, tmp_path, g, x), ref(x)
    assert got.shape == want.shape
    if any(name.startswith(p) for p in EXACT):
        assert np.array_equal(got, want), name
    else:
        np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-6, err_msg=name)


def test_per_channel_constants_division_depth_to_space_and_batch_norm(ctx, tmp_path):
    rng = np.random.default_rng(12)
    x = rng.standard_normal((10, 14, 8)).astype(np.float32)
    k = rng.uniform(0.5, 2.0, 8).astype(np.float32)
    g = gw.GraphBuilder()
    inp = g.placeholder("input", (-1, -1, -1, 8))
    cur = g.with_const("RealDiv", g.with_const("Mul", inp, k), 3.0)
    cur = g.depth_to_space(cur, 2)
    p = dict(scale=rng.uniform(0.5, 1.5, 2).astype(np.float32), offset=rng.standard_normal(2).astype(np.float32),
             mean=rng.standard_normal(2).astype(np.float32), variance=rng.uniform(0.5, 1.5, 2).astype(np.float32))
    cur = g.batch_norm(cur, p["scale"], p["offset"], p["mean"], p["variance"], 1e-3)
    g.unary("Relu6", cur, name="output")
    want = O.depth_to_space(((x * k) / np.float32(3.0)).astype(np.float32), 2)
    want = np.clip(O.batch_norm(want, p["scale"], p["offset"], p["mean"], p["variance"], 1e-3), 0, 6)
    got = run_graph(ctx, tmp_path, g, x)
    assert got.shape == (20, 28, 2)
    np.testing.assert_allclose(got, want, rtol=2e-6, atol=2e-6)


@pytest.mark.parametrize("up,down", [("nearest", "maxpool"), ("bilinear", "avgpool"), ("d2s", "conv")])
@pytest.mark.parametrize("precision", [api.MLX_PRECISION_FP32, api.MLX_PRECISION_BF16X3])
def test_unets_against_the_oracle(ctx, tmp_path, up, down, precision):
    path = tmp_path / "unet.pb"
    _, prog = gw.make_unet(path, channels=3, width=16, levels=2, seed=7, up=up, down=down)
    w, h = 72, 48
    x = np.random.default_rng(3).random((h, w, 3), dtype=np.float32)
    model, i, o = prepared_model(ctx, path, w, h)
    assert (o.width, o.height, o.channels) == (w, h, 3)         # shape negotiation through pool / resize / concat
    model.set_precision(precision)
    if precision == api.MLX_PRECISION_BF16X3:
        assert model.tensor_core_layers() >= 3                  # encoder + bottleneck 3x3 convs on tcgen05 (a decoder conv after a
                                                                # concat has 48 or 96 input channels: fp32 CUDA-core kernel)
    inp, out = ctx.create_image(w, h, 3), ctx.create_image(w, h, 3)
    inp.write(x)
    model.infer(inp, out)
    close(out.read(), O.run_program(x, prog), f"unet {up}/{down}")
    # in place (input image == output image): the first op reads the input element-wise, the residual is folded
    model.infer(inp, inp)
    close(inp.read(), O.run_program(x, prog), f"unet {up}/{down} in place")
    # a frame whose size is not a multiple of 4: the skip connection and the upsampled tensor differ -> loud failure
    # (on a fresh model: a dimension that has been set cannot be overridden, reference model.cpp:173)
    fresh = ctx.create_model(path)
    bad, _ = fresh.info()
    bad.width, bad.height = 70, 46
    with pytest.raises(api.MLError, match="differ in spatial size"):
        fresh.set_input_info(bad)
    for obj in (inp, out, model, fresh):
        obj.release()


def test_in_place_inference_through_a_single_convolution(ctx, tmp_path):
    """a graph whose last op is a convolution reading the placeholder must not read pixels it has already
    overwritten when the caller passes the same image as input and output."""
    rng = np.random.default_rng(5)
    w3 = (rng.standard_normal((3, 3, 3, 3)) * 0.3).astype(np.float32)
    g = gw.GraphBuilder()
    x = g.placeholder("input", (-1, -1, -1, 3))
    g.conv2d(x, w3, name="output")
    path = tmp_path / "one_conv.pb"
    path.write_bytes(g.finish())
    img = np.random.default_rng(6).random((40, 56, 3), dtype=np.float32)
    for precision in (api.MLX_PRECISION_FP32, api.MLX_PRECISION_BF16X3):
        model, i, o = prepared_model(ctx, path, 56, 40)
        model.set_precision(precision)
        a = ctx.create_image(56, 40, 3)
        a.write(img)
        model.infer(a, a)
        close(a.read(), O.conv2d(img, w3), "in-place single conv")
        a.release(); model.release()


def test_unet_through_the_unmodified_reference_test_app(tmp_path):
    """`test_app -m unet.pb`: an encoder / decoder .pb written by graphdef_writer.py runs through the reference's own
    client compiled unmodified (oracle/_ref/test_app) and matches the oracle within 1e-3."""
    exe = ROOT / "oracle" / "_ref" / "test_app"
    if not exe.exists():
        pytest.skip("oracle/_ref/test_app not built")
    model = tmp_path / "unet.pb"
    _, prog = gw.make_unet(model, channels=3, width=16, levels=2, seed=1, up="nearest", down="maxpool")
    x = np.random.default_rng(0).random((256, 256, 3), dtype=np.float32)
    (tmp_path / "in.bin").write_bytes(x.tobytes())
    r = subprocess.run([str(exe), "-m", str(model), "-w", "256", "-h", "256", "-i", str(tmp_path / "in.bin"),
                        "-o", str(tmp_path / "out.bin")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "Output: 256 x 256 x 3" in r.stderr
    y = np.frombuffer((tmp_path / "out.bin").read_bytes(), np.float32).reshape(256, 256, 3)
    close(y, O.run_program(x, prog), "test_app unet")


def test_golden_unet(ctx, tmp_path):
    """The committed encoder / decoder fixture (tests/golden/unet_24x32.npz: graph bytes, input, oracle output; the CPU
    suite pins the oracle and the writer against it) through the C ABI on the fp32 path."""
    g = np.load(ROOT / "tests" / "golden" / "unet_24x32.npz")
    path = tmp_path / "unet.pb"
    path.write_bytes(g["graphdef"].tobytes())
    model, i, o = prepared_model(ctx, path, 32, 24)
    assert (o.width, o.height, o.channels) == (32, 24, 3)
    model.set_precision(api.MLX_PRECISION_FP32)
    inp, out = ctx.create_image(32, 24, 3), ctx.create_image(32, 24, 3)
    inp.write(g["x"])
    model.infer(inp, out)
    close(out.read(), g["y"], "golden unet")
    for obj in (inp, out, model):
        obj.release()
