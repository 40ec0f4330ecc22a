"""CPU: the conv oracle against an independent implementation (torch CPU fp32)
and against the committed golden vector."""
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.nn.functional as F

import oracle_lib as O

GOLD = Path(__file__).resolve().parent / "golden"


def torch_conv(x, w, b, stride, dilation, padding, act, alpha):
    xt = torch.from_numpy(x).permute(2, 0, 1)[None]
    wt = torch.from_numpy(w).permute(3, 2, 0, 1).contiguous()
    H, W = x.shape[:2]
    KH, KW = w.shape[:2]
    if padding == "SAME":
        def pads(n, k):
            out = -(-n // stride)
            total = max((out - 1) * stride + (k - 1) * dilation + 1 - n, 0)
            return total // 2, total - total // 2
        pt, pb = pads(H, KH); pl, pr = pads(W, KW)
        xt = F.pad(xt, (pl, pr, pt, pb))
    y = F.conv2d(xt, wt, None if b is None else torch.from_numpy(b), stride=stride, dilation=dilation)
    if act == "Relu": y = F.relu(y)
    elif act == "LeakyRelu": y = F.leaky_relu(y, alpha)
    elif act == "Relu6": y = F.relu6(y)
    elif act == "Sigmoid": y = torch.sigmoid(y)
    elif act == "Tanh": y = torch.tanh(y)
    return y[0].permute(1, 2, 0).contiguous().numpy()


CASES = [
    # H, W, Cin, Cout, K, stride, dilation, padding, act
    (16, 16, 3, 8, 3, 1, 1, "SAME", "Relu"),
    (17, 23, 5, 7, 3, 1, 1, "SAME", None),
    (17, 23, 5, 7, 3, 1, 1, "VALID", "LeakyRelu"),
    (20, 18, 4, 6, 5, 1, 1, "SAME", "Tanh"),
    (20, 19, 4, 6, 3, 2, 1, "SAME", "Relu6"),
    (21, 19, 4, 6, 3, 2, 1, "VALID", "Sigmoid"),
    (15, 15, 3, 4, 3, 1, 2, "SAME", None),
    (9, 11, 8, 3, 1, 1, 1, "SAME", None),
    (1, 1, 2, 2, 3, 1, 1, "SAME", None),
]


@pytest.mark.parametrize("case", CASES)
def test_conv_oracle_matches_torch(case):
    H, W, Cin, Cout, K, stride, dil, padding, act = case
    rng = np.random.default_rng(hash(case) & 0xFFFF)
    x = rng.standard_normal((H, W, Cin)).astype(np.float32)
    w = (rng.standard_normal((K, K, Cin, Cout)) * 0.3).astype(np.float32)
    b = rng.standard_normal(Cout).astype(np.float32)
    got = O.conv2d(x, w, b, stride=stride, dilation=dil, padding=padding, act=act, alpha=0.1, threads=2)
    ref = torch_conv(x, w, b, stride, dil, padding, act, 0.1)
    assert got.shape == ref.shape
    np.testing.assert_allclose(got, ref, rtol=1e-4, atol=1e-5)


def test_conv_oracle_matches_golden_network():
    g = np.load(GOLD / "convnet_16.npz")
    from simpleraytracer_b200 import graphdef_writer as gw
    data, layers = gw.make_denoiser(channels=3, width=8, depth=3, seed=5)
    assert np.array_equal(np.frombuffer(data, np.uint8), g["graphdef"])   # writer is deterministic
    y = O.run_denoiser(g["x"], layers)
    assert np.array_equal(y, g["y"])


def test_oracle_matches_golden_unet():
    """The committed encoder / decoder fixture (tests/golden/make_golden.py::unet): conv + folded batch norm, max pool,
    bilinear resize, concat, 1x1 head.  Pins the writer (same bytes) and the oracle's program runner (same output)."""
    import sys
    sys.path.insert(0, str(GOLD))
    from make_golden import UNET
    from simpleraytracer_b200 import graphdef_writer as gw
    g = np.load(GOLD / "unet_24x32.npz")
    data, prog = gw.make_unet(None, **UNET)
    assert np.array_equal(np.frombuffer(data, np.uint8), g["graphdef"])
    y = O.run_program(g["x"], prog)
    assert y.shape == (24, 32, 3) and np.array_equal(y, g["y"])
    assert float(np.abs(y - g["x"]).max()) > 1e-3            # the net does something


def test_conv_oracle_threads_do_not_change_results():
    rng = np.random.default_rng(0)
    x = rng.standard_normal((33, 20, 6)).astype(np.float32)
    w = rng.standard_normal((3, 3, 6, 5)).astype(np.float32)
    assert np.array_equal(O.conv2d(x, w, threads=1), O.conv2d(x, w, threads=7))


# ---- the ops around the convolutions of an encoder / decoder graph, each against torch on the CPU -------------
def chw(x):
    return torch.from_numpy(np.ascontiguousarray(x)).permute(2, 0, 1)[None]


def hwc(t):
    return t[0].permute(1, 2, 0).contiguous().numpy()


@pytest.mark.parametrize("H,W,k,stride,padding", [(16, 16, 2, 2, "SAME"), (17, 23, 2, 2, "SAME"), (17, 23, 3, 2, "VALID"),
                                                  (9, 9, 3, 1, "SAME"), (8, 10, 2, 2, "VALID"), (5, 7, 3, 3, "SAME")])
@pytest.mark.parametrize("kind", ["max", "avg"])
def test_pool_oracle_matches_torch(H, W, k, stride, padding, kind):
    x = np.random.default_rng(H * W + k).standard_normal((H, W, 5)).astype(np.float32)
    got = O.pool2d(x, k, stride, padding, kind)
    xt = chw(x)
    if padding == "SAME":
        def pads(n):
            out = -(-n // stride)
            total = max((out - 1) * stride + k - n, 0)
            return total // 2, total - total // 2
        (pt, pb), (pl, pr) = pads(H), pads(W)
        if kind == "max":
            ref = F.max_pool2d(F.pad(xt, (pl, pr, pt, pb), value=float("-inf")), k, stride)
        else:       # TensorFlow averages over the in-bounds taps only: sum / count, both by an unpadded-aware pool
            ones = torch.ones_like(xt)
            s = F.avg_pool2d(F.pad(xt, (pl, pr, pt, pb)), k, stride) * (k * k)
            n = F.avg_pool2d(F.pad(ones, (pl, pr, pt, pb)), k, stride) * (k * k)
            ref = s / n
    else:
        ref = F.max_pool2d(xt, k, stride) if kind == "max" else F.avg_pool2d(xt, k, stride)
    ref = hwc(ref)
    assert got.shape == ref.shape
    np.testing.assert_allclose(got, ref, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("H,W,OH,OW", [(8, 8, 16, 16), (7, 5, 14, 10), (6, 9, 15, 20), (10, 10, 5, 5), (3, 4, 3, 4)])
def test_resize_oracle_matches_torch(H, W, OH, OW):
    x = np.random.default_rng(H + 31 * OW).standard_normal((H, W, 3)).astype(np.float32)
    xt = chw(x)
    # bilinear: half_pixel_centers = torch align_corners=False; align_corners = torch align_corners=True
    got = O.resize(x, OH, OW, bilinear=True, half_pixel_centers=True)
    np.testing.assert_allclose(got, hwc(F.interpolate(xt, (OH, OW), mode="bilinear", align_corners=False)), rtol=1e-5, atol=1e-5)
    got = O.resize(x, OH, OW, bilinear=True, align_corners=True)
    np.testing.assert_allclose(got, hwc(F.interpolate(xt, (OH, OW), mode="bilinear", align_corners=True)), rtol=1e-5, atol=1e-5)
    # nearest: TensorFlow's legacy mode (floor(dst * in/out)) is torch's "nearest"; half-pixel is "nearest-exact"
    assert np.array_equal(O.resize(x, OH, OW), hwc(F.interpolate(xt, (OH, OW), mode="nearest")))
    assert np.array_equal(O.resize(x, OH, OW, half_pixel_centers=True), hwc(F.interpolate(xt, (OH, OW), mode="nearest-exact")))
    # legacy bilinear (no half pixel, no align_corners): no torch twin; pin two defining properties instead
    leg = O.resize(x, OH, OW, bilinear=True)
    assert np.array_equal(leg[0, 0], x[0, 0])
    if OH == 2 * H and OW == 2 * W:
        assert np.array_equal(leg[::2, ::2], x)          # even output samples fall exactly on input samples


def test_depth_to_space_pad_and_batch_norm_oracles_match_torch():
    rng = np.random.default_rng(4)
    x = rng.standard_normal((6, 5, 12)).astype(np.float32)
    # TensorFlow DepthToSpace is the DCR order: channel = (i * bs + j) * Co + c; torch pixel_shuffle is CRD
    bs, Co = 2, 3
    got = O.depth_to_space(x, bs)
    crd = x.reshape(6, 5, bs, bs, Co).transpose(0, 1, 4, 2, 3).reshape(6, 5, Co * bs * bs)
    assert np.array_equal(got, hwc(F.pixel_shuffle(chw(crd), bs)))
    y = rng.standard_normal((7, 6, 2)).astype(np.float32)
    for mode, tmode in (("CONSTANT", "constant"), ("REFLECT", "reflect")):
        got = O.pad2d(y, 2, 1, 3, 2, mode)
        assert np.array_equal(got, hwc(F.pad(chw(y), (3, 2, 2, 1), mode=tmode)))
    assert np.array_equal(O.pad2d(y, 2, 1, 3, 2, "SYMMETRIC"), np.pad(y, ((2, 1), (3, 2), (0, 0)), mode="symmetric"))
    assert np.array_equal(O.pad2d(y, 2, 1, 3, 2, "REFLECT"), np.pad(y, ((2, 1), (3, 2), (0, 0)), mode="reflect"))
    g, b, m, v = (rng.uniform(0.5, 1.5, 12).astype(np.float32), rng.standard_normal(12).astype(np.float32),
                  rng.standard_normal(12).astype(np.float32), rng.uniform(0.5, 1.5, 12).astype(np.float32))
    got = O.batch_norm(x, g, b, m, v, 1e-3)
    ref = hwc(F.batch_norm(chw(x), torch.from_numpy(m), torch.from_numpy(v), torch.from_numpy(g), torch.from_numpy(b),
                           training=False, eps=1e-3))
    np.testing.assert_allclose(got, ref, rtol=1e-5, atol=1e-6)


def test_unet_program_oracle_matches_a_torch_restatement():
    """The whole U-Net program (graphdef_writer.make_unet) through the oracle interpreter against the same net
    written with torch ops: anchors the interpreter, not just the single ops."""
    from simpleraytracer_b200 import graphdef_writer as gw
    _, prog = gw.make_unet(channels=3, width=8, levels=2, seed=3, up="bilinear", down="avgpool")
    x = np.random.default_rng(1).random((16, 24, 3), dtype=np.float32)
    got = O.run_program(x, prog)
    env = {"input": chw(x)}
    for op, name, ins, kw in prog:
        a = [env[i] for i in ins]
        if op == "conv":
            w = torch.from_numpy(kw["w"]).permute(3, 2, 0, 1).contiguous()
            k, s = kw["w"].shape[0], kw.get("stride", 1)
            H, W = a[0].shape[2:]
            def pads(n):
                out = -(-n // s)
                total = max((out - 1) * s + k - n, 0)
                return total // 2, total - total // 2
            (pt, pb), (pl, pr) = pads(H), pads(W)
            r = F.conv2d(F.pad(a[0], (pl, pr, pt, pb)), w, torch.from_numpy(kw["b"]), stride=s)
        elif op == "bn":
            r = F.batch_norm(a[0], *(torch.from_numpy(kw[q]) for q in ("mean", "variance", "scale", "offset")), training=False,
                             eps=kw["epsilon"])
        elif op == "act":
            r = F.relu(a[0]) if kw["act"] == "Relu" else a[0]
        elif op == "pool":
            r = F.avg_pool2d(a[0], 2, 2)
        elif op == "resize":
            r = F.interpolate(a[0], scale_factor=2, mode="bilinear", align_corners=False)
        elif op == "concat":
            r = torch.cat(a, dim=1)
        elif op == "add":
            r = a[0] + a[1]
        elif op == "subc":
            r = a[0] - kw["c"]
        elif op == "mulc":
            r = a[0] * kw["c"]
        else:
            raise AssertionError(op)
        env[name] = r
    np.testing.assert_allclose(got, hwc(r), rtol=1e-4, atol=1e-5)
