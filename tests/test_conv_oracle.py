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


def test_conv_oracle_threads_do_not_change_results():
    rng = np.random.default_rng(0)
    x = rng.standard_normal((33, 20, 6)).astype(np.float32)
    w = rng.standard_normal((3, 3, 6, 5)).astype(np.float32)
    assert np.array_equal(O.conv2d(x, w, threads=1), O.conv2d(x, w, threads=7))
