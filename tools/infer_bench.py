#!/usr/bin/env python
"""Times mlInfer of the BUILD-DEFINED 1080p denoiser (config 5 network) with device-resident images.
    python tools/infer_bench.py [--width 64] [--depth 6] [--w 1920] [--h 1080] [--precision bf16|fp32] [--iters 10]
"""
import argparse
import sys
import tempfile
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from simpleraytracer_b200 import api, graphdef_writer as gw  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--width", type=int, default=64)
    ap.add_argument("--depth", type=int, default=6)
    ap.add_argument("--w", type=int, default=1920)
    ap.add_argument("--h", type=int, default=1080)
    ap.add_argument("--precision", default="bf16")
    ap.add_argument("--iters", type=int, default=10)
    a = ap.parse_args()
    ctx = api.Context()
    with tempfile.TemporaryDirectory() as d:
        path = Path(d) / "net.pb"
        gw.make_denoiser(path, channels=3, width=a.width, depth=a.depth, seed=0)
        model = ctx.create_model(path)
    i, o = model.info()
    i.width, i.height = a.w, a.h
    model.set_input_info(i)
    model.set_precision(api.MLX_PRECISION_BF16 if a.precision == "bf16" else api.MLX_PRECISION_FP32)
    inp = ctx.create_image(a.w, a.h, 3); out = ctx.create_image(a.w, a.h, 3)
    inp.write(np.random.default_rng(0).random((a.h, a.w, 3), dtype=np.float32))
    inp.device_pointer(False)          # upload once ...
    inp.device_pointer(True)           # ... and make the HBM copy authoritative: no H2D inside mlInfer
    for _ in range(3):
        model.infer(inp, out)
    t0 = time.perf_counter()
    for _ in range(a.iters):
        model.infer(inp, out)
    dt = (time.perf_counter() - t0) / a.iters
    px = a.w * a.h
    flop = px * 2 * 9 * (3 * a.width + (a.depth - 2) * a.width * a.width + a.width * 3)
    print(f"{a.precision} width={a.width} depth={a.depth} {a.w}x{a.h}: {dt*1e3:.3f} ms/image, {1/dt:.1f} images/s, "
          f"{flop/dt/1e12:.1f} TFLOP/s algorithmic, tc_layers={model.tensor_core_layers()} launches={model.launch_count()}")
    prof = model.profile(inp, out)
    print("per op ms:", " ".join(f"{ms:.4f}" for ms, _, _ in prof),
          "| TFLOP/s:", " ".join(f"{fl/ms/1e9:.0f}" for ms, fl, _ in prof if ms > 0),
          "| GB/s:", " ".join(f"{by/ms/1e6:.0f}" for ms, _, by in prof if ms > 0))


if __name__ == "__main__":
    main()
