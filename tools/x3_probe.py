"""Per-op times and error of the three inference precisions (fp32 CUDA cores, bf16x3, bf16) on a few
networks.  `python tools/x3_probe.py big` runs only the 1080p benchmark net (use with MLB200_TC_PROF=1
for the per-role cycle counters of the tensor-core kernels)."""
import sys, tempfile, numpy as np
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import oracle_lib as O
from simpleraytracer_b200 import api, graphdef_writer as gw
ctx = api.Context()
CASES = ((300, 37, 64, 4), (131, 5, 16, 3), (1920, 1080, 64, 6))
if len(sys.argv) > 1 and sys.argv[1] == "big":
    CASES = CASES[2:]
for (W, H, width, depth) in CASES:
    with tempfile.TemporaryDirectory() as d:
        path = Path(d) / "net.pb"
        _, layers = gw.make_denoiser(path, channels=3, width=width, depth=depth, seed=0)
        model = ctx.create_model(path)
    i, _ = model.info(); i.width, i.height = W, H; model.set_input_info(i)
    x = np.random.default_rng(3).random((H, W, 3), dtype=np.float32)
    inp = ctx.create_image(W, H, 3); out = ctx.create_image(W, H, 3); inp.write(x)
    ref = O.run_denoiser(x, layers, alpha=0.1) if W < 1000 else None
    model.infer(inp, out); f32 = out.read().copy()
    if ref is None: ref = f32
    scale = float(np.abs(ref).max())
    for prec, name in ((0, "fp32"), (2, "bf16x3"), (1, "bf16")):
        model.set_precision(prec)
        model.infer(inp, out)
        got = out.read().copy()
        ops = model.profile(inp, out)
        print(f"{W}x{H} w{width} d{depth} {name}: tc_layers={model.tensor_core_layers()} err={np.abs(got-ref).max()/scale:.3e} "
              f"ops_ms={[round(o[0],4) for o in ops]} sum={sum(o[0] for o in ops):.3f}", flush=True)
