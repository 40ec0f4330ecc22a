# quick probe: bf16x3 correctness + timing per op
import sys, tempfile, numpy as np
from pathlib import Path
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import oracle_lib as O
from simpleraytracer_b200 import api, graphdef_writer as gw
ctx = api.Context()
for (W, H, width, depth) in ((300, 37, 64, 4), (131, 5, 16, 3), (1920, 1080, 64, 6)):
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
