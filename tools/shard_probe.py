"""One rank's share of the config-3 frame on ONE GPU (interleaved 8-row blocks, stride N, phase P), rendered a few
times: lets `ncu --metrics gpu__time_duration.sum --cache-control none --clock-control none` time every launch of
the sharded frame -- the five grid-build launches and the work-list render -- without a multi-rank command.

    python tools/shard_probe.py [--stride 8] [--phase 0] [--frames 6] [--workload config3|config4]
prints the library's own event times of the last frame (grid build, render span)."""
import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--stride", type=int, default=8)
    ap.add_argument("--phase", type=int, default=0)
    ap.add_argument("--frames", type=int, default=6)
    ap.add_argument("--workload", default="config3")
    a = ap.parse_args()
    import torch
    from simpleraytracer_b200 import api
    W, H, spp, tris = (1920, 1080, 64, 100000) if a.workload == "config3" else (3840, 2160, 1, 1000000)
    ctx = api.Context(0)
    v, alb, cam, light = api.gen_mesh(1, tris, W, H)      # bench.py: mesh_seed 1, sample_seed 1
    scene = ctx.create_scene(v, alb)
    p = api.make_params(W, H, spp=spp, jitter=1, seed=1, shade=api.SRT_SHADE_SHADOW, accel=api.SRT_ACCEL_BVH,
                        background=(0.5, 0.7, 0.9), row_block_stride=a.stride if a.stride > 1 else 0,
                        row_block_phase=a.phase)
    rgb = torch.zeros(H, W, 3, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    for f in range(a.frames):
        scene.render_device(cam, light, p, rgb.data_ptr(), 0, 0, s)
        st = scene.stats()
        torch.cuda.synchronize()
    print({k: getattr(st, k) for k in ("kernel_ms", "grid_build_ms", "light_grid_build_ms")})
    scene.release()


if __name__ == "__main__":
    main()
