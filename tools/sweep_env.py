#!/usr/bin/env python
"""Runs bench.py once per value of an environment knob and prints one summary line each.
    python tools/sweep_env.py SRT_BVH_LEAF 1 2 4 8 -- --spp 16 --steps 3 --no-cpu-baseline --no-brute
"""
import json
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def main():
    args = sys.argv[1:]
    split = args.index("--") if "--" in args else len(args)
    knob, values, bench_args = args[0], args[1:split], args[split + 1:]
    for v in values:
        env = dict(os.environ)
        env[knob] = v
        r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--no-inference"] + bench_args, env=env, capture_output=True, text=True)
        try:
            d = json.loads(r.stdout.strip().splitlines()[-1])
            rl = d["roofline"]
            print(f"{knob}={v} kernel_ms={rl['kernel_ms']:.3f} Mrays/s={d['value']:.0f} box={rl['box_tests']:.3g} "
                  f"tri={rl['tri_tests']:.3g} nodes={d['config']['bvh_nodes']} build_ms={d['config']['bvh_build_ms']:.0f} "
                  f"light_grid_ms={d['config'].get('light_grid_build_ms', 0):.2f} e2e={d['e2e']['value']:.0f}",
                  flush=True)
        except Exception as e:  # noqa: BLE001
            print(f"{knob}={v} FAILED {e}: {r.stderr[-400:]}", flush=True)


if __name__ == "__main__":
    main()
