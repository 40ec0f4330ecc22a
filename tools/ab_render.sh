A="--steps 5 --warmup 3 --no-cpu-baseline --no-inference --no-brute"
P='import sys,json; d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith("{")][-1]); c4=d.get("config4") or {}; print(sys.argv[1], round(d["value"]), round(d["ms_per_step"],4), "grid_ms", round(d["config"]["screen_grid_build_ms"],4), "| config4", round(c4.get("value",0)), round(c4.get("ms_per_step",0),4), "grid_ms", round(c4.get("screen_grid_build_ms",0),4), "| config2 us", round((d.get("config2") or {}).get("closest_hit",{}).get("us_per_frame_median",0),1))'
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py $A 2>/dev/null | python -c "$P" default
