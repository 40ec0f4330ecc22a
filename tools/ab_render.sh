# render launch order at 1 and N ranks: default (work list when sharded), forced work list, natural order
N=${1:-2}
A="--steps 5 --warmup 3 --no-cpu-baseline --no-inference --no-brute --no-extra-configs"
B1="python bench.py $A"
BN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29547 bench.py --gpus $N $A"
P='import sys,json; d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith("{")][-1]); b=d["config"].get("n%d_breakdown_ms" % d["n_gpus"], {}); print(sys.argv[1], "N=%d" % d["n_gpus"], round(d["value"]), round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"]), "grid", round(d["config"]["screen_grid_build_ms"],3), "spans", [round(x,3) for x in b.get("per_rank_span", [])], d["frame_check"]["equal"] if d["frame_check"] else None)'
python -m pytest tests/test_render_gpu.py tests/test_multi_gpu.py -m gpu -x -q 2>&1 | tail -2
$B1 2>/dev/null | python -c "$P" default
SRT_BLOCK_ORDER=2 $B1 2>/dev/null | python -c "$P" forced_work_list
$BN 2>/dev/null | python -c "$P" default
SRT_BLOCK_TRACE=1 $BN 2>&1 >/dev/null | grep "srt trace\]" | tail -8 | cut -c1-260
