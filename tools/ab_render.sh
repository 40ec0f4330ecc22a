# e2e of a sharded frame: one work-list render + pipelined copy (BENCH_E2E_CHUNKS=1) against a 4-chunk render
N=${1:-2}
A="--steps 5 --warmup 3 --no-cpu-baseline --no-inference --no-brute"
BN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29547 bench.py --gpus $N $A"
P='import sys,json; d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith("{")][-1]); c4=d.get("config4") or {}; print(sys.argv[1], "N=%d" % d["n_gpus"], "dev", round(d["value"]), "e2e", round(d["e2e"]["value"]), "pinned", round(d["e2e"]["value_page_locked_buffer"]), d["e2e"]["frame_equal_single_gpu"], "| config4 dev", round(c4.get("value",0)), "e2e", round(c4.get("e2e",{}).get("value",0)), "pinned", round(c4.get("e2e",{}).get("value_page_locked_buffer",0)), c4.get("e2e",{}).get("frame_equal_single_gpu"))'
python -m pytest tests/test_multi_gpu.py tests/test_model_gpu.py -m gpu -x -q 2>&1 | tail -2
python bench.py $A 2>/dev/null | python -c "$P" N1
$BN 2>/dev/null | python -c "$P" chunks1
BENCH_E2E_CHUNKS=4 $BN 2>/dev/null | python -c "$P" chunks4
