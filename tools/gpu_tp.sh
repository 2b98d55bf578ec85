#!/bin/bash
# 2-GPU box: TP parity tests, then cfg4 (Falcon-7B) at tp=1 and tp=2
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_tp.py -m gpu -x -q > $O/pytest_tp.log 2>&1; echo "pytest tp exit $?"; tail -5 $O/pytest_tp.log
timeout 900 python bench.py --config 4 --steps 2 --warmup 3 --no-cpu-baseline > $O/bench_cfg4_tp1.json 2> $O/bench_cfg4_tp1.err; echo "cfg4 tp1 exit $?"; tail -3 $O/bench_cfg4_tp1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --config 4 --gpus 2 --steps 2 --warmup 3 --no-cpu-baseline > $O/bench_cfg4_tp2.json 2> $O/bench_cfg4_tp2.err; echo "cfg4 tp2 exit $?"; tail -3 $O/bench_cfg4_tp2.err
python - <<'PY'
import json
for f in ("bench_cfg4_tp1","bench_cfg4_tp2"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, "value",round(d["value"]),"e2e",round(d["e2e"]["value"]),d["phases"], d["roofline"])
        for k,v in d["kernels"].items(): print("  ",k,v)
    except Exception as e: print(f, "ERR", e)
PY
