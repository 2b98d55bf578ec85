#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q --durations=8 > $O/pytest.log 2>&1; echo "pytest exit $?"; tail -12 $O/pytest.log
timeout 600 python bench.py --steps 3 --warmup 3 > $O/bench_cfg2.json 2> $O/bench_cfg2.err; echo "bench exit $?"; tail -3 $O/bench_cfg2.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_cfg2.json").read().strip().splitlines()[-1])
print("value",round(d["value"]),"e2e",round(d["e2e"]["value"]),d["phases"]); print(d["roofline"]); print(d["parity"])
for k,v in d["kernels"].items(): print("  ",k,v)
PY
