#!/bin/bash
# quick GPU loop: tests + short bench; outputs under gpurun_out/
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest.log; tail -12 gpurun_out/pytest.log
timeout 400 python bench.py --steps 2 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -3 gpurun_out/bench.err
python - <<'PY'
import json
d=json.load(open("gpurun_out/bench.json"))
print("value",round(d["value"]),"e2e",round(d["e2e"]["value"]),d["phases"])
for k,v in d["kernels"].items(): print(k,v)
print(d["roofline"])
PY
