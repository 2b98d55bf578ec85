#!/bin/bash
# 1-GPU box: bench lines for the given configs (default 1 3 5)
O=gpurun_out; mkdir -p $O
for c in ${@:-1 3 5}; do
timeout 900 python bench.py --config $c --steps 2 --warmup 3 --no-cpu-baseline > $O/bench_cfg$c.json 2> $O/bench_cfg$c.err; echo "cfg$c exit $?"; tail -3 $O/bench_cfg$c.err
done
python - "$@" <<'PY'
import json,sys
for c in (sys.argv[1:] or ["1","3","5"]):
    f=f"bench_cfg{c}"
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, "value",round(d["value"]),"e2e",round(d["e2e"]["value"]),d["phases"], d["roofline"])
        for k,v in d["kernels"].items(): print("  ",k,v)
    except Exception as e: print(f, "ERR", e)
PY
