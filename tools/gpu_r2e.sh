#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench_cfg2_full.json 2> $O/bench_cfg2_full.err; echo "bench cfg2 full exit $?"; tail -3 $O/bench_cfg2_full.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_cfg2_full.json").read().strip().splitlines()[-1])
print("value",round(d["value"]),"e2e",round(d["e2e"]["value"])); print(d["phases"]); print(d["roofline"]); print("parity",d["parity"]); print("tp",d["tp"]); print("cpu",d["cpu_baseline"])
PY
for c in 1 3; do
timeout 600 python bench.py --config $c --steps 10 --warmup 3 --no-cpu-baseline > $O/bench_cfg$c.json 2> $O/bench_cfg$c.err; echo "cfg$c exit $?"; tail -2 $O/bench_cfg$c.err
done
python - <<'PY'
import json
for c in (1,3):
    d=json.loads(open(f"gpurun_out/bench_cfg{c}.json").read().strip().splitlines()[-1])
    print("cfg",c,"value",round(d["value"]),"e2e",round(d["e2e"]["value"])); print(d["phases"]); print(d["roofline"])
    for k,v in d["kernels"].items(): print("  ",k,v)
PY
