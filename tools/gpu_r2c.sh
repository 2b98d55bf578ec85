#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $O/bench_cfg2_default.json 2> $O/bench_cfg2_default.err; echo "bench default exit $?"; tail -3 $O/bench_cfg2_default.err
NVLLM_RESID_SPLITK=1 timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $O/bench_cfg2_splitk.json 2> $O/bench_cfg2_splitk.err; echo "bench splitk exit $?"; tail -3 $O/bench_cfg2_splitk.err
python - <<'PY'
import json
for f in ("bench_cfg2_default","bench_cfg2_splitk"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f,"value",round(d["value"]),"e2e",round(d["e2e"]["value"]),{k:(round(v,3) if isinstance(v,float) else v) for k,v in d["phases"].items()}); print(d["roofline"])
        for k,v in d["kernels"].items(): print("  ",k,v)
    except Exception as e: print(f,"ERR",e)
PY
