#!/bin/bash
# A/B of environment switches on one box: usage gpu_ab.sh "<name>:<ENV=..,ENV=..>" ...  [CFG=2]
O=gpurun_out; mkdir -p $O
CFG=${CFG:-2}
for spec in "$@"; do
  name=${spec%%:*}; envs=${spec#*:}; envs=${envs//,/ }
  env $envs timeout 600 python bench.py --config $CFG --steps 3 --warmup 3 --no-cpu-baseline --no-tp > $O/ab_$name.json 2> $O/ab_$name.err; echo "$name exit $?"; tail -2 $O/ab_$name.err
done
python - "$@" <<'PY'
import json,sys
for spec in sys.argv[1:]:
    f=spec.split(":")[0]
    try:
        d=json.loads(open(f"gpurun_out/ab_{f}.json").read().strip().splitlines()[-1])
        ph=d["phases"]
        print(f"{f:14s} value {d['value']:9.0f} e2e {d['e2e']['value']:9.0f} prefill_ms {ph['prefill_ms']:7.2f} decode_us/step {ph.get('decode_us_per_step') or 0:8.1f}  roofline {d['roofline']['kernel']} {d['roofline']['frac']:.3f}")
        print("     ", {k:(v['ms'],v['launches']) for k,v in d["kernels"].items() if k.startswith('decode')})
    except Exception as e: print(f,"ERR",e)
PY
