#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q --durations=5 > $O/pytest_final.log 2>&1; echo "pytest exit $?"; tail -10 $O/pytest_final.log
timeout 300 python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke exit $?"; tail -2 $O/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > $O/bench_cfg2_final.json 2> $O/bench_cfg2_final.err; echo "bench cfg2 exit $?"; tail -2 $O/bench_cfg2_final.err
timeout 600 python bench.py --config 5 --steps 10 --warmup 3 > $O/bench_cfg5_final.json 2> $O/bench_cfg5_final.err; echo "cfg5 exit $?"; tail -2 $O/bench_cfg5_final.err
python - <<'PY'
import json
for c in (2,5):
    try:
        d=json.loads(open(f"gpurun_out/bench_cfg{c}_final.json").read().strip().splitlines()[-1])
        ph=d["phases"]
        print(f"cfg{c} value {d['value']:.0f} e2e {d['e2e']['value']:.0f} ms/step {d['ms_per_step']:.2f} prefill_ms {ph['prefill_ms']:.2f} decode_us {ph.get('decode_us_per_step') or 0:.1f} roofline {d['roofline']['kernel']} {d['roofline']['frac']:.3f} parity {d.get('parity')}")
    except Exception as e: print("cfg",c,"ERR",e)
PY
