#!/bin/bash
# Final evidence run on one GPU: whole GPU test-suite, smoke(), bench lines of cfg1/2/3/5 with >= 10 steps
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q --durations=10 > $O/pytest_final.log 2>&1; echo "pytest exit $?"; tail -16 $O/pytest_final.log
timeout 300 python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke exit $?"; tail -4 $O/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > $O/bench_cfg2_final.json 2> $O/bench_cfg2_final.err; echo "bench cfg2 exit $?"; tail -2 $O/bench_cfg2_final.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo "reference exit $?"
for c in 1 3 5; do
timeout 600 python bench.py --config $c --steps 10 --warmup 3 > $O/bench_cfg${c}_final.json 2> $O/bench_cfg${c}_final.err; echo "cfg$c exit $?"; tail -2 $O/bench_cfg${c}_final.err
done
python - <<'PY'
import json
for c in (2,1,3,5):
    try:
        d=json.loads(open(f"gpurun_out/bench_cfg{c}_final.json").read().strip().splitlines()[-1])
        ph=d["phases"]
        print(f"cfg{c} value {d['value']:.0f} e2e {d['e2e']['value']:.0f} ms/step {d['ms_per_step']:.2f} prefill_ms {ph['prefill_ms']:.2f} decode_us {ph.get('decode_us_per_step') or 0:.1f} roofline {d['roofline']['kernel']} {d['roofline']['frac']:.3f} parity {d.get('parity')}")
        if c == 2: print("  tp", d["tp"])
    except Exception as e: print("cfg",c,"ERR",e)
PY
