#!/bin/bash
# end-of-round check at HEAD: smoke(), the default bench line (cfg2), the reference arm, cfg1
O=gpurun_out; mkdir -p $O
timeout 300 python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke exit $?"; tail -2 $O/smoke.log
timeout 900 python bench.py > $O/bench_cfg2_final.json 2> $O/bench_cfg2_final.err; echo "bench default exit $?"; tail -2 $O/bench_cfg2_final.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo "reference exit $?"; tail -c 400 $O/bench_reference.json
timeout 600 python bench.py --config 1 --steps 10 --warmup 3 --no-tp > $O/bench_cfg1_final.json 2> $O/bench_cfg1_final.err; echo "cfg1 exit $?"
python - <<'PY'
import json
for c in (2,1):
    try:
        d=json.loads(open(f"gpurun_out/bench_cfg{c}_final.json").read().strip().splitlines()[-1])
        ph=d["phases"]; p=d.get("parity") or {}
        print(f"cfg{c} value {d['value']:.0f} e2e {d['e2e']['value']:.0f} ms/step {d['ms_per_step']:.2f} prefill_ms {ph['prefill_ms']:.2f} decode_us {ph.get('decode_us_per_step') or 0:.1f} roofline {d['roofline']['kernel']} {d['roofline']['frac']:.3f} launches {d.get('gpu_launches')} parity err {p.get('max_rel_err')} tokens_equal {p.get('tokens_equal')} clocks {d.get('clocks')}")
    except Exception as e: print("cfg",c,"ERR",e)
PY
