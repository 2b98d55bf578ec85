#!/usr/bin/env python
"""Times prefill and the decode loop of the cfg2 job (device-resident) — development aid.
NVLLM_SKIP / NVLLM_PDL env vars select ablations."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import nvllm_pkg
C = nvllm_pkg.sub("config")
model = sys.argv[1] if len(sys.argv) > 1 else "llama-3.2-1b"
batch = int(sys.argv[2]) if len(sys.argv) > 2 else 64
plen = int(sys.argv[3]) if len(sys.argv) > 3 else 512
dsteps = int(sys.argv[4]) if len(sys.argv) > 4 else 256
cfg = C.PRESETS[model]()
job = bench.Job(cfg, batch, plen, dsteps, 4321)
run = bench.make_runner(cfg, job, 0)
run.set_graphs(bool(int(os.environ.get('NVLLM_GRAPHS', '1'))))
job.stage(run)
for _ in range(2):
    job.run_resident(run)
run.sync()
res = []
for _ in range(3):
    run.timer_start(); run.rerun(); tp = run.timer_stop()
    run.timer_start()
    for _ in range(job.D):
        run.run_resident()
    td = run.timer_stop()
    res.append((tp, td))
tp = min(r[0] for r in res); td = min(r[1] for r in res)
print(json.dumps({"skip": os.environ.get("NVLLM_SKIP", "0"), "pdl": os.environ.get("NVLLM_PDL", "1"),
                  "prefill_ms": round(tp, 3), "decode_ms": round(td, 3),
                  "decode_us_per_step": round(1e3 * td / job.D, 1)}))
