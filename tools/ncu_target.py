#!/usr/bin/env python
"""Smallest program that reaches steady-state decode steps of the cfg2 job: one prefill, then
N decode steps (device-resident).  Target for `ncu -k regex:... -s ... -c ...`."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
import nvllm_pkg
C = nvllm_pkg.sub("config")
model = sys.argv[1] if len(sys.argv) > 1 else "llama-3.2-1b"
batch = int(sys.argv[2]) if len(sys.argv) > 2 else 64
plen = int(sys.argv[3]) if len(sys.argv) > 3 else 512
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 3
cfg = C.PRESETS[model]()
job = bench.Job(cfg, batch, plen, steps, 4321)
run = bench.make_runner(cfg, job, 0)
# ncu serialises and replays kernels: capture-free launches keep -k / -s / -c counting simple
run.set_graphs(bool(int(os.environ.get('NVLLM_GRAPHS', '0'))))
job.stage(run)
for _ in range(steps):
    run.run_resident()
run.sync()
print("ok")
