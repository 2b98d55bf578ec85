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
if os.environ.get("NVLLM_NCU_LAYERS"):
    from dataclasses import replace
    cfg = replace(cfg, num_layers=int(os.environ["NVLLM_NCU_LAYERS"]))
job = bench.Job(cfg, batch, plen, steps, 4321)
run = bench.make_runner(cfg, job, 0, fast_weights=not os.environ.get("NVLLM_NCU_HOST_WEIGHTS"))
# ncu serialises and replays kernels: capture-free launches keep -k / -s / -c counting simple
run.set_graphs(bool(int(os.environ.get('NVLLM_GRAPHS', '0'))))
job.stage(run)
for _ in range(steps):
    run.run_resident()
run.sync()
if os.environ.get("NVLLM_NCU_SAMPLING"):
    # one nvllm_run decode step drawn by the device-side SampleWithHistory (top-k / top-p / penalty)
    import numpy as np
    run.set_sampling_params(temperature=0.8, top_p=0.9, top_k=40, repetition_penalty=1.2)
    toks = [[int(t) for t in job.prompts[i]] + [1] for i in range(job.n)]
    run.run_flat(toks, False, None, job.tables, [0.8] * job.n, [0.5] * job.n)
    run.sync()
print("ok")
