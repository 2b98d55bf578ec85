#!/usr/bin/env python
"""Device timeline of a few decode steps (nvllm_trace_*): per launch, when the first CTA
entered, when the dependency wait released, when the last CTA left.  Development aid; the
summary it prints is what profiles/*_timeline.md holds."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
import nvllm_pkg
C = nvllm_pkg.sub("config")
model = sys.argv[1] if len(sys.argv) > 1 else "llama-3.2-1b"
batch = int(sys.argv[2]) if len(sys.argv) > 2 else 64
plen = int(sys.argv[3]) if len(sys.argv) > 3 else 512
cfg = C.PRESETS[model]()
job = bench.Job(cfg, batch, plen, 16, 4321)
run = bench.make_runner(cfg, job, 0)
job.stage(run)
job.run_resident(run)
job.stage(run)
for _ in range(3):
    job.run_resident(run)
job.stage(run)
for _ in range(4):
    run.run_resident()
run.sync()
run.trace_begin(1 << 17)
for _ in range(2):
    run.run_resident()
rec = run.trace_end(1 << 17)
names = {1: "gemm", 2: "norm", 3: "rope_kv", 4: "attn_dec"}
aux = rec[rec[:, 1] == 9]
rec = rec[rec[:, 1] != 9]
seqs = sorted(set(int(x) for x in rec[:, 0]))
t00 = int(rec[:, 4].min())
rows = []
for sq in seqs:
    r = rec[rec[:, 0] == sq]
    kid = int(r[0, 1])
    ent0, ent1 = int(r[:, 4].min()) - t00, int(r[:, 4].max()) - t00
    w = r[:, 5][r[:, 5] > 0]
    wait0 = (int(w.min()) - t00) if len(w) else -1
    ex0, ex1 = int(r[:, 6].min()) - t00, int(r[:, 6].max()) - t00
    ex = r[:, 7]
    extra = ""
    if kid == 1:
        a = ((ex >> np.uint64(42)) & np.uint64(0x1fffff)).astype(np.int64)
        b = ((ex >> np.uint64(21)) & np.uint64(0x1fffff)).astype(np.int64)
        c = (ex & np.uint64(0x1fffff)).astype(np.int64)
        d = (r[:, 6].astype(np.int64) - r[:, 5].astype(np.int64))
        extra = (f" per-CTA median: mma_done {np.median(a)/1e3:.2f} epi_seen {np.median(b)/1e3:.2f} "
                 f"epi_done {np.median(c)/1e3:.2f} exit {np.median(d)/1e3:.2f}")
        ax = aux[aux[:, 0] == sq]
        if len(ax):
            e7 = ax[:, 7]
            extra += f" | first_full {np.median(ax[:,4].astype(np.int64))/1e3:.2f} last_full {np.median(ax[:,5].astype(np.int64))/1e3:.2f}"
            if os.environ.get("TIMELINE_CLUSTER") and len(r) in (96, 112):  # cluster epilogue stamps (16-bit ns)
                f16 = [((e7 >> np.uint64(sh)) & np.uint64(0xffff)).astype(np.int64) for sh in (48, 32, 16, 0)]
                extra += " cluster: " + " ".join(f"{nm} {np.median(v)/1e3:.2f}/{np.max(v)/1e3:.2f}" for nm, v in zip(("staged", "landed", "rows_done", "-"), f16)) + " (median/max)"
            else:
                extra += f" ldtm0 {np.median((e7 >> np.uint64(32)).astype(np.int64))/1e3:.2f} st0 {np.median((e7 & np.uint64(0xffffffff)).astype(np.int64))/1e3:.2f}"
    if kid == 4:
        d = (r[:, 6].astype(np.int64) - np.maximum(r[:, 5], r[:, 4]).astype(np.int64)) / 1e3
        extra = " per-CTA dur p10/50/90/max: " + "/".join(f"{np.percentile(d, q):.1f}" for q in (10, 50, 90, 100))
        extra += f" SMs used {len(set(int(x) for x in r[:,3]))}"
        a_ = ((ex >> np.uint64(42)) & np.uint64(0x1fffff)).astype(np.int64)
        b_ = ((ex >> np.uint64(21)) & np.uint64(0x1fffff)).astype(np.int64)
        c_ = (ex & np.uint64(0x1fffff)).astype(np.int64)
        extra += (f" | after wait: prologue_done {np.median(a_)/1e3:.2f} first_tile {np.median(b_)/1e3:.2f} "
                  f"head0_tiles_done {np.median(c_)/1e3:.2f}")
    if kid == 2:
        extra = (f" per-CTA median: loads_done {np.median((ex >> np.uint64(32)).astype(np.int64))/1e3:.2f} "
                 f"reduce_done {np.median((ex & np.uint64(0xffffffff)).astype(np.int64))/1e3:.2f} "
                 f"exit {np.median(r[:, 6].astype(np.int64) - r[:, 5].astype(np.int64))/1e3:.2f}")
    rows.append((sq, names.get(kid, str(kid)), len(r), ent0, ent1, wait0, ex0, ex1, extra))
print("seq kernel ctas | first_entry last_entry | wait_release | first_exit last_exit | dur_after_wait  (us)")
prev_exit = 0
for (sq, nm, n, e0, e1, w0, x0, x1, extra) in rows[:40]:
    print(f"{sq:4d} {nm:9s} {n:4d} | {e0/1e3:8.2f} {e1/1e3:8.2f} | {w0/1e3:8.2f} | {x0/1e3:8.2f} {x1/1e3:8.2f} | "
          f"{(x1-w0)/1e3:6.2f}  gap_from_prev_exit {(w0-prev_exit)/1e3:6.2f}{extra}")
    prev_exit = x1
