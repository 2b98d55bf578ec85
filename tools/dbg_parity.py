#!/usr/bin/env python
"""Debug aid: full-width parity variants, prints the per-step error instead of asserting."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from dataclasses import replace
import numpy as np
from helpers import C, W, make_oracle, oracle_mod, rel_err
import test_gpu_parity_full as T

def trial(tag, cfg, prec, B, pl, steps):
    prompts, tables = T._plain_batch(cfg, B, pl, steps)
    orc = make_oracle(cfg, round_bf16=(prec == C.PREC_BF16))
    rt, rl = T._oracle_batch(orc, prompts, steps)
    run = T._runner(cfg, prec, num_kv_blocks=max(max(t) for t in tables) + 1, max_num_seqs=max(B, 8))
    seqs = [[int(t) for t in p] for p in prompts]
    for i in range(steps):
        out, lg = run.run_flat(seqs, i == 0, [0]*B if i == 0 else None, tables, [0.0]*B, want_logits=True)
        errs = [rel_err(lg[s], rl[s][i]) for s in range(B)]
        print(tag, "step", i, "max err %.3e" % max(errs), "argmax bad seqs", [s for s in range(B) if errs[s] > 0.02][:8], flush=True)
        for s in range(B):
            seqs[s].append(int(rt[s][i]))
    run.close()

base = C.llama32_1b()
which = sys.argv[1:] or ["a", "b", "c", "d"]
if "a" in which: trial("L1 B1 bf16", replace(base, num_layers=1), C.PREC_BF16, 1, 16, 2)
if "b" in which: trial("L1 B8 bf16", replace(base, num_layers=1), C.PREC_BF16, 8, 16, 2)
if "c" in which: trial("L1 B8 fp32", replace(base, num_layers=1), C.PREC_FP32, 8, 16, 2)
if "d" in which: trial("L1 B64 bf16", replace(base, num_layers=1), C.PREC_BF16, 64, 16, 2)
if "e" in which: trial("L1 B8 bf16 smallV", replace(base, num_layers=1, vocab_size=32000), C.PREC_BF16, 8, 16, 2)
