#!/usr/bin/env python
"""Picks the prompt seeds of the free-running greedy parity tests (tests/test_gpu_parity*.py).

Random-init logits are nearly flat, so some prompts have greedy steps whose top-2 margin is
smaller than the bf16 tolerance (2e-2 of max|logit|): a correct bf16 implementation may then
legitimately pick the runner-up and a FREE-RUNNING comparison diverges.  Instead of forgiving
token flips, the tests use, per model, the first prompt seed >= 4321 for which the oracle's own
top-2 margin stays >= MARGIN * max|logit| on every one of the compared steps (fp32 weights and
bf16-rounded weights both).  This script prints that table; it is committed next to the
fixtures so the choice is reproducible:

    python tests/golden/find_margin_seeds.py            # tiny configs + full-size GPT-2 small
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import C, W, TINY, TINY_HD, make_oracle  # noqa: E402

MARGIN = 0.02


def min_margin(orc, prompt, steps):
    _t, logs = orc.greedy(prompt, steps)
    worst = 1e9
    for lg in logs:
        top = np.partition(lg, -2)[-2:]
        worst = min(worst, float(top[1] - top[0]) / float(np.abs(lg).max()))
    return worst


def find(cfg, prompt_len, steps, start=4321, tries=200):
    orcs = [make_oracle(cfg, round_bf16=False), make_oracle(cfg, round_bf16=True)]
    for seed in range(start, start + tries):
        p = W.random_prompt(cfg.vocab_size, prompt_len, seed)
        m = min(min_margin(o, p, steps) for o in orcs)
        if m >= MARGIN:
            return seed, m
    return None, 0.0


if __name__ == "__main__":
    cases = [(n, f(), 40, 32) for n, f in TINY.items()]
    cases += [(n, f(), 100, 12) for n, f in TINY_HD.items()]
    cases += [("gpt2_small_full", C.gpt2_small(), 32, 64)]
    for name, cfg, pl, steps in cases:
        seed, m = find(cfg, pl, steps)
        print(f"{name:18s} prompt_len {pl:4d} steps {steps:3d} -> seed {seed} (min margin {m:.4f} of max|logit|)")
