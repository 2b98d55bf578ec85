"""Shared test helpers: package access, the tiny configs of the golden fixtures, oracle glue."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
import nvllm_pkg  # noqa: E402

C = nvllm_pkg.sub("config")
W = nvllm_pkg.sub("weights")
from harness import go_engine as E  # noqa: E402  (restated Go scheduler / block manager: test + bench harness)
GOLDEN = os.path.join(ROOT, "tests", "golden")

# must match tests/golden/make_hf_golden.py::CASES
GOLDEN_CASES = {
    "gpt2": lambda: C.tiny_gpt2(vocab_size=320, hidden=64, head_dim=16, ffn_dim=256, max_seq_len=64),
    "llama": lambda: C.tiny_llama(vocab_size=320, hidden=64, num_heads=4, num_kv_heads=2,
                                  head_dim=16, ffn_dim=128, max_seq_len=64),
    "granite": lambda: C.tiny_granite(vocab_size=320, hidden=64, num_heads=4, num_kv_heads=2,
                                      head_dim=16, ffn_dim=32, max_seq_len=64),
    "falcon": lambda: C.tiny_falcon(vocab_size=320, hidden=5 * 16, num_heads=5, head_dim=16,
                                    ffn_dim=4 * 5 * 16, max_seq_len=64),
    # Granite-4 hybrid (Mamba2, attention, Mamba2, Mamba2) against GraniteMoeHybridForCausalLM
    "hybrid": lambda: C.tiny_hybrid(vocab_size=320, hidden=64, num_heads=4, num_kv_heads=2,
                                    head_dim=16, ffn_dim=96, max_seq_len=64, mamba_num_heads=8,
                                    mamba_head_dim=16, mamba_state_size=16, mamba_n_groups=2,
                                    mamba_conv_kernel=4, norm_eps=1e-5),
}

# parity-test configs (bigger than the goldens, still seconds on the CPU oracle)
TINY = {
    "gpt2": lambda: C.tiny_gpt2(),
    "llama": lambda: C.tiny_llama(),
    "granite": lambda: C.tiny_granite(),
    "falcon": lambda: C.tiny_falcon(),
}

# head_dim 64 / 128 siblings: in bf16 mode these take the tensor-core attention kernels
# (attention_mma.cu); G = query heads per kv head
TINY_HD = {
    "gpt2_hd64": lambda: C.tiny_gpt2(hidden=128, num_heads=2, num_kv_heads=2, head_dim=64),       # G=1
    "llama_hd64": lambda: C.tiny_llama(hidden=256, num_heads=4, num_kv_heads=1, head_dim=64),     # G=4
    "llama_hd128": lambda: C.tiny_llama(hidden=256, num_heads=2, num_kv_heads=1, head_dim=128),   # G=2
    "granite_hd64": lambda: C.tiny_granite(hidden=128, num_heads=2, num_kv_heads=1, head_dim=64),
    "falcon_g18": lambda: C.tiny_falcon(hidden=18 * 64, num_heads=18, head_dim=64,
                                        ffn_dim=2 * 18 * 64),                                    # G=18
}


def oracle_mod():
    from oracle import oracle as O
    return O


def make_oracle(cfg, seed=1234, round_bf16=False):
    O = oracle_mod()
    return O.OracleModel(cfg, W.random_weights(cfg, seed, round_bf16=round_bf16))


def rel_err(a, b):
    """max |a-b| / max |b| — the gate of BASELINE.json's north_star."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def ref_filter_probs(logits, history, temp, top_p, top_k, rep):
    """numpy float64 restatement of the filter pipeline — only used to measure how far a draw
    is from a CDF step / a filter cut (the token itself is checked against the C oracle)."""
    lc = logits.astype(np.float64).copy()
    if rep != 1.0 and len(history):
        cnt = {}
        for i, t in enumerate(history):
            cnt[t] = cnt.get(t, 0) + (3 if i >= len(history) - 10 else 1)
        for t, c in cnt.items():
            lc[t] = lc[t] / (rep * c) if lc[t] > 0 else lc[t] * (rep * c)
    if temp > 0 and temp != 1.0:
        lc /= temp
    p = np.exp(lc - lc.max())
    p /= p.sum()
    order = np.lexsort((np.arange(p.size), -p))
    if 0 < top_k < p.size:
        keep = order[:top_k]
        q = np.zeros_like(p)
        q[keep] = p[keep]
        p = q
        order = np.lexsort((np.arange(p.size), -p))
    if top_p < 1.0:
        cum = np.cumsum(p[order])
        cut = int(np.searchsorted(cum, top_p, side="left")) + 1
        q = np.zeros_like(p)
        q[order[:cut]] = p[order[:cut]]
        p = q
    return p / p.sum()
