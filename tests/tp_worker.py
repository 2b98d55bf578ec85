"""One rank of a tensor-parallel parity run (launched by tests/test_tp.py, once per GPU).

usage: tp_worker.py <rank> <tp> <id_file> <case> <prec>

Every rank is handed the same FULL random-init tensors and the same Run calls (that is the
contract of include/nvllm.h: the library shards on its own); every rank must return the same
tokens and logits, equal to the single-device CPU oracle within the precision's gate.
"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from helpers import C, W, make_oracle, nvllm_pkg, oracle_mod, rel_err  # noqa: E402

CASES = {
    # Falcon family (MQA, parallel block): 7 heads pad to 8 for tp 2 / 4 — the 71 -> 72 case
    "falcon": lambda: C.tiny_falcon(),
    "falcon_g18": lambda: C.tiny_falcon(hidden=18 * 64, num_heads=18, head_dim=64, ffn_dim=2 * 18 * 64),
    # sequential blocks: two sums over ranks per layer
    "llama": lambda: C.tiny_llama(hidden=256, num_heads=8, num_kv_heads=4, head_dim=32, ffn_dim=512),
    "gpt2": lambda: C.tiny_gpt2(hidden=128, num_heads=4, num_kv_heads=4, head_dim=32, ffn_dim=512),
}
TOL = {C.PREC_FP32: 1e-4, C.PREC_BF16: 2e-2}


def main():
    rank, tp, id_file, case, prec = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], sys.argv[4], int(sys.argv[5])
    L = nvllm_pkg.sub("lib")
    R = nvllm_pkg.sub("runner")
    if rank == 0:
        tid = L.tp_unique_id()
        with open(id_file + ".tmp", "wb") as f:
            f.write(tid)
        os.replace(id_file + ".tmp", id_file)
    else:
        t0 = time.time()
        while not os.path.exists(id_file):
            if time.time() - t0 > 120:
                raise TimeoutError("no tp id from rank 0")
            time.sleep(0.05)
        tid = open(id_file, "rb").read()
    cfg = CASES[case]()
    bf16 = prec == C.PREC_BF16
    run = R.CudaModelRunner(cfg, W.random_weights(cfg, 1234, round_bf16=bf16), precision=prec,
                            num_kv_blocks=16, device=rank, tp_size=tp, tp_rank=rank, tp_id=tid,
                            sampling=None)
    orc = make_oracle(cfg, round_bf16=bf16)
    O = oracle_mod()
    # two sequences of different length in one ragged batch, crossing a KV block boundary
    prompts = [W.random_prompt(cfg.vocab_size, 250, 4321), W.random_prompt(cfg.vocab_size, 37, 4322)]
    steps = 8
    bts = [[0, 1], [2]]
    if os.environ.get("TP_SINGLE"):
        # ONE sequence long enough for the two-chunk prefill pipeline (NVLLM_TP_PIPE_MIN_T)
        prompts, steps, bts = [W.random_prompt(cfg.vocab_size, 300, 4321)], 4, [[0, 1]]
    refs = [orc.greedy(p, steps) for p in prompts]
    toks = [[int(t) for t in p] for p in prompts]
    worst = 0.0
    for i in range(steps):
        out, logits = run.run_flat(toks, is_prefill=(i == 0), num_cached=[0] * len(toks), block_tables=bts,
                                   temperature=[0.0] * len(toks), want_logits=True)
        for s in range(len(toks)):
            ref_toks, ref_logits = refs[s]
            err = rel_err(logits[s], ref_logits[i])
            worst = max(worst, err)
            assert err <= TOL[prec], f"rank {rank} {case} seq {s} step {i}: rel err {err}"
            assert int(out[s]) == O.argmax(logits[s])
            if int(out[s]) != int(ref_toks[i]):
                margin = float(ref_logits[i][ref_toks[i]] - ref_logits[i][out[s]])
                scale = float(np.abs(ref_logits[i]).max())
                assert bf16 and margin <= 2 * err * scale + 1e-6, \
                    f"rank {rank} {case} seq {s} step {i}: token {out[s]} != {ref_toks[i]}"
            toks[s].append(int(ref_toks[i]))
    # temperature sampling through the gathered (rank-major) logits
    uni = [0.3, 0.9][:len(toks)]
    out, logits = run.run_flat(toks, is_prefill=False, num_cached=None, block_tables=bts,
                               temperature=[0.8, 1.3][:len(toks)], uniforms=uni, want_logits=True)
    for s in range(len(toks)):
        assert int(out[s]) == O.sample(logits[s], None, [0.8, 1.3][s], u=uni[s]), "sampled token differs"
    prof_ok = run.last_stats.kernel_launches > 0
    run.close()
    assert prof_ok
    print(f"TP_OK rank={rank} tp={tp} case={case} prec={prec} worst={worst:.3e}", flush=True)


if __name__ == "__main__":
    main()
