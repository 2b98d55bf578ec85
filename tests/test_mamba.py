"""SURVEY.md §8f rank 4 — Granite-4 hybrid (Mamba2 + attention blocks): the CUDA path against the
oracle's restatement of purego/tensor/mamba2.go (itself pinned to an independent numpy
restatement in tests/test_oracle.py).  Ragged batches, decode steps that carry the SSM state
through the per-KV-block snapshots, a sequence crossing a 256-token block boundary."""
import numpy as np
import pytest

from helpers import C, W, make_oracle, oracle_mod, rel_err
import nvllm_pkg

pytestmark = pytest.mark.gpu
TOL = {C.PREC_FP32: 1e-4, C.PREC_BF16: 2e-2}


def _run(cfg, prec, prompts, steps, tables, num_kv_blocks, **kw):
    R = nvllm_pkg.sub("runner")
    bf16 = prec == C.PREC_BF16
    orc = make_oracle(cfg, round_bf16=bf16)
    refs = [orc.greedy(p, steps) for p in prompts]
    run = R.CudaModelRunner(cfg, W.random_weights(cfg, 1234, round_bf16=bf16), precision=prec,
                            num_kv_blocks=num_kv_blocks, sampling=None, **kw)
    worst, flips = 0.0, 0
    try:
        seqs = [[int(t) for t in p] for p in prompts]
        n = len(seqs)
        for i in range(steps):
            out, logits = run.run_flat(seqs, i == 0, [0] * n if i == 0 else None, tables, [0.0] * n,
                                       want_logits=True)
            for s in range(n):
                ref = refs[s][1][i]
                err = rel_err(logits[s], ref)
                worst = max(worst, err)
                assert err <= TOL[prec], f"seq {s} step {i}: rel err {err}"
                assert int(out[s]) == oracle_mod().argmax(logits[s])
                if int(out[s]) != int(refs[s][0][i]):
                    margin = float(ref[refs[s][0][i]] - ref[out[s]])
                    assert bf16 and margin <= 2 * err * float(np.abs(ref).max()) + 1e-6
                    flips += 1
                seqs[s].append(int(refs[s][0][i]))
        assert flips <= 1
    finally:
        run.close()
    return worst


@pytest.mark.parametrize("prec", [C.PREC_FP32, C.PREC_BF16])
def test_hybrid_single_sequence(prec):
    cfg = C.tiny_hybrid()
    _run(cfg, prec, [W.random_prompt(cfg.vocab_size, 40, 1)], 6, [[0]], 2)


@pytest.mark.parametrize("prec", [C.PREC_FP32, C.PREC_BF16])
def test_hybrid_ragged_batch_crossing_a_block_boundary(prec):
    """Three sequences; the 250-token one reaches position 256 during the decode steps, so its SSM
    state moves from block to block through the snapshots."""
    cfg = C.tiny_hybrid()
    prompts = [W.random_prompt(cfg.vocab_size, n, 10 + i) for i, n in enumerate((250, 17, 300))]
    _run(cfg, prec, prompts, 10, [[0, 1], [2], [3, 4]], 6)


def test_hybrid_larger_state_and_groups():
    """state 64 / head_dim 64 / 1 group / 3 conv taps, batch of 12 decode rows (tcgen05 GEMM tiles)."""
    cfg = C.tiny_hybrid(mamba_num_heads=4, mamba_head_dim=64, mamba_state_size=64, mamba_n_groups=1,
                        mamba_conv_kernel=3, hybrid_layers=(True, True, False, True))
    prompts = [W.random_prompt(cfg.vocab_size, 20 + i, 30 + i) for i in range(12)]
    _run(cfg, C.PREC_BF16, prompts, 4, [[i] for i in range(12)], 12, max_num_seqs=16)


def test_hybrid_resident_decode_matches_stepwise():
    """nvllm_run_resident (device-side advance, CUDA graph) on a hybrid model."""
    R = nvllm_pkg.sub("runner")
    cfg = C.tiny_hybrid()
    prompt = [int(t) for t in W.random_prompt(cfg.vocab_size, 30, 3)]
    outs = []
    for resident in (False, True):
        run = R.CudaModelRunner(cfg, W.random_weights(cfg, 1234), precision=C.PREC_FP32, num_kv_blocks=2,
                                sampling=None)
        toks = list(prompt)
        o, _ = run.run_flat([toks], True, [0], [[0]], [0.0])
        toks.append(int(o[0]))
        o, _ = run.run_flat([toks], False, None, [[0]], [0.0])
        toks.append(int(o[0]))
        for _ in range(12):
            if resident:
                run.run_resident()
                o, _ms = run.sync(1)
            else:
                o, _ = run.run_flat([toks], False, None, [[0]], [0.0])
            toks.append(int(o[0]))
        run.close()
        outs.append(toks)
    assert outs[0] == outs[1]


def test_hybrid_rejects_multi_token_prefill_on_a_cached_prefix():
    R = nvllm_pkg.sub("runner")
    L = nvllm_pkg.sub("lib")
    cfg = C.tiny_hybrid()
    run = R.CudaModelRunner(cfg, W.random_weights(cfg, 1234), precision=C.PREC_FP32, num_kv_blocks=4,
                            sampling=None)
    try:
        p = [int(t) for t in W.random_prompt(cfg.vocab_size, 300, 2)]
        run.run_flat([p], True, [0], [[0, 1]], [0.0])
        with pytest.raises(L.NvllmError, match="Mamba2"):
            run.run_flat([p], True, [256], [[0, 1]], [0.0])
        run.run_flat([p + [5]], True, [300], [[0, 1]], [0.0])  # decode-shaped: fine
    finally:
        run.close()
