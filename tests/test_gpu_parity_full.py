"""GPU parity at the sizes BASELINE.json names (run with -m gpu on a B200).

  * configs[0] at FULL size — GPT-2 small 124M, 32-token prompt: fp32 mode <= 1e-4 on every
    step and 64 identical greedy tokens; bf16 mode <= 2e-2 and 32 identical greedy tokens.  Both
    FREE-RUNNING (the device's own tokens are fed back), on the prompt seed picked by
    tests/golden/find_margin_seeds.py (first seed whose oracle top-2 margin stays >= 2 % of
    max|logit|: random-init logits are nearly flat and a near-tie may legitimately flip in bf16).
  * configs[1..4] at FULL WIDTH and the benchmark's batch size, reduced depth and few tokens so
    the scalar CPU oracle finishes in seconds (it runs one sequence per host thread): these are
    the launch shapes bench.py times — split-K 8 tcgen05 GEMMs at M = 64 / 128, the LM head at
    N = 128 256, two kv heads per decode-attention CTA picked by the heuristic, the L2-prefetch
    paths, grouped expert GEMMs with the speculative prefetch, Falcon's 71-head MQA and
    K = 4544 / N = 18 176 GEMMs, head_dim-128 prefill with a prefix-cache hit in the same step.
    Logits within the gate for every sequence and step; the oracle's token is fed back, and a
    device token may differ from it only where the oracle's own top-2 margin is inside the
    measured error (counted and printed).
"""
from dataclasses import replace

import numpy as np
import pytest

from helpers import C, W, make_oracle, oracle_mod, rel_err
import nvllm_pkg

pytestmark = pytest.mark.gpu

TOL = {C.PREC_FP32: 1e-4, C.PREC_BF16: 2e-2}
GPT2_SMALL_SEED = 4323  # tests/golden/find_margin_seeds.py: min margin 0.078 of max|logit|


def _runner(cfg, prec, num_kv_blocks, **kw):
    R = nvllm_pkg.sub("runner")
    bf16 = prec == C.PREC_BF16
    return R.CudaModelRunner(cfg, W.random_weights(cfg, 1234, round_bf16=bf16), precision=prec,
                             num_kv_blocks=num_kv_blocks, sampling=None,
                             weight_dtype=C.DT_BF16 if bf16 else None, **kw)


# ------------------------------------------------------------------ configs[0], full size
@pytest.mark.parametrize("prec,steps", [(C.PREC_FP32, 64), (C.PREC_BF16, 32)])
def test_cfg1_gpt2_small_full_size_free_running(prec, steps):
    """cmd/ask/main.go:287-360,389-402 greedy loop on purego/tensor/config.go:125-148."""
    cfg = C.gpt2_small()
    orc = make_oracle(cfg, round_bf16=(prec == C.PREC_BF16))
    prompt = W.random_prompt(cfg.vocab_size, 32, GPT2_SMALL_SEED)
    ref_toks, ref_logits = orc.greedy(prompt, steps)
    run = _runner(cfg, prec, num_kv_blocks=4)
    try:
        toks = [int(t) for t in prompt]
        worst = 0.0
        for i in range(steps):
            out, logits = run.run_flat([toks], i == 0, [0], [[0]], [0.0], want_logits=True)
            err = rel_err(logits[0], ref_logits[i])
            worst = max(worst, err)
            assert err <= TOL[prec], f"step {i}: rel err {err}"
            assert int(out[0]) == oracle_mod().argmax(logits[0])
            assert int(out[0]) == int(ref_toks[i]), f"step {i}: token {out[0]} != oracle {ref_toks[i]}"
            toks.append(int(out[0]))
        print(f"gpt2-small full size prec={prec}: {steps} identical tokens, worst rel err {worst:.2e}")
    finally:
        run.close()


# ------------------------------------------------------------------ configs[1..4], full width
def _oracle_batch(orc, prompts, steps):
    """Per-sequence greedy runs of the oracle, one host thread each."""
    return orc.greedy_batch(prompts, steps)


def _batch_parity(cfg, prec, prompts, steps, tables, num_cached=None, num_kv_blocks=None):
    bf16 = prec == C.PREC_BF16
    n = len(prompts)
    orc = make_oracle(cfg, round_bf16=bf16)
    ref_toks, ref_logits = _oracle_batch(orc, prompts, steps)
    del orc
    run = _runner(cfg, prec, num_kv_blocks=num_kv_blocks or (max(max(t) for t in tables) + 1),
                  max_num_seqs=max(n, 8), max_batched_tokens=max(4096, sum(len(p) for p in prompts)))
    worst, ties, mism = 0.0, 0, 0
    try:
        seqs = [[int(t) for t in p] for p in prompts]
        for i in range(steps):
            out, logits = run.run_flat(seqs, i == 0, num_cached if i == 0 else None, tables,
                                       [0.0] * n, want_logits=True)
            for s in range(n):
                ref = ref_logits[s][i]
                err = rel_err(logits[s], ref)
                worst = max(worst, err)
                assert err <= TOL[prec], f"{cfg.name} seq {s} step {i}: rel err {err}"
                assert int(out[s]) == oracle_mod().argmax(logits[s])
                if int(out[s]) != int(ref_toks[s][i]):
                    margin = float(ref[ref_toks[s][i]] - ref[out[s]])
                    assert bf16 and margin <= 2 * err * float(np.abs(ref).max()) + 1e-6, \
                        f"{cfg.name} seq {s} step {i}: token {out[s]} != {ref_toks[s][i]} (margin {margin})"
                    ties += 1
                seqs[s].append(int(ref_toks[s][i]))
        print(f"{cfg.name} L={cfg.num_layers} B={n} prec={prec}: worst rel err {worst:.2e}, "
              f"{ties} near-tie token(s) of {n * steps}")
        assert ties <= max(2, (n * steps) // 20)
        return worst
    finally:
        run.close()


def _plain_batch(cfg, batch, prompt_len, steps, seed0=4321):
    prompts = [W.random_prompt(cfg.vocab_size, prompt_len, seed0 + i) for i in range(batch)]
    nb = (prompt_len + steps + 255) // 256
    tables = [list(range(i * nb, (i + 1) * nb)) for i in range(batch)]
    return prompts, tables


def test_cfg2_llama_1b_full_width_batch64():
    """Llama-3.2-1B width (H 2048, 32/8 heads, F 8192, V 128256), 4 of 16 layers, the bench's
    64 sequences: 1024-token prefill step (256-token GEMM tiles, several per weight tile) and
    M = 64 decode steps (`gemm_tc5<64,8>` split-K, fused-RoPE decode attention, LM head)."""
    cfg = replace(C.llama32_1b(), num_layers=4)
    prompts, tables = _plain_batch(cfg, 64, 16, 5)
    _batch_parity(cfg, C.PREC_BF16, prompts, 5, tables)


def test_cfg2_llama_1b_full_depth():
    """All 16 layers at full width: error growth through the whole stack, 2 sequences."""
    cfg = C.llama32_1b()
    prompts, tables = _plain_batch(cfg, 2, 12, 4)
    _batch_parity(cfg, C.PREC_BF16, prompts, 4, tables)


def test_cfg3_granite_moe_full_width_batch128():
    """Granite-3.0-1B-A400M width (32 experts, top-8, muP multipliers), 4 of 24 layers, the
    bench's 128 sequences: 1024 (token, expert) pairs per decode step through the grouped
    tcgen05 expert GEMMs, router fused into the norm, sort + gather, weighted combine."""
    cfg = replace(C.granite_moe_1b(), num_layers=4)
    prompts, tables = _plain_batch(cfg, 128, 16, 5)
    _batch_parity(cfg, C.PREC_BF16, prompts, 5, tables)


def test_cfg4_falcon_7b_full_width_two_layers():
    """Falcon-7B width (71-head MQA, H 4544, F 18176, untied LM head 65024), 2 of 32 layers, one
    sequence like the bench: 160-token prefill (256-row GEMM tiles, MQA prefill attention) and
    M = 1 decode steps (split-K GEMV, 71-head MQA decode attention)."""
    cfg = replace(C.falcon_7b(), num_layers=2)
    prompts, tables = _plain_batch(cfg, 1, 160, 5)
    _batch_parity(cfg, C.PREC_BF16, prompts, 5, tables)


def test_cfg5_llama_3b_full_width_prefix_cached_prompts():
    """Llama-3.2-3B width (H 3072, 24/8 heads of 128), 2 of 28 layers: four prompts sharing a
    256-token prefix (one KV block) + 48 unique tokens, scheduled in ONE prefill step like the
    bench: request 0 cold, the others prefix-cache hits (NumCachedTokens = 256) reading the
    block request 0 writes in the same step; then decode steps at head_dim 128."""
    cfg = replace(C.llama32_3b(), num_layers=2)
    common = W.random_prompt(cfg.vocab_size, 256, 99)
    prompts = [np.concatenate([common, W.random_prompt(cfg.vocab_size, 48, 4321 + i)]) for i in range(4)]
    # block 0 is the shared prefix; every sequence owns one block for its tail + decode tokens
    tables = [[0, 1 + i] for i in range(4)]
    _batch_parity(cfg, C.PREC_BF16, prompts, 4, tables, num_cached=[0, 256, 256, 256],
                  num_kv_blocks=8)


def test_cfg2_llama_1b_full_width_fp32_mode():
    """The fp32 parity mode (1e-4 gate) at full width: SIMT fp32 GEMM tiles + fp32 attention,
    8 sequences, 2 layers."""
    cfg = replace(C.llama32_1b(), num_layers=2)
    prompts, tables = _plain_batch(cfg, 8, 12, 3)
    _batch_parity(cfg, C.PREC_FP32, prompts, 3, tables)


@pytest.mark.parametrize("family,batch", [("llama", 12), ("gpt2", 40), ("falcon", 70)])
def test_chained_ffn_gemms_bit_identical_to_two_launches(family, batch, monkeypatch):
    """Decode batches of 9..128 rows run W1 (+ activation) and W2 as ONE chained launch
    (gemm_tcgen05.cu Tc5Chain): same tiles, same split-K slices, same summation order as the two
    separate launches (NVLLM_CHAIN=0), so logits must be bit-identical; and within the bf16
    gate of the oracle.  Covers SwiGLU (llama), GELU + biases (gpt2), the parallel block (falcon)
    and the three token-tile widths 32 / 64 / 128."""
    from helpers import TINY
    cfg = TINY[family]()
    prompts, tables = _plain_batch(cfg, batch, 9, 3)
    n = len(prompts)
    outs = {}
    for chain in ("1", "0"):
        monkeypatch.setenv("NVLLM_CHAIN", chain)
        run = _runner(cfg, C.PREC_BF16, num_kv_blocks=n, max_num_seqs=n, max_batched_tokens=4096)
        try:
            seqs = [[int(t) for t in p] for p in prompts]
            got = []
            for i in range(3):
                out, logits = run.run_flat(seqs, i == 0, None, tables, [0.0] * n, want_logits=True)
                got.append(logits.copy())
                for s in range(n):
                    seqs[s].append(int(out[s]))
            outs[chain] = (np.stack(got), [list(x) for x in seqs])
        finally:
            run.close()
    assert outs["1"][1] == outs["0"][1]
    assert np.array_equal(outs["1"][0], outs["0"][0])
    monkeypatch.setenv("NVLLM_CHAIN", "1")
    _batch_parity(cfg, C.PREC_BF16, prompts, 3, tables)


@pytest.mark.parametrize("family", ["llama", "gpt2", "falcon"])
@pytest.mark.parametrize("batch", [1, 3, 8])
def test_w2_gemv_builds_its_own_activation(family, batch, monkeypatch):
    """Steps of <= 8 rows: the W2 GEMV computes the SwiGLU / GELU of the W1 GEMV's split-K slices
    itself (GemvPre mode 2, gemm_simt.cu) instead of reading what the activation kernel stored.
    Same slices, same summation order, same rounding: logits are bit-identical with the
    stand-alone kernel (NVLLM_GEMV_PRE=0), and inside the oracle gates — prefill of a short
    prompt and decode steps, sequential and parallel blocks."""
    from helpers import TINY
    cfg = TINY[family]()
    prompts, tables = _plain_batch(cfg, batch, 5 if batch == 1 else 1, 3)
    n = len(prompts)
    outs = {}
    for pre in ("1", "0"):
        monkeypatch.setenv("NVLLM_GEMV_PRE", pre)
        run = _runner(cfg, C.PREC_FP32, num_kv_blocks=n, max_num_seqs=n, max_batched_tokens=4096)
        try:
            seqs = [[int(t) for t in p] for p in prompts]
            got = []
            for i in range(4):
                out, logits = run.run_flat(seqs, i == 0, None, tables, [0.0] * n, want_logits=True)
                got.append(logits.copy())
                for s in range(n):
                    seqs[s].append(int(out[s]))
            outs[pre] = (np.stack(got), [list(x) for x in seqs])
        finally:
            run.close()
    assert outs["1"][1] == outs["0"][1]
    assert np.array_equal(outs["1"][0], outs["0"][0])
    monkeypatch.setenv("NVLLM_GEMV_PRE", "1")
    _batch_parity(cfg, C.PREC_FP32, prompts, 3, tables)
    _batch_parity(cfg, C.PREC_BF16, prompts, 3, tables)


@pytest.mark.parametrize("hidden,heads,kv,ffn,batch", [
    (1024, 16, 4, 4096, 12),    # 32-token tile, 8 tiles x 8 splits (O: K 1024, W2: K 4096)
    (1024, 16, 4, 4096, 40),    # 64-token tile
    (1024, 16, 4, 4096, 64),    # full tile: 8 token rows per cluster rank
    (1088, 17, 1, 2560, 33),    # 8.5 column tiles (a half-empty one), uneven row shares
    (3072, 24, 8, 8192, 20),    # Llama-3B width: 24 tiles x 6 splits, clusters of 6
])
def test_decode_fused_residual_through_cluster_shared_memory(hidden, heads, kv, ffn, batch, monkeypatch):
    """Decode batches of 9..64 rows of a dense RMSNorm model (bf16): the O and W2 GEMMs run their
    k-splits as one thread-block cluster per output tile, reduce the slices through distributed
    shared memory and apply the residual update + the folded-norm outputs themselves
    (gemm_tcgen05.cu, TC5_EPI_RESID with S > 1) — no norm kernel in the layer.  Opt-in
    (NVLLM_RESID_CLUSTER=1: measured slower than partials + norm kernel).  Checked
    against the oracle (bf16 gate) and against the default path, whose logits must agree to bf16
    noise."""
    cfg = C.tiny_llama(vocab_size=2048, hidden=hidden, num_heads=heads, num_kv_heads=kv, head_dim=64,
                       ffn_dim=ffn, num_layers=2, max_seq_len=512)
    prompts, tables = _plain_batch(cfg, batch, 7, 3)
    n = len(prompts)
    outs, launches = {}, {}
    for mode in ("1", "0"):
        monkeypatch.setenv("NVLLM_RESID_CLUSTER", mode)
        run = _runner(cfg, C.PREC_BF16, num_kv_blocks=n, max_num_seqs=n, max_batched_tokens=4096)
        try:
            run.set_speculation(False)  # the profiled step below must launch, not be a speculative hit
            seqs = [[int(t) for t in p] for p in prompts]
            got = []
            for i in range(3):
                out, logits = run.run_flat(seqs, i == 0, None, tables, [0.0] * n, want_logits=True)
                got.append(logits.copy())
                for s in range(n):
                    seqs[s].append(int(out[s]))
            outs[mode] = np.stack(got)
            run.set_profile(True)
            run.run_flat(seqs, False, None, tables, [0.0] * n)
            launches[mode] = run.get_profile()["norm"]["launches"]
            run.set_profile(False)
        finally:
            run.close()
    # the decode step keeps the final norm only; the default path has two more per layer, less
    # layer 0's attention norm (folded: the embedding kernel writes bf16(x g) and the row sums)
    assert launches == {"1": 1, "0": 2 * cfg.num_layers}, launches
    scale = float(np.abs(outs["0"]).max())
    assert float(np.abs(outs["1"][:2] - outs["0"][:2]).max()) <= 2e-2 * scale
    monkeypatch.setenv("NVLLM_RESID_CLUSTER", "1")
    _batch_parity(cfg, C.PREC_BF16, prompts, 3, tables)
