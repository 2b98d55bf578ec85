"""GPU parity tests (run with -m gpu on a B200): the CUDA path, called through the C ABI
(include/nvllm.h via nano-vllm-go_b200/lib.py + runner.py), against the CPU oracle on the
same seeded random-init weights and prompts.

Gates (BASELINE.json north_star / SURVEY.md §8d):
  fp32 mode  max|dlogit| / max|logit| <= 1e-4, greedy tokens identical
  bf16 mode  (weights rounded once to bf16, the same values fed to the oracle as fp32,
              fp32 accumulate) <= 2e-2, greedy tokens identical for 32 steps
  token ids / argmax: bit-exact (first-max rule, cmd/ask/main.go:389-402)
"""
import numpy as np
import pytest

from helpers import C, W, E, TINY, TINY_HD, make_oracle, oracle_mod, rel_err, ref_filter_probs
import nvllm_pkg

pytestmark = pytest.mark.gpu

TOL = {C.PREC_FP32: 1e-4, C.PREC_BF16: 2e-2}


def _runner(cfg, prec, seed=1234, **kw):
    R = nvllm_pkg.sub("runner")
    return R.CudaModelRunner(cfg, W.random_weights(cfg, seed, round_bf16=(prec == C.PREC_BF16)),
                             precision=prec, num_kv_blocks=kw.pop("num_kv_blocks", 32), **kw)


def _blocks(n_tokens, first=0):
    nb = (n_tokens + 255) // 256
    return list(range(first, first + nb))


# ------------------------------------------------------------------ single operators
@pytest.mark.parametrize("M,N,K", [(1, 768, 768), (1, 1000, 256), (4, 515, 1024), (8, 64, 4544),
                                   (33, 200, 256), (130, 777, 512)])
def test_matmul_fp32_vs_oracle(M, N, K):
    O = oracle_mod()
    L = nvllm_pkg.sub("lib")
    rng = np.random.default_rng(M * 1000 + N)
    A = rng.standard_normal((M, K)).astype(np.float32)
    Wm = (rng.standard_normal((K, N)) * 0.05).astype(np.float32)
    bias = rng.standard_normal(N).astype(np.float32)
    ref = O.matmul(A, Wm) + bias
    got, _ = L.op_matmul(A, Wm, bias, prec=C.PREC_FP32)
    assert rel_err(got, ref) < 1e-5


@pytest.mark.parametrize("M,N,K", [
    (16, 128, 64), (9, 128, 128), (32, 256, 512), (64, 3072, 2048), (33, 200, 264),
    (130, 777, 512), (300, 1000, 1024), (64, 50257, 768), (128, 4672, 4544)])
def test_matmul_tcgen05_vs_oracle(M, N, K):
    """The tcgen05/TMEM/TMA GEMM (kernel=3): tile tails in M, N and K, split-K, LM-head and
    Falcon shapes.  bf16-exact inputs, fp32 accumulate: only summation order differs."""
    O = oracle_mod()
    L = nvllm_pkg.sub("lib")
    rng = np.random.default_rng(M * 7 + N * 3 + K)
    A = W.bf16_round(rng.standard_normal((M, K)).astype(np.float32))
    Wm = W.bf16_round((rng.standard_normal((K, N)) * 0.05).astype(np.float32))
    ref = A.astype(np.float64) @ Wm.astype(np.float64) if M * N * K > 2e9 else O.matmul(A, Wm)
    got, _ = L.op_matmul(A, Wm, None, prec=C.PREC_BF16, kernel=3)
    assert rel_err(got, ref) < 2e-5


@pytest.mark.parametrize("M,N,K", [(1, 768, 768), (8, 520, 2048), (64, 512, 512), (200, 264, 1024)])
def test_matmul_bf16_vs_oracle(M, N, K):
    O = oracle_mod()
    L = nvllm_pkg.sub("lib")
    rng = np.random.default_rng(M + N + K)
    A = W.bf16_round(rng.standard_normal((M, K)).astype(np.float32))
    Wm = W.bf16_round((rng.standard_normal((K, N)) * 0.05).astype(np.float32))
    ref = O.matmul(A, Wm)
    got, _ = L.op_matmul(A, Wm, None, prec=C.PREC_BF16)
    # inputs are exactly representable; only the fp32 summation order differs
    assert rel_err(got, ref) < 1e-5


@pytest.mark.parametrize("bias", [True, False])
def test_norm_vs_oracle(bias):
    O = oracle_mod()
    L = nvllm_pkg.sub("lib")
    rng = np.random.default_rng(3)
    x = rng.standard_normal((37, 768)).astype(np.float32) * 3 + 0.5
    w = (1 + 0.02 * rng.standard_normal(768)).astype(np.float32)
    b = (0.02 * rng.standard_normal(768)).astype(np.float32) if bias else None
    ref = O.layer_norm(x, w, b, 1e-5)
    got, _ = L.op_norm(x, w, b, 1e-5, prec=C.PREC_FP32)
    assert rel_err(got, ref) < 1e-5


def test_argmax_bit_exact_with_ties():
    O = oracle_mod()
    L = nvllm_pkg.sub("lib")
    rng = np.random.default_rng(5)
    logits = rng.standard_normal((9, 50257)).astype(np.float32)
    logits[0, 100] = logits[0, 40000] = 99.0  # tie: first index wins
    logits[1, :] = 0.0  # all equal -> 0
    logits[2, 50256] = 50.0
    got, _ = L.op_argmax(logits)
    ref = np.array([O.argmax(r) for r in logits], dtype=np.int32)
    np.testing.assert_array_equal(got, ref)
    assert got[0] == 100 and got[1] == 0 and got[2] == 50256


# ------------------------------------------------------------------ whole forward pass
def _greedy_parity(name, prec, prompt_len, steps, table=None):
    cfg = (table or TINY)[name]()
    bf16 = prec == C.PREC_BF16
    orc = make_oracle(cfg, round_bf16=bf16)
    run = _runner(cfg, prec)
    try:
        prompt = W.random_prompt(cfg.vocab_size, prompt_len, 4321)
        ref_toks, ref_logits = orc.greedy(prompt, steps)
        toks = list(int(t) for t in prompt)
        bt = _blocks(prompt_len + steps)
        worst = 0.0
        flips = 0
        for i in range(steps):
            out, logits = run.run_flat([toks], is_prefill=(i == 0), num_cached=[0],
                                       block_tables=[bt], temperature=[0.0], want_logits=True)
            err = rel_err(logits[0], ref_logits[i])
            worst = max(worst, err)
            assert err <= TOL[prec], f"{name} step {i}: rel err {err}"
            # the device argmax must be the first-max of the device logits, bit-exact
            assert int(out[0]) == oracle_mod().argmax(logits[0])
            if int(out[0]) != int(ref_toks[i]):
                # fp32 mode: never.  bf16 mode: only when the oracle's own top-2 margin is inside
                # the numerical tolerance (random-init logits are nearly flat, so near-ties exist)
                margin = float(ref_logits[i][ref_toks[i]] - ref_logits[i][out[0]])
                scale = float(np.abs(ref_logits[i]).max())
                assert prec == C.PREC_BF16 and margin <= 2 * err * scale + 1e-6, \
                    f"{name} step {i}: token {out[0]} != {ref_toks[i]} (margin {margin}, err {err})"
                flips += 1
            toks.append(int(ref_toks[i]))  # teacher-forced with the oracle's token
        assert flips <= 2, f"{name}: {flips} near-tie flips in {steps} steps"
        return worst
    finally:
        run.close()


@pytest.mark.parametrize("name", list(TINY))
def test_fp32_logits_and_greedy_tokens(name):
    _greedy_parity(name, C.PREC_FP32, prompt_len=40, steps=32)


@pytest.mark.parametrize("name", list(TINY))
def test_bf16_logits_and_greedy_tokens(name):
    _greedy_parity(name, C.PREC_BF16, prompt_len=40, steps=32)


@pytest.mark.parametrize("name", list(TINY_HD))
@pytest.mark.parametrize("prec", [C.PREC_FP32, C.PREC_BF16])
def test_head_dim_64_128_families(name, prec):
    """head_dim 64 / 128: bf16 mode runs the tensor-core attention kernels (prefill tiles of 64
    queries, decode with the query-head group as MMA rows, MQA group of 18 heads in 2 chunks)."""
    _greedy_parity(name, prec, prompt_len=100, steps=12, table=TINY_HD)


@pytest.mark.parametrize("name", ["llama_hd64", "llama_hd128", "falcon_g18"])
def test_bf16_long_context_split_kv(name):
    """700-token prompt: several KV tiles per CTA, split-KV decode (few CTAs -> splits),
    three KV blocks."""
    _greedy_parity(name, C.PREC_BF16, prompt_len=700, steps=6, table=TINY_HD)


PACKED = {
    # four kv heads so that 1, 2 or 4 of them can share a decode-attention CTA
    "llama_kv4_hd64": lambda: C.tiny_llama(hidden=512, num_heads=8, num_kv_heads=4, head_dim=64),
    "llama_kv4_hd128": lambda: C.tiny_llama(hidden=512, num_heads=4, num_kv_heads=4, head_dim=128),
}


@pytest.mark.parametrize("name", list(PACKED))
@pytest.mark.parametrize("kpc", [1, 2, 4])
def test_decode_attention_heads_per_cta(name, kpc, monkeypatch):
    """attention_dec.cu packs 1, 2 or 4 kv heads per CTA (4, 2 or 1 warps per head, tiles of the
    heads interleaved in the TMA ring); big batches pick 2 or 4 on their own, here the knob
    forces each variant.  300-token prompt: two KV blocks, the appended row crosses tiles."""
    monkeypatch.setenv("NVLLM_ATTN_T_KPC", str(kpc))
    _greedy_parity(name, C.PREC_BF16, prompt_len=300, steps=8, table=PACKED)


MOE_WIDE = {
    # grouped tensor-core expert GEMMs: several 128-row weight tiles per expert, experts with
    # more rows than one token tile (600 x 4 / 8 = 300 rows per expert on average), N not a
    # multiple of 128 (2 I = 192)
    "granite_wide": lambda: C.tiny_granite(hidden=256, num_heads=4, num_kv_heads=2, head_dim=64,
                                           ffn_dim=128, num_experts=8, num_experts_per_tok=4),
    "granite_odd": lambda: C.tiny_granite(hidden=192, num_heads=6, num_kv_heads=2, head_dim=32,
                                          ffn_dim=96, num_experts=6, num_experts_per_tok=3),
}


@pytest.mark.parametrize("name", list(MOE_WIDE))
@pytest.mark.parametrize("prompt_len", [40, 600])
def test_bf16_moe_grouped_tensor_core_gemm(name, prompt_len):
    _greedy_parity(name, C.PREC_BF16, prompt_len=prompt_len, steps=6, table=MOE_WIDE)


@pytest.mark.parametrize("name", ["gpt2", "llama"])
def test_context_crossing_block_boundaries(name):
    """Prompt of 250 tokens, decode across position 256 (second KV block)."""
    _greedy_parity(name, C.PREC_FP32, prompt_len=250, steps=12)


class OracleRunner:
    """TensorModelRunner restated over the oracle (tensor_model_runner.go:55-97): per-sequence
    loop, per-SeqID cache, prefill = whole prompt."""

    def __init__(self, model):
        self.m = model
        self.caches = {}
        self.logits = {}

    def run(self, seqs, is_prefill):
        out = []
        for s in seqs:
            if is_prefill or s.seq_id not in self.caches:
                self.caches[s.seq_id] = self.m.new_cache()
                lg = self.m.forward(s.token_ids, self.caches[s.seq_id], 0)
            else:
                lg = self.m.forward(s.token_ids[-1:], self.caches[s.seq_id], len(s.token_ids) - 1)
            self.logits.setdefault(s.seq_id, []).append(lg)
            out.append(oracle_mod().argmax(lg))
        return out

    def close(self):
        pass


@pytest.mark.parametrize("name", list(TINY))
def test_engine_batch_with_prefix_cache(name):
    """The restated Go engine (scheduler + block manager) drives the CUDA runner with a ragged
    batch: different prompt lengths, two prompts sharing a 256-token prefix (prefix-cache hit in
    the same step and in a later step), decode across a block boundary.  Tokens must equal the
    per-sequence oracle run."""
    cfg = TINY[name]()
    rng = np.random.default_rng(11)
    shared = rng.integers(0, cfg.vocab_size, 256).tolist()
    prompts = [
        rng.integers(0, cfg.vocab_size, 7).tolist(),
        shared + rng.integers(0, cfg.vocab_size, 9).tolist(),
        shared + rng.integers(0, cfg.vocab_size, 30).tolist(),  # same-step prefix hit
        rng.integers(0, cfg.vocab_size, 250).tolist(),          # crosses 256 while decoding
    ]
    late = shared + rng.integers(0, cfg.vocab_size, 3).tolist()  # later-step prefix hit
    steps = 10
    ecfg = dict(num_kvcache_blocks=32, max_model_len=cfg.max_seq_len,
                max_num_batched_tokens=max(4096, cfg.max_seq_len))

    def drive(runner):
        eng = E.LLMEngine(E.EngineConfig(**ecfg), runner)
        seqs = [eng.add_request(p, E.SamplingParams(max_tokens=steps, ignore_eos=True),
                                temperature=0.0) for p in prompts]
        cached = []
        n_steps = 0
        while not eng.is_finished():
            if n_steps == 3:
                seqs.append(eng.add_request(late, E.SamplingParams(max_tokens=steps, ignore_eos=True),
                                            temperature=0.0))
            eng.step()
            n_steps += 1
            cached.append([s.num_cached_tokens for s in seqs])
        return [s.completion_token_ids() for s in seqs], cached

    ref, cached_ref = drive(OracleRunner(make_oracle(cfg)))
    run = _runner(cfg, C.PREC_FP32)
    try:
        got, cached_got = drive(run)
    finally:
        run.close()
    assert cached_got == cached_ref
    assert any(256 in c for c in cached_got), "the prefix cache never hit"
    assert got == ref


def test_fully_cached_prompt_still_yields_logits():
    """A prompt that is an exact multiple of 256 and fully cached (NumCachedTokens == Len) must
    still compute its last token (SURVEY.md §7 hard part 2)."""
    cfg = TINY["llama"]()
    orc = make_oracle(cfg)
    run = _runner(cfg, C.PREC_FP32)
    try:
        prompt = W.random_prompt(cfg.vocab_size, 256, 77).tolist()
        ref = orc.forward(prompt, orc.new_cache())
        t1, l1 = run.run_flat([prompt], True, [0], [[3]], [0.0], want_logits=True)
        t2, l2 = run.run_flat([prompt], True, [256], [[3]], [0.0], want_logits=True)
        assert rel_err(l1[0], ref) <= 1e-4 and rel_err(l2[0], ref) <= 1e-4
        assert t1[0] == t2[0] == oracle_mod().argmax(ref)
    finally:
        run.close()


def test_temperature_sampling_with_injected_uniforms():
    """sampling.go:70-74,105-128,196-217 with rand.Float32() replaced by a given uniform."""
    O = oracle_mod()
    cfg = TINY["gpt2"]()
    run = _runner(cfg, C.PREC_FP32)
    try:
        prompts = [W.random_prompt(cfg.vocab_size, 12 + i, 50 + i).tolist() for i in range(6)]
        temps = [0.7, 1.0, 1.3, 0.0, 2.0, 0.5]
        us = [0.01, 0.35, 0.62, 0.5, 0.93, 0.999]
        bts = [[i] for i in range(6)]
        toks, logits = run.run_flat(prompts, True, [0] * 6, bts, temps, us, want_logits=True)
        exact = 0
        for i in range(6):
            if temps[i] == 0.0:
                assert toks[i] == O.argmax(logits[i])
                continue
            ref = O.sample(logits[i], None, temps[i], u=us[i])
            # identical up to fp32 summation order of the CDF: allow the neighbouring index when
            # u lands within rounding of a CDF step
            p = np.exp((logits[i] / temps[i]).astype(np.float64))
            cdf = np.cumsum(p / p.sum())
            assert abs(cdf[toks[i]] - cdf[ref]) < 1e-3 or toks[i] == ref
            exact += int(toks[i] == ref)
        assert exact >= 4
    finally:
        run.close()


@pytest.mark.parametrize("top_p,top_k,rep", [(1.0, 0, 1.2), (1.0, 40, 1.0), (0.9, 0, 1.0),
                                              (0.85, 50, 1.2), (0.5, 5, 1.5)])
def test_sample_with_history_filters(top_p, top_k, rep):
    """sampling.go:33-217 SampleWithHistory on the device: repetition penalty over the token
    history (x3 for the last ten), temperature, top-k, top-p, renormalise, inverse-CDF draw —
    rand.Float32() replaced by injected uniforms.  Tokens must equal the oracle's; where u
    lands within fp32 rounding of a CDF step the neighbouring kept token is accepted."""
    O = oracle_mod()
    cfg = TINY["gpt2"]()
    run = _runner(cfg, C.PREC_FP32)
    try:
        run.set_sampling_params(1.0, top_p=top_p, top_k=top_k, repetition_penalty=rep)
        rng = np.random.default_rng(7)
        prompts = []
        for i in range(8):
            pr = W.random_prompt(cfg.vocab_size, 14 + 3 * i, 90 + i).tolist()
            pr[-3] = pr[2]  # a repeat inside the last ten tokens (weight 3 + 1)
            pr[5] = pr[1]
            prompts.append(pr)
        temps = [0.7, 1.0, 1.3, 0.0, 2.0, 0.5, 1.0, 0.9]
        us = [float(u) for u in rng.uniform(0.02, 0.98, size=8)]
        bts = [[i] for i in range(8)]
        toks, logits = run.run_flat(prompts, True, [0] * 8, bts, temps, us, want_logits=True)
        exact = 0
        for i in range(8):
            if temps[i] == 0.0:
                assert toks[i] == O.argmax(logits[i])  # greedy rows are not filtered
                continue
            ref = O.sample(logits[i], prompts[i], temps[i], top_p=top_p, top_k=top_k,
                           rep_penalty=rep, u=us[i])
            p = ref_filter_probs(logits[i], prompts[i], temps[i], top_p, top_k, rep)
            assert p[toks[i]] > 0, f"row {i}: token {toks[i]} was filtered out by the reference"
            if toks[i] != ref:
                cdf = np.cumsum(p)
                assert abs(cdf[toks[i]] - cdf[ref]) < 1e-3 or abs(cdf[min(toks[i], ref)] - us[i]) < 1e-3, \
                    f"row {i}: token {toks[i]} != oracle {ref}"
            exact += int(toks[i] == ref)
        assert exact >= 6
        # decode step: the history is the whole sequence so far
        seqs = [pr + [int(t)] for pr, t in zip(prompts, toks)]
        toks2, logits2 = run.run_flat(seqs, False, None, bts, temps, us, want_logits=True)
        same = sum(int(toks2[i] == O.sample(logits2[i], seqs[i], temps[i], top_p=top_p, top_k=top_k,
                                            rep_penalty=rep, u=us[i]))
                   for i in range(8) if temps[i] > 0)
        assert same >= 6
        # the device-resident loop has no host-side history: it must refuse, not ignore the knobs
        L = nvllm_pkg.sub("lib")
        with pytest.raises(L.NvllmError):
            run.run_resident()
    finally:
        run.close()


def test_error_paths_return_codes_not_aborts():
    L = nvllm_pkg.sub("lib")
    cfg = C.tiny_llama(max_seq_len=64)
    run = _runner(cfg, C.PREC_FP32)
    try:
        with pytest.raises(L.NvllmError) as e:  # token outside the vocabulary
            run.run_flat([[cfg.vocab_size]], True, [0], [[0]], [0.0])
        assert e.value.code == L.ERR_RANGE
        with pytest.raises(L.NvllmError) as e:  # RoPE position past max_seq_len (rope.go:84)
            run.run_flat([[1] * 65], True, [0], [[0]], [0.0])
        assert e.value.code == L.ERR_RANGE
        with pytest.raises(L.NvllmError) as e:  # block table too short
            run.run_flat([[1] * 10], True, [0], [[]], [0.0])
        assert e.value.code == L.ERR_INVALID
        with pytest.raises(L.NvllmError) as e:  # block id outside the pool
            run.run_flat([[1] * 10], True, [0], [[9999]], [0.0])
        assert e.value.code == L.ERR_RANGE
        for bad in ((0.0, 0, 1.0), (1.0, -1, 1.0), (1.0, 0, 0.0)):  # top_p, top_k, penalty
            with pytest.raises(L.NvllmError) as e:
                run.set_sampling_params(1.0, top_p=bad[0], top_k=bad[1], repetition_penalty=bad[2])
            assert e.value.code == L.ERR_INVALID
        # the handle is still usable after errors (and after a speculative step was thrown away
        # by the failing calls above)
        toks, _ = run.run_flat([[1, 2, 3]], True, [0], [[0]], [0.0])
        assert 0 <= toks[0] < cfg.vocab_size
        toks2, _ = run.run_flat([[1, 2, 3, int(toks[0])]], False, None, [[0]], [0.0])
        assert 0 <= toks2[0] < cfg.vocab_size
    finally:
        run.close()


def test_missing_weight_is_reported():
    L = nvllm_pkg.sub("lib")
    R = nvllm_pkg.sub("runner")
    cfg = TINY["llama"]()
    ws = [w for w in W.random_weights(cfg) if not (w[0] == C.W_O and w[1] == 1)]
    with pytest.raises(L.NvllmError) as e:
        R.CudaModelRunner(cfg, ws, precision=C.PREC_FP32, num_kv_blocks=4)
    assert e.value.code == L.ERR_INVALID and "NVLLM_W_O" in str(e.value)


def test_resident_decode_matches_stepwise():
    """nvllm_run_resident (device-resident decode loop used by bench.py's `value` leg) must
    produce the same tokens as stepping through nvllm_run."""
    cfg = TINY["llama"]()
    prompts = [W.random_prompt(cfg.vocab_size, 20 + 5 * i, 200 + i).tolist() for i in range(5)]
    steps = 12
    bts = [_blocks(len(p) + steps + 2, first=2 * i) for i, p in enumerate(prompts)]
    run = _runner(cfg, C.PREC_FP32)
    try:
        seqs = [list(p) for p in prompts]
        out, _ = run.run_flat(seqs, True, [0] * 5, bts, [0.0] * 5)
        for s, t in zip(seqs, out):
            s.append(int(t))
        ref = []
        for _ in range(steps):
            out, _ = run.run_flat(seqs, False, None, bts, [0.0] * 5)
            ref.append(out.copy())
            for s, t in zip(seqs, out):
                s.append(int(t))
        # again, resident
        seqs = [list(p) for p in prompts]
        out, _ = run.run_flat(seqs, True, [0] * 5, bts, [0.0] * 5)
        first_out = out.copy()
        for s, t in zip(seqs, out):
            s.append(int(t))
        out, _ = run.run_flat(seqs, False, None, bts, [0.0] * 5)
        np.testing.assert_array_equal(out, ref[0])
        for i in range(1, steps):
            run.run_resident()
            got, _ = run.sync(5)
            np.testing.assert_array_equal(got, ref[i])
            # steady-state decode steps replay one captured CUDA graph
            assert run.last_stats.used_graph == 1
        # the same loop with graphs off launches kernel by kernel and gives the same tokens
        run.set_graphs(False)
        seqs = [list(p) + [int(t)] for p, t in zip(prompts, first_out)]
        out, _ = run.run_flat(seqs, False, None, bts, [0.0] * 5)
        np.testing.assert_array_equal(out, ref[0])
        for i in range(1, 4):
            run.run_resident()
            got, _ = run.sync(5)
            np.testing.assert_array_equal(got, ref[i])
            assert run.last_stats.used_graph == 0
        run.set_graphs(True)
        # resident straight after a prefill step (decode state derived on the device), twice:
        # the second time from nvllm_rerun's replay of the staged prefill
        first, _ = run.run_flat([list(p) for p in prompts], True, [0] * 5, bts, [0.0] * 5)
        for rep in range(2):
            if rep == 1:
                run.rerun()
                got, _ = run.sync(5)
                np.testing.assert_array_equal(got, first)
            for i in range(steps):
                run.run_resident()
                got, _ = run.sync(5)
                np.testing.assert_array_equal(got, ref[i])
    finally:
        run.close()


def test_speculative_next_step_hits_and_misses():
    """nvllm_run ends by issuing the step the engine asks for next; a call that matches it copies
    nothing to the device and must return what a non-speculating runner returns, a call that
    does not match (another token, another batch) must be unaffected by the discarded step."""
    cfg = TINY["llama"]()
    prompts = [W.random_prompt(cfg.vocab_size, 18 + 7 * i, 300 + i).tolist() for i in range(4)]
    bts = [_blocks(len(p) + 20, first=2 * i) for i, p in enumerate(prompts)]

    def drive(spec, force_at=None):
        run = _runner(cfg, C.PREC_FP32)
        try:
            run.set_speculation(spec)
            seqs = [list(p) for p in prompts]
            out, _ = run.run_flat(seqs, True, [0] * 4, bts, [0.0] * 4)
            toks, hits, logs = [out.copy()], 0, []
            for step in range(10):
                for i, (s, t) in enumerate(zip(seqs, out)):
                    # a forced token differs from what the device predicted: a miss
                    s.append(7 if (force_at == step and i == 1) else int(t))
                if step == 6:  # another batch in between (the first two sequences only): a miss
                    o2, _ = run.run_flat(seqs[:2], False, None, bts[:2], [0.0] * 2)
                    logs.append(o2.copy())
                    for s, t in zip(seqs[:2], o2):
                        s.append(int(t))
                out, lg = run.run_flat(seqs, False, None, bts, [0.0] * 4, want_logits=True)
                hits += int(run.last_stats.h2d_bytes == 0)
                toks.append(out.copy())
                logs.append(lg.copy())
            return toks, hits, logs
        finally:
            run.close()

    for force_at in (None, 3):
        ref_t, ref_h, ref_l = drive(False, force_at)
        got_t, got_h, got_l = drive(True, force_at)
        assert ref_h == 0
        # steps 0..9: every step hits except the forced one, the one after the interleaved
        # batch, and the one call after each miss that does not speculate
        assert got_h >= 5, got_h
        for a, b in zip(ref_t, got_t):
            np.testing.assert_array_equal(a, b)
        for a, b in zip(ref_l, got_l):
            np.testing.assert_array_equal(a, b)  # same kernels, same order: bit-identical


def test_profile_reports_every_kernel_class():
    cfg = TINY["granite"]()
    run = _runner(cfg, C.PREC_BF16)
    try:
        prompt = W.random_prompt(cfg.vocab_size, 40, 1).tolist()
        run.set_profile(True)
        run.run_flat([prompt], True, [0], [[0]], [0.0])
        prof = run.get_profile()
        # (the router runs inside the FFN norm; sort + gather are one "moe_route" launch at this
        # size and the weighted combine is part of the next norm: no "moe_misc" launches)
        for k in ("embed", "norm", "gemm_tc5", "rope_kv", "attn_prefill", "moe_route", "moe_gemm",
                  "sample"):
            assert k in prof and prof[k]["ms"] > 0 and prof[k]["bytes"] > 0, k
        assert prof["gemm_tc5"]["launches"] == 2 * cfg.num_layers
        run.set_profile(False)
    finally:
        run.close()
