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
                             precision=prec, num_kv_blocks=kw.pop("num_kv_blocks", 32),
                             sampling=kw.pop("sampling", None), **kw)


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
# Prompt seeds of the strict (free-running, identical-token) cases: the first seed >= 4321 whose
# oracle top-2 margin is >= 2 % of max|logit| on every compared step, for fp32 and bf16-rounded
# weights alike (tests/golden/find_margin_seeds.py prints this table).  Random-init logits are
# nearly flat; on other seeds a near-tie lets a correct bf16 run pick the runner-up.
STRICT_SEEDS = {"gpt2": 4322, "llama": 4321, "granite": 4321, "falcon": 4391, "gpt2_hd64": 4321,
                "llama_hd64": 4321, "llama_hd128": 4321, "granite_hd64": 4321, "falcon_g18": 4323}


def _greedy_parity(name, prec, prompt_len, steps, table=None, strict_seed=None):
    """strict_seed given: FREE-RUNNING — the device's own tokens are fed back, every one of them
    must equal the oracle's (north_star: identical greedy tokens), logits within the gate at every
    step.  Otherwise (kernel-variant coverage on prompts whose margins were not screened): the
    oracle's token is fed back, logits within the gate, and a token may differ only where the
    oracle's own top-2 margin is inside the measured numerical error (reported, never > 2)."""
    cfg = (table or TINY)[name]()
    bf16 = prec == C.PREC_BF16
    orc = make_oracle(cfg, round_bf16=bf16)
    run = _runner(cfg, prec)
    try:
        prompt = W.random_prompt(cfg.vocab_size, prompt_len, strict_seed or 4321)
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
            if strict_seed is not None:
                assert int(out[0]) == int(ref_toks[i]), \
                    f"{name} step {i}: token {out[0]} != oracle {ref_toks[i]} (free-running)"
                toks.append(int(out[0]))
                continue
            if int(out[0]) != int(ref_toks[i]):
                margin = float(ref_logits[i][ref_toks[i]] - ref_logits[i][out[0]])
                scale = float(np.abs(ref_logits[i]).max())
                assert prec == C.PREC_BF16 and margin <= 2 * err * scale + 1e-6, \
                    f"{name} step {i}: token {out[0]} != {ref_toks[i]} (margin {margin}, err {err})"
                flips += 1
            toks.append(int(ref_toks[i]))
        assert flips <= 2, f"{name}: {flips} near-tie flips in {steps} steps"
        if flips:
            print(f"{name}: {flips} near-tie token(s) inside the numerical error")
        return worst
    finally:
        run.close()


@pytest.mark.parametrize("name", list(TINY))
def test_fp32_logits_and_greedy_tokens(name):
    _greedy_parity(name, C.PREC_FP32, prompt_len=40, steps=32, strict_seed=STRICT_SEEDS[name])


@pytest.mark.parametrize("name", list(TINY))
def test_bf16_logits_and_greedy_tokens(name):
    """north_star: <= 2e-2 and identical greedy tokens for the first 32 steps, free-running."""
    _greedy_parity(name, C.PREC_BF16, prompt_len=40, steps=32, strict_seed=STRICT_SEEDS[name])


@pytest.mark.parametrize("name", list(TINY_HD))
@pytest.mark.parametrize("prec", [C.PREC_FP32, C.PREC_BF16])
def test_head_dim_64_128_families(name, prec):
    """head_dim 64 / 128: bf16 mode runs the tensor-core attention kernels (prefill tiles of 64
    queries, decode with the query-head group as MMA rows, MQA group of 18 heads in 2 chunks).
    Free-running, identical tokens."""
    _greedy_parity(name, prec, prompt_len=100, steps=12, table=TINY_HD,
                   strict_seed=STRICT_SEEDS[name])


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


def _safe_uniform(p, target):
    """A uniform whose inverse-CDF draw from the distribution p is unambiguous: the midpoint of
    a CDF bin at least 4e-4 wide, nearest to `target` — >= 2e-4 away from every CDF step, far
    outside what fp32 summation order can move (the oracle test uses the same exclusion,
    tests/test_oracle.py)."""
    cdf = np.cumsum(p.astype(np.float64))
    lo = np.concatenate([[0.0], cdf[:-1]])
    ok = np.nonzero((cdf - lo) >= 4e-4)[0]
    assert len(ok), "distribution too flat for an unambiguous draw"
    mid = 0.5 * (lo[ok] + cdf[ok])
    j = int(np.argmin(np.abs(mid - target)))
    return float(np.float32(mid[j])), int(ok[j])


def test_temperature_sampling_with_injected_uniforms():
    """sampling.go:70-74,105-128,196-217 with rand.Float32() replaced by a given uniform: index
    work, so the drawn tokens must EQUAL the oracle's.  A first greedy pass yields the logits, the
    uniforms are then placed mid-bin of each row's CDF."""
    O = oracle_mod()
    cfg = TINY["gpt2"]()
    run = _runner(cfg, C.PREC_FP32)
    try:
        prompts = [W.random_prompt(cfg.vocab_size, 12 + i, 50 + i).tolist() for i in range(6)]
        temps = [0.7, 1.0, 1.3, 0.0, 2.0, 0.5]
        targets = [0.01, 0.35, 0.62, 0.5, 0.93, 0.999]
        bts = [[i] for i in range(6)]
        _t, logits0 = run.run_flat(prompts, True, [0] * 6, bts, [0.0] * 6, want_logits=True)
        us, expect = [], []
        for i in range(6):
            if temps[i] == 0.0:
                us.append(0.5)
                expect.append(O.argmax(logits0[i]))
                continue
            u, tok = _safe_uniform(ref_filter_probs(logits0[i], [], temps[i], 1.0, 0, 1.0), targets[i])
            us.append(u)
            expect.append(tok)
        toks, logits = run.run_flat(prompts, True, [0] * 6, bts, temps, us, want_logits=True)
        np.testing.assert_array_equal(logits, logits0)  # same step, same kernels: bit-identical
        for i in range(6):
            ref = O.argmax(logits[i]) if temps[i] == 0.0 else O.sample(logits[i], None, temps[i], u=us[i])
            assert ref == expect[i], f"row {i}: oracle {ref} vs float64 restatement {expect[i]}"
            assert int(toks[i]) == ref, f"row {i}: device drew {toks[i]}, oracle {ref}"
    finally:
        run.close()


@pytest.mark.parametrize("top_p,top_k,rep", [(1.0, 0, 1.2), (1.0, 40, 1.0), (0.9, 0, 1.0),
                                              (0.85, 50, 1.2), (0.5, 5, 1.5)])
def test_sample_with_history_filters(top_p, top_k, rep):
    """sampling.go:33-217 SampleWithHistory on the device: repetition penalty over the token
    history (x3 for the last ten), temperature, top-k, top-p, renormalise, inverse-CDF draw —
    rand.Float32() replaced by injected uniforms placed mid-bin of the filtered CDF.  Tokens
    must EQUAL the oracle's on the prefill step and on the decode step after it."""
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
        targets = [float(u) for u in rng.uniform(0.02, 0.98, size=8)]
        bts = [[i] for i in range(8)]

        def step(seqs, prefill):
            # pass 1 (greedy rows) for the logits, pass 2 with uniforms placed mid-bin
            nc = [0] * 8 if prefill else None
            _t, lg0 = run.run_flat(seqs, prefill, nc, bts, [0.0] * 8, want_logits=True)
            us = []
            for i in range(8):
                if temps[i] == 0.0:
                    us.append(0.5)
                    continue
                p = ref_filter_probs(lg0[i], seqs[i], temps[i], top_p, top_k, rep)
                us.append(_safe_uniform(p, targets[i])[0])
            toks, lg = run.run_flat(seqs, prefill, nc, bts, temps, us, want_logits=True)
            np.testing.assert_array_equal(lg, lg0)
            for i in range(8):
                if temps[i] == 0.0:
                    assert toks[i] == O.argmax(lg[i])  # greedy rows are not filtered
                    continue
                ref = O.sample(lg[i], seqs[i], temps[i], top_p=top_p, top_k=top_k, rep_penalty=rep,
                               u=us[i])
                assert int(toks[i]) == ref, f"row {i}: device drew {toks[i]}, oracle {ref} (u={us[i]})"
            return toks

        toks = step(prompts, True)
        # decode step: the history is the whole sequence so far
        step([pr + [int(t)] for pr, t in zip(prompts, toks)], False)
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
        R.CudaModelRunner(cfg, ws, precision=C.PREC_FP32, num_kv_blocks=4, sampling=None)
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


def test_resident_ragged_batch_stops_at_each_sequences_own_table():
    """in a ragged batch the shorter table is padded with block 0 on the device; the
    resident loop must refuse to step a sequence past ITS OWN table (NVLLM_ERR_RANGE) instead of
    appending its K/V into pool block 0, which here belongs to the other sequence."""
    L = nvllm_pkg.sub("lib")
    cfg = TINY["llama"]()
    pa = W.random_prompt(cfg.vocab_size, 250, 901).tolist()   # one block: positions 0..255
    pb = W.random_prompt(cfg.vocab_size, 40, 902).tolist()    # owns pool block 0
    steps = 12

    def stepwise():
        run = _runner(cfg, C.PREC_FP32)
        try:
            run.set_speculation(False)
            seqs, bts = [list(pa), list(pb)], [[5, 6], [0, 1]]
            out, _ = run.run_flat(seqs, True, [0, 0], bts, [0.0, 0.0])
            toks = [out.copy()]
            for _ in range(steps):
                for s_, t in zip(seqs, out):
                    s_.append(int(t))
                out, _ = run.run_flat(seqs, False, None, bts, [0.0, 0.0])
                toks.append(out.copy())
            return toks
        finally:
            run.close()

    ref = stepwise()
    run = _runner(cfg, C.PREC_FP32)
    try:
        run.set_speculation(False)
        seqs = [list(pa), list(pb)]
        out, _ = run.run_flat(seqs, True, [0, 0], [[5], [0, 1]], [0.0, 0.0])
        np.testing.assert_array_equal(out, ref[0])
        for s_, t in zip(seqs, out):
            s_.append(int(t))
        # 250-token context + 6 appended tokens fill block 5 exactly
        for i in range(1, 7):
            run.run_resident()
            got, _ = run.sync(2)
            np.testing.assert_array_equal(got, ref[i])
            for s_, t in zip(seqs, got):
                s_.append(int(t))
        with pytest.raises(L.NvllmError) as e:
            run.run_resident()
        assert e.value.code == L.ERR_RANGE and "sequence 0" in str(e.value)
        # the engine allocates the next block and carries on: nothing of sequence 1 was damaged
        for i in range(7, steps + 1):
            out, _ = run.run_flat(seqs, False, None, [[5, 6], [0, 1]], [0.0, 0.0])
            np.testing.assert_array_equal(out, ref[i])
            for s_, t in zip(seqs, out):
                s_.append(int(t))
    finally:
        run.close()


@pytest.mark.parametrize("batch", [1, 64])
def test_resident_and_speculative_loops_1000_steps(batch):
    """the device-resident loop (advance_kernel -> forward pass, captured in
    a CUDA graph, every kernel released by programmatic dependent launch) against stepping
    through nvllm_run with speculation off, token for token over 1000 steps; then the same loop
    through nvllm_run with the speculative next step on."""
    cfg = TINY["llama"]()
    steps = 1000
    prompts = [W.random_prompt(cfg.vocab_size, 8 + (i % 13), 7000 + i).tolist() for i in range(batch)]
    nb = (8 + 12 + steps + 2 + 255) // 256
    bts = [list(range(i * nb, (i + 1) * nb)) for i in range(batch)]
    zeros = [0.0] * batch

    def drive(mode):
        run = _runner(cfg, C.PREC_FP32, num_kv_blocks=batch * nb)
        try:
            run.set_speculation(mode == "spec")
            seqs = [list(p) for p in prompts]
            out, _ = run.run_flat(seqs, True, [0] * batch, bts, zeros)
            hist = [out.copy()]
            for s_, t in zip(seqs, out):
                s_.append(int(t))
            out, _ = run.run_flat(seqs, False, None, bts, zeros)
            hist.append(out.copy())
            hits = 0
            for _ in range(steps - 1):
                if mode == "resident":
                    run.run_resident()
                    out, _ = run.sync(batch)
                else:
                    for s_, t in zip(seqs, out):
                        s_.append(int(t))
                    out, _ = run.run_flat(seqs, False, None, bts, zeros)
                    hits += int(run.last_stats.h2d_bytes == 0)
                hist.append(out.copy())
            return np.stack(hist), hits
        finally:
            run.close()

    ref, _ = drive("stepwise")
    got, _ = drive("resident")
    np.testing.assert_array_equal(got, ref)
    got, hits = drive("spec")
    np.testing.assert_array_equal(got, ref)
    # every step hits except the ones that need a new KV block's first position staged by the host
    assert hits >= steps - 16, hits
