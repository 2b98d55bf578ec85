"""SURVEY.md §8f rank 3 (engine-side fixes): mixed prefill / decode steps with chunked prompts
through the unchanged C ABI, outputs by request, KV-pool sizing from free HBM, block-size check.
The scheduler logic runs on the CPU against a scripted runner; the GPU test drives the library
and demands the oracle's greedy tokens."""
import numpy as np
import pytest

from helpers import C, W, TINY, make_oracle
import nvllm_pkg
from harness import go_engine as G
from harness import engine_fixes as F


class ScriptedRunner:
    """Next token = (sum of the sequence's tokens) % 97: depends on the whole history, so any
    bookkeeping slip (wrong chunk, wrong order, lost token) changes the completions."""

    def __init__(self):
        self.calls = []

    def run_flat(self, token_lists, is_prefill, num_cached=None, block_tables=(), temperature=None,
                 uniforms=None, want_logits=False):
        self.calls.append([(len(t), c, list(b)) for t, c, b in zip(token_lists, num_cached, block_tables)])
        return [int(sum(t) % 97) for t in token_lists], None


def expected(prompt, n):
    toks = list(prompt)
    for _ in range(n):
        toks.append(sum(toks) % 97)
    return toks[len(prompt):]


def test_mixed_steps_chunk_long_prompts_and_keep_decoding():
    cfg = G.EngineConfig(max_num_batched_tokens=4096, max_model_len=4096, num_kvcache_blocks=64)
    run = ScriptedRunner()
    eng = F.MixedEngine(cfg, run, chunk_tokens=300)
    rng = np.random.default_rng(0)
    prompts = [list(map(int, rng.integers(0, 90, n))) for n in (40, 1000, 17, 513)]
    out = eng.generate(prompts, 9)
    assert out == [expected(p, 9) for p in prompts]          # by request, not completion order
    # the 1000-token prompt took 4 chunks of <= 300; decode rows of the short requests rode along
    assert any(d > 0 and c > 0 for d, c in eng.steps), eng.steps
    assert max(c for _d, c in eng.steps) <= 4096
    for call in run.calls:
        for n_tok, cached, bt in call:
            assert 0 <= cached < n_tok and len(bt) >= (n_tok + 255) // 256


def test_token_budget_is_respected_and_a_prompt_longer_than_the_budget_still_runs():
    cfg = G.EngineConfig(max_num_batched_tokens=4096, max_model_len=4096, num_kvcache_blocks=64)
    eng = F.MixedEngine(cfg, ScriptedRunner(), chunk_tokens=4096)
    eng.scheduler.max_num_batched_tokens = 256  # the reference's Schedule would never admit 1500 tokens
    p = list(range(1500))
    assert eng.generate([p], 3) == [expected(p, 3)]
    assert all(d + c <= 256 for d, c in eng.steps)


def test_prefix_hit_on_a_block_whose_chunk_has_not_run_is_recomputed():
    cfg = G.EngineConfig(max_num_batched_tokens=4096, max_model_len=4096, num_kvcache_blocks=64)
    run = ScriptedRunner()
    eng = F.MixedEngine(cfg, run, chunk_tokens=256)
    shared = list(range(600))
    a, b = shared + [1, 2, 3], shared + [4, 5]
    out = eng.generate([a, b], 4)
    assert out == [expected(a, 4), expected(b, 4)]
    # request b was admitted while a had computed less than its two full shared blocks: it must not
    # claim them as cached in that step
    first_b = next(c for call in run.calls for c in call if c[0] <= len(b) and c[2] and c[2][0] == run.calls[0][0][2][0] and c is not call[0])
    assert first_b[1] % 256 == 0


def test_preemption_under_a_tiny_pool_still_completes():
    cfg = G.EngineConfig(max_num_batched_tokens=4096, max_model_len=4096, num_kvcache_blocks=4)
    eng = F.MixedEngine(cfg, ScriptedRunner(), chunk_tokens=128)
    rng = np.random.default_rng(1)
    prompts = [list(map(int, rng.integers(0, 90, n))) for n in (250, 250, 250)]
    assert eng.generate(prompts, 12) == [expected(p, 12) for p in prompts]


def test_pool_sizing_and_block_size_check():
    # Llama-3.2-1B bf16: 16 layers x 8 kv heads x 64 -> 8 MiB per 256-token block
    n = F.num_kv_blocks_from_free_hbm(192 * 2**30, 3 * 2**30, 2 * 2**30, 16, 8, 64, 2, 0.9)
    assert n == int((0.9 * 192 - 5) * 2**30) // (8 * 2**20)
    assert F.num_kv_blocks_from_free_hbm(8 * 2**30, 9 * 2**30, 0, 16, 8, 64, 2) == 0
    with pytest.raises(ValueError):
        F.check_block_size(G.EngineConfig(kvcache_block_size=512))


@pytest.mark.gpu
@pytest.mark.parametrize("family", ["llama", "gpt2", "granite"])
def test_mixed_engine_on_the_library_matches_the_oracle(family):
    """fp32 mode: every request's greedy completion equals the oracle's, with prompts chunked at
    boundaries that are NOT multiples of 256 and decode rows riding in prefill steps."""
    R = nvllm_pkg.sub("runner")
    cfg = TINY[family]()
    run = R.CudaModelRunner(cfg, W.random_weights(cfg, 1234), precision=C.PREC_FP32, num_kv_blocks=24,
                            max_num_seqs=8, max_batched_tokens=1024, sampling=None)
    try:
        ecfg = G.EngineConfig(max_num_batched_tokens=4096, max_model_len=4096, num_kvcache_blocks=24)
        eng = F.MixedEngine(ecfg, run, chunk_tokens=100)
        eng.scheduler.max_num_batched_tokens = 160
        lens = (33, 300, 12, 270) if cfg.max_seq_len >= 320 else (33, 100, 12, 90)
        prompts = [[int(t) for t in W.random_prompt(cfg.vocab_size, n, 50 + i)] for i, n in enumerate(lens)]
        out = eng.generate(prompts, 6)
        orc = make_oracle(cfg)
        for p, got in zip(prompts, out):
            ref, _ = orc.greedy(np.asarray(p, dtype=np.int32), 6)
            assert got == [int(t) for t in ref]
        assert any(d > 0 and c > 0 for d, c in eng.steps)
    finally:
        run.close()


@pytest.mark.parametrize("seed", range(40))
def test_randomised_arrivals_pools_and_shared_prefixes(seed):
    """Fuzz of the mixed scheduler on the scripted runner: random prompt lengths (some sharing
    whole-block prefixes so that the prefix cache hits, some identical), requests arriving while
    others run, pools from roomy to barely enough for one sequence, random chunk sizes and token
    budgets.  Every request must finish with exactly the tokens its own history implies, and every
    call must respect the budget, the sequence limit and the block tables."""
    rng = np.random.default_rng(1000 + seed)
    # long completions cross block boundaries: with a tight pool that forces preemption
    max_new = int(rng.integers(1, 20)) if seed % 2 else int(rng.integers(200, 400))
    n_req = int(rng.integers(1, 9))
    base = list(map(int, rng.integers(0, 90, 700)))
    prompts = []
    for _ in range(n_req):
        kind = rng.integers(0, 4)
        n = int(rng.integers(1, 700))
        if kind == 0 and prompts:
            prompts.append(list(prompts[int(rng.integers(0, len(prompts)))]))           # identical
        elif kind == 1:
            prompts.append(base[:n])                                                     # shared prefix
        elif kind == 2:
            prompts.append(base[:256 * int(rng.integers(1, 3))] + list(map(int, rng.integers(0, 90, n % 50))))
        else:
            prompts.append(list(map(int, rng.integers(0, 90, n))))
    prompts = [p if p else [1] for p in prompts]
    longest = max(len(p) for p in prompts) + max_new
    need_one = (longest + 255) // 256
    blocks = int(need_one + rng.integers(0, (3 * need_one + 2) if seed % 4 else 2))
    cfg = G.EngineConfig(max_num_batched_tokens=4096, max_model_len=4096, num_kvcache_blocks=blocks,
                         max_num_seqs=int(rng.integers(1, 6)))
    run = ScriptedRunner()
    eng = F.MixedEngine(cfg, run, chunk_tokens=int(rng.choice([64, 100, 256, 300, 1024])))
    budget = int(rng.choice([64, 256, 700, 4096]))
    eng.scheduler.max_num_batched_tokens = budget
    seqs, arrivals = [], sorted(int(x) for x in rng.integers(0, 12, n_req))
    step = 0
    while len(seqs) < n_req or not eng.scheduler.is_finished():
        while len(seqs) < n_req and arrivals[len(seqs)] <= step:
            seqs.append(eng.add_request(prompts[len(seqs)], max_new))
        if not eng.scheduler.is_finished():
            eng.step()
        step += 1
        assert step < 20000, "no progress"
    for p, s in zip(prompts, seqs):
        assert s.completion_token_ids() == expected(p, max_new)
    for (d, c), call in zip(eng.steps, run.calls):
        assert d + c <= budget and len(call) <= cfg.max_num_seqs
        for n_tok, cached, bt in call:
            assert 0 <= cached < n_tok and len(bt) >= (n_tok + 255) // 256
            assert all(0 <= b < blocks for b in bt)
    # everything was given back
    assert len(eng.scheduler.block_manager.free_block_ids) == blocks
