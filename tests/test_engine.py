"""Ports of the reference's own unit tests for the bookkeeping that produces the hot path's
integer inputs (block tables, NumCachedTokens): nanovllm/block_manager_test.go and
nanovllm/sequence_test.go, against the restatement in harness/go_engine.py; plus
scheduler behaviour (prefill-first, preemption) driven with the MockModelRunner
(nanovllm/model_runner.go:19-55)."""
import struct

import pytest
import xxhash

from helpers import E


def _seq(tokens, **kw):
    return E.Sequence(tokens, E.SamplingParams(**kw))


# ---- block_manager_test.go -------------------------------------------------
def test_block_manager_creation():  # :7-21
    bm = E.BlockManager(100, 256)
    assert len(bm.blocks) == 100 and len(bm.free_block_ids) == 100 and bm.block_size == 256


def test_block_manager_allocate():  # :23-47
    bm = E.BlockManager(100, 256)
    seq = _seq(list(range(300)))
    assert bm.can_allocate(seq)
    bm.allocate(seq)
    assert len(seq.block_table) == 2
    assert len(bm.free_block_ids) == 98


def test_block_manager_deallocate():  # :49-73
    bm = E.BlockManager(100, 256)
    seq = _seq(list(range(300)))
    bm.allocate(seq)
    bm.deallocate(seq)
    assert seq.block_table == [] and len(bm.free_block_ids) == 100 and seq.num_cached_tokens == 0


def test_block_manager_prefix_caching():  # :75-110
    bm = E.BlockManager(100, 256)
    s1, s2 = _seq(list(range(256))), _seq(list(range(256)))
    bm.allocate(s1)
    free1 = len(bm.free_block_ids)
    bm.allocate(s2)
    assert s2.num_cached_tokens == 256
    assert s2.block_table == s1.block_table
    assert len(bm.free_block_ids) == free1  # shared, ref-counted
    assert bm.blocks[s1.block_table[0]].ref_count == 2


def test_block_manager_compute_hash():  # :112-129 + the XXH64 definition of :72-88
    h1 = E.BlockManager.compute_hash([1, 2, 3, 4, 5], 0)
    assert h1 == E.BlockManager.compute_hash([1, 2, 3, 4, 5], 0)
    assert h1 != E.BlockManager.compute_hash([1, 2, 3, 4, 6], 0)
    assert h1 == xxhash.xxh64(struct.pack("<5I", 1, 2, 3, 4, 5), seed=0).intdigest()
    hp = E.BlockManager.compute_hash([7, 8], 0xDEADBEEF)
    assert hp == xxhash.xxh64(struct.pack("<Q", 0xDEADBEEF) + struct.pack("<2I", 7, 8)).intdigest()


# ---- sequence_test.go ------------------------------------------------------
def test_sequence_creation():  # :7-31
    seq = _seq([1, 2, 3, 4, 5], temperature=0.8, max_tokens=100)
    assert len(seq) == 5 and seq.num_prompt_tokens == 5 and seq.num_completion_tokens() == 0
    assert seq.status == E.STATUS_WAITING


def test_sequence_append_token():  # :33-51
    seq = _seq([1, 2, 3])
    seq.append_token(4)
    assert len(seq) == 4 and seq.last_token == 4 and seq.num_completion_tokens() == 1


def test_sequence_blocks():  # :53-78
    seq = _seq(list(range(600)))
    assert seq.num_blocks() == 3
    assert len(seq.block(0)) == 256
    assert len(seq.block(2)) == 600 - 2 * 256
    assert seq.last_block_num_tokens() == 88


def test_sampling_params():  # :80-98
    sp = E.SamplingParams(temperature=0.7, max_tokens=128, ignore_eos=True)
    assert sp.temperature == 0.7 and sp.max_tokens == 128 and sp.ignore_eos


def test_sampling_params_validation():  # :100-109 (Go panics)
    with pytest.raises(ValueError):
        E.SamplingParams(temperature=0.0)


# ---- scheduler / engine (scheduler.go, llm_engine.go) with the mock backend -------------
def test_engine_config_validation():  # config.go:52-70
    with pytest.raises(ValueError):
        E.EngineConfig(kvcache_block_size=100)
    with pytest.raises(ValueError):
        E.EngineConfig(tensor_parallel_size=9)
    with pytest.raises(ValueError):
        E.EngineConfig(max_num_batched_tokens=100, max_model_len=200)


def test_scheduler_prefill_first_then_decode():
    eng = E.LLMEngine(E.EngineConfig(num_kvcache_blocks=16), E.MockModelRunner())
    a = eng.add_request(list(range(10)), E.SamplingParams(max_tokens=3))
    b = eng.add_request(list(range(300)), E.SamplingParams(max_tokens=3))
    seqs, is_prefill = eng.scheduler.schedule()
    assert is_prefill and [s.seq_id for s in seqs] == [a.seq_id, b.seq_id]
    assert len(a.block_table) == 1 and len(b.block_table) == 2
    eng.scheduler.postprocess(seqs, [1, 2])
    seqs, is_prefill = eng.scheduler.schedule()
    assert not is_prefill and len(seqs) == 2


def test_scheduler_token_budget_splits_prefill():  # scheduler.go:51-56
    cfg = E.EngineConfig(max_num_batched_tokens=4096, max_model_len=4096, num_kvcache_blocks=64)
    eng = E.LLMEngine(cfg, E.MockModelRunner())
    for i in range(3):
        eng.add_request([i] * 2000, E.SamplingParams(max_tokens=2))
    seqs, is_prefill = eng.scheduler.schedule()
    assert is_prefill and len(seqs) == 2  # the third would exceed 4096 batched tokens
    eng.scheduler.postprocess(seqs, [5, 5])
    seqs, is_prefill = eng.scheduler.schedule()
    assert is_prefill and len(seqs) == 1


def test_scheduler_preemption_when_pool_is_full():  # scheduler.go:74-119
    cfg = E.EngineConfig(num_kvcache_blocks=2)
    eng = E.LLMEngine(cfg, E.MockModelRunner())
    a = eng.add_request(list(range(256)), E.SamplingParams(max_tokens=8, ignore_eos=True))
    b = eng.add_request(list(range(1000, 1256)), E.SamplingParams(max_tokens=8, ignore_eos=True))
    seqs, is_prefill = eng.scheduler.schedule()
    assert is_prefill and len(seqs) == 2
    eng.scheduler.postprocess(seqs, [1, 1])  # both now have 257 tokens and need a second block
    seqs, is_prefill = eng.scheduler.schedule()
    # no free block: the last running sequence is preempted, then a itself
    assert not is_prefill or len(seqs) >= 1
    assert b.status == E.STATUS_WAITING or a.status == E.STATUS_WAITING


def test_generate_with_mock_runner_finishes():
    eng = E.LLMEngine(E.EngineConfig(num_kvcache_blocks=32, eos=-1), E.MockModelRunner())
    outs = eng.generate([[1, 2, 3], list(range(40))], E.SamplingParams(max_tokens=5))
    assert [len(o) for o in outs] == [5, 5]
    assert eng.is_finished()
    assert len(eng.scheduler.block_manager.free_block_ids) == 32
