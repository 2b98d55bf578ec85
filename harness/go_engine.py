"""Host-side stand-in for the Go engine that calls ``ModelRunner.Run``.

The north-star keeps the engine (scheduler, block manager, sequence lifecycle)
in Go and replaces only what sits behind ``ModelRunner``.  There is no Go
toolchain in this image, so to drive the C-ABI library with exactly the inputs
the Go engine would produce (block tables, NumCachedTokens, prefill-first
batches, preemption) this module restates that bookkeeping in Python:

  Sequence        /root/reference nanovllm/sequence.go:15-114
  SamplingParams  nanovllm/sampling_params.go
  EngineConfig    nanovllm/config.go:9-70
  BlockManager    nanovllm/block_manager.go  (xxhash64 prefix cache, ref-counts)
  Scheduler       nanovllm/scheduler.go      (prefill-first, decode w/ preemption)
  LLMEngine       nanovllm/llm_engine.go:25-97 (AddRequest / Step / Generate, no tokenizer)

It computes nothing on tensors (neither does nanovllm/); it is not part of the
hot path and has no CUDA in it.  The reference's own unit tests for this
bookkeeping (block_manager_test.go, sequence_test.go) are ported in tests/.
"""
from __future__ import annotations

import itertools
import struct
from collections import deque
from dataclasses import dataclass, field
from typing import Deque, Dict, List, Optional, Sequence as Seq, Tuple

import xxhash  # cespare/xxhash/v2 v2.3.0 == XXH64 seed 0 (go.mod:6)

STATUS_WAITING, STATUS_RUNNING, STATUS_FINISHED = 0, 1, 2


@dataclass
class SamplingParams:  # sampling_params.go:4-40
    temperature: float = 1.0
    max_tokens: int = 64
    ignore_eos: bool = False

    def __post_init__(self):
        if self.temperature <= 1e-10:
            # the Go constructor panics: greedy is not reachable through SamplingParams
            raise ValueError("greedy sampling is not permitted")


@dataclass
class EngineConfig:  # config.go:9-70
    max_num_batched_tokens: int = 16384
    max_num_seqs: int = 512
    max_model_len: int = 4096
    tensor_parallel_size: int = 1
    eos: int = -1
    kvcache_block_size: int = 256
    num_kvcache_blocks: int = -1

    def __post_init__(self):
        if self.kvcache_block_size % 256 != 0:
            raise ValueError("kvcache_block_size must be divisible by 256")
        if not 1 <= self.tensor_parallel_size <= 8:
            raise ValueError("tensor_parallel_size must be between 1 and 8")
        if self.max_num_batched_tokens < self.max_model_len:
            raise ValueError("max_num_batched_tokens must be >= max_model_len")


_seq_counter = itertools.count()


class Sequence:  # sequence.go:15-114
    def __init__(self, token_ids: Seq[int], sampling: Optional[SamplingParams] = None,
                 temperature: Optional[float] = None):
        sp = sampling or SamplingParams()
        self.seq_id = next(_seq_counter)
        self.status = STATUS_WAITING
        self.token_ids: List[int] = [int(t) for t in token_ids]
        self.last_token = self.token_ids[-1]
        self.num_tokens = len(self.token_ids)
        self.num_prompt_tokens = len(self.token_ids)
        self.num_cached_tokens = 0
        self.block_table: List[int] = []
        # `temperature` override lets tests/bench request greedy (0.0), which the Go
        # SamplingParams cannot express but cmd/ask does (main.go:79-85).
        self.temperature = sp.temperature if temperature is None else temperature
        self.max_tokens = sp.max_tokens
        self.ignore_eos = sp.ignore_eos
        self.block_size = 256  # sequence.go:52

    def __len__(self):
        return self.num_tokens

    def is_finished(self):
        return self.status == STATUS_FINISHED

    def num_completion_tokens(self):
        return self.num_tokens - self.num_prompt_tokens

    def completion_token_ids(self):
        return self.token_ids[self.num_prompt_tokens:]

    def num_cached_blocks(self):
        return self.num_cached_tokens // self.block_size

    def num_blocks(self):
        return (self.num_tokens + self.block_size - 1) // self.block_size

    def last_block_num_tokens(self):
        return self.num_tokens - (self.num_blocks() - 1) * self.block_size

    def block(self, i: int) -> Optional[List[int]]:
        if i < 0 or i >= self.num_blocks():
            return None
        return self.token_ids[i * self.block_size:(i + 1) * self.block_size]

    def append_token(self, token_id: int):
        self.token_ids.append(int(token_id))
        self.last_token = int(token_id)
        self.num_tokens += 1


@dataclass
class Block:  # block_manager.go:10-39
    block_id: int
    ref_count: int = 0
    hash: int = 0
    token_ids: List[int] = field(default_factory=list)

    def update(self, h: int, token_ids: Seq[int]):
        self.hash = h
        self.token_ids = list(token_ids)

    def reset(self):
        self.ref_count = 1
        self.hash = 0
        self.token_ids = []


class BlockManager:  # block_manager.go:41-263
    def __init__(self, num_blocks: int, block_size: int):
        self.block_size = block_size
        self.blocks = [Block(i) for i in range(num_blocks)]
        self.hash_to_block_id: Dict[int, int] = {}
        self.free_block_ids: List[int] = list(range(num_blocks))
        self.used_block_ids: set = set()

    @staticmethod
    def compute_hash(token_ids: Seq[int], prefix_hash: int = 0) -> int:
        """XXH64(seed 0) over [LE64 prefix hash if != 0] || LE32 tokens (block_manager.go:72-88)."""
        h = xxhash.xxh64()
        if prefix_hash != 0:
            h.update(struct.pack("<Q", prefix_hash))
        h.update(struct.pack(f"<{len(token_ids)}I", *[t & 0xFFFFFFFF for t in token_ids]))
        return h.intdigest()

    def _allocate_block(self, block_id: int) -> Block:
        block = self.blocks[block_id]
        if block.ref_count != 0:
            raise RuntimeError("block is already allocated")
        block.reset()
        self.free_block_ids.remove(block_id)
        self.used_block_ids.add(block_id)
        return block

    def _deallocate_block(self, block_id: int):
        if self.blocks[block_id].ref_count != 0:
            raise RuntimeError("block still has references")
        self.used_block_ids.discard(block_id)
        self.free_block_ids.append(block_id)

    def can_allocate(self, seq: Sequence) -> bool:
        return len(self.free_block_ids) >= seq.num_blocks()

    def allocate(self, seq: Sequence):  # block_manager.go:128-203
        if seq.block_table:
            raise RuntimeError("sequence already has blocks allocated")
        h = 0
        cache_miss = False
        for i in range(seq.num_blocks()):
            token_ids = seq.block(i)
            h = self.compute_hash(token_ids, h) if len(token_ids) == self.block_size else 0
            block_id = self.hash_to_block_id.get(h, -1) if h != 0 else -1
            if block_id != -1 and self.blocks[block_id].token_ids != token_ids:
                block_id = -1
            if block_id == -1:
                cache_miss = True
            if cache_miss:
                block_id = self.free_block_ids[0]
                self._allocate_block(block_id)
            else:
                seq.num_cached_tokens += self.block_size
                if block_id in self.used_block_ids:
                    self.blocks[block_id].ref_count += 1
                else:
                    self._allocate_block(block_id)
            if h != 0:
                self.blocks[block_id].update(h, token_ids)
                self.hash_to_block_id[h] = block_id
            seq.block_table.append(block_id)

    def deallocate(self, seq: Sequence):  # :206-219
        for block_id in reversed(seq.block_table):
            block = self.blocks[block_id]
            block.ref_count -= 1
            if block.ref_count == 0:
                self._deallocate_block(block_id)
        seq.num_cached_tokens = 0
        seq.block_table = []

    def can_append(self, seq: Sequence) -> bool:  # :222-228
        if len(seq) % self.block_size == 1:
            return len(self.free_block_ids) >= 1
        return True

    def may_append(self, seq: Sequence):  # :231-263
        last_block = self.blocks[seq.block_table[-1]]
        m = len(seq) % self.block_size
        if m == 1:
            if last_block.hash == 0:
                raise RuntimeError("last block should have a hash")
            block_id = self.free_block_ids[0]
            self._allocate_block(block_id)
            seq.block_table.append(block_id)
        elif m == 0:
            if last_block.hash != 0:
                raise RuntimeError("last block should not have a hash")
            token_ids = seq.block(seq.num_blocks() - 1)
            prefix = self.blocks[seq.block_table[-2]].hash if len(seq.block_table) > 1 else 0
            h = self.compute_hash(token_ids, prefix)
            last_block.update(h, token_ids)
            self.hash_to_block_id[h] = last_block.block_id
        else:
            if last_block.hash != 0:
                raise RuntimeError("last block should not have a hash")


class Scheduler:  # scheduler.go:6-140
    def __init__(self, cfg: EngineConfig):
        num_blocks = cfg.num_kvcache_blocks if cfg.num_kvcache_blocks != -1 else 1024
        self.max_num_seqs = cfg.max_num_seqs
        self.max_num_batched_tokens = cfg.max_num_batched_tokens
        self.eos = cfg.eos
        self.block_manager = BlockManager(num_blocks, cfg.kvcache_block_size)
        self.waiting: Deque[Sequence] = deque()
        self.running: Deque[Sequence] = deque()

    def is_finished(self):
        return not self.waiting and not self.running

    def add(self, seq: Sequence):
        self.waiting.append(seq)

    def schedule(self) -> Tuple[List[Sequence], bool]:  # :45-112
        scheduled: List[Sequence] = []
        num_seqs = 0
        num_batched = 0
        bm = self.block_manager
        while self.waiting and num_seqs < self.max_num_seqs:
            seq = self.waiting[0]
            if num_batched + len(seq) > self.max_num_batched_tokens or not bm.can_allocate(seq):
                break
            num_seqs += 1
            bm.allocate(seq)
            num_batched += len(seq) - seq.num_cached_tokens
            seq.status = STATUS_RUNNING
            self.waiting.popleft()
            self.running.append(seq)
            scheduled.append(seq)
        if scheduled:
            return scheduled, True

        while self.running and num_seqs < self.max_num_seqs:
            seq = self.running.popleft()
            while not bm.can_append(seq):
                if self.running:
                    self._preempt(self.running.pop())
                else:
                    self._preempt(seq)
                    break
            if seq.status == STATUS_RUNNING:
                num_seqs += 1
                bm.may_append(seq)
                scheduled.append(seq)
        if not scheduled:
            raise RuntimeError("no sequences scheduled")
        self.running.extendleft(reversed(scheduled))
        return scheduled, False

    def _preempt(self, seq: Sequence):  # :115-119
        seq.status = STATUS_WAITING
        self.block_manager.deallocate(seq)
        self.waiting.appendleft(seq)

    def postprocess(self, seqs: List[Sequence], token_ids: Seq[int]):  # :122-140
        for seq, tok in zip(seqs, token_ids):
            seq.append_token(int(tok))
            if (not seq.ignore_eos and tok == self.eos) or \
                    seq.num_completion_tokens() == seq.max_tokens:
                seq.status = STATUS_FINISHED
                self.block_manager.deallocate(seq)
                for s in self.running:
                    if s.seq_id == seq.seq_id:
                        self.running.remove(s)
                        break


class MockModelRunner:  # model_runner.go:19-55
    def __init__(self, eos: int = -1, vocab: int = 32000):
        self.eos, self.vocab = eos, vocab

    def run(self, seqs: List[Sequence], is_prefill: bool) -> List[int]:
        out = []
        for s in seqs:
            t = (s.seq_id + s.num_tokens) % self.vocab
            n = s.num_completion_tokens()
            if n > 10 and n % 20 == 0:
                t = self.eos
            out.append(t)
        return out

    def close(self):
        pass


class LLMEngine:  # llm_engine.go:25-203 without the tokenizer / progress bar
    def __init__(self, cfg: EngineConfig, model_runner):
        self.cfg = cfg
        self.model_runner = model_runner
        self.scheduler = Scheduler(cfg)

    def add_request(self, token_ids: Seq[int], sampling: Optional[SamplingParams] = None,
                    temperature: Optional[float] = None) -> Sequence:
        seq = Sequence(token_ids, sampling, temperature)
        self.scheduler.add(seq)
        return seq

    def step(self):  # :62-97
        seqs, is_prefill = self.scheduler.schedule()
        token_ids = self.model_runner.run(seqs, is_prefill)
        self.scheduler.postprocess(seqs, token_ids)
        finished = [s for s in seqs if s.is_finished()]
        num_tokens = sum(len(s) - s.num_cached_tokens for s in seqs) if is_prefill else -len(seqs)
        return finished, num_tokens

    def is_finished(self):
        return self.scheduler.is_finished()

    def generate(self, prompts: Seq[Seq[int]], sampling: Optional[SamplingParams] = None,
                 temperature: Optional[float] = None) -> List[List[int]]:
        seqs = [self.add_request(p, sampling, temperature) for p in prompts]
        while not self.is_finished():
            self.step()
        return [s.completion_token_ids() for s in seqs]

    def close(self):
        self.model_runner.close()
