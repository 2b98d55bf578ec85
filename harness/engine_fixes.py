"""SURVEY.md §8f rank 3 — the engine-side fixes that limit batch efficiency, restated on top of
the Go bookkeeping of harness/go_engine.py (the Go versions are go/engine_fixes.go; this image
has no Go toolchain, so the Python mirror is what the GPU tests drive):

  1. MixedScheduler: ONE step carries decode rows and prompt chunks together ("chunked
     prefill").  The reference's Scheduler.Schedule (nanovllm/scheduler.go:45-112) runs a
     prefill-only step whenever a request is waiting, so every arrival stalls all running
     decodes, and a prompt longer than MaxNumBatchedTokens can never be scheduled at all.  The
     library needs nothing new for this: nvllm_run(is_prefill = 1) computes tokens
     [num_cached, len) of every sequence, and num_cached may be ANY count of tokens whose K/V
     is already in the pool — a decode row is simply num_cached = len - 1.
  2. generate() returns the outputs BY REQUEST (seq id), not in completion order
     (llm_engine.go:169-181 appends outputs as sequences finish, so the i-th output is not the
     i-th prompt's).
  3. num_kv_blocks_from_free_hbm: sizes NumKVCacheBlocks from the device memory that is left
     after the weights (Config.GPUMemoryUtilization exists, config.go:17, and is never read;
     scheduler.go:17-21 hard-wires 1024 blocks).
  4. The block size: Sequence hard-wires 256 (sequence.go:52) while Config.KVCacheBlockSize only
     has to be a multiple of 256 (config.go:57) — a config of 512 makes BlockManager and
     Sequence disagree.  The library's pool uses 256-token blocks (NVLLM_KV_BLOCK_SIZE), so the
     fix is to reject anything else up front: check_block_size().
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Sequence as Seq, Tuple

from . import go_engine as G


@dataclass
class StepItem:
    seq: G.Sequence
    num_cached: int   # tokens of the sequence whose K/V is already in the pool
    num_new: int      # tokens this step computes: [num_cached, num_cached + num_new)
    emits: bool       # the step's sampled token is the sequence's next token


class MixedScheduler(G.Scheduler):
    """Decode rows first (they never wait for a prompt), then prompt chunks of running
    sequences, then newly admitted prompts — all inside one MaxNumBatchedTokens budget."""

    def __init__(self, cfg: G.EngineConfig, chunk_tokens: int = 512):
        super().__init__(cfg)
        self.chunk_tokens = chunk_tokens
        self.computed: Dict[int, int] = {}  # seq id -> tokens with K/V in the pool
        self.alloc_len: Dict[int, int] = {}  # seq id -> tokens the admission allocated blocks for
        # blocks whose 256 tokens all have their K/V in the pool.  Allocate() publishes a prompt's
        # block hashes immediately (block_manager.go:128-203); with chunked prompts a later request
        # could otherwise "hit" a block whose chunk has not run yet.
        self.block_ready: set = set()

    def schedule_mixed(self) -> List[StepItem]:
        bm = self.block_manager
        items: List[StepItem] = []
        budget = self.max_num_batched_tokens
        # 1. sequences already running: a decode row, or the next chunk of their prompt
        for seq in list(self.running):
            if len(items) >= self.max_num_seqs or budget <= 0:
                break
            if seq.status != G.STATUS_RUNNING:
                continue  # preempted a moment ago to make room for an earlier sequence
            done = self.computed[seq.seq_id]
            if done >= self.alloc_len[seq.seq_id]:
                # decode (every token Allocate() sized the block table for has been computed): may need a new block (block_manager.go:222-263); preempt from the
                # back like scheduler.go:82-95 when the pool is full
                while not bm.can_append(seq):
                    victim = self.running[-1]
                    if victim is seq:
                        break
                    self.running.pop()
                    self._preempt_mixed(victim)
                    items = [it for it in items if it.seq is not victim]
                if not bm.can_append(seq):
                    self.running.remove(seq)
                    self._preempt_mixed(seq)
                    continue
                nb = len(seq.block_table)
                bm.may_append(seq)
                if len(seq.block_table) > nb:
                    self.block_ready.discard(seq.block_table[-1])
                items.append(StepItem(seq, len(seq) - 1, 1, True))
                budget -= 1
            else:
                take = min(len(seq) - done, self.chunk_tokens, budget)
                items.append(StepItem(seq, done, take, done + take == len(seq)))
                budget -= take
        # 2. admit waiting prompts while the budget and the pool allow (the whole block table is
        # allocated up front, exactly as Schedule does, so prefix-cache hits work unchanged)
        while self.waiting and len(items) < self.max_num_seqs and budget > 0:
            seq = self.waiting[0]
            if not bm.can_allocate(seq):
                break
            bm.allocate(seq)
            seq.status = G.STATUS_RUNNING
            self.waiting.popleft()
            self.running.append(seq)
            hit = seq.num_cached_tokens // 256
            for i, b in enumerate(seq.block_table):
                if i < hit and b not in self.block_ready:
                    hit = i  # the owner's chunk has not reached this block yet: compute it here too
                if i >= hit:
                    self.block_ready.discard(b)
            done = min(hit * 256, len(seq) - 1)  # a fully cached prompt recomputes its last token
            self.computed[seq.seq_id] = done
            self.alloc_len[seq.seq_id] = len(seq)
            take = min(len(seq) - done, self.chunk_tokens, budget)
            items.append(StepItem(seq, done, take, done + take == len(seq)))
            budget -= take
        if not items:
            raise RuntimeError("no sequences scheduled")
        return items

    def _preempt_mixed(self, seq: G.Sequence):
        self._preempt(seq)
        self.computed.pop(seq.seq_id, None)
        self.alloc_len.pop(seq.seq_id, None)

    def postprocess_mixed(self, items: Seq[StepItem], token_ids: Seq[int]):
        for it, tok in zip(items, token_ids):
            seq = it.seq
            self.computed[seq.seq_id] = it.num_cached + it.num_new
            for i in range((it.num_cached + it.num_new) // 256):
                self.block_ready.add(seq.block_table[i])
            if not it.emits:
                continue  # a prompt chunk that does not reach the prompt's end: its sample is discarded
            seq.append_token(int(tok))
            if (not seq.ignore_eos and tok == self.eos) or seq.num_completion_tokens() == seq.max_tokens:
                seq.status = G.STATUS_FINISHED
                self.block_manager.deallocate(seq)
                self.running.remove(seq)
                self.computed.pop(seq.seq_id, None)
                self.alloc_len.pop(seq.seq_id, None)


class MixedEngine:
    """LLMEngine.Step / Generate (llm_engine.go:62-70,151-181) over the mixed scheduler.  `runner`
    is a CudaModelRunner (or anything with run_flat)."""

    def __init__(self, cfg: G.EngineConfig, runner, chunk_tokens: int = 512):
        check_block_size(cfg)
        self.scheduler = MixedScheduler(cfg, chunk_tokens)
        self.runner = runner
        self.steps: List[Tuple[int, int]] = []  # (decode rows, prompt-chunk tokens) per step

    def add_request(self, token_ids: Seq[int], max_tokens: int, temperature: float = 0.0) -> G.Sequence:
        seq = G.Sequence(token_ids, G.SamplingParams(max_tokens=max_tokens, ignore_eos=True),
                         temperature=temperature)
        self.scheduler.add(seq)
        return seq

    def step(self):
        items = self.scheduler.schedule_mixed()
        toks, _ = self.runner.run_flat(
            [it.seq.token_ids[:it.num_cached + it.num_new] for it in items], True,
            num_cached=[it.num_cached for it in items],
            block_tables=[it.seq.block_table for it in items],
            temperature=[it.seq.temperature for it in items])
        self.steps.append((sum(1 for it in items if it.num_cached == len(it.seq) - 1 and it.num_new == 1),
                           sum(it.num_new for it in items if not (it.num_cached == len(it.seq) - 1 and it.num_new == 1))))
        self.scheduler.postprocess_mixed(items, [int(t) for t in toks])

    def generate(self, prompts: Seq[Seq[int]], max_tokens: int) -> List[List[int]]:
        """Completions in REQUEST order (fix 2)."""
        seqs = [self.add_request(p, max_tokens) for p in prompts]
        while not self.scheduler.is_finished():
            self.step()
        return [s.completion_token_ids() for s in seqs]


def check_block_size(cfg: G.EngineConfig):
    """Fix 4: the block size the engine plans with must be the one the runner's pool uses."""
    if cfg.kvcache_block_size != 256:
        raise ValueError(f"kvcache_block_size {cfg.kvcache_block_size}: Sequence (sequence.go:52) and the "
                         "CUDA runner's paged pool use 256-token blocks")


def num_kv_blocks_from_free_hbm(total_bytes: int, weight_bytes: int, workspace_bytes: int,
                                num_layers: int, num_kv_heads: int, head_dim: int,
                                kv_elem_bytes: int, gpu_memory_utilization: float = 0.9,
                                block_size: int = 256) -> int:
    """Fix 3: blocks = (utilization * total - weights - workspace) / bytes per block, where one
    block holds K and V of block_size tokens for every layer (the pool is
    [L][2][blocks][n_kv][256][hd])."""
    per_block = 2 * num_layers * num_kv_heads * block_size * head_dim * kv_elem_bytes
    free = int(total_bytes * gpu_memory_utilization) - weight_bytes - workspace_bytes
    return max(0, free // per_block)
