// Package nanovllm — engine-side fixes that let the CUDA runner be used at its batch
// efficiency (SURVEY.md §8f rank 3).  Drop this file next to scheduler.go / llm_engine.go.
// Like cuda_model_runner.go it is shipped UNCOMPILED (no Go toolchain in the build image);
// harness/engine_fixes.py is the same logic in Python, exercised by tests/test_engine_fixes.py
// (scripted runner on the CPU, the real library on the GPU), and tests/test_go_shim.py checks
// every selector used below against the reference's declarations.
//
//  1. ScheduleMixed / PostprocessMixed: ONE step carries decode rows and prompt chunks
//     together.  Scheduler.Schedule (scheduler.go:45-112) runs a prefill-only step whenever a
//     request waits — every arrival stalls all running decodes — and can never schedule a
//     prompt longer than MaxNumBatchedTokens.  The runner needs nothing new: nvllm_run with
//     is_prefill = 1 computes tokens [NumCachedTokens, len) of every sequence and NumCachedTokens
//     may be any count of tokens whose K/V is in the pool (a decode row is len - 1).
//  2. GenerateByRequest: outputs in request order (LLMEngine.Generate, llm_engine.go:169-181,
//     appends them in completion order).
//  3. NumKVBlocksFromFreeHBM: sizes NumKVCacheBlocks from what the device has left after the
//     weights (Config.GPUMemoryUtilization is never read; scheduler.go:17-21 hard-wires 1024).
//  4. CheckBlockSize: Sequence hard-wires 256-token blocks (sequence.go:52) while
//     Config.KVCacheBlockSize may be any multiple of 256 (config.go:57); the CUDA runner's pool
//     uses 256.
package nanovllm

import "fmt"

// StepItem is one row group of a mixed step.
type StepItem struct {
	Seq       *Sequence
	NumCached int  // tokens whose K/V is already in the pool
	NumNew    int  // tokens this step computes: [NumCached, NumCached+NumNew)
	Emits     bool // the step's sampled token is the sequence's next token
}

// MixedScheduler wraps the reference Scheduler (same queues, same BlockManager).
type MixedScheduler struct {
	*Scheduler
	ChunkTokens int
	computed    map[int64]int // SeqID -> tokens with K/V in the pool
	allocLen    map[int64]int // SeqID -> tokens Allocate sized the block table for
	blockReady  map[int]bool  // block id -> all 256 tokens have their K/V in the pool
}

func NewMixedScheduler(config *Config, chunkTokens int) *MixedScheduler {
	return &MixedScheduler{Scheduler: NewScheduler(config), ChunkTokens: chunkTokens,
		computed: map[int64]int{}, allocLen: map[int64]int{}, blockReady: map[int]bool{}}
}

func minInt(a, b int) int {
	if a < b {
		return a
	}
	return b
}

func (m *MixedScheduler) preemptMixed(seq *Sequence) {
	m.preempt(seq)
	delete(m.computed, seq.SeqID)
	delete(m.allocLen, seq.SeqID)
}

// ScheduleMixed: decode rows first (they never wait for a prompt), then the next chunk of
// every running prompt, then newly admitted prompts — inside one MaxNumBatchedTokens budget.
func (m *MixedScheduler) ScheduleMixed() []StepItem {
	s := m.Scheduler
	items := make([]StepItem, 0)
	budget := s.maxNumBatchedTokens
	running := make([]*Sequence, 0, s.running.Len())
	for e := s.running.Front(); e != nil; e = e.Next() {
		running = append(running, e.Value.(*Sequence))
	}
	remove := func(seq *Sequence) {
		for e := s.running.Front(); e != nil; e = e.Next() {
			if e.Value.(*Sequence).SeqID == seq.SeqID {
				s.running.Remove(e)
				return
			}
		}
	}
	for _, seq := range running {
		if len(items) >= s.maxNumSeqs || budget <= 0 {
			break
		}
		if seq.Status != StatusRunning {
			continue // preempted a moment ago to make room for an earlier sequence
		}
		done := m.computed[seq.SeqID]
		if done >= m.allocLen[seq.SeqID] {
			// decode row; may need a new block (block_manager.go:222-263).  Preempt from the
			// back like scheduler.go:82-95 when the pool is full.
			for !s.blockManager.CanAppend(seq) {
				last := s.running.Back().Value.(*Sequence)
				if last.SeqID == seq.SeqID {
					break
				}
				s.running.Remove(s.running.Back())
				m.preemptMixed(last)
				kept := items[:0]
				for _, it := range items {
					if it.Seq.SeqID != last.SeqID {
						kept = append(kept, it)
					}
				}
				items = kept
			}
			if !s.blockManager.CanAppend(seq) {
				remove(seq)
				m.preemptMixed(seq)
				continue
			}
			nb := len(seq.BlockTable)
			s.blockManager.MayAppend(seq)
			if len(seq.BlockTable) > nb {
				delete(m.blockReady, seq.BlockTable[len(seq.BlockTable)-1])
			}
			items = append(items, StepItem{seq, seq.Len() - 1, 1, true})
			budget--
		} else {
			take := minInt(minInt(seq.Len()-done, m.ChunkTokens), budget)
			items = append(items, StepItem{seq, done, take, done+take == seq.Len()})
			budget -= take
		}
	}
	for s.waiting.Len() > 0 && len(items) < s.maxNumSeqs && budget > 0 {
		elem := s.waiting.Front()
		seq := elem.Value.(*Sequence)
		if !s.blockManager.CanAllocate(seq) {
			break
		}
		s.blockManager.Allocate(seq)
		seq.Status = StatusRunning
		s.waiting.Remove(elem)
		s.running.PushBack(seq)
		// Allocate publishes the prompt's block hashes at once (block_manager.go:128-203); a hit
		// on a block whose chunk has not run yet is recomputed here instead of trusted
		hit := seq.NumCachedTokens / 256
		for i, b := range seq.BlockTable {
			if i < hit && !m.blockReady[b] {
				hit = i
			}
			if i >= hit {
				delete(m.blockReady, b)
			}
		}
		done := minInt(hit*256, seq.Len()-1) // a fully cached prompt recomputes its last token
		m.computed[seq.SeqID] = done
		m.allocLen[seq.SeqID] = seq.Len()
		take := minInt(minInt(seq.Len()-done, m.ChunkTokens), budget)
		items = append(items, StepItem{seq, done, take, done+take == seq.Len()})
		budget -= take
	}
	if len(items) == 0 {
		panic("no sequences scheduled")
	}
	return items
}

// PostprocessMixed mirrors Scheduler.Postprocess (scheduler.go:122-140) for a mixed step.
func (m *MixedScheduler) PostprocessMixed(items []StepItem, tokenIDs []int) {
	s := m.Scheduler
	for i, it := range items {
		seq := it.Seq
		m.computed[seq.SeqID] = it.NumCached + it.NumNew
		for b := 0; b < (it.NumCached+it.NumNew)/256; b++ {
			m.blockReady[seq.BlockTable[b]] = true
		}
		if !it.Emits {
			continue // a prompt chunk that does not reach the prompt's end: its sample is discarded
		}
		tokenID := tokenIDs[i]
		seq.AppendToken(tokenID)
		if (!seq.IgnoreEOS && tokenID == s.eos) || seq.NumCompletionTokens() == seq.MaxTokens {
			seq.Status = StatusFinished
			s.blockManager.Deallocate(seq)
			for e := s.running.Front(); e != nil; e = e.Next() {
				if e.Value.(*Sequence).SeqID == seq.SeqID {
					s.running.Remove(e)
					break
				}
			}
			delete(m.computed, seq.SeqID)
			delete(m.allocLen, seq.SeqID)
		}
	}
}

// RunMixed flattens a mixed step for CudaModelRunner: every row group is a "prefill" of tokens
// [NumCached, NumCached+NumNew) — the sequence is presented truncated to the chunk's end with
// NumCachedTokens set to what the pool already holds.
func (r *CudaModelRunner) RunMixed(items []StepItem) ([]int, error) {
	views := make([]*Sequence, len(items))
	for i, it := range items {
		v := *it.Seq
		v.TokenIDs = it.Seq.TokenIDs[:it.NumCached+it.NumNew]
		v.NumTokens = it.NumCached + it.NumNew
		v.NumCachedTokens = it.NumCached
		views[i] = &v
	}
	return r.Run(views, true)
}

// GenerateByRequest drives a mixed scheduler to completion and returns the completions in the
// order of `seqs` (request order), whatever order they finish in.
func GenerateByRequest(m *MixedScheduler, r *CudaModelRunner, seqs []*Sequence) ([][]int, error) {
	for _, seq := range seqs {
		m.Add(seq)
	}
	for !m.IsFinished() {
		items := m.ScheduleMixed()
		toks, err := r.RunMixed(items)
		if err != nil {
			return nil, err
		}
		m.PostprocessMixed(items, toks)
	}
	out := make([][]int, len(seqs))
	for i, seq := range seqs {
		out[i] = seq.CompletionTokenIDs()
	}
	return out, nil
}

// CheckBlockSize rejects a KVCacheBlockSize the runner's pool does not use.
func CheckBlockSize(config *Config) error {
	if config.KVCacheBlockSize != 256 {
		return fmt.Errorf("KVCacheBlockSize %d: Sequence and the CUDA runner's paged pool use 256-token blocks",
			config.KVCacheBlockSize)
	}
	return nil
}

// NumKVBlocksFromFreeHBM: blocks = (utilization * total - weights - workspace) / bytes per
// block; one block holds K and V of 256 tokens for every layer.
func NumKVBlocksFromFreeHBM(config *Config, totalBytes, weightBytes, workspaceBytes uint64,
	numLayers, numKVHeads, headDim, kvElemBytes int) int {
	perBlock := uint64(2 * numLayers * numKVHeads * 256 * headDim * kvElemBytes)
	budget := uint64(float64(totalBytes) * float64(config.GPUMemoryUtilization))
	if budget <= weightBytes+workspaceBytes {
		return 0
	}
	return int((budget - weightBytes - workspaceBytes) / perBlock)
}
