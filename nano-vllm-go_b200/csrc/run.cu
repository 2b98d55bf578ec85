// run.cu — nvllm_run and its device-resident siblings: one engine step (ModelRunner.Run,
// reference nanovllm/tensor_model_runner.go:55-97).  The caller's flattened Sequence records
// are validated, turned into the step's integer arrays (new-token ids, positions, KV slots
// through the block tables, ragged offsets, attention tiles) in ONE pinned buffer, copied
// with ONE cudaMemcpyAsync, and the forward pass (forward.cu) runs on them.
#include <string.h>

#include <algorithm>
#include <vector>

#include "common.cuh"
#include "handle.h"

using namespace nvllm;

static int ensure_stage(nvllm_handle* h, size_t words) {
  if (words <= h->stage_words) return NVLLM_OK;
  NV_CHECK(h, cudaStreamSynchronize(h->stream));
  if (h->h_stage) cudaFreeHost(h->h_stage);
  if (h->d_stage) cudaFree(h->d_stage);
  h->h_stage = nullptr;
  h->d_stage = nullptr;
  h->stage_words = 0;
  h->resident_ok = false;
  h->run_lay_ok = false;
  size_t cap = std::max(words * 2, (size_t)4096);
  NV_CHECK(h, cudaMallocHost((void**)&h->h_stage, cap * 4));
  NV_CHECK(h, cudaMalloc((void**)&h->d_stage, cap * 4));
  h->stage_words = cap;
  return NVLLM_OK;
}

// offsets of a step's arrays inside its staging buffer (every array 16-byte aligned)
static void plan_layout(StepLayout& L, int T, int n, int max_blocks, int n_qtiles) {
  L.T = T;
  L.n = n;
  L.max_blocks = max_blocks;
  L.n_qtiles = n_qtiles;
  size_t o = 0;
  auto take = [&](size_t cnt) {
    size_t r = o;
    o += (cnt + 3) / 4 * 4;
    return r;
  };
  L.tokens = take(T);
  L.positions = take(T);
  L.slots = take(T);
  L.cu_q = take(n + 1);
  L.seq_start = take(n);
  L.seq_len = take(n);
  L.last_row = take(n);
  L.block_tables = take((size_t)n * max_blocks);
  L.tile_seq = take(n_qtiles);
  L.tile_q0 = take(n_qtiles);
  L.temperature = take(n);
  L.uniforms = take(n);
  L.words = o;
}

// splitmix64 -> uniform in [0,1): the library's own stream when the caller passes no uniforms
static float next_uniform(nvllm_handle* h) {
  uint64_t z = (h->seed + 0x9E3779B97F4A7C15ull * (++h->rng_counter));
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  return (float)(z >> 40) * (1.0f / 16777216.0f);
}

static void fill_stats(nvllm_handle* h, nvllm_run_stats* stats, int T, uint64_t h2d, uint64_t d2h,
                       bool timed) {
  if (!stats) return;
  memset(stats, 0, sizeof *stats);
  if (timed) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, h->ev0, h->ev1) == cudaSuccess) stats->device_ms = ms;
  }
  stats->kernel_launches = h->launches;
  stats->new_tokens = T;
  stats->h2d_bytes = h2d;
  stats->d2h_bytes = d2h;
}

// ---- CUDA graph replay of a decode step.  Every kernel argument of a decode step is a device
// pointer or a constant of the launch shape (the step state itself lives in device memory), so
// the ~100 launches are captured once — programmatic dependent-launch edges included — and
// replayed with one cudaGraphLaunch until the shape changes (staging buffer, batch,
// block-table width, KV split count).  slot 0 = nvllm_run_resident (`pre` launches the advance
// kernel), slot 1 = nvllm_run's decode step (`pre` does nothing; the H2D copy stays outside).
static bool graph_usable(const nvllm_handle* h, int n) {
  return h->graphs_enabled && !h->profile && !trace_active() && h->tp == 1 && h->skip_mask == 0 &&
         h->hist_n == 0 && h->ws.T_cap >= n && h->ws.n_cap >= n;
}
template <typename Pre>
static int graphed_step(nvllm_handle* h, int slot, Pre pre, bool* used) {
  const StepLayout& L = h->lay;
  cudaStream_t st = h->stream;
  nvllm_handle::GraphSig sig;
  sig.d = L.d;
  sig.n = L.n;
  sig.max_blocks = L.max_blocks;
  sig.n_splits = attn_decode_splits(h->adt, h->kvdt, L.n, h->nh, h->nkv, h->hd, L.max_len,
                                    h->sm_count);
  if (!h->step_graph[slot] || !(sig == h->graph_sig[slot])) {
    if (h->step_graph[slot]) cudaGraphExecDestroy(h->step_graph[slot]);
    h->step_graph[slot] = nullptr;
    NV_CHECK(h, cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    int rc = pre();
    if (!rc) rc = forward_step(h);
    cudaGraph_t g = nullptr;
    cudaError_t e = cudaStreamEndCapture(st, &g);
    if (rc) {
      if (g) cudaGraphDestroy(g);
      return rc;
    }
    if (e != cudaSuccess) return set_err(h, NVLLM_ERR_CUDA, "graph capture: %s", cudaGetErrorString(e));
    e = cudaGraphInstantiate(&h->step_graph[slot], g, 0);
    cudaGraphDestroy(g);
    if (e != cudaSuccess) {
      h->step_graph[slot] = nullptr;
      return set_err(h, NVLLM_ERR_CUDA, "graph instantiate: %s", cudaGetErrorString(e));
    }
    h->graph_sig[slot] = sig;
    h->graph_launches[slot] = h->launches;
    h->graph_logits[slot] = h->ws.logits_final;
  }
  NV_CHECK(h, cudaGraphLaunch(h->step_graph[slot], st));
  h->launches = h->graph_launches[slot];
  h->ws.logits_final = h->graph_logits[slot];
  *used = true;
  return NVLLM_OK;
}

// ---- speculative next step.  ModelRunner.Run is synchronous (the engine needs the tokens to
// schedule the next step), so between two calls the GPU would idle for the host's turnaround
// (sync, Go/cgo, staging, launch: ~100 us of a ~1.1 ms cfg2 step).  A decode batch is almost
// always followed by the same batch one token longer, so nvllm_run ends by issuing exactly that
// step on the device (what nvllm_run_resident does).  The next call either asks for it — same
// sequences, one token longer, last token = the one just returned, same block tables, greedy —
// and only has to wait for it, or asks for something else and restages everything as usual.
// A speculative step that is thrown away has only written K/V at the next position of each
// sequence, a slot nothing reads before its owner writes it.
static int resident_step(nvllm_handle* h, nvllm_run_stats* stats, bool* used_graph = nullptr);
static void remember_call(nvllm_handle* h, int n, const int32_t* cu_seqlens, const int32_t* block_tables,
                          const int32_t* cu_blocks) {
  h->spec_len.resize(n);
  h->spec_nb.resize(n);
  for (int i = 0; i < n; i++) {
    h->spec_len[i] = cu_seqlens[i + 1] - cu_seqlens[i];
    h->spec_nb[i] = cu_blocks[i + 1] - cu_blocks[i];
  }
  h->spec_bt.assign(block_tables + cu_blocks[0], block_tables + cu_blocks[n]);
  h->res_len = h->spec_len;
  h->res_nb = h->spec_nb;
}
static void try_speculate(nvllm_handle* h, int n, const float* temperature) {
  h->spec_pending = false;
  if (!h->spec_enabled || !graph_usable(h, n)) return;
  if (h->spec_skip > 0) {
    h->spec_skip--;
    return;
  }
  for (int i = 0; i < n; i++) {
    if (temperature && temperature[i] > 0.f) return;  // a draw needs the next call's uniform
    if (h->spec_len[i] >= h->spec_nb[i] * KV_BLOCK) return;  // next position needs a new block
    if (h->use_rope && h->spec_len[i] + 1 > h->cfg.max_seq_len) return;
  }
  // a failed speculative step must leave no trace: layout, per-sequence extents and the
  // caller-visible error string are restored
  const int saved = h->launches;
  const StepLayout lay0 = h->lay;
  const std::vector<int32_t> len0 = h->res_len;
  const std::string err0 = h->err;
  h->spec_used_graph = false;
  if (resident_step(h, nullptr, &h->spec_used_graph) == NVLLM_OK) {
    h->spec_pending = true;
  } else {
    h->lay = lay0;
    h->res_len = len0;
    h->err = err0;
    h->resident_ok = false;  // the device state may be half advanced: restage on the next call
  }
  h->launches = saved;
}
static bool speculation_hit(const nvllm_handle* h, int n, int is_prefill, const int32_t* token_ids,
                            const int32_t* cu_seqlens, const int32_t* block_tables,
                            const int32_t* cu_blocks, const float* temperature) {
  if (is_prefill || n != (int)h->spec_len.size() || h->filters_active()) return false;
  size_t o = 0;
  for (int i = 0; i < n; i++) {
    const int len = cu_seqlens[i + 1] - cu_seqlens[i], nb = cu_blocks[i + 1] - cu_blocks[i];
    if (len != h->spec_len[i] + 1 || nb != h->spec_nb[i]) return false;
    if (token_ids[cu_seqlens[i + 1] - 1] != h->h_out[i]) return false;
    if (temperature && temperature[i] > 0.f) return false;
    if (memcmp(block_tables + cu_blocks[i], h->spec_bt.data() + o, (size_t)nb * 4) != 0) return false;
    o += nb;
  }
  return true;
}

extern "C" int nvllm_run(nvllm_handle* h, int n, int is_prefill, const int32_t* token_ids,
                         const int32_t* cu_seqlens, const int32_t* num_cached,
                         const int32_t* block_tables, const int32_t* cu_blocks,
                         const float* temperature, const float* uniforms, int32_t* out_tokens,
                         float* out_logits, nvllm_run_stats* stats) {
  if (!h) return NVLLM_ERR_INVALID;
  if (!h->finalized) return set_err(h, NVLLM_ERR_STATE, "nvllm_run before nvllm_finalize");
  if (n < 0 || (n > 0 && (!token_ids || !cu_seqlens || !block_tables || !cu_blocks || !out_tokens)))
    return set_err(h, NVLLM_ERR_INVALID, "null argument");
  if (stats) memset(stats, 0, sizeof *stats);
  if (n == 0) return NVLLM_OK;
  NV_CHECK(h, cudaSetDevice(h->dev));
  const nvllm_config& c = h->cfg;
  if (h->spec_pending) {
    h->spec_pending = false;
    if (speculation_hit(h, n, is_prefill, token_ids, cu_seqlens, block_tables, cu_blocks, temperature)) {
      // the requested step is already running (or done): collect it
      cudaStream_t st = h->stream;
      NV_CHECK(h, cudaMemcpyAsync(h->h_out, h->ws.out_tokens, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
      uint64_t d2h = (uint64_t)n * 4;
      if (out_logits) {
        NV_CHECK(h, cudaMemcpy2DAsync(out_logits, (size_t)h->V * 4, h->ws.logits_final,
                                      (size_t)h->ldl * 4, (size_t)h->V * 4, n, cudaMemcpyDeviceToHost, st));
        d2h += (uint64_t)n * h->V * 4;
      }
      cudaError_t e = cudaStreamSynchronize(st);
      if (e != cudaSuccess)
        return set_err(h, NVLLM_ERR_CUDA, "forward pass failed: %s", cudaGetErrorString(e));
      memcpy(out_tokens, h->h_out, (size_t)n * 4);
      h->run_lay = h->lay;
      h->run_lay_ok = true;
      fill_stats(h, stats, n, 0, d2h, true);
      if (stats) stats->used_graph = h->spec_used_graph ? 1 : 0;
      remember_call(h, n, cu_seqlens, block_tables, cu_blocks);
      try_speculate(h, n, temperature);
      return NVLLM_OK;
    }
    h->spec_skip = 1;  // missed: the caller is not stepping one batch; back off once
  }
  h->resident_ok = false;
  h->run_lay_ok = false;

  // ---- pass 1: sizes and validation
  int T = 0, max_blocks = 0, n_qtiles = 0, max_len = 0;
  const int qtile = attn_prefill_qtile(h->adt, h->kvdt, h->hd);
  long long sum_len = 0;
  double pairs = 0;
  for (int i = 0; i < n; i++) {
    const int len = cu_seqlens[i + 1] - cu_seqlens[i];
    const int nb = cu_blocks[i + 1] - cu_blocks[i];
    if (len <= 0) return set_err(h, NVLLM_ERR_INVALID, "sequence %d is empty", i);
    if (nb < (len + KV_BLOCK - 1) / KV_BLOCK)
      return set_err(h, NVLLM_ERR_INVALID, "sequence %d: %d tokens need %d blocks, table has %d", i,
                     len, (len + KV_BLOCK - 1) / KV_BLOCK, nb);
    if (h->use_rope && len > c.max_seq_len)  // rope.go:84-86 panics
      return set_err(h, NVLLM_ERR_RANGE, "sequence %d: position %d >= max_seq_len %d", i, len - 1,
                     c.max_seq_len);
    int start = len - 1;
    if (is_prefill) {
      int nc = num_cached ? num_cached[i] : 0;
      if (nc < 0) return set_err(h, NVLLM_ERR_INVALID, "sequence %d: num_cached < 0", i);
      start = std::min(nc, len - 1);  // a fully cached prompt still computes its last token
      // Mamba2 layers: the reference's causal conv restarts (zero padding) at the first token of
      // every forward call (mamba2.go:209-270 never reads ConvCache), so a prompt computed in two
      // calls differs from the same prompt in one.  A cached prefix is therefore only accepted
      // when the call is decode-shaped (one new token, the reference's own decode semantics).
      if (h->n_mamba && start > 0 && len - start > 1)
        return set_err(h, NVLLM_ERR_INVALID,
                       "sequence %d: hybrid (Mamba2) models cannot prefill %d tokens on top of %d cached "
                       "ones; disable prefix caching / chunked prefill for them", i, len - start, start);
    }
    const int n_new = len - start;
    T += n_new;
    n_qtiles += (n_new + qtile - 1) / qtile;
    max_blocks = std::max(max_blocks, nb);
    max_len = std::max(max_len, len);
    sum_len += len;
    pairs += 0.5 * ((double)len * (len + 1) - (double)start * (start + 1));
  }
  StepLayout& L = h->lay;
  plan_layout(L, T, n, max_blocks, is_prefill ? n_qtiles : 0);
  L.prefill = is_prefill != 0;
  L.max_len = max_len;
  L.sum_len = sum_len;
  L.attn_pairs = pairs;
  int rc;
  if ((rc = ensure_stage(h, L.words))) return rc;
  L.d = h->d_stage;
  if ((rc = ensure_workspace(h, T, n))) return rc;
  if (!h->h_out || n > h->ws.n_cap_out) {
    NV_CHECK(h, cudaStreamSynchronize(h->stream));
    if (h->h_out) cudaFreeHost(h->h_out);
    h->h_out = nullptr;
    NV_CHECK(h, cudaMallocHost((void**)&h->h_out, (size_t)std::max(n, 512) * 4));
    h->ws.n_cap_out = std::max(n, 512);
  }
  // ---- pass 2: fill
  int32_t* s = h->h_stage;
  float* s_temp = reinterpret_cast<float*>(s + L.temperature);
  float* s_uni = reinterpret_cast<float*>(s + L.uniforms);
  int row = 0, tile = 0;
  for (int i = 0; i < n; i++) {
    const int len = cu_seqlens[i + 1] - cu_seqlens[i];
    const int32_t* toks = token_ids + cu_seqlens[i];
    const int32_t* bt = block_tables + cu_blocks[i];
    const int nb = cu_blocks[i + 1] - cu_blocks[i];
    int start = len - 1;
    if (is_prefill) start = std::min(num_cached ? num_cached[i] : 0, len - 1);
    for (int b = 0; b < nb; b++) {
      if (bt[b] < 0 || bt[b] >= h->num_blocks)
        return set_err(h, NVLLM_ERR_RANGE, "sequence %d: block id %d outside the pool of %d", i,
                       bt[b], h->num_blocks);
      s[L.block_tables + (size_t)i * max_blocks + b] = bt[b];
    }
    for (int b = nb; b < max_blocks; b++) s[L.block_tables + (size_t)i * max_blocks + b] = 0;
    s[L.cu_q + i] = row;
    s[L.seq_start + i] = start;
    s[L.seq_len + i] = len;
    for (int p = start; p < len; p++, row++) {
      const int tok = toks[p];
      if (tok < 0 || tok >= h->V)
        return set_err(h, NVLLM_ERR_RANGE, "sequence %d: token id %d outside vocab %d", i, tok, h->V);
      s[L.tokens + row] = tok;
      s[L.positions + row] = p;
      s[L.slots + row] = bt[p / KV_BLOCK] * KV_BLOCK + p % KV_BLOCK;
    }
    s[L.last_row + i] = row - 1;
    if (is_prefill)
      for (int q0 = 0; q0 < len - start; q0 += qtile, tile++) {
        s[L.tile_seq + tile] = i;
        s[L.tile_q0 + tile] = q0;
      }
    const float t = temperature ? temperature[i] : 0.f;
    s_temp[i] = t;
    s_uni[i] = (t > 0.f) ? (uniforms ? uniforms[i] : next_uniform(h)) : 0.f;
  }
  s[L.cu_q + n] = row;

  cudaStream_t st = h->stream;
  // SampleWithHistory filters need every sequence's full token history on the device
  // (previousTokens = seq.TokenIDs, tensor_model_runner.go:94)
  h->hist_n = 0;
  if (h->filters_active()) {
    bool any = false;
    for (int i = 0; i < n && !any; i++) any = temperature && temperature[i] > 0.f;
    if (any) {
      const size_t total = (size_t)(cu_seqlens[n] - cu_seqlens[0]);
      const size_t words = (size_t)(n + 1) + total;
      if (words > h->hist_words) {
        NV_CHECK(h, cudaStreamSynchronize(st));
        if (h->d_hist) cudaFree(h->d_hist);
        h->d_hist = nullptr;
        h->hist_words = 0;
        NV_CHECK(h, cudaMalloc((void**)&h->d_hist, words * 2 * 4));
        h->hist_words = words * 2;
      }
      const size_t pf = (size_t)n * h->V;
      if (pf > h->probs_floats) {
        NV_CHECK(h, cudaStreamSynchronize(st));
        if (h->d_probs) cudaFree(h->d_probs);
        h->d_probs = nullptr;
        h->probs_floats = 0;
        NV_CHECK(h, cudaMalloc((void**)&h->d_probs, pf * 4));
        h->probs_floats = pf;
      }
      // offsets relative to the first token, then the tokens themselves (the caller's buffers
      // are pageable: these copies are staged by the driver before the call returns)
      std::vector<int32_t> cu(n + 1);
      for (int i = 0; i <= n; i++) cu[i] = cu_seqlens[i] - cu_seqlens[0];
      NV_CHECK(h, cudaMemcpyAsync(h->d_hist, cu.data(), (size_t)(n + 1) * 4, cudaMemcpyHostToDevice, st));
      NV_CHECK(h, cudaMemcpyAsync(h->d_hist + (n + 1), token_ids + cu_seqlens[0], total * 4,
                                  cudaMemcpyHostToDevice, st));
      NV_CHECK(h, cudaStreamSynchronize(st));  // `cu` is a stack temporary
      h->hist_n = n;
    }
  }
  NV_CHECK(h, cudaMemcpyAsync(h->d_stage, h->h_stage, L.words * 4, cudaMemcpyHostToDevice, st));
  NV_CHECK(h, cudaEventRecord(h->ev0, st));
  h->launches = 0;
  bool used_graph = false;
  if (!is_prefill && graph_usable(h, n)) {
    if ((rc = graphed_step(h, 1, [] { return (int)NVLLM_OK; }, &used_graph))) return rc;
  } else if ((rc = is_prefill ? forward_step_auto(h) : forward_step(h))) {
    return rc;
  }
  NV_CHECK(h, cudaEventRecord(h->ev1, st));
  NV_CHECK(h, cudaMemcpyAsync(h->h_out, h->ws.out_tokens, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
  uint64_t d2h = (uint64_t)n * 4;
  if (out_logits) {
    // rows are ldl apart on the device; copy them packed [n, V]
    if (h->tp > 1) {  // gathered rank-major vocabulary slices [tp][n][Vl]
      for (int r = 0; r < h->tp; r++) {
        const int c0 = r * h->Vl, cn = std::min(h->Vl, h->V - c0);
        if (cn <= 0) break;
        NV_CHECK(h, cudaMemcpy2DAsync(out_logits + c0, (size_t)h->V * 4,
                                      h->ws.logits_final + (size_t)r * n * h->Vl,
                                      (size_t)h->Vl * 4, (size_t)cn * 4, n,
                                      cudaMemcpyDeviceToHost, st));
      }
    } else {
      NV_CHECK(h, cudaMemcpy2DAsync(out_logits, (size_t)h->V * 4, h->ws.logits_final,
                                    (size_t)h->ldl * 4, (size_t)h->V * 4, n,
                                    cudaMemcpyDeviceToHost, st));
    }
    d2h += (uint64_t)n * h->V * 4;
  }
  cudaError_t e = cudaStreamSynchronize(st);
  if (e != cudaSuccess)
    return set_err(h, NVLLM_ERR_CUDA, "forward pass failed: %s", cudaGetErrorString(e));
  memcpy(out_tokens, h->h_out, (size_t)n * 4);
  h->resident_ok = true;
  h->run_lay = L;
  h->run_lay_ok = true;
  fill_stats(h, stats, T, L.words * 4, d2h, true);
  if (stats) stats->used_graph = used_graph ? 1 : 0;
  remember_call(h, n, cu_seqlens, block_tables, cu_blocks);
  try_speculate(h, n, temperature);
  return NVLLM_OK;
}

// One more decode step for the sequences of the previous step, entirely on the device: the
// sampled tokens become the inputs, positions advance by one, KV slots come from the block
// tables already resident.  No host<->device copy, no host sync.  Works after a decode step
// (state advanced in place) and after a prefill step (decode state derived by a kernel).
static int resident_step(nvllm_handle* h, nvllm_run_stats* stats, bool* used_graph) {
  if (!h->resident_ok)
    return set_err(h, NVLLM_ERR_STATE, "nvllm_run_resident needs a preceding nvllm_run");
  if (h->hist_n)
    return set_err(h, NVLLM_ERR_STATE,
                   "repetition penalty / top-k / top-p need the token history: use nvllm_run");
  NV_CHECK(h, cudaSetDevice(h->dev));
  StepLayout& L = h->lay;
  const int n = L.n;
  if (L.max_len + 1 > L.max_blocks * KV_BLOCK || (h->use_rope && L.max_len + 1 > h->cfg.max_seq_len))
    return set_err(h, NVLLM_ERR_RANGE, "resident decode ran past the block tables / max_seq_len");
  // ragged batches: every sequence against ITS OWN table (entries past it are padding)
  if ((int)h->res_len.size() != n || (int)h->res_nb.size() != n)
    return set_err(h, NVLLM_ERR_STATE, "nvllm_run_resident: no per-sequence state for this batch");
  for (int i = 0; i < n; i++)
    if (h->res_len[i] + 1 > h->res_nb[i] * KV_BLOCK)
      return set_err(h, NVLLM_ERR_RANGE,
                     "resident decode: sequence %d would grow to %d tokens, its block table holds %d",
                     i, h->res_len[i] + 1, h->res_nb[i] * KV_BLOCK);
  for (int i = 0; i < n; i++) h->res_len[i] += 1;
  cudaStream_t st = h->stream;
  int rc;
  NV_CHECK(h, cudaEventRecord(h->ev0, st));
  h->launches = 0;
  if (L.prefill) {
    StepLayout D;
    plan_layout(D, n, n, L.max_blocks, 0);
    if (D.words > h->stage2_words) {
      NV_CHECK(h, cudaStreamSynchronize(st));
      if (h->d_stage2) cudaFree(h->d_stage2);
      h->d_stage2 = nullptr;
      h->stage2_words = 0;
      NV_CHECK(h, cudaMalloc((void**)&h->d_stage2, D.words * 2 * 4));
      h->stage2_words = D.words * 2;
    }
    D.d = h->d_stage2;
    const int32_t* s = L.d;
    int32_t* d = D.d;
    cudaError_t e = launch_make_decode_state(
        st, n, L.max_blocks, h->ws.out_tokens, s + L.seq_len, s + L.block_tables, s + L.temperature,
        s + L.uniforms, d + D.tokens, d + D.positions, d + D.slots, d + D.cu_q, d + D.seq_start,
        d + D.seq_len, d + D.last_row, d + D.block_tables, d + D.temperature, d + D.uniforms);
    h->launches++;
    if (e != cudaSuccess)
      return set_err(h, NVLLM_ERR_CUDA, "make_decode_state: %s", cudaGetErrorString(e));
    D.prefill = false;
    D.max_len = L.max_len + 1;
    D.sum_len = L.sum_len + n;
    D.attn_pairs = (double)D.sum_len;
    L = D;
  } else {
    int32_t* d = L.d;
    L.max_len += 1;
    L.sum_len += n;
    L.attn_pairs = (double)L.sum_len;
    auto advance = [&]() -> int {
      cudaError_t e = launch_advance(st, n, h->ws.out_tokens, d + L.tokens, d + L.positions,
                                     d + L.slots, d + L.seq_len, d + L.block_tables, L.max_blocks,
                                     L.max_blocks * KV_BLOCK);
      h->launches++;
      if (e != cudaSuccess) return set_err(h, NVLLM_ERR_CUDA, "advance: %s", cudaGetErrorString(e));
      return NVLLM_OK;
    };
    if (graph_usable(h, n)) {
      bool used = false;
      if ((rc = graphed_step(h, 0, advance, &used))) return rc;
      NV_CHECK(h, cudaEventRecord(h->ev1, st));
      fill_stats(h, stats, n, 0, 0, false);
      if (stats) stats->used_graph = 1;
      if (used_graph) *used_graph = true;
      return NVLLM_OK;
    }
    if ((rc = advance())) return rc;
  }
  if ((rc = forward_step(h))) return rc;
  NV_CHECK(h, cudaEventRecord(h->ev1, st));
  fill_stats(h, stats, n, 0, 0, false);
  return NVLLM_OK;
}

extern "C" int nvllm_run_resident(nvllm_handle* h, nvllm_run_stats* stats) {
  if (!h) return NVLLM_ERR_INVALID;
  if (h->spec_pending) {  // the step asked for is the one already in flight
    h->spec_pending = false;
    fill_stats(h, stats, h->lay.n, 0, 0, false);
    if (stats) stats->used_graph = h->spec_used_graph ? 1 : 0;
    return NVLLM_OK;
  }
  return resident_step(h, stats);
}

// Re-execute the last nvllm_run step from its staged device inputs.
extern "C" int nvllm_rerun(nvllm_handle* h, nvllm_run_stats* stats) {
  if (!h) return NVLLM_ERR_INVALID;
  if (!h->run_lay_ok) return set_err(h, NVLLM_ERR_STATE, "nvllm_rerun needs a preceding nvllm_run");
  NV_CHECK(h, cudaSetDevice(h->dev));
  if (!h->run_lay.prefill)
    return set_err(h, NVLLM_ERR_STATE,
                   "nvllm_rerun replays prefill steps only (a decode step's staged state is "
                   "advanced in place by nvllm_run_resident)");
  h->lay = h->run_lay;
  h->spec_pending = false;
  h->res_len = h->spec_len;  // the device state is the replayed step's again
  h->res_nb = h->spec_nb;
  NV_CHECK(h, cudaEventRecord(h->ev0, h->stream));
  h->launches = 0;
  int rc = forward_step_auto(h);
  if (rc) return rc;
  NV_CHECK(h, cudaEventRecord(h->ev1, h->stream));
  h->resident_ok = true;
  fill_stats(h, stats, h->lay.T, 0, 0, false);
  return NVLLM_OK;
}

// Block until the steps issued so far are done; copies the last sampled tokens out.
extern "C" int nvllm_sync(nvllm_handle* h, int32_t* out_tokens, int n, float* last_step_ms) {
  if (!h) return NVLLM_ERR_INVALID;
  NV_CHECK(h, cudaSetDevice(h->dev));
  cudaError_t e = cudaStreamSynchronize(h->stream);
  if (e != cudaSuccess)
    return set_err(h, NVLLM_ERR_CUDA, "stream failed: %s", cudaGetErrorString(e));
  if (out_tokens && n > 0) {
    if (n > h->lay.n) n = h->lay.n;
    if (h->spec_pending)  // ws.out_tokens already holds the speculative step's tokens
      memcpy(out_tokens, h->h_out, (size_t)n * 4);
    else
      NV_CHECK(h, cudaMemcpy(out_tokens, h->ws.out_tokens, (size_t)n * 4, cudaMemcpyDeviceToHost));
  }
  if (last_step_ms) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, h->ev0, h->ev1) != cudaSuccess) ms = 0.f;
    *last_step_ms = ms;
  }
  return NVLLM_OK;
}

// Device timer on the compute stream, for callers that time several calls as one region
// (bench.py): which = 0 records the start event, 1 the stop event.
extern "C" int nvllm_timer_record(nvllm_handle* h, int which) {
  if (!h) return NVLLM_ERR_INVALID;
  NV_CHECK(h, cudaSetDevice(h->dev));
  if (!h->tm0) {
    NV_CHECK(h, cudaEventCreate(&h->tm0));
    NV_CHECK(h, cudaEventCreate(&h->tm1));
  }
  NV_CHECK(h, cudaEventRecord(which ? h->tm1 : h->tm0, h->stream));
  return NVLLM_OK;
}
// Synchronises the stream and returns stop - start in milliseconds.
extern "C" int nvllm_timer_elapsed(nvllm_handle* h, float* ms) {
  if (!h || !ms) return NVLLM_ERR_INVALID;
  NV_CHECK(h, cudaSetDevice(h->dev));
  if (!h->tm0) return set_err(h, NVLLM_ERR_STATE, "timer was never started");
  cudaError_t e = cudaStreamSynchronize(h->stream);
  if (e != cudaSuccess) return set_err(h, NVLLM_ERR_CUDA, "stream failed: %s", cudaGetErrorString(e));
  NV_CHECK(h, cudaEventElapsedTime(ms, h->tm0, h->tm1));
  return NVLLM_OK;
}

// ---- diagnostics: device timeline (see include/nvllm.h)
namespace nvllm {
void trace_set(unsigned long long* buf);
}
static unsigned long long* g_trace_dev = nullptr;
static int g_trace_cap = 0;
extern "C" int nvllm_trace_begin(nvllm_handle* h, int capacity) {
  if (!h || capacity <= 0) return NVLLM_ERR_INVALID;
  NV_CHECK(h, cudaSetDevice(h->dev));
  NV_CHECK(h, cudaStreamSynchronize(h->stream));
  if (g_trace_dev) cudaFree(g_trace_dev);
  g_trace_dev = nullptr;
  const size_t words = 8 + (size_t)capacity * 8;
  NV_CHECK(h, cudaMalloc((void**)&g_trace_dev, words * 8));
  NV_CHECK(h, cudaMemset(g_trace_dev, 0, words * 8));
  unsigned long long cap = (unsigned long long)capacity;
  NV_CHECK(h, cudaMemcpy(g_trace_dev + 1, &cap, 8, cudaMemcpyHostToDevice));
  g_trace_cap = capacity;
  trace_set(g_trace_dev);
  return NVLLM_OK;
}
extern "C" int nvllm_trace_end(nvllm_handle* h, uint64_t* out, int max_records, int* n_records) {
  if (!h || !n_records) return NVLLM_ERR_INVALID;
  NV_CHECK(h, cudaSetDevice(h->dev));
  trace_set(nullptr);
  if (!g_trace_dev) return set_err(h, NVLLM_ERR_STATE, "trace was not started");
  NV_CHECK(h, cudaStreamSynchronize(h->stream));
  unsigned long long n = 0;
  NV_CHECK(h, cudaMemcpy(&n, g_trace_dev, 8, cudaMemcpyDeviceToHost));
  if (n > (unsigned long long)g_trace_cap) n = g_trace_cap;
  if ((int)n > max_records) n = max_records;
  if (out && n) NV_CHECK(h, cudaMemcpy(out, g_trace_dev + 8, n * 64, cudaMemcpyDeviceToHost));
  *n_records = (int)n;
  cudaFree(g_trace_dev);
  g_trace_dev = nullptr;
  return NVLLM_OK;
}
