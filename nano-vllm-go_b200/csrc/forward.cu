// forward.cu — nvllm_run: one engine step (ModelRunner.Run, reference
// nanovllm/tensor_model_runner.go:55-97) as a sequence of sm_100a kernels on one stream.
//
// Where the reference loops over sequences and runs ForwardWithCache on each
// (generic_model.go:276-480), here all scheduled sequences are one ragged batch of T new
// tokens: embedding -> per layer [norm -> fused QKV GEMM -> RoPE + paged KV append ->
// paged attention -> O GEMM -> norm -> FFN / MoE] -> final norm of the last token of each
// sequence -> LM head on those rows only -> on-device argmax / temperature sampling.
// Residual adds are folded into the following norm kernel; GEMMs emit fp32 split-K
// partials that the consumer kernel sums.
#include <string.h>

#include <algorithm>

#include "common.cuh"
#include "handle.h"

using namespace nvllm;

// ------------------------------------------------------------------ workspace
namespace nvllm {
void free_workspace(nvllm_handle* h) {
  Workspace& w = h->ws;
  void* ptrs[] = {w.x, w.xn, w.xn_f32, w.ws_a, w.ws_b, w.ws_c, w.q, w.attn_out, w.hbuf, w.xn_last,
                  w.logits_p, w.logits, w.attn_ws, w.out_tokens, w.topk_idx, w.topk_w,
                  w.sort.counts, w.sort.offsets, w.sort.cursor, w.sort.sorted_pair, w.sort.pair_row,
                  w.sort.a_rows, w.sort.tile_expert, w.sort.tile_row0, w.sort.tile_rows,
                  w.sort.n_tiles, w.moe_Y, w.moe_y};
  for (void* p : ptrs)
    if (p) cudaFree(p);
  w = Workspace();
  h->ws_bytes = 0;
}
}  // namespace nvllm

static int ensure_workspace(nvllm_handle* h, int T, int n) {
  Workspace& w = h->ws;
  if (T <= w.T_cap && n <= w.n_cap) return NVLLM_OK;
  NV_CHECK(h, cudaStreamSynchronize(h->stream));
  int Tc = std::max(T, w.T_cap), nc = std::max(n, w.n_cap);
  nvllm::free_workspace(h);
  w.T_cap = Tc;
  w.n_cap = nc;
  const size_t H = h->H, asz = dt_size(h->adt);
  const size_t rows = std::max(Tc, 512);  // room for split-K slices of skinny GEMMs
  uint64_t* acc = &h->ws_bytes;
  auto A = [&](void** p, size_t bytes) -> cudaError_t {
    *p = nullptr;
    if (!bytes) return cudaSuccess;
    bytes = (bytes + 255) / 256 * 256;
    cudaError_t e = cudaMalloc(p, bytes);
    if (e == cudaSuccess) *acc += bytes;
    return e;
  };
#define WS(ptr, bytes)                                                                    \
  {                                                                                       \
    cudaError_t _e = A((void**)&(ptr), (bytes));                                          \
    if (_e != cudaSuccess) {                                                              \
      nvllm::free_workspace(h);                                                           \
      return set_err(h, NVLLM_ERR_NOMEM, "workspace for %d tokens: %s", Tc,               \
                     cudaGetErrorString(_e));                                             \
    }                                                                                     \
  }
  WS(w.x, Tc * H * 4);
  WS(w.xn, Tc * H * asz);
  size_t a_cols = std::max(h->qkv_cols, h->moe ? 0 : h->w1_cols);
  w.ws_a_floats = rows * a_cols;
  if (h->moe) w.ws_a_floats = std::max(w.ws_a_floats, (size_t)Tc * h->topk * 2 * h->F);
  w.ws_b_floats = rows * H;
  WS(w.ws_a, w.ws_a_floats * 4);
  WS(w.ws_b, w.ws_b_floats * 4);
  if (h->parallel) WS(w.ws_c, w.ws_b_floats * 4);
  WS(w.q, (size_t)Tc * h->q_cols * asz);
  WS(w.attn_out, (size_t)Tc * h->q_cols * asz);
  WS(w.hbuf, h->moe ? (size_t)Tc * h->topk * h->F * asz : (size_t)Tc * h->F * asz);
  WS(w.xn_last, nc * H * asz);
  w.logits_p_floats = (size_t)std::max(nc, 64) * h->ldl;
  WS(w.logits_p, w.logits_p_floats * 4);
  WS(w.logits, (size_t)nc * h->ldl * 4);
  w.attn_ws_floats = (size_t)nc * h->nh * 32 * (h->hd + 2);
  WS(w.attn_ws, w.attn_ws_floats * 4);
  WS(w.out_tokens, (size_t)nc * 4);
  if (h->moe) {
    const size_t P = (size_t)Tc * h->topk;
    w.max_tiles = (int)(P / 64 + h->E);
    WS(w.xn_f32, Tc * H * 4);
    WS(w.topk_idx, P * 4);
    WS(w.topk_w, P * 4);
    WS(w.sort.counts, h->E * 4);
    WS(w.sort.offsets, (h->E + 1) * 4);
    WS(w.sort.cursor, h->E * 4);
    WS(w.sort.sorted_pair, P * 4);
    WS(w.sort.pair_row, P * 4);
    WS(w.sort.a_rows, P * 4);
    WS(w.sort.tile_expert, w.max_tiles * 4);
    WS(w.sort.tile_row0, w.max_tiles * 4);
    WS(w.sort.tile_rows, w.max_tiles * 4);
    WS(w.sort.n_tiles, 4);
    WS(w.moe_Y, P * H * 4);
    WS(w.moe_y, Tc * H * 4);
  }
#undef WS
  return build_activation_maps(h);
}

// ------------------------------------------------------------------ step staging
static int ensure_stage(nvllm_handle* h, size_t words, int n) {
  if (words > h->stage_words) {
    NV_CHECK(h, cudaStreamSynchronize(h->stream));
    if (h->h_stage) cudaFreeHost(h->h_stage);
    if (h->d_stage) cudaFree(h->d_stage);
    h->h_stage = nullptr;
    h->d_stage = nullptr;
    h->stage_words = 0;
    size_t cap = std::max(words * 2, (size_t)4096);
    NV_CHECK(h, cudaMallocHost((void**)&h->h_stage, cap * 4));
    NV_CHECK(h, cudaMalloc((void**)&h->d_stage, cap * 4));
    h->stage_words = cap;
    h->resident_ok = false;
  }
  (void)n;
  return NVLLM_OK;
}

// ------------------------------------------------------------------ kernel glue
struct Pending {  // a GEMM result still to be added to the residual stream
  const float* p = nullptr;
  int S = 0;
  size_t stride = 0;
  float m = 1.f;
  const float* bias = nullptr;
};

#define LAUNCH(h, expr)                                                                       \
  do {                                                                                        \
    cudaError_t _e = (expr);                                                                  \
    (h)->launches++;                                                                          \
    if (_e != cudaSuccess)                                                                    \
      return set_err(h, NVLLM_ERR_CUDA, "%s: %s (%s:%d)", #expr, cudaGetErrorString(_e),      \
                     __FILE__, __LINE__);                                                     \
  } while (0)

// C = A @ W^T into fp32 split-K partials at `out`
static int run_gemm(nvllm_handle* h, const void* A, int lda, const TmaDesc* tmA, const void* W,
                    const TmaDesc* tmW, int M, int N, int K, float alpha, float* out, size_t cap,
                    int ldo, Pending* res) {
  res->p = out;
  res->stride = (size_t)M * ldo;
  if (M <= 8) {
    GemmArgs g{};
    g.A = A; g.adt = h->adt; g.lda = lda; g.W = W; g.wdt = h->wdt;
    g.M = M; g.N = N; g.K = K; g.out = out; g.ldo = ldo; g.split_stride = res->stride;
    g.alpha = alpha; g.out_capacity = cap;
    LAUNCH(h, launch_gemv(h->stream, g));
    res->S = g.S;
    return NVLLM_OK;
  }
  if (h->wdt == DT_BF16 && h->use_tc5 && tmA && tmW && tc5_supported(M, N, K)) {
    Tc5Args t{};
    t.tmA = tmA; t.tmW = tmW; t.M = M; t.N = N; t.K = K; t.out = out; t.ldo = ldo;
    t.split_stride = res->stride; t.alpha = alpha; t.out_capacity = cap;
    t.S = tc5_splits(M, N, K, h->sm_count);
    while (t.S > 1 && (size_t)t.S * M * ldo > cap) t.S--;
    LAUNCH(h, launch_gemm_tc5(h->stream, t, h->sm_count));
    res->S = t.S;
    return NVLLM_OK;
  }
  GemmArgs g{};
  g.A = A; g.adt = h->adt; g.lda = lda; g.W = W; g.wdt = h->wdt;
  g.M = M; g.N = N; g.K = K; g.out = out; g.ldo = ldo; g.split_stride = res->stride;
  g.alpha = alpha; g.out_capacity = cap;
  LAUNCH(h, launch_gemm_simt(h->stream, g));
  res->S = g.S;
  return NVLLM_OK;
}

static int run_norm(nvllm_handle* h, float* x, Pending* p1, Pending* p2, const float* w,
                    const float* b, void* y, float* y_f32, const int* row_idx, int n_rows) {
  NormArgs a{};
  a.x = x;
  if (p1 && p1->p) { a.p1 = p1->p; a.S1 = p1->S; a.stride1 = p1->stride; a.m1 = p1->m; a.bias1 = p1->bias; }
  if (p2 && p2->p) { a.p2 = p2->p; a.S2 = p2->S; a.stride2 = p2->stride; a.m2 = p2->m; a.bias2 = p2->bias; }
  a.w = w; a.b = b; a.eps = h->cfg.norm_eps; a.y = y; a.ydt = h->adt; a.y_f32 = y_f32;
  a.row_idx = row_idx; a.n_rows = n_rows; a.H = h->H;
  LAUNCH(h, launch_norm(h->stream, a));
  if (p1) *p1 = Pending();
  if (p2) *p2 = Pending();
  return NVLLM_OK;
}

// ------------------------------------------------------------------ forward pass
// Inputs are the device arrays of h->lay (already uploaded).  max_len = longest context
// (tokens incl. the new ones) of any sequence in the step.
static int forward_step(nvllm_handle* h, bool is_prefill, int max_len) {
  const StepLayout& L = h->lay;
  Workspace& w = h->ws;
  const int T = L.T, n = L.n, H = h->H;
  const int32_t* d = h->d_stage;
  const int* d_tokens = d + L.tokens;
  const int* d_pos = d + L.positions;
  const int* d_slots = d + L.slots;
  const nvllm_config& c = h->cfg;
  cudaStream_t st = h->stream;
  int rc;

  LAUNCH(h, launch_embed(st, h->wdt, d_tokens, d_pos, h->wte, h->wpe, c.max_seq_len,
                         c.embedding_multiplier, w.x, T, H));
  // generic_model.go:359-365: res + rm * x when ResidualMultiplier != 0, else plain add
  const float m_res = (c.residual_multiplier != 0.f) ? c.residual_multiplier : 1.f;
  // generic_model.go:405-416 (parallel block): both multipliers or neither
  const bool par_scaled = c.attention_multiplier != 0.f && c.residual_multiplier != 0.f;
  Pending p1, p2;
  const size_t kv_layer_bytes = h->kv_layer_elems * dt_size(h->kvdt);

  AttnArgs at{};
  at.q = w.q; at.out = w.attn_out; at.qdt = h->adt; at.kvdt = h->kvdt;
  at.n_seqs = n; at.nh = h->nh; at.nkv = h->nkv; at.hd = h->hd; at.scale = h->attn_scale;
  at.block_tables = d + L.block_tables; at.max_blocks = L.max_blocks;
  at.cu_q = d + L.cu_q; at.seq_start = d + L.seq_start;
  at.tile_seq = d + L.tile_seq; at.tile_q0 = d + L.tile_q0; at.n_qtiles = L.n_qtiles;
  at.seq_len = d + L.seq_len;
  at.ws = w.attn_ws;
  if (!is_prefill) {
    const int G = h->nh / h->nkv;
    int gc = 1;
    for (int cc = 4; cc >= 1; cc--)
      if (G % cc == 0) { gc = cc; break; }
    const int groups = n * (h->nh / gc);
    int ns = (2 * h->sm_count + groups - 1) / groups;
    ns = std::min(ns, (max_len + 255) / 256);
    ns = std::max(1, std::min(ns, 32));
    at.n_splits = ns;
  }

  for (int l = 0; l < h->L; l++) {
    const LayerW& lw = h->layers[l];
    // ---- attention sub-block
    if ((rc = run_norm(h, w.x, &p1, &p2, lw.attn_norm_w, lw.attn_norm_b, w.xn, nullptr, nullptr, T)))
      return rc;
    Pending qkv;
    if ((rc = run_gemm(h, w.xn, H, &w.tm_xn, lw.wqkv, &lw.tm_wqkv, T, h->qkv_cols, H, 1.f, w.ws_a,
                       w.ws_a_floats, h->qkv_cols, &qkv)))
      return rc;
    void* kp = (char*)h->kpool + (size_t)l * kv_layer_bytes;
    void* vp = (char*)h->vpool + (size_t)l * kv_layer_bytes;
    LAUNCH(h, launch_rope_kv(st, h->adt, h->kvdt, T, h->nh, h->nkv, h->hd, qkv.p, qkv.S, qkv.stride,
                             lw.bqkv, d_pos, d_slots, h->use_rope ? h->cosT : nullptr,
                             h->use_rope ? h->sinT : nullptr, w.q, kp, vp));
    at.kpool = kp;
    at.vpool = vp;
    if (is_prefill)
      LAUNCH(h, launch_attn_prefill(st, at));
    else {
      LAUNCH(h, launch_attn_decode(st, at));
      if (at.n_splits > 1) h->launches++;
    }
    if ((rc = run_gemm(h, w.attn_out, h->q_cols, &w.tm_attn_out, lw.wo, &lw.tm_wo, T, H, h->q_cols,
                       1.f, w.ws_b, w.ws_b_floats, H, &p1)))
      return rc;
    p1.bias = lw.bo;
    p1.m = h->parallel ? (par_scaled ? c.attention_multiplier : 1.f) : m_res;
    // ---- feed-forward sub-block
    float* ffn_out = w.ws_b;
    if (h->parallel) {
      ffn_out = w.ws_c;  // FFN reads the same normed input (generic_model.go:395-418)
    } else {
      if ((rc = run_norm(h, w.x, &p1, nullptr, lw.ffn_norm_w, lw.ffn_norm_b, w.xn,
                         h->moe ? w.xn_f32 : nullptr, nullptr, T)))
        return rc;
    }
    Pending& pf = h->parallel ? p2 : p1;
    if (h->moe) {
      const int k = h->topk, I = h->F, P = T * k;
      LAUNCH(h, launch_moe_router(st, w.xn_f32, lw.router, T, H, h->E, k, w.topk_idx, w.topk_w));
      LAUNCH(h, launch_moe_sort(st, w.topk_idx, T, k, h->E, w.sort, w.max_tiles));
      GemmArgs g{};
      g.A = w.xn; g.adt = h->adt; g.lda = H; g.W = lw.moe_in; g.wdt = h->wdt;
      g.M = P; g.N = 2 * I; g.K = H; g.out = w.ws_a; g.ldo = 2 * I; g.alpha = 1.f;
      g.a_rows = w.sort.a_rows; g.tile_expert = w.sort.tile_expert; g.tile_row0 = w.sort.tile_row0;
      g.tile_rows = w.sort.tile_rows; g.n_tiles = w.sort.n_tiles;
      g.expert_stride = (size_t)2 * I * H; g.max_tiles = w.max_tiles;
      LAUNCH(h, launch_gemm_simt(st, g));
      LAUNCH(h, launch_moe_glu(st, h->adt, P, I, w.ws_a, w.hbuf));
      GemmArgs g2{};
      g2.A = w.hbuf; g2.adt = h->adt; g2.lda = I; g2.W = lw.moe_out; g2.wdt = h->wdt;
      g2.M = P; g2.N = H; g2.K = I; g2.out = w.moe_Y; g2.ldo = H; g2.alpha = 1.f;
      g2.tile_expert = w.sort.tile_expert; g2.tile_row0 = w.sort.tile_row0;
      g2.tile_rows = w.sort.tile_rows; g2.n_tiles = w.sort.n_tiles;
      g2.expert_stride = (size_t)H * I; g2.max_tiles = w.max_tiles;
      LAUNCH(h, launch_gemm_simt(st, g2));
      LAUNCH(h, launch_moe_combine(st, T, k, H, w.moe_Y, w.sort.pair_row, w.topk_w, w.moe_y));
      pf.p = w.moe_y; pf.S = 1; pf.stride = 0; pf.bias = nullptr;
    } else {
      Pending h1;
      if ((rc = run_gemm(h, w.xn, H, &w.tm_xn, lw.w1, &lw.tm_w1, T, h->w1_cols, H, 1.f, w.ws_a,
                         w.ws_a_floats, h->w1_cols, &h1)))
        return rc;
      LAUNCH(h, launch_act(st, h->adt, T, h->F, h1.p, h1.S, h1.stride, lw.b1, h->swiglu ? 1 : 0,
                           w.hbuf));
      if ((rc = run_gemm(h, w.hbuf, h->F, &w.tm_hbuf, lw.w2, &lw.tm_w2, T, H, h->F, 1.f, ffn_out,
                         w.ws_b_floats, H, &pf)))
        return rc;
      pf.bias = lw.b2;
    }
    pf.m = h->parallel ? (par_scaled ? c.residual_multiplier : 1.f) : m_res;
  }
  // final norm on the last token of each sequence only (the reference norms and projects all
  // T rows, generic_model.go:464-470, then keeps row T-1, :595-604)
  if ((rc = run_norm(h, w.x, &p1, &p2, h->lnf_w, h->lnf_b, w.xn_last, nullptr, d + L.last_row, n)))
    return rc;
  const float alpha = (c.logits_scaling != 0.f) ? 1.0f / c.logits_scaling : 1.f;
  Pending lg;
  if ((rc = run_gemm(h, w.xn_last, H, &w.tm_xn_last, h->lm_head, &h->tm_lm_head, n, h->V, H, alpha,
                     w.logits_p, w.logits_p_floats, h->ldl, &lg)))
    return rc;
  const float* logits = lg.p;
  if (lg.S > 1) {
    LAUNCH(h, launch_reduce_partials(st, lg.p, lg.S, lg.stride, (size_t)n * h->ldl, w.logits));
    logits = w.logits;
  }
  const float* d_temp = reinterpret_cast<const float*>(d + L.temperature);
  const float* d_uni = reinterpret_cast<const float*>(d + L.uniforms);
  LAUNCH(h, launch_sample(st, logits, h->ldl, h->V, n, d_temp, d_uni, w.out_tokens));
  h->ws.logits_final = logits;
  return NVLLM_OK;
}

// splitmix64 -> uniform in [0,1): the library's own stream when the caller passes no uniforms
static float next_uniform(nvllm_handle* h) {
  uint64_t z = (h->seed + 0x9E3779B97F4A7C15ull * (++h->rng_counter));
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  return (float)(z >> 40) * (1.0f / 16777216.0f);
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
  h->resident_ok = false;

  // ---- pass 1: sizes and validation
  int T = 0, max_blocks = 0, n_qtiles = 0, max_len = 0;
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
    }
    const int n_new = len - start;
    T += n_new;
    n_qtiles += (n_new + ATTN_PREFILL_QTILE - 1) / ATTN_PREFILL_QTILE;
    max_blocks = std::max(max_blocks, nb);
    max_len = std::max(max_len, len);
  }
  // ---- layout of the staging buffer
  StepLayout& L = h->lay;
  L.T = T; L.n = n; L.max_blocks = max_blocks; L.n_qtiles = is_prefill ? n_qtiles : 0;
  size_t o = 0;
  auto take = [&](size_t cnt) { size_t r = o; o += (cnt + 3) / 4 * 4; return r; };  // 16 B aligned
  L.tokens = take(T); L.positions = take(T); L.slots = take(T);
  L.cu_q = take(n + 1); L.seq_start = take(n); L.seq_len = take(n); L.last_row = take(n);
  L.block_tables = take((size_t)n * max_blocks);
  L.tile_seq = take(L.n_qtiles); L.tile_q0 = take(L.n_qtiles);
  L.temperature = take(n); L.uniforms = take(n);
  L.words = o;
  int rc;
  if ((rc = ensure_stage(h, L.words, n))) return rc;
  if ((rc = ensure_workspace(h, T, n))) return rc;
  if (!h->h_out || n > h->ws.n_cap_out) {
    if (h->h_out) cudaFreeHost(h->h_out);
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
      for (int q0 = 0; q0 < len - start; q0 += ATTN_PREFILL_QTILE, tile++) {
        s[L.tile_seq + tile] = i;
        s[L.tile_q0 + tile] = q0;
      }
    const float t = temperature ? temperature[i] : 0.f;
    s_temp[i] = t;
    s_uni[i] = (t > 0.f) ? (uniforms ? uniforms[i] : next_uniform(h)) : 0.f;
  }
  s[L.cu_q + n] = row;

  cudaStream_t st = h->stream;
  NV_CHECK(h, cudaMemcpyAsync(h->d_stage, h->h_stage, L.words * 4, cudaMemcpyHostToDevice, st));
  NV_CHECK(h, cudaEventRecord(h->ev0, st));
  h->launches = 0;
  if ((rc = forward_step(h, is_prefill != 0, max_len))) return rc;
  NV_CHECK(h, cudaEventRecord(h->ev1, st));
  NV_CHECK(h, cudaMemcpyAsync(h->h_out, h->ws.out_tokens, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
  uint64_t d2h = (uint64_t)n * 4;
  if (out_logits) {
    // rows are ldl apart on the device; copy them packed [n, V]
    NV_CHECK(h, cudaMemcpy2DAsync(out_logits, (size_t)h->V * 4, h->ws.logits_final,
                                  (size_t)h->ldl * 4, (size_t)h->V * 4, n, cudaMemcpyDeviceToHost,
                                  st));
    d2h += (uint64_t)n * h->V * 4;
  }
  cudaError_t e = cudaStreamSynchronize(st);
  if (e != cudaSuccess)
    return set_err(h, NVLLM_ERR_CUDA, "forward pass failed: %s", cudaGetErrorString(e));
  memcpy(out_tokens, h->h_out, (size_t)n * 4);
  h->resident_ok = !is_prefill;
  h->resident_max_len = max_len;
  if (stats) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, h->ev0, h->ev1);
    stats->device_ms = ms;
    stats->kernel_launches = h->launches;
    stats->new_tokens = T;
    stats->used_graph = 0;
    stats->h2d_bytes = L.words * 4;
    stats->d2h_bytes = d2h;
  }
  return NVLLM_OK;
}

// One more decode step for the sequences of the previous nvllm_run(decode) call, entirely on
// the device: the sampled tokens become the inputs, positions advance by one, KV slots come
// from the block tables already resident.  No host<->device copy, no host sync.
extern "C" int nvllm_run_resident(nvllm_handle* h, nvllm_run_stats* stats) {
  if (!h) return NVLLM_ERR_INVALID;
  if (!h->resident_ok)
    return set_err(h, NVLLM_ERR_STATE, "nvllm_run_resident needs a preceding decode nvllm_run");
  NV_CHECK(h, cudaSetDevice(h->dev));
  StepLayout& L = h->lay;
  const int n = L.n;
  if (h->resident_max_len + 1 > L.max_blocks * KV_BLOCK ||
      (h->use_rope && h->resident_max_len + 1 > h->cfg.max_seq_len))
    return set_err(h, NVLLM_ERR_RANGE, "resident decode ran past the block tables / max_seq_len");
  int32_t* d = h->d_stage;
  cudaStream_t st = h->stream;
  NV_CHECK(h, cudaEventRecord(h->ev0, st));
  h->launches = 0;
  LAUNCH(h, launch_advance(st, n, h->ws.out_tokens, d + L.tokens, d + L.positions, d + L.slots,
                           d + L.seq_len, d + L.block_tables, L.max_blocks,
                           L.max_blocks * KV_BLOCK));
  h->resident_max_len += 1;
  int rc = forward_step(h, false, h->resident_max_len);
  if (rc) return rc;
  NV_CHECK(h, cudaEventRecord(h->ev1, st));
  if (stats) {
    memset(stats, 0, sizeof *stats);
    stats->kernel_launches = h->launches;
    stats->new_tokens = n;
  }
  return NVLLM_OK;
}

// Block until the resident steps issued so far are done; copies the last sampled tokens out.
extern "C" int nvllm_sync(nvllm_handle* h, int32_t* out_tokens, int n, float* last_step_ms) {
  if (!h) return NVLLM_ERR_INVALID;
  NV_CHECK(h, cudaSetDevice(h->dev));
  cudaError_t e = cudaStreamSynchronize(h->stream);
  if (e != cudaSuccess)
    return set_err(h, NVLLM_ERR_CUDA, "stream failed: %s", cudaGetErrorString(e));
  if (out_tokens && n > 0) {
    if (n > h->lay.n) n = h->lay.n;
    NV_CHECK(h, cudaMemcpy(out_tokens, h->ws.out_tokens, (size_t)n * 4, cudaMemcpyDeviceToHost));
  }
  if (last_step_ms) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, h->ev0, h->ev1) != cudaSuccess) ms = 0.f;
    *last_step_ms = ms;
  }
  return NVLLM_OK;
}
