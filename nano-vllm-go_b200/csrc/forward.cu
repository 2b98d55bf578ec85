// forward.cu — the forward pass of one engine step as a sequence of sm_100a kernels on one
// stream (what TransformerModel.ForwardWithCache does per sequence in the reference,
// purego/tensor/generic_model.go:276-480, here for a ragged batch of T new tokens):
// embedding -> per layer [norm -> fused QKV GEMM -> RoPE + paged KV append -> paged
// attention -> O GEMM -> norm -> FFN / MoE] -> final norm of the last token of each sequence
// -> LM head on those rows only -> on-device argmax / temperature sampling.
// Residual adds are folded into the following norm kernel; GEMMs emit fp32 split-K partials
// that the consumer kernel sums.
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "common.cuh"
#include "handle.h"

namespace nvllm {


// ------------------------------------------------------------------ workspace
void free_workspace(nvllm_handle* h) {
  Workspace& w = h->ws;
  h->drop_graphs();  // the captured decode steps hold these pointers
  void* ptrs[] = {w.tp_buf, w.x, w.xn, w.xn_f32, w.ws_a, w.ws_b, w.ws_c, w.q, w.attn_out, w.hbuf, w.xn_last,
                  w.xb, w.ssq, w.chain_ctr, w.m_xbc, w.m_dt, w.m_y, w.m_yn, w.logits_p, w.logits, w.attn_ws, w.out_tokens, w.amax, w.amax_ticket, w.topk_idx, w.topk_w,
                  w.sort.counts, w.sort.offsets, w.sort.cursor, w.sort.sorted_pair, w.sort.pair_row,
                  w.sort.a_rows, w.sort.tile_expert, w.sort.tile_row0, w.sort.tile_rows,
                  w.sort.n_tiles, w.moe_Y, w.moe_y, w.moe_xs};
  for (void* p : ptrs)
    if (p) cudaFree(p);
  int keep = w.n_cap_out;
  w = Workspace();
  w.n_cap_out = keep;
  h->ws_bytes = 0;
}

int ensure_workspace(nvllm_handle* h, int T, int n) {
  Workspace& w = h->ws;
  if (T <= w.T_cap && n <= w.n_cap) return NVLLM_OK;
  NV_CHECK(h, cudaStreamSynchronize(h->stream));
  int Tc = std::max(T, w.T_cap), nc = std::max(n, w.n_cap);
  free_workspace(h);
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
      free_workspace(h);                                                                  \
      return set_err(h, NVLLM_ERR_NOMEM, "workspace for %d tokens: %s", Tc,               \
                     cudaGetErrorString(_e));                                             \
    }                                                                                     \
  }
  WS(w.x, Tc * H * 4);
  WS(w.xn, Tc * H * asz);
  size_t a_cols = std::max(h->qkv_cols, h->moe ? 0 : h->w1_cols);
  if (h->n_mamba) a_cols = std::max(a_cols, (size_t)h->m_P);
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
  WS(w.chain_ctr, 256);
  if (h->n_mamba) {
    WS(w.m_xbc, (size_t)Tc * h->m_conv * 4);
    WS(w.m_dt, (size_t)Tc * h->m_nh * 4);
    WS(w.m_y, (size_t)Tc * h->m_E * 4);
    WS(w.m_yn, (size_t)Tc * h->m_E * asz);
  }
  if (cudaMemsetAsync(w.chain_ctr, 0, 256, h->stream) != cudaSuccess)
    return set_err(h, NVLLM_ERR_CUDA, "memset(chain counters) failed");
  if (h->fold) {
    WS(w.xb, Tc * H * 2);
    WS(w.ssq, (size_t)ceil_div(h->H, 128) * Tc * 4);
  }
  if (h->tp > 1) WS(w.tp_buf, Tc * H * 4);
  // tp > 1: this rank's LM-head slice is Vl wide and the gathered logits are [tp][n][Vl]
  const size_t lcols = h->tp > 1 ? (size_t)h->Vl : (size_t)h->ldl;
  w.logits_p_floats = (size_t)std::max(nc, 64) * lcols;
  WS(w.logits_p, w.logits_p_floats * 4);
  WS(w.logits, (size_t)nc * lcols * h->tp * 4);
  w.attn_ws_floats = (size_t)nc * h->nh * 32 * (h->hd + 2);
  WS(w.attn_ws, w.attn_ws_floats * 4);
  WS(w.out_tokens, (size_t)nc * 4);
  WS(w.amax, (size_t)nc * 8);
  WS(w.amax_ticket, (size_t)nc * 4);
  if (cudaMemsetAsync(w.amax, 0, (size_t)nc * 8, h->stream) != cudaSuccess ||
      cudaMemsetAsync(w.amax_ticket, 0, (size_t)nc * 4, h->stream) != cudaSuccess)
    return set_err(h, NVLLM_ERR_CUDA, "memset(argmax scratch) failed");
  if (h->moe) {
    const size_t P = (size_t)Tc * h->topk;
    w.max_tiles = (int)(P / 32 + h->E);  // token tiles are >= 32 rows
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
    if (h->adt == DT_BF16) WS(w.moe_xs, P * H * 2);
    WS(w.moe_y, Tc * H * 4);
  }
#undef WS
  return NVLLM_OK;
}

// ------------------------------------------------------------------ profiling
static cudaEvent_t prof_event(nvllm_handle* h) {
  if (!h->prof_pool.empty()) {
    cudaEvent_t e = h->prof_pool.back();
    h->prof_pool.pop_back();
    return e;
  }
  cudaEvent_t e = nullptr;
  cudaEventCreate(&e);
  return e;
}
// Two instruments behind nvllm_set_profile:
//  mode 1  CUDA events around every launch.  Serialises the chain (no programmatic-dependent-
//          launch overlap, no graph replay): per-launch times are an upper bound.  The only mode
//          that sees the NCCL calls of a tensor-parallel step.
//  mode 2  device timeline: every kernel records min(entry) / min(dependency released) /
//          max(exit) of its CTAs in %globaltimer ns (Trace::agg); launches stay back to back on
//          the stream with their PDL edges, exactly as in a real step.  A launch is charged the
//          time between its predecessor's last exit and its own last exit, so the classes add up
//          to the device time of the profiled window.
// `bytes` are ALGORITHMIC bytes only (weights, K/V, logits, one pass over each activation the
// reference's dataflow has); `impl` are the extra bytes of this implementation (split-K
// partials, staged activations).
struct ProfScope {
  nvllm_handle* h;
  int idx = -1;
  ProfScope(nvllm_handle* hh, int cls, double bytes, double flops, double impl = 0.0) : h(hh) {
    h->cur_cls = cls;
    if (!h->profile) return;
    nvllm_handle::ProfRec r{cls, bytes, flops, nullptr, nullptr, impl, 0, 0};
    if (h->profile == 2) {
      r.seq0 = trace_seq();
    } else {
      r.e0 = prof_event(h);
      r.e1 = prof_event(h);
      cudaEventRecord(r.e0, h->stream);
    }
    idx = (int)h->prof_open.size();
    h->prof_open.push_back(r);
  }
  ~ProfScope() {
    if (idx < 0) return;
    if (h->profile == 2) h->prof_open[idx].seq1 = trace_seq();
    else cudaEventRecord(h->prof_open[idx].e1, h->stream);
  }
};
constexpr int PROF_AGG_CAP = 1 << 18;  // launches per profiled window (32 B each)
static int prof_agg_reset(nvllm_handle* h) {
  std::vector<unsigned long long> init((size_t)PROF_AGG_CAP * 4);
  for (size_t i = 0; i < init.size(); i += 4) {
    init[i] = init[i + 1] = ~0ull;
    init[i + 2] = init[i + 3] = 0ull;
  }
  NV_CHECK(h, cudaMemcpy(h->prof_agg, init.data(), init.size() * 8, cudaMemcpyHostToDevice));
  trace_agg_set(h->prof_agg, PROF_AGG_CAP);
  return NVLLM_OK;
}
int prof_collect(nvllm_handle* h) {
  if (h->prof_open.empty()) return NVLLM_OK;
  NV_CHECK(h, cudaStreamSynchronize(h->stream));
  if (h->profile == 2) {
    const int nseq = std::min(trace_seq(), PROF_AGG_CAP);
    std::vector<unsigned long long> agg((size_t)nseq * 4);
    if (nseq) NV_CHECK(h, cudaMemcpy(agg.data(), h->prof_agg, agg.size() * 8, cudaMemcpyDeviceToHost));
    unsigned long long prev_exit = 0;
    for (auto& r : h->prof_open) {
      unsigned long long first = ~0ull, exit = 0;
      for (int sq = r.seq0; sq < r.seq1 && sq < nseq; sq++) {
        const unsigned long long* a = &agg[(size_t)sq * 4];
        if (a[2] == 0) continue;  // the launch had no traced CTA
        first = std::min(first, a[1] != ~0ull ? a[1] : a[0]);
        exit = std::max(exit, a[2]);
      }
      h->prof[r.cls].bytes += r.bytes;
      h->prof[r.cls].impl_bytes += r.impl;
      h->prof[r.cls].flops += r.flops;
      h->prof[r.cls].launches += 1;
      if (exit == 0) continue;  // un-instrumented kernel: its time goes to the next record
      const unsigned long long start = prev_exit ? prev_exit : first;
      if (exit > start) h->prof[r.cls].ms += (double)(exit - start) * 1e-6;
      prev_exit = std::max(prev_exit, exit);
    }
    h->prof_open.clear();
    return prof_agg_reset(h);
  }
  for (auto& r : h->prof_open) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, r.e0, r.e1) == cudaSuccess) {
      h->prof[r.cls].ms += ms;
      h->prof[r.cls].bytes += r.bytes;
      h->prof[r.cls].impl_bytes += r.impl;
      h->prof[r.cls].flops += r.flops;
      h->prof[r.cls].launches += 1;
    }
    h->prof_pool.push_back(r.e0);
    h->prof_pool.push_back(r.e1);
  }
  h->prof_open.clear();
  return NVLLM_OK;
}

// ------------------------------------------------------------------ kernel glue
struct Pending {  // a GEMM result still to be added to the residual stream
  const float* p = nullptr;
  int S = 0;
  size_t stride = 0;
  float m = 1.f;
  const float* bias = nullptr;
  bool bf16 = false;  // p is one bf16 slice (all-reduced over the tensor-parallel ranks)
  // mixture-of-experts: p holds the expert outputs by sorted row; the addend of token r is
  // sum_s comb_w[r, s] * p[comb_rows[r, s], :]  (fused into the consuming norm kernel)
  const int* comb_rows = nullptr;
  const float* comb_w = nullptr;
  int comb_k = 0;
};

// Folded RMSNorm, consumer side: the GEMM reads the un-normalised bf16 residual stream and scales
// accumulator row m by 1 / sqrt(sum_p ssq[p][m] / H + eps)
struct RowScale {
  const float* ssq = nullptr;
  int parts = 0;
};
// Folded RMSNorm, producer side: the GEMM's epilogue performs the residual update
// x += m * (acc + bias) and leaves bf16(x) + the row sums of squares for the next consumer
struct Resid {
  float m = 1.f;
  const float* bias = nullptr;
  int site = 0;  // 0 = attention output projection, 1 = FFN down projection (counter sets)
  const float* g = nullptr;  // weight of the norm that reads x next (folded into the bf16 copy)
};

struct RouterArgs {  // expert routing fused into the FFN norm of a MoE layer
  const float* R = nullptr;
  int E = 0, k = 0;
  int* topk_idx = nullptr;
  float* topk_w = nullptr;
};

#define LAUNCH(h, expr)                                                                       \
  do {                                                                                        \
    if (((h)->skip_mask >> (h)->cur_cls) & 1) break; /* NVLLM_SKIP timing ablation */         \
    cudaError_t _e = (expr);                                                                  \
    (h)->launches++;                                                                          \
    if (_e != cudaSuccess)                                                                    \
      return set_err(h, NVLLM_ERR_CUDA, "%s: %s (%s:%d)", #expr, cudaGetErrorString(_e),      \
                     __FILE__, __LINE__);                                                     \
  } while (0)

// C = A @ W^T into fp32 split-K partials at `out`
// epi != 0 asks for the FFN activation to be fused into the GEMM epilogue (bf16 hidden state
// written to epi_out); granted only on the tcgen05 path when the GEMM needs no split-K, in
// which case res->S comes back 0 (no partials to reduce).
static int run_gemm(nvllm_handle* h, const void* A, int lda, const void* W, int M, int N, int K,
                    float alpha, float* out, size_t cap, int ldo, Pending* res, int epi = 0,
                    void* epi_out = nullptr, int epi_ld = 0, const float* epi_bias = nullptr,
                    const void* pf_ptr = nullptr, size_t pf_bytes = 0, double alg_out = 0.0,
                    const RowScale* rs = nullptr, const Resid* rd = nullptr,
                    const GemvPre* gp = nullptr) {
  res->p = out;
  res->stride = (size_t)M * ldo;
  const double flops = 2.0 * M * N * K;
  // algorithmic: the weights once (+ the logits the LM head must produce, alg_out);
  // implementation: the staged activations and the fp32 split-K partials
  const double io = (double)N * K * dt_size(h->wdt) + alg_out;
  const double a_in = (double)M * K * dt_size(h->adt) - alg_out;
  if (M <= 8) {
    GemmArgs g{};
    g.A = A; g.adt = h->adt; g.lda = lda; g.W = W; g.wdt = h->wdt;
    g.M = M; g.N = N; g.K = K; g.out = out; g.ldo = ldo; g.split_stride = res->stride;
    g.alpha = alpha; g.out_capacity = cap;
    // a fused pre-stage reads what the activation kernel would have read
    double pre_alg = 0.0, pre_impl = 0.0;
    if (gp && gp->mode == 2) {
      pre_alg = (double)M * gp->hld * 4 - (double)M * K * dt_size(h->adt);
      pre_impl = (double)M * gp->hld * 4 * (gp->Sh - 1);
    }
    ProfScope ps(h, NVLLM_K_GEMV, io + pre_alg, flops, a_in + pre_impl + 4.0 * gemv_splits(M, N, K) * M * N);
    LAUNCH(h, launch_gemv(h->stream, g, gp));
    res->S = g.S;
    return NVLLM_OK;
  }
  if (h->wdt == DT_BF16 && h->use_tc5 && lda == K && tc5_supported(M, N, K)) {
    Tc5Args t{};
    t.A = A; t.W = W; t.M = M; t.N = N; t.K = K; t.out = out; t.ldo = ldo;
    t.split_stride = res->stride; t.alpha = alpha; t.out_capacity = cap;
    t.S = tc5_splits(M, N, K, h->sm_count);
    t.max_ctas = h->gemm_max_ctas;
    if (rs && rs->ssq) {
      t.rs_ssq = rs->ssq; t.rs_parts = rs->parts; t.rs_ld = h->ws.T_cap;
      t.rs_inv_h = 1.0f / (float)h->H; t.rs_eps = h->cfg.norm_eps;
    }
    if (rd) {
      // x += m (A W^T + bias) inside the GEMM; nothing is left pending for a norm kernel
      // split-K shapes (decode batches): the k-splits of a tile are one thread-block cluster and
      // meet in shared memory — no partials in global memory either way
      t.S = tc5_resid_splits(M, N, K, h->sm_count);
      t.epi = TC5_EPI_RESID;
      t.res_x = h->ws.x; t.res_xb = h->ws.xb; t.res_ssq = h->ws.ssq; t.res_ssq_ld = h->ws.T_cap;
      t.res_bias = rd->bias; t.res_m = rd->m; t.res_g = rd->g;
      ProfScope ps(h, NVLLM_K_GEMM_TC5, io + (double)M * N * 4 * 2, flops, a_in + (double)M * N * 2);
      LAUNCH(h, launch_gemm_tc5(h->stream, t, nullptr));
      *res = Pending();
      return NVLLM_OK;
    }
    if (M <= 128) {  // weight-streaming regime: warm L2 with the next big matrix
      t.pf_ptr = pf_ptr;
      t.pf_bytes = pf_bytes;
    }
    const long tiles = (long)ceil_div(N, 128) * ceil_div(M, tc5_box_rows_act(M));
    if (epi && (t.S == 1 || 2 * tiles >= h->sm_count)) {
      t.S = 1;
      t.epi = epi; t.epi_out = epi_out; t.epi_ld = epi_ld; t.epi_bias = epi_bias;
      const double obytes = (double)M * epi_ld * 2;
      ProfScope ps(h, NVLLM_K_GEMM_TC5, io, flops, a_in + obytes);
      LAUNCH(h, launch_gemm_tc5(h->stream, t, nullptr));
      res->S = 0;
      return NVLLM_OK;
    }
    while (t.S > 1 && (size_t)t.S * M * ldo > cap) t.S--;
    ProfScope ps(h, NVLLM_K_GEMM_TC5, io, flops, a_in + 4.0 * t.S * M * N);
    LAUNCH(h, launch_gemm_tc5(h->stream, t, &res->S));
    return NVLLM_OK;
  }
  GemmArgs g{};
  g.A = A; g.adt = h->adt; g.lda = lda; g.W = W; g.wdt = h->wdt;
  g.M = M; g.N = N; g.K = K; g.out = out; g.ldo = ldo; g.split_stride = res->stride;
  g.alpha = alpha; g.out_capacity = cap;
  ProfScope ps(h, NVLLM_K_GEMM_SIMT, io, flops, a_in + 4.0 * M * N);
  LAUNCH(h, launch_gemm_simt(h->stream, g));
  res->S = g.S;
  return NVLLM_OK;
}

// can (M, N, K) with this activation stride take the tcgen05 path (and with it the folded-norm
// epilogues)?
static bool fold_gemm_ok(const nvllm_handle* h, int M, int N, int K, int lda) {
  return h->fold && M > 8 && h->wdt == DT_BF16 && h->use_tc5 && lda == K && tc5_supported(M, N, K);
}
static bool resid_gemm_ok(const nvllm_handle* h, int M, int N, int K, int lda) {
  if (!fold_gemm_ok(h, M, N, K, lda) || N != h->H) return false;
  const int S = tc5_resid_splits(M, N, K, h->sm_count);
  if (S <= 1) return S == 1 && tc5_splits(M, N, K, h->sm_count) == 1;
  // Split-K shapes (decode batches) can run the k-splits of an output tile as one thread-block
  // cluster that reduces through distributed shared memory and finishes the residual update
  // itself (gemm_tcgen05.cu) — no norm kernel behind the O and W2 GEMMs.  Parity-green, but
  // measured SLOWER on cfg2 decode (1150 -> 1195 us/step, DESIGN.md "rejected"): B200 places at
  // most 15 clusters of 8 such CTAs (16 tiles need a second wave) so the split drops to 6 = 96
  // CTAs streaming the weights instead of 128, and the in-cluster exchange + reduce costs more
  // than the norm launch it removes.  Opt-in: NVLLM_RESID_CLUSTER=1 (read per step).
  const char* ce = getenv("NVLLM_RESID_CLUSTER");
  return ce && atoi(ce) != 0;
}

static int run_norm(nvllm_handle* h, float* x, Pending* p1, Pending* p2, const float* w,
                    const float* b, void* y, float* y_f32, const int* row_idx, int n_rows,
                    const RouterArgs* route = nullptr) {
  NormArgs a{};
  a.x = x;
  // algorithmic: x in, y out, and (when a GEMM result is pending) that result once + x back
  double bytes = (double)n_rows * h->H * (4 + dt_size(h->adt)), impl = 0.0;
  if (p1 && p1->p) {
    a.p1 = p1->p; a.S1 = p1->S; a.stride1 = p1->stride; a.m1 = p1->m; a.bias1 = p1->bias;
    a.p1_bf16 = p1->bf16 ? 1 : 0;
    a.comb_rows = p1->comb_rows; a.comb_w = p1->comb_w; a.comb_k = p1->comb_k;
    bytes += (double)n_rows * h->H * 4 * 2;
    impl += (double)n_rows * h->H * 4 * ((p1->comb_rows ? p1->comb_k : p1->S) - 1);
  }
  if (route && route->R) {
    a.router = route->R; a.E = route->E; a.topk = route->k;
    a.topk_idx = route->topk_idx; a.topk_w = route->topk_w;
    bytes += (double)route->E * h->H * 4;
  }
  if (p2 && p2->p && p2->comb_rows)
    return set_err(h, NVLLM_ERR_INVALID, "MoE layers in a parallel block are not supported");
  if (p2 && p2->p) {
    a.p2 = p2->p; a.S2 = p2->S; a.stride2 = p2->stride; a.m2 = p2->m; a.bias2 = p2->bias;
    bytes += (double)n_rows * h->H * 4;
    impl += (double)n_rows * h->H * 4 * (p2->S - 1);
  }
  a.w = w; a.b = b; a.eps = h->cfg.norm_eps; a.y = y; a.ydt = h->adt; a.y_f32 = y_f32;
  a.row_idx = row_idx; a.n_rows = n_rows; a.H = h->H;
  a.row_bias = row_idx ? h->view_r0 : 0;  // row_idx holds rows of the whole step
  ProfScope ps(h, NVLLM_K_NORM, bytes, 0, impl);
  LAUNCH(h, launch_norm(h->stream, a));
  if (p1) *p1 = Pending();
  if (p2) *p2 = Pending();
  return NVLLM_OK;
}

// Tensor parallel: the pending row-parallel GEMM results are this rank's partial sums.  Fold
// them (split-K slices, multipliers, rank-0 biases) into one [T, H] buffer, sum it over the
// ranks and hand it to the norm as a single finished addend.
static int tp_reduce_pending(nvllm_handle* h, Pending* p1, Pending* p2, int T) {
  if (h->tp <= 1) return NVLLM_OK;
  Pending* a = (p1 && p1->p) ? p1 : nullptr;
  Pending* b = (p2 && p2->p) ? p2 : nullptr;
  if (!a) { a = b; b = nullptr; }
  if (!a) return NVLLM_OK;
  Workspace& w = h->ws;
  // large (prefill) messages travel as bf16 in bf16 mode: half the bytes of the exposed
  // all-reduce; the sum over ranks of the per-rank addends is rounded once more, the residual
  // stream itself stays fp32.  NVLLM_TP_BF16=0 keeps fp32 on the wire.
  static const bool wire_bf16_ok = !(getenv("NVLLM_TP_BF16") && atoi(getenv("NVLLM_TP_BF16")) == 0);
  // the peer-memory kernel keeps one flag / epoch state per handle: the two chunks of a pipelined
  // prefill must not use it concurrently
  const bool p2p = h->pipe_role == 0 && tp_p2p_fits(h, (size_t)T * h->H);
  const bool wire_bf16 = !p2p && wire_bf16_ok && h->adt == DT_BF16 && ((size_t)T * h->H % 8) == 0;
  {
    const double n = (double)T * h->H * 4;
    ProfScope ps(h, NVLLM_K_COMM, (wire_bf16 ? 0.5 : 1.0) * n, 0, n * (a->S + (b ? b->S : 0) + 1));  // bytes on the wire per rank
    if (p2p) {
      // one kernel: partial sums + all-reduce over NVLink peer memory (tp.cu)
      LAUNCH(h, launch_tp_allreduce_p2p(h, T, h->H, a->p, a->S, a->stride, a->m, a->bias,
                                        b ? b->p : nullptr, b ? b->S : 0, b ? b->stride : 0,
                                        b ? b->m : 0.f, b ? b->bias : nullptr, w.tp_buf));
    } else {
      LAUNCH(h, launch_tp_partial(h->stream, T, h->H, a->p, a->S, a->stride, a->m, a->bias,
                                  b ? b->p : nullptr, b ? b->S : 0, b ? b->stride : 0,
                                  b ? b->m : 0.f, b ? b->bias : nullptr, w.tp_buf, wire_bf16));
      int rc = tp_all_reduce(h, w.tp_buf, (size_t)T * h->H, wire_bf16);
      if (rc) return rc;
      h->launches++;
    }
  }
  if (p1) *p1 = Pending();
  if (p2) *p2 = Pending();
  Pending& o = p1 ? *p1 : *p2;
  o.p = w.tp_buf; o.S = 1; o.stride = 0; o.m = 1.f; o.bias = nullptr;
  o.bf16 = wire_bf16;
  return NVLLM_OK;
}

// ------------------------------------------------------------------ forward pass
// Inputs are the device arrays of h->lay (already on the device).
int forward_step(nvllm_handle* h) {
  const StepLayout& L = h->lay;
  Workspace& w = h->ws;
  const bool is_prefill = L.prefill;
  const int T = L.T, n = L.n, H = h->H;
  const int32_t* d = L.d;
  const int* d_tokens = d + L.tokens;
  const int* d_pos = d + L.positions;
  const int* d_slots = d + L.slots;
  const nvllm_config& c = h->cfg;
  cudaStream_t st = h->stream;
  const double asz = dt_size(h->adt), kvsz = dt_size(h->kvdt);
  int rc;

  {
    ProfScope ps(h, NVLLM_K_EMBED, (double)T * H * (4 + dt_size(h->wdt)), 0);
    LAUNCH(h, launch_embed(st, h->wdt, d_tokens, d_pos, h->wte, h->wpe, c.max_seq_len,
                           c.embedding_multiplier, w.x, T, H, h->fold ? w.xb : nullptr,
                           h->fold ? w.ssq : nullptr, h->fold ? h->layers[0].attn_norm_w : nullptr));
  }
  // Folded RMSNorm (h->fold): while `xb_parts` > 0, w.xb holds bf16(x) of the CURRENT residual
  // stream and w.ssq its row sums of squares in xb_parts column-tile slices — the next GEMM
  // that would read a normed activation reads w.xb instead and the norm kernel is not launched.
  // read per step (not cached): the parity tests switch it between runs
  const char* chain_env = getenv("NVLLM_CHAIN");
  const bool chain_on = !(chain_env && atoi(chain_env) == 0);
  // Steps of <= 8 rows (GEMV shapes): the FFN activation is built inside the W2 GEMV — one
  // launch per layer less on a chain of ~100 dependent launches.  NVLLM_GEMV_PRE=0 keeps the
  // stand-alone activation kernel (read per step: the parity tests switch it).
  const char* gpre_env = getenv("NVLLM_GEMV_PRE");
  const bool gpre_act = !(gpre_env && atoi(gpre_env) == 0) && T <= 8 && !h->moe;
  int xb_parts = h->fold ? 1 : 0;
  const int h_tiles = ceil_div(H, 128);
  // generic_model.go:359-365: res + rm * x when ResidualMultiplier != 0, else plain add
  const float m_res = (c.residual_multiplier != 0.f) ? c.residual_multiplier : 1.f;
  // generic_model.go:405-416 (parallel block): both multipliers or neither
  const bool par_scaled = c.attention_multiplier != 0.f && c.residual_multiplier != 0.f;
  Pending p1, p2;
  const size_t kv_layer_bytes = h->kv_layer_elems * dt_size(h->kvdt);

  AttnArgs at{};
  // the attention kernels index rows of the WHOLE step (cu_q + q0): a chunk's view of the
  // workspace starts view_r0 rows in
  at.q = (char*)w.q - (size_t)h->view_r0 * h->q_cols * dt_size(h->adt);
  at.out = (char*)w.attn_out - (size_t)h->view_r0 * h->q_cols * dt_size(h->adt);
  at.qdt = h->adt; at.kvdt = h->kvdt;
  at.n_seqs = n; at.nh = h->nh; at.nkv = h->nkv; at.hd = h->hd; at.scale = h->attn_scale;
  at.block_tables = d + L.block_tables; at.max_blocks = L.max_blocks;
  at.cu_q = d + L.cu_q; at.seq_start = d + L.seq_start;
  at.tile_seq = d + L.tile_seq; at.tile_q0 = d + L.tile_q0; at.n_qtiles = L.n_qtiles;
  at.seq_len = d + L.seq_len;
  at.ws = w.attn_ws;
  at.pool_blocks = h->num_blocks;
  at.total_rows = h->view_T > 0 ? h->view_T : T;
  at.sm_count = h->sm_count;
  if (!is_prefill)
    at.n_splits = attn_decode_splits(h->adt, h->kvdt, n, h->nh, h->nkv, h->hd, L.max_len, h->sm_count);
  const double qo_bytes = 2.0 * T * h->q_cols * asz;
  const double attn_flops = 4.0 * h->nh * h->hd * L.attn_pairs;
  // algorithmic: every K/V row of the contexts once (+ the new rows appended by a fused decode
  // prologue); implementation: Q / O staging (and the QKV partials a fused prologue sums)
  const double attn_bytes = 2.0 * L.sum_len * h->nkv * h->hd * kvsz;

  for (int l = 0; l < h->L; l++) {
    const LayerW& lw = h->layers[l];
    if (!h->is_mamba.empty() && h->is_mamba[l]) {
      // ---- Mamba2 sub-block (generic_model.go:162-181, mamba2.go:72-207): norm -> in_proj GEMM
      // -> conv + SiLU + softplus -> selective scan -> gate + RMS norm -> out_proj GEMM, whose
      // result is added to the residual stream by the next norm like an attention output
      if ((rc = run_norm(h, w.x, &p1, &p2, lw.attn_norm_w, lw.attn_norm_b, w.xn, nullptr, nullptr, T)))
        return rc;
      Pending pr;
      if ((rc = run_gemm(h, w.xn, H, lw.m_in_proj, T, h->m_P, H, 1.f, w.ws_a, w.ws_a_floats, h->m_P, &pr)))
        return rc;
      MambaArgs ma{};
      ma.T = T; ma.n_seqs = n;
      ma.nh = h->m_nh; ma.hd = h->m_hd; ma.N = h->m_N; ma.G = h->m_G; ma.K = h->m_K; ma.E = h->m_E;
      ma.conv = h->m_conv; ma.P = h->m_P;
      ma.proj = pr.p; ma.S = pr.S; ma.stride = pr.stride;
      ma.conv_w = lw.m_conv_w; ma.conv_b = lw.m_conv_b; ma.dt_bias = lw.m_dt_bias; ma.a_log = lw.m_a_log;
      ma.d_skip = lw.m_d; ma.norm_w = lw.m_norm;
      ma.xbc = w.m_xbc; ma.dt = w.m_dt; ma.y = w.m_y;
      ma.cu_q = d + L.cu_q; ma.positions = d_pos; ma.block_tables = d + L.block_tables;
      ma.max_blocks = L.max_blocks;
      ma.ssm_pool = h->ssm_pool; ma.n_mamba = h->n_mamba; ma.layer = h->mamba_idx[l];
      {
        const double state_bytes = 2.0 * n * h->m_nh * h->m_hd * h->m_N * 4;  // SSM state in + out
        ProfScope ps(h, NVLLM_K_MISC, (double)T * (h->m_P * 4.0 + h->m_E * asz) + state_bytes, 0,
                     (double)T * (h->m_conv + h->m_nh + 2.0 * h->m_E) * 4 + (double)T * h->m_P * 4 * (pr.S - 1));
        LAUNCH(h, launch_mamba_conv(st, ma));
        LAUNCH(h, launch_mamba_scan(st, ma));
        LAUNCH(h, launch_mamba_gate_norm(st, ma, h->adt, w.m_yn));
      }
      if ((rc = run_gemm(h, w.m_yn, h->m_E, lw.m_out_proj, T, H, h->m_E, 1.f, w.ws_b, w.ws_b_floats, H, &p1)))
        return rc;
      p1.bias = nullptr;
      p1.m = m_res;
    } else {
    // ---- attention sub-block
    if ((rc = tp_reduce_pending(h, &p1, &p2, T))) return rc;
    Pending qkv;
    if (xb_parts > 0 && fold_gemm_ok(h, T, h->qkv_cols, H, H)) {
      RowScale rsc{w.ssq, xb_parts};
      if ((rc = run_gemm(h, w.xb, H, lw.wqkv, T, h->qkv_cols, H, 1.f, w.ws_a, w.ws_a_floats,
                         h->qkv_cols, &qkv, 0, nullptr, 0, nullptr, nullptr, 0, 0.0, &rsc)))
        return rc;
    } else {
      if ((rc = run_norm(h, w.x, &p1, &p2, lw.attn_norm_w, lw.attn_norm_b, w.xn,
                         nullptr, nullptr, T)))
        return rc;
      if ((rc = run_gemm(h, w.xn, H, lw.wqkv, T, h->qkv_cols, H, 1.f, w.ws_a, w.ws_a_floats,
                         h->qkv_cols, &qkv)))
        return rc;
    }
    xb_parts = 0;
    void* kp = (char*)h->kpool + (size_t)l * kv_layer_bytes;
    void* vp = (char*)h->vpool + (size_t)l * kv_layer_bytes;
    const bool fuse_rope = !is_prefill && attn_decode_fuses_rope(h->adt, h->kvdt, h->hd);
    if (fuse_rope) {
      at.qkv_p = qkv.p; at.qkv_S = qkv.S; at.qkv_stride = qkv.stride; at.qkv_ld = h->qkv_cols;
      at.qkv_bias = lw.bqkv;
      at.cosT = h->use_rope ? h->cosT : nullptr;
      at.sinT = h->use_rope ? h->sinT : nullptr;
    } else {
      ProfScope ps(h, NVLLM_K_ROPE_KV,
                   (double)T * h->qkv_cols * 4 + (double)T * h->q_cols * asz +
                       2.0 * T * h->kv_cols * kvsz, 0, (double)T * h->qkv_cols * 4 * (qkv.S - 1));
      LAUNCH(h, launch_rope_kv(st, h->adt, h->kvdt, T, h->nh, h->nkv, h->hd, qkv.p, qkv.S,
                               qkv.stride, lw.bqkv, d_pos, d_slots,
                               h->use_rope ? h->cosT : nullptr, h->use_rope ? h->sinT : nullptr, w.q,
                               kp, vp));
      if (h->pipe_role == 1) NV_CHECK(h, cudaEventRecord(h->pipe_kv[l], st));
    }
    at.kpool = kp;
    at.vpool = vp;
    if (is_prefill) {
      // second chunk: its queries attend the first chunk's keys of this layer
      if (h->pipe_role == 2) NV_CHECK(h, cudaStreamWaitEvent(st, h->pipe_kv[l], 0));
      ProfScope ps(h, NVLLM_K_ATTN_PREFILL, attn_bytes, attn_flops, qo_bytes);
      LAUNCH(h, launch_attn_prefill(st, at));
    } else {
      ProfScope ps(h, NVLLM_K_ATTN_DECODE,
                   attn_bytes + (fuse_rope ? 2.0 * T * h->kv_cols * kvsz : 0.0), attn_flops,
                   fuse_rope ? (double)T * h->qkv_cols * 4 * qkv.S + (double)T * h->q_cols * asz
                             : qo_bytes);
      LAUNCH(h, launch_attn_decode(st, at));
      if (at.n_splits > 1) h->launches++;
    }
    if (resid_gemm_ok(h, T, H, h->q_cols, h->q_cols)) {
      Resid rd{m_res, lw.bo, 0, lw.ffn_norm_w};
      if ((rc = run_gemm(h, w.attn_out, h->q_cols, lw.wo, T, H, h->q_cols, 1.f, w.ws_b,
                         w.ws_b_floats, H, &p1, 0, nullptr, 0, nullptr, nullptr, 0, 0.0, nullptr, &rd)))
        return rc;
      xb_parts = h_tiles;
    } else {
      if ((rc = run_gemm(h, w.attn_out, h->q_cols, lw.wo, T, H, h->q_cols, 1.f, w.ws_b,
                         w.ws_b_floats, H, &p1)))
        return rc;
      p1.bias = lw.bo;
      p1.m = h->parallel ? (par_scaled ? c.attention_multiplier : 1.f) : m_res;
    }
    }  // attention / Mamba2 sub-block
    // ---- feed-forward sub-block
    float* ffn_out = w.ws_b;
    if (h->parallel) {
      ffn_out = w.ws_c;  // FFN reads the same normed input (generic_model.go:395-418)
    } else {
      if ((rc = tp_reduce_pending(h, &p1, nullptr, T))) return rc;
      RouterArgs ra;
      if (h->moe) {
        ra.R = lw.router; ra.E = h->E; ra.k = h->topk;
        ra.topk_idx = w.topk_idx; ra.topk_w = w.topk_w;
      }
      const bool skip_norm = xb_parts > 0 && !h->moe && fold_gemm_ok(h, T, h->w1_cols, H, H);
      if (!skip_norm) {
        if ((rc = run_norm(h, w.x, &p1, nullptr, lw.ffn_norm_w, lw.ffn_norm_b,
                           w.xn, nullptr, nullptr, T, &ra)))
          return rc;
        xb_parts = 0;
      }
    }
    Pending& pf = h->parallel ? p2 : p1;
    if (h->moe) {
      const int k = h->topk, I = h->F, P = T * k;
      // bf16 mode: grouped tcgen05 GEMMs over the expert-sorted rows; fp32 parity mode (and
      // shapes TMA cannot address): the SIMT fp32 tiles
      const bool tc = h->adt == DT_BF16 && h->wdt == DT_BF16 && tc5_supported(P, 2 * I, H) &&
                      tc5_supported(P, H, I) && (I % 8) == 0;
      const int tile = tc ? tc5_group_tile((P + h->E - 1) / h->E) : 64;
      // the routing itself (softmax, top-k) ran inside the FFN norm kernel above
      const bool sg = tc && moe_sort_gather_fits(T, k, h->E, H);
      {
        ProfScope ps(h, NVLLM_K_MOE_ROUTE, (double)P * 4 * 4 + (sg ? (double)P * H * 2 * 2 : 0.0), 0);
        if (sg)
          LAUNCH(h, launch_moe_sort_gather(st, w.topk_idx, T, k, h->E, w.sort, w.max_tiles, tile, H,
                                           w.xn, w.moe_xs));
        else
          LAUNCH(h, launch_moe_sort(st, w.topk_idx, T, k, h->E, w.sort, w.max_tiles, tile));
      }
      const double wsz = dt_size(h->wdt);
      const int e_touched = std::min(h->E, P);
      if (tc) {
        if (!sg) {
          ProfScope ps(h, NVLLM_K_MOE_MISC, (double)P * H * 2 * 2, 0);
          LAUNCH(h, launch_moe_gather(st, P, H, w.xn, w.sort.a_rows, w.moe_xs));
        }
        Tc5Args t{};
        t.A = w.moe_xs; t.W = lw.moe_in; t.M = P; t.N = 2 * I; t.K = H;
        t.S = 1; t.alpha = 1.f; t.epi = TC5_EPI_SWIGLU; t.epi_out = w.hbuf; t.epi_ld = I;
        t.g_experts = h->E; t.g_tile = tile; t.g_max_tiles = w.max_tiles;
        t.g_tile_expert = w.sort.tile_expert; t.g_tile_row0 = w.sort.tile_row0;
        t.g_tile_rows = w.sort.tile_rows; t.g_n_tiles = w.sort.n_tiles;
        {
          ProfScope ps(h, NVLLM_K_MOE_GEMM, (double)e_touched * 2 * I * H * wsz, 2.0 * P * 2 * I * H,
                       (double)P * H * asz + (double)P * I * asz);
          LAUNCH(h, launch_gemm_tc5(st, t, nullptr));
        }
        Tc5Args t2 = t;
        t2.A = w.hbuf; t2.W = lw.moe_out; t2.N = H; t2.K = I;
        t2.epi = TC5_EPI_PARTIAL; t2.epi_out = nullptr; t2.out = w.moe_Y; t2.ldo = H;
        t2.split_stride = 0; t2.out_capacity = 0;
        {
          ProfScope ps(h, NVLLM_K_MOE_GEMM, (double)e_touched * H * I * wsz, 2.0 * P * H * I,
                       (double)P * I * asz + (double)P * H * 4);
          LAUNCH(h, launch_gemm_tc5(st, t2, nullptr));
        }
      } else {
        GemmArgs g{};
        g.A = w.xn; g.adt = h->adt; g.lda = H; g.W = lw.moe_in; g.wdt = h->wdt;
        g.M = P; g.N = 2 * I; g.K = H; g.out = w.ws_a; g.ldo = 2 * I; g.alpha = 1.f;
        g.a_rows = w.sort.a_rows; g.tile_expert = w.sort.tile_expert; g.tile_row0 = w.sort.tile_row0;
        g.tile_rows = w.sort.tile_rows; g.n_tiles = w.sort.n_tiles;
        g.expert_stride = (size_t)2 * I * H; g.max_tiles = w.max_tiles;
        {
          ProfScope ps(h, NVLLM_K_MOE_GEMM,
                       (double)e_touched * 2 * I * H * wsz + (double)P * H * asz + (double)P * 2 * I * 4,
                       2.0 * P * 2 * I * H);
          LAUNCH(h, launch_gemm_simt(st, g));
        }
        {
          ProfScope ps(h, NVLLM_K_MOE_MISC, (double)P * I * (8 + asz), 0);
          LAUNCH(h, launch_moe_glu(st, h->adt, P, I, w.ws_a, w.hbuf));
        }
        GemmArgs g2{};
        g2.A = w.hbuf; g2.adt = h->adt; g2.lda = I; g2.W = lw.moe_out; g2.wdt = h->wdt;
        g2.M = P; g2.N = H; g2.K = I; g2.out = w.moe_Y; g2.ldo = H; g2.alpha = 1.f;
        g2.tile_expert = w.sort.tile_expert; g2.tile_row0 = w.sort.tile_row0;
        g2.tile_rows = w.sort.tile_rows; g2.n_tiles = w.sort.n_tiles;
        g2.expert_stride = (size_t)H * I; g2.max_tiles = w.max_tiles;
        {
          ProfScope ps(h, NVLLM_K_MOE_GEMM,
                       (double)e_touched * H * I * wsz + (double)P * I * asz + (double)P * H * 4,
                       2.0 * P * H * I);
          LAUNCH(h, launch_gemm_simt(st, g2));
        }
      }
      // the weighted combine (moe.go:116-118) happens inside the norm kernel that consumes pf
      pf.p = w.moe_Y; pf.S = 1; pf.stride = 0; pf.bias = nullptr;
      pf.comb_rows = w.sort.pair_row; pf.comb_w = w.topk_w; pf.comb_k = k;
    } else {
      // decode-sized steps: W1 (+ fused activation) and W2 in ONE launch — W2's weight stream
      // starts under W1's epilogues and the launch / drain / release gap between the two
      // GEMMs is gone (gemm_tcgen05.cu, Tc5Chain).  NVLLM_CHAIN=0 keeps two launches.
      const long w1_tiles = ceil_div(h->w1_cols, 128);
      if (chain_on && T > 8 && T <= 128 && xb_parts == 0 && h->wdt == DT_BF16 && h->use_tc5 &&
          tc5_supported(T, h->w1_cols, H) && tc5_supported(T, H, h->F) &&
          (tc5_splits(T, h->w1_cols, H, h->sm_count) == 1 || 2 * w1_tiles >= h->sm_count)) {
        Tc5Args t1{};
        t1.A = w.xn; t1.W = lw.w1; t1.M = T; t1.N = h->w1_cols; t1.K = H; t1.S = 1; t1.alpha = 1.f;
        t1.epi = h->swiglu ? TC5_EPI_SWIGLU : TC5_EPI_GELU; t1.epi_out = w.hbuf; t1.epi_ld = h->F;
        t1.epi_bias = lw.b1;
        Tc5Args t2{};
        t2.A = w.hbuf; t2.W = lw.w2; t2.M = T; t2.N = H; t2.K = h->F; t2.alpha = 1.f;
        t2.out = ffn_out; t2.ldo = H; t2.split_stride = (size_t)T * H; t2.out_capacity = w.ws_b_floats;
        t2.S = tc5_splits(T, H, h->F, h->sm_count);
        while (t2.S > 1 && (size_t)t2.S * T * H > w.ws_b_floats) t2.S--;
        int S2 = 1;
        {
          const double wsz = dt_size(h->wdt);
          ProfScope ps(h, NVLLM_K_GEMM_TC5, ((double)h->w1_cols * H + (double)H * h->F) * wsz,
                       2.0 * T * ((double)h->w1_cols * H + (double)H * h->F),
                       (double)T * H * asz + 2.0 * T * h->F * asz + 4.0 * t2.S * T * H);
          LAUNCH(h, launch_gemm_tc5_chain(st, t1, t2, w.chain_ctr, &S2));
        }
        pf.p = ffn_out; pf.S = S2; pf.stride = (size_t)T * H; pf.bias = lw.b2;
        pf.m = h->parallel ? (par_scaled ? c.residual_multiplier : 1.f) : m_res;
        continue;
      }
      Pending h1;
      RowScale rsc{xb_parts > 0 ? w.ssq : nullptr, xb_parts};
      if ((rc = run_gemm(h, xb_parts > 0 ? w.xb : w.xn, H, lw.w1, T, h->w1_cols, H, 1.f, w.ws_a,
                         w.ws_a_floats, h->w1_cols, &h1, h->swiglu ? TC5_EPI_SWIGLU : TC5_EPI_GELU,
                         w.hbuf, h->F, lw.b1, nullptr, 0, 0.0, &rsc)))
        return rc;
      xb_parts = 0;
      if (h1.S > 0 && gpre_act) {
        GemvPre ga{};
        ga.mode = 2; ga.hp = h1.p; ga.Sh = h1.S; ga.hstride = h1.stride; ga.hld = h->w1_cols;
        ga.hbias = lw.b1; ga.swiglu = h->swiglu ? 1 : 0;
        if ((rc = run_gemm(h, w.hbuf, h->F, lw.w2, T, H, h->F, 1.f, ffn_out, w.ws_b_floats, H, &pf, 0,
                           nullptr, 0, nullptr, nullptr, 0, 0.0, nullptr, nullptr, &ga)))
          return rc;
        pf.bias = lw.b2;
        pf.m = h->parallel ? (par_scaled ? c.residual_multiplier : 1.f) : m_res;
        continue;
      }
      if (h1.S > 0) {
        ProfScope ps(h, NVLLM_K_ACT, (double)T * h->w1_cols * 4 + (double)T * h->F * asz, 0,
                     (double)T * h->w1_cols * 4 * (h1.S - 1));
        LAUNCH(h, launch_act(st, h->adt, T, h->F, h1.p, h1.S, h1.stride, lw.b1, h->swiglu ? 1 : 0,
                             w.hbuf));
      }
      if (!h->parallel && resid_gemm_ok(h, T, H, h->F, h->F)) {
        // the next reader of x is layer l + 1's attention norm (the final norm reads fp32 x)
        Resid rd{m_res, lw.b2, 1, l + 1 < h->L ? h->layers[l + 1].attn_norm_w : nullptr};
        if ((rc = run_gemm(h, w.hbuf, h->F, lw.w2, T, H, h->F, 1.f, ffn_out, w.ws_b_floats, H, &pf, 0,
                           nullptr, 0, nullptr, nullptr, 0, 0.0, nullptr, &rd)))
          return rc;
        xb_parts = h_tiles;
      } else {
        if ((rc = run_gemm(h, w.hbuf, h->F, lw.w2, T, H, h->F, 1.f, ffn_out, w.ws_b_floats, H, &pf)))
          return rc;
        pf.bias = lw.b2;
      }
    }
    if (pf.p) pf.m = h->parallel ? (par_scaled ? c.residual_multiplier : 1.f) : m_res;
  }
  if (h->skip_head) return NVLLM_OK;  // no sequence ends in this chunk: its last residual update is never read
  // final norm on the last token of each sequence only (the reference norms and projects all
  // T rows, generic_model.go:464-470, then keeps row T-1, :595-604)
  if ((rc = tp_reduce_pending(h, &p1, &p2, T))) return rc;
  if ((rc = run_norm(h, w.x, &p1, &p2, h->lnf_w, h->lnf_b, w.xn_last, nullptr, d + L.last_row, n)))
    return rc;
  const float alpha = (c.logits_scaling != 0.f) ? 1.0f / c.logits_scaling : 1.f;
  Pending lg;
  const float* logits = nullptr;
  int vsplit = 0;
  if (h->tp > 1) {
    // vocabulary-parallel LM head: this rank's Vl_valid rows -> slot `rank` of [tp][n][Vl],
    // then one all-gather; every rank samples the same token from the same gathered logits
    const int Vl = h->Vl;
    float* slot = w.logits + (size_t)h->rank * n * Vl;
    if (h->Vl_valid > 0) {
      if ((rc = run_gemm(h, w.xn_last, H, h->lm_head, n, h->Vl_valid, H, alpha, w.logits_p,
                         w.logits_p_floats, Vl, &lg)))
        return rc;
      ProfScope ps(h, NVLLM_K_MISC, (double)n * Vl * 4 * (lg.S + 1), 0);
      LAUNCH(h, launch_reduce_partials(st, lg.p, lg.S, lg.stride, (size_t)n * Vl, slot));
    }
    {
      ProfScope ps(h, NVLLM_K_COMM, (double)n * Vl * 4 * h->tp, 0);
      if ((rc = tp_all_gather_f32(h, w.logits, (size_t)n * Vl))) return rc;
      h->launches++;
    }
    logits = w.logits;
    vsplit = Vl;
  } else {
    if ((rc = run_gemm(h, w.xn_last, H, h->lm_head, n, h->V, H, alpha, w.logits_p,
                       w.logits_p_floats, h->ldl, &lg, 0, nullptr, 0, nullptr, nullptr, 0,
                       (double)n * h->V * 4)))
      return rc;
    logits = lg.p;
    if (lg.S > 1) {
      ProfScope ps(h, NVLLM_K_MISC, 0, 0, (double)n * h->ldl * 4 * (lg.S + 1));
      LAUNCH(h, launch_reduce_partials(st, lg.p, lg.S, lg.stride, (size_t)n * h->ldl, w.logits));
      logits = w.logits;
    }
  }
  const float* d_temp = reinterpret_cast<const float*>(d + L.temperature);
  const float* d_uni = reinterpret_cast<const float*>(d + L.uniforms);
  {
    ProfScope ps(h, NVLLM_K_SAMPLE, (double)n * h->V * 4, 0);
    LAUNCH(h, launch_sample(st, logits, h->ldl, h->V, n, d_temp, d_uni, w.out_tokens, w.amax,
                            w.amax_ticket, vsplit));
    if (h->filters_active() && h->hist_n == n) {
      // TensorModelRunner's sampler (tensor_model_runner.go:92-94): the rows with
      // temperature > 0 are redrawn with the repetition penalty / top-k / top-p applied
      LAUNCH(h, launch_sample_filtered(st, logits, h->ldl, h->V, n, vsplit, d_temp, d_uni,
                                       h->d_hist + (n + 1), h->d_hist, h->top_p, h->top_k,
                                       h->rep_penalty, h->d_probs, w.out_tokens));
    }
  }
  h->ws.logits_final = logits;
  return NVLLM_OK;
}

// Tensor-parallel prefill of ONE long sequence as two token chunks in flight at once: chunk A
// (rows [0, r0)) on the handle's stream and communicator, chunk B (rows [r0, T)) on stream2 with
// the duplicate communicator.  Each chunk is a complete forward pass over its rows — causal
// attention only needs the earlier chunk's K / V of the same layer, handed over by one event
// per layer — so B's GEMMs run while A's per-layer all-reduce is on the wire and vice versa
// (SURVEY.md 8e / hard part 3: "token-chunk pipelining").  The persistent GEMM grids leave 16
// SMs to the collective kernels.  Only B holds the sequence's last token: A skips the head.
// Opt-in: NVLLM_TP_PIPE=1 (read at tp_init; measured slower at tp 2, see tp.cu), threshold
// NVLLM_TP_PIPE_MIN_T.
int forward_step_auto(nvllm_handle* h) {
  const StepLayout L = h->lay;
  static const int min_T = getenv("NVLLM_TP_PIPE_MIN_T") ? atoi(getenv("NVLLM_TP_PIPE_MIN_T")) : 1024;
  const int qtile = attn_prefill_qtile(h->adt, h->kvdt, h->hd);
  const int ts = L.n_qtiles / 2, r0 = ts * qtile;
  if (!(h->tp > 1 && h->nccl2 && L.prefill && L.n == 1 && !h->profile && !h->moe && L.T >= min_T &&
        ts >= 1 && r0 < L.T && (r0 % 8) == 0))
    return forward_step(h);
  const Workspace W0 = h->ws;
  cudaStream_t s0 = h->stream;
  static bool said = false;
  if (!said && getenv("NVLLM_TP_DEBUG")) {
    said = true;
    fprintf(stderr, "[nvllm] rank %d/%d: two-chunk prefill pipeline, rows [0, %d) + [%d, %d)\n", h->rank,
            h->tp, r0, r0, L.T);
  }
  NV_CHECK(h, cudaEventRecord(h->pipe_fork, s0));
  NV_CHECK(h, cudaStreamWaitEvent(h->stream2, h->pipe_fork, 0));
  const size_t asz = dt_size(h->adt);
  int rc = NVLLM_OK;
  for (int role = 1; role <= 2 && rc == NVLLM_OK; role++) {
    const int rb = role == 1 ? 0 : r0, Tc = role == 1 ? r0 : L.T - r0;
    Workspace v = W0;
    v.x = W0.x + (size_t)rb * h->H;
    v.xn = (char*)W0.xn + (size_t)rb * h->H * asz;
    v.q = (char*)W0.q + (size_t)rb * h->q_cols * asz;
    v.attn_out = (char*)W0.attn_out + (size_t)rb * h->q_cols * asz;
    v.hbuf = (char*)W0.hbuf + (size_t)rb * h->F * asz;
    if (W0.tp_buf) v.tp_buf = W0.tp_buf + (size_t)rb * h->H;
    // split-K partial buffers: one half each
    v.ws_a_floats = W0.ws_a_floats / 2;
    v.ws_b_floats = W0.ws_b_floats / 2;
    if (role == 2) {
      v.ws_a = W0.ws_a + v.ws_a_floats;
      v.ws_b = W0.ws_b + v.ws_b_floats;
      if (W0.ws_c) v.ws_c = W0.ws_c + v.ws_b_floats;
    }
    StepLayout Lv = L;
    Lv.T = Tc;
    Lv.tokens += rb;
    Lv.positions += rb;
    Lv.slots += rb;
    Lv.tile_seq += role == 1 ? 0 : ts;
    Lv.tile_q0 += role == 1 ? 0 : ts;
    Lv.n_qtiles = role == 1 ? ts : L.n_qtiles - ts;
    h->ws = v;
    h->lay = Lv;
    h->stream = role == 1 ? s0 : h->stream2;
    h->pipe_role = role;
    h->view_r0 = rb;
    h->view_T = L.T;
    h->skip_head = role == 1;
    h->gemm_max_ctas = std::max(1, h->sm_count - 16);
    rc = forward_step(h);
    const float* lf = h->ws.logits_final;
    h->ws = W0;
    h->ws.logits_final = lf;
  }
  h->lay = L;
  h->stream = s0;
  h->pipe_role = 0;
  h->view_r0 = 0;
  h->view_T = 0;
  h->skip_head = false;
  h->gemm_max_ctas = 0;
  if (rc) return rc;
  NV_CHECK(h, cudaEventRecord(h->pipe_join, h->stream2));
  NV_CHECK(h, cudaStreamWaitEvent(s0, h->pipe_join, 0));
  return NVLLM_OK;
}

}  // namespace nvllm

using namespace nvllm;

extern "C" int nvllm_set_profile(nvllm_handle* h, int enabled) {
  if (!h) return NVLLM_ERR_INVALID;
  NV_CHECK(h, cudaSetDevice(h->dev));
  if (enabled < 0 || enabled > 2) return set_err(h, NVLLM_ERR_INVALID, "profile mode %d", enabled);
  int rc = prof_collect(h);
  if (rc) return rc;
  memset(h->prof, 0, sizeof h->prof);
  if (h->profile == 2 && enabled != 2) trace_agg_set(nullptr, 0);
  if (enabled == 2) {
    if (!h->prof_agg) NV_CHECK(h, cudaMalloc((void**)&h->prof_agg, (size_t)PROF_AGG_CAP * 32));
    h->profile = 2;
    return prof_agg_reset(h);
  }
  h->profile = enabled;
  return NVLLM_OK;
}

extern "C" int nvllm_get_profile(nvllm_handle* h, nvllm_profile_entry* out) {
  if (!h || !out) return NVLLM_ERR_INVALID;
  NV_CHECK(h, cudaSetDevice(h->dev));
  int rc = prof_collect(h);
  if (rc) return rc;
  memcpy(out, h->prof, sizeof h->prof);
  return NVLLM_OK;
}
