// weights.cu — nvllm_set_weight / nvllm_finalize.
//
// The caller hands tensors in the reference's POST-LOAD layouts (what TransformerModel holds
// after purego/tensor/generic_loader.go:184-265,353-604): every linear [in, out] row-major.
// The device layout is [out, in] (K contiguous) so that GEMV rows are contiguous streams and
// the tcgen05 GEMM sees two K-major operands; Q, K, V (or Falcon's fused [K|V],
// generic_loader.go:750-765) are stacked into one [Q heads | K heads | V heads] matrix so
// the three projections are one GEMM.  Conversion (fp32/bf16/f16 -> fp32 or bf16) and the
// transpose run on the device; the host buffer is only read during the call.
#include <math.h>

#include <vector>

#include "common.cuh"
#include "handle.h"

using namespace nvllm;

static size_t api_dt_size(int dt) { return dt == NVLLM_DT_F32 ? 4 : 2; }

// stage the caller's buffer on the device
static int stage_in(nvllm_handle* h, const void* host, size_t bytes, void** dev) {
  NV_CHECK(h, cudaMalloc(dev, bytes));
  cudaError_t e = cudaMemcpyAsync(*dev, host, bytes, cudaMemcpyHostToDevice, h->stream);
  if (e != cudaSuccess) {
    cudaFree(*dev);
    return set_err(h, NVLLM_ERR_CUDA, "cudaMemcpy(weight): %s", cudaGetErrorString(e));
  }
  return NVLLM_OK;
}

extern "C" int nvllm_set_weight(nvllm_handle* h, int id, int layer, int dtype, const void* host,
                                uint64_t numel) {
  if (!h) return NVLLM_ERR_INVALID;
  if (!host) return set_err(h, NVLLM_ERR_INVALID, "null weight pointer");
  if (dtype < 0 || dtype > 2) return set_err(h, NVLLM_ERR_INVALID, "bad dtype %d", dtype);
  if (h->finalized) return set_err(h, NVLLM_ERR_STATE, "set_weight after finalize");
  NV_CHECK(h, cudaSetDevice(h->dev));
  const int H = h->H, V = h->V, F = h->F;
  const bool model_level = id < 16;
  if (!model_level && (layer < 0 || layer >= h->L))
    return set_err(h, NVLLM_ERR_INVALID, "layer %d out of range for weight %d", layer, id);
  if (id < 0 || id >= 64) return set_err(h, NVLLM_ERR_INVALID, "unknown weight id %d", id);
  LayerW* w = model_level ? nullptr : &h->layers[layer];

  // what to do with this tensor
  enum { FLAT_DT, FLAT_F32, TRANSPOSE_DT, TRANSPOSE_F32 } mode = FLAT_F32;
  int glu_F = 0;  // > 0: interleave the [gate | up] halves row by row
  void* dst = nullptr;
  size_t expect = 0;
  int K = 0, N = 0, ld = 0;  // transpose: src [K, N] -> dst [N, ld]
  const bool mqa = h->cfg.attention_type == NVLLM_ATTN_MQA;
  switch (id) {
    case NVLLM_W_TOKEN_EMBEDDING: mode = FLAT_DT; dst = h->wte; expect = (size_t)V * H; break;
    case NVLLM_W_POS_EMBEDDING:
      if (!h->wpe) return set_err(h, NVLLM_ERR_INVALID, "model has no learned positions");
      mode = FLAT_DT; dst = h->wpe; expect = (size_t)h->cfg.max_seq_len * H; break;
    case NVLLM_W_FINAL_NORM_W: dst = h->lnf_w; expect = H; break;
    case NVLLM_W_FINAL_NORM_B: dst = h->lnf_b; expect = H; break;
    case NVLLM_W_LM_HEAD:
      if (!h->lm_head_owned)
        return set_err(h, NVLLM_ERR_INVALID, "tied_embedding is set: LM head is the embedding");
      mode = TRANSPOSE_DT; dst = h->lm_head; K = H; N = V; ld = H; expect = (size_t)H * V; break;
    case NVLLM_W_ATTN_NORM_W: dst = w->attn_norm_w; expect = H; break;
    case NVLLM_W_ATTN_NORM_B: dst = w->attn_norm_b; expect = H; break;
    case NVLLM_W_FFN_NORM_W: dst = w->ffn_norm_w; expect = H; break;
    case NVLLM_W_FFN_NORM_B: dst = w->ffn_norm_b; expect = H; break;
    case NVLLM_W_Q:
      mode = TRANSPOSE_DT; dst = w->wqkv; K = H; N = h->q_cols; ld = H; expect = (size_t)H * N; break;
    case NVLLM_W_K:
    case NVLLM_W_V: {
      if (mqa) return set_err(h, NVLLM_ERR_INVALID, "MQA takes the fused NVLLM_W_KV");
      mode = TRANSPOSE_DT; K = H; N = h->kv_cols; ld = H; expect = (size_t)H * N;
      size_t row0 = h->q_cols + (id == NVLLM_W_V ? h->kv_cols : 0);
      dst = (char*)w->wqkv + row0 * H * dt_size(h->wdt);
      break;
    }
    case NVLLM_W_KV:
      if (!mqa) return set_err(h, NVLLM_ERR_INVALID, "NVLLM_W_KV is MQA only");
      mode = TRANSPOSE_DT; K = H; N = 2 * h->hd; ld = H; expect = (size_t)H * N;
      dst = (char*)w->wqkv + (size_t)h->q_cols * H * dt_size(h->wdt);
      break;
    case NVLLM_W_O:
      mode = TRANSPOSE_DT; dst = w->wo; K = h->q_cols; N = H; ld = h->q_cols;
      expect = (size_t)h->q_cols * H; break;
    case NVLLM_W_Q_BIAS: dst = w->bqkv; expect = h->q_cols; w->has_bq = true; break;
    case NVLLM_W_K_BIAS: dst = w->bqkv + h->q_cols; expect = h->kv_cols; w->has_bk = true; break;
    case NVLLM_W_V_BIAS:
      dst = w->bqkv + h->q_cols + h->kv_cols; expect = h->kv_cols; w->has_bv = true; break;
    case NVLLM_W_O_BIAS: dst = w->bo; expect = H; break;
    case NVLLM_W_FFN_W1:
      if (h->moe) return set_err(h, NVLLM_ERR_INVALID, "MoE model takes MOE_INPUT/MOE_OUTPUT");
      mode = TRANSPOSE_DT; dst = w->w1; K = H; N = h->w1_cols; ld = H; expect = (size_t)H * N;
      if (h->swiglu) glu_F = F;
      break;
    case NVLLM_W_FFN_B1:
      if (h->moe || h->swiglu)  // transformer.go:50-66: the SwiGLU branch never adds B1
        return set_err(h, NVLLM_ERR_INVALID, "FFN_B1 only exists on the GELU branch");
      dst = w->b1; expect = F; break;
    case NVLLM_W_FFN_W2:
      if (h->moe) return set_err(h, NVLLM_ERR_INVALID, "MoE model takes MOE_INPUT/MOE_OUTPUT");
      mode = TRANSPOSE_DT; dst = w->w2; K = F; N = H; ld = F; expect = (size_t)F * H; break;
    case NVLLM_W_FFN_B2:
      if (h->moe) return set_err(h, NVLLM_ERR_INVALID, "MoE layers have no B2");
      dst = w->b2; expect = H; break;
    case NVLLM_W_MOE_ROUTER:
      if (!h->moe) return set_err(h, NVLLM_ERR_INVALID, "not a MoE model");
      mode = TRANSPOSE_F32; dst = w->router; K = H; N = h->E; ld = H; expect = (size_t)H * N; break;
    case NVLLM_W_MOE_INPUT:
      if (!h->moe) return set_err(h, NVLLM_ERR_INVALID, "not a MoE model");
      mode = FLAT_DT; dst = w->moe_in; expect = (size_t)h->E * 2 * F * H; break;
    case NVLLM_W_MOE_OUTPUT:
      if (!h->moe) return set_err(h, NVLLM_ERR_INVALID, "not a MoE model");
      mode = FLAT_DT; dst = w->moe_out; expect = (size_t)h->E * H * F; break;
    default: return set_err(h, NVLLM_ERR_INVALID, "unknown weight id %d", id);
  }
  if (numel != expect)
    return set_err(h, NVLLM_ERR_INVALID, "weight %d (layer %d): numel %llu, expected %llu", id,
                   layer, (unsigned long long)numel, (unsigned long long)expect);
  void* stage = nullptr;
  int rc = stage_in(h, host, numel * api_dt_size(dtype), &stage);
  if (rc) return rc;
  cudaError_t e;
  switch (mode) {
    case FLAT_DT: e = launch_convert(h->stream, dtype, stage, numel, h->wdt, dst); break;
    case FLAT_F32: e = launch_convert(h->stream, dtype, stage, numel, DT_F32, dst); break;
    case TRANSPOSE_DT:
      e = launch_transpose_convert(h->stream, dtype, stage, K, N, h->wdt, dst, ld, glu_F);
      break;
    default: e = launch_transpose_convert(h->stream, dtype, stage, K, N, DT_F32, dst, ld); break;
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);  // host buffer is free on return
  cudaFree(stage);
  if (e != cudaSuccess)
    return set_err(h, NVLLM_ERR_CUDA, "weight repack failed: %s", cudaGetErrorString(e));
  h->have[(size_t)(model_level ? 0 : layer + 1) * 64 + id] = 1;
  return NVLLM_OK;
}

extern "C" int nvllm_finalize(nvllm_handle* h) {
  if (!h) return NVLLM_ERR_INVALID;
  if (h->finalized) return set_err(h, NVLLM_ERR_STATE, "already finalized");
  NV_CHECK(h, cudaSetDevice(h->dev));
  auto have = [&](int layer, int id) { return h->have[(size_t)(layer + 1) * 64 + id] != 0; };
  auto need = [&](int layer, int id, const char* name) {
    if (have(layer, id)) return 0;
    return set_err(h, NVLLM_ERR_INVALID, "missing weight %s (id %d, layer %d)", name, id, layer);
  };
  const nvllm_config& c = h->cfg;
  const bool mqa = c.attention_type == NVLLM_ATTN_MQA;
  int rc;
#define NEED(layer, id) \
  if ((rc = need(layer, id, #id)) != 0) return rc;
  NEED(-1, NVLLM_W_TOKEN_EMBEDDING);
  if (c.position_type == NVLLM_POS_LEARNED) NEED(-1, NVLLM_W_POS_EMBEDDING);
  NEED(-1, NVLLM_W_FINAL_NORM_W);
  if (!c.tied_embedding) NEED(-1, NVLLM_W_LM_HEAD);
  // tensor.go:197: a norm with no bias is an RMSNorm
  if (!have(-1, NVLLM_W_FINAL_NORM_B)) {
    cudaFree(h->lnf_b);
    h->lnf_b = nullptr;
  }
  for (int l = 0; l < h->L; l++) {
    LayerW& w = h->layers[l];
    NEED(l, NVLLM_W_ATTN_NORM_W);
    if (!have(l, NVLLM_W_ATTN_NORM_B)) { cudaFree(w.attn_norm_b); w.attn_norm_b = nullptr; }
    if (!h->parallel) NEED(l, NVLLM_W_FFN_NORM_W);
    if (!have(l, NVLLM_W_FFN_NORM_B)) { cudaFree(w.ffn_norm_b); w.ffn_norm_b = nullptr; }
    NEED(l, NVLLM_W_Q);
    if (mqa) {
      NEED(l, NVLLM_W_KV);
    } else {
      NEED(l, NVLLM_W_K);
      NEED(l, NVLLM_W_V);
    }
    NEED(l, NVLLM_W_O);
    // biases: MHA adds Q/K/V/O biases (attention.go:66-86), MQA adds Q and O biases when present
    // (mqa.go:101-114,88-96), GQA has none
    if (!(w.has_bq || w.has_bk || w.has_bv)) { cudaFree(w.bqkv); w.bqkv = nullptr; }
    if (!have(l, NVLLM_W_O_BIAS)) { cudaFree(w.bo); w.bo = nullptr; }
    if (h->moe) {
      NEED(l, NVLLM_W_MOE_ROUTER);
      NEED(l, NVLLM_W_MOE_INPUT);
      NEED(l, NVLLM_W_MOE_OUTPUT);
    } else {
      NEED(l, NVLLM_W_FFN_W1);
      NEED(l, NVLLM_W_FFN_W2);
      if (!have(l, NVLLM_W_FFN_B1)) { cudaFree(w.b1); w.b1 = nullptr; }
      if (!have(l, NVLLM_W_FFN_B2)) { cudaFree(w.b2); w.b2 = nullptr; }
    }
  }
#undef NEED

  // RoPE tables, computed in double exactly like rope.go:18-50 (cos/sin of pos * base^(-2i/hd)),
  // stored un-duplicated [max_seq_len, hd/2]
  if (h->use_rope) {
    const int half = h->hd / 2, n = c.max_seq_len;
    const double base = mqa ? 10000.0 : c.rope_base;  // mqa.go:35 / generic_model.go:92
    std::vector<float> cs((size_t)n * half), sn((size_t)n * half);
    for (int i = 0; i < half; i++) {
      const double freq = 1.0 / pow(base, (double)(2 * i) / (double)h->hd);
      for (int pos = 0; pos < n; pos++) {
        const double ang = (double)pos * freq;
        cs[(size_t)pos * half + i] = (float)cos(ang);
        sn[(size_t)pos * half + i] = (float)sin(ang);
      }
    }
    NV_CHECK(h, cudaMalloc((void**)&h->cosT, cs.size() * 4));
    NV_CHECK(h, cudaMalloc((void**)&h->sinT, sn.size() * 4));
    NV_CHECK(h, cudaMemcpy(h->cosT, cs.data(), cs.size() * 4, cudaMemcpyHostToDevice));
    NV_CHECK(h, cudaMemcpy(h->sinT, sn.data(), sn.size() * 4, cudaMemcpyHostToDevice));
    h->weight_bytes += 2 * cs.size() * 4;
  }

  // paged KV pool: K and V each [L][num_blocks][nkv][256][hd]
  h->kv_layer_elems = (size_t)h->num_blocks * h->nkv * KV_BLOCK * h->hd;
  const size_t pool_bytes = h->kv_layer_elems * h->L * dt_size(h->kvdt);
  cudaError_t e = cudaMalloc(&h->kpool, pool_bytes);
  if (e == cudaSuccess) e = cudaMalloc(&h->vpool, pool_bytes);
  if (e != cudaSuccess)
    return set_err(h, NVLLM_ERR_NOMEM, "KV pool of %d blocks (%.1f GB) does not fit: %s",
                   h->num_blocks, 2.0 * pool_bytes / 1e9, cudaGetErrorString(e));
  NV_CHECK(h, cudaMemsetAsync(h->kpool, 0, pool_bytes, h->stream));
  NV_CHECK(h, cudaMemsetAsync(h->vpool, 0, pool_bytes, h->stream));
  h->kv_bytes = 2 * pool_bytes;

  NV_CHECK(h, cudaStreamSynchronize(h->stream));
  h->finalized = true;
  return NVLLM_OK;
}
