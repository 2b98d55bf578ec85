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
#include <stdlib.h>

#include <algorithm>
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

// One device-side repack of (a slice of) the staged tensor.
struct Repack {
  bool transpose;   // true: dst[n*row_mul+row_add, k] = src[k, n]; false: flat convert
  bool to_f32;      // destination dtype fp32 (norms, biases, router) instead of the weight dtype
  size_t src_off;   // elements into the staged tensor
  size_t ld_src;    // transpose: source row stride
  int K, N;         // transpose: extent taken; flat: N = element count
  void* dst;
  int ld_dst, row_mul, row_add;
};

extern "C" int nvllm_set_weight(nvllm_handle* h, int id, int layer, int dtype, const void* host,
                                uint64_t numel) {
  return nvllm::set_weight_src(h, id, layer, dtype, host, numel, nvllm::WeightSrc());
}

// The one implementation behind nvllm_set_weight (reference post-load layout, [in, out]) and the
// safetensors loader (safetensors.cu), which hands the tensors as the file holds them: PyTorch
// linears are [out, in] — already the device orientation, so the "transpose" of the tables
// below becomes a strided row copy — and fused tensors (GPT-2 c_attn [H, 3H], its bias) are
// addressed as a column slice of a wider matrix (src.ld / src.col0).
int nvllm::set_weight_src(nvllm_handle* h, int id, int layer, int dtype, const void* host,
                          uint64_t numel, const WeightSrc& src) {
  if (!h) return NVLLM_ERR_INVALID;
  if (!host) return set_err(h, NVLLM_ERR_INVALID, "null weight pointer");
  if (dtype < 0 || dtype > 2) return set_err(h, NVLLM_ERR_INVALID, "bad dtype %d", dtype);
  if (h->finalized) return set_err(h, NVLLM_ERR_STATE, "set_weight after finalize");
  NV_CHECK(h, cudaSetDevice(h->dev));
  const int H = h->H, V = h->V;
  const bool model_level = id < 16;
  if (!model_level && (layer < 0 || layer >= h->L))
    return set_err(h, NVLLM_ERR_INVALID, "layer %d out of range for weight %d", layer, id);
  if (id < 0 || id >= 64) return set_err(h, NVLLM_ERR_INVALID, "unknown weight id %d", id);
  LayerW* w = model_level ? nullptr : &h->layers[layer];
  const bool mqa = h->cfg.attention_type == NVLLM_ATTN_MQA;
  const size_t wsz = dt_size(h->wdt);

  // The caller always hands the FULL tensor (reference post-load layout); with tp > 1 this rank
  // keeps its slice: Q / K / V / W1 columns and O / W2 rows by head / FFN unit (Megatron
  // column -> row split), LM head by vocabulary rows.  Padding heads (Falcon 71 -> 72) keep the
  // zeros their buffers were allocated with.
  const int rank = h->rank;
  const int hd = h->hd;
  const int qg = h->nh_g * hd;                       // real global Q columns
  const int q0 = rank * h->q_cols;                   // this rank's first Q column
  const int qn = std::max(0, std::min(h->q_cols, qg - q0));  // real (non-padding) columns here
  const int kvg = h->nkv_g * hd;
  const bool kv_split = h->nkv_g > 1 && h->tp > 1;   // MQA's single KV head is replicated
  const int kv0 = kv_split ? rank * h->kv_cols : 0;
  const int Fg = h->F_g, Fl = h->F, f0 = rank * h->F;
  Repack ops[2];
  int n_ops = 1;
  size_t expect = 0;
  Repack& r = ops[0];
  r = Repack{false, true, 0, 0, 0, 0, nullptr, 0, 1, 0};
  auto flat_f32 = [&](float* dst, size_t off, size_t cnt, size_t full) {
    r.dst = dst; r.src_off = off; r.N = (int)cnt; expect = full;
  };
  auto tr = [&](Repack& o, void* dst, size_t off, size_t ld_src, int K, int N, int ld_dst,
                int mul = 1, int add = 0) {
    o = Repack{true, false, off, ld_src, K, N, dst, ld_dst, mul, add};
  };
  switch (id) {
    case NVLLM_W_TOKEN_EMBEDDING:
      r.to_f32 = false; r.dst = h->wte; r.N = 0; expect = (size_t)V * H;
      r.N = -1;  // whole tensor (count may exceed int)
      break;
    case NVLLM_W_POS_EMBEDDING:
      if (!h->wpe) return set_err(h, NVLLM_ERR_INVALID, "model has no learned positions");
      r.to_f32 = false; r.dst = h->wpe; r.N = -1; expect = (size_t)h->cfg.max_seq_len * H; break;
    case NVLLM_W_FINAL_NORM_W: flat_f32(h->lnf_w, 0, H, H); break;
    case NVLLM_W_FINAL_NORM_B: flat_f32(h->lnf_b, 0, H, H); break;
    case NVLLM_W_LM_HEAD:
      if (!h->lm_head_owned)
        return set_err(h, NVLLM_ERR_INVALID, "tied_embedding is set: LM head is the embedding");
      expect = (size_t)H * V;
      tr(r, h->lm_head, (size_t)rank * h->Vl * (h->tp > 1), V, H, h->Vl_valid, H);
      break;
    case NVLLM_W_ATTN_NORM_W: flat_f32(w->attn_norm_w, 0, H, H); break;
    case NVLLM_W_ATTN_NORM_B: flat_f32(w->attn_norm_b, 0, H, H); break;
    case NVLLM_W_FFN_NORM_W: flat_f32(w->ffn_norm_w, 0, H, H); break;
    case NVLLM_W_FFN_NORM_B: flat_f32(w->ffn_norm_b, 0, H, H); break;
    case NVLLM_W_Q:
      expect = (size_t)H * qg;
      tr(r, w->wqkv, q0, qg, H, qn, H);
      break;
    case NVLLM_W_K:
    case NVLLM_W_V: {
      if (mqa) return set_err(h, NVLLM_ERR_INVALID, "MQA takes the fused NVLLM_W_KV");
      expect = (size_t)H * kvg;
      size_t row0 = h->q_cols + (id == NVLLM_W_V ? h->kv_cols : 0);
      tr(r, (char*)w->wqkv + row0 * H * wsz, kv0, kvg, H, h->kv_cols, H);
      break;
    }
    case NVLLM_W_KV:
      if (!mqa) return set_err(h, NVLLM_ERR_INVALID, "NVLLM_W_KV is MQA only");
      expect = (size_t)H * 2 * hd;
      tr(r, (char*)w->wqkv + (size_t)h->q_cols * H * wsz, 0, 2 * hd, H, 2 * hd, H);
      break;
    case NVLLM_W_O:
      expect = (size_t)qg * H;
      tr(r, w->wo, (size_t)q0 * H, H, qn, H, h->q_cols);
      break;
    case NVLLM_W_Q_BIAS: flat_f32(w->bqkv, q0, qn, qg); w->has_bq = true; break;
    case NVLLM_W_K_BIAS: flat_f32(w->bqkv + h->q_cols, kv0, h->kv_cols, kvg); w->has_bk = true; break;
    case NVLLM_W_V_BIAS:
      flat_f32(w->bqkv + h->q_cols + h->kv_cols, kv0, h->kv_cols, kvg); w->has_bv = true; break;
    case NVLLM_W_O_BIAS: flat_f32(w->bo, 0, H, H); break;
    case NVLLM_W_FFN_W1:
      if (h->moe) return set_err(h, NVLLM_ERR_INVALID, "MoE model takes MOE_INPUT/MOE_OUTPUT");
      if (h->swiglu) {  // [gate | up] columns -> interleaved rows (gate j -> 2j, up j -> 2j + 1)
        expect = (size_t)H * 2 * Fg;
        tr(ops[0], w->w1, f0, 2 * (size_t)Fg, H, Fl, H, 2, 0);
        tr(ops[1], w->w1, (size_t)Fg + f0, 2 * (size_t)Fg, H, Fl, H, 2, 1);
        n_ops = 2;
      } else {
        expect = (size_t)H * Fg;
        tr(r, w->w1, f0, Fg, H, Fl, H);
      }
      break;
    case W_INT_FFN_GATE:  // one half of a SwiGLU W1 (Llama gate_proj / up_proj): [H, F]
    case W_INT_FFN_UP:
      if (h->moe || !h->swiglu) return set_err(h, NVLLM_ERR_INVALID, "gate / up halves need a SwiGLU FFN");
      expect = (size_t)H * Fg;
      tr(r, w->w1, f0, Fg, H, Fl, H, 2, id == W_INT_FFN_UP ? 1 : 0);
      break;
    case NVLLM_W_FFN_B1:
      if (h->moe || h->swiglu)  // transformer.go:50-66: the SwiGLU branch never adds B1
        return set_err(h, NVLLM_ERR_INVALID, "FFN_B1 only exists on the GELU branch");
      flat_f32(w->b1, f0, Fl, Fg); break;
    case NVLLM_W_FFN_W2:
      if (h->moe) return set_err(h, NVLLM_ERR_INVALID, "MoE model takes MOE_INPUT/MOE_OUTPUT");
      expect = (size_t)Fg * H;
      tr(r, w->w2, (size_t)f0 * H, H, Fl, H, Fl);
      break;
    case NVLLM_W_FFN_B2:
      if (h->moe) return set_err(h, NVLLM_ERR_INVALID, "MoE layers have no B2");
      flat_f32(w->b2, 0, H, H); break;
    case NVLLM_W_MOE_ROUTER:
      if (!h->moe) return set_err(h, NVLLM_ERR_INVALID, "not a MoE model");
      expect = (size_t)H * h->E;
      tr(r, w->router, 0, h->E, H, h->E, H);
      r.to_f32 = true;
      break;
    case NVLLM_W_MOE_INPUT:
      if (!h->moe) return set_err(h, NVLLM_ERR_INVALID, "not a MoE model");
      r.to_f32 = false; r.dst = w->moe_in; r.N = -1; expect = (size_t)h->E * 2 * Fg * H; break;
    case NVLLM_W_MOE_OUTPUT:
      if (!h->moe) return set_err(h, NVLLM_ERR_INVALID, "not a MoE model");
      r.to_f32 = false; r.dst = w->moe_out; r.N = -1; expect = (size_t)h->E * H * Fg; break;
    // Mamba2Layer tensors (mamba2.go:9-28): the reference keeps in_proj / out_proj as loaded,
    // [out, in] (it multiplies by their transpose, :87,:178) — already the device orientation
    case NVLLM_W_MAMBA_IN_PROJ:
    case NVLLM_W_MAMBA_OUT_PROJ:
    case NVLLM_W_MAMBA_CONV_W:
    case NVLLM_W_MAMBA_CONV_B:
    case NVLLM_W_MAMBA_DT_BIAS:
    case NVLLM_W_MAMBA_A_LOG:
    case NVLLM_W_MAMBA_D:
    case NVLLM_W_MAMBA_NORM: {
      if (h->is_mamba.empty() || !h->is_mamba[layer])
        return set_err(h, NVLLM_ERR_INVALID, "layer %d is not a Mamba2 layer (nvllm_set_layer_types)", layer);
      r.N = -1;
      switch (id) {
        case NVLLM_W_MAMBA_IN_PROJ: r.to_f32 = false; r.dst = w->m_in_proj; expect = (size_t)h->m_P * H; break;
        case NVLLM_W_MAMBA_OUT_PROJ: r.to_f32 = false; r.dst = w->m_out_proj; expect = (size_t)H * h->m_E; break;
        case NVLLM_W_MAMBA_CONV_W: r.dst = w->m_conv_w; expect = (size_t)h->m_conv * h->m_K; break;
        case NVLLM_W_MAMBA_CONV_B: r.dst = w->m_conv_b; expect = h->m_conv; break;
        case NVLLM_W_MAMBA_DT_BIAS: r.dst = w->m_dt_bias; expect = h->m_nh; break;
        case NVLLM_W_MAMBA_A_LOG: r.dst = w->m_a_log; expect = h->m_nh; break;
        case NVLLM_W_MAMBA_D: r.dst = w->m_d; expect = h->m_nh; break;
        default: r.dst = w->m_norm; expect = h->m_E; break;
      }
      break;
    }
    default: return set_err(h, NVLLM_ERR_INVALID, "unknown weight id %d", id);
  }
  // `numel` is what the caller's buffer holds: the tensor itself, or the wider matrix it is a
  // column slice of (src.ld > 0: [rows, src.ld] with this tensor at columns [col0, col0 + width))
  uint64_t want = expect;
  if (src.ld > 0) {
    const size_t width = ops[0].transpose ? ops[0].ld_src : expect;  // natural row length
    if (src.col0 + width > src.ld)
      return set_err(h, NVLLM_ERR_INVALID, "weight %d: column slice [%zu, +%zu) outside ld %zu", id,
                     src.col0, width, src.ld);
    want = (expect / width) * src.ld;
  }
  if (numel != want)
    return set_err(h, NVLLM_ERR_INVALID, "weight %d (layer %d): numel %llu, expected %llu", id,
                   layer, (unsigned long long)numel, (unsigned long long)want);
  void* stage = nullptr;
  const size_t ssz = api_dt_size(dtype);
  int rc = stage_in(h, host, numel * ssz, &stage);
  if (rc) return rc;
  cudaError_t e = cudaSuccess;
  for (int i = 0; i < n_ops && e == cudaSuccess; i++) {
    Repack o = ops[i];
    const int ddt = o.to_f32 ? DT_F32 : h->wdt;
    if (o.transpose) {
      // element (k, n) of the logical [in, out] slice: k < K input features, n < N outputs
      const size_t r0 = o.src_off / o.ld_src, c0 = o.src_off % o.ld_src;
      if (src.out_in) {
        // the buffer is [out, in]: (k, n) lives at (c0 + n) * in_full + r0 + k — rows of the
        // destination are contiguous runs of the source
        const size_t in_full = (size_t)(expect / o.ld_src);
        const void* sp = (const char*)stage + (c0 * in_full + r0) * ssz;
        e = launch_copy_rows(h->stream, dtype, sp, in_full, o.N, o.K, ddt, o.dst, o.ld_dst,
                             o.row_mul, o.row_add);
      } else {
        const size_t ld = src.ld > 0 ? src.ld : o.ld_src;
        const void* sp = (const char*)stage + (r0 * ld + src.col0 + c0) * ssz;
        e = launch_transpose_convert(h->stream, dtype, sp, ld, o.K, o.N, ddt, o.dst, o.ld_dst,
                                     o.row_mul, o.row_add);
      }
      continue;
    }
    const void* sp = (const char*)stage + (o.src_off + src.col0) * ssz;
    e = launch_convert(h->stream, dtype, sp, o.N < 0 ? (size_t)numel : (size_t)o.N, ddt, o.dst,
                       id == NVLLM_W_MOE_INPUT ? Fg : 0, H);
  }
  // row-parallel output biases are added once: ranks other than 0 hold zeros
  if (e == cudaSuccess && h->rank != 0 && (id == NVLLM_W_O_BIAS || id == NVLLM_W_FFN_B2))
    e = cudaMemsetAsync(ops[0].dst, 0, (size_t)H * 4, h->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);  // host buffer is free on return
  cudaFree(stage);
  if (e != cudaSuccess)
    return set_err(h, NVLLM_ERR_CUDA, "weight repack failed: %s", cudaGetErrorString(e));
  h->have[(size_t)(model_level ? 0 : layer + 1) * 64 + id] = 1;
  // both halves of a SwiGLU W1 make the whole
  if (id == W_INT_FFN_GATE || id == W_INT_FFN_UP) {
    const size_t base = (size_t)(layer + 1) * 64;
    if (h->have[base + W_INT_FFN_GATE] && h->have[base + W_INT_FFN_UP]) h->have[base + NVLLM_W_FFN_W1] = 1;
  }
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
    if (!h->is_mamba.empty() && h->is_mamba[l]) {
      for (int id = NVLLM_W_MAMBA_IN_PROJ; id <= NVLLM_W_MAMBA_OUT_PROJ; id++)
        if ((rc = need(l, id, "NVLLM_W_MAMBA_*")) != 0) return rc;
    } else {
      NEED(l, NVLLM_W_Q);
      if (mqa) {
        NEED(l, NVLLM_W_KV);
      } else {
        NEED(l, NVLLM_W_K);
        NEED(l, NVLLM_W_V);
      }
      NEED(l, NVLLM_W_O);
    }
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

  // RMSNorm folded into the GEMMs (forward.cu, gemm_tcgen05.cu): y = (x * rstd * g) W^T =
  // rstd * ((x * g) W^T) — the kernel that PRODUCES x (embedding, O / FFN-down GEMM epilogue)
  // also writes bf16(x * g) with the weight g of the norm that reads x next and the row's sum
  // of squares; the consuming GEMM reads that matrix and scales its accumulator rows by rstd.
  // The weights stay exactly the bf16 values the reference loader would hold, and the
  // activation is rounded to bf16 once, as on the un-fused path.
  {
    const char* env = getenv("NVLLM_FOLD");
    bool rms = !have(-1, NVLLM_W_FINAL_NORM_B);
    for (int l = 0; l < h->L; l++) rms = rms && !have(l, NVLLM_W_ATTN_NORM_B) && !have(l, NVLLM_W_FFN_NORM_B);
    h->fold = rms && !h->moe && !h->parallel && h->n_mamba == 0 && h->tp == 1 && h->wdt == DT_BF16 && h->adt == DT_BF16 &&
              h->use_tc5 && (h->H % 8) == 0 && !(env && atoi(env) == 0);
  }

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
  if (h->n_mamba) {
    // SSM state snapshot of every (KV block, Mamba2 layer): see handle.h
    const size_t ssm_bytes = (size_t)h->num_blocks * h->n_mamba * h->m_nh * h->m_hd * h->m_N * 4;
    e = cudaMalloc((void**)&h->ssm_pool, ssm_bytes);
    if (e != cudaSuccess)
      return set_err(h, NVLLM_ERR_NOMEM, "SSM state pool of %d blocks (%.1f GB) does not fit: %s",
                     h->num_blocks, ssm_bytes / 1e9, cudaGetErrorString(e));
    NV_CHECK(h, cudaMemsetAsync(h->ssm_pool, 0, ssm_bytes, h->stream));
    h->kv_bytes += ssm_bytes;
  }

  NV_CHECK(h, cudaStreamSynchronize(h->stream));
  h->finalized = true;
  return NVLLM_OK;
}
