// nvllm.cu — C-ABI entry points of libnvllm_b200.so: handle lifecycle, weight intake
// (the reference's post-load layouts -> device layouts) and the KV pool.  The forward
// pass itself is forward.cu; single-operator entry points are ops.cu.
//
// Replaces, behind the same boundary, what NewTensorModelRunner + TensorModelRunner.Run do
// in the reference (nanovllm/tensor_model_runner.go:21-97).
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>

#include "common.cuh"
#include "handle.h"

namespace nvllm {

thread_local std::string g_create_error;

int set_err(nvllm_handle* h, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (h)
    h->err = buf;
  else
    g_create_error = buf;
  return code;
}

static size_t round_up(size_t a, size_t b) { return (a + b - 1) / b * b; }

template <typename T>
static cudaError_t dev_alloc(nvllm_handle* h, T** p, size_t count, uint64_t* account, bool zero = false) {
  *p = nullptr;
  if (count == 0) return cudaSuccess;
  size_t bytes = round_up(count * sizeof(T), 256);
  cudaError_t e = cudaMalloc((void**)p, bytes);
  if (e != cudaSuccess) return e;
  if (zero) e = cudaMemsetAsync(*p, 0, bytes, h->stream);
  if (account) *account += bytes;
  return e;
}
static cudaError_t dev_alloc_dt(nvllm_handle* h, void** p, size_t count, int dt, uint64_t* account,
                                bool zero = false) {
  return dev_alloc<uint8_t>(h, (uint8_t**)p, count * dt_size(dt), account, zero);
}

}  // namespace nvllm

using namespace nvllm;

extern "C" {

int nvllm_abi_version(void) { return NVLLM_ABI_VERSION; }

const char* nvllm_last_error(const nvllm_handle* h) {
  return h ? h->err.c_str() : g_create_error.c_str();
}

int nvllm_device_info(int device, int* sm_count, int* cc_major, int* cc_minor, uint64_t* total_mem,
                      char name[256]) {
  cudaDeviceProp p;
  cudaError_t e = cudaGetDeviceProperties(&p, device);
  if (e != cudaSuccess)
    return set_err(nullptr, NVLLM_ERR_CUDA, "cudaGetDeviceProperties(%d): %s", device,
                   cudaGetErrorString(e));
  if (sm_count) *sm_count = p.multiProcessorCount;
  if (cc_major) *cc_major = p.major;
  if (cc_minor) *cc_minor = p.minor;
  if (total_mem) *total_mem = p.totalGlobalMem;
  if (name) {
    strncpy(name, p.name, 255);
    name[255] = 0;
  }
  return NVLLM_OK;
}

int nvllm_create(const nvllm_config* c, const uint8_t* tp_id, nvllm_handle** out) {
  if (!c || !out) return set_err(nullptr, NVLLM_ERR_INVALID, "null argument");
  *out = nullptr;
  if (c->abi_version != NVLLM_ABI_VERSION)
    return set_err(nullptr, NVLLM_ERR_INVALID, "abi_version %d != %d", c->abi_version,
                   NVLLM_ABI_VERSION);
  auto bad = [&](const char* what) {
    return set_err(nullptr, NVLLM_ERR_INVALID, "invalid config: %s", what);
  };
  if (c->vocab_size <= 0 || c->hidden <= 0 || c->num_layers <= 0 || c->num_heads <= 0)
    return bad("non-positive dimension");
  if (c->hidden % 8) return bad("hidden must be a multiple of 8");
  if (c->head_dim != 16 && c->head_dim != 32 && c->head_dim != 64 && c->head_dim != 128)
    return bad("head_dim must be 16, 32, 64 or 128");
  if (c->ffn_dim <= 0 || c->ffn_dim % 8) return bad("ffn_dim must be a positive multiple of 8");
  if (c->attention_type < 0 || c->attention_type > 2) return bad("attention_type");
  if (c->precision != NVLLM_PREC_FP32 && c->precision != NVLLM_PREC_BF16) return bad("precision");
  if (c->max_seq_len <= 0) return bad("max_seq_len");
  int nkv = c->attention_type == NVLLM_ATTN_MQA ? 1
            : c->attention_type == NVLLM_ATTN_MHA ? c->num_heads : c->num_kv_heads;
  if (nkv <= 0 || c->num_heads % nkv) return bad("num_heads must be a multiple of num_kv_heads");
  if (c->use_moe) {
    if (c->num_experts <= 0 || c->num_experts > 256) return bad("num_experts must be 1..256");
    if (c->num_experts_per_tok <= 0 || c->num_experts_per_tok > 32 ||
        c->num_experts_per_tok > c->num_experts)
      return bad("num_experts_per_tok");
    if (c->activation_type != NVLLM_ACT_SWIGLU) return bad("MoE experts are GLU (moe.go:203-212)");
  }
  const int tp = c->tp_size > 1 ? c->tp_size : 1;
  const int rank = tp > 1 ? c->tp_rank : 0;
  int nh_pad = c->num_heads;
  if (tp > 1) {
    if (!tp_id) return bad("tp_size > 1 needs the NCCL unique id of the group (nvllm_tp_unique_id)");
    if (rank < 0 || rank >= tp) return bad("tp_rank outside [0, tp_size)");
    if (c->use_moe) return bad("tensor parallelism is not offered for MoE models");
    if (nkv == 1) {
      nh_pad = (c->num_heads + tp - 1) / tp * tp;  // Falcon: 71 -> 72 zero-weight heads
    } else if (c->num_heads % tp || nkv % tp) {
      return bad("num_heads and num_kv_heads must be multiples of tp_size");
    }
    if (c->ffn_dim % tp || (c->ffn_dim / tp) % 8) return bad("ffn_dim / tp_size must be a multiple of 8");
  }

  cudaError_t e = cudaSetDevice(c->device);
  if (e != cudaSuccess)
    return set_err(nullptr, NVLLM_ERR_CUDA, "cudaSetDevice(%d): %s (no CPU fallback exists)",
                   c->device, cudaGetErrorString(e));
  cudaDeviceProp prop;
  e = cudaGetDeviceProperties(&prop, c->device);
  if (e != cudaSuccess)
    return set_err(nullptr, NVLLM_ERR_CUDA, "cudaGetDeviceProperties: %s", cudaGetErrorString(e));
  if (prop.major != 10)
    return set_err(nullptr, NVLLM_ERR_CUDA,
                   "device %d is sm_%d%d; this library contains sm_100a code only", c->device,
                   prop.major, prop.minor);

  if (const char* env = getenv("NVLLM_PDL")) set_pdl_enabled(atoi(env) != 0);  // debugging aid
  nvllm_handle* h = new nvllm_handle();
  h->cfg = *c;
  if (const char* env = getenv("NVLLM_SKIP")) h->skip_mask = (unsigned)strtoul(env, nullptr, 0);
  h->cfg.num_kv_heads = nkv;
  if (h->cfg.num_kv_blocks <= 0) h->cfg.num_kv_blocks = 1024;  // scheduler.go:17-21
  if (h->cfg.max_num_seqs <= 0) h->cfg.max_num_seqs = 512;
  if (h->cfg.max_batched_tokens <= 0) h->cfg.max_batched_tokens = 16384;
  h->cfg.tp_size = tp;
  h->cfg.tp_rank = rank;
  h->tp = tp;
  h->rank = rank;
  h->dev = c->device;
  h->sm_count = prop.multiProcessorCount;
  h->wdt = h->adt = h->kvdt = (c->precision == NVLLM_PREC_BF16) ? DT_BF16 : DT_F32;
  h->H = c->hidden;
  h->V = c->vocab_size;
  h->L = c->num_layers;
  h->nh_g = c->num_heads;
  h->nkv_g = nkv;
  h->F_g = c->ffn_dim;
  h->nh = nh_pad / tp;
  h->nkv = (nkv == 1) ? 1 : nkv / tp;
  h->hd = c->head_dim;
  h->F = c->ffn_dim / tp;
  h->moe = c->use_moe != 0;
  h->E = h->moe ? c->num_experts : 0;
  h->topk = h->moe ? c->num_experts_per_tok : 0;
  h->swiglu = c->activation_type == NVLLM_ACT_SWIGLU;
  h->parallel = (c->attention_type == NVLLM_ATTN_MQA && c->block_style == NVLLM_BLOCK_PARALLEL);
  h->q_cols = h->nh * h->hd;
  h->kv_cols = h->nkv * h->hd;
  h->qkv_cols = h->q_cols + 2 * h->kv_cols;
  h->w1_cols = h->swiglu ? 2 * h->F : h->F;
  h->ldl = (int)round_up(h->V, 8);
  h->Vl = tp > 1 ? (int)round_up((h->V + tp - 1) / tp, 8) : h->ldl;
  h->Vl_valid = tp > 1 ? std::max(0, std::min(h->Vl, h->V - rank * h->Vl)) : h->V;
  // RoPE: always for MQA (mqa.go:65-67), for GQA when PositionRoPE, never for MHA
  h->use_rope = (c->attention_type == NVLLM_ATTN_MQA) ||
                (c->attention_type == NVLLM_ATTN_GQA && c->position_type == NVLLM_POS_ROPE);
  // attention.go:361-364 (GQA honours AttentionScale); MHA / MQA are always 1/sqrt(hd)
  if (c->attention_type == NVLLM_ATTN_GQA && c->attention_multiplier != 0.f)
    h->attn_scale = c->attention_multiplier;
  else
    h->attn_scale = (float)(1.0 / sqrt((double)h->hd));
  h->num_blocks = h->cfg.num_kv_blocks;
  h->max_blocks_per_seq = (c->max_seq_len + NVLLM_KV_BLOCK_SIZE - 1) / NVLLM_KV_BLOCK_SIZE;
  h->layers.resize(h->L);
  h->have.assign((size_t)(h->L + 1) * 64, 0);

  auto fail = [&](cudaError_t ce, const char* what) {
    int code = (ce == cudaErrorMemoryAllocation) ? NVLLM_ERR_NOMEM : NVLLM_ERR_CUDA;
    set_err(nullptr, code, "%s: %s", what, cudaGetErrorString(ce));
    nvllm_destroy(h);
    return code;
  };
  if ((e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking)) != cudaSuccess)
    return fail(e, "cudaStreamCreate");
  if ((e = cudaEventCreate(&h->ev0)) != cudaSuccess) return fail(e, "cudaEventCreate");
  if ((e = cudaEventCreate(&h->ev1)) != cudaSuccess) return fail(e, "cudaEventCreate");

  // ---- weights (device layouts; see handle.h) ----
  uint64_t* wb = &h->weight_bytes;
  const int H = h->H, wdt = h->wdt;
#define A_DT(ptr, count)                                                                       \
  if ((e = dev_alloc_dt(h, (void**)&(ptr), (size_t)(count), wdt, wb, tp > 1)) != cudaSuccess) \
    return fail(e, "cudaMalloc(weights)");
#define A_F32(ptr, count, zero)                                                         \
  if ((e = dev_alloc<float>(h, &(ptr), (size_t)(count), wb, zero)) != cudaSuccess)      \
    return fail(e, "cudaMalloc(weights)");
  A_DT(h->wte, (size_t)h->V * H);
  if (c->position_type == NVLLM_POS_LEARNED) A_DT(h->wpe, (size_t)c->max_seq_len * H);
  A_F32(h->lnf_w, H, false);
  A_F32(h->lnf_b, H, false);
  if (c->tied_embedding) {
    // [V, H] is already the [out, in] layout the GEMMs want; a TP rank reads its row slice
    h->lm_head = (char*)h->wte + (size_t)rank * (tp > 1 ? h->Vl : 0) * H * dt_size(wdt);
  } else {
    A_DT(h->lm_head, (size_t)(tp > 1 ? h->Vl : h->V) * H);
    h->lm_head_owned = true;
  }
  for (int l = 0; l < h->L; l++) {
    LayerW& w = h->layers[l];
    A_F32(w.attn_norm_w, H, false);
    A_F32(w.attn_norm_b, H, false);
    A_F32(w.ffn_norm_w, H, false);
    A_F32(w.ffn_norm_b, H, false);
    A_DT(w.wqkv, (size_t)h->qkv_cols * H);
    A_F32(w.bqkv, h->qkv_cols, true);
    A_DT(w.wo, (size_t)H * h->q_cols);
    A_F32(w.bo, H, false);
    if (h->moe) {
      A_F32(w.router, (size_t)h->E * H, false);
      A_DT(w.moe_in, (size_t)h->E * 2 * h->F * H);
      A_DT(w.moe_out, (size_t)h->E * H * h->F);
    } else {
      A_DT(w.w1, (size_t)h->w1_cols * H);
      A_F32(w.b1, h->F, false);
      A_DT(w.w2, (size_t)H * h->F);
      A_F32(w.b2, H, false);
    }
  }
#undef A_DT
#undef A_F32
  if (tp > 1) {
    int rc = tp_init(h, tp_id);
    if (rc) {
      g_create_error = h->err;
      nvllm_destroy(h);
      return rc;
    }
  }
  *out = h;
  return NVLLM_OK;
}

// ModelConfig.HybridLayers (config.go:113): which blocks are Mamba2.  Allocates their tensors.
int nvllm_set_layer_types(nvllm_handle* h, const int32_t* types, int n_layers) {
  if (!h || !types) return NVLLM_ERR_INVALID;
  if (h->finalized) return set_err(h, NVLLM_ERR_STATE, "set_layer_types after finalize");
  if (n_layers != h->L) return set_err(h, NVLLM_ERR_INVALID, "%d layer types for %d layers", n_layers, h->L);
  const nvllm_config& c = h->cfg;
  bool any = false;
  for (int l = 0; l < n_layers; l++) any = any || types[l] != 0;
  if (!any) return NVLLM_OK;
  if (h->n_mamba) return set_err(h, NVLLM_ERR_STATE, "layer types are already set");
  if (c.mamba_num_heads <= 0 || c.mamba_head_dim <= 0 || c.mamba_state_size <= 0 || c.mamba_n_groups <= 0 ||
      c.mamba_conv_kernel <= 0 || c.mamba_conv_kernel > 8 || c.mamba_state_size > 128 ||
      c.mamba_head_dim > 128 || c.mamba_num_heads % c.mamba_n_groups != 0 ||
      (c.mamba_num_heads * c.mamba_head_dim) % 8 != 0)
    return set_err(h, NVLLM_ERR_INVALID, "Mamba2 layers need valid mamba_* config fields");
  if (h->tp > 1 || h->moe || h->parallel)
    return set_err(h, NVLLM_ERR_INVALID, "Mamba2 layers: sequential dense blocks on one GPU only");
  NV_CHECK(h, cudaSetDevice(h->dev));
  h->m_nh = c.mamba_num_heads; h->m_hd = c.mamba_head_dim; h->m_N = c.mamba_state_size;
  h->m_G = c.mamba_n_groups; h->m_K = c.mamba_conv_kernel;
  h->m_E = h->m_nh * h->m_hd;
  h->m_conv = h->m_E + 2 * h->m_G * h->m_N;
  h->m_P = h->m_E + h->m_conv + h->m_nh;
  h->is_mamba.assign(h->L, 0);
  h->mamba_idx.assign(h->L, -1);
  const size_t wsz = dt_size(h->wdt);
  for (int l = 0; l < h->L; l++) {
    if (!types[l]) continue;
    h->is_mamba[l] = 1;
    h->mamba_idx[l] = h->n_mamba++;
    LayerW& w = h->layers[l];
    const size_t sizes[8] = {(size_t)h->m_P * h->H * wsz, (size_t)h->H * h->m_E * wsz,
                             (size_t)h->m_conv * h->m_K * 4, (size_t)h->m_conv * 4, (size_t)h->m_nh * 4,
                             (size_t)h->m_nh * 4, (size_t)h->m_nh * 4, (size_t)h->m_E * 4};
    void** ptrs[8] = {&w.m_in_proj, &w.m_out_proj, (void**)&w.m_conv_w, (void**)&w.m_conv_b,
                      (void**)&w.m_dt_bias, (void**)&w.m_a_log, (void**)&w.m_d, (void**)&w.m_norm};
    for (int i = 0; i < 8; i++) {
      cudaError_t e = cudaMalloc(ptrs[i], sizes[i]);
      if (e != cudaSuccess) return set_err(h, NVLLM_ERR_NOMEM, "Mamba2 weights: %s", cudaGetErrorString(e));
      h->weight_bytes += sizes[i];
    }
  }
  h->spec_enabled = 0;  // a discarded speculative step would leave its SSM state in the block
  return NVLLM_OK;
}

int nvllm_destroy(nvllm_handle* h) {
  if (!h) return NVLLM_OK;
  cudaSetDevice(h->dev);
  if (h->stream) cudaStreamSynchronize(h->stream);
  auto F = [](void* p) {
    if (p) cudaFree(p);
  };
  F(h->wte);
  F(h->wpe);
  F(h->lnf_w);
  F(h->lnf_b);
  if (h->lm_head_owned) F(h->lm_head);
  for (auto& w : h->layers) {
    F(w.attn_norm_w); F(w.attn_norm_b); F(w.ffn_norm_w); F(w.ffn_norm_b);
    F(w.wqkv); F(w.bqkv); F(w.wo); F(w.bo); F(w.w1); F(w.b1); F(w.w2); F(w.b2);
    F(w.m_in_proj); F(w.m_out_proj); F(w.m_conv_w); F(w.m_conv_b); F(w.m_dt_bias); F(w.m_a_log);
    F(w.m_d); F(w.m_norm);
    F(w.router); F(w.moe_in); F(w.moe_out);
  }
  F(h->cosT);
  F(h->sinT);
  F(h->kpool);
  F(h->vpool);
  F(h->ssm_pool);
  F(h->d_stage);
  h->drop_graphs();
  F(h->d_stage2);
  F(h->d_hist);
  F(h->d_probs);
  for (auto& r : h->prof_open) {
    if (r.e0) cudaEventDestroy(r.e0);
    if (r.e1) cudaEventDestroy(r.e1);
  }
  for (auto e : h->prof_pool) cudaEventDestroy(e);
  if (h->prof_agg) {
    if (h->profile == 2) nvllm::trace_agg_set(nullptr, 0);
    cudaFree(h->prof_agg);
  }
  if (h->h_stage) cudaFreeHost(h->h_stage);
  if (h->h_out) cudaFreeHost(h->h_out);
  if (h->h_logits) cudaFreeHost(h->h_logits);
  nvllm::free_workspace(h);
  nvllm::tp_destroy(h);
  if (h->ev0) cudaEventDestroy(h->ev0);
  if (h->ev1) cudaEventDestroy(h->ev1);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
  return NVLLM_OK;
}

int nvllm_get_config(const nvllm_handle* h, nvllm_config* out) {
  if (!h || !out) return NVLLM_ERR_INVALID;
  *out = h->cfg;
  return NVLLM_OK;
}

int nvllm_memory_info(const nvllm_handle* h, uint64_t* weight_bytes, uint64_t* kv_bytes,
                      uint64_t* workspace_bytes) {
  if (!h) return NVLLM_ERR_INVALID;
  if (weight_bytes) *weight_bytes = h->weight_bytes;
  if (kv_bytes) *kv_bytes = h->kv_bytes;
  if (workspace_bytes) *workspace_bytes = h->ws_bytes;
  return NVLLM_OK;
}

int nvllm_set_sampling(nvllm_handle* h, float top_p, int top_k, float repetition_penalty) {
  if (!h) return NVLLM_ERR_INVALID;
  if (!(top_p > 0.f) || top_k < 0 || !(repetition_penalty > 0.f))
    return set_err(h, NVLLM_ERR_INVALID, "sampling: need top_p > 0, top_k >= 0, repetition_penalty > 0");
  h->top_p = top_p;
  h->top_k = top_k;
  h->rep_penalty = repetition_penalty;
  return NVLLM_OK;
}
int nvllm_set_seed(nvllm_handle* h, uint64_t seed) {
  if (!h) return NVLLM_ERR_INVALID;
  h->seed = seed;
  h->rng_counter = 0;
  return NVLLM_OK;
}
int nvllm_set_graphs(nvllm_handle* h, int enabled) {
  if (!h) return NVLLM_ERR_INVALID;
  h->graphs_enabled = enabled;
  return NVLLM_OK;
}

int nvllm_set_speculation(nvllm_handle* h, int enabled) {
  if (!h) return NVLLM_ERR_INVALID;
  h->spec_enabled = (enabled && h->n_mamba == 0) ? 1 : 0;  // see nvllm_set_layer_types
  return NVLLM_OK;
}

int nvllm_tp_unique_id(uint8_t out_id[128]) {
  if (!out_id) return set_err(nullptr, NVLLM_ERR_INVALID, "null argument");
  return tp_unique_id(out_id);
}

}  // extern "C"
