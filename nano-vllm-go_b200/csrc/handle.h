// handle.h — the library handle behind `nvllm_handle*` (include/nvllm.h) and the
// host-side helpers shared by nvllm.cu (lifecycle + forward pass) and ops.cu.
#pragma once
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>

#include <string>
#include <vector>

#include "../../include/nvllm.h"
#include "kernels.h"

namespace nvllm {

struct LayerW {
  float *attn_norm_w = nullptr, *attn_norm_b = nullptr, *ffn_norm_w = nullptr, *ffn_norm_b = nullptr;
  void* wqkv = nullptr;   // [(nh + 2 nkv) hd, H]  rows = Q heads | K heads | V heads
  float* bqkv = nullptr;  // [(nh + 2 nkv) hd] or null
  void* wo = nullptr;     // [H, nh hd]
  float* bo = nullptr;
  void* w1 = nullptr;     // [F or 2F, H]
  float* b1 = nullptr;
  void* w2 = nullptr;     // [H, F]
  float* b2 = nullptr;
  float* router = nullptr;  // [E, H] fp32
  void* moe_in = nullptr;   // [E, 2I, H]
  void* moe_out = nullptr;  // [E, H, I]
  bool has_bq = false, has_bk = false, has_bv = false;
  TmaDesc tm_wqkv, tm_wo, tm_w1, tm_w2;
};

// one engine step's integer inputs, packed in a single pinned host buffer and copied with
// one cudaMemcpyAsync; the device pointers below point into the device mirror
struct StepLayout {
  int T = 0, n = 0, max_blocks = 0, n_qtiles = 0;
  size_t words = 0;  // 4-byte words in use
  // offsets (in words) of each array
  size_t tokens, positions, slots, cu_q, seq_start, seq_len, last_row, block_tables, tile_seq,
      tile_q0, temperature, uniforms;
};

struct Workspace {
  int T_cap = 0, n_cap = 0;
  float* x = nullptr;        // [T, H] fp32 residual stream
  void* xn = nullptr;        // [T, H] act dtype
  float* xn_f32 = nullptr;   // [T, H] (MoE router input)
  float *ws_a = nullptr, *ws_b = nullptr, *ws_c = nullptr;
  size_t ws_a_floats = 0, ws_b_floats = 0;
  void* q = nullptr;         // [T, nh hd]
  void* attn_out = nullptr;  // [T, nh hd]
  void* hbuf = nullptr;      // [T, F] or [T k, I]
  void* xn_last = nullptr;   // [n, H]
  float* logits_p = nullptr; // [S][n][ldl]
  float* logits = nullptr;   // [n][ldl]
  const float* logits_final = nullptr;  // where the last step's summed logits live
  int n_cap_out = 0;
  size_t logits_p_floats = 0;
  float* attn_ws = nullptr;
  size_t attn_ws_floats = 0;
  int* out_tokens = nullptr;  // [n]
  // moe
  int* topk_idx = nullptr;
  float* topk_w = nullptr;
  MoeSortBufs sort{};
  int max_tiles = 0;
  float* moe_Y = nullptr;  // [T k, H]
  float* moe_y = nullptr;  // [T, H]
  // TMA maps of the activation buffers (rebuilt when the workspace grows)
  TmaDesc tm_xn, tm_attn_out, tm_hbuf, tm_xn_last;
  int tm_box_rows = 0;
};

}  // namespace nvllm

struct nvllm_handle {
  nvllm_config cfg{};
  int dev = 0, sm_count = 148;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  std::string err;
  bool finalized = false;
  int wdt = 0, adt = 0, kvdt = 0;
  // dims
  int H = 0, V = 0, L = 0, nh = 0, nkv = 0, hd = 0, F = 0, E = 0, topk = 0;
  int q_cols = 0, kv_cols = 0, qkv_cols = 0, w1_cols = 0, ldl = 0;
  bool swiglu = false, moe = false, parallel = false, use_rope = false;
  float attn_scale = 0.f;
  // weights
  void *wte = nullptr, *wpe = nullptr, *lm_head = nullptr;
  bool lm_head_owned = false;
  float *lnf_w = nullptr, *lnf_b = nullptr;
  std::vector<nvllm::LayerW> layers;
  std::vector<uint8_t> have;  // [L+1][64] flags, index (layer+1)*64 + id
  nvllm::TmaDesc tm_lm_head;
  float *cosT = nullptr, *sinT = nullptr;
  // paged KV pool: K and V each [L][num_blocks][nkv][256][hd]
  void *kpool = nullptr, *vpool = nullptr;
  size_t kv_layer_elems = 0;
  int num_blocks = 0, max_blocks_per_seq = 0;
  // step staging
  int32_t* h_stage = nullptr;  // pinned
  int32_t* d_stage = nullptr;
  size_t stage_words = 0;
  int32_t* h_out = nullptr;  // pinned [n_cap]
  float* h_logits = nullptr;  // pinned, grown on demand
  size_t h_logits_floats = 0;
  nvllm::StepLayout lay;
  nvllm::Workspace ws;
  // sampling
  float top_p = 1.f, rep_penalty = 1.f;
  int top_k = 0;
  uint64_t seed = 0, rng_counter = 0;
  int graphs_enabled = 1;
  int use_tc5 = 1;
  // accounting
  uint64_t weight_bytes = 0, kv_bytes = 0, ws_bytes = 0;
  int launches = 0;  // kernels launched by the current step
  bool resident_ok = false;  // a decode step's state is on the device (nvllm_run_resident)
  int resident_max_len = 0;
};

namespace nvllm {

extern thread_local std::string g_create_error;

void free_workspace(nvllm_handle* h);
int build_weight_maps(nvllm_handle* h);      // gemm_tcgen05.cu: TMA maps of the packed weights
int build_activation_maps(nvllm_handle* h);  // gemm_tcgen05.cu: TMA maps of the workspace

int set_err(nvllm_handle* h, int code, const char* fmt, ...);

#define NV_CHECK(h, expr)                                                                   \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess)                                                                  \
      return set_err(h, NVLLM_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                     __FILE__, __LINE__);                                                   \
  } while (0)

}  // namespace nvllm
