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
  // Mamba2 block (mamba.cu): in_proj [P, H] / out_proj [H, E] in the weight dtype, the rest fp32
  void *m_in_proj = nullptr, *m_out_proj = nullptr;
  float *m_conv_w = nullptr, *m_conv_b = nullptr, *m_dt_bias = nullptr, *m_a_log = nullptr,
        *m_d = nullptr, *m_norm = nullptr;
};

// one engine step's integer inputs, packed in a single pinned host buffer and copied with
// one cudaMemcpyAsync; the device pointers below point into the device mirror
struct StepLayout {
  int T = 0, n = 0, max_blocks = 0, n_qtiles = 0;
  bool prefill = false;
  int max_len = 0;        // longest context (incl. new tokens) in the step
  long long sum_len = 0;  // sum of contexts (attention bytes accounting)
  double attn_pairs = 0;  // sum over new tokens of (position + 1) (attention flops accounting)
  int32_t* d = nullptr;   // device buffer the offsets below refer to
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
  float* m_xbc = nullptr;    // [T, conv_dim] Mamba2: conv + SiLU output (x | B | C)
  float* m_dt = nullptr;     // [T, heads]    softplus(dt + bias)
  float* m_y = nullptr;      // [T, E] fp32 scan output
  void* m_yn = nullptr;      // [T, E] gated + normed, act dtype (out_proj input)
  // folded RMSNorm (forward.cu): bf16(x * g) of the residual stream (g = the next norm's weight), the
  // per-column-tile row sums of squares [parts][T_cap] of it
  void* xb = nullptr;
  float* ssq = nullptr;
  unsigned* chain_ctr = nullptr;  // [2] counters of the chained W1 -> W2 GEMM launch (zero between launches)
  float* logits_p = nullptr; // [S][n][ldl]
  float* logits = nullptr;   // [n][ldl]  (tp > 1: [tp][n][Vl], rank-major slices)
  float* tp_buf = nullptr;   // [T, H] fp32 partial sums handed to the all-reduce (tp > 1)
  const float* logits_final = nullptr;  // where the last step's summed logits live
  int n_cap_out = 0;
  size_t logits_p_floats = 0;
  float* attn_ws = nullptr;
  size_t attn_ws_floats = 0;
  int* out_tokens = nullptr;  // [n]
  unsigned long long* amax = nullptr;  // [n] argmax merge scratch (zero between kernels)
  int* amax_ticket = nullptr;          // [n]
  // moe
  int* topk_idx = nullptr;
  float* topk_w = nullptr;
  MoeSortBufs sort{};
  int max_tiles = 0;
  float* moe_Y = nullptr;  // [T k, H]
  float* moe_y = nullptr;  // [T, H]
  void* moe_xs = nullptr;  // [T k, H] bf16: normed activations in expert-sorted order
};

}  // namespace nvllm

struct nvllm_handle {
  nvllm_config cfg{};
  int dev = 0, sm_count = 148;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  cudaEvent_t tm0 = nullptr, tm1 = nullptr;  // nvllm_timer_record
  std::string err;
  bool finalized = false;
  int wdt = 0, adt = 0, kvdt = 0;
  // dims
  int H = 0, V = 0, L = 0, nh = 0, nkv = 0, hd = 0, F = 0, E = 0, topk = 0;
  int q_cols = 0, kv_cols = 0, qkv_cols = 0, w1_cols = 0, ldl = 0;
  bool swiglu = false, moe = false, parallel = false, use_rope = false;
  float attn_scale = 0.f;
  // tensor parallel (tp.cu): nh / nkv / F / q_cols / ... above are this rank's LOCAL extents;
  // *_g are the model's.  Heads are padded to a multiple of tp (Falcon 71 -> 72, zero weights).
  int tp = 1, rank = 0;
  int nh_g = 0, nkv_g = 0, F_g = 0;
  int Vl = 0;        // LM-head rows per rank (slot width of the gathered logits), multiple of 8
  int Vl_valid = 0;  // rows of this rank's slice that exist (last rank may be short)
  void* nccl = nullptr;  // ncclComm_t
  // two-chunk prefill pipeline of a tensor-parallel step (forward.cu): the second chunk runs on
  // stream2 with its own communicator, so its GEMMs hide the first chunk's all-reduces
  void* nccl2 = nullptr;
  cudaStream_t stream2 = nullptr;
  cudaEvent_t pipe_fork = nullptr, pipe_join = nullptr;
  std::vector<cudaEvent_t> pipe_kv;  // [L]: the first chunk's K / V rows of layer l are in the pool
  int pipe_role = 0;       // 0 = plain step, 1 = first chunk (records pipe_kv), 2 = second (waits)
  int view_r0 = 0;         // first row of the current chunk within the step
  int view_T = 0;          // rows of the whole step (0 = no chunking)
  bool skip_head = false;  // the chunk holds no sequence end: no final norm / LM head / sampling
  int gemm_max_ctas = 0;   // > 0: persistent GEMM grids leave SMs to the collective kernels
  void* nccl_cur() const { return (pipe_role == 2 && nccl2) ? nccl2 : nccl; }
  // NVLink peer-memory all-reduce (tp.cu): this rank's region, the peers' mapped regions
  float* p2p_peer[8] = {nullptr};  // [rank] base pointers (own entry = p2p_own)
  float* p2p_own = nullptr;
  size_t p2p_cap = 0;              // floats per data buffer
  uint32_t p2p_epoch = 0;
  bool p2p_ok = false;
  // Granite-4 hybrid: Mamba2 layers (nvllm_set_layer_types).  SSM state is kept PER KV BLOCK:
  // ssm_pool[block][mamba layer][heads][head_dim][state] holds the state after the last token
  // written into that block, so it is addressed through the block table like K / V — no
  // per-sequence bookkeeping at the boundary, and a prefix-cache hit finds the state of the
  // cached prefix in its last block.
  std::vector<uint8_t> is_mamba;  // [L]
  std::vector<int> mamba_idx;     // [L] index among the Mamba2 layers, -1 for attention layers
  int n_mamba = 0;
  int m_nh = 0, m_hd = 0, m_N = 0, m_G = 0, m_K = 0, m_E = 0, m_conv = 0, m_P = 0;
  float* ssm_pool = nullptr;
  // weights
  void *wte = nullptr, *wpe = nullptr, *lm_head = nullptr;
  bool lm_head_owned = false;
  float *lnf_w = nullptr, *lnf_b = nullptr;
  std::vector<nvllm::LayerW> layers;
  std::vector<uint8_t> have;  // [L+1][64] flags, index (layer+1)*64 + id
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
  // SampleWithHistory filters (sampling.cu): the step's flattened token history
  // [cu_hist (n + 1) | tokens] and the per-row probability scratch, grown on demand
  int32_t* d_hist = nullptr;
  size_t hist_words = 0;
  int hist_n = 0;           // sequences of the uploaded history (0 = none for this step)
  float* d_probs = nullptr;
  size_t probs_floats = 0;
  bool filters_active() const { return rep_penalty != 1.f || top_k > 0 || top_p < 1.f; }
  uint64_t seed = 0, rng_counter = 0;
  int graphs_enabled = 1;
  int use_tc5 = 1;
  // RMSNorm folded into the GEMMs that consume it: whoever produces x also writes bf16(x * g),
  // the row statistic is applied to the accumulators (dense RMSNorm models, bf16, tp == 1;
  // NVLLM_FOLD=0 switches it off)
  bool fold = false;
  // accounting
  uint64_t weight_bytes = 0, kv_bytes = 0, ws_bytes = 0;
  int launches = 0;  // kernels launched by the current step
  int cur_cls = 0;            // kernel class of the launch being issued
  unsigned skip_mask = 0;     // NVLLM_SKIP: classes not launched (timing ablation only)
  bool resident_ok = false;  // the last step's state is on the device (nvllm_run_resident)
  nvllm::StepLayout run_lay;  // layout of the last nvllm_run (nvllm_rerun replays it)
  bool run_lay_ok = false;
  // CUDA graph of one device-resident decode step (advance + forward pass), replayed while
  // the launch shape stays the same (run.cu)
  // slot 0: nvllm_run_resident (advance + forward pass), slot 1: nvllm_run decode step
  cudaGraphExec_t step_graph[2] = {nullptr, nullptr};
  struct GraphSig {
    const int32_t* d = nullptr;
    int n = 0, max_blocks = 0, n_splits = 0;
    bool operator==(const GraphSig& o) const {
      return d == o.d && n == o.n && max_blocks == o.max_blocks && n_splits == o.n_splits;
    }
  } graph_sig[2];
  int graph_launches[2] = {0, 0};
  // where the captured step leaves its summed logits (forward_step picks the buffer per launch
  // shape; a replay does not run the host code that records it)
  const float* graph_logits[2] = {nullptr, nullptr};
  void drop_graphs() {
    for (auto& g : step_graph) {
      if (g) cudaGraphExecDestroy(g);
      g = nullptr;
    }
  }
  // Speculative next decode step (run.cu): issued at the end of nvllm_run so that the GPU is
  // not idle while the host turns the sampled tokens around; consumed by the next nvllm_run /
  // nvllm_run_resident when it asks for exactly that step, overwritten otherwise
  int spec_enabled = 1;
  bool spec_pending = false;
  bool spec_used_graph = false;  // the in-flight speculative step replayed a graph
  int spec_skip = 0;  // calls that do not speculate (set after a miss)
  std::vector<int32_t> spec_len, spec_nb, spec_bt;  // per-sequence lengths / block tables of the last call
  // device-resident decode state as the host knows it: context length (incl. the token the last
  // step appended) and block-table extent of every sequence — nvllm_run_resident refuses to step
  // a sequence past its own table (the padded entries of a ragged batch are not its blocks)
  std::vector<int32_t> res_len, res_nb;
  int32_t* d_stage2 = nullptr;  // decode layout derived on the device from a prefill step
  size_t stage2_words = 0;
  // per-kernel-class profile (nvllm_set_profile / nvllm_get_profile)
  struct ProfRec {
    int cls;
    double bytes, flops;
    cudaEvent_t e0, e1;
    double impl;
    int seq0, seq1;  // timeline mode: the traced launches issued inside the scope
  };
  int profile = 0;  // 0 off, 1 events around every launch, 2 device timeline (forward.cu)
  unsigned long long* prof_agg = nullptr;
  std::vector<ProfRec> prof_open;
  std::vector<cudaEvent_t> prof_pool;
  nvllm_profile_entry prof[NVLLM_K_COUNT] = {};
};

namespace nvllm {

extern thread_local std::string g_create_error;

void free_workspace(nvllm_handle* h);
int ensure_workspace(nvllm_handle* h, int T, int n);
int forward_step(nvllm_handle* h);  // runs h->lay
int forward_step_auto(nvllm_handle* h);  // forward_step, or the two-chunk tensor-parallel prefill pipeline
int prof_collect(nvllm_handle* h);

int set_err(nvllm_handle* h, int code, const char* fmt, ...);

// weights.cu: how the caller's buffer holds a tensor (nvllm_set_weight: all defaults)
struct WeightSrc {
  bool out_in = false;  // linears are [out, in] (PyTorch / safetensors) instead of [in, out]
  size_t ld = 0;        // > 0: the tensor is columns [col0, col0 + width) of a [rows, ld] matrix
  size_t col0 = 0;      //      (vectors: elements [col0, col0 + n) of a longer vector of ld)
};
// internal weight ids (never cross the C ABI): the two halves of a SwiGLU W1, each [H, F]
enum { W_INT_FFN_GATE = 60, W_INT_FFN_UP = 61 };
int set_weight_src(nvllm_handle* h, int id, int layer, int dtype, const void* host, uint64_t numel,
                   const WeightSrc& src);

// tp.cu — tensor-parallel group over NCCL (loaded with dlopen: no link-time dependency)
int tp_unique_id(uint8_t out_id[128]);
int tp_init(nvllm_handle* h, const uint8_t* id);  // ncclCommInitRank(tp, id, rank)
void tp_destroy(nvllm_handle* h);
int tp_all_reduce(nvllm_handle* h, void* buf, size_t count, bool bf16_wire);  // in place, sum
// peer-memory path: true when the fused partial-sum + all-reduce kernel can take T x H floats
bool tp_p2p_fits(const nvllm_handle* h, size_t count);
cudaError_t launch_tp_allreduce_p2p(nvllm_handle* h, int T, int H, const float* p1, int S1,
                                    size_t stride1, float m1, const float* bias1, const float* p2,
                                    int S2, size_t stride2, float m2, const float* bias2,
                                    float* out);
int tp_all_gather_f32(nvllm_handle* h, float* buf, size_t count_per_rank);  // in place, slot = rank

#define NV_CHECK(h, expr)                                                                   \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess)                                                                  \
      return set_err(h, NVLLM_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                     __FILE__, __LINE__);                                                   \
  } while (0)

}  // namespace nvllm
