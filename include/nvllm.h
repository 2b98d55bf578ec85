/*
 * nvllm.h — C ABI of libnvllm_b200.so, the B200 (sm_100a) forward-pass backend
 * for nano-vllm-go.
 *
 * This is the drop-in boundary: everything below `ModelRunner.Run`
 * (reference nanovllm/model_runner.go:9-16, today implemented by
 * nanovllm/tensor_model_runner.go:55-97 ->
 * purego/tensor/generic_model.go:276-480) is replaced by the calls declared
 * here.  Plain pointers and sizes only; no torch / C++ types.  A cgo shim
 * (go/cuda_model_runner.go) or the ctypes mirror in
 * nano-vllm-go_b200/runner.py binds exactly these symbols.
 *
 * Rules of the boundary:
 *  - the caller owns every host buffer; the library copies what it needs
 *    before returning and retains no caller pointer (cgo rule);
 *  - the library owns device memory, streams, CUDA graphs, NCCL comms;
 *  - a handle is NOT re-entrant (one Run at a time, like LLMEngine.Step,
 *    reference nanovllm/llm_engine.go:62-70); one handle per GPU;
 *  - every entry point returns an int status (0 = ok); the message is
 *    available from nvllm_last_error(); nothing aborts the process;
 *  - there is no CPU fallback: without a usable sm_100 device nvllm_create
 *    fails with NVLLM_ERR_CUDA.
 */
#ifndef NVLLM_H_
#define NVLLM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NVLLM_ABI_VERSION 1

/* KV block size is hard-wired to 256 tokens in the reference
 * (nanovllm/sequence.go:52, nanovllm/config.go:57). */
#define NVLLM_KV_BLOCK_SIZE 256

enum nvllm_status {
  NVLLM_OK = 0,
  NVLLM_ERR_INVALID = 1,   /* bad argument / shape / missing weight          */
  NVLLM_ERR_CUDA = 2,      /* CUDA runtime / driver / no sm_100 device       */
  NVLLM_ERR_STATE = 3,     /* call order (e.g. run before finalize)          */
  NVLLM_ERR_RANGE = 4,     /* position >= max_seq_len, block id out of pool  */
  NVLLM_ERR_NCCL = 5,
  NVLLM_ERR_NOMEM = 6
};

/* purego/tensor/config.go:14-55 */
enum nvllm_attention_type { NVLLM_ATTN_MHA = 0, NVLLM_ATTN_GQA = 1, NVLLM_ATTN_MQA = 2 };
enum nvllm_norm_type { NVLLM_NORM_LAYER = 0, NVLLM_NORM_RMS = 1 };
enum nvllm_position_type { NVLLM_POS_LEARNED = 0, NVLLM_POS_ROPE = 1, NVLLM_POS_NONE = 2 };
enum nvllm_activation_type { NVLLM_ACT_GELU = 0, NVLLM_ACT_SWIGLU = 1 };
enum nvllm_block_style { NVLLM_BLOCK_SEQUENTIAL = 0, NVLLM_BLOCK_PARALLEL = 1 };

/* Arithmetic mode of the whole forward pass (BASELINE.json north_star):
 *  FP32: fp32 weights / activations / KV, fp32 FMA everywhere. Parity gate
 *        1e-4 relative on logits against the reference's fp32 path.
 *  BF16: bf16 weights and KV cache, fp32 accumulate, fp32 residual stream;
 *        tcgen05 tensor-core GEMMs for prefill and batched decode. Gate 2e-2. */
enum nvllm_precision { NVLLM_PREC_FP32 = 0, NVLLM_PREC_BF16 = 1 };

/* dtype of a host weight buffer handed to nvllm_set_weight
 * (the reference loader accepts the same three: generic_loader.go:645-663). */
enum nvllm_dtype { NVLLM_DT_F32 = 0, NVLLM_DT_BF16 = 1, NVLLM_DT_F16 = 2 };

/* Mirrors the fields of tensor.ModelConfig that affect the math
 * (purego/tensor/config.go:58-122) plus engine-side pool sizing
 * (nanovllm/config.go:9-20, nanovllm/scheduler.go:17-21). */
typedef struct nvllm_config {
  int32_t abi_version;        /* = NVLLM_ABI_VERSION */
  int32_t vocab_size;
  int32_t hidden;
  int32_t num_layers;
  int32_t num_heads;
  int32_t num_kv_heads;       /* 1 for MQA, = num_heads for MHA */
  int32_t head_dim;
  int32_t ffn_dim;            /* expert intermediate size when use_moe */
  int32_t max_seq_len;        /* RoPE table / wpe extent; pos >= this is an error (rope.go:84) */
  int32_t attention_type;     /* nvllm_attention_type */
  int32_t norm_type;          /* nvllm_norm_type */
  int32_t position_type;      /* nvllm_position_type */
  int32_t activation_type;    /* nvllm_activation_type */
  int32_t block_style;        /* nvllm_block_style */
  int32_t tied_embedding;     /* LM head = token embedding^T (generic_loader.go:255-259) */
  int32_t use_moe;
  int32_t num_experts;
  int32_t num_experts_per_tok;
  float   norm_eps;
  float   embedding_multiplier;  /* 0 = unset (generic_model.go:298) */
  float   attention_multiplier;  /* 0 = 1/sqrt(head_dim) (attention.go:361-364) */
  float   residual_multiplier;   /* 0 = plain residual add (generic_model.go:359) */
  float   logits_scaling;        /* 0 = unset; logits are DIVIDED by it (generic_model.go:473) */
  double  rope_base;             /* GQA only; MQA is hard-wired 10000 (mqa.go:35) */
  /* engine / pool */
  int32_t precision;          /* nvllm_precision */
  int32_t num_kv_blocks;      /* Config.NumKVCacheBlocks; <=0 -> 1024 (scheduler.go:17-21) */
  int32_t max_num_seqs;       /* Config.MaxNumSeqs (512) */
  int32_t max_batched_tokens; /* Config.MaxNumBatchedTokens (16384) */
  /* tensor parallel (Falcon): this handle is rank tp_rank of tp_size */
  int32_t tp_size;            /* 1 = no TP */
  int32_t tp_rank;
  int32_t device;             /* CUDA device ordinal for this handle */
  /* Mamba2 layers of the Granite-4 hybrid (purego/tensor/mamba2.go:30-66, config.go:100-115);
   * all 0 = the model has none.  expand * hidden = mamba_num_heads * mamba_head_dim. */
  int32_t mamba_num_heads;
  int32_t mamba_head_dim;
  int32_t mamba_state_size;
  int32_t mamba_n_groups;
  int32_t mamba_conv_kernel;
} nvllm_config;

/* Weight identifiers. Layouts are the reference's POST-LOAD layouts
 * (what TransformerModel holds after generic_loader.go:184-265,353-604):
 * every linear is [in, out] row-major.  `layer` is ignored (pass -1) for the
 * model-level ids. */
enum nvllm_weight_id {
  NVLLM_W_TOKEN_EMBEDDING = 0, /* [V, H]           generic_model.go:8 */
  NVLLM_W_POS_EMBEDDING = 1,   /* [max_pos, H]     learned positions only */
  NVLLM_W_FINAL_NORM_W = 2,    /* [H] */
  NVLLM_W_FINAL_NORM_B = 3,    /* [H] optional (absent => RMSNorm, tensor.go:197) */
  NVLLM_W_LM_HEAD = 4,         /* [H, V]; omit when tied_embedding */
  /* per layer */
  NVLLM_W_ATTN_NORM_W = 16,    /* AttnLN (sequential) or InputLN (parallel) */
  NVLLM_W_ATTN_NORM_B = 17,
  NVLLM_W_FFN_NORM_W = 18,     /* FFNLN (sequential blocks only) */
  NVLLM_W_FFN_NORM_B = 19,
  NVLLM_W_Q = 20,              /* [H, n_heads*hd] */
  NVLLM_W_K = 21,              /* [H, n_kv*hd]          MHA / GQA */
  NVLLM_W_V = 22,              /* [H, n_kv*hd]          MHA / GQA */
  NVLLM_W_KV = 23,             /* [H, 2*hd] = [K | V]   MQA (mqa.go:15, generic_loader.go:750-765) */
  NVLLM_W_O = 24,              /* [n_heads*hd, H] */
  NVLLM_W_Q_BIAS = 25,         /* optional */
  NVLLM_W_K_BIAS = 26,
  NVLLM_W_V_BIAS = 27,
  NVLLM_W_O_BIAS = 28,
  NVLLM_W_FFN_W1 = 29,         /* [H, F] (GELU) or [H, 2F] = [gate | up] (SwiGLU, generic_loader.go:531-536) */
  NVLLM_W_FFN_B1 = 30,         /* optional, GELU branch only (transformer.go:67-75) */
  NVLLM_W_FFN_W2 = 31,         /* [F, H] */
  NVLLM_W_FFN_B2 = 32,         /* optional */
  NVLLM_W_MOE_ROUTER = 33,     /* [H, E]            (generic_loader.go:576) */
  NVLLM_W_MOE_INPUT = 34,      /* [E, 2I, H] HF layout, un-transposed (generic_loader.go:578-586) */
  NVLLM_W_MOE_OUTPUT = 35,     /* [E, H, I]  HF layout */
  /* Mamba2 layers (Mamba2Layer, mamba2.go:9-28; loaded un-transposed, generic_loader.go:461-498).
   * E = mamba_num_heads * mamba_head_dim, conv_dim = E + 2 * n_groups * state_size. */
  NVLLM_W_MAMBA_IN_PROJ = 36,  /* [E + conv_dim + heads, H]  rows = [gate | xBC | dt] (mamba2.go:80-98) */
  NVLLM_W_MAMBA_CONV_W = 37,   /* [conv_dim, 1, conv_kernel] */
  NVLLM_W_MAMBA_CONV_B = 38,   /* [conv_dim] */
  NVLLM_W_MAMBA_DT_BIAS = 39,  /* [heads] */
  NVLLM_W_MAMBA_A_LOG = 40,    /* [heads] */
  NVLLM_W_MAMBA_D = 41,        /* [heads] */
  NVLLM_W_MAMBA_NORM = 42,     /* [E]  RMSNormGated weight */
  NVLLM_W_MAMBA_OUT_PROJ = 43  /* [H, E] */
};

typedef struct nvllm_handle nvllm_handle;

/* Per-call counters filled by nvllm_run (all optional). */
typedef struct nvllm_run_stats {
  float    device_ms;        /* CUDA-event time of the forward+sample on the compute stream */
  int32_t  kernel_launches;  /* kernels of this library launched (graph nodes count) */
  int32_t  new_tokens;       /* tokens actually computed (prefill skips cached prefix) */
  int32_t  used_graph;       /* 1 if the step replayed a CUDA graph */
  uint64_t h2d_bytes;
  uint64_t d2h_bytes;
} nvllm_run_stats;

/* ---- lifetime ----------------------------------------------------------- */

int nvllm_abi_version(void);

/* 128-byte NCCL unique id for a TP group; rank 0 generates, the host code
 * distributes it (Go: in-process; python: torch.distributed / file). */
int nvllm_tp_unique_id(uint8_t out_id[128]);

/* Create a handle on cfg->device.  tp_id may be NULL when tp_size == 1.
 * Replaces tensor.LoadModelFromDirectory inside NewTensorModelRunner
 * (tensor_model_runner.go:21-32) together with set_weight/finalize. */
int nvllm_create(const nvllm_config* cfg, const uint8_t* tp_id, nvllm_handle** out);

/* Hybrid models (ModelConfig.HybridLayers, config.go:113): types[l] = 1 makes layer l a Mamba2
 * block (norm -> Mamba2 -> residual, norm -> FFN -> residual; generic_model.go:162-202), 0 an
 * attention block.  Call before the first nvllm_set_weight; needs the mamba_* config fields. */
int nvllm_set_layer_types(nvllm_handle* h, const int32_t* types, int n_layers);

/* Hand one tensor in the reference's post-load layout; the library converts /
 * repacks / TP-shards on its own and keeps no pointer.  numel is checked. */
int nvllm_set_weight(nvllm_handle* h, int weight_id, int layer, int dtype,
                     const void* host_data, uint64_t numel);

/* Checkpoint families the reference maps (purego/tensor/generic_loader.go:60-182). */
enum nvllm_arch {
  NVLLM_ARCH_GPT2 = 0,        /* GetGPT2Mapping: Conv1D [in,out] tensors, fused c_attn */
  NVLLM_ARCH_FALCON = 1,      /* GetFalconMapping: fused query_key_value, parallel block */
  NVLLM_ARCH_LLAMA = 2,       /* GetLlamaMapping: separate q/k/v, gate_proj + up_proj */
  NVLLM_ARCH_GRANITE_MOE = 3  /* GetGraniteMoEMapping: router + stacked expert tensors */
};

/* Load every weight of the model straight from `model_dir` — model.safetensors, or the shards
 * named by model.safetensors.index.json — instead of one nvllm_set_weight call per tensor.
 * Replaces the weight half of tensor.LoadModelFromDirectory (generic_loader.go:1016-1163,
 * 184-265): the files are mmap'ed, tensors go to the device in their file dtype (F32 / F16 /
 * BF16; nothing is widened to fp32 on the host) and the loader's transposes, the GPT-2 /
 * Falcon fused-QKV splits and the [gate | up] concatenation happen in the device-side repack.
 * The handle's config (created from the same config.json the Go side parses with
 * LoadModelConfig) says what to expect; call nvllm_finalize afterwards.  Tensor-parallel
 * handles each read the full files and keep their shard.  bytes_read / tensors_loaded are
 * optional. */
int nvllm_load_safetensors(nvllm_handle* h, const char* model_dir, int arch,
                           uint64_t* bytes_read, int32_t* tensors_loaded);

/* Host-only checkpoint probe (no handle, no device): finds `name` the way the loader does
 * ("transformer." retry, sharded index) and reports dtype (nvllm_dtype, -1 = unsupported),
 * shape, byte size and the tensor's first 8 bytes.  NVLLM_ERR_INVALID when absent; the reason
 * is in nvllm_last_error(NULL). */
int nvllm_safetensors_find(const char* model_dir, const char* name, int32_t* dtype,
                           int32_t* ndim, int64_t shape[8], uint64_t* nbytes,
                           uint64_t* first_bytes);

/* Pack weights (bf16 [out,in] K-major for TMA), build tensor maps, allocate
 * the paged KV pool [L][2][num_blocks][n_kv][256][hd] and workspaces. */
int nvllm_finalize(nvllm_handle* h);

/* ⇔ ModelRunner.Close (model_runner.go:15). */
int nvllm_destroy(nvllm_handle* h);

const char* nvllm_last_error(const nvllm_handle* h); /* h may be NULL: create errors */

/* ---- the hot path: ModelRunner.Run -------------------------------------- */

/*
 * One engine step for n_seqs scheduled sequences (uniformly prefill or decode,
 * scheduler.go:69-71,111).  Flattened Sequence records (sequence.go:15-28):
 *   token_ids   all tokens of every sequence back to back (TokenIDs, incl. the
 *               last sampled one); cu_seqlens[n_seqs+1] delimits them
 *   num_cached  NumCachedTokens per sequence (multiple of 256; prefill only,
 *               may be NULL => 0).  Prefill computes tokens [num_cached, len)
 *               and always at least the last one.
 *   block_tables / cu_blocks[n_seqs+1]   BlockTable per sequence; block i
 *               holds positions [256 i, 256 i + 256)
 *   temperature per sequence; <= 0 => greedy argmax with the first-max rule of
 *               cmd/ask/main.go:389-402; NULL => all greedy
 *   uniforms    optional per-sequence u in [0,1) replacing rand.Float32()
 *               (sampling.go:205) so temperature sampling is reproducible;
 *               NULL => library Philox stream seeded by nvllm_set_seed
 * Outputs:
 *   out_tokens[n_seqs]          next token per sequence, same order
 *   out_logits[n_seqs * V]      optional (NULL to skip): last-token logits
 *                               (GetLogitsForLastToken, generic_model.go:595)
 * Decode = the last token of each sequence at position len-1
 * (tensor_model_runner.go:78-80).
 */
int nvllm_run(nvllm_handle* h, int n_seqs, int is_prefill,
              const int32_t* token_ids, const int32_t* cu_seqlens,
              const int32_t* num_cached,
              const int32_t* block_tables, const int32_t* cu_blocks,
              const float* temperature, const float* uniforms,
              int32_t* out_tokens, float* out_logits,
              nvllm_run_stats* stats);

/* Device-resident variant for throughput measurement: identical work, but the
 * step's inputs are taken from the previous nvllm_run's device staging
 * buffers (no host<->device copies, no host sync); out_tokens stay on device.
 * Used by bench.py for the `value` leg; nvllm_run is the `e2e` leg. */
int nvllm_run_resident(nvllm_handle* h, nvllm_run_stats* stats);

/* Re-execute the last nvllm_run step from its device-resident staged inputs (no copies, no
 * host sync).  With nvllm_run_resident this lets a whole prefill + decode job be repeated
 * with every input already in HBM. */
int nvllm_rerun(nvllm_handle* h, nvllm_run_stats* stats);

/* Per-kernel-class device timing.  mode 1: every launch is bracketed by CUDA events on the
 * compute stream (serialises the chain: an upper bound; the only mode that times the NCCL calls
 * of a tensor-parallel step).  mode 2: device timeline — every kernel records the first entry /
 * dependency release / last exit of its CTAs in %globaltimer ns while the launches stay back to
 * back with their programmatic-dependent-launch edges; a launch is charged the time from its
 * predecessor's last exit to its own, so the classes add up to the window's device time (CUDA
 * graph replay is off while profiling; up to 2^18 launches per window).  mode 0: off.
 * `bytes` / `flops` are the ALGORITHMIC figures of each launch, `impl_bytes` what this
 * implementation moves on top.  nvllm_set_profile resets the counters;
 * nvllm_get_profile synchronises and fills out[NVLLM_K_COUNT]. */
enum nvllm_kernel_class {
  NVLLM_K_EMBED = 0,
  NVLLM_K_NORM = 1,
  NVLLM_K_GEMM_TC5 = 2,     /* tcgen05/TMEM/TMA bf16 GEMM */
  NVLLM_K_GEMV = 3,         /* split-K GEMV, M <= 8 */
  NVLLM_K_GEMM_SIMT = 4,    /* fp32 parity-mode GEMM */
  NVLLM_K_ROPE_KV = 5,      /* RoPE + paged KV append */
  NVLLM_K_ATTN_PREFILL = 6,
  NVLLM_K_ATTN_DECODE = 7,
  NVLLM_K_ACT = 8,          /* SwiGLU / GELU */
  NVLLM_K_MOE_ROUTE = 9,    /* router + top-k + grouping */
  NVLLM_K_MOE_GEMM = 10,    /* grouped expert GEMMs */
  NVLLM_K_MOE_MISC = 11,    /* GLU + weighted combine */
  NVLLM_K_SAMPLE = 12,
  NVLLM_K_MISC = 13,
  NVLLM_K_COMM = 14,        /* tensor-parallel partial sum + NCCL all-reduce / all-gather */
  NVLLM_K_COUNT = 15
};
typedef struct nvllm_profile_entry {
  double ms;
  double bytes;       /* algorithmic bytes (weights, K/V, logits, one pass over activations) */
  double flops;
  int64_t launches;
  double impl_bytes;  /* extra bytes of this implementation (split-K partials, staging) */
} nvllm_profile_entry;
int nvllm_set_profile(nvllm_handle* h, int mode);
int nvllm_get_profile(nvllm_handle* h, nvllm_profile_entry* out);

/* Diagnostics: device timeline of the instrumented kernels (GEMM, norm, RoPE+KV, decode
 * attention).  nvllm_trace_begin arms a buffer of `capacity` records; every CTA of those
 * kernels then appends {launch seq, kernel id, cta, sm, t_entry, t_after_dependency_wait,
 * t_exit, extra} (8 x uint64, ns of %globaltimer).  nvllm_trace_end synchronises, disarms and
 * copies up to max_records out.  Process-global; not for production runs. */
int nvllm_trace_begin(nvllm_handle* h, int capacity);
int nvllm_trace_end(nvllm_handle* h, uint64_t* out, int max_records, int* n_records);

/* Wait for the resident steps issued so far; optionally read back the last sampled tokens
 * (n <= n_seqs of the last nvllm_run) and the device time of the last step. */
int nvllm_sync(nvllm_handle* h, int32_t* out_tokens, int n, float* last_step_ms);

/* Device timer on the handle's compute stream, for timing several calls as one region:
 * which = 0 records the start event, 1 the stop event; nvllm_timer_elapsed synchronises the
 * stream and returns stop - start in milliseconds. */
int nvllm_timer_record(nvllm_handle* h, int which);
int nvllm_timer_elapsed(nvllm_handle* h, float* ms);

/* SetSamplingParamsWithRepetition (tensor_model_runner.go:35-52):
 * repetition_penalty 1.0 disables; top_k 0 disables; top_p 1.0 disables.
 * With any of them enabled, rows with temperature > 0 are drawn by SampleWithHistory
 * (sampling.go:33-217: penalty over the sequence's token_ids with the x3 weight of the last
 * ten, temperature, softmax, top-k, top-p, renormalise, inverse CDF with the step's uniform);
 * rows with temperature <= 0 stay greedy.  The history comes from nvllm_run's token_ids, so
 * nvllm_run_resident returns NVLLM_ERR_STATE after such a step.  Invalid values
 * (top_p <= 0, top_k < 0, repetition_penalty <= 0) return NVLLM_ERR_INVALID. */
int nvllm_set_sampling(nvllm_handle* h, float top_p, int top_k, float repetition_penalty);
int nvllm_set_seed(nvllm_handle* h, uint64_t seed);

/* Enable/disable CUDA-graph replay of decode steps (default on). */
int nvllm_set_graphs(nvllm_handle* h, int enabled);

/* Speculative next decode step (default on; needs graphs on, greedy rows, tp_size 1): at the
 * end of nvllm_run the library issues the step the engine will most likely ask for next — the
 * same sequences one token longer (llm_engine.go:62-70 loops Step over an unchanged batch) — so
 * the GPU does not idle during the host's turnaround.  The next nvllm_run that asks for exactly
 * that step (same lengths + 1, last token = the one returned, same block tables) only waits
 * for it and copies nothing to the device; any other call restages everything as usual.
 * After one miss the next call does not speculate. */
int nvllm_set_speculation(nvllm_handle* h, int enabled);

/* GetVocabSize / GetModelConfig (purego/generic_runner.go:184-196). */
int nvllm_get_config(const nvllm_handle* h, nvllm_config* out);

/* Bytes of HBM the handle holds: weights, KV pool, workspaces. */
int nvllm_memory_info(const nvllm_handle* h, uint64_t* weight_bytes,
                      uint64_t* kv_bytes, uint64_t* workspace_bytes);

/* ---- single-operator entry points (parity tests + roofline timing) ------ */
/* Each takes HOST buffers, runs the same kernel the forward pass uses on
 * cfg-independent shapes, and returns the average device time over `iters`
 * launches (CUDA events on the launching stream; L2 flushed between launches
 * when flush_l2 != 0).  prec selects the fp32 or bf16 kernel family. */

/* C[M,N] = A[M,K] @ W[K,N] (+ bias[N]); W given [in,out] like tensor.MatMul
 * (tensor.go:62-88).  kernel: 0 auto, 1 gemv/split-K, 2 simt tiled, 3 tcgen05. */
int nvllm_op_matmul(int device, int prec, int kernel, int M, int N, int K,
                    const float* A, const float* W, const float* bias, float* C,
                    int iters, int flush_l2, float* avg_ms);

/* y = LayerNorm / RMSNorm (bias == NULL) of x[rows, H]   (tensor.go:193-250) */
int nvllm_op_norm(int device, int prec, int rows, int H, const float* x,
                  const float* w, const float* b, float eps, float* y,
                  int iters, float* avg_ms);

/* first-max argmax per row (cmd/ask/main.go:389-402) */
int nvllm_op_argmax(int device, int rows, int V, const float* logits, int32_t* out,
                    int iters, float* avg_ms);

/* Allocation-free split-K GEMV: C[M,N] (fp32) = A[M,K] @ W[N,K]^T on DEVICE pointers, launched on
 * `stream` (a cudaStream_t, NULL = default stream); asynchronous, nothing is synchronised.
 * No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (CUDA graph capture, tight serving loops).  It calls no cudaMalloc /
 * cudaFree / allocating library routine: the split-K slices live in a static device array of
 * NVLLM_NOALLOC_MAX_SCRATCH_FLOATS floats (4 MiB, held by every GPU that loads the library
 * until exit).  prec NVLLM_PREC_FP32: A and W are fp32; NVLLM_PREC_BF16: both bf16; W is
 * [out, in] row-major as the library keeps weights.  1 <= M <= 8, K % 8 == 0, A and W 16-byte
 * aligned, and slices(M,N,K) * M * N <= NVLLM_NOALLOC_MAX_SCRATCH_FLOATS, else
 * NVLLM_ERR_INVALID with the sizes in nvllm_last_error(NULL).  M = N = K = 0 only performs the
 * per-device set-up (the first call on a device looks the array's address up: make it outside a
 * capture).  All calls share the one array: calls that overlap on the GPU race. */
#define NVLLM_NOALLOC_MAX_SCRATCH_FLOATS (1 << 20)
int nvllm_op_gemv_noalloc(int device, int prec, int M, int N, int K, const void* d_A,
                          const void* d_W, float* d_C, void* stream);

/* Device properties the host side wants for roofline reporting. */
int nvllm_device_info(int device, int* sm_count, int* cc_major, int* cc_minor,
                      uint64_t* total_mem, char name[256]);

#ifdef __cplusplus
}
#endif
#endif /* NVLLM_H_ */
