// kernels.h — internal launcher interface between the C-ABI orchestration
// (nvllm.cu) and the kernel translation units.  Everything here is device
// pointers + a stream; errors come back as cudaError_t.
//
// Conventions
//  * residual stream x: fp32 [T, H]
//  * "act" tensors (GEMM inputs): dtype DT (fp32 or bf16), row-major, K contiguous
//  * weights: [N, K] K contiguous ("out,in"), dtype DT
//  * GEMM outputs: fp32 "partials" out[s][row][col], s < S (split-K); consumers
//    sum the S slices (deterministic, no atomics) and add the bias
//  * KV pool of one layer: K and V each [num_blocks][n_kv][256][hd], dtype DT
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

namespace nvllm {

enum DType { DT_F32 = 0, DT_BF16 = 1 };
static inline size_t dt_size(int dt) { return dt == DT_F32 ? 4 : 2; }

constexpr int KV_BLOCK = 256;

// ---------------------------------------------------------------- elementwise.cu
// x[t,:] = (wte[tok[t],:] (+ wpe[pos[t],:] if wpe && pos < max_pos)) * mult   (mult==0 => 1)
// xb / ssq / g (all or none): also write bf16(row * g) — g = the weight of the RMSNorm that reads
// the row next — and the row's sum of squares ssq[t] (what a GEMM with a folded RMSNorm reads
// instead of a normed activation)
cudaError_t launch_embed(cudaStream_t st, int wdt, const int* tokens, const int* positions,
                         const void* wte, const void* wpe, int max_pos, float mult, float* x, int T,
                         int H, void* xb = nullptr, float* ssq = nullptr, const float* g = nullptr);

struct NormArgs {
  float* x;            // [*, H] residual stream, updated in place when p1/p2 given
  const float* p1;     // partials [S1][rows_total][H] or null
  int p1_bf16;         // p1 is ONE slice of bf16 (a tensor-parallel all-reduce result)
  int S1;
  size_t stride1;      // elements between slices
  float m1;            // x += m1 * (sum_s p1 + bias1)
  const float* bias1;  // [H] or null
  const float* p2;     // second addend (Falcon parallel block: attn + ffn)
  int S2;
  size_t stride2;
  float m2;
  const float* bias2;
  const float* w;      // [H]; null = all ones (the weight is folded into the next GEMM)
  const float* b;      // [H] or null => RMSNorm (tensor.go:197)
  float eps;
  void* y;             // [n_rows, H] dtype ydt (may be null: residual update only)
  int ydt;
  float* y_f32;        // optional extra fp32 copy of y (MoE router input)
  const int* row_idx;  // optional gather: output row i reads x row row_idx[i] - row_bias
  int row_bias;
  int n_rows;
  int H;
  // mixture-of-experts fusions (moe.cu holds the stand-alone kernels)
  //  * comb_rows != null: the first addend is the weighted combine of the expert outputs,
  //    sum_s comb_w[r, s] * p1[comb_rows[r, s], :] in slot order (moe.go:116-118), instead of a
  //    split-K sum (S1 / stride1 unused)
  const int* comb_rows;   // [*, comb_k] row of p1 holding (token r, slot s)
  const float* comb_w;    // [*, comb_k]
  int comb_k;
  //  * router != null: route the normed row (fp32, before rounding) — softmax over all E
  //    experts, top-k, renormalise (moe.go:63-103) — and write topk_idx / topk_w [n_rows, k]
  const float* router;    // [E, H] fp32
  int E, topk;
  int* topk_idx;
  float* topk_w;
};
cudaError_t launch_norm(cudaStream_t st, const NormArgs& a);

// SwiGLU: out[t,j] = silu(h[t,j]) * h[t,F+j]   (transformer.go:50-66, mamba2.go:360-367)
// GELU:   out[t,j] = gelu_tanh(h[t,j] + b1[j]) (transformer.go:67-78, tensor.go:181-190)
cudaError_t launch_act(cudaStream_t st, int odt, int T, int F, const float* h, int S, size_t stride,
                       const float* bias, int swiglu, void* out);

// Sum split-K partials (+bias), RoPE rotate_half on Q and K (rope.go:153-205), write Q
// [T, nh*hd] (dtype qdt) and scatter K,V into the paged pool (dtype kvdt) at slots[t].
// cos/sin: [max_pos, hd/2] fp32 computed in double on the host exactly like rope.go:18-50;
// null => no RoPE (GPT-2).  qkv columns: [Q heads | K heads | V heads].
cudaError_t launch_rope_kv(cudaStream_t st, int qdt, int kvdt, int T, int nh, int nkv, int hd,
                           const float* qkv, int S, size_t stride, const float* bias,
                           const int* positions, const int* slots, const float* cosT,
                           const float* sinT, void* q_out, void* kpool, void* vpool);

// logits [n, ldl] fp32 (first V columns valid).  temperature[i] <= 0: first-max argmax
// (cmd/ask/main.go:389-402).  Otherwise softmax(l/T) and inverse-CDF draw with uniforms[i]
// (sampling.go:70-74,105-128,196-217).
// vsplit > 0: the logits are vocabulary slices gathered rank-major, [V / vsplit][n][vsplit]
// (tensor-parallel LM head); logit v of row s lives at (v / vsplit) * n * vsplit + s * vsplit +
// v % vsplit.  vsplit = 0: plain [n, ldl].
// sampling.cu: SampleWithHistory (sampling.go:33-217) for the rows with temperature > 0 —
// repetition penalty over hist[cu_hist[s] .. cu_hist[s+1]), temperature, softmax, top-k, top-p,
// renormalise, inverse-CDF draw with uniforms[s]; probs = [n, V] scratch.  Overwrites
// out_tokens[s] of those rows only.
cudaError_t launch_sample_filtered(cudaStream_t st, const float* logits, int ldl, int V, int n,
                                   int vsplit, const float* temperature, const float* uniforms,
                                   const int* hist, const int* cu_hist, float top_p, int top_k,
                                   float rep, float* probs, int* out_tokens);
cudaError_t launch_sample(cudaStream_t st, const float* logits, int ldl, int V, int n,
                          const float* temperature, const float* uniforms, int* out_tokens,
                          unsigned long long* amax = nullptr, int* ticket = nullptr, int vsplit = 0);

// advance device-resident decode state by one token (nvllm_run_resident)
cudaError_t launch_advance(cudaStream_t st, int n, const int* sampled, int* tokens, int* positions,
                           int* slots, int* seq_len, const int* block_tables, int max_blocks,
                           int max_pos);

// decode-step state from the prefill step that just sampled `sampled` (temperature / uniforms
// are copied bit-wise)
cudaError_t launch_make_decode_state(cudaStream_t st, int n, int mb, const int* sampled,
                                     const int* src_seq_len, const int* src_bt, const int* src_temp,
                                     const int* src_uni, int* tokens, int* positions, int* slots,
                                     int* cu_q, int* seq_start, int* seq_len, int* last_row, int* bt,
                                     int* temp, int* uni);

// dst[n*row_mul + row_add, k] (dtype ddt) = src[k, n] (fp32|bf16|f16, rows ld_src apart) for
// k < K, n < N — weight repack (and tensor-parallel column / row slicing) at load
cudaError_t launch_transpose_convert(cudaStream_t st, int sdt_api, const void* src, size_t ld_src,
                                     int K, int N, int ddt, void* dst, int ld_dst, int row_mul = 1,
                                     int row_add = 0);
// dst (dtype ddt) = convert(src)  flat copy; glu_I > 0: rows of `cols` elements, each
// [2 glu_I, cols] matrix of the stack rewritten with gate / up rows interleaved
cudaError_t launch_convert(cudaStream_t st, int sdt_api, const void* src, size_t n, int ddt,
                           void* dst, int glu_I = 0, int cols = 0);
// dst[(n*row_mul + row_add), k] (dtype ddt) = src[n, k] for n < N, k < K (src rows ld_src apart)
cudaError_t launch_copy_rows(cudaStream_t st, int sdt_api, const void* src, size_t ld_src, int N,
                             int K, int ddt, void* dst, int ld_dst, int row_mul = 1,
                             int row_add = 0);
// out[i] = sum_s p[s*stride + i]
cudaError_t launch_reduce_partials(cudaStream_t st, const float* p, int S, size_t stride, size_t n,
                                   float* out);
// a bigger-than-L2 write, for cold-cache timing
cudaError_t launch_fill(cudaStream_t st, float* p, size_t n, float v);

// ---------------------------------------------------------------- tp.cu
// out[t,:] = m1 * (sum_s p1 + bias1) + m2 * (sum_s p2 + bias2): a rank's share of a residual
// update, summed over ranks by the all-reduce that follows (p2 may be null)
// out_bf16: the addend goes on the wire as bf16 (half the all-reduce bytes of a prefill step)
cudaError_t launch_tp_partial(cudaStream_t st, int T, int H, const float* p1, int S1, size_t stride1,
                              float m1, const float* bias1, const float* p2, int S2, size_t stride2,
                              float m2, const float* bias2, void* out, bool out_bf16 = false);

// ---------------------------------------------------------------- gemm_simt.cu
struct GemmArgs {
  const void* A;  // [M, K] dtype adt (row gather through a_rows if set)
  int adt;
  int lda;
  const void* W;  // [N, K] dtype wdt
  int wdt;
  int M, N, K;
  float* out;         // [S][M][ldo]
  int ldo;
  size_t split_stride;
  int S;              // filled by the launcher: number of split-K slices written
  size_t out_capacity;  // floats available at `out`; the launcher keeps S*M*ldo within it (0 = unchecked)
  float alpha;        // out = alpha * acc
  // grouped (MoE) extension: tile t covers rows [tile_row0[t], tile_row0[t]+tile_rows[t]) of the
  // sorted pair list with expert weights W + tile_expert[t]*expert_stride
  const int* a_rows;       // A row index per output row (gather), or null
  const int* tile_expert;  // per 64-row tile, or null
  const int* tile_row0;
  const int* tile_rows;
  const int* n_tiles;      // device scalar: number of valid tiles
  size_t expert_stride;    // elements
  int max_tiles;
};
// Single-sequence decode (M <= 8) is a chain of ~100 dependent launches of ~3 us: the W2 GEMV
// builds its own activation from the W1 GEMV's split-K slices instead of reading what an
// activation kernel stored (one launch per layer less; every CTA only touches its own k-chunk,
// so nothing is read twice).
//  mode 2: A = act(sum_s hp[s] + hbias): SwiGLU over interleaved gate / up columns, or tanh-GELU,
//          rounded to `adt` exactly like the stand-alone kernel rounds what it stores.
// (The same trick for the norms — every CTA re-deriving the row statistics from x + the pending
// slices — was measured on cfg1 and lost: 313 -> 334 us per decode step; removed.)
struct GemvPre {
  int mode;  // 0 = none, 2 = activation
  const float* hp;  // [Sh][M][hld] split-K slices of the FFN up projection
  int Sh;
  size_t hstride;
  int hld;
  const float* hbias;  // [K] (GELU branch) or null
  int swiglu;
};
// split-K slices a skinny GEMV will produce for (M,N,K); workspace must hold S*M*N floats
int gemv_splits(int M, int N, int K);
cudaError_t launch_gemv(cudaStream_t st, GemmArgs& g, const GemvPre* pre = nullptr);  // M <= 8
cudaError_t launch_gemm_simt(cudaStream_t st, GemmArgs& g);  // any M (fp32 parity mode, MoE)

// ---------------------------------------------------------------- gemm_tcgen05.cu
struct TmaDesc {
  alignas(64) unsigned char bytes[128];  // CUtensorMap
};
// 2-D bf16 tensor map: rows x cols (cols contiguous), box = box_rows x 64 cols, 128B swizzle
int make_tma_2d_bf16(TmaDesc* out, const void* base, uint64_t rows, uint64_t cols,
                     uint64_t row_stride_elems, uint32_t box_rows, std::string* err);
// memoised tensor map (pure function of pointer, shape, box); null on failure
const TmaDesc* tma_map_cached(const void* base, int rows, int cols, int box_rows);
struct Tc5Args {
  const void* A;  // activations [M, K] bf16, K contiguous
  const void* W;  // weights     [N, K] bf16, K contiguous
  int M, N, K;
  float* out;          // [S][M][ldo] fp32 partials
  int ldo;
  size_t split_stride;
  int S;               // requested split-K slices (tc5_splits)
  float alpha;
  size_t out_capacity;
  // fused epilogue (needs S == 1): the activation is applied to the accumulator and the bf16
  // FFN hidden state is written directly, no fp32 partials and no act kernel
  int epi;              // TC5_EPI_*
  void* epi_out;        // bf16 [M, epi_ld]
  int epi_ld;
  const float* epi_bias;  // [N] or null (GELU)
  // ---- RMSNorm folded into the GEMMs (forward.cu "fold"): the producer of x writes
  // bf16(x * g) (g = norm weight), the row statistic is applied to the accumulator.
  // consumer side, rs_ssq != null: A holds bf16(x * g), not yet divided by the row's rms, and every
  // accumulator row m is multiplied by 1 / sqrt(sum_p rs_ssq[p * rs_ld + m] * rs_inv_h + rs_eps)
  // before the epilogue proper (partials, SwiGLU, GELU)
  const float* rs_ssq;
  int rs_parts, rs_ld;
  float rs_inv_h, rs_eps;
  // producer side, epi == TC5_EPI_RESID (N = hidden size): x[m, n] += res_m * (acc + res_bias[n])
  // in place (fp32), the same value rounded to bf16 into res_xb, and the row's sum of squares
  // over this CTA's 128 columns into res_ssq[n_tile * res_ssq_ld + m] — the next GEMM's rs_ssq.
  // With S > 1 (token tile <= 64 rows) the S CTAs of an output tile are launched as ONE
  // thread-block cluster: the slices meet in the owners' shared memory (DSMEM), nothing goes
  // through `out`, and each CTA finishes its share of the tile's token rows.
  float* res_x;
  void* res_xb;
  float* res_ssq;
  int res_ssq_ld;
  const float* res_bias;
  float res_m;
  const float* res_g;    // [N] weight of the norm that consumes x next: res_xb = bf16(x * g)
  int max_ctas;  // > 0: cap of the persistent grid (leave SMs to a concurrent collective)
  // weights of a later GEMM to pull into L2 while this one waits / drains (decode only)
  const void* pf_ptr;
  size_t pf_bytes;
  // grouped (mixture-of-experts) GEMM, g_experts > 0: A is the expert-sorted activation matrix
  // [M, K]; W stacks g_experts matrices of [N, K]; the token tiles (g_tile rows each, device
  // lists written by moe_sort_kernel) say which rows meet which expert.  S must be 1.
  int g_experts;
  int g_tile;            // rows per token tile the sort used: 32 / 64 / 128 / 256
  int g_max_tiles;       // capacity of the tile lists
  const int* g_tile_expert;
  const int* g_tile_row0;
  const int* g_tile_rows;
  const int* g_n_tiles;  // [1] on the device
};
enum { TC5_EPI_PARTIAL = 0, TC5_EPI_SWIGLU = 1, TC5_EPI_GELU = 2, TC5_EPI_RESID = 3 };
// split-K slices for a TC5_EPI_RESID GEMM: like tc5_splits, but every unit must own an SM
// (the CTAs of a tile wait for each other); 0 = the shape cannot take the fused epilogue
int tc5_resid_splits(int M, int N, int K, int sm_count);
// token-tile rows for a grouped GEMM averaging `rows_per_expert` rows per expert
int tc5_group_tile(int rows_per_expert);
bool tc5_supported(int M, int N, int K);
int tc5_splits(int M, int N, int K, int sm_count);
// *S_out = slices actually written (<= a.S)
cudaError_t launch_gemm_tc5(cudaStream_t st, const Tc5Args& a, int* S_out);
int tc5_box_rows_act(int M);  // token-tile rows (UMMA N) the kernel uses for M tokens
// Two GEMMs in one launch (decode FFN: W1 + fused activation, then W2 on its output): a2's
// weight stream starts under a1's epilogues, no launch / drain gap in between.  M <= 128.
// ctr = two zeroed device counters kept by the caller.  *S2_out = slices of a2 actually written.
cudaError_t launch_gemm_tc5_chain(cudaStream_t st, const Tc5Args& a1, const Tc5Args& a2,
                                  unsigned* ctr, int* S2_out);

// ---------------------------------------------------------------- attention.cu
struct AttnArgs {
  const void* q;  // [T, nh*hd] dtype qdt
  void* out;      // [T, nh*hd] dtype qdt
  int qdt;
  const void* kpool;  // one layer
  const void* vpool;
  int kvdt;
  int n_seqs, nh, nkv, hd;
  float scale;
  const int* block_tables;  // [n_seqs, max_blocks]
  int max_blocks;
  // prefill
  const int* cu_q;       // [n_seqs+1] row offsets of the new tokens
  const int* seq_start;  // [n_seqs] position of the first new token
  int max_new;           // max new tokens of any sequence (grid sizing)
  const int* tile_seq;   // [n_qtiles] sequence of each query tile (prefill)
  const int* tile_q0;    // [n_qtiles] first new-token index of each query tile
  int n_qtiles;
  // decode
  const int* seq_len;  // [n_seqs] tokens incl. the new one
  int n_splits;
  float* ws;           // [n_seqs*nh*n_splits*(hd+2)] split partials
  // decode with RoPE + KV append fused in (attention_mma.cu), qkv_p != null: the kernel builds
  // Q and the new K / V row (position seq_len - 1) itself from the QKV GEMM's split-K partials
  // [S][n_seqs][qkv_ld] (columns [Q heads | K heads | V heads]) — `q` is not read and the
  // rope_kv kernel is not launched
  const float* qkv_p;
  int qkv_S;
  size_t qkv_stride;
  int qkv_ld;
  const float* qkv_bias;  // [qkv_ld] or null
  const float* cosT;      // [max_pos, hd/2] or null (no RoPE)
  const float* sinT;
  int newkv_off;          // set by the launcher: smem offset of the new K / V rows
  // stream-K decode (attention_mma.cu): ticket per (seq, kv head, head chunk), zero between
  // kernels; total_tiles = host LOWER bound of the key tiles in the step (caps the grid)
  int* tickets;        // unused (stream-K experiment removed)
  int pool_blocks;     // blocks in the layer's K / V pool (tensor-map extent)
  int total_rows;      // rows of q / out in this step (tensor-map extent; tcgen05 prefill kernel)
  int sm_count;
  long long total_tiles;
};
constexpr int ATTN_PREFILL_QTILE = 32;  // SIMT (fp32 parity) kernel; the mma kernel uses 64
// queries per tile of the kernel launch_attn_prefill will pick: the host builds tile_seq /
// tile_q0 with it
int attn_prefill_qtile(int qdt, int kvdt, int hd);
// KV splits for a decode step of n_seqs sequences whose longest context is max_len
int attn_decode_splits(int qdt, int kvdt, int n_seqs, int nh, int nkv, int hd, int max_len,
                       int sm_count);
cudaError_t launch_attn_prefill(cudaStream_t st, const AttnArgs& a);
cudaError_t launch_attn_decode(cudaStream_t st, const AttnArgs& a);
bool attn_decode_fuses_rope(int qdt, int kvdt, int hd);
// attention_mma.cu (bf16, hd 64 / 128)
bool attn_mma_supported(int qdt, int kvdt, int hd);
int attn_decode_mma_tile(int hd);
cudaError_t launch_attn_prefill_mma(cudaStream_t st, const AttnArgs& a);
cudaError_t launch_attn_decode_mma(cudaStream_t st, const AttnArgs& a);
// attention_tc5.cu: tcgen05 / TMEM / TMA prefill kernel (bf16, hd 64 / 128; 128-query tiles;
// needs AttnArgs::pool_blocks and total_rows).  NVLLM_ATTN_PREFILL_TC5=0 keeps the mma.sync one.
bool attn_prefill_tc5_enabled();
cudaError_t launch_attn_prefill_tc5(cudaStream_t st, const AttnArgs& a);
// attention_dec.cu: decode kernel for query-head groups of <= 8 (needs AttnArgs::pool_blocks)
bool attn_decode_t_supported(const AttnArgs& a);
cudaError_t launch_attn_decode_t(cudaStream_t st, const AttnArgs& a);

// ---------------------------------------------------------------- mamba.cu
// Mamba2 block between its two projections (mamba2.go:72-351), ragged batch
struct MambaArgs {
  int T, n_seqs;
  int nh, hd, N, G, K, E, conv, P;  // heads, head dim, state size, groups, conv taps, E = nh hd,
                                    // conv = E + 2 G N, P = E + conv + nh
  const float* proj;   // [S][T][P] split-K slices of x @ in_proj^T: [gate | xBC | dt]
  int S;
  size_t stride;
  const float *conv_w, *conv_b, *dt_bias, *a_log, *d_skip, *norm_w;
  float* xbc;          // [T, conv]  SiLU(conv1d(xBC))
  float* dt;           // [T, nh]    softplus(dt + bias)
  float* y;            // [T, E]     scan output
  const int* cu_q;         // [n_seqs + 1] rows of each sequence
  const int* positions;    // [T]
  const int* block_tables; // [n_seqs, max_blocks]
  int max_blocks;
  float* ssm_pool;     // [blocks][n_mamba][nh][hd][N]: state after the last token of each block
  int n_mamba, layer;  // Mamba2 layers in the model, index of this one among them
};
cudaError_t launch_mamba_conv(cudaStream_t st, const MambaArgs& a);
cudaError_t launch_mamba_scan(cudaStream_t st, const MambaArgs& a);
cudaError_t launch_mamba_gate_norm(cudaStream_t st, const MambaArgs& a, int odt, void* out);

// ---------------------------------------------------------------- moe.cu
// Router: probs = softmax(x @ R) over all E, top-k (descending), weights renormalised
// (moe.go:63-103).  x fp32 [T,H]; R fp32 [E,H].
cudaError_t launch_moe_router(cudaStream_t st, const float* x, const float* R, int T, int H, int E,
                              int k, int* topk_idx, float* topk_w);
// Group the T*k (token,slot) pairs by expert into `tile`-row tiles.
struct MoeSortBufs {
  int* counts;       // [E]
  int* offsets;      // [E+1]
  int* cursor;       // [E]
  int* sorted_pair;  // [T*k] pair id (t*k+s) grouped by expert
  int* pair_row;     // [T*k] row of pair in the sorted order
  int* a_rows;       // [T*k] token index per sorted row
  int* tile_expert;  // [max_tiles]
  int* tile_row0;
  int* tile_rows;
  int* n_tiles;      // [1]
};
cudaError_t launch_moe_sort(cudaStream_t st, const int* topk_idx, int T, int k, int E,
                            const MoeSortBufs& b, int max_tiles, int tile);
// sort + gather in one launch for decode-sized batches (moe_sort_gather_fits)
bool moe_sort_gather_fits(int T, int k, int E, int H);
cudaError_t launch_moe_sort_gather(cudaStream_t st, const int* topk_idx, int T, int k, int E,
                                   const MoeSortBufs& b, int max_tiles, int tile, int H,
                                   const void* x, void* xs);
// xs[row, :] = x[a_rows[row], :]  (bf16 rows of H elements)
cudaError_t launch_moe_gather(cudaStream_t st, int rows, int H, const void* x, const int* a_rows,
                              void* xs);
// g[row, j] = silu(h[row, 2j]) * h[row, 2j + 1] (gate / up interleaved at load) with the fp32
// sigmoid of moe.go:203-212
cudaError_t launch_moe_glu(cudaStream_t st, int odt, int rows, int I, const float* h, void* g);
// y[t,:] = sum_s w[t,s] * Y[pair_row[t*k+s], :]   (moe.go:116-118, slot order = descending prob)
cudaError_t launch_moe_combine(cudaStream_t st, int T, int k, int H, const float* Y,
                               const int* pair_row, const float* topk_w, float* y);

}  // namespace nvllm
