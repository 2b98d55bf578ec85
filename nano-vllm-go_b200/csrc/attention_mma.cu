// attention_mma.cu — bf16 paged attention on the tensor cores (warp-level mma.sync
// m16n8k16, fp32 accumulate, online softmax in registers), K/V tiles staged in shared memory
// with cp.async straight out of the paged pool through the block tables.
//
// Same semantics as attention.cu (reference attention.go:354-421, mqa.go:184-243): scores =
// (q.k)*scale, causal mask key_pos > query_pos, softmax, P.V; q head h uses kv head
// h / (n_heads / n_kv_heads).  bf16 mode only (gate 2e-2): P is rounded to bf16 for the
// P.V product; the fp32 parity mode keeps the SIMT kernels of attention.cu.
//
//  * attn_prefill_mma_kernel: CTA = 64 queries of one sequence x one head (4 warps x 16
//    rows), K/V tiles of 64 keys double-buffered.  Tensor-bound.
//  * attn_decode_mma_kernel: CTA = (sequence, kv head, 16-row chunk of its query-head
//    group, KV split).  The query heads that share a kv head are the M rows of the MMA,
//    so each K/V byte is read once per group.  HBM-bound: 3-4 cp.async stages keep
//    >= 48 KB per SM in flight.
#include "common.cuh"
#include "kernels.h"

namespace nvllm {

// ------------------------------------------------------------------ primitives
__device__ __forceinline__ uint32_t smem_addr(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
  const int sz = valid ? 16 : 0;  // src-size 0 => 16 bytes of zeros
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() {
  asm volatile("cp.async.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2,
                                        uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2,
                                          uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
__device__ __forceinline__ void mma16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                         uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
      "{%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
// byte offset of 16-byte chunk `c` of row `r` in a [rows][HD] bf16 tile; the low 3 chunk bits
// are XORed with the row so that the 8 rows of an ldmatrix hit 8 different bank groups
template <int HD>
__device__ __forceinline__ uint32_t swz(int r, int c) {
  return (uint32_t)(r * (HD * 2) + ((c >> 3) << 7) + (((c & 7) ^ (r & 7)) << 4));
}

// One 16-row x BN-key block of attention for one warp: S = Q K^T (Q fragments in registers),
// mask, online softmax update of (m, l, O), O += P V.
//   qf      Q fragments [HD/16][4]
//   sK, sV  shared addresses of the K / V tiles ([keys][HD], swizzled), key0 = first key row
//           of this warp inside the tile, NT = number of 8-key n-tiles this warp covers
//   kpos0   position of key row `key0`; limit[i] = last visible key position of row i (g, g+8)
template <int HD, int NT>
__device__ __forceinline__ void attn_block(const uint32_t (&qf)[HD / 16][4], uint32_t sK,
                                           uint32_t sV, int key0, int kpos0, const int (&limit)[2],
                                           bool need_mask, float scale_log2, float (&m)[2],
                                           float (&l)[2], float (&o)[HD / 8][4]) {
  const int lane = threadIdx.x & 31;
  const int t = lane & 3;
  const int mi = lane >> 3, lr = lane & 7;
  float s[NT][4];
#pragma unroll
  for (int nt = 0; nt < NT; nt++)
#pragma unroll
    for (int i = 0; i < 4; i++) s[nt][i] = 0.f;
#pragma unroll
  for (int ks = 0; ks < HD / 16; ks++) {
#pragma unroll
    for (int nt = 0; nt < NT; nt += 2) {
      uint32_t b0, b1, b2, b3;
      ldsm_x4(sK + swz<HD>(key0 + nt * 8 + lr + (mi >> 1) * 8, ks * 2 + (mi & 1)), b0, b1, b2, b3);
      mma16816(s[nt], qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3], b0, b1);
      mma16816(s[nt + 1], qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3], b2, b3);
    }
  }
  if (need_mask) {
#pragma unroll
    for (int nt = 0; nt < NT; nt++) {
      const int kp = kpos0 + nt * 8 + 2 * t;
      if (kp > limit[0]) s[nt][0] = -INFINITY;
      if (kp + 1 > limit[0]) s[nt][1] = -INFINITY;
      if (kp > limit[1]) s[nt][2] = -INFINITY;
      if (kp + 1 > limit[1]) s[nt][3] = -INFINITY;
    }
  }
#pragma unroll
  for (int i = 0; i < 2; i++) {
    float mx = -INFINITY;
#pragma unroll
    for (int nt = 0; nt < NT; nt++) mx = fmaxf(mx, fmaxf(s[nt][2 * i], s[nt][2 * i + 1]));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
    const float m_new = fmaxf(m[i], mx);
    const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
    const float corr = exp2f((m[i] - m_use) * scale_log2);  // m = -inf => 0
    m[i] = m_new;
    l[i] *= corr;
#pragma unroll
    for (int dn = 0; dn < HD / 8; dn++) {
      o[dn][2 * i] *= corr;
      o[dn][2 * i + 1] *= corr;
    }
    float rs = 0.f;
#pragma unroll
    for (int nt = 0; nt < NT; nt++) {
      const float p0 = exp2f((s[nt][2 * i] - m_use) * scale_log2);
      const float p1 = exp2f((s[nt][2 * i + 1] - m_use) * scale_log2);
      s[nt][2 * i] = p0;
      s[nt][2 * i + 1] = p1;
      rs += p0 + p1;
    }
    l[i] += rs;  // per-thread partial; the quad is reduced once at the end
  }
#pragma unroll
  for (int kt = 0; kt < NT / 2; kt++) {
    const uint32_t a0 = pack_bf16(s[2 * kt][0], s[2 * kt][1]);
    const uint32_t a1 = pack_bf16(s[2 * kt][2], s[2 * kt][3]);
    const uint32_t a2 = pack_bf16(s[2 * kt + 1][0], s[2 * kt + 1][1]);
    const uint32_t a3 = pack_bf16(s[2 * kt + 1][2], s[2 * kt + 1][3]);
#pragma unroll
    for (int dn = 0; dn < HD / 8; dn += 2) {
      uint32_t b0, b1, b2, b3;
      ldsm_x4_t(sV + swz<HD>(key0 + kt * 16 + lr + (mi & 1) * 8, dn + (mi >> 1)), b0, b1, b2, b3);
      mma16816(o[dn], a0, a1, a2, a3, b0, b1);
      mma16816(o[dn + 1], a0, a1, a2, a3, b2, b3);
    }
  }
}

// cooperative cp.async of `rows` x HD bf16 (global rows contiguous, HD*2 bytes each) into a
// swizzled tile; rows >= valid_rows are zero-filled
template <int HD, int THREADS>
__device__ __forceinline__ void load_tile(uint32_t dst, const bf16* src, int rows, int valid_rows,
                                          size_t src_row_stride) {
  constexpr int CPR = HD / 8;
  for (int i = threadIdx.x; i < rows * CPR; i += THREADS) {
    const int r = i / CPR, c = i % CPR;
    const bool ok = r < valid_rows;
    cp_async16(dst + swz<HD>(r, c), ok ? (const void*)(src + (size_t)r * src_row_stride + c * 8)
                                       : (const void*)src, ok);
  }
}

// ------------------------------------------------------------------ prefill
template <int HD>
__global__ void __launch_bounds__(128)
attn_prefill_mma_kernel(AttnArgs a) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int BM = 64, BN = 64;
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t sQ = smem_addr(smem);
  const uint32_t sK = sQ + BM * HD * 2;
  const uint32_t sV = sK + 2 * BN * HD * 2;
  const int tile = blockIdx.x, head = blockIdx.y;
  const int seq = a.tile_seq[tile], q0 = a.tile_q0[tile];
  const int row0 = a.cu_q[seq];
  const int n_new = a.cu_q[seq + 1] - row0;
  const int start = a.seq_start[seq];
  const int kvh = head / (a.nh / a.nkv);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int* bt = a.block_tables + (size_t)seq * a.max_blocks;
  const bf16* kpool = (const bf16*)a.kpool;
  const bf16* vpool = (const bf16*)a.vpool;
  const int rows_valid = min(BM, n_new - q0);
  const int kv_end = start + q0 + rows_valid;  // keys [0, kv_end) are visible to this tile
  const int n_tiles = (kv_end + BN - 1) / BN;

  auto issue_kv = [&](int j, int buf) {
    const int k0 = j * BN;
    const int nk = min(BN, kv_end - k0);
    const size_t base = (((size_t)bt[k0 / KV_BLOCK] * a.nkv + kvh) * KV_BLOCK + (k0 % KV_BLOCK)) * HD;
    load_tile<HD, 128>(sK + buf * BN * HD * 2, kpool + base, BN, nk, HD);
    load_tile<HD, 128>(sV + buf * BN * HD * 2, vpool + base, BN, nk, HD);
  };
  load_tile<HD, 128>(sQ, (const bf16*)a.q + (size_t)(row0 + q0) * a.nh * HD + head * HD, BM,
                     rows_valid, (size_t)a.nh * HD);
  issue_kv(0, 0);
  cp_async_commit();

  uint32_t qf[HD / 16][4];
  float o[HD / 8][4];
#pragma unroll
  for (int dn = 0; dn < HD / 8; dn++)
#pragma unroll
    for (int i = 0; i < 4; i++) o[dn][i] = 0.f;
  float m[2] = {-INFINITY, -INFINITY}, l[2] = {0.f, 0.f};
  const int qpos = start + q0 + warp * 16 + g;
  const int limit[2] = {qpos, qpos + 8};
  const float scale_log2 = a.scale * 1.4426950408889634f;

  for (int j = 0; j < n_tiles; j++) {
    const int buf = j & 1;
    cp_async_wait<0>();
    __syncthreads();  // tile j landed for everyone; everyone is done with buffer buf^1
    if (j + 1 < n_tiles) issue_kv(j + 1, buf ^ 1);
    cp_async_commit();
    if (j == 0) {
      const int mi = lane >> 3, lr = lane & 7;
#pragma unroll
      for (int ks = 0; ks < HD / 16; ks++)
        ldsm_x4(sQ + swz<HD>(warp * 16 + lr + (mi & 1) * 8, ks * 2 + (mi >> 1)), qf[ks][0], qf[ks][1],
                qf[ks][2], qf[ks][3]);
    }
    const int k0 = j * BN;
    // the warp's rows see every key of the tile when k0 + BN - 1 <= its first row's position
    const bool need_mask = (k0 + BN - 1) > (start + q0 + warp * 16);
    attn_block<HD, BN / 8>(qf, sK + buf * BN * HD * 2, sV + buf * BN * HD * 2, 0, k0, limit,
                           need_mask, scale_log2, m, l, o);
  }
#pragma unroll
  for (int i = 0; i < 2; i++) {
    l[i] += __shfl_xor_sync(0xffffffffu, l[i], 1);
    l[i] += __shfl_xor_sync(0xffffffffu, l[i], 2);
  }
  bf16* out = (bf16*)a.out;
#pragma unroll
  for (int i = 0; i < 2; i++) {
    const int r = warp * 16 + g + 8 * i;
    if (r >= rows_valid) continue;
    const float inv = 1.0f / l[i];
    bf16* op = out + (size_t)(row0 + q0 + r) * a.nh * HD + head * HD + 2 * t;
#pragma unroll
    for (int dn = 0; dn < HD / 8; dn++)
      *reinterpret_cast<__nv_bfloat162*>(op + dn * 8) =
          __floats2bfloat162_rn(o[dn][2 * i] * inv, o[dn][2 * i + 1] * inv);
  }
}

// ------------------------------------------------------------------ decode
// split partial record (shared with attention.cu's combine kernel): [hd] unnormalised
// accumulator, then m (in score units, already multiplied by scale), l
template <int HD, int BN, int STAGES>
__global__ void __launch_bounds__(128)
attn_decode_mma_kernel(AttnArgs a, int G, int chunks) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int TILE_BYTES = BN * HD * 2;
  constexpr int WK = BN / 4;  // keys per warp per tile
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t sQ = smem_addr(smem);  // [16][HD]
  const uint32_t sK = sQ + 16 * HD * 2;
  const uint32_t sV = sK + STAGES * TILE_BYTES;
  const int seq = blockIdx.x;
  const int kvh = blockIdx.y / chunks, chunk = blockIdx.y % chunks;
  const int split = blockIdx.z;
  const int h0 = kvh * G + chunk * 16;           // first query head of this CTA
  const int rows = min(16, G - chunk * 16);      // valid query heads
  const int len = a.seq_len[seq];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const int* bt = a.block_tables + (size_t)seq * a.max_blocks;
  const bf16* kpool = (const bf16*)a.kpool;
  const bf16* vpool = (const bf16*)a.vpool;
  // this split's key range, in whole tiles
  const int tiles_total = (len + BN - 1) / BN;
  const int tps = (tiles_total + a.n_splits - 1) / a.n_splits;
  const int tile_begin = split * tps;
  const int n_tiles = max(0, min(tps, tiles_total - tile_begin));

  auto issue_kv = [&](int j, int stage) {
    const int k0 = (tile_begin + j) * BN;
    const int nk = min(BN, len - k0);
    const size_t base = (((size_t)bt[k0 / KV_BLOCK] * a.nkv + kvh) * KV_BLOCK + (k0 % KV_BLOCK)) * HD;
    load_tile<HD, 128>(sK + stage * TILE_BYTES, kpool + base, BN, nk, HD);
    load_tile<HD, 128>(sV + stage * TILE_BYTES, vpool + base, BN, nk, HD);
  };
  load_tile<HD, 128>(sQ, (const bf16*)a.q + (size_t)seq * a.nh * HD + (size_t)h0 * HD, 16, rows, HD);
#pragma unroll
  for (int s = 0; s < STAGES - 1; s++) {
    if (s < n_tiles) issue_kv(s, s);
    cp_async_commit();
  }

  uint32_t qf[HD / 16][4];
  float o[HD / 8][4];
#pragma unroll
  for (int dn = 0; dn < HD / 8; dn++)
#pragma unroll
    for (int i = 0; i < 4; i++) o[dn][i] = 0.f;
  float m[2] = {-INFINITY, -INFINITY}, l[2] = {0.f, 0.f};
  const int limit[2] = {len - 1, len - 1};
  const float scale_log2 = a.scale * 1.4426950408889634f;

  for (int j = 0; j < n_tiles; j++) {
    const int stage = j % STAGES;
    cp_async_wait<STAGES - 2>();
    __syncthreads();
    if (j + STAGES - 1 < n_tiles) issue_kv(j + STAGES - 1, (j + STAGES - 1) % STAGES);
    cp_async_commit();
    if (j == 0) {
      const int mi = lane >> 3, lr = lane & 7;
#pragma unroll
      for (int ks = 0; ks < HD / 16; ks++)
        ldsm_x4(sQ + swz<HD>(lr + (mi & 1) * 8, ks * 2 + (mi >> 1)), qf[ks][0], qf[ks][1], qf[ks][2],
                qf[ks][3]);
    }
    const int k0 = (tile_begin + j) * BN;
    const bool need_mask = k0 + BN > len;
    attn_block<HD, WK / 8>(qf, sK + stage * TILE_BYTES, sV + stage * TILE_BYTES, warp * WK,
                           k0 + warp * WK, limit, need_mask, scale_log2, m, l, o);
  }
  cp_async_wait<0>();
#pragma unroll
  for (int i = 0; i < 2; i++) {
    l[i] += __shfl_xor_sync(0xffffffffu, l[i], 1);
    l[i] += __shfl_xor_sync(0xffffffffu, l[i], 2);
  }
  // merge the 4 warps (each covered different keys) through shared memory
  __syncthreads();  // pipeline buffers are free now
  float* red = reinterpret_cast<float*>(smem + 16 * HD * 2);  // [4][16][HD + 2]
  constexpr int RS = HD + 2;
#pragma unroll
  for (int i = 0; i < 2; i++) {
    float* rr = red + ((size_t)warp * 16 + g + 8 * i) * RS;
#pragma unroll
    for (int dn = 0; dn < HD / 8; dn++) {
      rr[dn * 8 + 2 * t] = o[dn][2 * i];
      rr[dn * 8 + 2 * t + 1] = o[dn][2 * i + 1];
    }
    if (t == 0) {
      rr[HD] = m[i];
      rr[HD + 1] = l[i];
    }
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < rows * HD; idx += 128) {
    const int r = idx / HD, d = idx % HD;
    float mn = -INFINITY;
#pragma unroll
    for (int w = 0; w < 4; w++) mn = fmaxf(mn, red[((size_t)w * 16 + r) * RS + HD]);
    float lt = 0.f, at = 0.f;
#pragma unroll
    for (int w = 0; w < 4; w++) {
      const float* rr = red + ((size_t)w * 16 + r) * RS;
      const float c = (rr[HD] == -INFINITY) ? 0.f : exp2f((rr[HD] - mn) * scale_log2);
      lt += rr[HD + 1] * c;
      at += rr[d] * c;
    }
    const int h = h0 + r;
    if (a.n_splits == 1) {
      ((bf16*)a.out)[(size_t)seq * a.nh * HD + (size_t)h * HD + d] = __float2bfloat16_rn(at / lt);
    } else {
      float* rec = a.ws + (((size_t)seq * a.nh + h) * a.n_splits + split) * (HD + 2);
      rec[d] = at;
      if (d == 0) {
        rec[HD] = (mn == -INFINITY) ? -INFINITY : mn * a.scale;  // combine kernel works in exp()
        rec[HD + 1] = lt;
      }
    }
  }
}

// ------------------------------------------------------------------ launchers
bool attn_mma_supported(int qdt, int kvdt, int hd) {
  return qdt == DT_BF16 && kvdt == DT_BF16 && (hd == 64 || hd == 128);
}

template <int HD>
static cudaError_t prefill_launch(cudaStream_t st, const AttnArgs& a) {
  constexpr size_t smem = (size_t)(64 + 4 * 64) * HD * 2;
  cudaError_t e = cudaFuncSetAttribute(attn_prefill_mma_kernel<HD>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  (void)launch_k(attn_prefill_mma_kernel<HD>, dim3(dim3(a.n_qtiles, a.nh)), dim3(128), smem, st, a);
  return cudaGetLastError();
}
cudaError_t launch_attn_prefill_mma(cudaStream_t st, const AttnArgs& a) {
  if (a.n_qtiles <= 0) return cudaSuccess;
  return a.hd == 64 ? prefill_launch<64>(st, a) : prefill_launch<128>(st, a);
}

template <int HD, int BN, int STAGES>
static cudaError_t decode_launch(cudaStream_t st, const AttnArgs& a, int G, int chunks) {
  constexpr size_t pipe = (size_t)2 * STAGES * BN * HD * 2;
  constexpr size_t red = (size_t)4 * 16 * (HD + 2) * 4;
  constexpr size_t smem = 16 * HD * 2 + (pipe > red ? pipe : red);
  cudaError_t e = cudaFuncSetAttribute(attn_decode_mma_kernel<HD, BN, STAGES>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  (void)launch_k(attn_decode_mma_kernel<HD, BN, STAGES>, dim3(dim3(a.n_seqs, a.nkv * chunks, a.n_splits)), dim3(128), smem, st, a, G, chunks);
  return cudaGetLastError();
}
int attn_decode_mma_tile(int hd) { return hd == 64 ? 128 : 64; }
cudaError_t launch_attn_decode_mma(cudaStream_t st, const AttnArgs& a) {
  if (a.n_seqs <= 0) return cudaSuccess;
  const int G = a.nh / a.nkv, chunks = (G + 15) / 16;
  if (a.hd == 64) return decode_launch<64, 128, 3>(st, a, G, chunks);
  return decode_launch<128, 64, 3>(st, a, G, chunks);
}

}  // namespace nvllm
