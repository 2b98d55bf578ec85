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
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"
#include "tma.cuh"
#include "mma.cuh"

namespace nvllm {

// One 16-row x BN-key block of attention for one warp: S = Q K^T (Q fragments in registers),
// mask, online softmax update of (m, l, O), O += P V.
//   qf      Q fragments [HD/16][4]
//   sK, sV  shared addresses of the K / V tiles ([keys][HD], swizzled), key0 = first key row
//           of this warp inside the tile, NT = number of 8-key n-tiles this warp covers
//   kpos0   position of key row `key0`; limit[i] = last visible key position of row i (g, g+8)
// TROWS == 0: K/V tiles use swz<HD> (cp.async-staged, rows HD*2 bytes apart); TROWS > 0: tiles
// were written by TMA with SWIZZLE_128B as HD/64 panels of [TROWS keys][128 B] (tswz)
template <int HD, int TROWS>
__device__ __forceinline__ uint32_t tile_off(int r, int c) {
  if constexpr (TROWS == 0) return swz<HD>(r, c);
  else return (uint32_t)((c >> 3) * (TROWS * 128) + r * 128 + (((c & 7) ^ (r & 7)) << 4));
}
template <int HD, int NT, int TROWS = 0>
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
      ldsm_x4(sK + tile_off<HD, TROWS>(key0 + nt * 8 + lr + (mi >> 1) * 8, ks * 2 + (mi & 1)), b0, b1,
              b2, b3);
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
      ldsm_x4_t(sV + tile_off<HD, TROWS>(key0 + kt * 16 + lr + (mi & 1) * 8, dn + (mi >> 1)), b0, b1,
                b2, b3);
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
attn_prefill_mma_kernel(AttnArgs a, Trace tr) {
  TRACE_ENTRY(tr);
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
  TRACE_EXIT(tr);
}

// ------------------------------------------------------------------ decode
// split partial record (shared with attention.cu's combine kernel): [hd] unnormalised
// accumulator, then m (in score units, already multiplied by scale), l
//
// One CTA = (sequence, `kpc` consecutive kv heads [or one 16-row chunk of a big head group], KV
// split).  The fixed cost of a CTA (launch, first-tile latency, Q build, warp merge) is ~5 us
// against ~1 us per 128-key tile, so when the step has more (sequence, kv head) pairs than
// resident CTA slots the launcher packs kpc heads into one CTA: the K/V tile ring runs straight
// through the head boundaries and the grid fits one wave.
template <int HD, int BN, int STAGES>
__global__ void __launch_bounds__(128, (HD == 64 && BN == 64) ? 4 : 2)
attn_decode_mma_kernel(AttnArgs a, int G, int chunks, int kpc, Trace tr) {
  const unsigned long long t_entry = trace_on(tr) ? gtime() : 0ull;
  pdl_launch_dependents();
  constexpr int TILE_BYTES = BN * HD * 2;
  constexpr int WK = BN / 4;  // keys per warp per tile
  constexpr int RS = HD + 2;
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t sQ = smem_addr(smem);  // [kpc][16][HD]
  const uint32_t sK = sQ + kpc * 16 * HD * 2;
  const uint32_t sV = sK + STAGES * TILE_BYTES;
  // warp-merge scratch [4][min(16, G)][HD + 2] fp32: its own region when heads are packed (the ring is
  // still streaming the next head), aliased on the drained ring otherwise
  float* red = reinterpret_cast<float*>(
      smem + kpc * 16 * HD * 2 + (kpc > 1 ? 2 * STAGES * TILE_BYTES : 0));
  bf16* s_newkv = reinterpret_cast<bf16*>(smem + a.newkv_off);  // [kpc][2][HD]
  const int seq = blockIdx.x;
  const int kvh0 = (blockIdx.y / chunks) * kpc, chunk = blockIdx.y % chunks;
  const int split = blockIdx.z;
  const int rows = min(16, G - chunk * 16);      // valid query heads per kv head
  const int rrows = min(16, G);                  // rows of the warp-merge scratch
  const int len = a.seq_len[seq];                // step state: stable before the wait
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
  const int F = n_tiles * kpc;  // flat (head, tile) work list

  auto issue_kv = [&](int f, int stage) {
    const int hh = f / n_tiles, j = f - hh * n_tiles;
    const int k0 = (tile_begin + j) * BN;
    const int nk = min(BN, len - k0);
    const size_t base =
        (((size_t)bt[k0 / KV_BLOCK] * a.nkv + kvh0 + hh) * KV_BLOCK + (k0 % KV_BLOCK)) * HD;
    load_tile<HD, 128>(sK + stage * TILE_BYTES, kpool + base, BN, nk, HD);
    load_tile<HD, 128>(sV + stage * TILE_BYTES, vpool + base, BN, nk, HD);
  };
  // The only K/V row the previous kernel writes for this sequence is position len - 1: tiles
  // that end before it were written by earlier steps, so their loads start BEFORE the
  // dependency resolves.
  auto tile_is_old = [&](int f) { return (tile_begin + f % n_tiles + 1) * BN <= len - 1; };
#pragma unroll
  for (int s = 0; s < STAGES - 1; s++) {
    if (s < F && tile_is_old(s)) issue_kv(s, s);
    cp_async_commit();
  }
  // fused mode: this CTA's key range ends with the token being decoded, whose K / V row does
  // not exist in the pool yet — it is built below and patched into the last tile of each head
  const bool fused = a.qkv_p != nullptr;
  const bool has_new = fused && n_tiles > 0 && tile_begin + n_tiles == tiles_total;
  pdl_wait();
  const unsigned long long t_wait = trace_on(tr) ? gtime() : 0ull;
  if (fused) {  // the remaining tiles stream while the prologue below builds Q / K / V
#pragma unroll
    for (int s = 0; s < STAGES - 1; s++)
      if (s < F && !tile_is_old(s)) issue_kv(s, s);
  }
  if (!fused) {
    for (int hh = 0; hh < kpc; hh++)
      load_tile<HD, 128>(sQ + hh * 16 * HD * 2,
                         (const bf16*)a.q + (size_t)seq * a.nh * HD +
                             (size_t)((kvh0 + hh) * G + chunk * 16) * HD,
                         16, rows, HD);
  } else {
    // split-K sum + bias, rotate_half RoPE at position len - 1 (rope.go:153-205), bf16: what
    // rope_kv_kernel does for a decode token, for this CTA's query heads and its kv heads
    constexpr int HALF = HD / 2, VPH = HALF / 8;
    const int pos = len - 1;
    const float* rowp = a.qkv_p + (size_t)seq * a.qkv_ld;
    const int q_items = 16 * VPH, kv_items = has_new ? 2 * VPH : 0;
    const int per_head = q_items + kv_items;
    for (int it0 = threadIdx.x; it0 < per_head * kpc; it0 += 128) {
      const int hh = it0 / per_head, it = it0 - hh * per_head;
      const int kvh = kvh0 + hh;
      const bool is_q = it < q_items;
      const int r = is_q ? it / VPH : 0;
      const int which = is_q ? 0 : (it - q_items) / VPH;  // 0 = K, 1 = V
      const int j = (is_q ? it % VPH : (it - q_items) % VPH) * 8;
      uint4 lo = make_uint4(0, 0, 0, 0), hi = lo;
      if (!is_q || r < rows) {
        const int col = (is_q ? kvh * G + chunk * 16 + r : a.nh + which * a.nkv + kvh) * HD + j;
        float x[8], y[8], c[8], sn[8];
        const bool rot = a.cosT != nullptr && (is_q || which == 0);
        if (rot) {  // issued first: independent of the partial sums
          load8(a.cosT + (size_t)pos * HALF + j, c);
          load8(a.sinT + (size_t)pos * HALF + j, sn);
        }
        if (a.qkv_S <= 8) {
          // every slice of both halves in flight at once: one L2 round trip, not 2 S / 4
          float px[8][8], py[8][8];
#pragma unroll
          for (int q = 0; q < 8; q++)
            if (q < a.qkv_S) {
              load8(rowp + (size_t)q * a.qkv_stride + col, px[q]);
              load8(rowp + (size_t)q * a.qkv_stride + col + HALF, py[q]);
            }
#pragma unroll
          for (int i = 0; i < 8; i++) x[i] = y[i] = 0.f;
#pragma unroll
          for (int q = 0; q < 8; q++)
            if (q < a.qkv_S) {
#pragma unroll
              for (int i = 0; i < 8; i++) {
                x[i] += px[q][i];
                y[i] += py[q][i];
              }
            }
        } else {
          sum_partials8(rowp + col, a.qkv_S, a.qkv_stride, x);
          sum_partials8(rowp + col + HALF, a.qkv_S, a.qkv_stride, y);
        }
        if (a.qkv_bias) {
          float bb[8];
          load8(a.qkv_bias + col, bb);
#pragma unroll
          for (int i = 0; i < 8; i++) x[i] += bb[i];
          load8(a.qkv_bias + col + HALF, bb);
#pragma unroll
          for (int i = 0; i < 8; i++) y[i] += bb[i];
        }
        if (rot) {
#pragma unroll
          for (int i = 0; i < 8; i++) {
            const float ra = x[i] * c[i] - y[i] * sn[i];
            const float rb = y[i] * c[i] + x[i] * sn[i];
            x[i] = ra;
            y[i] = rb;
          }
        }
        lo = make_uint4(pack_bf16(x[0], x[1]), pack_bf16(x[2], x[3]), pack_bf16(x[4], x[5]),
                        pack_bf16(x[6], x[7]));
        hi = make_uint4(pack_bf16(y[0], y[1]), pack_bf16(y[2], y[3]), pack_bf16(y[4], y[5]),
                        pack_bf16(y[6], y[7]));
      }
      if (is_q) {
        const uint32_t q = sQ + hh * 16 * HD * 2;
        st_shared16(q + swz<HD>(r, j / 8), lo);
        st_shared16(q + swz<HD>(r, (j + HALF) / 8), hi);
      } else {
        bf16* nk = s_newkv + ((size_t)hh * 2 + which) * HD;
        *reinterpret_cast<uint4*>(nk + j) = lo;
        *reinterpret_cast<uint4*>(nk + j + HALF) = hi;
        if (chunk == 0) {  // one CTA per (sequence, kv head) appends to the paged pool
          bf16* pool = which ? (bf16*)a.vpool : (bf16*)a.kpool;
          bf16* dst = pool + (((size_t)bt[pos / KV_BLOCK] * a.nkv + kvh) * KV_BLOCK + (pos % KV_BLOCK)) * HD + j;
          *reinterpret_cast<uint4*>(dst) = lo;
          *reinterpret_cast<uint4*>(dst + HALF) = hi;
        }
      }
    }
  }
  if (!fused) {
#pragma unroll
    for (int s = 0; s < STAGES - 1; s++)
      if (s < F && !tile_is_old(s)) issue_kv(s, s);
  }
  cp_async_commit();

  uint32_t qf[HD / 16][4];
  float o[HD / 8][4];
  float m[2], l[2];
  const int limit[2] = {len - 1, len - 1};
  const float scale_log2 = a.scale * 1.4426950408889634f;
  auto reset_head = [&]() {
#pragma unroll
    for (int dn = 0; dn < HD / 8; dn++)
#pragma unroll
      for (int i = 0; i < 4; i++) o[dn][i] = 0.f;
    m[0] = m[1] = -INFINITY;
    l[0] = l[1] = 0.f;
  };
  // merge the 4 warps (each covered different keys) through shared memory and write head
  // group hh's output (or its split record)
  auto finish_head = [&](int hh) {
#pragma unroll
    for (int i = 0; i < 2; i++) {
      l[i] += __shfl_xor_sync(0xffffffffu, l[i], 1);
      l[i] += __shfl_xor_sync(0xffffffffu, l[i], 2);
    }
    if (kpc == 1) {
      cp_async_wait<0>();
      __syncthreads();  // the ring is drained: `red` aliases it
    }
#pragma unroll
    for (int i = 0; i < 2; i++) {
      if (g + 8 * i >= rrows) continue;  // padding rows of the 16-row MMA tile
      float* rr = red + ((size_t)warp * rrows + g + 8 * i) * RS;
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
    const int h0 = (kvh0 + hh) * G + chunk * 16;
    for (int idx = threadIdx.x; idx < rows * HD; idx += 128) {
      const int r = idx / HD, d = idx % HD;
      float mn = -INFINITY;
#pragma unroll
      for (int w = 0; w < 4; w++) mn = fmaxf(mn, red[((size_t)w * rrows + r) * RS + HD]);
      float lt = 0.f, at = 0.f;
#pragma unroll
      for (int w = 0; w < 4; w++) {
        const float* rr = red + ((size_t)w * rrows + r) * RS;
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
  };

  reset_head();
  const unsigned long long t_prol = tr.buf ? gtime() : 0ull;
  unsigned long long t_first = 0, t_fin0 = 0;
  for (int f = 0; f < F; f++) {
    const int stage = f % STAGES;
    const int hh = f / n_tiles, j = f - hh * n_tiles;
    if (f == 0) cp_async_wait<0>();  // Q and the post-wait tiles are in the newest group
    else cp_async_wait<STAGES - 2>();
    __syncthreads();
    if (tr.buf && f == 0) t_first = gtime();
    if (f + STAGES - 1 < F) issue_kv(f + STAGES - 1, (f + STAGES - 1) % STAGES);
    cp_async_commit();
    if (j == 0) {
      const int mi = lane >> 3, lr = lane & 7;
      const uint32_t q = sQ + hh * 16 * HD * 2;
#pragma unroll
      for (int ks = 0; ks < HD / 16; ks++)
        ldsm_x4(q + swz<HD>(lr + (mi & 1) * 8, ks * 2 + (mi >> 1)), qf[ks][0], qf[ks][1], qf[ks][2],
                qf[ks][3]);
    }
    if (has_new && j == n_tiles - 1) {
      // the tile just landed with whatever the pool held at position len - 1: overwrite that
      // row with the K / V built in the prologue
      const int row = len - 1 - (tile_begin + j) * BN;
      if (threadIdx.x < 2 * (HD / 8)) {
        const int which = threadIdx.x / (HD / 8), c = threadIdx.x % (HD / 8);
        st_shared16((which ? sV : sK) + stage * TILE_BYTES + swz<HD>(row, c),
                    *reinterpret_cast<const uint4*>(s_newkv + ((size_t)hh * 2 + which) * HD + c * 8));
      }
      __syncthreads();
    }
    const int k0 = (tile_begin + j) * BN;
    const bool need_mask = k0 + BN > len;
    attn_block<HD, WK / 8>(qf, sK + stage * TILE_BYTES, sV + stage * TILE_BYTES, warp * WK,
                           k0 + warp * WK, limit, need_mask, scale_log2, m, l, o);
    if (j == n_tiles - 1) {
      if (tr.buf && hh == 0) t_fin0 = gtime();
      finish_head(hh);
      reset_head();
    }
  }
  if (n_tiles == 0) finish_head(0);  // empty split: (0, -inf, 0) records
  cp_async_wait<0>();
  if (threadIdx.x == 0 && trace_on(tr))
    trace_rec(tr, t_entry, t_wait,
              (((t_prol - t_wait) & 0x1fffff) << 42) | (((t_first - t_wait) & 0x1fffff) << 21) |
                  ((t_fin0 - t_wait) & 0x1fffff));
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
  (void)launch_k(attn_prefill_mma_kernel<HD>, dim3(dim3(a.n_qtiles, a.nh)), dim3(128), smem, st, a, next_trace(13));
  return cudaGetLastError();
}
cudaError_t launch_attn_prefill_mma(cudaStream_t st, const AttnArgs& a) {
  if (a.n_qtiles <= 0) return cudaSuccess;
  return a.hd == 64 ? prefill_launch<64>(st, a) : prefill_launch<128>(st, a);
}

template <int HD, int BN, int STAGES>
static cudaError_t decode_launch(cudaStream_t st, AttnArgs a, int G, int chunks) {
  constexpr size_t pipe = (size_t)2 * STAGES * BN * HD * 2;
  const size_t red = (size_t)4 * (G < 16 ? G : 16) * (HD + 2) * 4;
  constexpr size_t smem_cap = 113 * 1024;  // two CTAs per SM
  // pack kpc kv heads per CTA while the grid still covers ~1.6 CTAs per SM (one wave of the
  // 2-per-SM residency) — only for small head groups, unsplit contexts
  int kpc = 1;
  if (chunks == 1 && a.n_splits == 1) {
    int sm = a.sm_count > 0 ? a.sm_count : 148;
    while (kpc < 4 && a.nkv % (kpc * 2) == 0 &&
           (long)a.n_seqs * (a.nkv / (kpc * 2)) * 10 >= (long)sm * 16 &&
           (size_t)(kpc * 2) * (16 * HD * 2 + 2 * HD * 2) + pipe + red <= smem_cap)
      kpc *= 2;
  }
  const size_t q_bytes = (size_t)kpc * 16 * HD * 2;
  const size_t body = kpc > 1 ? pipe + red : (pipe > red ? pipe : red);
  a.newkv_off = (int)(q_bytes + body);
  const size_t smem = q_bytes + body + (size_t)kpc * 2 * HD * 2;
  static size_t smem_set = 0;
  if (smem > smem_set) {
    cudaError_t e = cudaFuncSetAttribute(attn_decode_mma_kernel<HD, BN, STAGES>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    smem_set = smem;
  }
  (void)launch_k(attn_decode_mma_kernel<HD, BN, STAGES>,
                 dim3(a.n_seqs, (a.nkv / kpc) * chunks, a.n_splits), dim3(128), smem, st, a, G,
                 chunks, kpc, next_trace(4));
  return cudaGetLastError();
}
int attn_decode_mma_tile(int hd) {
  return hd == 64 ? 128 : 64;  // the BN of decode_launch below
}
cudaError_t launch_attn_decode_mma(cudaStream_t st, const AttnArgs& a_in) {
  if (a_in.n_seqs <= 0) return cudaSuccess;
  AttnArgs a = a_in;
  const int G = a.nh / a.nkv, chunks = (G + 15) / 16;
  if (a.hd == 64) return decode_launch<64, 128, 3>(st, a, G, chunks);
  return decode_launch<128, 64, 3>(st, a, G, chunks);
}

}  // namespace nvllm
