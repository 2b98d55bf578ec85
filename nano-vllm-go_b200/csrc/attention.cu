// attention.cu — paged attention (MHA / GQA / MQA) reading K/V straight through
// the scheduler's block tables (block = 256 positions, nanovllm/sequence.go:52).
//
// Semantics follow the reference: scores = (q.k) * scale, causal mask
// j > kv_len - q_len + i (attention.go:164-169,374-380, mqa.go:216-222), row
// softmax, P.V.  The reference writes -1e10 into masked scores; exp(-1e10 - max)
// is exactly 0 in fp32, so skipping those keys is identical.  q head h uses kv
// head h / (n_heads / n_kv_heads) (attention.go:337-339).
//
//  * attn_prefill_kernel: flash-style (online softmax, no S x S score tensor).
//    CTA = 32 queries of one sequence x one head; 4 threads per query each owning
//    hd/4 dims; K/V tiles of 64 keys staged in shared memory as fp32.
//  * attn_decode_kernel: one new token per sequence.  CTA = (sequence, group of
//    GC query heads sharing a kv head, KV split).  HBM-bound: each K/V row is
//    read once with 16-byte loads by hd/8 lanes; online softmax in registers;
//    split partials merged by attn_combine_kernel.
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"

namespace nvllm {

constexpr int PF_KT = 64;  // keys per shared-memory tile (divides KV_BLOCK)

template <typename QT, typename KT, int HD>
__global__ void __launch_bounds__(128)
attn_prefill_kernel(AttnArgs a) {
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ float smem[];
  float* Ks = smem;               // [PF_KT][HD]
  float* Vs = smem + PF_KT * HD;  // [PF_KT][HD]
  constexpr int DPT = HD / 4;     // dims per thread
  constexpr int NV = DPT / 4;     // float4 per thread
  const int tile = blockIdx.x;
  const int head = blockIdx.y;
  const int seq = a.tile_seq[tile];
  const int q0 = a.tile_q0[tile];
  const int n_new = a.cu_q[seq + 1] - a.cu_q[seq];
  const int start = a.seq_start[seq];
  const int kvh = head / (a.nh / a.nkv);
  const int qi = threadIdx.x >> 2, sub = threadIdx.x & 3;
  const int qidx = q0 + qi;  // index among the new tokens of this sequence
  const bool qvalid = qidx < n_new;
  const int qpos = start + qidx;
  const int row = a.cu_q[seq] + qidx;
  const int* bt = a.block_tables + (size_t)seq * a.max_blocks;

  float q[DPT], acc[DPT];
#pragma unroll
  for (int i = 0; i < DPT; i++) acc[i] = 0.f;
  if (qvalid) {
    const QT* qp = (const QT*)a.q + (size_t)row * a.nh * HD + head * HD;
#pragma unroll
    for (int v = 0; v < NV; v++)
#pragma unroll
      for (int j = 0; j < 4; j++) q[v * 4 + j] = to_f32(qp[v * 16 + sub * 4 + j]) * a.scale;
  } else {
#pragma unroll
    for (int i = 0; i < DPT; i++) q[i] = 0.f;
  }
  float m = -INFINITY, l = 0.f;
  const int kv_end = min(start + q0 + ATTN_PREFILL_QTILE, start + n_new);  // keys [0, kv_end)
  const KT* kpool = (const KT*)a.kpool;
  const KT* vpool = (const KT*)a.vpool;

  for (int k0 = 0; k0 < kv_end; k0 += PF_KT) {
    const int nk = min(PF_KT, kv_end - k0);
    const int blk = bt[k0 / KV_BLOCK];
    const size_t base = (((size_t)blk * a.nkv + kvh) * KV_BLOCK + (k0 % KV_BLOCK)) * HD;
    __syncthreads();
    for (int i = threadIdx.x * 8; i < nk * HD; i += blockDim.x * 8) {
      float v[8];
      load8(kpool + base + i, v);
      store8(Ks + i, v);
      load8(vpool + base + i, v);
      store8(Vs + i, v);
    }
    __syncthreads();
    // causal: key position <= query position.  The trip count must be warp-uniform (the
    // shuffles below name all 32 lanes), so every lane runs to the warp's largest jmax and
    // masks its own tail.
    const int jmax = qvalid ? max(0, min(nk, qpos - k0 + 1)) : 0;
    const int jlim = __reduce_max_sync(0xffffffffu, jmax);
    for (int j0 = 0; j0 < jlim; j0 += 8) {
      float s[8];
      float cmax = -INFINITY;
#pragma unroll
      for (int jj = 0; jj < 8; jj++) {
        const int j = j0 + jj;
        float d = 0.f;
        if (j < jmax) {
          const float* kr = Ks + j * HD + sub * 4;
#pragma unroll
          for (int v = 0; v < NV; v++) {
            float4 k4 = *reinterpret_cast<const float4*>(kr + v * 16);
            d = fmaf(q[v * 4 + 0], k4.x, d);
            d = fmaf(q[v * 4 + 1], k4.y, d);
            d = fmaf(q[v * 4 + 2], k4.z, d);
            d = fmaf(q[v * 4 + 3], k4.w, d);
          }
        }
        d += __shfl_xor_sync(0xffffffffu, d, 1);
        d += __shfl_xor_sync(0xffffffffu, d, 2);
        s[jj] = (j < jmax) ? d : -INFINITY;
        cmax = fmaxf(cmax, s[jj]);
      }
      if (cmax == -INFINITY) continue;  // nothing visible to this query in the chunk (no shuffles below)
      const float m_new = fmaxf(m, cmax);
      const float corr = __expf(m - m_new);  // m = -inf on the first chunk -> 0
      l *= corr;
#pragma unroll
      for (int i = 0; i < DPT; i++) acc[i] *= corr;
#pragma unroll
      for (int jj = 0; jj < 8; jj++) {
        const int j = j0 + jj;
        if (j < jmax) {
          const float p = expf(s[jj] - m_new);
          l += p;
          const float* vr = Vs + j * HD + sub * 4;
#pragma unroll
          for (int v = 0; v < NV; v++) {
            float4 v4 = *reinterpret_cast<const float4*>(vr + v * 16);
            acc[v * 4 + 0] = fmaf(p, v4.x, acc[v * 4 + 0]);
            acc[v * 4 + 1] = fmaf(p, v4.y, acc[v * 4 + 1]);
            acc[v * 4 + 2] = fmaf(p, v4.z, acc[v * 4 + 2]);
            acc[v * 4 + 3] = fmaf(p, v4.w, acc[v * 4 + 3]);
          }
        }
      }
      m = m_new;
    }
  }
  if (qvalid) {
    const float inv = 1.0f / l;
    QT* op = (QT*)a.out + (size_t)row * a.nh * HD + head * HD;
#pragma unroll
    for (int v = 0; v < NV; v++)
#pragma unroll
      for (int j = 0; j < 4; j++) op[v * 16 + sub * 4 + j] = from_f32<QT>(acc[v * 4 + j] * inv);
  }
}

template <typename QT, typename KT>
static cudaError_t prefill_dispatch(cudaStream_t st, const AttnArgs& a) {
  dim3 grid(a.n_qtiles, a.nh);
  size_t smem = (size_t)2 * PF_KT * a.hd * sizeof(float);
#define PF(HD)                                                                               \
  {                                                                                          \
    if (smem > 48 * 1024)                                                                    \
      cudaFuncSetAttribute(attn_prefill_kernel<QT, KT, HD>,                                  \
                           cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);          \
    (void)launch_k(attn_prefill_kernel<QT, KT, HD>, dim3(grid), dim3(128), smem, st, a);                             \
  }
  switch (a.hd) {
    case 16: PF(16); break;
    case 32: PF(32); break;
    case 64: PF(64); break;
    case 128: PF(128); break;
    default: return cudaErrorInvalidValue;
  }
#undef PF
  return cudaGetLastError();
}

int attn_prefill_qtile(int qdt, int kvdt, int hd) {
  if (attn_mma_supported(qdt, kvdt, hd)) return attn_prefill_tc5_enabled() ? 128 : 64;
  return ATTN_PREFILL_QTILE;
}

int attn_decode_splits(int qdt, int kvdt, int n_seqs, int nh, int nkv, int hd, int max_len,
                       int sm_count) {
  int ctas, max_splits;
  if (attn_mma_supported(qdt, kvdt, hd)) {
    const int G = nh / nkv;
    ctas = n_seqs * nkv * ((G + 15) / 16);
    max_splits = (max_len + attn_decode_mma_tile(hd) - 1) / attn_decode_mma_tile(hd);
  } else {
    const int G = nh / nkv;
    int gc = 1;
    for (int c = 4; c >= 1; c--)
      if (G % c == 0) {
        gc = c;
        break;
      }
    ctas = n_seqs * (nh / gc);
    max_splits = (max_len + 255) / 256;
  }
  int ns = (2 * sm_count + ctas - 1) / ctas;
  ns = min(ns, max_splits);
  return max(1, min(ns, 32));
}

cudaError_t launch_attn_prefill(cudaStream_t st, const AttnArgs& a) {
  if (a.n_qtiles <= 0) return cudaSuccess;
  if (attn_mma_supported(a.qdt, a.kvdt, a.hd)) {
    if (attn_prefill_tc5_enabled() && a.pool_blocks > 0 && a.total_rows > 0)
      return launch_attn_prefill_tc5(st, a);
    return launch_attn_prefill_mma(st, a);
  }
  if (a.qdt == DT_F32 && a.kvdt == DT_F32) return prefill_dispatch<float, float>(st, a);
  if (a.qdt == DT_F32 && a.kvdt == DT_BF16) return prefill_dispatch<float, bf16>(st, a);
  if (a.qdt == DT_BF16 && a.kvdt == DT_BF16) return prefill_dispatch<bf16, bf16>(st, a);
  return cudaErrorInvalidValue;
}

// ------------------------------------------------------------------- decode
// split partial record: [hd] unnormalised accumulator, then m, l
template <typename QT, typename KT, int HD, int GC>
__global__ void __launch_bounds__(128)
attn_decode_kernel(AttnArgs a) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int LPT = HD / 8;    // lanes per token row
  constexpr int TPI = 32 / LPT;  // token rows per warp iteration
  constexpr int NW = 4;
  __shared__ float sm_acc[NW][GC][HD];
  __shared__ float sm_ml[NW][GC][2];
  const int seq = blockIdx.x;
  const int hg = blockIdx.y;  // group of GC consecutive q heads
  const int split = blockIdx.z;
  const int h0 = hg * GC;
  const int kvh = h0 / (a.nh / a.nkv);
  const int len = a.seq_len[seq];
  int chunk = (len + a.n_splits - 1) / a.n_splits;
  chunk = ((chunk + 63) / 64) * 64;
  const int t0 = split * chunk;
  const int t1 = min(len, t0 + chunk);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int slot = lane / LPT, sub = lane % LPT;
  const int* bt = a.block_tables + (size_t)seq * a.max_blocks;
  const KT* kpool = (const KT*)a.kpool;
  const KT* vpool = (const KT*)a.vpool;

  float q[GC][8], acc[GC][8], m[GC], l[GC];
  const QT* qp = (const QT*)a.q + (size_t)seq * a.nh * HD + h0 * HD + sub * 8;
#pragma unroll
  for (int g = 0; g < GC; g++) {
    load8(qp + g * HD, q[g]);
#pragma unroll
    for (int i = 0; i < 8; i++) {
      q[g][i] *= a.scale;
      acc[g][i] = 0.f;
    }
    m[g] = -INFINITY;
    l[g] = 0.f;
  }
  // warp-uniform trip count (the shuffles name all 32 lanes); lanes past t1 are masked
  for (int tb = t0 + warp * TPI; tb < t1; tb += NW * TPI) {
    const int t = tb + slot;
    const bool valid = t < t1;
    const int tc = valid ? t : t1 - 1;
    const int blk = bt[tc / KV_BLOCK];
    const size_t off = (((size_t)blk * a.nkv + kvh) * KV_BLOCK + (tc % KV_BLOCK)) * HD + sub * 8;
    float k[8], v[8];
    load8(kpool + off, k);
    load8(vpool + off, v);
#pragma unroll
    for (int g = 0; g < GC; g++) {
      float d = 0.f;
#pragma unroll
      for (int i = 0; i < 8; i++) d = fmaf(q[g][i], k[i], d);
#pragma unroll
      for (int o = 1; o < LPT; o <<= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
      if (!valid) continue;
      if (d > m[g]) {  // rescale only when the running max moves
        const float corr = __expf(m[g] - d);
        l[g] *= corr;
#pragma unroll
        for (int i = 0; i < 8; i++) acc[g][i] *= corr;
        m[g] = d;
      }
      const float p = expf(d - m[g]);
      l[g] += p;
#pragma unroll
      for (int i = 0; i < 8; i++) acc[g][i] = fmaf(p, v[i], acc[g][i]);
    }
  }
  // merge the TPI token slots of the warp (lanes with equal `sub`)
#pragma unroll
  for (int g = 0; g < GC; g++) {
#pragma unroll
    for (int o = LPT; o < 32; o <<= 1) {
      const float mo = __shfl_xor_sync(0xffffffffu, m[g], o);
      const float lo = __shfl_xor_sync(0xffffffffu, l[g], o);
      const float mn = fmaxf(m[g], mo);
      const float c0 = (m[g] == -INFINITY) ? 0.f : __expf(m[g] - mn);
      const float c1 = (mo == -INFINITY) ? 0.f : __expf(mo - mn);
      l[g] = l[g] * c0 + lo * c1;
#pragma unroll
      for (int i = 0; i < 8; i++) {
        const float ao = __shfl_xor_sync(0xffffffffu, acc[g][i], o);
        acc[g][i] = acc[g][i] * c0 + ao * c1;
      }
      m[g] = mn;
    }
    if (slot == 0) {
#pragma unroll
      for (int i = 0; i < 8; i++) sm_acc[warp][g][sub * 8 + i] = acc[g][i];
      if (sub == 0) {
        sm_ml[warp][g][0] = m[g];
        sm_ml[warp][g][1] = l[g];
      }
    }
  }
  __syncthreads();
  // merge the 4 warps; thread d handles dim d of every head of the group
  for (int idx = threadIdx.x; idx < GC * HD; idx += blockDim.x) {
    const int g = idx / HD, d = idx % HD;
    float mn = -INFINITY;
#pragma unroll
    for (int w = 0; w < NW; w++) mn = fmaxf(mn, sm_ml[w][g][0]);
    float lt = 0.f, at = 0.f;
#pragma unroll
    for (int w = 0; w < NW; w++) {
      const float mw = sm_ml[w][g][0];
      const float c = (mw == -INFINITY) ? 0.f : __expf(mw - mn);
      lt += sm_ml[w][g][1] * c;
      at += sm_acc[w][g][d] * c;
    }
    const int h = h0 + g;
    if (a.n_splits == 1) {
      ((QT*)a.out)[(size_t)seq * a.nh * HD + h * HD + d] = from_f32<QT>(at / lt);
    } else {
      float* rec = a.ws + (((size_t)seq * a.nh + h) * a.n_splits + split) * (HD + 2);
      rec[d] = at;
      if (d == 0) {
        rec[HD] = mn;
        rec[HD + 1] = lt;
      }
    }
  }
}

template <typename QT>
__global__ void attn_combine_kernel(AttnArgs a, Trace tr) {
  TRACE_ENTRY(tr);
  pdl_launch_dependents();
  pdl_wait();
  const int seq = blockIdx.x, h = blockIdx.y, HD = a.hd;
  const float* rec0 = a.ws + (((size_t)seq * a.nh + h) * a.n_splits) * (HD + 2);
  float mn = -INFINITY;
  for (int s = 0; s < a.n_splits; s++) mn = fmaxf(mn, rec0[(size_t)s * (HD + 2) + HD]);
  for (int d = threadIdx.x; d < HD; d += blockDim.x) {
    float lt = 0.f, at = 0.f;
    for (int s = 0; s < a.n_splits; s++) {
      const float* rec = rec0 + (size_t)s * (HD + 2);
      const float ms = rec[HD];
      const float c = (ms == -INFINITY) ? 0.f : __expf(ms - mn);
      lt += rec[HD + 1] * c;
      at += rec[d] * c;
    }
    ((QT*)a.out)[(size_t)seq * a.nh * HD + h * HD + d] = from_f32<QT>(at / lt);
  }
  TRACE_EXIT(tr);
}

template <typename QT, typename KT, int HD>
static cudaError_t decode_dispatch_gc(cudaStream_t st, const AttnArgs& a) {
  const int G = a.nh / a.nkv;
  int gc = 1;
  for (int c = 4; c >= 1; c--)
    if (G % c == 0) {
      gc = c;
      break;
    }
  dim3 grid(a.n_seqs, a.nh / gc, a.n_splits);
  switch (gc) {
    case 4: (void)launch_k(attn_decode_kernel<QT, KT, HD, 4>, dim3(grid), dim3(128), 0, st, a); break;
    case 3: (void)launch_k(attn_decode_kernel<QT, KT, HD, 3>, dim3(grid), dim3(128), 0, st, a); break;
    case 2: (void)launch_k(attn_decode_kernel<QT, KT, HD, 2>, dim3(grid), dim3(128), 0, st, a); break;
    default: (void)launch_k(attn_decode_kernel<QT, KT, HD, 1>, dim3(grid), dim3(128), 0, st, a); break;
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  if (a.n_splits > 1) {
    (void)launch_k(attn_combine_kernel<QT>, dim3(dim3(a.n_seqs, a.nh)), dim3(min(128, a.hd)), 0, st, a, next_trace(14));
    e = cudaGetLastError();
  }
  return e;
}
template <typename QT, typename KT>
static cudaError_t decode_dispatch(cudaStream_t st, const AttnArgs& a) {
  switch (a.hd) {
    case 16: return decode_dispatch_gc<QT, KT, 16>(st, a);
    case 32: return decode_dispatch_gc<QT, KT, 32>(st, a);
    case 64: return decode_dispatch_gc<QT, KT, 64>(st, a);
    case 128: return decode_dispatch_gc<QT, KT, 128>(st, a);
  }
  return cudaErrorInvalidValue;
}

// true: launch_attn_decode takes the QKV GEMM partials (AttnArgs::qkv_p) and does the RoPE +
// KV append of the decoded token itself; the caller then skips launch_rope_kv
bool attn_decode_fuses_rope(int qdt, int kvdt, int hd) {
  static const bool on = [] {
    const char* e = getenv("NVLLM_FUSE_ROPE");
    return e ? atoi(e) != 0 : true;
  }();
  return on && attn_mma_supported(qdt, kvdt, hd);
}

cudaError_t launch_attn_decode(cudaStream_t st, const AttnArgs& a) {
  if (a.n_seqs <= 0) return cudaSuccess;
  if (a.qkv_p && !attn_decode_fuses_rope(a.qdt, a.kvdt, a.hd)) return cudaErrorInvalidValue;
  if (attn_mma_supported(a.qdt, a.kvdt, a.hd)) {
    // small head groups: TMA-fed, keys-as-M-rows kernel (attention_dec.cu); Falcon's 71-head
    // MQA group keeps the 16-query-row kernel
    cudaError_t e = attn_decode_t_supported(a) ? launch_attn_decode_t(st, a)
                                               : launch_attn_decode_mma(st, a);
    if (e != cudaSuccess || a.n_splits == 1) return e;
    (void)launch_k(attn_combine_kernel<bf16>, dim3(dim3(a.n_seqs, a.nh)), dim3(min(128, a.hd)), 0, st, a, next_trace(14));
    return cudaGetLastError();
  }
  if (a.qdt == DT_F32 && a.kvdt == DT_F32) return decode_dispatch<float, float>(st, a);
  if (a.qdt == DT_F32 && a.kvdt == DT_BF16) return decode_dispatch<float, bf16>(st, a);
  if (a.qdt == DT_BF16 && a.kvdt == DT_BF16) return decode_dispatch<bf16, bf16>(st, a);
  return cudaErrorInvalidValue;
}

}  // namespace nvllm
