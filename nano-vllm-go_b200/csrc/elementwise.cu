// elementwise.cu — HBM-bound kernels of the forward pass: embedding gather,
// fused residual-add + LayerNorm/RMSNorm, SwiGLU/GELU, RoPE + paged KV append,
// on-device sampling, weight repack.  All accesses are 16-byte vectors over the
// contiguous (hidden) dimension; one CTA per token row.
#include <cuda_fp16.h>

#include "common.cuh"
#include "moe_route.cuh"
#include "kernels.h"

namespace nvllm {

// ------------------------------------------------------------------ embedding
// generic_model.go:567-592 embedWithOffset + :298-302 EmbeddingMultiplier
template <typename WT>
__global__ void embed_kernel(const int* __restrict__ tokens, const int* __restrict__ positions,
                             const WT* __restrict__ wte, const WT* __restrict__ wpe, int max_pos,
                             float mult, float* __restrict__ x, int H, bf16* __restrict__ xb,
                             float* __restrict__ ssq, const float* __restrict__ g, Trace tr) {
  TRACE_ENTRY(tr);
  pdl_wait();  // step prologue: later kernels may read the step state before their own wait
  pdl_launch_dependents();
  __shared__ float red[32];
  float lsq = 0.f;
  const int t = blockIdx.x;
  const int tok = tokens[t];
  const int pos = positions[t];
  const WT* er = wte + (size_t)tok * H;
  const bool add_pos = (wpe != nullptr) && (pos < max_pos);
  const WT* pr = add_pos ? wpe + (size_t)pos * H : nullptr;
  for (int i = threadIdx.x * 8; i < H; i += blockDim.x * 8) {
    float v[8];
    load8(er + i, v);
    if (add_pos) {
      float p[8];
      load8(pr + i, p);
#pragma unroll
      for (int j = 0; j < 8; j++) v[j] += p[j];
    }
    if (mult != 0.f) {
#pragma unroll
      for (int j = 0; j < 8; j++) v[j] *= mult;
    }
    store8(x + (size_t)t * H + i, v);
    if (xb) {
      float gv[8], o[8];
      load8(g + i, gv);
#pragma unroll
      for (int j = 0; j < 8; j++) {
        lsq += v[j] * v[j];
        o[j] = v[j] * gv[j];
      }
      store8(xb + (size_t)t * H + i, o);
    }
  }
  if (xb) {
    const float tot = block_sum(lsq, red);
    if (threadIdx.x == 0) ssq[t] = tot;
  }
  TRACE_EXIT(tr);
}

cudaError_t launch_embed(cudaStream_t st, int wdt, const int* tokens, const int* positions,
                         const void* wte, const void* wpe, int max_pos, float mult, float* x, int T,
                         int H, void* xb, float* ssq, const float* g) {
  if (T <= 0) return cudaSuccess;
  if ((xb == nullptr) != (ssq == nullptr) || (xb && !g)) return cudaErrorInvalidValue;
  int threads = min(256, max(32, ((H / 8 + 31) / 32) * 32));
  if (wdt == DT_F32)
    (void)launch_k(embed_kernel<float>, dim3(T), dim3(threads), 0, st, tokens, positions, (const float*)wte,
                                               (const float*)wpe, max_pos, mult, x, H, (bf16*)xb, ssq, g, next_trace(5));
  else
    (void)launch_k(embed_kernel<bf16>, dim3(T), dim3(threads), 0, st, tokens, positions, (const bf16*)wte,
                                              (const bf16*)wpe, max_pos, mult, x, H, (bf16*)xb, ssq, g, next_trace(5));
  return cudaGetLastError();
}

// ------------------------------------------------- fused residual add + norm
// tensor.go:193-250 LayerNorm (RMSNorm when bias == nil), fused with the
// residual combine of generic_model.go:320-326 / :359-365 / :405-416:
//   x <- x + m1 * (sum_s p1[s] + bias1) + m2 * (sum_s p2[s] + bias2);  y = norm(x)
// The updated row is staged in shared memory so x is read once and written once.
template <typename YT>
__global__ void __launch_bounds__(512) norm_kernel(NormArgs a, Trace tr) {
  const unsigned long long t_entry = trace_on(tr) ? gtime() : 0ull;
  pdl_launch_dependents();
  extern __shared__ float row[];
  __shared__ float red[32];
  const int H = a.H;
  // norm weights / biases are immutable: fetch the first vector's worth before the
  // dependency resolves (they come from HBM on first touch)
  const int i0 = threadIdx.x * 8;
  float w0[8], b0[8], bias1_0[8], bias2_0[8];
#pragma unroll
  for (int j = 0; j < 8; j++) w0[j] = b0[j] = bias1_0[j] = bias2_0[j] = 0.f;
  if (i0 < H) {
    if (a.y) {
      if (a.w) load8(a.w + i0, w0);
      else
#pragma unroll
        for (int j = 0; j < 8; j++) w0[j] = 1.f;
      if (a.b) load8(a.b + i0, b0);
    }
    if (a.p1 && a.bias1) load8(a.bias1 + i0, bias1_0);
    if (a.p2 && a.bias2) load8(a.bias2 + i0, bias2_0);
  }
  pdl_wait();
  const unsigned long long t_wait = trace_on(tr) ? gtime() : 0ull;
  const int orow = blockIdx.x;
  const int r = a.row_idx ? a.row_idx[orow] - a.row_bias : orow;
  float* xr = a.x + (size_t)r * H;
  const bool upd = (a.p1 != nullptr) || (a.p2 != nullptr);
  float lsum = 0.f;
  for (int i = i0; i < H; i += blockDim.x * 8) {
    float v[8];
    load8(xr + i, v);
    if (a.p1) {
      float acc[8];
      if (a.comb_rows) {
#pragma unroll
        for (int j = 0; j < 8; j++) acc[j] = 0.f;
        // slot order (moe.go:116-118), four slots at a time: the index / weight loads of all
        // four first, then all four row loads, then the ordered accumulation — two L2 round
        // trips per four slots instead of two per slot
        for (int s0 = 0; s0 < a.comb_k; s0 += 4) {
          int prow[4];
          float cw[4], pv[4][8];
#pragma unroll
          for (int q = 0; q < 4; q++)
            if (s0 + q < a.comb_k) {
              prow[q] = a.comb_rows[(size_t)r * a.comb_k + s0 + q];
              cw[q] = a.comb_w[(size_t)r * a.comb_k + s0 + q];
            }
#pragma unroll
          for (int q = 0; q < 4; q++)
            if (s0 + q < a.comb_k) load8(a.p1 + (size_t)prow[q] * H + i, pv[q]);
#pragma unroll
          for (int q = 0; q < 4; q++)
            if (s0 + q < a.comb_k) {
#pragma unroll
              for (int j = 0; j < 8; j++) acc[j] += cw[q] * pv[q][j];
            }
        }
      } else if (a.p1_bf16) {
        load8(reinterpret_cast<const bf16*>(a.p1) + (size_t)r * H + i, acc);
      } else {
        sum_partials8(a.p1 + (size_t)r * H + i, a.S1, a.stride1, acc);
      }
      if (a.bias1) {
        float p[8];
        if (i == i0) {
#pragma unroll
          for (int j = 0; j < 8; j++) p[j] = bias1_0[j];
        } else {
          load8(a.bias1 + i, p);
        }
#pragma unroll
        for (int j = 0; j < 8; j++) acc[j] += p[j];
      }
#pragma unroll
      for (int j = 0; j < 8; j++) v[j] = v[j] + a.m1 * acc[j];
    }
    if (a.p2) {
      float acc[8];
      sum_partials8(a.p2 + (size_t)r * H + i, a.S2, a.stride2, acc);
      if (a.bias2) {
        float p[8];
        if (i == i0) {
#pragma unroll
          for (int j = 0; j < 8; j++) p[j] = bias2_0[j];
        } else {
          load8(a.bias2 + i, p);
        }
#pragma unroll
        for (int j = 0; j < 8; j++) acc[j] += p[j];
      }
#pragma unroll
      for (int j = 0; j < 8; j++) v[j] = v[j] + a.m2 * acc[j];
    }
    if (upd) store8(xr + i, v);
    store8(row + i, v);
    if (a.b) {
#pragma unroll
      for (int j = 0; j < 8; j++) lsum += v[j];
    } else {
#pragma unroll
      for (int j = 0; j < 8; j++) lsum += v[j] * v[j];
    }
  }
  if (a.y == nullptr) return;
  const unsigned long long t_loads = tr.buf ? gtime() : 0ull;
  float tot = block_sum(lsum, red);
  const unsigned long long t_red = tr.buf ? gtime() : 0ull;
  float mean = 0.f, inv;
  if (a.b) {
    mean = tot / (float)H;
    float lv = 0.f;
    for (int i = i0; i < H; i += blockDim.x * 8) {
#pragma unroll
      for (int j = 0; j < 8; j++) {
        float d = row[i + j] - mean;
        lv += d * d;
      }
    }
    float var = block_sum(lv, red) / (float)H;
    inv = 1.0f / sqrtf(var + a.eps);
  } else {
    inv = 1.0f / sqrtf(tot / (float)H + a.eps);
  }
  YT* yr = (YT*)a.y + (size_t)orow * H;
  for (int i = i0; i < H; i += blockDim.x * 8) {
    float w[8], bb[8], o[8];
    if (i == i0) {
#pragma unroll
      for (int j = 0; j < 8; j++) {
        w[j] = w0[j];
        bb[j] = b0[j];
      }
    } else {
      if (a.w) load8(a.w + i, w);
      else
#pragma unroll
        for (int j = 0; j < 8; j++) w[j] = 1.f;
      if (a.b) load8(a.b + i, bb);
    }
    if (a.b) {
#pragma unroll
      for (int j = 0; j < 8; j++) o[j] = (row[i + j] - mean) * inv * w[j] + bb[j];
    } else {
#pragma unroll
      for (int j = 0; j < 8; j++) o[j] = row[i + j] * inv * w[j];
    }
    store8(yr + i, o);
    if (a.y_f32) store8(a.y_f32 + (size_t)orow * H + i, o);
    if (a.router) store8(row + i, o);  // each thread overwrites only what it alone reads
  }
  if (a.router) {
    __shared__ float probs[MOE_MAX_E];
    __syncthreads();
    moe_route_row(row, a.router, H, a.E, a.topk, probs, (size_t)orow, a.topk_idx, a.topk_w);
  }
  if (threadIdx.x == 0 && trace_on(tr))
    trace_rec(tr, t_entry, t_wait, ((t_loads - t_wait) << 32) | (t_red - t_wait));
}

cudaError_t launch_norm(cudaStream_t st, const NormArgs& a) {
  if (a.n_rows <= 0) return cudaSuccess;
  int threads = min(512, max(64, ((a.H / 8 + 31) / 32) * 32));
  // fused router, decode-sized batches (latency-bound: fewer row CTAs than two waves): 16 warps
  // so that each takes two experts of a 32-expert layer in one pass; prefill keeps the
  // occupancy of the narrow CTA (measured: 512 threads cost cfg3 prefill 35 -> 44 ms)
  if (a.router && a.n_rows <= 1024) threads = 512;
  size_t smem = (size_t)a.H * sizeof(float);
  if (a.ydt == DT_F32)
    (void)launch_k(norm_kernel<float>, dim3(a.n_rows), dim3(threads), smem, st, a, next_trace(2));
  else
    (void)launch_k(norm_kernel<bf16>, dim3(a.n_rows), dim3(threads), smem, st, a, next_trace(2));
  return cudaGetLastError();
}

// ---------------------------------------------------------------- activations
template <typename OT>
__global__ void act_kernel(int T, int F, const float* __restrict__ h, int S, size_t stride,
                           const float* __restrict__ bias, int swiglu, OT* __restrict__ out, Trace tr) {
  TRACE_ENTRY(tr);
  pdl_launch_dependents();
  pdl_wait();
  const int ld = swiglu ? 2 * F : F;
  const size_t nvec = (size_t)T * (F / 8);
  for (size_t v = (size_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec;
       v += (size_t)gridDim.x * blockDim.x) {
    const int t = (int)(v / (F / 8));
    const int j = (int)(v % (F / 8)) * 8;
    float g[8], o[8];
#pragma unroll
    for (int i = 0; i < 8; i++) g[i] = 0.f;
    if (!swiglu)
      for (int s = 0; s < S; s++) {
        float p[8];
        load8(h + (size_t)s * stride + (size_t)t * ld + j, p);
#pragma unroll
        for (int i = 0; i < 8; i++) g[i] += p[i];
      }
    if (swiglu) {
      // W1 rows are stored interleaved (weights.cu): column 2j = gate j, 2j + 1 = up j
      float a[8], b[8];
#pragma unroll
      for (int i = 0; i < 8; i++) a[i] = b[i] = 0.f;
      for (int s = 0; s < S; s++) {
        float p[8];
        const float* base = h + (size_t)s * stride + (size_t)t * ld + 2 * j;
        load8(base, p);
#pragma unroll
        for (int i = 0; i < 8; i++) a[i] += p[i];
        load8(base + 8, p);
#pragma unroll
        for (int i = 0; i < 8; i++) b[i] += p[i];
      }
#pragma unroll
      for (int i = 0; i < 4; i++) {
        o[i] = silu(a[2 * i]) * a[2 * i + 1];
        o[4 + i] = silu(b[2 * i]) * b[2 * i + 1];
      }
    } else {
      if (bias) {
        float p[8];
        load8(bias + j, p);
#pragma unroll
        for (int i = 0; i < 8; i++) g[i] += p[i];
      }
#pragma unroll
      for (int i = 0; i < 8; i++) o[i] = gelu_tanh(g[i]);
    }
    store8(out + (size_t)t * F + j, o);
  }
  TRACE_EXIT(tr);
}

cudaError_t launch_act(cudaStream_t st, int odt, int T, int F, const float* h, int S, size_t stride,
                       const float* bias, int swiglu, void* out) {
  if (T <= 0) return cudaSuccess;
  size_t nvec = (size_t)T * (F / 8);
  int threads = 256;
  int blocks = (int)min((size_t)148 * 16, (nvec + threads - 1) / threads);
  if (odt == DT_F32)
    (void)launch_k(act_kernel<float>, dim3(blocks), dim3(threads), 0, st, T, F, h, S, stride, bias, swiglu, (float*)out, next_trace(6));
  else
    (void)launch_k(act_kernel<bf16>, dim3(blocks), dim3(threads), 0, st, T, F, h, S, stride, bias, swiglu, (bf16*)out, next_trace(6));
  return cudaGetLastError();
}

// ------------------------------------------------------- RoPE + paged KV append
// One CTA per token.  Work items are (head, 8-wide vector of the first half):
// rotate_half pairs element i with i + hd/2 (rope.go:187-201).
template <typename QT, typename KT>
__global__ void rope_kv_kernel(int nh, int nkv, int hd, const float* __restrict__ qkv, int S,
                               size_t stride, const float* __restrict__ bias,
                               const int* __restrict__ positions, const int* __restrict__ slots,
                               const float* __restrict__ cosT, const float* __restrict__ sinT,
                               QT* __restrict__ q_out, KT* __restrict__ kpool,
                               KT* __restrict__ vpool, Trace tr) {
  const unsigned long long t_entry = trace_on(tr) ? gtime() : 0ull;
  pdl_launch_dependents();
  const int t = blockIdx.x;
  const int half = hd / 2;
  const int vph = half / 8;  // vectors per half head
  const int ncols = (nh + 2 * nkv) * hd;
  // positions / slots were written by the step's first kernel (which triggers its dependents
  // only after its own wait) or by the host copy: stable here.  Fetch them and the first
  // item's RoPE factors before the dependency on the QKV GEMM resolves.
  const int pos = positions[t];
  const int slot = slots[t];
  const int blk = slot / KV_BLOCK, off = slot % KV_BLOCK;
  const int items = (nh + 2 * nkv) * vph;
  float c0v[8], s0v[8];
#pragma unroll
  for (int i = 0; i < 8; i++) c0v[i] = s0v[i] = 0.f;
  const int it_first = blockIdx.y * blockDim.x + threadIdx.x;
  if (cosT != nullptr && it_first < items) {
    const int j = (it_first % vph) * 8;
    load8(cosT + (size_t)pos * half + j, c0v);
    load8(sinT + (size_t)pos * half + j, s0v);
  }
  pdl_wait();
  const unsigned long long t_wait = trace_on(tr) ? gtime() : 0ull;
  for (int it = blockIdx.y * blockDim.x + threadIdx.x; it < items; it += gridDim.y * blockDim.x) {
    const int head = it / vph;
    const int j = (it % vph) * 8;
    const int c0 = head * hd + j;
    float a[8], b[8];
    const float* base = qkv + (size_t)t * ncols + c0;
    sum_partials8(base, S, stride, a);
    sum_partials8(base + half, S, stride, b);
    if (bias) {
      float p[8];
      load8(bias + c0, p);
#pragma unroll
      for (int i = 0; i < 8; i++) a[i] += p[i];
      load8(bias + c0 + half, p);
#pragma unroll
      for (int i = 0; i < 8; i++) b[i] += p[i];
    }
    const bool is_v = head >= nh + nkv;
    if (cosT != nullptr && !is_v) {
      float c[8], sn[8];
      if (it == it_first) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
          c[i] = c0v[i];
          sn[i] = s0v[i];
        }
      } else {
        load8(cosT + (size_t)pos * half + j, c);
        load8(sinT + (size_t)pos * half + j, sn);
      }
#pragma unroll
      for (int i = 0; i < 8; i++) {
        float ra = a[i] * c[i] - b[i] * sn[i];
        float rb = b[i] * c[i] + a[i] * sn[i];
        a[i] = ra;
        b[i] = rb;
      }
    }
    if (head < nh) {
      QT* dst = q_out + (size_t)t * nh * hd + head * hd + j;
      store8(dst, a);
      store8(dst + half, b);
    } else {
      const int kvh = is_v ? head - nh - nkv : head - nh;
      KT* pool = is_v ? vpool : kpool;
      KT* dst = pool + (((size_t)blk * nkv + kvh) * KV_BLOCK + off) * hd + j;
      store8(dst, a);
      store8(dst + half, b);
    }
  }
  if (threadIdx.x == 0) trace_rec(tr, t_entry, t_wait, 0);
}

cudaError_t launch_rope_kv(cudaStream_t st, int qdt, int kvdt, int T, int nh, int nkv, int hd,
                           const float* qkv, int S, size_t stride, const float* bias,
                           const int* positions, const int* slots, const float* cosT,
                           const float* sinT, void* q_out, void* kpool, void* vpool) {
  if (T <= 0) return cudaSuccess;
  int items = (nh + 2 * nkv) * (hd / 16);
  // few tokens (decode): split each token's heads over several CTAs — a CTA pulls S partial
  // slices of its columns out of L2 and one SM's intake would pace the kernel
  int groups = 1;
  while (T * groups < 256 && items / (groups * 2) >= 32) groups *= 2;
  int per = (items + groups - 1) / groups;
  int threads = min(256, max(32, ((per + 31) / 32) * 32));
#define RK(QT, KT)                                                                              \
  (void)launch_k(rope_kv_kernel<QT, KT>, dim3(T, groups), dim3(threads), 0, st, nh, nkv, hd, qkv, S, stride, bias, positions,   \
                                                slots, cosT, sinT, (QT*)q_out, (KT*)kpool,      \
                                                (KT*)vpool, next_trace(3))
  if (qdt == DT_F32 && kvdt == DT_F32) RK(float, float);
  else if (qdt == DT_F32 && kvdt == DT_BF16) RK(float, bf16);
  else if (qdt == DT_BF16 && kvdt == DT_BF16) RK(bf16, bf16);
  else return cudaErrorInvalidValue;
#undef RK
  return cudaGetLastError();
}

// ------------------------------------------------------------------- sampling
struct ArgMax {
  float v;
  int i;
};
__device__ __forceinline__ ArgMax argmax_merge(ArgMax a, ArgMax b) {
  // larger value wins; equal values keep the lower index (first-max rule)
  if (b.v > a.v || (b.v == a.v && b.i < a.i)) return b;
  return a;
}

// order-preserving float -> uint32 key; the low word ~index makes atomicMax keep the LOWEST
// index among equal values (first-max rule of cmd/ask/main.go:389-402)
__device__ __forceinline__ unsigned long long argmax_key(float v, int i) {
  uint32_t b = __float_as_uint(v);
  b = (b & 0x80000000u) ? ~b : (b | 0x80000000u);
  return ((unsigned long long)b << 32) | (uint32_t)(~(uint32_t)i);
}

// Tensor-parallel LM head: the row's logits are V / vsplit slices gathered rank-major.  One
// CTA (blockIdx.y == 0) per row; same arithmetic and scan order as the plain path below.
__device__ __noinline__ void sample_split_rows(const float* __restrict__ logits, int V, int n,
                                               int vsplit, const float* __restrict__ temperature,
                                               const float* __restrict__ uniforms,
                                               int* __restrict__ out_tokens, float* red,
                                               ArgMax* amred, float* scan, int* winner) {
  if (blockIdx.y != 0) return;
  const int s = blockIdx.x;
  auto at = [&](int v) -> float {
    const int r = v / vsplit;
    return logits[(size_t)r * n * vsplit + (size_t)s * vsplit + (v - r * vsplit)];
  };
  const float temp = temperature ? temperature[s] : 0.f;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  if (!(temp > 0.f)) {
    ArgMax am = {-INFINITY, 0x7fffffff};
    for (int i = threadIdx.x; i < V; i += blockDim.x) {
      const float v = at(i);
      if (v > am.v) { am.v = v; am.i = i; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      ArgMax ot = {__shfl_xor_sync(0xffffffffu, am.v, o), __shfl_xor_sync(0xffffffffu, am.i, o)};
      am = argmax_merge(am, ot);
    }
    if (lane == 0) amred[w] = am;
    __syncthreads();
    if (w == 0) {
      ArgMax r = (lane < nw) ? amred[lane] : ArgMax{-INFINITY, 0x7fffffff};
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        ArgMax ot = {__shfl_xor_sync(0xffffffffu, r.v, o), __shfl_xor_sync(0xffffffffu, r.i, o)};
        r = argmax_merge(r, ot);
      }
      if (lane == 0) out_tokens[s] = (r.i == 0x7fffffff) ? 0 : r.i;
    }
    return;
  }
  const int chunk = (V + blockDim.x - 1) / blockDim.x;
  const int lo = min(V, (int)threadIdx.x * chunk), hi = min(V, lo + chunk);
  float mx = -INFINITY;
  for (int i = lo; i < hi; i++) mx = fmaxf(mx, at(i) / temp);
  mx = block_max(mx, red);
  float csum = 0.f;
  for (int i = lo; i < hi; i++) csum += expf(at(i) / temp - mx);
  scan[threadIdx.x] = csum;
  __syncthreads();
  for (int o = 1; o < (int)blockDim.x; o <<= 1) {
    float add = (threadIdx.x >= (unsigned)o) ? scan[threadIdx.x - o] : 0.f;
    __syncthreads();
    scan[threadIdx.x] += add;
    __syncthreads();
  }
  const float total = scan[blockDim.x - 1];
  const float r = uniforms[s] * total;
  if (threadIdx.x == 0) *winner = V - 1;
  __syncthreads();
  const float incl = scan[threadIdx.x];
  const float excl = incl - csum;
  const bool mine = (incl >= r) && (threadIdx.x == 0 || scan[threadIdx.x - 1] < r) && (hi > lo);
  if (mine) {
    float run = excl;
    int pick = hi - 1;
    for (int i = lo; i < hi; i++) {
      run += expf(at(i) / temp - mx);
      if (run >= r) {
        pick = i;
        break;
      }
    }
    atomicMin(winner, pick);
  }
  __syncthreads();
  if (threadIdx.x == 0) out_tokens[s] = *winner;
}

// grid (n_seqs, chunks).  Greedy rows: every CTA reduces its slice of the vocabulary and
// merges with one 64-bit atomicMax; the last CTA of the row (atomic ticket) publishes the
// token and re-arms the scratch, so the kernel needs no separate init.  Temperature rows: CTA
// chunk 0 runs the whole softmax / inverse-CDF draw.
__global__ void sample_kernel(const float* __restrict__ logits, int ldl, int V,
                              const float* __restrict__ temperature,
                              const float* __restrict__ uniforms, int* __restrict__ out_tokens,
                              unsigned long long* __restrict__ amax, int* __restrict__ ticket,
                              int vsplit, Trace tr) {
  TRACE_ENTRY(tr);
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float red[32];
  __shared__ ArgMax amred[32];
  __shared__ float scan[1024];
  __shared__ int winner;
  const int s = blockIdx.x;
  if (vsplit > 0) {
    sample_split_rows(logits, V, gridDim.x, vsplit, temperature, uniforms, out_tokens, red, amred,
                      scan, &winner);
    TRACE_EXIT(tr);
    return;
  }
  const float* l = logits + (size_t)s * ldl;
  const float temp = temperature ? temperature[s] : 0.f;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  if (!(temp > 0.f)) {
    const int chunks = gridDim.y;
    const int per = ((V + chunks - 1) / chunks + 3) & ~3;
    const int lo = blockIdx.y * per, hi = min(V, lo + per);
    ArgMax am = {-INFINITY, 0x7fffffff};
    // the forward pass has ldl % 8 == 0 and `per` % 4 == 0 (16-byte aligned slices); the
    // single-operator entry point may pass any row stride
    const bool vec = ((reinterpret_cast<uintptr_t>(l + lo)) & 15) == 0;
    for (int i = lo + threadIdx.x * 4; i < hi; i += blockDim.x * 4) {
      if (vec && i + 3 < hi) {
        const float4 v = *reinterpret_cast<const float4*>(l + i);
        if (v.x > am.v) { am.v = v.x; am.i = i; }
        if (v.y > am.v) { am.v = v.y; am.i = i + 1; }
        if (v.z > am.v) { am.v = v.z; am.i = i + 2; }
        if (v.w > am.v) { am.v = v.w; am.i = i + 3; }
      } else {
        for (int j = i; j < min(hi, i + 4); j++) {
          const float v = l[j];
          if (v > am.v) { am.v = v; am.i = j; }
        }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      ArgMax ot = {__shfl_xor_sync(0xffffffffu, am.v, o), __shfl_xor_sync(0xffffffffu, am.i, o)};
      am = argmax_merge(am, ot);
    }
    if (lane == 0) amred[w] = am;
    __syncthreads();
    if (w == 0) {
      ArgMax r = (lane < nw) ? amred[lane] : ArgMax{-INFINITY, 0x7fffffff};
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        ArgMax ot = {__shfl_xor_sync(0xffffffffu, r.v, o), __shfl_xor_sync(0xffffffffu, r.i, o)};
        r = argmax_merge(r, ot);
      }
      if (lane == 0) {
        if (chunks == 1) {
          out_tokens[s] = (r.i == 0x7fffffff) ? 0 : r.i;
        } else {
          if (r.i != 0x7fffffff) atomicMax(amax + s, argmax_key(r.v, r.i));
          __threadfence();
          if (atomicAdd(ticket + s, 1) == chunks - 1) {
            __threadfence();
            const unsigned long long k = atomicExch(amax + s, 0ull);
            out_tokens[s] = k ? (int)(~(uint32_t)(k & 0xffffffffu)) : 0;
            ticket[s] = 0;
          }
        }
      }
    }
    TRACE_EXIT(tr);
    return;
  }
  if (blockIdx.y != 0) return;
  // temperature sampling: sampling.go:70-74 (divide), :105-128 (softmax), :196-217 (inverse CDF).
  // Each thread owns a contiguous chunk so that prefix sums follow index order.
  const int chunk = (V + blockDim.x - 1) / blockDim.x;
  const int lo = min(V, (int)threadIdx.x * chunk), hi = min(V, lo + chunk);
  float mx = -INFINITY;
  for (int i = lo; i < hi; i++) mx = fmaxf(mx, l[i] / temp);
  mx = block_max(mx, red);
  float csum = 0.f;
  for (int i = lo; i < hi; i++) csum += expf(l[i] / temp - mx);
  scan[threadIdx.x] = csum;
  __syncthreads();
  // inclusive scan over <= 1024 chunk sums (Hillis-Steele)
  for (int o = 1; o < (int)blockDim.x; o <<= 1) {
    float add = (threadIdx.x >= (unsigned)o) ? scan[threadIdx.x - o] : 0.f;
    __syncthreads();
    scan[threadIdx.x] += add;
    __syncthreads();
  }
  const float total = scan[blockDim.x - 1];
  const float r = uniforms[s] * total;
  if (threadIdx.x == 0) winner = V - 1;  // sampling.go:212-214 clamp
  __syncthreads();
  const float incl = scan[threadIdx.x];
  const float excl = incl - csum;
  // first chunk whose inclusive prefix reaches r
  const bool mine = (incl >= r) && (threadIdx.x == 0 || scan[threadIdx.x - 1] < r) && (hi > lo);
  if (mine) {
    float run = excl;
    int pick = hi - 1;
    for (int i = lo; i < hi; i++) {
      run += expf(l[i] / temp - mx);
      if (run >= r) {
        pick = i;
        break;
      }
    }
    atomicMin(&winner, pick);
  }
  __syncthreads();
  if (threadIdx.x == 0) out_tokens[s] = winner;
  TRACE_EXIT(tr);
}

cudaError_t launch_sample(cudaStream_t st, const float* logits, int ldl, int V, int n,
                          const float* temperature, const float* uniforms, int* out_tokens,
                          unsigned long long* amax, int* ticket, int vsplit) {
  if (n <= 0) return cudaSuccess;
  // ~8K logits per CTA keeps >= 2 waves of CTAs in flight for a large vocabulary
  const int chunks = (amax && ticket && vsplit == 0) ? max(1, min(64, V / 8192)) : 1;
  (void)launch_k(sample_kernel, dim3(n, chunks), dim3(chunks > 1 ? 256 : 1024), 0, st, logits, ldl,
                 V, temperature, uniforms, out_tokens, amax, ticket, vsplit, next_trace(7));
  return cudaGetLastError();
}

__global__ void advance_kernel(int n, const int* __restrict__ sampled, int* tokens, int* positions,
                               int* slots, int* seq_len, const int* __restrict__ block_tables,
                               int max_blocks, int max_pos, Trace tr) {
  TRACE_ENTRY(tr);
  // Step prologue: later kernels read the step state BEFORE their own griddepcontrol.wait, so
  // the dependents are released only after this kernel's writes (the embedding kernel that
  // follows additionally waits for this grid's completion before it releases its own).
  pdl_wait();
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    tokens[i] = sampled[i];
    int pos = positions[i] + 1;
    if (pos >= max_pos) pos = max_pos - 1;  // the host bounds every sequence; never index past the table
    int bi = pos / KV_BLOCK;
    if (bi >= max_blocks) bi = max_blocks - 1;
    positions[i] = pos;
    seq_len[i] = pos + 1;
    slots[i] = block_tables[(size_t)i * max_blocks + bi] * KV_BLOCK + pos % KV_BLOCK;
  }
  __threadfence();
  pdl_launch_dependents();
  TRACE_EXIT(tr);
}

cudaError_t launch_advance(cudaStream_t st, int n, const int* sampled, int* tokens, int* positions,
                           int* slots, int* seq_len, const int* block_tables, int max_blocks,
                           int max_pos) {
  if (n <= 0) return cudaSuccess;
  (void)launch_k(advance_kernel, dim3((n + 127) / 128), dim3(128), 0, st, n, sampled, tokens, positions, slots, seq_len,
                                                 block_tables, max_blocks, max_pos, next_trace(8));
  return cudaGetLastError();
}

// ------------------------------------------------------------- weight repack
template <typename ST> __device__ __forceinline__ float src_to_f32(ST v);
template <> __device__ __forceinline__ float src_to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float src_to_f32<bf16>(bf16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float src_to_f32<__half>(__half v) { return __half2float(v); }

// dst[n * row_mul + row_add, k] = src[k, n] for k < K, n < N; src rows are ld_src apart (a
// column slice of a wider matrix: tensor-parallel shards).  32x32 tiles through shared memory
// so both sides are coalesced.  row_mul = 2: [gate | up] halves become interleaved rows (gate j
// -> 2j, up j -> 2j + 1) so that a GEMM tile holds matching gate / up rows for the fused
// SwiGLU epilogue.
template <typename ST, typename DT>
__global__ void transpose_convert_kernel(const ST* __restrict__ src, size_t ld_src, int K, int N,
                                         DT* __restrict__ dst, int ld_dst, int row_mul,
                                         int row_add) {
  __shared__ float tile[32][33];
  const int n0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    int k = k0 + r, n = n0 + threadIdx.x;
    tile[r][threadIdx.x] = (k < K && n < N) ? src_to_f32<ST>(src[(size_t)k * ld_src + n]) : 0.f;
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    int n = n0 + r, k = k0 + threadIdx.x;
    if (n < N && k < K)
      dst[((size_t)n * row_mul + row_add) * ld_dst + k] = from_f32<DT>(tile[threadIdx.x][r]);
  }
}

cudaError_t launch_transpose_convert(cudaStream_t st, int sdt_api, const void* src, size_t ld_src,
                                     int K, int N, int ddt, void* dst, int ld_dst, int row_mul,
                                     int row_add) {
  if (K <= 0 || N <= 0) return cudaSuccess;
  dim3 grid((N + 31) / 32, (K + 31) / 32), block(32, 8);
#define TC(ST, DT)                                                                            \
  transpose_convert_kernel<ST, DT><<<grid, block, 0, st>>>((const ST*)src, ld_src, K, N, (DT*)dst, \
                                                           ld_dst, row_mul, row_add)
  if (ddt == DT_F32) {
    if (sdt_api == 0) TC(float, float);
    else if (sdt_api == 1) TC(bf16, float);
    else TC(__half, float);
  } else {
    if (sdt_api == 0) TC(float, bf16);
    else if (sdt_api == 1) TC(bf16, bf16);
    else TC(__half, bf16);
  }
#undef TC
  return cudaGetLastError();
}

// glu_I > 0: src is a stack of [2 glu_I, cols] matrices whose first glu_I rows are GLU gates and
// last glu_I rows the matching ups (HF MoE input_linear, generic_loader.go:578-586); dst
// interleaves them (gate j -> row 2j, up j -> row 2j + 1) so that a 128-row weight tile holds
// complete gate / up pairs for the fused SwiGLU epilogue.
template <typename ST, typename DT>
__global__ void convert_kernel(const ST* __restrict__ src, size_t n, DT* __restrict__ dst, int glu_I,
                               int cols) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x) {
    size_t o = i;
    if (glu_I > 0) {
      const size_t row = i / cols, c = i - row * cols;
      const size_t e = row / (2 * (size_t)glu_I), q = row - e * 2 * glu_I;
      const size_t g = q / glu_I, j = q - g * glu_I;
      o = (e * 2 * glu_I + 2 * j + g) * cols + c;
    }
    dst[o] = from_f32<DT>(src_to_f32<ST>(src[i]));
  }
}

cudaError_t launch_convert(cudaStream_t st, int sdt_api, const void* src, size_t n, int ddt,
                           void* dst, int glu_I, int cols) {
  if (n == 0) return cudaSuccess;
  int blocks = (int)min((size_t)148 * 8, (n + 255) / 256);
#define CV(ST, DT) \
  convert_kernel<ST, DT><<<blocks, 256, 0, st>>>((const ST*)src, n, (DT*)dst, glu_I, cols)
  if (ddt == DT_F32) {
    if (sdt_api == 0) CV(float, float);
    else if (sdt_api == 1) CV(bf16, float);
    else CV(__half, float);
  } else {
    if (sdt_api == 0) CV(float, bf16);
    else if (sdt_api == 1) CV(bf16, bf16);
    else CV(__half, bf16);
  }
#undef CV
  return cudaGetLastError();
}

// dst[(n * row_mul + row_add) * ld_dst + k] = src[n * ld_src + k] for n < N rows, k < K: a row
// range of a [out, in] (PyTorch / safetensors) linear taken as it lies — the device layout has
// the same orientation, so loading is a strided copy + dtype conversion, no transpose
template <typename ST, typename DT>
__global__ void copy_rows_kernel(const ST* __restrict__ src, size_t ld_src, int N, int K,
                                 DT* __restrict__ dst, int ld_dst, int row_mul, int row_add) {
  const size_t total = (size_t)N * K;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (size_t)gridDim.x * blockDim.x) {
    const size_t n = i / K;
    const int k = (int)(i - n * K);
    dst[(n * row_mul + row_add) * ld_dst + k] = from_f32<DT>(src_to_f32<ST>(src[n * ld_src + k]));
  }
}
cudaError_t launch_copy_rows(cudaStream_t st, int sdt_api, const void* src, size_t ld_src, int N,
                             int K, int ddt, void* dst, int ld_dst, int row_mul, int row_add) {
  if (N <= 0 || K <= 0) return cudaSuccess;
  const size_t total = (size_t)N * K;
  int blocks = (int)min((size_t)148 * 8, (total + 255) / 256);
#define CR(ST, DT) \
  copy_rows_kernel<ST, DT><<<blocks, 256, 0, st>>>((const ST*)src, ld_src, N, K, (DT*)dst, ld_dst, row_mul, row_add)
  if (ddt == DT_F32) {
    if (sdt_api == 0) CR(float, float);
    else if (sdt_api == 1) CR(bf16, float);
    else CR(__half, float);
  } else {
    if (sdt_api == 0) CR(float, bf16);
    else if (sdt_api == 1) CR(bf16, bf16);
    else CR(__half, bf16);
  }
#undef CR
  return cudaGetLastError();
}

__global__ void fill_kernel(float* p, size_t n, float v) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x)
    p[i] = v;
}
cudaError_t launch_fill(cudaStream_t st, float* p, size_t n, float v) {
  fill_kernel<<<148 * 8, 256, 0, st>>>(p, n, v);
  return cudaGetLastError();
}

// out[i] = sum_s p[s*stride + i]   (split-K partials of the LM head -> logits)
__global__ void reduce_partials_kernel(const float* __restrict__ p, int S, size_t stride, size_t n,
                                       float* __restrict__ out, Trace tr) {
  TRACE_ENTRY(tr);
  pdl_launch_dependents();
  pdl_wait();
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (size_t)gridDim.x * blockDim.x) {
    float acc = 0.f;
    for (int s = 0; s < S; s++) acc += p[(size_t)s * stride + i];
    out[i] = acc;
  }
  TRACE_EXIT(tr);
}
cudaError_t launch_reduce_partials(cudaStream_t st, const float* p, int S, size_t stride, size_t n,
                                   float* out) {
  if (n == 0) return cudaSuccess;
  int blocks = (int)min((size_t)148 * 8, (n + 255) / 256);
  (void)launch_k(reduce_partials_kernel, dim3(blocks), dim3(256), 0, st, p, S, stride, n, out, next_trace(10));
  return cudaGetLastError();
}

// Build the decode-step state of n sequences from the state of the prefill step that just
// sampled `sampled`: next position = seq_len, KV slot through the block table.
__global__ void make_decode_state_kernel(int n, int mb, const int* __restrict__ sampled,
                                         const int* __restrict__ src_seq_len,
                                         const int* __restrict__ src_bt,
                                         const int* __restrict__ src_temp,
                                         const int* __restrict__ src_uni, int* tokens,
                                         int* positions, int* slots, int* cu_q, int* seq_start,
                                         int* seq_len, int* last_row, int* bt, int* temp, int* uni,
                                         Trace tr) {
  TRACE_ENTRY(tr);
  // step prologue: dependents are released after the writes below (see advance_kernel)
  pdl_wait();
  const int i = blockIdx.x;
  for (int b = threadIdx.x; b < mb; b += blockDim.x) bt[(size_t)i * mb + b] = src_bt[(size_t)i * mb + b];
  if (threadIdx.x == 0) {
    const int pos = src_seq_len[i];
    int bi = pos / KV_BLOCK;
    if (bi >= mb) bi = mb - 1;  // the host checks the range before launching
    tokens[i] = sampled[i];
    positions[i] = pos;
    slots[i] = src_bt[(size_t)i * mb + bi] * KV_BLOCK + pos % KV_BLOCK;
    cu_q[i] = i;
    if (i == n - 1) cu_q[n] = n;
    seq_start[i] = pos;
    seq_len[i] = pos + 1;
    last_row[i] = i;
    temp[i] = src_temp[i];
    uni[i] = src_uni[i];
  }
  __threadfence();
  __syncthreads();
  pdl_launch_dependents();
  TRACE_EXIT(tr);
}
cudaError_t launch_make_decode_state(cudaStream_t st, int n, int mb, const int* sampled,
                                     const int* src_seq_len, const int* src_bt, const int* src_temp,
                                     const int* src_uni, int* tokens, int* positions, int* slots,
                                     int* cu_q, int* seq_start, int* seq_len, int* last_row, int* bt,
                                     int* temp, int* uni) {
  if (n <= 0) return cudaSuccess;
  (void)launch_k(make_decode_state_kernel, dim3(n), dim3(64), 0, st, n, mb, sampled, src_seq_len, src_bt, src_temp, src_uni,
                                             tokens, positions, slots, cu_q, seq_start, seq_len,
                                             last_row, bt, temp, uni, next_trace(8));
  return cudaGetLastError();
}

}  // namespace nvllm
