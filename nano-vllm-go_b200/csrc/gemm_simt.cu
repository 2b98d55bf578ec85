// gemm_simt.cu — CUDA-core matrix kernels.
//
//  * gemv_kernel: skinny (M <= 8) weight-streaming split-K GEMV for single-sequence decode.
//    Weights [N,K] are read exactly once with 16-byte streaming loads (L1 no-allocate), one
//    warp per R output rows, activations staged in shared memory as fp32, fp32 accumulate.
//    HBM-bound: algorithmic bytes = N*K*sizeof(W).
//  * gemm_simt_kernel: 64x64x16 register-tiled fp32-FMA GEMM.  This is the fp32 PARITY-mode
//    engine (1e-4 gate: TF32/bf16 tensor cores are not accurate enough) and the grouped
//    (MoE expert) GEMM with row gather.  The bf16 throughput path is gemm_tcgen05.cu.
//
// Both compute C[m,n] = alpha * sum_k A[m,k] * W[n,k]  (reference tensor.go:62-88 MatMul with
// the weight held [out,in]) and write fp32 split-K partials.
#include "common.cuh"
#include "kernels.h"

namespace nvllm {

// ------------------------------------------------------------------ GEMV
struct GemvPlan {
  int warps, R, KC, S;
};
static GemvPlan gemv_plan(int M, int N, int K) {
  GemvPlan p;
  p.warps = 8;
  p.R = (M > 2) ? 4 : 2;
  int kc_max = (M > 4) ? 1024 : 2048;  // fp32 activations in smem: M*KC*4 <= 32 KB
  p.KC = kc_max;
  auto ctas = [&](int R, int KC) { return ceil_div(N, p.warps * R) * ceil_div(K, KC); };
  // fill 148 SMs x 2 CTAs when the matrix is small: first shrink the K chunk, then R
  while (ctas(p.R, p.KC) < 296 && p.KC > 512) p.KC /= 2;
  while (ctas(p.R, p.KC) < 296 && p.R > 1) p.R /= 2;
  if (p.KC > K) p.KC = ceil_div(K, 256) * 256;
  p.S = ceil_div(K, p.KC);
  return p;
}
int gemv_splits(int M, int N, int K) { return gemv_plan(M, N, K).S; }

// 8 weight elements as loaded (the conversion to fp32 is deferred to the use)
template <typename WT> struct Raw8;
template <> struct Raw8<bf16> {
  uint4 v;
  __device__ __forceinline__ void ld(const bf16* p) {
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  }
  __device__ __forceinline__ void cvt(float (&o)[8]) const {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
    for (int i = 0; i < 4; i++) {
      float2 f = __bfloat1622float2(h[i]);
      o[2 * i] = f.x;
      o[2 * i + 1] = f.y;
    }
  }
};
template <> struct Raw8<float> {
  float4 a, b;
  __device__ __forceinline__ void ld(const float* p) {
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w) : "l"(p));
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p + 4));
  }
  __device__ __forceinline__ void cvt(float (&o)[8]) const {
    o[0] = a.x; o[1] = a.y; o[2] = a.z; o[3] = a.w;
    o[4] = b.x; o[5] = b.y; o[6] = b.z; o[7] = b.w;
  }
};

// The weights are immutable: every warp issues the loads of its first GEMV_PF k-iterations
// BEFORE griddepcontrol.wait, so the HBM stream of this GEMV runs under the tail of the kernel
// in front of it (a single-sequence decode step is a chain of ~100 such dependent launches).
constexpr int GEMV_PF = 2;
template <typename AT, typename WT, int M, int R>
__global__ void __launch_bounds__(256)
gemv_kernel(const AT* __restrict__ A, int lda, const WT* __restrict__ W, int N, int K, int KC,
            float* __restrict__ out, int ldo, size_t split_stride, float alpha, GemvPre pre, Trace tr) {
  TRACE_ENTRY(tr);
  pdl_launch_dependents();
  extern __shared__ float a_s[];  // [M][KC]
  const int ks = blockIdx.y * KC;
  const int ke = min(K, ks + KC);
  const int klen = ke - ks;
  Raw8<WT> wpre[GEMV_PF][R];
  {
    const int lane_ = threadIdx.x & 31, warp_ = threadIdx.x >> 5, nw_ = blockDim.x >> 5;
    const int n0_ = (blockIdx.x * nw_ + warp_) * R;
    if (n0_ < N) {
#pragma unroll
      for (int it = 0; it < GEMV_PF; it++) {
        const int k = lane_ * 8 + it * 256;
        if (k < klen) {
#pragma unroll
          for (int r = 0; r < R; r++) wpre[it][r].ld(W + (size_t)min(n0_ + r, N - 1) * K + ks + k);
        }
      }
    }
  }
  pdl_wait();
  if (pre.mode == 0) {
    for (int i = threadIdx.x * 8; i < M * KC; i += blockDim.x * 8) {
      const int m = i / KC, k = i % KC;
      float v[8];
      if (k < klen) {
        load8(A + (size_t)m * lda + ks + k, v);
      } else {
#pragma unroll
        for (int j = 0; j < 8; j++) v[j] = 0.f;
      }
      store8(a_s + i, v);
    }
  } else {
    // the activation kernel's work (act_kernel): SwiGLU over interleaved gate / up columns or
    // tanh-GELU of the FFN up projection's split-K slices (transformer.go:50-78)
    for (int i = threadIdx.x; i < M * KC; i += blockDim.x) {
      const int m = i / KC, k = i % KC;
      float y = 0.f;
      if (k < klen) {
        const int col = ks + k;
        if (pre.swiglu) {
          float g = 0.f, u = 0.f;
          for (int sl = 0; sl < pre.Sh; sl++) {
            const float* hrow = pre.hp + (size_t)sl * pre.hstride + (size_t)m * pre.hld + 2 * col;
            g += hrow[0];
            u += hrow[1];
          }
          y = silu(g) * u;
        } else {
          float hsum = 0.f;
          for (int sl = 0; sl < pre.Sh; sl++) hsum += pre.hp[(size_t)sl * pre.hstride + (size_t)m * pre.hld + col];
          if (pre.hbias) hsum += pre.hbias[col];
          y = gelu_tanh(hsum);
        }
        y = to_f32(from_f32<AT>(y));
      }
      a_s[i] = y;
    }
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const int n0 = (blockIdx.x * nwarps + warp) * R;
  if (n0 >= N) return;
  float acc[R][M];
#pragma unroll
  for (int r = 0; r < R; r++)
#pragma unroll
    for (int m = 0; m < M; m++) acc[r][m] = 0.f;
  const WT* wrow[R];
#pragma unroll
  for (int r = 0; r < R; r++) wrow[r] = W + (size_t)min(n0 + r, N - 1) * K + ks;

  auto fma_block = [&](const float (&w)[R][8], int k) {
#pragma unroll
    for (int m = 0; m < M; m++) {
      float a[8];
      load8(a_s + m * KC + k, a);
#pragma unroll
      for (int r = 0; r < R; r++)
#pragma unroll
        for (int j = 0; j < 8; j++) acc[r][m] = fmaf(w[r][j], a[j], acc[r][m]);
    }
  };
#pragma unroll
  for (int it = 0; it < GEMV_PF; it++) {
    const int k = lane * 8 + it * 256;
    if (k < klen) {
      float w[R][8];
#pragma unroll
      for (int r = 0; r < R; r++) wpre[it][r].cvt(w[r]);
      fma_block(w, k);
    }
  }
#pragma unroll 2
  for (int k = lane * 8 + GEMV_PF * 256; k < klen; k += 256) {
    float w[R][8];
#pragma unroll
    for (int r = 0; r < R; r++) load8_stream(wrow[r] + k, w[r]);
#pragma unroll
    for (int m = 0; m < M; m++) {
      float a[8];
      load8(a_s + m * KC + k, a);
#pragma unroll
      for (int r = 0; r < R; r++)
#pragma unroll
        for (int j = 0; j < 8; j++) acc[r][m] = fmaf(w[r][j], a[j], acc[r][m]);
    }
  }
#pragma unroll
  for (int r = 0; r < R; r++)
#pragma unroll
    for (int m = 0; m < M; m++) {
      float v = warp_sum(acc[r][m]);
      if (lane == 0 && n0 + r < N)
        out[(size_t)blockIdx.y * split_stride + (size_t)m * ldo + n0 + r] = alpha * v;
    }
  TRACE_EXIT(tr);
}

template <typename AT, typename WT, int M>
static cudaError_t gemv_dispatch_r(cudaStream_t st, const GemmArgs& g, const GemvPlan& p, const GemvPre& pre) {
  dim3 grid(ceil_div(g.N, p.warps * p.R), p.S);
  size_t smem = (size_t)M * p.KC * sizeof(float);
#define GV(RR)                                                                                     \
  (void)launch_k(gemv_kernel<AT, WT, M, RR>, dim3(grid), dim3(p.warps * 32), smem, st,                                     \
      (const AT*)g.A, g.lda, (const WT*)g.W, g.N, g.K, p.KC, g.out, g.ldo, g.split_stride, g.alpha, \
      pre, next_trace(11))
  if (p.R == 4) GV(4);
  else if (p.R == 2) GV(2);
  else GV(1);
#undef GV
  return cudaGetLastError();
}
template <typename AT, typename WT>
static cudaError_t gemv_dispatch_m(cudaStream_t st, const GemmArgs& g, const GemvPlan& p, const GemvPre& pre) {
  switch (g.M) {
    case 1: return gemv_dispatch_r<AT, WT, 1>(st, g, p, pre);
    case 2: return gemv_dispatch_r<AT, WT, 2>(st, g, p, pre);
    case 3: return gemv_dispatch_r<AT, WT, 3>(st, g, p, pre);
    case 4: return gemv_dispatch_r<AT, WT, 4>(st, g, p, pre);
    case 5: return gemv_dispatch_r<AT, WT, 5>(st, g, p, pre);
    case 6: return gemv_dispatch_r<AT, WT, 6>(st, g, p, pre);
    case 7: return gemv_dispatch_r<AT, WT, 7>(st, g, p, pre);
    case 8: return gemv_dispatch_r<AT, WT, 8>(st, g, p, pre);
  }
  return cudaErrorInvalidValue;
}

cudaError_t launch_gemv(cudaStream_t st, GemmArgs& g, const GemvPre* pre_in) {
  GemvPre pre{};
  if (pre_in) pre = *pre_in;
  if (g.M < 1 || g.M > 8 || (g.K % 8) != 0) return cudaErrorInvalidValue;
  if (pre.mode != 0 && pre.mode != 2) return cudaErrorInvalidValue;
  GemvPlan p = gemv_plan(g.M, g.N, g.K);
  g.S = p.S;
  if (g.out_capacity && (size_t)p.S * g.split_stride > g.out_capacity) return cudaErrorInvalidValue;
  if (g.adt == DT_F32 && g.wdt == DT_F32) return gemv_dispatch_m<float, float>(st, g, p, pre);
  if (g.adt == DT_F32 && g.wdt == DT_BF16) return gemv_dispatch_m<float, bf16>(st, g, p, pre);
  if (g.adt == DT_BF16 && g.wdt == DT_BF16) return gemv_dispatch_m<bf16, bf16>(st, g, p, pre);
  return cudaErrorInvalidValue;
}

// ------------------------------------------------------------------ tiled SIMT GEMM
constexpr int SBM = 64, SBN = 64, SBK = 16;

template <typename T>
__device__ __forceinline__ void load4(const T* p, float (&v)[4]);
template <>
__device__ __forceinline__ void load4<float>(const float* p, float (&v)[4]) {
  float4 a = *reinterpret_cast<const float4*>(p);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
}
template <>
__device__ __forceinline__ void load4<bf16>(const bf16* p, float (&v)[4]) {
  uint2 r = *reinterpret_cast<const uint2*>(p);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
  float2 a = __bfloat1622float2(h[0]), b = __bfloat1622float2(h[1]);
  v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}

template <typename AT, typename WT>
__global__ void __launch_bounds__(256)
gemm_simt_kernel(GemmArgs g, Trace tr) {
  TRACE_ENTRY(tr);
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float As[2][SBK][SBM + 4];
  __shared__ float Ws[2][SBK][SBN + 4];
  int row0, rows_valid;
  const WT* W = (const WT*)g.W;
  if (g.tile_expert) {
    if ((int)blockIdx.x >= *g.n_tiles) return;
    row0 = g.tile_row0[blockIdx.x];
    rows_valid = g.tile_rows[blockIdx.x];
    W += (size_t)g.tile_expert[blockIdx.x] * g.expert_stride;
  } else {
    row0 = blockIdx.x * SBM;
    rows_valid = min(SBM, g.M - row0);
  }
  const int col0 = blockIdx.y * SBN;
  const AT* A = (const AT*)g.A;
  const int tid = threadIdx.x;
  // loader mapping: 64 rows x 16 k = 256 threads x 4 consecutive k
  const int lr = tid >> 2, lk = (tid & 3) * 4;
  int arow = -1;
  if (lr < rows_valid) arow = g.a_rows ? g.a_rows[row0 + lr] : row0 + lr;
  const int wrow = (col0 + lr < g.N) ? col0 + lr : -1;
  const AT* ap = (arow >= 0) ? A + (size_t)arow * g.lda : nullptr;
  const WT* wp = (wrow >= 0) ? W + (size_t)wrow * g.K : nullptr;
  const int ty = tid >> 4, tx = tid & 15;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) acc[i][j] = 0.f;

  float ra[4], rw[4];
  auto fetch = [&](int k0) {
#pragma unroll
    for (int j = 0; j < 4; j++) ra[j] = rw[j] = 0.f;
    if (k0 + lk < g.K) {  // K % 4 == 0 is guaranteed (K % 8 == 0)
      if (ap) load4<AT>(ap + k0 + lk, ra);
      if (wp) load4<WT>(wp + k0 + lk, rw);
    }
  };
  auto stash = [&](int buf) {
#pragma unroll
    for (int j = 0; j < 4; j++) {
      As[buf][lk + j][lr] = ra[j];
      Ws[buf][lk + j][lr] = rw[j];
    }
  };
  const int nk = ceil_div(g.K, SBK);
  fetch(0);
  stash(0);
  __syncthreads();
  for (int kb = 0; kb < nk; kb++) {
    const int buf = kb & 1;
    if (kb + 1 < nk) fetch((kb + 1) * SBK);
#pragma unroll
    for (int k = 0; k < SBK; k++) {
      float4 a4 = *reinterpret_cast<const float4*>(&As[buf][k][ty * 4]);
      float4 b4 = *reinterpret_cast<const float4*>(&Ws[buf][k][tx * 4]);
      const float a[4] = {a4.x, a4.y, a4.z, a4.w};
      const float b[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (kb + 1 < nk) {
      stash(buf ^ 1);
      __syncthreads();
    }
  }
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const int r = ty * 4 + i;
    if (r >= rows_valid) continue;
    float* orow = g.out + (size_t)(row0 + r) * g.ldo;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int c = col0 + tx * 4 + j;
      if (c < g.N) orow[c] = g.alpha * acc[i][j];
    }
  }
  TRACE_EXIT(tr);
}

cudaError_t launch_gemm_simt(cudaStream_t st, GemmArgs& g) {
  if ((g.K % 8) != 0) return cudaErrorInvalidValue;
  g.S = 1;
  dim3 grid(g.tile_expert ? g.max_tiles : ceil_div(g.M, SBM), ceil_div(g.N, SBN));
  if (grid.x == 0 || grid.y == 0) return cudaSuccess;
  if (g.adt == DT_F32 && g.wdt == DT_F32) (void)launch_k(gemm_simt_kernel<float, float>, dim3(grid), dim3(256), 0, st, g, next_trace(12));
  else if (g.adt == DT_F32 && g.wdt == DT_BF16) (void)launch_k(gemm_simt_kernel<float, bf16>, dim3(grid), dim3(256), 0, st, g, next_trace(12));
  else if (g.adt == DT_BF16 && g.wdt == DT_BF16) (void)launch_k(gemm_simt_kernel<bf16, bf16>, dim3(grid), dim3(256), 0, st, g, next_trace(12));
  else return cudaErrorInvalidValue;
  return cudaGetLastError();
}

}  // namespace nvllm
