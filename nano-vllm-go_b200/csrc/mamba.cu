// mamba.cu — Mamba2 block of the Granite-4 hybrid (SURVEY.md §8f rank 4;
// purego/tensor/mamba2.go:72-351), for a ragged batch of sequences.  The two projections are
// ordinary GEMMs (forward.cu); this file holds what sits between them:
//   mamba_conv_kernel   sum of the in_proj split-K slices, causal conv1d over the xBC columns
//                       (zero-padded at the start of every CALL, like causalConv1d :209-270 — the
//                       reference never reads its ConvCache), SiLU, and dt = softplus(dt + bias)
//   mamba_scan_kernel   the selective scan (:273-351), sequential over a sequence's new tokens,
//                       one CTA per (sequence, head), one thread per head_dim row holding its
//                       state_size states in registers; the SSM state lives per KV BLOCK
//                       (handle.h) and is read / written through the block table
//   mamba_gate_norm_kernel  y * silu(gate), RMS norm over E with eps 1e-5, * Norm (:139-173)
// Correctness-first kernels: HBM traffic is small next to the projections, and the scan's
// token loop is latency-bound by construction (a chunked scan is the obvious next step).
#include "common.cuh"
#include "kernels.h"

namespace nvllm {

namespace {

__device__ __forceinline__ int seq_of_row(const int* __restrict__ cu_q, int n_seqs, int row) {
  int lo = 0, hi = n_seqs - 1;  // last s with cu_q[s] <= row
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (cu_q[mid] <= row) lo = mid;
    else hi = mid - 1;
  }
  return lo;
}

__device__ __forceinline__ float sum_slices(const float* __restrict__ p, int S, size_t stride) {
  float acc = 0.f;
  for (int s = 0; s < S; s++) acc += p[(size_t)s * stride];
  return acc;
}

}  // namespace

// grid (T, ceil(conv / 256)); proj: [S][T][P] split-K slices of x @ in_proj^T, columns
// [gate (E) | xBC (conv) | dt (heads)]
__global__ void __launch_bounds__(256)
mamba_conv_kernel(MambaArgs a, Trace tr) {
  TRACE_ENTRY(tr);
  pdl_launch_dependents();
  pdl_wait();
  const int t = blockIdx.x;
  const int s = seq_of_row(a.cu_q, a.n_seqs, t);
  const int li = t - a.cu_q[s];  // index of the token inside this call's rows of its sequence
  const int c = blockIdx.y * blockDim.x + threadIdx.x;
  if (c < a.conv) {
    float sum = 0.f;
    for (int k = 0; k < a.K; k++) {
      const int back = a.K - 1 - k;  // tap k reads the token `back` rows earlier
      if (li - back >= 0)
        sum += sum_slices(a.proj + (size_t)(t - back) * a.P + a.E + c, a.S, a.stride) * a.conv_w[c * a.K + k];
    }
    if (a.conv_b) sum += a.conv_b[c];
    a.xbc[(size_t)t * a.conv + c] = silu(sum);
  }
  if (blockIdx.y == 0 && threadIdx.x < a.nh) {
    const int h = threadIdx.x;
    float dt = sum_slices(a.proj + (size_t)t * a.P + a.E + a.conv + h, a.S, a.stride);
    if (a.dt_bias) dt += a.dt_bias[h];
    a.dt[(size_t)t * a.nh + h] = (float)log(1.0 + exp((double)dt));  // Softplus, mamba2.go:369-376
  }
  TRACE_EXIT(tr);
}

// grid (n_seqs, heads), block = head_dim threads (thread d owns state[d][0..N))
template <int NMAX>
__global__ void __launch_bounds__(128)
mamba_scan_kernel(MambaArgs a, Trace tr) {
  TRACE_ENTRY(tr);
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float sB[NMAX], sC[NMAX];
  const int s = blockIdx.x, h = blockIdx.y, d = threadIdx.x;
  const int r0 = a.cu_q[s], r1 = a.cu_q[s + 1];
  const int start = a.positions[r0];  // position of the first new token
  const int* bt = a.block_tables + (size_t)s * a.max_blocks;
  const size_t per_layer = (size_t)a.nh * a.hd * a.N;
  const size_t per_block = (size_t)a.n_mamba * per_layer;
  float st[NMAX];
#pragma unroll
  for (int i = 0; i < NMAX; i++) st[i] = 0.f;
  if (start > 0 && d < a.hd) {
    // the state after position start - 1 sits in that position's block
    const float* src = a.ssm_pool + (size_t)bt[(start - 1) / KV_BLOCK] * per_block + (size_t)a.layer * per_layer +
                       ((size_t)h * a.hd + d) * a.N;
#pragma unroll
    for (int i = 0; i < NMAX; i++)
      if (i < a.N) st[i] = src[i];
  }
  int grp = h * a.G / a.nh;
  if (grp >= a.G) grp = a.G - 1;
  const float A = a.a_log ? -(float)exp((double)a.a_log[h]) : 0.f;
  const float Dh = a.d_skip ? a.d_skip[h] : 0.f;
  for (int t = r0; t < r1; t++) {
    __syncthreads();  // the previous token's sB / sC have been consumed
    for (int i = threadIdx.x; i < a.N; i += blockDim.x) {
      sB[i] = a.xbc[(size_t)t * a.conv + a.E + (size_t)grp * a.N + i];
      sC[i] = a.xbc[(size_t)t * a.conv + a.E + (size_t)a.G * a.N + (size_t)grp * a.N + i];
    }
    __syncthreads();
    if (d < a.hd) {
      const float dt = a.dt[(size_t)t * a.nh + h];
      const float abar = a.a_log ? (float)exp((double)(A * dt)) : 1.0f;
      const float u = a.xbc[(size_t)t * a.conv + (size_t)h * a.hd + d];
      float sum = 0.f;
#pragma unroll
      for (int i = 0; i < NMAX; i++)
        if (i < a.N) {
          st[i] = abar * st[i] + dt * sB[i] * u;  // mamba2.go:325-331
          sum += sC[i] * st[i];                   // :335-340, s ascending
        }
      a.y[(size_t)t * a.E + (size_t)h * a.hd + d] = sum + Dh * u;
      const int pos = start + (t - r0);
      if ((pos + 1) % KV_BLOCK == 0 || t == r1 - 1) {
        // snapshot: the state after the last token written into this block
        float* dst = a.ssm_pool + (size_t)bt[pos / KV_BLOCK] * per_block + (size_t)a.layer * per_layer +
                     ((size_t)h * a.hd + d) * a.N;
#pragma unroll
        for (int i = 0; i < NMAX; i++)
          if (i < a.N) dst[i] = st[i];
      }
    }
  }
  TRACE_EXIT(tr);
}

// one CTA per token row: y[t, :] * silu(gate[t, :]) -> RMS norm (eps 1e-5, mamba2.go:150-166)
// -> * Norm -> act dtype (the out_proj GEMM's input)
template <typename OT>
__global__ void __launch_bounds__(256)
mamba_gate_norm_kernel(MambaArgs a, OT* __restrict__ out, Trace tr) {
  TRACE_ENTRY(tr);
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ float row[];
  __shared__ float red[32];
  const int t = blockIdx.x;
  float lsq = 0.f;
  for (int i = threadIdx.x; i < a.E; i += blockDim.x) {
    const float g = sum_slices(a.proj + (size_t)t * a.P + i, a.S, a.stride);
    const float v = a.y[(size_t)t * a.E + i] * silu(g);
    row[i] = v;
    lsq += v * v;
  }
  const float tot = block_sum(lsq, red);
  const float rms = 1.0f / sqrtf(tot / (float)a.E + 1e-5f);
  for (int i = threadIdx.x; i < a.E; i += blockDim.x)
    out[(size_t)t * a.E + i] = from_f32<OT>(row[i] * rms * (a.norm_w ? a.norm_w[i] : 1.f));
  TRACE_EXIT(tr);
}

cudaError_t launch_mamba_conv(cudaStream_t st, const MambaArgs& a) {
  if (a.T <= 0) return cudaSuccess;
  (void)launch_k(mamba_conv_kernel, dim3(a.T, ceil_div(a.conv, 256)), dim3(256), 0, st, a, next_trace(20));
  return cudaGetLastError();
}

cudaError_t launch_mamba_scan(cudaStream_t st, const MambaArgs& a) {
  if (a.n_seqs <= 0) return cudaSuccess;
  const int threads = max(32, ((a.hd + 31) / 32) * 32);
  if (threads > 128 || a.N > 128) return cudaErrorInvalidValue;
  const dim3 grid(a.n_seqs, a.nh);
  if (a.N <= 16) (void)launch_k(mamba_scan_kernel<16>, grid, dim3(threads), 0, st, a, next_trace(21));
  else if (a.N <= 32) (void)launch_k(mamba_scan_kernel<32>, grid, dim3(threads), 0, st, a, next_trace(21));
  else if (a.N <= 64) (void)launch_k(mamba_scan_kernel<64>, grid, dim3(threads), 0, st, a, next_trace(21));
  else (void)launch_k(mamba_scan_kernel<128>, grid, dim3(threads), 0, st, a, next_trace(21));
  return cudaGetLastError();
}

cudaError_t launch_mamba_gate_norm(cudaStream_t st, const MambaArgs& a, int odt, void* out) {
  if (a.T <= 0) return cudaSuccess;
  const size_t smem = (size_t)a.E * sizeof(float);
  if (odt == DT_F32)
    (void)launch_k(mamba_gate_norm_kernel<float>, dim3(a.T), dim3(256), smem, st, a, (float*)out, next_trace(22));
  else
    (void)launch_k(mamba_gate_norm_kernel<bf16>, dim3(a.T), dim3(256), smem, st, a, (bf16*)out, next_trace(22));
  return cudaGetLastError();
}

}  // namespace nvllm
