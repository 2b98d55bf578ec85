// moe.cu — Granite top-k mixture-of-experts plumbing around the grouped expert GEMMs:
// router (matmul + softmax over ALL experts + top-k + renormalise, moe.go:63-103), grouping
// of the (token, slot) pairs by expert into 64-row tiles, the GLU between the two expert
// GEMMs (moe.go:203-212) and the weighted combine (moe.go:116-118).
#include <algorithm>

#include "common.cuh"
#include "kernels.h"
#include "moe_route.cuh"

namespace nvllm {

// One CTA (4 warps) per token: the stand-alone router (fp32 parity mode and callers that did
// not fuse it into the norm).
__global__ void __launch_bounds__(128)
moe_router_kernel(const float* __restrict__ x, const float* __restrict__ R, int H, int E, int k,
                  int* __restrict__ topk_idx, float* __restrict__ topk_w) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float probs[MOE_MAX_E];
  const int t = blockIdx.x;
  moe_route_row(x + (size_t)t * H, R, H, E, k, probs, (size_t)t, topk_idx, topk_w);
}

cudaError_t launch_moe_router(cudaStream_t st, const float* x, const float* R, int T, int H, int E,
                              int k, int* topk_idx, float* topk_w) {
  if (T <= 0) return cudaSuccess;
  if (E > MOE_MAX_E || k > 32 || k > E || (H % 8) != 0) return cudaErrorInvalidValue;
  (void)launch_k(moe_router_kernel, dim3(T), dim3(128), 0, st, x, R, H, E, k, topk_idx, topk_w);
  return cudaGetLastError();
}

// Single-CTA counting sort of the T*k pairs by expert + tile list.  Rows inside an expert are
// in arbitrary order; every consumer addresses them through pair_row / a_rows, so results do
// not depend on it.
__global__ void __launch_bounds__(1024)
moe_sort_kernel(const int* __restrict__ topk_idx, int n_pairs, int k, int E, MoeSortBufs b,
                int max_tiles, int tile, Trace tr) {
  TRACE_ENTRY(tr);
  pdl_launch_dependents();
  pdl_wait();
  __shared__ int cnt[MOE_MAX_E];
  __shared__ int off[MOE_MAX_E + 1];
  for (int e = threadIdx.x; e < E; e += blockDim.x) cnt[e] = 0;
  __syncthreads();
  for (int p = threadIdx.x; p < n_pairs; p += blockDim.x) atomicAdd(&cnt[topk_idx[p]], 1);
  __syncthreads();
  if (threadIdx.x == 0) {
    int run = 0, nt = 0;
    for (int e = 0; e < E; e++) {
      off[e] = run;
      b.counts[e] = cnt[e];
      b.offsets[e] = run;
      for (int r = 0; r < cnt[e] && nt < max_tiles; r += tile) {
        b.tile_expert[nt] = e;
        b.tile_row0[nt] = run + r;
        b.tile_rows[nt] = min(tile, cnt[e] - r);
        nt++;
      }
      run += cnt[e];
    }
    off[E] = run;
    b.offsets[E] = run;
    *b.n_tiles = nt;
  }
  __syncthreads();
  for (int e = threadIdx.x; e < E; e += blockDim.x) cnt[e] = 0;  // reuse as cursors
  __syncthreads();
  for (int p = threadIdx.x; p < n_pairs; p += blockDim.x) {
    const int e = topk_idx[p];
    const int row = off[e] + atomicAdd(&cnt[e], 1);
    b.sorted_pair[row] = p;
    b.pair_row[p] = row;
    b.a_rows[row] = p / k;
  }
  TRACE_EXIT(tr);
}

cudaError_t launch_moe_sort(cudaStream_t st, const int* topk_idx, int T, int k, int E,
                            const MoeSortBufs& b, int max_tiles, int tile) {
  if (T <= 0) return cudaSuccess;
  if (E > MOE_MAX_E || tile <= 0) return cudaErrorInvalidValue;
  (void)launch_k(moe_sort_kernel, dim3(1), dim3(1024), 0, st, topk_idx, T * k, k, E, b, max_tiles,
                 tile, next_trace(15));
  return cudaGetLastError();
}

// xs[row, :] = x[a_rows[row], :] — the expert-sorted copy of the normed activations that the
// grouped tensor-core GEMM reads with plain 2-D TMA boxes (bf16, H % 8 == 0)
__global__ void moe_gather_kernel(size_t nvec, int vpr, const uint4* __restrict__ x,
                                  const int* __restrict__ a_rows, uint4* __restrict__ xs,
                                  Trace tr) {
  TRACE_ENTRY(tr);
  pdl_launch_dependents();
  pdl_wait();
  for (size_t v = (size_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec;
       v += (size_t)gridDim.x * blockDim.x) {
    const size_t row = v / vpr;
    const int j = (int)(v - row * vpr);
    xs[v] = x[(size_t)a_rows[row] * vpr + j];
  }
  TRACE_EXIT(tr);
}

cudaError_t launch_moe_gather(cudaStream_t st, int rows, int H, const void* x, const int* a_rows,
                              void* xs) {
  if (rows <= 0) return cudaSuccess;
  if (H % 8) return cudaErrorInvalidValue;
  const int vpr = H / 8;
  const size_t nvec = (size_t)rows * vpr;
  const int blocks = (int)min((size_t)148 * 8, (nvec + 255) / 256);
  (void)launch_k(moe_gather_kernel, dim3(blocks), dim3(256), 0, st, nvec, vpr, (const uint4*)x,
                 a_rows, (uint4*)xs, next_trace(16));
  return cudaGetLastError();
}

// Decode-sized batches (<= MOE_SG_MAX_P pairs, <= MOE_SG_MAX_E experts): sort and gather in ONE
// launch.  Every CTA recomputes the same stable counting sort of the pairs in shared memory
// (row of pair p = offset of its expert + number of earlier pairs routed to that expert —
// deterministic, so all CTAs agree), CTA 0 publishes the lists the grouped GEMMs and the
// combine read, and each CTA copies its share of the sorted activation rows.
constexpr int MOE_SG_MAX_P = 4096, MOE_SG_MAX_E = 64;
__global__ void __launch_bounds__(1024)
moe_sort_gather_kernel(const int* __restrict__ topk_idx, int n_pairs, int k, int E, MoeSortBufs b,
                       int max_tiles, int tile, int vpr, const uint4* __restrict__ x,
                       uint4* __restrict__ xs, Trace tr) {
  TRACE_ENTRY(tr);
  pdl_launch_dependents();
  __shared__ int s_cnt[MOE_SG_MAX_E];
  __shared__ int s_off[MOE_SG_MAX_E + 1];
  __shared__ int s_base[MOE_SG_MAX_E];
  __shared__ int s_wcnt[32][MOE_SG_MAX_E];
  __shared__ int s_inv[MOE_SG_MAX_P];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid < MOE_SG_MAX_E) s_cnt[tid] = s_base[tid] = 0;
  pdl_wait();
  __syncthreads();
  for (int p = tid; p < n_pairs; p += 1024) atomicAdd(&s_cnt[topk_idx[p]], 1);
  __syncthreads();
  if (warp == 0) {
    // expert offsets and the tile list by warp scans (E <= 64: lane holds experts lane and
    // lane + 32) — the serial loop over the experts cost ~2 us on the single publishing thread
    const bool pub = blockIdx.x == 0;
    const int e0 = lane, e1 = lane + 32;
    const int c0 = e0 < E ? s_cnt[e0] : 0, c1 = e1 < E ? s_cnt[e1] : 0;
    const int t0 = (c0 + tile - 1) / tile, t1 = (c1 + tile - 1) / tile;
    int ic0 = c0, ic1 = c1, it0 = t0, it1 = t1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int a0 = __shfl_up_sync(0xffffffffu, ic0, o), a1 = __shfl_up_sync(0xffffffffu, ic1, o);
      const int b0 = __shfl_up_sync(0xffffffffu, it0, o), b1 = __shfl_up_sync(0xffffffffu, it1, o);
      if (lane >= o) {
        ic0 += a0;
        ic1 += a1;
        it0 += b0;
        it1 += b1;
      }
    }
    const int tot_c0 = __shfl_sync(0xffffffffu, ic0, 31), tot_t0 = __shfl_sync(0xffffffffu, it0, 31);
    const int tot_c = tot_c0 + __shfl_sync(0xffffffffu, ic1, 31);
    const int tot_t = tot_t0 + __shfl_sync(0xffffffffu, it1, 31);
    const int off0 = ic0 - c0, off1 = tot_c0 + ic1 - c1;
    const int tb0 = it0 - t0, tb1 = tot_t0 + it1 - t1;
    if (e0 < E) s_off[e0] = off0;
    if (e1 < E) s_off[e1] = off1;
    if (lane == 0) s_off[E] = tot_c;
    if (pub) {
      if (e0 < E) {
        b.counts[e0] = c0;
        b.offsets[e0] = off0;
        for (int r = 0; r < t0 && tb0 + r < max_tiles; r++) {
          b.tile_expert[tb0 + r] = e0;
          b.tile_row0[tb0 + r] = off0 + r * tile;
          b.tile_rows[tb0 + r] = min(tile, c0 - r * tile);
        }
      }
      if (e1 < E) {
        b.counts[e1] = c1;
        b.offsets[e1] = off1;
        for (int r = 0; r < t1 && tb1 + r < max_tiles; r++) {
          b.tile_expert[tb1 + r] = e1;
          b.tile_row0[tb1 + r] = off1 + r * tile;
          b.tile_rows[tb1 + r] = min(tile, c1 - r * tile);
        }
      }
      if (lane == 0) {
        b.offsets[E] = tot_c;
        *b.n_tiles = min(tot_t, max_tiles);
      }
    }
  }
  for (int c0 = 0; c0 < n_pairs; c0 += 1024) {
    for (int i = tid; i < 32 * MOE_SG_MAX_E; i += 1024) (&s_wcnt[0][0])[i] = 0;
    __syncthreads();
    const int p = c0 + tid;
    const int e = p < n_pairs ? topk_idx[p] : -1;
    const unsigned same = __match_any_sync(0xffffffffu, e);
    const int rank = __popc(same & ((1u << lane) - 1));
    if (e >= 0 && rank == 0) s_wcnt[warp][e] = __popc(same);
    __syncthreads();
    if (tid < E) {  // exclusive prefix over the warps, continuing from the earlier chunks
      int run = s_base[tid];
      for (int w = 0; w < 32; w++) {
        const int n = s_wcnt[w][tid];
        s_wcnt[w][tid] = run;
        run += n;
      }
      s_base[tid] = run;
    }
    __syncthreads();
    if (e >= 0) {
      const int row = s_off[e] + s_wcnt[warp][e] + rank;
      s_inv[row] = p;
      if (blockIdx.x == 0) {
        b.sorted_pair[row] = p;
        b.pair_row[p] = row;
        b.a_rows[row] = p / k;
      }
    }
    __syncthreads();
  }
  const int rpc = (n_pairs + gridDim.x - 1) / gridDim.x;
  const int r0 = blockIdx.x * rpc;
  for (int idx = tid; idx < rpc * vpr; idx += 1024) {
    const int r = r0 + idx / vpr, j = idx % vpr;
    if (r < n_pairs) xs[(size_t)r * vpr + j] = x[(size_t)(s_inv[r] / k) * vpr + j];
  }
  TRACE_EXIT(tr);
}

bool moe_sort_gather_fits(int T, int k, int E, int H) {
  return (long)T * k <= MOE_SG_MAX_P && E <= MOE_SG_MAX_E && (H % 8) == 0;
}
cudaError_t launch_moe_sort_gather(cudaStream_t st, const int* topk_idx, int T, int k, int E,
                                   const MoeSortBufs& b, int max_tiles, int tile, int H,
                                   const void* x, void* xs) {
  if (T <= 0) return cudaSuccess;
  if (!moe_sort_gather_fits(T, k, E, H) || tile <= 0) return cudaErrorInvalidValue;
  const int vpr = H / 8;
  const long work = (long)T * k * vpr;
  const int grid = (int)std::max(1L, std::min(148L, work / 2048));
  (void)launch_k(moe_sort_gather_kernel, dim3(grid), dim3(1024), 0, st, topk_idx, T * k, k, E, b,
                 max_tiles, tile, vpr, (const uint4*)x, (uint4*)xs, next_trace(17));
  return cudaGetLastError();
}

// moe.go:203-212: sigmoid with an fp32 divide; gate j = proj[j], up j = proj[I + j].  The
// expert input weights are repacked at load with gate / up rows interleaved (weights.cu), so
// h[row, 2j] is gate j and h[row, 2j + 1] is up j.
template <typename OT>
__global__ void moe_glu_kernel(size_t nvec, int I, const float* __restrict__ h,
                               OT* __restrict__ g) {
  pdl_launch_dependents();
  pdl_wait();
  const int vpr = I / 8;
  for (size_t v = (size_t)blockIdx.x * blockDim.x + threadIdx.x; v < nvec;
       v += (size_t)gridDim.x * blockDim.x) {
    const size_t row = v / vpr;
    const int j = (int)(v % vpr) * 8;
    float lo[8], hi[8], o[8];
    load8(h + row * 2 * I + 2 * j, lo);
    load8(h + row * 2 * I + 2 * j + 8, hi);
#pragma unroll
    for (int i = 0; i < 8; i++) {
      const float a = i < 4 ? lo[2 * i] : hi[2 * i - 8];
      const float u = i < 4 ? lo[2 * i + 1] : hi[2 * i - 7];
      const float sig = 1.0f / (1.0f + expf(-a));
      o[i] = a * sig * u;
    }
    store8(g + row * I + j, o);
  }
}

cudaError_t launch_moe_glu(cudaStream_t st, int odt, int rows, int I, const float* h, void* g) {
  if (rows <= 0) return cudaSuccess;
  if ((I % 8) != 0) return cudaErrorInvalidValue;
  size_t nvec = (size_t)rows * (I / 8);
  int blocks = (int)min((size_t)148 * 16, (nvec + 255) / 256);
  if (odt == DT_F32)
    (void)launch_k(moe_glu_kernel<float>, dim3(blocks), dim3(256), 0, st, nvec, I, h, (float*)g);
  else
    (void)launch_k(moe_glu_kernel<bf16>, dim3(blocks), dim3(256), 0, st, nvec, I, h, (bf16*)g);
  return cudaGetLastError();
}

// y[t,:] = sum_s w[t,s] * Y[pair_row[t*k+s],:]  in slot order (descending probability)
__global__ void moe_combine_kernel(int k, int H, const float* __restrict__ Y,
                                   const int* __restrict__ pair_row,
                                   const float* __restrict__ topk_w, float* __restrict__ y) {
  pdl_launch_dependents();
  pdl_wait();
  const int t = blockIdx.x;
  for (int i = threadIdx.x * 8; i < H; i += blockDim.x * 8) {
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; j++) acc[j] = 0.f;
    for (int s = 0; s < k; s++) {
      const int row = pair_row[(size_t)t * k + s];
      const float w = topk_w[(size_t)t * k + s];
      float v[8];
      load8(Y + (size_t)row * H + i, v);
#pragma unroll
      for (int j = 0; j < 8; j++) acc[j] += w * v[j];
    }
    store8(y + (size_t)t * H + i, acc);
  }
}

cudaError_t launch_moe_combine(cudaStream_t st, int T, int k, int H, const float* Y,
                               const int* pair_row, const float* topk_w, float* y) {
  if (T <= 0) return cudaSuccess;
  int threads = min(256, max(32, ((H / 8 + 31) / 32) * 32));
  (void)launch_k(moe_combine_kernel, dim3(T), dim3(threads), 0, st, k, H, Y, pair_row, topk_w, y);
  return cudaGetLastError();
}

}  // namespace nvllm
