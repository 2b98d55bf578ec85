// moe_route.cuh — one token's expert routing (moe.go:63-103), shared by the stand-alone router
// kernel (moe.cu) and the norm kernel that fuses it (elementwise.cu).
#pragma once
#include "common.cuh"

namespace nvllm {

constexpr int MOE_MAX_E = 256;

// All warps of the CTA: warp w computes the logits of experts w, w + nw, ... from the
// normed row y (shared or global memory, fp32) and R [E, H]; warp 0 then does the softmax
// over ALL E experts (tensor.go:128-155), k rounds of first-max selection (ties -> lower
// index, which is what a stable descending sort gives; the reference's sort.Slice is unstable,
// ties have measure zero with random-init weights) and the renormalisation over the chosen k
// (moe.go:88-103).  probs: shared scratch of >= E floats.  Contains __syncthreads.
__device__ __forceinline__ void moe_route_row(const float* y, const float* __restrict__ R, int H,
                                              int E, int k, float* probs, size_t t,
                                              int* __restrict__ topk_idx,
                                              float* __restrict__ topk_w) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  // Two experts per pass with the k loop unrolled four deep: up to 8 independent 32-byte loads
  // of R per lane in flight, so the E / nw passes cost one L2 round trip each instead of one
  // per 256 columns per expert (measured 10 us per row CTA with the rolled loop at E = 32).
  // The per-expert accumulation order (ascending i, then the warp tree) is unchanged.
  for (int e = warp; e < E; e += 2 * nw) {
    const int e2 = e + nw;
    const bool two = e2 < E;
    const float* r1 = R + (size_t)e * H;
    const float* r2 = R + (size_t)(two ? e2 : e) * H;
    float acc1 = 0.f, acc2 = 0.f;
    for (int i0 = lane * 8; i0 < H; i0 += 1024) {
      float b1[4][8], b2[4][8];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const int i = i0 + u * 256;
        if (i < H) {
          load8(r1 + i, b1[u]);
          if (two) load8(r2 + i, b2[u]);
        }
      }
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const int i = i0 + u * 256;
        if (i < H) {
          float a[8];
          load8(y + i, a);
#pragma unroll
          for (int j = 0; j < 8; j++) acc1 = fmaf(a[j], b1[u][j], acc1);
          if (two) {
#pragma unroll
            for (int j = 0; j < 8; j++) acc2 = fmaf(a[j], b2[u][j], acc2);
          }
        }
      }
    }
    acc1 = warp_sum(acc1);
    acc2 = warp_sum(acc2);
    if (lane == 0) {
      probs[e] = acc1;
      if (two) probs[e2] = acc2;
    }
  }
  __syncthreads();
  if (warp != 0) return;
  float mx = -INFINITY;
  for (int e = lane; e < E; e += 32) mx = fmaxf(mx, probs[e]);
  mx = warp_max(mx);
  float sum = 0.f;
  for (int e = lane; e < E; e += 32) {
    float p = expf(probs[e] - mx);
    probs[e] = p;
    sum += p;
  }
  sum = warp_sum(sum);
  __syncwarp();
  for (int e = lane; e < E; e += 32) probs[e] = probs[e] / sum;
  __syncwarp();
  float sel_p = 0.f;  // lane s keeps the s-th selection (k <= 32)
  int sel_i = 0;
  float ksum = 0.f;
  for (int s = 0; s < k; s++) {
    float bv = -1.f;
    int bi = 0x7fffffff;
    for (int e = lane; e < E; e += 32) {
      float p = probs[e];
      if (p > bv) {
        bv = p;
        bi = e;
      }
    }
    // warp arg-max in two redux operations instead of five shuffle rounds: probabilities are
    // >= 0, so their bit patterns order like the values (+1 keeps a genuine 0.0 above the -1 of
    // a lane with no candidate); ties go to the lower expert index
    const unsigned key = bv < 0.f ? 0u : __float_as_uint(bv) + 1u;
    const unsigned kmax = __reduce_max_sync(0xffffffffu, key);
    bi = (int)__reduce_min_sync(0xffffffffu, key == kmax ? (unsigned)bi : 0x7fffffffu);
    bv = kmax ? __uint_as_float(kmax - 1u) : -1.f;
    if (lane == s) {
      sel_p = bv;
      sel_i = bi;
    }
    ksum += bv;  // moe.go:98-103: sum in descending order
    if (lane == 0) probs[bi] = -2.f;
    __syncwarp();
  }
  if (lane < k) {
    topk_idx[t * k + lane] = sel_i;
    topk_w[t * k + lane] = sel_p / ksum;
  }
}

}  // namespace nvllm
