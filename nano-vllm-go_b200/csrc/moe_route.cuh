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
  for (int e = warp; e < E; e += nw) {
    const float* rr = R + (size_t)e * H;
    float acc = 0.f;
    for (int i = lane * 8; i < H; i += 256) {
      float a[8], b[8];
      load8(y + i, a);
      load8(rr + i, b);
#pragma unroll
      for (int j = 0; j < 8; j++) acc = fmaf(a[j], b[j], acc);
    }
    acc = warp_sum(acc);
    if (lane == 0) probs[e] = acc;
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
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > bv || (ov == bv && oi < bi)) {
        bv = ov;
        bi = oi;
      }
    }
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
