// sampling.cu — SampleWithHistory on the device (reference purego/tensor/sampling.go:33-217,
// driven by TensorModelRunner.Run, nanovllm/tensor_model_runner.go:92-94): repetition penalty
// over the sequence's token history with the x3 weight of the last ten tokens (:43-67),
// temperature (:70-74), softmax (:105-128), top-k (:131-155), top-p (:158-193),
// renormalisation (:90-98) and the inverse-CDF draw (:196-217) with rand.Float32() replaced by
// the step's uniform.  One CTA of 1024 threads per sampled row (rows with temperature <= 0
// stay greedy and are not touched).
//
// The reference sorts all V probabilities twice (sort.Slice); here the top-k / top-p cut is a
// four-pass radix select on the fp32 bit pattern of the probabilities (non-negative floats
// order like their bit patterns): per 8-bit digit a shared-memory histogram of counts and of
// probability mass, walked from the top digit down until the k-th element / the top_p mass is
// reached.  Ties at the threshold are broken by ascending token index (what a stable
// descending sort gives; the oracle pins its qsort the same way).  Mass is accumulated in
// fp64, the reference's running fp32 sum can differ in the last ulp — the parity tests keep u
// and top_p away from a CDF step by more than that.
#include "common.cuh"
#include "kernels.h"

namespace nvllm {

constexpr int SF_THREADS = 1024;
constexpr int SF_HIST_SMEM = 4096;  // history tokens kept in shared memory

struct SfRadix {
  unsigned cnt[256];
  double mass[256];
  unsigned sel_bucket, found;
  unsigned cnt_gt;
  double mass_gt;
};

// Walks the keys of p[0..V) in descending order until (by_mass ? mass >= target : count >= k).
// Outputs tau (threshold key), n_gt / mass_gt of the keys above it, n_eq keys equal to it.
// Returns false (all threads) when the walk never gets there (by_mass with total < target).
__device__ bool sf_radix_select(const float* __restrict__ p, int V, bool by_mass, unsigned k,
                                double target, SfRadix& R, unsigned& tau, unsigned& n_gt,
                                double& mass_gt, unsigned& n_eq) {
  unsigned prefix = 0, mask = 0;
  if (threadIdx.x == 0) {
    R.cnt_gt = 0;
    R.mass_gt = 0.0;
  }
  for (int shift = 24; shift >= 0; shift -= 8) {
    for (int b = threadIdx.x; b < 256; b += blockDim.x) {
      R.cnt[b] = 0;
      R.mass[b] = 0.0;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < V; i += blockDim.x) {
      const float v = p[i];
      const unsigned key = __float_as_uint(v);
      if ((key & mask) == prefix) {
        const unsigned b = (key >> shift) & 255u;
        atomicAdd(&R.cnt[b], 1u);
        if (by_mass) atomicAdd(&R.mass[b], (double)v);
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned c = R.cnt_gt;
      double m = R.mass_gt;
      R.found = 0;
      for (int b = 255; b >= 0; b--) {
        if (R.cnt[b] == 0) continue;
        const bool hit = by_mass ? (m + R.mass[b] >= target) : (c + R.cnt[b] >= k);
        if (hit) {
          R.sel_bucket = (unsigned)b;
          R.found = 1;
          break;
        }
        c += R.cnt[b];
        m += R.mass[b];
      }
      R.cnt_gt = c;
      R.mass_gt = m;
    }
    __syncthreads();
    if (!R.found) return false;
    prefix |= R.sel_bucket << shift;
    mask |= 255u << shift;
    if (shift == 0) n_eq = R.cnt[R.sel_bucket];
    __syncthreads();
  }
  tau = prefix;
  n_gt = R.cnt_gt;
  mass_gt = R.mass_gt;
  return true;
}

// Keeps p[i] with key > tau and the first `need_eq` (by index) with key == tau; zeroes the rest.
// Each thread owns a contiguous index chunk so that ranks follow index order.
__device__ void sf_apply_cut(float* __restrict__ p, int V, unsigned tau, unsigned need_eq,
                             unsigned* scan) {
  const int chunk = (V + blockDim.x - 1) / blockDim.x;
  const int lo = min(V, (int)threadIdx.x * chunk), hi = min(V, lo + chunk);
  unsigned eq = 0;
  for (int i = lo; i < hi; i++) eq += __float_as_uint(p[i]) == tau;
  scan[threadIdx.x] = eq;
  __syncthreads();
  for (int o = 1; o < (int)blockDim.x; o <<= 1) {
    const unsigned add = (threadIdx.x >= (unsigned)o) ? scan[threadIdx.x - o] : 0u;
    __syncthreads();
    scan[threadIdx.x] += add;
    __syncthreads();
  }
  unsigned rank = scan[threadIdx.x] - eq;  // equals before this chunk
  for (int i = lo; i < hi; i++) {
    const unsigned key = __float_as_uint(p[i]);
    if (key > tau) continue;
    if (key == tau && rank++ < need_eq) continue;
    p[i] = 0.f;
  }
  __syncthreads();
}

__global__ void __launch_bounds__(SF_THREADS)
sample_filtered_kernel(const float* __restrict__ logits, int ldl, int V, int n, int vsplit,
                       const float* __restrict__ temperature, const float* __restrict__ uniforms,
                       const int* __restrict__ hist, const int* __restrict__ cu_hist, float top_p,
                       int top_k, float rep, float* __restrict__ probs, int* __restrict__ out_tokens) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float red[32];
  __shared__ SfRadix R;
  __shared__ unsigned scan_u[SF_THREADS];
  __shared__ int s_hist[SF_HIST_SMEM];
  __shared__ int winner;
  const int s = blockIdx.x;
  const float temp = temperature[s];
  if (!(temp > 0.f)) return;  // greedy rows were decided by sample_kernel
  float* p = probs + (size_t)s * V;
  // tensor parallel: the gathered logits are rank-major vocabulary slices [tp][n][vsplit]
  auto at = [&](int i) -> float {
    if (vsplit > 0) return logits[((size_t)(i / vsplit) * n + s) * vsplit + (i % vsplit)];
    return logits[(size_t)s * ldl + i];
  };
  for (int i = threadIdx.x; i < V; i += blockDim.x) p[i] = at(i);
  // ---- repetition penalty (:43-67): count = sum of weights of the token's occurrences, the
  // last ten history positions weigh 3; the FIRST occurrence applies the penalty once
  const int h0 = cu_hist[s], hn = cu_hist[s + 1] - h0;
  if (rep != 1.0f && hn > 0) {
    const bool in_smem = hn <= SF_HIST_SMEM;
    if (in_smem)
      for (int i = threadIdx.x; i < hn; i += blockDim.x) s_hist[i] = hist[h0 + i];
    __syncthreads();  // also orders the copy of p above before the read-modify-write below
    const int* hs = in_smem ? s_hist : hist + h0;
    for (int i = threadIdx.x; i < hn; i += blockDim.x) {
      const int tok = hs[i];
      int cnt = 0;
      bool first = true;
      for (int j = 0; j < hn; j++) {
        if (hs[j] == tok) {
          cnt += (j >= hn - 10) ? 3 : 1;
          if (j < i) first = false;
        }
      }
      if (first && tok >= 0 && tok < V) {
        const float pen = rep * (float)cnt;
        const float v = p[tok];
        p[tok] = v > 0.f ? v / pen : v * pen;
      }
    }
  }
  __syncthreads();
  // ---- temperature (:70-74) + softmax (:105-128; exp through float64 like the reference)
  const bool scale = temp != 1.0f;
  float mx = -INFINITY;
  for (int i = threadIdx.x; i < V; i += blockDim.x) {
    float v = p[i];
    if (scale) {
      v = v / temp;
      p[i] = v;
    }
    mx = fmaxf(mx, v);
  }
  mx = block_max(mx, red);
  float lsum = 0.f;
  for (int i = threadIdx.x; i < V; i += blockDim.x) {
    const float e = (float)exp((double)(p[i] - mx));
    p[i] = e;
    lsum += e;
  }
  const float sum = block_sum(lsum, red);
  for (int i = threadIdx.x; i < V; i += blockDim.x) p[i] = p[i] / sum;
  __syncthreads();
  // ---- top-k (:131-155)
  if (top_k > 0 && top_k < V) {
    unsigned tau, n_gt, n_eq;
    double m_gt;
    if (sf_radix_select(p, V, false, (unsigned)top_k, 0.0, R, tau, n_gt, m_gt, n_eq))
      sf_apply_cut(p, V, tau, (unsigned)top_k - n_gt, scan_u);
  }
  // ---- top-p (:158-193): the shortest descending prefix whose mass reaches top_p
  if (top_p < 1.0f) {
    unsigned tau, n_gt, n_eq;
    double m_gt;
    if (sf_radix_select(p, V, true, 0u, (double)top_p, R, tau, n_gt, m_gt, n_eq)) {
      const double tv = (double)__uint_as_float(tau);
      unsigned need = 1;
      if (tv > 0.0) {
        const double m = ceil(((double)top_p - m_gt) / tv);
        need = m < 1.0 ? 1u : (m > (double)n_eq ? n_eq : (unsigned)m);
      } else {
        need = n_eq;
      }
      sf_apply_cut(p, V, tau, need, scan_u);
    }
  }
  // ---- renormalise (:90-98) + inverse CDF (:196-217).  Dividing every p by the total and
  // drawing r = u * sum(p / total) is the same walk as r = u * total over the unnormalised p.
  float* scan = reinterpret_cast<float*>(scan_u);
  const int chunk = (V + blockDim.x - 1) / blockDim.x;
  const int lo = min(V, (int)threadIdx.x * chunk), hi = min(V, lo + chunk);
  float csum = 0.f;
  for (int i = lo; i < hi; i++) csum += p[i];
  scan[threadIdx.x] = csum;
  __syncthreads();
  for (int o = 1; o < (int)blockDim.x; o <<= 1) {
    const float add = (threadIdx.x >= (unsigned)o) ? scan[threadIdx.x - o] : 0.f;
    __syncthreads();
    scan[threadIdx.x] += add;
    __syncthreads();
  }
  const float total = scan[blockDim.x - 1];
  const float r = uniforms[s] * total;
  if (threadIdx.x == 0) winner = V - 1;  // sampling.go:212-214 clamp
  __syncthreads();
  const float incl = scan[threadIdx.x];
  const bool mine = (incl >= r) && (threadIdx.x == 0 || scan[threadIdx.x - 1] < r) && (hi > lo);
  if (mine) {
    float run = incl - csum;
    int pick = hi - 1;
    for (int i = lo; i < hi; i++) {
      run += p[i];
      if (run >= r) {
        pick = i;
        break;
      }
    }
    atomicMin(&winner, pick);
  }
  __syncthreads();
  if (threadIdx.x == 0) out_tokens[s] = winner;
}

cudaError_t launch_sample_filtered(cudaStream_t st, const float* logits, int ldl, int V, int n,
                                   int vsplit, const float* temperature, const float* uniforms,
                                   const int* hist, const int* cu_hist, float top_p, int top_k,
                                   float rep, float* probs, int* out_tokens) {
  if (n <= 0) return cudaSuccess;
  if (!temperature || !uniforms || !hist || !cu_hist || !probs) return cudaErrorInvalidValue;
  (void)launch_k(sample_filtered_kernel, dim3(n), dim3(SF_THREADS), 0, st, logits, ldl, V, n, vsplit,
                 temperature, uniforms, hist, cu_hist, top_p, top_k, rep, probs, out_tokens);
  return cudaGetLastError();
}

}  // namespace nvllm
