// common.cuh — shared device helpers for libnvllm_b200 (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <utility>

#define NVLLM_WARP 32

namespace nvllm {

typedef __nv_bfloat16 bf16;

// ---- error plumbing: kernels never abort the process; launchers return cudaError_t
#define NV_CUDA(expr)                                        \
  do {                                                       \
    cudaError_t _e = (expr);                                 \
    if (_e != cudaSuccess) return _e;                        \
  } while (0)

// ---- type conversion ------------------------------------------------------
__device__ __forceinline__ float to_f32(float v) { return v; }
__device__ __forceinline__ float to_f32(bf16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f32<bf16>(float v) { return __float2bfloat16_rn(v); }

// 8 consecutive elements -> fp32 registers (16 B load for bf16, 2x16 B for fp32)
__device__ __forceinline__ void load8(const float* p, float (&v)[8]) {
  float4 a = *reinterpret_cast<const float4*>(p);
  float4 b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
  v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void load8(const bf16* p, float (&v)[8]) {
  uint4 r = *reinterpret_cast<const uint4*>(p);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; i++) {
    float2 f = __bfloat1622float2(h[i]);
    v[2 * i] = f.x;
    v[2 * i + 1] = f.y;
  }
}
// streaming variant for weights read exactly once (do not pollute L1)
__device__ __forceinline__ void load8_stream(const bf16* p, float (&v)[8]) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; i++) {
    float2 f = __bfloat1622float2(h[i]);
    v[2 * i] = f.x;
    v[2 * i + 1] = f.y;
  }
}
__device__ __forceinline__ void load8_stream(const float* p, float (&v)[8]) {
  float4 a, b;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w)
               : "l"(p));
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
               : "l"(p + 4));
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
  v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void store8(float* p, const float (&v)[8]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}
__device__ __forceinline__ void store8(bf16* p, const float (&v)[8]) {
  uint4 r;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; i++) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
  *reinterpret_cast<uint4*>(p) = r;
}

// ---- activations (shared by act_kernel and the GEMM's fused epilogue) ---------
__device__ __forceinline__ float gelu_tanh(float x) {  // tensor.go:181-190
  float x3 = x * x * x;
  float inner = 0.7978845608028654f * (x + 0.044715f * x3);
  return 0.5f * x * (1.0f + tanhf(inner));
}
__device__ __forceinline__ float silu(float x) {  // mamba2.go:360-367
  return x / (1.0f + expf(-x));
}

// ---- reductions -------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// block-wide sum; `red` is >= 32 floats of shared memory; all threads get the result
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();  // protect `red` from a previous use
  if (lane == 0) red[w] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : 0.f;
  r = warp_sum(r);
  return r;
}
__device__ __forceinline__ float block_max(float v, float* red) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : -INFINITY;
  r = warp_max(r);
  return r;
}

// Programmatic dependent launch (PDL): let the next kernel's prologue start early,
// and wait for the previous kernel's memory to be visible.
__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// Every kernel of the forward pass is launched with the programmatic-stream-serialization
// attribute: it may start while its predecessor drains, runs its independent prologue
// (barrier init, TMEM alloc, weight prefetch) and blocks in griddepcontrol.wait until the
// predecessor's memory is visible.  Kernels call pdl_launch_dependents() at the top and
// pdl_wait() before the first dependent global access.
bool pdl_enabled();
void set_pdl_enabled(bool on);
bool carveout_known(const void* fn);  // true once (fn, current device) has been configured
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                   cudaStream_t st, Args&&... args) {
  // Every kernel of the chain asks for the same (maximum) shared-memory carve-out: an SM only
  // changes its L1/shared split when idle, so a kernel with a different preference could not
  // become co-resident with its predecessor and the PDL overlap would be lost.
  if (!carveout_known((const void*)kernel)) {
    cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                         cudaSharedmemCarveoutMaxShared);
    (void)cudaGetLastError();
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(std::forward<Args>(args))...);
}
// the same as a thread-block-cluster launch: `cluster_x` consecutive CTAs (grid.x must be a
// multiple) are co-scheduled on one GPC and can address each other's shared memory
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_k_cluster(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                           cudaStream_t st, int cluster_x, Args&&... args) {
  if (!carveout_known((const void*)kernel)) {
    cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                         cudaSharedmemCarveoutMaxShared);
    (void)cudaGetLastError();
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = (unsigned)cluster_x;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(std::forward<Args>(args))...);
}

// ---- device timeline trace (diagnostics; nvllm_trace_begin / nvllm_trace_end) -----------
// Record = 8 x u64: launch seq, kernel id, CTA, SM, t_entry, t_after_wait, t_exit, extra
// (ns of %globaltimer).  buf[0] = records written, buf[1] = capacity; records start at buf+8.
// `agg` (nvllm_set_profile mode 2) is the per-LAUNCH summary the timeline profile is built from:
// agg[4 seq + {0,1,2}] = {min t_entry, min t_after_wait, max t_exit} over the launch's CTAs.
struct Trace {
  unsigned long long* buf;
  int seq, kid;
  unsigned long long* agg;
};
bool trace_active();       // a device timeline is being recorded: kernel arguments carry its buffer
Trace next_trace(int kid);  // {nullptr} unless tracing is on (host side, gemm_tcgen05.cu)
int trace_seq();            // launches traced so far (host side)
void trace_agg_set(unsigned long long* agg, int capacity);
__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ bool trace_on(const Trace& t) { return t.buf != nullptr || t.agg != nullptr; }
__device__ __forceinline__ void trace_rec(const Trace& t, unsigned long long t0,
                                          unsigned long long t1, unsigned long long extra) {
  if (t.agg) {
    unsigned long long* a = t.agg + 4 * (size_t)t.seq;
    atomicMin(a, t0);
    if (t1) atomicMin(a + 1, t1);
    atomicMax(a + 2, gtime());
  }
  if (!t.buf) return;
  const unsigned long long t2 = gtime();
  const unsigned long long slot = atomicAdd(t.buf, 1ull);
  if (slot >= t.buf[1]) return;
  unsigned sm;
  asm volatile("mov.u32 %0, %%smid;" : "=r"(sm));
  unsigned long long* r = t.buf + 8 + slot * 8;
  r[0] = t.seq; r[1] = t.kid;
  r[2] = blockIdx.x + (unsigned long long)gridDim.x * (blockIdx.y + (unsigned long long)gridDim.y * blockIdx.z);
  r[3] = sm; r[4] = t0; r[5] = t1; r[6] = t2; r[7] = extra;
}
// kernels whose only trace points are entry and exit
#define TRACE_ENTRY(tr) const unsigned long long _tr_entry = trace_on(tr) ? gtime() : 0ull
#define TRACE_EXIT(tr)                                                              \
  do {                                                                              \
    if (threadIdx.x == 0 && threadIdx.y == 0 && trace_on(tr)) trace_rec(tr, _tr_entry, 0ull, 0ull); \
  } while (0)

__host__ __device__ static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
static inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }

// sum of S split-K slices of 8 consecutive floats, slices `stride` apart; the loads of four
// slices are issued together (one L2 round trip per group instead of one per slice)
__device__ __forceinline__ void sum_partials8(const float* __restrict__ p, int S, size_t stride,
                                              float (&acc)[8]) {
#pragma unroll
  for (int j = 0; j < 8; j++) acc[j] = 0.f;
  int s = 0;
  for (; s + 8 <= S; s += 8) {
    float a[8][8];
#pragma unroll
    for (int q = 0; q < 8; q++) load8(p + (size_t)(s + q) * stride, a[q]);
#pragma unroll
    for (int q = 0; q < 8; q++)
#pragma unroll
      for (int j = 0; j < 8; j++) acc[j] += a[q][j];
  }
  for (; s + 4 <= S; s += 4) {
    float a[8], b[8], c[8], d[8];
    load8(p + (size_t)s * stride, a);
    load8(p + (size_t)(s + 1) * stride, b);
    load8(p + (size_t)(s + 2) * stride, c);
    load8(p + (size_t)(s + 3) * stride, d);
#pragma unroll
    for (int j = 0; j < 8; j++) acc[j] = (((acc[j] + a[j]) + b[j]) + c[j]) + d[j];
  }
  for (; s < S; s++) {
    float a[8];
    load8(p + (size_t)s * stride, a);
#pragma unroll
    for (int j = 0; j < 8; j++) acc[j] += a[j];
  }
}


}  // namespace nvllm
