// noalloc.cu — ONE allocation-free C-ABI entry point: the split-K GEMV (gemm_simt.cu, the M <= 8
// kernel of single-sequence decode) with its fp32 split-K slices in a file-scope device array
// instead of a cudaMalloc'ed workspace.  For callers that cannot allocate: code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc.  No counterpart in the
// reference.
//
//   C[M, N] (fp32) = A[M, K] @ W[N, K]^T     A, W, C are DEVICE pointers the caller owns.
//
// The scratch array is part of the library image: every GPU that loads the library holds its
// 4 MiB until the process exits.  All calls share it — it keeps the last call's contents, and
// two calls that overlap on the GPU (different streams, or one stream captured into graphs that
// run concurrently) race on it.  Nothing in it has to start at zero: the reduce kernel reads
// exactly the S x M x N floats the GEMV of the same call wrote before it on the same stream.
#include <mutex>

#include "common.cuh"
#include "handle.h"

using namespace nvllm;

constexpr size_t NOALLOC_FLOATS = (size_t)NVLLM_NOALLOC_MAX_SCRATCH_FLOATS;
// the split-K slices [S][M][N]
static __device__ __align__(256) float g_gemv_scratch[NOALLOC_FLOATS];

namespace {

constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
float* g_scratch_addr[MAX_DEVICES] = {nullptr};  // the array's DEVICE address, per device

// In host code `g_gemv_scratch` is only a stand-in for the symbol, not a device address: the
// address is looked up once per device, with that device current.
int scratch_for_device(int device, float** out) {
  std::lock_guard<std::mutex> lock(g_mu);
  if (!g_scratch_addr[device]) {
    float* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_gemv_scratch);
    if (e != cudaSuccess || !p)
      return set_err(nullptr, NVLLM_ERR_CUDA, "gemv_noalloc: cudaGetSymbolAddress on device %d: %s",
                     device, cudaGetErrorString(e));
    g_scratch_addr[device] = p;
  }
  *out = g_scratch_addr[device];
  return NVLLM_OK;
}

}  // namespace

extern "C" int nvllm_op_gemv_noalloc(int device, int prec, int M, int N, int K, const void* d_A,
                                     const void* d_W, float* d_C, void* stream) {
  if (device < 0 || device >= MAX_DEVICES)
    return set_err(nullptr, NVLLM_ERR_INVALID, "gemv_noalloc: device %d out of range", device);
  if (prec != NVLLM_PREC_FP32 && prec != NVLLM_PREC_BF16)
    return set_err(nullptr, NVLLM_ERR_INVALID, "gemv_noalloc: precision %d", prec);
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess)
    return set_err(nullptr, NVLLM_ERR_CUDA, "gemv_noalloc: cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
  float* scratch = nullptr;
  int rc = scratch_for_device(device, &scratch);
  if (rc) return rc;
  if (M == 0 && N == 0 && K == 0) return NVLLM_OK;  // set-up only (do this outside a stream capture)
  if (M < 1 || M > 8 || N < 1 || K < 8 || (K % 8) != 0)
    return set_err(nullptr, NVLLM_ERR_INVALID,
                   "gemv_noalloc: shape M=%d N=%d K=%d (1 <= M <= 8, N >= 1, K a multiple of 8)", M, N, K);
  if (!d_A || !d_W || !d_C || ((uintptr_t)d_A & 15) || ((uintptr_t)d_W & 15) || ((uintptr_t)d_C & 3))
    return set_err(nullptr, NVLLM_ERR_INVALID, "gemv_noalloc: null or misaligned pointer (A, W: 16 bytes)");
  const size_t need = (size_t)gemv_splits(M, N, K) * M * N;
  if (need > NOALLOC_FLOATS)
    return set_err(nullptr, NVLLM_ERR_INVALID,
                   "gemv_noalloc: M=%d N=%d K=%d needs %zu scratch floats (%d split-K slices), the "
                   "static scratch holds %zu; use nvllm_op_matmul or a handle",
                   M, N, K, need, gemv_splits(M, N, K), NOALLOC_FLOATS);
  cudaStream_t st = (cudaStream_t)stream;
  const int dt = prec == NVLLM_PREC_BF16 ? DT_BF16 : DT_F32;
  GemmArgs g{};
  g.A = d_A; g.adt = dt; g.lda = K; g.W = d_W; g.wdt = dt; g.M = M; g.N = N; g.K = K;
  g.out = scratch; g.ldo = N; g.split_stride = (size_t)M * N; g.alpha = 1.f;
  g.out_capacity = NOALLOC_FLOATS;
  e = launch_gemv(st, g);
  if (e != cudaSuccess)
    return set_err(nullptr, NVLLM_ERR_CUDA, "gemv_noalloc: gemv launch: %s", cudaGetErrorString(e));
  e = launch_reduce_partials(st, scratch, g.S, (size_t)M * N, (size_t)M * N, d_C);
  if (e != cudaSuccess)
    return set_err(nullptr, NVLLM_ERR_CUDA, "gemv_noalloc: reduce launch: %s", cudaGetErrorString(e));
  return NVLLM_OK;
}
