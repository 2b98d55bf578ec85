// ops.cu — single-operator C-ABI entry points: the same kernels the forward pass uses, on
// host buffers, for parity tests against the oracle and for roofline timing.
#include <vector>

#include "common.cuh"
#include "handle.h"

using namespace nvllm;

namespace {
struct Scratch {  // frees everything on scope exit
  std::vector<void*> ptrs;
  cudaStream_t st = nullptr;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  ~Scratch() {
    for (void* p : ptrs) cudaFree(p);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    if (st) cudaStreamDestroy(st);
  }
  cudaError_t alloc(void** p, size_t bytes) {
    cudaError_t e = cudaMalloc(p, bytes ? bytes : 16);
    if (e == cudaSuccess) ptrs.push_back(*p);
    return e;
  }
};
#define OP_CHECK(expr)                                                                      \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess)                                                                  \
      return set_err(nullptr, NVLLM_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(_e));     \
  } while (0)

int op_begin(Scratch& s, int device) {
  OP_CHECK(cudaSetDevice(device));
  OP_CHECK(cudaStreamCreateWithFlags(&s.st, cudaStreamNonBlocking));
  OP_CHECK(cudaEventCreate(&s.e0));
  OP_CHECK(cudaEventCreate(&s.e1));
  return 0;
}
constexpr size_t FLUSH_FLOATS = (size_t)64 << 20;  // 256 MB > 126 MB L2
}  // namespace

extern "C" int nvllm_op_matmul(int device, int prec, int kernel, int M, int N, int K,
                               const float* A, const float* W, const float* bias, float* C,
                               int iters, int flush_l2, float* avg_ms) {
  if (M <= 0 || N <= 0 || K <= 0 || (K % 8) != 0 || !A || !W || !C)
    return set_err(nullptr, NVLLM_ERR_INVALID, "op_matmul: bad shape (K must be a multiple of 8)");
  Scratch s;
  int rc = op_begin(s, device);
  if (rc) return rc;
  const int dt = prec == NVLLM_PREC_BF16 ? DT_BF16 : DT_F32;
  if (kernel == 0) kernel = (M <= 8) ? 1 : ((dt == DT_BF16 && tc5_supported(M, N, K)) ? 3 : 2);
  if (kernel == 1 && M > 8) return set_err(nullptr, NVLLM_ERR_INVALID, "gemv kernel needs M <= 8");
  if (kernel == 3 && (dt != DT_BF16 || !tc5_supported(M, N, K)))
    return set_err(nullptr, NVLLM_ERR_INVALID, "tcgen05 kernel needs bf16 and a supported shape");
  float *dA32, *dW32, *dOut, *dC, *dFlush = nullptr;
  void *dA, *dWt;
  const int ldo = (N + 7) / 8 * 8;
  OP_CHECK(s.alloc((void**)&dA32, (size_t)M * K * 4));
  OP_CHECK(s.alloc((void**)&dW32, (size_t)K * N * 4));
  OP_CHECK(s.alloc(&dA, (size_t)M * K * dt_size(dt)));
  OP_CHECK(s.alloc(&dWt, (size_t)N * K * dt_size(dt)));
  const size_t cap = (size_t)std::max(M, 512) * ldo;
  OP_CHECK(s.alloc((void**)&dOut, cap * 4));
  OP_CHECK(s.alloc((void**)&dC, (size_t)M * ldo * 4));
  if (flush_l2) OP_CHECK(s.alloc((void**)&dFlush, FLUSH_FLOATS * 4));
  OP_CHECK(cudaMemcpyAsync(dA32, A, (size_t)M * K * 4, cudaMemcpyHostToDevice, s.st));
  OP_CHECK(cudaMemcpyAsync(dW32, W, (size_t)K * N * 4, cudaMemcpyHostToDevice, s.st));
  OP_CHECK(launch_convert(s.st, NVLLM_DT_F32, dA32, (size_t)M * K, dt, dA));
  OP_CHECK(launch_transpose_convert(s.st, NVLLM_DT_F32, dW32, (size_t)N, K, N, dt, dWt, K));
  int S = 1;
  auto launch = [&]() -> cudaError_t {
    if (kernel == 3) {
      Tc5Args t{};
      t.A = dA; t.W = dWt; t.M = M; t.N = N; t.K = K; t.out = dOut; t.ldo = ldo;
      t.split_stride = (size_t)M * ldo; t.alpha = 1.f; t.out_capacity = cap;
      t.S = tc5_splits(M, N, K, 148);
      while (t.S > 1 && (size_t)t.S * M * ldo > cap) t.S--;
      return launch_gemm_tc5(s.st, t, &S);
    }
    GemmArgs g{};
    g.A = dA; g.adt = dt; g.lda = K; g.W = dWt; g.wdt = dt; g.M = M; g.N = N; g.K = K;
    g.out = dOut; g.ldo = ldo; g.split_stride = (size_t)M * ldo; g.alpha = 1.f;
    g.out_capacity = cap;
    cudaError_t e = (kernel == 1) ? launch_gemv(s.st, g) : launch_gemm_simt(s.st, g);
    S = g.S;
    return e;
  };
  OP_CHECK(launch());
  OP_CHECK(cudaStreamSynchronize(s.st));
  float total = 0.f;
  for (int it = 0; it < iters; it++) {
    if (flush_l2) OP_CHECK(launch_fill(s.st, dFlush, FLUSH_FLOATS, (float)it));
    OP_CHECK(cudaEventRecord(s.e0, s.st));
    OP_CHECK(launch());
    OP_CHECK(cudaEventRecord(s.e1, s.st));
    OP_CHECK(cudaStreamSynchronize(s.st));
    float ms;
    OP_CHECK(cudaEventElapsedTime(&ms, s.e0, s.e1));
    total += ms;
  }
  if (avg_ms) *avg_ms = iters > 0 ? total / iters : 0.f;
  // sum the split-K slices and add the bias the way the consumer kernels do
  OP_CHECK(launch_reduce_partials(s.st, dOut, S, (size_t)M * ldo, (size_t)M * ldo, dC));
  OP_CHECK(cudaStreamSynchronize(s.st));
  std::vector<float> hc((size_t)M * ldo);
  OP_CHECK(cudaMemcpy(hc.data(), dC, hc.size() * 4, cudaMemcpyDeviceToHost));
  for (int m = 0; m < M; m++)
    for (int n = 0; n < N; n++) C[(size_t)m * N + n] = hc[(size_t)m * ldo + n] + (bias ? bias[n] : 0.f);
  return NVLLM_OK;
}

extern "C" int nvllm_op_norm(int device, int prec, int rows, int H, const float* x, const float* w,
                             const float* b, float eps, float* y, int iters, float* avg_ms) {
  if (rows <= 0 || H <= 0 || (H % 8) != 0 || !x || !w || !y)
    return set_err(nullptr, NVLLM_ERR_INVALID, "op_norm: bad shape (H must be a multiple of 8)");
  Scratch s;
  int rc = op_begin(s, device);
  if (rc) return rc;
  const int dt = prec == NVLLM_PREC_BF16 ? DT_BF16 : DT_F32;
  float *dx, *dw, *db = nullptr, *dy32;
  void* dy;
  OP_CHECK(s.alloc((void**)&dx, (size_t)rows * H * 4));
  OP_CHECK(s.alloc((void**)&dw, (size_t)H * 4));
  OP_CHECK(s.alloc(&dy, (size_t)rows * H * dt_size(dt)));
  OP_CHECK(s.alloc((void**)&dy32, (size_t)rows * H * 4));
  OP_CHECK(cudaMemcpyAsync(dx, x, (size_t)rows * H * 4, cudaMemcpyHostToDevice, s.st));
  OP_CHECK(cudaMemcpyAsync(dw, w, (size_t)H * 4, cudaMemcpyHostToDevice, s.st));
  if (b) {
    OP_CHECK(s.alloc((void**)&db, (size_t)H * 4));
    OP_CHECK(cudaMemcpyAsync(db, b, (size_t)H * 4, cudaMemcpyHostToDevice, s.st));
  }
  NormArgs a{};
  a.x = dx; a.w = dw; a.b = db; a.eps = eps; a.y = dy; a.ydt = dt; a.n_rows = rows; a.H = H;
  OP_CHECK(launch_norm(s.st, a));
  OP_CHECK(cudaStreamSynchronize(s.st));
  float total = 0.f;
  for (int it = 0; it < iters; it++) {
    OP_CHECK(cudaEventRecord(s.e0, s.st));
    OP_CHECK(launch_norm(s.st, a));
    OP_CHECK(cudaEventRecord(s.e1, s.st));
    OP_CHECK(cudaStreamSynchronize(s.st));
    float ms;
    OP_CHECK(cudaEventElapsedTime(&ms, s.e0, s.e1));
    total += ms;
  }
  if (avg_ms) *avg_ms = iters > 0 ? total / iters : 0.f;
  OP_CHECK(launch_convert(s.st, dt == DT_BF16 ? NVLLM_DT_BF16 : NVLLM_DT_F32, dy, (size_t)rows * H,
                          DT_F32, dy32));
  OP_CHECK(cudaStreamSynchronize(s.st));
  OP_CHECK(cudaMemcpy(y, dy32, (size_t)rows * H * 4, cudaMemcpyDeviceToHost));
  return NVLLM_OK;
}

extern "C" int nvllm_op_argmax(int device, int rows, int V, const float* logits, int32_t* out,
                               int iters, float* avg_ms) {
  if (rows <= 0 || V <= 0 || !logits || !out)
    return set_err(nullptr, NVLLM_ERR_INVALID, "op_argmax: bad shape");
  Scratch s;
  int rc = op_begin(s, device);
  if (rc) return rc;
  float* dl;
  int* dout;
  OP_CHECK(s.alloc((void**)&dl, (size_t)rows * V * 4));
  OP_CHECK(s.alloc((void**)&dout, (size_t)rows * 4));
  OP_CHECK(cudaMemcpyAsync(dl, logits, (size_t)rows * V * 4, cudaMemcpyHostToDevice, s.st));
  OP_CHECK(launch_sample(s.st, dl, V, V, rows, nullptr, nullptr, dout));
  OP_CHECK(cudaStreamSynchronize(s.st));
  float total = 0.f;
  for (int it = 0; it < iters; it++) {
    OP_CHECK(cudaEventRecord(s.e0, s.st));
    OP_CHECK(launch_sample(s.st, dl, V, V, rows, nullptr, nullptr, dout));
    OP_CHECK(cudaEventRecord(s.e1, s.st));
    OP_CHECK(cudaStreamSynchronize(s.st));
    float ms;
    OP_CHECK(cudaEventElapsedTime(&ms, s.e0, s.e1));
    total += ms;
  }
  if (avg_ms) *avg_ms = iters > 0 ? total / iters : 0.f;
  OP_CHECK(cudaMemcpy(out, dout, (size_t)rows * 4, cudaMemcpyDeviceToHost));
  return NVLLM_OK;
}
