// tp.cu — tensor-parallel plumbing (Falcon-7B): one handle per GPU and per
// process (or per thread), joined in an NCCL communicator over NVLink.  The reference has no
// distributed code at all (Config.TensorParallelSize is validated and never read,
// nanovllm/config.go:61-63); the split is Megatron's: Q/K/V and W1 by columns, O and W2 by
// rows, one sum over ranks per residual update, LM head by vocabulary rows + all-gather.
//
// NCCL is resolved at run time with dlopen("libnccl.so.2"): a process that already loaded a
// libnccl (torch's bundled one) shares it, a plain C / cgo caller gets the system library, and a
// single-GPU user never needs NCCL to be installed.
#include <dlfcn.h>
#include <nccl.h>
#include <string.h>

#include <mutex>

#include "common.cuh"
#include "handle.h"

namespace nvllm {

namespace {
struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                            cudaStream_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t,
                            cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  std::string why;
};

NcclApi* nccl_api() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
      api.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
      if (api.lib) break;
    }
    if (!api.lib) {
      api.why = std::string("cannot load libnccl.so.2: ") + dlerror();
      return;
    }
#define SYM(field, name)                                              \
  api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.lib, name)); \
  if (!api.field) {                                                   \
    api.why = std::string("libnccl lacks ") + name;                   \
    api.lib = nullptr;                                                \
    return;                                                           \
  }
    SYM(GetUniqueId, "ncclGetUniqueId");
    SYM(CommInitRank, "ncclCommInitRank");
    SYM(CommDestroy, "ncclCommDestroy");
    SYM(AllReduce, "ncclAllReduce");
    SYM(AllGather, "ncclAllGather");
    SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
  });
  return &api;
}
}  // namespace

int tp_unique_id(uint8_t out_id[128]) {
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  memset(out_id, 0, 128);
  NcclApi* a = nccl_api();
  if (!a->lib) return set_err(nullptr, NVLLM_ERR_NCCL, "%s", a->why.c_str());
  ncclUniqueId id;
  ncclResult_t r = a->GetUniqueId(&id);
  if (r != ncclSuccess)
    return set_err(nullptr, NVLLM_ERR_NCCL, "ncclGetUniqueId: %s", a->GetErrorString(r));
  memcpy(out_id, &id, 128);
  return NVLLM_OK;
}

int tp_init(nvllm_handle* h, const uint8_t* id_bytes) {
  NcclApi* a = nccl_api();
  if (!a->lib) return set_err(h, NVLLM_ERR_NCCL, "%s", a->why.c_str());
  ncclUniqueId id;
  memcpy(&id, id_bytes, 128);
  ncclComm_t comm = nullptr;
  ncclResult_t r = a->CommInitRank(&comm, h->tp, id, h->rank);
  if (r != ncclSuccess)
    return set_err(h, NVLLM_ERR_NCCL, "ncclCommInitRank(rank %d of %d): %s", h->rank, h->tp,
                   a->GetErrorString(r));
  h->nccl = comm;
  return NVLLM_OK;
}

void tp_destroy(nvllm_handle* h) {
  if (!h->nccl) return;
  NcclApi* a = nccl_api();
  if (a->lib) a->CommDestroy((ncclComm_t)h->nccl);
  h->nccl = nullptr;
}

int tp_all_reduce_f32(nvllm_handle* h, float* buf, size_t count) {
  NcclApi* a = nccl_api();
  ncclResult_t r = a->AllReduce(buf, buf, count, ncclFloat32, ncclSum, (ncclComm_t)h->nccl, h->stream);
  if (r != ncclSuccess)
    return set_err(h, NVLLM_ERR_NCCL, "ncclAllReduce: %s", a->GetErrorString(r));
  return NVLLM_OK;
}

int tp_all_gather_f32(nvllm_handle* h, float* buf, size_t count_per_rank) {
  NcclApi* a = nccl_api();
  ncclResult_t r = a->AllGather(buf + (size_t)h->rank * count_per_rank, buf, count_per_rank,
                                ncclFloat32, (ncclComm_t)h->nccl, h->stream);
  if (r != ncclSuccess)
    return set_err(h, NVLLM_ERR_NCCL, "ncclAllGather: %s", a->GetErrorString(r));
  return NVLLM_OK;
}

// ---------------------------------------------------------------------------------------
// This rank's contribution to a residual update, ready for the sum over ranks:
//   out[t, :] = m1 * (sum_s p1[s][t, :] + bias1) + m2 * (sum_s p2[s][t, :] + bias2)
// (generic_model.go:320-326 / :359-365 / :405-416 — the update itself is applied by the norm
// kernel after the all-reduce.)  Row-parallel biases are zero on ranks other than 0.
__global__ void tp_partial_kernel(size_t n8, int H, const float* __restrict__ p1, int S1,
                                  size_t stride1, float m1, const float* __restrict__ bias1,
                                  const float* __restrict__ p2, int S2, size_t stride2, float m2,
                                  const float* __restrict__ bias2, float* __restrict__ out) {
  pdl_launch_dependents();
  pdl_wait();
  for (size_t v = (size_t)blockIdx.x * blockDim.x + threadIdx.x; v < n8;
       v += (size_t)gridDim.x * blockDim.x) {
    const size_t i = v * 8;
    const int col = (int)(i % H);
    float acc[8], o[8];
#pragma unroll
    for (int j = 0; j < 8; j++) acc[j] = 0.f;
    for (int s = 0; s < S1; s++) {
      float a[8];
      load8(p1 + (size_t)s * stride1 + i, a);
#pragma unroll
      for (int j = 0; j < 8; j++) acc[j] += a[j];
    }
    if (bias1) {
      float b[8];
      load8(bias1 + col, b);
#pragma unroll
      for (int j = 0; j < 8; j++) acc[j] += b[j];
    }
#pragma unroll
    for (int j = 0; j < 8; j++) o[j] = m1 * acc[j];
    if (p2) {
#pragma unroll
      for (int j = 0; j < 8; j++) acc[j] = 0.f;
      for (int s = 0; s < S2; s++) {
        float a[8];
        load8(p2 + (size_t)s * stride2 + i, a);
#pragma unroll
        for (int j = 0; j < 8; j++) acc[j] += a[j];
      }
      if (bias2) {
        float b[8];
        load8(bias2 + col, b);
#pragma unroll
        for (int j = 0; j < 8; j++) acc[j] += b[j];
      }
#pragma unroll
      for (int j = 0; j < 8; j++) o[j] += m2 * acc[j];
    }
    store8(out + i, o);
  }
}

cudaError_t launch_tp_partial(cudaStream_t st, int T, int H, const float* p1, int S1, size_t stride1,
                              float m1, const float* bias1, const float* p2, int S2, size_t stride2,
                              float m2, const float* bias2, float* out) {
  const size_t n8 = (size_t)T * H / 8;
  if (!n8) return cudaSuccess;
  const int blocks = (int)std::min<size_t>((size_t)148 * 8, (n8 + 255) / 256);
  return launch_k(tp_partial_kernel, dim3(blocks), dim3(256), 0, st, n8, H, p1, S1, stride1, m1,
                  bias1, p2, S2, stride2, m2, bias2, out);
}

}  // namespace nvllm
