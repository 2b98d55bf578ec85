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
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <vector>

#include "common.cuh"
#include "handle.h"

namespace nvllm {

// peer-memory region of one rank (see the all-reduce kernels below):
//   part[2][cap] | res[2][cap] fp32 | f1[8][128] | f2[8][128] u32 | ll[2][8][P2P_LL_CAP] x 8 B
constexpr size_t P2P_FLAG_BYTES = (size_t)2 * 8 * 128 * 4;
constexpr size_t P2P_LL_CAP = 256 * 1024;  // floats per (parity, source rank) LL lane
constexpr size_t P2P_LL_BYTES = (size_t)2 * 8 * P2P_LL_CAP * 8;

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
  // optional (NCCL >= 2.18): a second communicator over the same ranks for the pipelined prefill
  ncclResult_t (*CommSplit)(ncclComm_t, int, int, ncclComm_t*, ncclConfig_t*) = nullptr;
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
    api.CommSplit = reinterpret_cast<decltype(api.CommSplit)>(dlsym(api.lib, "ncclCommSplit"));
  });
  return &api;
}
}  // namespace

// Peer-memory region exchange: every rank allocates its region, the 64-byte cudaIpc handles
// travel through one ncclAllGather, every rank maps the others.  All ranks must agree on the
// outcome (a kernel that spins on a peer that fell back to NCCL would hang), hence the final
// min-reduce of the success flag.  NVLLM_TP_P2P=0 forces the NCCL path.
static void p2p_setup(nvllm_handle* h) {
  NcclApi* a = nccl_api();
  const char* env = getenv("NVLLM_TP_P2P");
  int ok = (env && atoi(env) == 0) ? 0 : 1;
  if (h->tp > 8) ok = 0;
  const size_t cap = ((size_t)h->cfg.max_batched_tokens * h->H + 7) & ~(size_t)7;
  const size_t bytes = 4 * cap * 4 + P2P_FLAG_BYTES + P2P_LL_BYTES;
  float* own = nullptr;
  if (ok && cudaMalloc((void**)&own, bytes) != cudaSuccess) {
    (void)cudaGetLastError();
    ok = 0;
    own = nullptr;
  }
  if (own) cudaMemset(own, 0, bytes);
  cudaIpcMemHandle_t mine;
  memset(&mine, 0, sizeof(mine));
  if (ok && cudaIpcGetMemHandle(&mine, own) != cudaSuccess) {
    (void)cudaGetLastError();
    ok = 0;
  }
  // gather handles + flags: [tp][64 + 64] bytes
  constexpr size_t REC = 128;
  uint8_t* d_all = nullptr;
  std::vector<uint8_t> all((size_t)h->tp * REC, 0);
  if (cudaMalloc((void**)&d_all, all.size()) != cudaSuccess) {
    if (own) cudaFree(own);
    return;
  }
  uint8_t rec[REC] = {0};
  memcpy(rec, &mine, sizeof(mine));
  rec[64] = (uint8_t)ok;
  cudaMemcpy(d_all + (size_t)h->rank * REC, rec, REC, cudaMemcpyHostToDevice);
  ncclResult_t r = a->AllGather(d_all + (size_t)h->rank * REC, d_all, REC, ncclUint8,
                                (ncclComm_t)h->nccl, h->stream);
  if (r != ncclSuccess || cudaStreamSynchronize(h->stream) != cudaSuccess) ok = 0;
  cudaMemcpy(all.data(), d_all, all.size(), cudaMemcpyDeviceToHost);
  for (int q = 0; q < h->tp; q++)
    if (!all[(size_t)q * REC + 64]) ok = 0;
  float* peer[8] = {nullptr};
  if (ok) {
    for (int q = 0; q < h->tp && ok; q++) {
      if (q == h->rank) {
        peer[q] = own;
        continue;
      }
      cudaIpcMemHandle_t hd;
      memcpy(&hd, &all[(size_t)q * REC], sizeof(hd));
      void* ptr = nullptr;
      if (cudaIpcOpenMemHandle(&ptr, hd, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
        (void)cudaGetLastError();
        ok = 0;
      } else {
        peer[q] = (float*)ptr;
      }
    }
  }
  // second round: did every rank map every peer?
  rec[64] = (uint8_t)ok;
  cudaMemcpy(d_all + (size_t)h->rank * REC, rec, REC, cudaMemcpyHostToDevice);
  r = a->AllGather(d_all + (size_t)h->rank * REC, d_all, REC, ncclUint8, (ncclComm_t)h->nccl,
                   h->stream);
  if (r != ncclSuccess || cudaStreamSynchronize(h->stream) != cudaSuccess) ok = 0;
  cudaMemcpy(all.data(), d_all, all.size(), cudaMemcpyDeviceToHost);
  for (int q = 0; q < h->tp; q++)
    if (!all[(size_t)q * REC + 64]) ok = 0;
  cudaFree(d_all);
  if (!ok) {
    for (int q = 0; q < h->tp; q++)
      if (peer[q] && peer[q] != own) cudaIpcCloseMemHandle(peer[q]);
    if (own) cudaFree(own);
    return;
  }
  for (int q = 0; q < h->tp; q++) h->p2p_peer[q] = peer[q];
  h->p2p_own = own;
  h->p2p_cap = cap;
  h->p2p_epoch = 0;
  h->p2p_ok = true;
  if (getenv("NVLLM_TP_DEBUG"))
    fprintf(stderr, "[nvllm] rank %d/%d: NVLink peer-memory all-reduce enabled (%zu MB region)\n",
            h->rank, h->tp, bytes >> 20);
}

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
  // second communicator (same ranks) + stream + events for the two-chunk prefill pipeline
  // (forward.cu): two chunks may then have collectives in flight at the same time.  Every rank
  // takes the same branch (same library, same environment).
  // OFF by default (NVLLM_TP_PIPE=1 turns it on): measured at tp 2 on Falcon-7B the pipelined
  // 2048-token prefill is slower (19.9 vs 17.7 ms) — two chunks' persistent GEMMs contend for the
  // SMs (the second kernel's first CTAs land on the SMs meant for the collective), which costs
  // more than the 15 % of exposed all-reduce it hides.
  if (a->CommSplit && getenv("NVLLM_TP_PIPE") && atoi(getenv("NVLLM_TP_PIPE")) != 0) {
    ncclComm_t c2 = nullptr;
    if (a->CommSplit(comm, 0, h->rank, &c2, nullptr) == ncclSuccess && c2) {
      h->nccl2 = c2;
      bool ok = cudaStreamCreateWithFlags(&h->stream2, cudaStreamNonBlocking) == cudaSuccess &&
                cudaEventCreateWithFlags(&h->pipe_fork, cudaEventDisableTiming) == cudaSuccess &&
                cudaEventCreateWithFlags(&h->pipe_join, cudaEventDisableTiming) == cudaSuccess;
      h->pipe_kv.resize(h->L, nullptr);
      for (int l = 0; ok && l < h->L; l++)
        ok = cudaEventCreateWithFlags(&h->pipe_kv[l], cudaEventDisableTiming) == cudaSuccess;
      if (!ok) {
        (void)cudaGetLastError();
        a->CommDestroy(c2);
        h->nccl2 = nullptr;
      }
    }
  }
  p2p_setup(h);  // optional: on failure the NCCL all-reduce stays in use
  return NVLLM_OK;
}

void tp_destroy(nvllm_handle* h) {
  for (int r = 0; r < 8; r++)
    if (h->p2p_peer[r] && h->p2p_peer[r] != h->p2p_own) cudaIpcCloseMemHandle(h->p2p_peer[r]);
  if (h->p2p_own) cudaFree(h->p2p_own);
  h->p2p_own = nullptr;
  h->p2p_ok = false;
  for (auto e : h->pipe_kv)
    if (e) cudaEventDestroy(e);
  h->pipe_kv.clear();
  if (h->pipe_fork) cudaEventDestroy(h->pipe_fork);
  if (h->pipe_join) cudaEventDestroy(h->pipe_join);
  if (h->stream2) cudaStreamDestroy(h->stream2);
  h->pipe_fork = h->pipe_join = nullptr;
  h->stream2 = nullptr;
  if (!h->nccl) return;
  NcclApi* a = nccl_api();
  if (a->lib && h->nccl2) a->CommDestroy((ncclComm_t)h->nccl2);
  h->nccl2 = nullptr;
  if (a->lib) a->CommDestroy((ncclComm_t)h->nccl);
  h->nccl = nullptr;
}

int tp_all_reduce(nvllm_handle* h, void* buf, size_t count, bool bf16_wire) {
  NcclApi* a = nccl_api();
  ncclResult_t r = a->AllReduce(buf, buf, count, bf16_wire ? ncclBfloat16 : ncclFloat32, ncclSum,
                                (ncclComm_t)h->nccl_cur(), h->stream);
  if (r != ncclSuccess)
    return set_err(h, NVLLM_ERR_NCCL, "ncclAllReduce: %s", a->GetErrorString(r));
  return NVLLM_OK;
}

int tp_all_gather_f32(nvllm_handle* h, float* buf, size_t count_per_rank) {
  NcclApi* a = nccl_api();
  ncclResult_t r = a->AllGather(buf + (size_t)h->rank * count_per_rank, buf, count_per_rank,
                                ncclFloat32, (ncclComm_t)h->nccl_cur(), h->stream);
  if (r != ncclSuccess)
    return set_err(h, NVLLM_ERR_NCCL, "ncclAllGather: %s", a->GetErrorString(r));
  return NVLLM_OK;
}

// ---------------------------------------------------------------------------------------
// All-reduce over NVLink peer memory, fused with the partial-sum kernel that feeds it.
//
// Every rank owns one peer-mapped region (cudaIpc handles exchanged through the NCCL
// communicator at init):  part[2][cap] | res[2][cap] | f1[tp][P2P_GRID] | f2[tp][P2P_GRID]
// (fp32 data, u32 flags; [2] = double buffer by call parity).  One kernel per residual update:
//   1. CTA c folds this rank's split-K slices / multipliers / biases for sub-range c of EVERY
//      rank's slice into part[par] (local stores), fences, and pushes flag f1[me][c] = epoch
//      into every peer's region;
//   2. waits until all peers' f1[q][c] reached the epoch, then sums sub-range c of ITS OWN
//      slice over the ranks in rank order (P2P loads; the same order everywhere, so every rank
//      holds bit-identical sums) — one-shot mode (small messages, decode) does this for all
//      slices and is done;
//   3. two-shot mode (prefill): the reduced slice goes to res[par] + flag f2, and sub-range c of
//      every other rank's reduced slice is pulled from that rank once its f2[q][c] is up.
// A rank can run at most one call ahead of a peer (it needs the peer's flags of call k to
// finish call k), so two buffers per kind are enough and flags only ever grow.
constexpr int P2P_GRID = 128;      // CTAs; all co-resident (<= 148 SMs), same on every rank
constexpr int P2P_THREADS = 512;

struct P2PPeers {
  float* base[8];
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// coherent 32-byte read of (possibly remote) memory another GPU has just written
__device__ __forceinline__ void load8_sys(const float* p, float (&v)[8]) {
  float4 a, b;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w) : "l"(p) : "memory");
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p + 4) : "memory");
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
  v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}

// ---- small messages (decode): one-shot all-reduce with NCCL-LL-style flagged words.
// Every 8 bytes on the wire are {fp32 value, epoch}: a rank PUSHES its addend into lane `me` of
// every peer's LL buffer with 16-byte stores (two flagged words each) and then polls its own
// buffer until each peer's words carry this call's epoch — one NVLink one-way latency, no
// fences, no separate flag round trip.  Sums are taken in rank order on every rank.
__global__ void __launch_bounds__(256)
tp_allreduce_ll_kernel(P2PPeers peers, int tp, int me, size_t ll_off_bytes, uint32_t epoch,
                       size_t n8, int H, const float* __restrict__ p1, int S1, size_t stride1,
                       float m1, const float* __restrict__ bias1, const float* __restrict__ p2,
                       int S2, size_t stride2, float m2, const float* __restrict__ bias2,
                       float* __restrict__ out) {
  pdl_launch_dependents();
  const int par = epoch & 1;
  pdl_wait();
  for (size_t v = (size_t)blockIdx.x * blockDim.x + threadIdx.x; v < n8;
       v += (size_t)gridDim.x * blockDim.x) {
    const size_t i = v * 8;
    const int col = (int)(i % H);
    float acc[8], o[8];
    sum_partials8(p1 + i, S1, stride1, acc);
    if (bias1) {
      float b[8];
      load8(bias1 + col, b);
#pragma unroll
      for (int j = 0; j < 8; j++) acc[j] += b[j];
    }
#pragma unroll
    for (int j = 0; j < 8; j++) o[j] = m1 * acc[j];
    if (p2) {
      sum_partials8(p2 + i, S2, stride2, acc);
      if (bias2) {
        float b[8];
        load8(bias2 + col, b);
#pragma unroll
        for (int j = 0; j < 8; j++) acc[j] += b[j];
      }
#pragma unroll
      for (int j = 0; j < 8; j++) o[j] += m2 * acc[j];
    }
    // push: lane `me` of every rank's buffer (own included), element i
    const size_t lane_off = ((size_t)par * 8 + me) * P2P_LL_CAP + i;  // in 8-byte words
    for (int q = 0; q < tp; q++) {
      uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(peers.base[q]) +
                                            ll_off_bytes) + lane_off / 2;
#pragma unroll
      for (int k = 0; k < 4; k++)
        asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(dst + k),
                     "r"(__float_as_uint(o[2 * k])), "r"(epoch), "r"(__float_as_uint(o[2 * k + 1])),
                     "r"(epoch)
                     : "memory");
    }
    // pull from the local buffer in rank order
#pragma unroll
    for (int j = 0; j < 8; j++) acc[j] = 0.f;
    const long long t0 = clock64();
    for (int r = 0; r < tp; r++) {
      const uint4* src = reinterpret_cast<const uint4*>(reinterpret_cast<const uint8_t*>(peers.base[me]) +
                                                        ll_off_bytes) +
                         (((size_t)par * 8 + r) * P2P_LL_CAP + i) / 2;
#pragma unroll
      for (int k = 0; k < 4; k++) {
        uint4 w;
        do {
          asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                       : "=r"(w.x), "=r"(w.y), "=r"(w.z), "=r"(w.w)
                       : "l"(src + k)
                       : "memory");
        } while ((w.y != epoch || w.w != epoch) && clock64() - t0 < 10000000000ll);
        acc[2 * k] += __uint_as_float(w.x);
        acc[2 * k + 1] += __uint_as_float(w.z);
      }
    }
    store8(out + i, acc);
  }
}

// wait until a peer's flag reaches `epoch`; gives up after 5 s (a peer that died must not hang
// this GPU: the step then returns garbage and the host sees the peer's error)
__device__ __forceinline__ void p2p_spin(const uint32_t* f, uint32_t epoch) {
  const unsigned long long t0 = gtime();
  while ((int32_t)(ld_acquire_sys(f) - epoch) < 0) {
    if (gtime() - t0 > 5000000000ull) break;
  }
}

__global__ void __launch_bounds__(P2P_THREADS, 1)
tp_allreduce_p2p_kernel(P2PPeers peers, int tp, int me, size_t cap, uint32_t epoch, int two_shot,
                        size_t n8, int H, const float* __restrict__ p1, int S1, size_t stride1,
                        float m1, const float* __restrict__ bias1, const float* __restrict__ p2,
                        int S2, size_t stride2, float m2, const float* __restrict__ bias2,
                        float* __restrict__ out) {
  pdl_launch_dependents();
  const int par = epoch & 1;
  const int c = blockIdx.x;
  float* my = peers.base[me];
  float* part = my + (size_t)par * cap;
  float* res = my + (size_t)(2 + par) * cap;
  const size_t flag_off = 4 * cap;  // in 4-byte words from the region base
  // vectors (8 floats) per rank slice and per CTA sub-range of a slice
  const size_t spv = (n8 + tp - 1) / tp;
  const size_t cpv = (spv + P2P_GRID - 1) / P2P_GRID;
  pdl_wait();
  // ---- 1. this rank's addend for sub-range c of every slice
  for (int q = 0; q < tp; q++) {
    const size_t lo = (size_t)q * spv + (size_t)c * cpv;
    const size_t hi = min(min(lo + cpv, (size_t)(q + 1) * spv), n8);
    for (size_t v = lo + threadIdx.x; v < hi; v += P2P_THREADS) {
      const size_t i = v * 8;
      const int col = (int)(i % H);
      float acc[8], o[8];
      sum_partials8(p1 + i, S1, stride1, acc);
      if (bias1) {
        float b[8];
        load8(bias1 + col, b);
#pragma unroll
        for (int j = 0; j < 8; j++) acc[j] += b[j];
      }
#pragma unroll
      for (int j = 0; j < 8; j++) o[j] = m1 * acc[j];
      if (p2) {
        sum_partials8(p2 + i, S2, stride2, acc);
        if (bias2) {
          float b[8];
          load8(bias2 + col, b);
#pragma unroll
          for (int j = 0; j < 8; j++) acc[j] += b[j];
        }
#pragma unroll
        for (int j = 0; j < 8; j++) o[j] += m2 * acc[j];
      }
      store8(part + i, o);
    }
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x < tp)
    st_release_sys(reinterpret_cast<uint32_t*>(peers.base[threadIdx.x]) + flag_off +
                       (size_t)me * P2P_GRID + c, epoch);
  // ---- 2. wait for sub-range c of every peer, reduce in rank order
  if (threadIdx.x < tp) {
    const uint32_t* f = reinterpret_cast<const uint32_t*>(my) + flag_off +
                        (size_t)threadIdx.x * P2P_GRID + c;
    p2p_spin(f, epoch);
  }
  __syncthreads();
  const int q_lo = two_shot ? me : 0, q_hi = two_shot ? me + 1 : tp;
  float* dst = two_shot ? res : out;
  for (int q = q_lo; q < q_hi; q++) {
    const size_t lo = (size_t)q * spv + (size_t)c * cpv;
    const size_t hi = min(min(lo + cpv, (size_t)(q + 1) * spv), n8);
    for (size_t v = lo + threadIdx.x; v < hi; v += P2P_THREADS) {
      const size_t i = v * 8;
      float acc[8];
#pragma unroll
      for (int j = 0; j < 8; j++) acc[j] = 0.f;
      for (int r = 0; r < tp; r++) {
        float a[8];
        load8_sys(peers.base[r] + (size_t)par * cap + i, a);
#pragma unroll
        for (int j = 0; j < 8; j++) acc[j] += a[j];
      }
      store8(dst + i, acc);
      if (two_shot) store8(out + i, acc);
    }
  }
  if (!two_shot) return;
  // ---- 3. publish the reduced slice, pull the others
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x < tp)
    st_release_sys(reinterpret_cast<uint32_t*>(peers.base[threadIdx.x]) + flag_off +
                       (size_t)(tp + me) * P2P_GRID + c, epoch);
  if (threadIdx.x < tp) {
    const uint32_t* f = reinterpret_cast<const uint32_t*>(my) + flag_off +
                        (size_t)(tp + threadIdx.x) * P2P_GRID + c;
    p2p_spin(f, epoch);
  }
  __syncthreads();
  for (int q = 0; q < tp; q++) {
    if (q == me) continue;
    const size_t lo = (size_t)q * spv + (size_t)c * cpv;
    const size_t hi = min(min(lo + cpv, (size_t)(q + 1) * spv), n8);
    const float* src = peers.base[q] + (size_t)(2 + par) * cap;
    for (size_t v = lo + threadIdx.x; v < hi; v += P2P_THREADS) {
      float a[8];
      load8_sys(src + v * 8, a);
      store8(out + v * 8, a);
    }
  }
}

// ---------------------------------------------------------------------------------------
// This rank's contribution to a residual update, ready for the sum over ranks:
//   out[t, :] = m1 * (sum_s p1[s][t, :] + bias1) + m2 * (sum_s p2[s][t, :] + bias2)
// (generic_model.go:320-326 / :359-365 / :405-416 — the update itself is applied by the norm
// kernel after the all-reduce.)  Row-parallel biases are zero on ranks other than 0.
template <typename OT>
__global__ void tp_partial_kernel(size_t n8, int H, const float* __restrict__ p1, int S1,
                                  size_t stride1, float m1, const float* __restrict__ bias1,
                                  const float* __restrict__ p2, int S2, size_t stride2, float m2,
                                  const float* __restrict__ bias2, OT* __restrict__ out) {
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
                              float m2, const float* bias2, void* out, bool out_bf16) {
  const size_t n8 = (size_t)T * H / 8;
  if (!n8) return cudaSuccess;
  const int blocks = (int)std::min<size_t>((size_t)148 * 8, (n8 + 255) / 256);
  if (out_bf16)
    return launch_k(tp_partial_kernel<bf16>, dim3(blocks), dim3(256), 0, st, n8, H, p1, S1, stride1,
                    m1, bias1, p2, S2, stride2, m2, bias2, (bf16*)out);
  return launch_k(tp_partial_kernel<float>, dim3(blocks), dim3(256), 0, st, n8, H, p1, S1, stride1, m1,
                  bias1, p2, S2, stride2, m2, bias2, (float*)out);
}

// NVLLM_TP_TWOSHOT=1 routes EVERY size through the two-shot peer-memory kernel (experiments and
// its parity test).  Default: LL one-shot up to P2P_LL_CAP floats (decode and short prefills);
// larger messages go to ncclAllReduce, which is faster than the two-shot kernel today
// (measured at tp 2, 37 MB: 120 us against 178 us per residual update).
static bool p2p_two_shot_forced() {
  static const bool on = getenv("NVLLM_TP_TWOSHOT") && atoi(getenv("NVLLM_TP_TWOSHOT")) != 0;
  return on;
}
bool tp_p2p_fits(const nvllm_handle* h, size_t count) {
  if (!h->p2p_ok || (count % 8) != 0) return false;
  return p2p_two_shot_forced() ? count <= h->p2p_cap : count <= P2P_LL_CAP;
}

cudaError_t launch_tp_allreduce_p2p(nvllm_handle* h, int T, int H, const float* p1, int S1,
                                    size_t stride1, float m1, const float* bias1, const float* p2,
                                    int S2, size_t stride2, float m2, const float* bias2,
                                    float* out) {
  const size_t n = (size_t)T * H;
  if (!n) return cudaSuccess;
  P2PPeers peers{};
  for (int q = 0; q < h->tp; q++) peers.base[q] = h->p2p_peer[q];
  const uint32_t epoch = ++h->p2p_epoch;
  if (!p2p_two_shot_forced()) {
    const size_t n8 = n / 8;
    const int grid = (int)std::min<size_t>(148, (n8 + 255) / 256);
    return launch_k(tp_allreduce_ll_kernel, dim3(grid), dim3(256), 0, h->stream, peers, h->tp,
                    h->rank, 4 * h->p2p_cap * 4 + P2P_FLAG_BYTES, epoch, n8, H, p1, S1, stride1, m1,
                    bias1, p2, S2, stride2, m2, bias2, out);
  }
  const int two_shot = 1;
  return launch_k(tp_allreduce_p2p_kernel, dim3(P2P_GRID), dim3(P2P_THREADS), 0, h->stream, peers,
                  h->tp, h->rank, h->p2p_cap, epoch, two_shot, n / 8, H, p1, S1, stride1, m1, bias1,
                  p2, S2, stride2, m2, bias2, out);
}

}  // namespace nvllm
