// gemm_tcgen05.cu — bf16 tensor-core GEMM for sm_100a: tcgen05.mma with the accumulator in
// TMEM, operands staged by TMA (cp.async.bulk.tensor, 128-byte swizzle) through an mbarrier
// ring, warp-specialised (1 TMA warp, 1 MMA warp, 4 epilogue warps).
//
// Computes out[s][m][n] = alpha * sum_{k in split s} A[m,k] * W[n,k]   (fp32 split-K partials)
// for activations A [M,K] and weights W [N,K], both K-contiguous bf16 — the reference's
// MatMul (tensor.go:62-88) with the weight held [out,in].
//
// Orientation ("swap-AB"): the WEIGHT tile is the UMMA M operand (128 weight rows per CTA)
// and the token tile is the UMMA N operand (BT = 32..256 tokens).  The accumulator in TMEM is
// therefore D[n][m]: TMEM lane = weight row n, column = token m.  Consequences:
//   * skinny batched-decode GEMMs (M = 16..128 tokens) still issue full 128-row MMAs and
//     stream each weight byte exactly once: the kernel is HBM-bound on W;
//   * in the epilogue each warp owns 32 consecutive n for a given token, so every global
//     store instruction writes 128 contiguous bytes of out[m][n..n+31].
#include <cuda.h>

#include <map>
#include <mutex>
#include <tuple>

#include "common.cuh"
#include "kernels.h"

namespace nvllm {

constexpr int TC_BN = 128;  // weight rows per CTA (UMMA M)
constexpr int TC_BK = 64;   // k elements per stage = one 128-byte swizzle row of bf16
constexpr int TC_THREADS = 192;

// ------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const void* map, uint32_t bar, int c0,
                                            int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d_hint(uint32_t dst, const void* map, uint32_t bar,
                                                 int c0, int c1, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar)
               : "memory");
}
// D[tmem] (+)= A[smem desc] * B[smem desc]^T, bf16 inputs, fp32 accumulate
__device__ __forceinline__ void tc_mma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc,
                                            uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// K-major, 128-byte swizzle shared-memory matrix descriptor (cute/arch/mma_sm100_desc.hpp
// layout: start>>4 [0,14), LBO>>4 [16,30), SBO>>4 [32,46), version=1 [46,48), layout [61,64)).
// Rows are 128 B apart, 8-row swizzle atoms 1024 B apart (SBO); LBO is unused for swizzled
// K-major operands (1).
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t addr) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;  // SWIZZLE_128B
  return d;
}
// 32 lanes x 32 consecutive columns of fp32 -> 32 registers per thread
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
        "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// ------------------------------------------------------------------ kernel
// Persistent: grid = min(units, #SM) CTAs, one per SM (the operand ring takes ~192 KB of
// shared memory).  A unit is (token tile, weight tile, k split); CTA c runs units c, c + grid,
// ... with the TMA ring never draining between units and two TMEM accumulators so the
// epilogue of unit i overlaps the MMAs of unit i + 1.
//
// Programmatic dependent launch: the weight tiles of the first STAGES k-blocks do not depend
// on the previous kernel, so the producer issues them BEFORE griddepcontrol.wait — the HBM
// stream of this GEMM starts while the elementwise kernel in front of it is still running.
// Only the activation tiles (and every global store) wait for the dependency.
struct Tc5Sched {
  int units, n_tiles, S, kbps, nkb_total;
};
__device__ __forceinline__ void tc5_unit(const Tc5Sched& sc, int u, int BT, int& n0, int& m0,
                                         int& kb0, int& nkb, int& split) {
  split = u % sc.S;
  const int r = u / sc.S;
  n0 = (r % sc.n_tiles) * TC_BN;
  m0 = (r / sc.n_tiles) * BT;
  kb0 = split * sc.kbps;
  nkb = min(sc.kbps, sc.nkb_total - kb0);
}

template <int BT, int STAGES>
__global__ void __launch_bounds__(TC_THREADS, 1)
gemm_tc5_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmA,
                int M, int N, Tc5Sched sc, float* __restrict__ out, int ldo, size_t split_stride,
                float alpha, int w_reused) {
  constexpr uint32_t W_BYTES = TC_BN * TC_BK * 2;
  constexpr uint32_t A_BYTES = BT * TC_BK * 2;
  constexpr uint32_t STAGE_BYTES = W_BYTES + A_BYTES;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t tiles = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B needs 1024 B
  const uint32_t bars = tiles + STAGES * STAGE_BYTES;
  // full[s] at bars + 8 s, empty[s] at bars + 8 (STAGES + s), then tmem_full[2], tmem_empty[2]
  const uint32_t bar_tfull = bars + 16 * STAGES;
  const uint32_t bar_tempty = bar_tfull + 16;
  const uint32_t tmem_slot = bar_tempty + 16;
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(
      smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_launch_dependents();

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < STAGES; s++) {
      mbar_init(bars + 8 * s, 1);
      mbar_init(bars + 8 * (STAGES + s), 1);
    }
    for (int b = 0; b < 2; b++) {
      mbar_init(bar_tfull + 8 * b, 1);
      mbar_init(bar_tempty + 8 * b, 4);  // one arrival per epilogue warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmW) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot),
                 "r"((uint32_t)(2 * BT))
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  const int grid = gridDim.x;

  if (warp == 0) {
    if (lane == 0) {
      // weights read once per step must not displace activations in L2; weights re-read by
      // other token tiles (prefill) stay
      uint64_t pol_w, pol_a;
      if (w_reused)
        asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_w));
      else
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_w));
      asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_a));
      int n0, m0, kb0, nkb, split;
      // (1) weight tiles of the first STAGES k-blocks, before the dependency resolves
      int pre = 0;
      {
        int u = blockIdx.x, kb = 0;
        while (pre < STAGES && u < sc.units) {
          tc5_unit(sc, u, BT, n0, m0, kb0, nkb, split);
          const uint32_t full = bars + 8 * pre;
          mbar_expect_tx(full, STAGE_BYTES);
          tma_load_2d_hint(tiles + pre * STAGE_BYTES, &tmW, full, (kb0 + kb) * TC_BK, n0, pol_w);
          pre++;
          if (++kb == nkb) {
            kb = 0;
            u += grid;
          }
        }
      }
      pdl_wait();
      // (2) their activation tiles, then (3) the steady state
      int u = blockIdx.x, kb = 0, i = 0;
      while (u < sc.units) {
        tc5_unit(sc, u, BT, n0, m0, kb0, nkb, split);
        for (; kb < nkb; kb++, i++) {
          const int s = i % STAGES;
          const uint32_t full = bars + 8 * s;
          const uint32_t dst = tiles + s * STAGE_BYTES;
          const int k = (kb0 + kb) * TC_BK;
          if (i >= pre) {
            mbar_wait(bars + 8 * (STAGES + s), ((i / STAGES) & 1) ^ 1);
            mbar_expect_tx(full, STAGE_BYTES);
            tma_load_2d_hint(dst, &tmW, full, k, n0, pol_w);
          }
          tma_load_2d_hint(dst + W_BYTES, &tmA, full, k, m0, pol_a);
        }
        kb = 0;
        u += grid;
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // instruction descriptor (cute/arch/mma_sm100_desc.hpp): fp32 accumulate, bf16 x bf16,
      // both K-major, N = BT, M = 128
      constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BT >> 3) << 17) |
                                 ((uint32_t)(TC_BN >> 4) << 24);
      int n0, m0, kb0, nkb, split;
      int i = 0, lu = 0;
      for (int u = blockIdx.x; u < sc.units; u += grid, lu++) {
        tc5_unit(sc, u, BT, n0, m0, kb0, nkb, split);
        const int buf = lu & 1;
        mbar_wait(bar_tempty + 8 * buf, ((lu >> 1) & 1) ^ 1);  // epilogue drained this buffer
        tc_fence_after();
        const uint32_t tacc = tmem_base + (uint32_t)(buf * BT);
        for (int kb = 0; kb < nkb; kb++, i++) {
          const int s = i % STAGES;
          mbar_wait(bars + 8 * s, (i / STAGES) & 1);
          tc_fence_after();
          const uint32_t wt = tiles + s * STAGE_BYTES;
          const uint64_t wdesc = make_smem_desc(wt);
          const uint64_t adesc = make_smem_desc(wt + W_BYTES);
#pragma unroll
          for (int kk = 0; kk < TC_BK / 16; kk++) {
            // advancing 16 bf16 = 32 bytes inside the swizzle row: +2 in the >>4 address field
            tc_mma_bf16(tacc, wdesc + 2 * kk, adesc + 2 * kk, idesc, (kb | kk) != 0);
          }
          tc_commit(bars + 8 * (STAGES + s));  // frees the stage when these MMAs retire
        }
        tc_commit(bar_tfull + 8 * buf);
      }
    }
  } else {
    // epilogue: warp w may touch TMEM lanes [32 (w % 4), +32)
    const int q = warp & 3;
    pdl_wait();  // global stores below must follow the previous kernel's completion
    int n0, m0, kb0, nkb, split;
    int lu = 0;
    for (int u = blockIdx.x; u < sc.units; u += grid, lu++) {
      tc5_unit(sc, u, BT, n0, m0, kb0, nkb, split);
      const int buf = lu & 1;
      mbar_wait(bar_tfull + 8 * buf, (lu >> 1) & 1);
      tc_fence_after();
      const int n = n0 + q * 32 + lane;
      float* obase = out + (size_t)split * split_stride + n;
      const bool n_ok = n < N;
#pragma unroll 1
      for (int c0 = 0; c0 < BT; c0 += 32) {
        if (m0 + c0 >= M) break;
        uint32_t r[32];
        tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BT + c0), r);
        if (n_ok) {
#pragma unroll
          for (int c = 0; c < 32; c++) {
            const int m = m0 + c0 + c;
            if (m < M) obase[(size_t)m * ldo] = alpha * __uint_as_float(r[c]);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_tempty + 8 * buf);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                 "r"((uint32_t)(2 * BT))
                 : "memory");
  }
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (EncodeTiledFn)p;
  }();
  return fn;
}

int make_tma_2d_bf16(TmaDesc* out, const void* base, uint64_t rows, uint64_t cols,
                     uint64_t row_stride_elems, uint32_t box_rows, std::string* err) {
  static_assert(sizeof(CUtensorMap) == sizeof(TmaDesc), "CUtensorMap is 128 bytes");
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    if (err) *err = "cuTensorMapEncodeTiled is not available from this driver";
    return -1;
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {row_stride_elems * 2};
  cuuint32_t box[2] = {(cuuint32_t)TC_BK, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(reinterpret_cast<CUtensorMap*>(out), CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                  const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    if (err) *err = "cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r);
    return -1;
  }
  return 0;
}

// tensor maps are pure functions of (pointer, shape, box): memoise them
static const TmaDesc* cached_map(const void* base, int rows, int cols, int box_rows) {
  static std::mutex mu;
  static std::map<std::tuple<const void*, int, int, int>, TmaDesc> cache;
  std::lock_guard<std::mutex> g(mu);
  auto key = std::make_tuple(base, rows, cols, box_rows);
  auto it = cache.find(key);
  if (it != cache.end()) return &it->second;
  if (cache.size() > 65536) cache.clear();
  TmaDesc d;
  if (make_tma_2d_bf16(&d, base, rows, cols, cols, box_rows, nullptr)) return nullptr;
  return &cache.emplace(key, d).first->second;
}

bool tc5_supported(int M, int N, int K) { return M >= 1 && N >= 1 && K >= 8 && (K % 8) == 0; }

static int pick_bt(int M) { return M <= 32 ? 32 : (M <= 64 ? 64 : (M <= 128 ? 128 : 256)); }
int tc5_box_rows_act(int M) { return pick_bt(M); }

// Split-K factor minimising the makespan of the persistent schedule: rounds of units over the
// SMs x (k-blocks per unit + a fixed per-unit cost for the accumulator hand-off / epilogue),
// plus what the consumer pays to re-read S partial slices.
int tc5_splits(int M, int N, int K, int sm_count) {
  const int bt = pick_bt(M);
  const long tiles = (long)ceil_div(N, TC_BN) * ceil_div(M, bt);
  const int nkb = ceil_div(K, TC_BK);
  if (tiles >= 2L * sm_count) return 1;
  int best = 1;
  double best_cost = 1e30;
  for (int s = 1; s <= 16 && s <= nkb; s++) {
    const int kbps = ceil_div(nkb, s);
    if (s > 1 && kbps < 2) break;
    const int s_eff = ceil_div(nkb, kbps);
    if (s_eff != s) continue;
    const long units = tiles * s;
    const long rounds = (units + sm_count - 1) / sm_count;
    const double cost = (double)rounds * (kbps + 1.5) + 0.35 * s;
    if (cost < best_cost - 1e-9) {
      best_cost = cost;
      best = s;
    }
  }
  return best;
}

static bool g_pdl = true;
void set_pdl_enabled(bool on) { g_pdl = on; }
bool pdl_enabled() { return g_pdl; }

template <int BT, int STAGES>
static cudaError_t launch_bt(cudaStream_t st, const Tc5Args& a, const TmaDesc* tw,
                             const TmaDesc* ta, int S) {
  constexpr size_t smem = (size_t)STAGES * (TC_BN + BT) * TC_BK * 2 + 1024 + 16 * STAGES + 64;
  static_assert(smem <= 232448, "shared memory budget");
  static uint64_t attr_mask = 0;  // per device
  static int sm_counts[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!(attr_mask >> dev & 1)) {
    cudaError_t e = cudaFuncSetAttribute(gemm_tc5_kernel<BT, STAGES>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    cudaDeviceGetAttribute(&sm_counts[dev], cudaDevAttrMultiProcessorCount, dev);
    attr_mask |= 1ull << dev;
  }
  Tc5Sched sc;
  sc.nkb_total = ceil_div(a.K, TC_BK);
  sc.kbps = ceil_div(sc.nkb_total, S);
  sc.S = ceil_div(sc.nkb_total, sc.kbps);
  sc.n_tiles = ceil_div(a.N, TC_BN);
  const int m_tiles = ceil_div(a.M, BT);
  sc.units = sc.n_tiles * m_tiles * sc.S;
  const int grid = min(sc.units, max(1, sm_counts[dev]));
  return launch_k(gemm_tc5_kernel<BT, STAGES>, dim3(grid), dim3(TC_THREADS), smem, st,
                  *reinterpret_cast<const CUtensorMap*>(tw),
                  *reinterpret_cast<const CUtensorMap*>(ta), a.M, a.N, sc, a.out, a.ldo,
                  a.split_stride, a.alpha, m_tiles > 1 ? 1 : 0);
}

// a.S is the requested number of splits; the number actually written is returned in *S_out
cudaError_t launch_gemm_tc5(cudaStream_t st, const Tc5Args& a, int* S_out) {
  if (!tc5_supported(a.M, a.N, a.K)) return cudaErrorInvalidValue;
  const int bt = pick_bt(a.M);
  const TmaDesc* tw = cached_map(a.W, a.N, a.K, TC_BN);
  const TmaDesc* ta = cached_map(a.A, a.M, a.K, bt);
  if (!tw || !ta) return cudaErrorInvalidValue;
  const int nkb = ceil_div(a.K, TC_BK);
  int S = max(1, min(a.S, nkb));
  const int kbps = ceil_div(nkb, S);
  S = ceil_div(nkb, kbps);
  if (S_out) *S_out = S;
  if (a.out_capacity && (size_t)S * a.split_stride > a.out_capacity) return cudaErrorInvalidValue;
  switch (bt) {
    case 32: return launch_bt<32, 10>(st, a, tw, ta, S);
    case 64: return launch_bt<64, 8>(st, a, tw, ta, S);
    case 128: return launch_bt<128, 6>(st, a, tw, ta, S);
    default: return launch_bt<256, 4>(st, a, tw, ta, S);
  }
}

}  // namespace nvllm
