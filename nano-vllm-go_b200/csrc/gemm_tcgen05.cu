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
#include <stdlib.h>

#include <map>
#include <mutex>
#include <tuple>

#include "common.cuh"
#include "tma.cuh"
#include "tc5.cuh"
#include "kernels.h"

namespace nvllm {

constexpr int TC_BN = 128;  // weight rows per CTA (UMMA M)
constexpr int TC_BK = 64;   // k elements per stage = one 128-byte swizzle row of bf16
constexpr int TC_THREADS = 192;

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
  // grouped (mixture-of-experts) mode, g_expert != nullptr: token tile t covers rows
  // [g_row0[t], g_row0[t] + g_rows[t]) of the expert-sorted activations and multiplies expert
  // g_expert[t]'s [g_N, K] weight matrix (experts stacked row-wise in W); the tile list lives
  // on the device (moe_sort_kernel), so units = *g_ntiles * n_tiles is read after the wait
  const int* g_expert;
  const int* g_row0;
  const int* g_rows;
  const int* g_ntiles;
  int g_N;
  // decode-sized grouped GEMMs: raw weight base / bytes per weight row / expert count for the
  // speculative L2 prefetch before the dependency resolves (null = off)
  const uint8_t* g_wbase;
  int g_row_bytes, g_E;
};
// the RMSNorm-fold arguments of Tc5Args (kernels.h), as the kernel takes them
struct Tc5Fuse {
  const float* rs_ssq;
  int rs_parts, rs_ld;
  float rs_inv_h, rs_eps;
  float* x;
  bf16* xb;
  float* ssq;
  int ssq_ld;
  const float* bias;
  float m;
  const float* g;  // weight of the norm that reads x next: xb = bf16(x * g); null = ones
  int cl;          // > 1: the S k-splits of an output tile are one thread-block cluster of `cl` CTAs
};
// A second GEMM chained behind the first inside ONE launch (the FFN pair of a decode step:
// W1 + activation, then W2): every CTA runs its units of problem 1, then its units of problem 2.
// Problem 2 reads what problem 1's epilogues wrote, so its activation tiles wait for
// ctr[0] == units of problem 1 — but its WEIGHT tiles are requested as soon as ring slots free
// up, i.e. under problem 1's epilogues and the wait, and there is no launch / drain / release
// gap between the two GEMMs.  ctr[1] counts the CTAs past the wait; the last one re-arms both.
struct Tc5Chain {
  Tc5Sched sc;  // units == 0: no chain
  int N;
  float* out;
  int ldo;
  size_t split_stride;
  float alpha;
  unsigned* ctr;
  int n_waiters;  // CTAs owning at least one unit of problem 2
};
__device__ __forceinline__ void named_bar(int id, int n) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// thread-block cluster primitives (the fused-residual split-K epilogue)
__device__ __forceinline__ void cluster_arrive_relaxed() {
  asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
}
__device__ __forceinline__ void cluster_arrive_release() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
}
__device__ __forceinline__ void cluster_wait_acquire() {
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_map(uint32_t smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
  return r;
}
// bulk copy from this CTA's shared memory into a peer's (both 16-byte aligned, bytes % 16 == 0);
// completion is signalled as `bytes` of transaction count on the PEER's mbarrier
__device__ __forceinline__ void dsmem_bulk_copy(uint32_t dst_cluster, uint32_t src_cta, uint32_t bytes,
                                                uint32_t bar_cluster) {
  asm volatile(
      "cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      ::"r"(dst_cluster), "r"(src_cta), "r"(bytes), "r"(bar_cluster)
      : "memory");
}
__device__ __forceinline__ void st_shared_f32(uint32_t addr, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ float4 ld_shared_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
// v[c] of every lane -> lane l gets sum over lanes of v[l]  (31 shuffles instead of 32 x 5)
__device__ __forceinline__ float warp_transpose_sum(float (&v)[32], int lane) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < off; i++) {
      const float send = up ? v[i] : v[i + off];
      const float keep = up ? v[i + off] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  return v[0];
}
// rot: CTAs working on the same k range start at different k-blocks (and wrap), so that at any
// instant they pull DIFFERENT activation tiles out of L2 — otherwise every CTA of a skinny
// GEMM requests the same 8 KB A tile at the same time and the few L2 slices holding it pace
// the whole kernel (measured: ~390 ns per k-block with W already in shared memory).
struct Tc5Unit {
  int n0;     // first output column (weight row inside the expert)
  int wrow;   // first weight row in the stacked weight tensor (TMA coordinate)
  int m0;     // first token row
  int m_end;  // one past the last valid token row of this tile
  int kb0, nkb, split, rot;
};
__device__ __forceinline__ Tc5Unit tc5_unit(const Tc5Sched& sc, int u, int BT, int M) {
  Tc5Unit t;
  t.split = u % sc.S;
  const int r = u / sc.S;
  const int nt = r % sc.n_tiles, mt = r / sc.n_tiles;
  t.n0 = nt * TC_BN;
  if (sc.g_expert) {
    t.m0 = sc.g_row0[mt];
    t.m_end = t.m0 + sc.g_rows[mt];
    t.wrow = sc.g_expert[mt] * sc.g_N + t.n0;
  } else {
    t.m0 = mt * BT;
    t.m_end = M;
    t.wrow = t.n0;
  }
  t.kb0 = t.split * sc.kbps;
  t.nkb = min(sc.kbps, sc.nkb_total - t.kb0);
  t.rot = (nt + 5 * mt) % t.nkb;
  return t;
}
__device__ __forceinline__ int tc5_kb(int kb0, int kb, int rot, int nkb) {
  const int k = kb + rot;
  return kb0 + (k >= nkb ? k - nkb : k);
}

template <int BT, int STAGES>
__global__ void __launch_bounds__(TC_THREADS, 1)
gemm_tc5_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmA,
                const __grid_constant__ CUtensorMap tmW2, const __grid_constant__ CUtensorMap tmA2,
                int M, int N_a, Tc5Sched sc, float* __restrict__ out_a, int ldo_a, size_t split_stride_a,
                float alpha_a, int w_reused, int epi_a, bf16* __restrict__ epi_out, int epi_ld,
                const float* __restrict__ epi_bias, int prefetch, const uint8_t* __restrict__ pf_ptr,
                unsigned long long pf_bytes, Tc5Fuse fz, Tc5Chain ch, Trace tr) {
  const unsigned long long t_entry = trace_on(tr) ? gtime() : 0ull;
  unsigned long long t_wait = 0;
  __shared__ unsigned long long tr_mma, tr_epi0, tr_epi1, tr_x[6], tr_c[4];
  __shared__ __align__(16) uint8_t epi_stage[4][32 * 48];
  __shared__ float rstd_s[2][BT];       // row scales of the unit in flight (double-buffered)
  __shared__ float ssq_s[2][4][32];     // RESID: per-warp row sums of squares of a 32-token chunk
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
  // cluster mode only (the launcher adds the bytes):  xch [source rank][per][128] fp32, the rows
  // this CTA owns as the peers' bulk copies deliver them — never part of the TMA ring, a peer may
  // be ahead of this CTA's main loop;  stg [BT][128], this CTA's partial tile (the copies'
  // source): written after this CTA's own last MMA has retired, so it reuses the ring's first
  // stages;  bar_x counts the delivered bytes
  const uint32_t bar_x = tmem_slot + 8;
  const uint32_t xch = (tmem_slot + 16 + 127u) & ~127u;
  const uint32_t stg = tiles;
  static_assert((size_t)BT * TC_BN * 4 <= (size_t)STAGES * STAGE_BYTES, "staging tile fits the ring");
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(
      smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool grouped = sc.g_expert != nullptr;
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
    if (fz.cl > 1) mbar_init(bar_x, 1);
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
  // cluster phase A ("this CTA runs": its shared memory may be written by peers), waited for by
  // each warp right before its phase-B arrival; phase B = "my slices are in the owners' buffers"
  if (fz.cl > 1) cluster_arrive_relaxed();

  // The producer and MMA warps run their loops with ALL 32 lanes (warp-uniform control flow)
  // and elect one lane only around the issue instructions: TMA and tcgen05 instructions take
  // uniform-register operands, and inside a divergent `if (lane == 0)` region the compiler
  // wraps every one of them in an elect/branch loop plus R2UR moves (~120 cycles per MMA
  // measured on B200 — enough to make a skinny GEMM issue-bound instead of HBM-bound).
  if (warp == 0) {
    // weights read once per step must not displace activations in L2; weights re-read by
    // other token tiles (prefill) stay
    uint64_t pol_w, pol_a;
    if (w_reused)
      asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_w));
    else
      asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_w));
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_a));
    // (1) weight tiles of the first STAGES k-blocks, before the dependency resolves (dense
    // mode only: a grouped GEMM learns which expert a tile multiplies from the previous kernel)
    int pre = 0;
    {
      int u = blockIdx.x, kb = 0;
      while ((prefetch & 1) && !grouped && pre < STAGES && u < sc.units) {
        const Tc5Unit t = tc5_unit(sc, u, BT, M);
        const uint32_t full = bars + 8 * pre;
        if (elect_one()) {
          mbar_expect_tx(full, (prefetch & 2) ? W_BYTES : STAGE_BYTES);
          tma_load_2d_hint(tiles + pre * STAGE_BYTES, &tmW, full,
                           tc5_kb(t.kb0, kb, t.rot, t.nkb) * TC_BK, t.wrow, pol_w);
        }
        __syncwarp();
        pre++;
        if (++kb == t.nkb) {
          kb = 0;
          u += grid;
        }
      }
    }
    // (1a') grouped mode cannot know its experts before the sort kernel has run, but a decode
    // batch routes a handful of rows to EVERY expert: one token tile per expert, in expert
    // order, so unit u multiplies expert u / n_tiles.  Pull that guess into L2 now (a wrong
    // guess only costs bandwidth); the real TMA loads after the wait then hit L2.
    if (grouped && sc.g_wbase && (prefetch & 1)) {
      for (int u = blockIdx.x; u < sc.g_E * sc.n_tiles; u += grid) {
        const int nt = u % sc.n_tiles, e = u / sc.n_tiles;
        const int rows = min(TC_BN, sc.g_N - nt * TC_BN);
        const uint8_t* base = sc.g_wbase + ((size_t)e * sc.g_N + (size_t)nt * TC_BN) * sc.g_row_bytes;
        const uint32_t bytes = (uint32_t)rows * sc.g_row_bytes;
        constexpr uint32_t PIECE = 8192;
        for (uint32_t o = lane * PIECE; o < bytes; o += 32 * PIECE) {
          const uint32_t sz = (bytes - o < PIECE ? bytes - o : PIECE) & ~15u;
          if (sz)
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(base + o), "r"(sz)
                         : "memory");
        }
      }
    }
    // (1b) L2 prefetch of the weights a LATER kernel streams (immutable, so legal before the
    // wait): this kernel's own stream leaves HBM idle while it waits for its dependency and
    // during its epilogue; the next big GEMM then reads L2 instead of HBM.
    if (pf_bytes) {
      const unsigned long long per = ((pf_bytes + gridDim.x - 1) / gridDim.x + 127) & ~127ull;
      const unsigned long long lo = per * blockIdx.x;
      const unsigned long long hi = lo + per < pf_bytes ? lo + per : pf_bytes;
      constexpr unsigned long long PIECE = 16384;
      for (unsigned long long o = lo + (unsigned long long)lane * PIECE; o < hi; o += 32 * PIECE) {
        const uint32_t sz = (uint32_t)((hi - o < PIECE ? hi - o : PIECE) & ~15ull);
        if (sz)
          asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(pf_ptr + o), "r"(sz)
                       : "memory");
      }
    }
    pdl_wait();
    if (trace_on(tr)) t_wait = gtime();
    // (2) their activation tiles, then (3) the steady state
    int i = 0;
    int ahead_lo = 0, ahead_hi = pre;  // k-block range [lo, hi) whose weight tile is already in flight
    for (int p = 0; p < 2; p++) {
      const Tc5Sched& scp = p ? ch.sc : sc;
      const CUtensorMap* mW = p ? &tmW2 : &tmW;
      const CUtensorMap* mA = p ? &tmA2 : &tmA;
      const int units = p ? ch.sc.units : (grouped ? *sc.g_ntiles * sc.n_tiles : sc.units);
      if (p == 1 && (int)blockIdx.x < units) {
        // chained GEMM: weight tiles of the first k-blocks now, activations after problem 1 is done
        const Tc5Unit t = tc5_unit(scp, blockIdx.x, BT, M);
        const int na = min(STAGES, t.nkb);
        for (int kb = 0; kb < na; kb++) {
          const int ii = i + kb, s = ii % STAGES;
          mbar_wait(bars + 8 * (STAGES + s), ((ii / STAGES) & 1) ^ 1);
          if (elect_one()) {
            mbar_expect_tx(bars + 8 * s, STAGE_BYTES);
            tma_load_2d_hint(tiles + s * STAGE_BYTES, mW, bars + 8 * s,
                             tc5_kb(t.kb0, kb, t.rot, t.nkb) * TC_BK, t.wrow, pol_w);
          }
          __syncwarp();
        }
        ahead_lo = i;
        ahead_hi = i + na;
        if (elect_one()) {
          while (ld_acquire_u32(ch.ctr) < (unsigned)sc.units) {
          }
          if (atomicAdd(ch.ctr + 1, 1u) == (unsigned)ch.n_waiters - 1) {  // everyone is past the wait
            ch.ctr[0] = 0;
            ch.ctr[1] = 0;
          }
        }
        __syncwarp();
        // the activations were written through the generic proxy by other CTAs; TMA reads them
        asm volatile("fence.proxy.async;" ::: "memory");
      }
      for (int u = blockIdx.x; u < units; u += grid) {
        const Tc5Unit t = tc5_unit(scp, u, BT, M);
        for (int kb = 0; kb < t.nkb; kb++, i++) {
          const int s = i % STAGES;
          const uint32_t full = bars + 8 * s;
          const uint32_t dst = tiles + s * STAGE_BYTES;
          const int k = tc5_kb(t.kb0, kb, t.rot, t.nkb) * TC_BK;
          const bool w_ahead = i >= ahead_lo && i < ahead_hi;
          if (!w_ahead) mbar_wait(bars + 8 * (STAGES + s), ((i / STAGES) & 1) ^ 1);
          if (elect_one()) {
            if (!w_ahead) {
              mbar_expect_tx(full, (prefetch & 2) ? W_BYTES : STAGE_BYTES);
              tma_load_2d_hint(dst, mW, full, k, t.wrow, pol_w);
            }
            if (!(prefetch & 2)) tma_load_2d_hint(dst + W_BYTES, mA, full, k, t.m0, pol_a);
          }
          __syncwarp();
        }
      }
    }
    if (fz.cl > 1) {
      cluster_wait_acquire();
      cluster_arrive_release();
      cluster_wait_acquire();
    }
  } else if (warp == 1) {
    // instruction descriptor (cute/arch/mma_sm100_desc.hpp): fp32 accumulate, bf16 x bf16,
    // both K-major, N = BT, M = 128
    constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BT >> 3) << 17) |
                               ((uint32_t)(TC_BN >> 4) << 24);
    if (grouped) pdl_wait();  // the tile list is the previous kernel's output
    int i = 0, lu = 0;
    for (int p = 0; p < 2; p++) {
    const Tc5Sched& scp = p ? ch.sc : sc;
    const int units = p ? ch.sc.units : (grouped ? *sc.g_ntiles * sc.n_tiles : sc.units);
    for (int u = blockIdx.x; u < units; u += grid, lu++) {
      const int nkb = tc5_unit(scp, u, BT, M).nkb;
      const int buf = lu & 1;
      mbar_wait(bar_tempty + 8 * buf, ((lu >> 1) & 1) ^ 1);  // epilogue drained this buffer
      tc_fence_after();
      const uint32_t tacc = tmem_base + (uint32_t)(buf * BT);
      for (int kb = 0; kb < nkb; kb++, i++) {
        const int s = i % STAGES;
        mbar_wait(bars + 8 * s, (i / STAGES) & 1);
        tc_fence_after();
        if (tr.buf && lane == 0) {
          if (i == 0) tr_x[0] = gtime();
          tr_x[1] = gtime();
        }
        const uint32_t wt = tiles + s * STAGE_BYTES;
        const uint64_t wdesc = make_smem_desc(wt);
        const uint64_t adesc = make_smem_desc(wt + W_BYTES);
        if (elect_one()) {
#pragma unroll
          for (int kk = 0; kk < TC_BK / 16; kk++) {
            // advancing 16 bf16 = 32 bytes inside the swizzle row: +2 in the >>4 address field
            tc_mma_bf16(tacc, wdesc + 2 * kk, adesc + 2 * kk, idesc, (kb | kk) != 0);
          }
          tc_commit(bars + 8 * (STAGES + s));  // frees the stage when these MMAs retire
          if (kb == nkb - 1) tc_commit(bar_tfull + 8 * buf);
        }
        __syncwarp();
      }
      if (tr.buf && lane == 0) tr_mma = gtime();
    }
    }
    if (fz.cl > 1) {
      cluster_wait_acquire();
      cluster_arrive_release();
      cluster_wait_acquire();
    }
  } else {
    // epilogue: warp w may touch TMEM lanes [32 (w % 4), +32)
    const int q = warp & 3;
    pdl_wait();  // global stores below must follow the previous kernel's completion
    int lu = 0;
    for (int p = 0; p < 2; p++) {
    const Tc5Sched& scp = p ? ch.sc : sc;
    const int units = p ? ch.sc.units : (grouped ? *sc.g_ntiles * sc.n_tiles : sc.units);
    // problem 2 of a chain always writes plain fp32 split-K partials
    const int N = p ? ch.N : N_a, ldo = p ? ch.ldo : ldo_a, epi = p ? (int)TC5_EPI_PARTIAL : epi_a;
    float* const out = p ? ch.out : out_a;
    const size_t split_stride = p ? ch.split_stride : split_stride_a;
    const float alpha = p ? ch.alpha : alpha_a;
    for (int u = blockIdx.x; u < units; u += grid, lu++) {
      const Tc5Unit t = tc5_unit(scp, u, BT, M);
      const int n0 = t.n0, m0 = t.m0, split = t.split;
      const int M = t.m_end;  // rows past the tile's extent belong to the next expert / are padding
      const int buf = lu & 1;
      const int et = threadIdx.x - 64;  // 0..127 over the four epilogue warps
      const float* rs = nullptr;
      if (fz.rs_ssq && p == 0) {
        // folded RMSNorm: 1 / sqrt(mean(x^2) + eps) per token row (tensor.go:211-224), from the
        // per-column-tile sums the producing GEMM (or the embedding kernel) left; computed while
        // the MMAs of this unit run
        float* w = rstd_s[lu & 1];
        for (int c = et; c < BT; c += 128) {
          const int mm = m0 + c;
          float ssum = 0.f;
          if (mm < M) {
#pragma unroll 8
            for (int p = 0; p < fz.rs_parts; p++) ssum += __ldcg(fz.rs_ssq + (size_t)p * fz.rs_ld + mm);
          }
          w[c] = 1.0f / sqrtf(ssum * fz.rs_inv_h + fz.rs_eps);
        }
        named_bar(1, 128);
        rs = w;
      }
      // cluster split-K fused residual: what this CTA, as the owner of token rows [r0, r1) of the
      // tile, needs from global memory (its x rows, bias, norm weight) depends on no peer and on
      // no MMA: fetched NOW, under the main loop — by the time the epilogue runs, the next
      // kernel's weight prefetch has the memory system queued up and a load takes microseconds
      constexpr int OWN = BT > 64 ? 1 : BT / 16;  // rows one warp finishes: ceil(BT / S / 4) with S >= 4
      const bool cl_resid = p == 0 && epi == TC5_EPI_RESID && scp.S > 1;
      const int valid = min(BT, M - m0);
      const int per = (valid + scp.S - 1) / scp.S;
      const int r0 = split * per, r1 = min(valid, r0 + per);
      const int ncol = n0 + lane * 4;
      const bool c_ok = ncol + 3 < N;  // N % 4 == 0 is checked by the launcher
      float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f), g4 = make_float4(1.f, 1.f, 1.f, 1.f);
      float4 xv[OWN];
      if (cl_resid) {
        if (fz.bias && c_ok) b4 = *reinterpret_cast<const float4*>(fz.bias + ncol);
        if (fz.g && c_ok) g4 = *reinterpret_cast<const float4*>(fz.g + ncol);
#pragma unroll
        for (int j = 0; j < OWN; j++) {
          const int row = r0 + q + 4 * j;
          xv[j] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (row < r1 && c_ok)
            xv[j] = __ldcg(reinterpret_cast<const float4*>(fz.x + (size_t)(m0 + row) * N + ncol));
        }
      }
      mbar_wait(bar_tfull + 8 * buf, (lu >> 1) & 1);
      tc_fence_after();
      if (tr.buf && warp == 2 && lane == 0) tr_epi0 = gtime();
      const int n = n0 + q * 32 + lane;
      const bool n_ok = n < N;
      if (epi == TC5_EPI_PARTIAL) {
        float* obase = out + (size_t)split * split_stride + n;
#pragma unroll 1
        for (int c0 = 0; c0 < BT; c0 += 32) {
          if (m0 + c0 >= M) break;
          uint32_t r[32];
          tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BT + c0), r);
          if (tr.buf && warp == 2 && lane == 0 && c0 == 0) tr_x[2] = gtime();
          if (rs) {
#pragma unroll
            for (int c = 0; c < 32; c++) r[c] = __float_as_uint(__uint_as_float(r[c]) * rs[c0 + c]);
          }
          if (n_ok) {
            float* op = obase + (size_t)(m0 + c0) * ldo;
            if (m0 + c0 + 32 <= M) {  // full chunk: straight-line stores, no per-row branch
#pragma unroll
              for (int c = 0; c < 32; c++) op[(size_t)c * ldo] = alpha * __uint_as_float(r[c]);
            } else {
#pragma unroll
              for (int c = 0; c < 32; c++)
                if (m0 + c0 + c < M) op[(size_t)c * ldo] = alpha * __uint_as_float(r[c]);
            }
          }
          if (tr.buf && warp == 2 && lane == 0 && c0 == 0) tr_x[3] = gtime();
        }
      } else if (epi == TC5_EPI_SWIGLU) {
        // rows are interleaved gate/up (weights.cu): even lane = gate j, odd lane = up j.  The
        // pair splits the 32 tokens: the even lane finishes tokens 0..15, the odd lane 16..31
        // (transformer.go:50-66: silu(gate) * up)
        // Outputs are staged per warp in shared memory ([32 tokens][16 columns] bf16, rows
        // 48 B apart: conflict-free 16-byte reads) so that each lane then writes 16 contiguous
        // bytes of one token row instead of sixteen 2-byte stores.
        const int odd = lane & 1, p = lane >> 1;
        uint8_t* stg = epi_stage[q];
        const int j0 = (n0 + q * 32) >> 1;
#pragma unroll 1
        for (int c0 = 0; c0 < BT; c0 += 32) {
          if (m0 + c0 >= M) break;
          uint32_t r[32];
          tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BT + c0), r);
          if (rs) {
#pragma unroll
            for (int c = 0; c < 32; c++) r[c] = __float_as_uint(__uint_as_float(r[c]) * rs[c0 + c]);
          }
#pragma unroll
          for (int c = 0; c < 16; c++) {
            const float mine = __uint_as_float(odd ? r[c + 16] : r[c]);
            const float send = __uint_as_float(odd ? r[c] : r[c + 16]);
            const float got = __shfl_xor_sync(0xffffffffu, send, 1);
            const float g = alpha * (odd ? got : mine), u = alpha * (odd ? mine : got);
            const float v = __fdividef(g, 1.0f + __expf(-g)) * u;
            *reinterpret_cast<bf16*>(stg + (c + (odd ? 16 : 0)) * 48 + p * 2) =
                __float2bfloat16_rn(v);
          }
          __syncwarp();
          const int m = m0 + c0 + lane;
          if (m < M) {
            bf16* orow = epi_out + (size_t)m * epi_ld + j0;
#pragma unroll
            for (int hh = 0; hh < 2; hh++)
              if (2 * (j0 + hh * 8) < N)
                *reinterpret_cast<uint4*>(orow + hh * 8) =
                    *reinterpret_cast<const uint4*>(stg + lane * 48 + hh * 16);
          }
          __syncwarp();
        }
      } else if (epi == TC5_EPI_RESID) {
        // residual update fused into the projection that feeds it (generic_model.go:359-365,
        // 383-389: x = x + m * (proj + bias)), plus what the FOLDED RMSNorm of the next
        // sub-block needs: the row rounded to bf16 and its sum of squares over this column tile
        const float b = (fz.bias && n_ok) ? fz.bias[n] : 0.f;
        const float gn = (fz.g && n_ok) ? fz.g[n] : 1.f;
        const int nt = n0 / TC_BN;
        if (scp.S == 1) {
          int chunk = 0;
#pragma unroll 1
          for (int c0 = 0; c0 < BT; c0 += 32, chunk++) {
            if (m0 + c0 >= M) break;
            uint32_t r[32];
            tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BT + c0), r);
            float sq[32];
            const size_t base = (size_t)(m0 + c0) * N + n;
            if (n_ok && m0 + c0 + 32 <= M) {
              float xv[32];
#pragma unroll
              for (int c = 0; c < 32; c++) xv[c] = fz.x[base + (size_t)c * N];
#pragma unroll
              for (int c = 0; c < 32; c++) {
                const float v = xv[c] + fz.m * (alpha * __uint_as_float(r[c]) + b);
                fz.x[base + (size_t)c * N] = v;
                fz.xb[base + (size_t)c * N] = __float2bfloat16_rn(v * gn);
                sq[c] = v * v;
              }
            } else {
#pragma unroll
              for (int c = 0; c < 32; c++) {
                sq[c] = 0.f;
                if (n_ok && m0 + c0 + c < M) {
                  const float v = fz.x[base + (size_t)c * N] + fz.m * (alpha * __uint_as_float(r[c]) + b);
                  fz.x[base + (size_t)c * N] = v;
                  fz.xb[base + (size_t)c * N] = __float2bfloat16_rn(v * gn);
                  sq[c] = v * v;
                }
              }
            }
            const float tot = warp_transpose_sum(sq, lane);  // token c0 + lane over this warp's 32 columns
            ssq_s[chunk & 1][q][lane] = tot;
            named_bar(2, 128);
            if (q == 0 && m0 + c0 + lane < M)
              fz.ssq[(size_t)nt * fz.ssq_ld + m0 + c0 + lane] =
                  ((ssq_s[chunk & 1][0][lane] + ssq_s[chunk & 1][1][lane]) + ssq_s[chunk & 1][2][lane]) +
                  ssq_s[chunk & 1][3][lane];
          }
        } else {
          // Split-K: the S CTAs of this output tile are ONE thread-block cluster (rank == split).
          // Token rows are dealt out to the ranks; every CTA stages its partial tile in its own
          // shared memory and bulk-copies each owner's rows into that owner's shared memory
          // (DSMEM, completion counted on the owner's mbarrier); each owner then sums the S slices
          // of its rows in rank order and finishes them — no global round trip, no counters, no
          // norm kernel behind this GEMM.  (Per-lane st.shared::cluster stores instead of bulk
          // copies were measured at ~14 us per 32 KB tile: the SM-to-SM path is transaction-bound.)
          // (1) the partial tile into the staging buffer, rows grouped by owner
          {
            const uint32_t col = (uint32_t)((q * 32 + lane) * 4);
#pragma unroll 1
            for (int c0 = 0; c0 < BT; c0 += 32) {
              if (c0 >= valid) break;
              uint32_t r[32];
              tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BT + c0), r);
#pragma unroll
              for (int c = 0; c < 32; c++)
                if (c0 + c < valid)  // row index == token index: owner o holds rows [o per, o per + per)
                  st_shared_f32(stg + (uint32_t)((c0 + c) * TC_BN * 4) + col, alpha * __uint_as_float(r[c]));
            }
          }
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic writes -> bulk-copy reads
          named_bar(2, 128);
          if (tr.buf && warp == 2 && lane == 0) tr_c[0] = gtime();
          cluster_wait_acquire();  // phase A: every peer is running and has initialised its bar_x
          // (2) one bulk copy per peer: its rows of my slice -> slot `split` of its xch
          if (q == 0) {
            const int o = lane;
            if (o < scp.S && o != split) {
              const int o0 = o * per, o1 = min(valid, o0 + per);
              if (o1 > o0)
                dsmem_bulk_copy(cluster_map(xch, (uint32_t)o) + (uint32_t)(split * per * TC_BN * 4),
                                stg + (uint32_t)(o0 * TC_BN * 4), (uint32_t)((o1 - o0) * TC_BN * 4),
                                cluster_map(bar_x, (uint32_t)o));
            }
            if (lane == 0) mbar_expect_tx(bar_x, (uint32_t)((scp.S - 1) * max(r1 - r0, 0) * TC_BN * 4));
          }
          mbar_wait(bar_x, 0);  // (3) the S - 1 peer slices of my rows have landed
          if (tr.buf && warp == 2 && lane == 0) tr_c[1] = gtime();
          const int nt_ = nt;
#pragma unroll
          for (int j = 0; j < OWN; j++) {
            const int row = r0 + q + 4 * j;
            if (row >= r1) break;
            float4 pv[8];  // all S shared-memory loads in flight before the first add
#pragma unroll
            for (int sp = 0; sp < 8; sp++)
              if (sp < scp.S)
                pv[sp] = ld_shared_f4(sp == split ? stg + (uint32_t)((row * TC_BN + lane * 4) * 4)
                                                  : xch + (uint32_t)(((sp * per + (row - r0)) * TC_BN + lane * 4) * 4));
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (c_ok) {
              float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
              for (int sp = 0; sp < 8; sp++)
                if (sp < scp.S) {
                  acc.x += pv[sp].x; acc.y += pv[sp].y; acc.z += pv[sp].z; acc.w += pv[sp].w;
                }
              const size_t xi = (size_t)(m0 + row) * N + ncol;
              v.x = xv[j].x + fz.m * (acc.x + b4.x);
              v.y = xv[j].y + fz.m * (acc.y + b4.y);
              v.z = xv[j].z + fz.m * (acc.z + b4.z);
              v.w = xv[j].w + fz.m * (acc.w + b4.w);
              *reinterpret_cast<float4*>(fz.x + xi) = v;
              __nv_bfloat162 lo = __floats2bfloat162_rn(v.x * g4.x, v.y * g4.y),
                             hi = __floats2bfloat162_rn(v.z * g4.z, v.w * g4.w);
              uint2 pk;
              pk.x = *reinterpret_cast<uint32_t*>(&lo);
              pk.y = *reinterpret_cast<uint32_t*>(&hi);
              *reinterpret_cast<uint2*>(fz.xb + xi) = pk;
            }
            const float tot = warp_sum((v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w));
            if (lane == 0) fz.ssq[(size_t)nt_ * fz.ssq_ld + m0 + row] = tot;
          }
          if (tr.buf && warp == 2 && lane == 0) {
            tr_c[2] = gtime();
            tr_c[3] = tr_c[2];
          }
          // phase B: once every CTA of the cluster is past its own wait, every copy OUT of this
          // CTA's staging buffer has completed too — only then may the CTA exit
          cluster_arrive_release();
          cluster_wait_acquire();
        }
      } else {
        // GELU branch: gelu_tanh(h + b1)   (transformer.go:67-78)
        const float b = (epi_bias && n_ok) ? epi_bias[n] : 0.f;
        bf16* obase = epi_out + n;
#pragma unroll 1
        for (int c0 = 0; c0 < BT; c0 += 32) {
          if (m0 + c0 >= M) break;
          uint32_t r[32];
          tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(buf * BT + c0), r);
          if (rs) {
#pragma unroll
            for (int c = 0; c < 32; c++) r[c] = __float_as_uint(__uint_as_float(r[c]) * rs[c0 + c]);
          }
          if (n_ok) {
            bf16* op = obase + (size_t)(m0 + c0) * epi_ld;
            if (m0 + c0 + 32 <= M) {
#pragma unroll
              for (int c = 0; c < 32; c++)
                op[(size_t)c * epi_ld] =
                    __float2bfloat16_rn(gelu_tanh(alpha * __uint_as_float(r[c]) + b));
            } else {
#pragma unroll
              for (int c = 0; c < 32; c++)
                if (m0 + c0 + c < M)
                  op[(size_t)c * epi_ld] =
                      __float2bfloat16_rn(gelu_tanh(alpha * __uint_as_float(r[c]) + b));
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (tr.buf && warp == 2 && lane == 0) tr_epi1 = gtime();
      if (lane == 0) mbar_arrive(bar_tempty + 8 * buf);
      if (p == 0 && ch.sc.units > 0) {  // publish this tile of problem 1 to the chained GEMM
        __threadfence();
        named_bar(3, 128);
        if (et == 0) atomicAdd(ch.ctr, 1u);
      }
    }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0 && trace_on(tr)) {
    // extra: (last MMA commit issued, epilogue saw the accumulator, epilogue stores issued)
    // as 21-bit ns offsets from the dependency release
    const unsigned long long a = (tr_mma - t_wait) & 0x1fffff, b = (tr_epi0 - t_wait) & 0x1fffff,
                             c = (tr_epi1 - t_wait) & 0x1fffff;
    trace_rec(tr, t_entry, t_wait, (a << 42) | (b << 21) | c);
    Trace t2 = tr;
    t2.kid = 9;
    t2.agg = nullptr;
    if (fz.cl > 1) {
      // cluster epilogue: staged, phase A passed, peer slices landed, rows finished (16-bit ns after the wait)
      unsigned long long pk = 0;
      for (int i = 0; i < 4; i++) pk = (pk << 16) | ((tr_c[i] - t_wait) & 0xffffull);
      trace_rec(t2, tr_x[0] - t_wait, tr_x[1] - t_wait, pk);
    } else {
      trace_rec(t2, tr_x[0] - t_wait, tr_x[1] - t_wait, ((tr_x[2] - t_wait) << 32) | (tr_x[3] - t_wait));
    }
  }
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

const TmaDesc* tma_map_cached(const void* base, int rows, int cols, int box_rows) {
  return cached_map(base, rows, cols, box_rows);
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
  // The consumer of the partials sums up to 8 slices with one batch of loads (sum_partials8); a
  // 9th slice costs it a second L2 round trip, more than the GEMM gains from the finer split
  // (cfg2 decode, W2 GEMM K = 8192: S = 9 -> 1105 us/step, capped at 8 -> 1079; 6 -> 1099 /
  // 1219 with the fused residual; 4 -> 1371).
  constexpr int smax = 8;
  for (int s = 1; s <= smax && s <= nkb; s++) {
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

// ring depths of the cluster variant (the receive buffer takes 36 / 20 KB of the ring's room)
constexpr int CL_STAGES_32 = 9, CL_STAGES_64 = 7;
constexpr size_t tc5_smem_bytes(int BT, int STAGES, bool CL) {
  return (size_t)STAGES * (TC_BN + BT) * TC_BK * 2 + 1024 + 16 * STAGES + 64 +
         (CL ? (size_t)(BT + 8) * TC_BN * 4 + 128 : 0);
}
// how many clusters of S CTAs of the fused-residual GEMM can be resident at once on this device
// (1 CTA per SM, all CTAs of a cluster on one GPC); 0 when there is no device
template <int BT, int STAGES>
static int max_clusters(int S) {
  static int cache[64][9];
  static bool known[64][9] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64 || S < 2 || S > 8) return 0;
  if (known[dev][S]) return cache[dev][S];
  const size_t smem = tc5_smem_bytes(BT, STAGES, true);
  int n = 0;
  if (cudaFuncSetAttribute(gemm_tc5_kernel<BT, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           (int)smem) == cudaSuccess) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(S * 16);
    cfg.blockDim = dim3(TC_THREADS);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)S;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (cudaOccupancyMaxActiveClusters(&n, gemm_tc5_kernel<BT, STAGES>, &cfg) != cudaSuccess) n = 0;
  }
  (void)cudaGetLastError();
  cache[dev][S] = n;
  known[dev][S] = true;
  return n;
}

// Split count of the fused-residual GEMM (TC5_EPI_RESID): 1 = enough output tiles on their own;
// 2..8 = the k-splits of a tile run as one thread-block cluster — the largest count whose clusters
// are ALL resident at once (a second wave of clusters doubles the kernel: B200 places 15 clusters
// of 8 such CTAs, one short of the 16 tiles of a 2048-wide model); 0 = not available for this shape
int tc5_resid_splits(int M, int N, int K, int sm_count) {
  const int bt = pick_bt(M);
  const long tiles = (long)ceil_div(N, TC_BN) * ceil_div(M, bt);
  if (tiles >= sm_count) return 1;
  int S = tc5_splits(M, N, K, sm_count);
  if (S == 1) return 1;
  if (bt > 64 || (N & 3)) return 0;
  const int nkb = ceil_div(K, TC_BK);
  for (; S >= 4; S--) {  // >= 4: an owner warp finishes at most BT / 16 rows (kernel, OWN)
    if (tiles * S > sm_count) continue;
    if (ceil_div(nkb, ceil_div(nkb, S)) != S) continue;  // no empty tail slice
    const int fit = bt == 32 ? max_clusters<32, CL_STAGES_32>(S) : max_clusters<64, CL_STAGES_64>(S);
    if (tiles <= fit) return S;
  }
  return 0;
}

bool carveout_known(const void* fn) {
  static std::mutex mu;
  static std::map<std::pair<const void*, int>, bool> seen;
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> g(mu);
  auto key = std::make_pair(fn, dev);
  if (seen.count(key)) return true;
  seen[key] = true;
  return false;
}

// bit 0: weight prefetch before the dependency wait (default on); bit 1: timing experiment
// only — skip the activation loads (results are garbage)
static int prefetch_flags() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("NVLLM_PREFETCH");
    v = e ? atoi(e) : 1;
  }
  return v;
}
static unsigned long long* g_trace_buf = nullptr;
static unsigned long long* g_trace_agg = nullptr;
static int g_trace_agg_cap = 0;
static int g_trace_seq = 0;
void trace_set(unsigned long long* buf) {
  g_trace_buf = buf;
  g_trace_seq = 0;
}
void trace_agg_set(unsigned long long* agg, int capacity) {
  g_trace_agg = agg;
  g_trace_agg_cap = capacity;
  g_trace_seq = 0;
}
bool trace_active() { return g_trace_buf != nullptr || g_trace_agg != nullptr; }
int trace_seq() { return g_trace_seq; }
Trace next_trace(int kid) {
  Trace t{g_trace_buf, 0, kid, nullptr};
  if (g_trace_buf || g_trace_agg) {
    t.seq = g_trace_seq++;
    if (g_trace_agg && t.seq < g_trace_agg_cap) t.agg = g_trace_agg;
  }
  return t;
}
static bool g_pdl = true;
void set_pdl_enabled(bool on) { g_pdl = on; }
bool pdl_enabled() { return g_pdl; }

// CL: the fused-residual split-K variant — the k-splits of a tile form a thread-block cluster and
// exchange their slices through (BT + 8) x 128 fp32 of extra shared memory per CTA
template <int BT, int STAGES, bool CL = false>
static cudaError_t launch_bt(cudaStream_t st, const Tc5Args& a, const TmaDesc* tw,
                             const TmaDesc* ta, int S, const Tc5Args* a2 = nullptr,
                             const TmaDesc* tw2 = nullptr, const TmaDesc* ta2 = nullptr, int S2 = 0,
                             unsigned* chain_ctr = nullptr) {
  constexpr size_t smem = tc5_smem_bytes(BT, STAGES, CL);
  static_assert(smem + 8192 + (BT > 64 ? 2560 : 1024) <= 232448,
                "shared memory budget (dynamic + the kernel's static arrays)");
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
  Tc5Sched sc{};
  sc.nkb_total = ceil_div(a.K, TC_BK);
  sc.kbps = ceil_div(sc.nkb_total, S);
  sc.S = ceil_div(sc.nkb_total, sc.kbps);
  sc.n_tiles = ceil_div(a.N, TC_BN);
  int m_tiles = ceil_div(a.M, BT);
  if (a.g_experts > 0) {
    sc.g_expert = a.g_tile_expert;
    sc.g_row0 = a.g_tile_row0;
    sc.g_rows = a.g_tile_rows;
    sc.g_ntiles = a.g_n_tiles;
    sc.g_N = a.N;
    // average rows per expert within one token tile: the one-tile-per-expert guess is plausible
    if (a.M <= (long)BT * a.g_experts && (((size_t)a.K * 2) % 16) == 0) {
      sc.g_wbase = (const uint8_t*)a.W;
      sc.g_row_bytes = a.K * 2;
      sc.g_E = a.g_experts;
    }
    m_tiles = min(a.g_max_tiles, a.M / BT + a.g_experts);  // upper bound; the kernel reads the count
  }
  sc.units = sc.n_tiles * m_tiles * sc.S;
  int grid = min(sc.units, max(1, sm_counts[dev]));
  if (a.max_ctas > 0 && a.epi != TC5_EPI_RESID) grid = max(1, min(grid, a.max_ctas));
  Tc5Fuse fz{a.rs_ssq, a.rs_parts, a.rs_ld, a.rs_inv_h, a.rs_eps, a.res_x, (bf16*)a.res_xb, a.res_ssq,
             a.res_ssq_ld, a.res_bias, a.res_m, a.res_g, CL ? sc.S : 0};
  if (a.epi == TC5_EPI_RESID) {
    if (!a.res_x || !a.res_xb || !a.res_ssq || a.g_experts > 0) return cudaErrorInvalidValue;
    // split-K only as a cluster: one CTA per unit, one token tile, cluster = the S splits of a tile
    if ((sc.S > 1) != CL) return cudaErrorInvalidValue;
    if (CL && (sc.S > 8 || sc.S < 4 || m_tiles != 1 || (a.N & 3) || a2)) return cudaErrorInvalidValue;
  } else if (CL) {
    return cudaErrorInvalidValue;
  }
  if (CL) {
    return launch_k_cluster(gemm_tc5_kernel<BT, STAGES>, dim3(sc.units), dim3(TC_THREADS), smem, st, sc.S,
                            *reinterpret_cast<const CUtensorMap*>(tw),
                            *reinterpret_cast<const CUtensorMap*>(ta),
                            *reinterpret_cast<const CUtensorMap*>(tw),
                            *reinterpret_cast<const CUtensorMap*>(ta), a.M, a.N, sc, a.out, a.ldo,
                            a.split_stride, a.alpha, 0, a.epi, (bf16*)a.epi_out, a.epi_ld, a.epi_bias,
                            prefetch_flags(), (const uint8_t*)nullptr, 0ull, fz, Tc5Chain{}, next_trace(1));
  }
  const bool w_reused = a.g_experts > 0 ? a.M > 2 * BT * a.g_experts : m_tiles > 1;
  Tc5Chain ch{};
  int grid_l = grid;
  if (a2) {
    // chained second GEMM: same token tile, plain split-K partials; its k-splits wait inside
    // the kernel for problem 1, so every CTA must be resident (grid <= #SM holds by construction)
    if (a.g_experts > 0 || m_tiles != 1 || !chain_ctr || a2->M != a.M) return cudaErrorInvalidValue;
    ch.sc.nkb_total = ceil_div(a2->K, TC_BK);
    ch.sc.kbps = ceil_div(ch.sc.nkb_total, S2);
    ch.sc.S = ceil_div(ch.sc.nkb_total, ch.sc.kbps);
    ch.sc.n_tiles = ceil_div(a2->N, TC_BN);
    ch.sc.units = ch.sc.n_tiles * ch.sc.S;
    ch.N = a2->N;
    ch.out = a2->out;
    ch.ldo = a2->ldo;
    ch.split_stride = a2->split_stride;
    ch.alpha = a2->alpha;
    ch.ctr = chain_ctr;
    grid_l = min(max(sc.units, ch.sc.units), max(1, sm_counts[dev]));
    ch.n_waiters = min(grid_l, ch.sc.units);
  }
  return launch_k(gemm_tc5_kernel<BT, STAGES>, dim3(grid_l), dim3(TC_THREADS), smem, st,
                  *reinterpret_cast<const CUtensorMap*>(tw),
                  *reinterpret_cast<const CUtensorMap*>(ta),
                  *reinterpret_cast<const CUtensorMap*>(a2 ? tw2 : tw),
                  *reinterpret_cast<const CUtensorMap*>(a2 ? ta2 : ta), a.M, a.N, sc, a.out, a.ldo,
                  a.split_stride, a.alpha, w_reused ? 1 : 0, a.epi, (bf16*)a.epi_out, a.epi_ld,
                  a.epi_bias, prefetch_flags(),
                  (const uint8_t*)((prefetch_flags() & 4) ? nullptr : a.pf_ptr),
                  (unsigned long long)((prefetch_flags() & 4) ? 0 : a.pf_bytes), fz, ch, next_trace(1));
}

// a1: GEMM with a fused activation epilogue (S = 1) writing the bf16 matrix that is a2's A;
// a2: plain split-K GEMM.  Both for the same M <= 128 tokens.  ctr: two zeroed counters.
cudaError_t launch_gemm_tc5_chain(cudaStream_t st, const Tc5Args& a1, const Tc5Args& a2,
                                  unsigned* ctr, int* S2_out) {
  if (!tc5_supported(a1.M, a1.N, a1.K) || !tc5_supported(a2.M, a2.N, a2.K)) return cudaErrorInvalidValue;
  if (a1.g_experts > 0 || a2.g_experts > 0 || a1.M != a2.M || a1.M > 128) return cudaErrorInvalidValue;
  if ((a1.epi != TC5_EPI_SWIGLU && a1.epi != TC5_EPI_GELU) || !a1.epi_out || a2.epi != TC5_EPI_PARTIAL)
    return cudaErrorInvalidValue;
  if (a1.epi == TC5_EPI_SWIGLU && (a1.N & 1)) return cudaErrorInvalidValue;
  const int bt = pick_bt(a1.M);
  const TmaDesc* tw = cached_map(a1.W, a1.N, a1.K, TC_BN);
  const TmaDesc* ta = cached_map(a1.A, a1.M, a1.K, bt);
  const TmaDesc* tw2 = cached_map(a2.W, a2.N, a2.K, TC_BN);
  const TmaDesc* ta2 = cached_map(a2.A, a2.M, a2.K, bt);
  if (!tw || !ta || !tw2 || !ta2) return cudaErrorInvalidValue;
  const int nkb2 = ceil_div(a2.K, TC_BK);
  int S2 = max(1, min(a2.S, nkb2));
  S2 = ceil_div(nkb2, ceil_div(nkb2, S2));
  if (S2_out) *S2_out = S2;
  if (a2.out_capacity && (size_t)S2 * a2.split_stride > a2.out_capacity) return cudaErrorInvalidValue;
  switch (bt) {
    case 32: return launch_bt<32, 10>(st, a1, tw, ta, 1, &a2, tw2, ta2, S2, ctr);
    case 64: return launch_bt<64, 8>(st, a1, tw, ta, 1, &a2, tw2, ta2, S2, ctr);
    default: return launch_bt<128, 6>(st, a1, tw, ta, 1, &a2, tw2, ta2, S2, ctr);
  }
}

// a.S is the requested number of splits; the number actually written is returned in *S_out
int tc5_group_tile(int rows_per_expert) { return pick_bt(rows_per_expert + rows_per_expert / 4); }

cudaError_t launch_gemm_tc5(cudaStream_t st, const Tc5Args& a, int* S_out) {
  if (!tc5_supported(a.M, a.N, a.K)) return cudaErrorInvalidValue;
  const bool grouped = a.g_experts > 0;
  if (grouped && (!a.g_tile_expert || !a.g_tile_row0 || !a.g_tile_rows || !a.g_n_tiles ||
                  (a.g_tile != 32 && a.g_tile != 64 && a.g_tile != 128 && a.g_tile != 256)))
    return cudaErrorInvalidValue;
  const int bt = grouped ? a.g_tile : pick_bt(a.M);
  const TmaDesc* tw = cached_map(a.W, grouped ? a.N * a.g_experts : a.N, a.K, TC_BN);
  const TmaDesc* ta = cached_map(a.A, a.M, a.K, bt);
  if (!tw || !ta) return cudaErrorInvalidValue;
  const int nkb = ceil_div(a.K, TC_BK);
  int S = grouped ? 1 : max(1, min(a.S, nkb));
  const int kbps = ceil_div(nkb, S);
  S = ceil_div(nkb, kbps);
  if (a.epi != TC5_EPI_PARTIAL && a.epi != TC5_EPI_RESID) {
    if (!a.epi_out || (a.epi == TC5_EPI_SWIGLU && (a.N & 1))) return cudaErrorInvalidValue;
    S = 1;
  }
  if (S_out) *S_out = S;
  if (a.epi == TC5_EPI_PARTIAL && a.out_capacity && (size_t)S * a.split_stride > a.out_capacity)
    return cudaErrorInvalidValue;
  if (a.epi == TC5_EPI_RESID && S > 1) {
    switch (bt) {
      case 32: return launch_bt<32, CL_STAGES_32, true>(st, a, tw, ta, S);
      case 64: return launch_bt<64, CL_STAGES_64, true>(st, a, tw, ta, S);
      default: return cudaErrorInvalidValue;
    }
  }
  switch (bt) {
    case 32: return launch_bt<32, 10>(st, a, tw, ta, S);
    case 64: return launch_bt<64, 8>(st, a, tw, ta, S);
    case 128: return launch_bt<128, 6>(st, a, tw, ta, S);
    default: return launch_bt<256, 4>(st, a, tw, ta, S);
  }
}

}  // namespace nvllm
