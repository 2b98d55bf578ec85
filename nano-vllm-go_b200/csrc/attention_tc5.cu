// attention_tc5.cu — paged causal prefill attention on the 5th-generation tensor cores
// (tcgen05.mma, accumulators in tensor memory, operands by TMA), bf16, head dim 64 / 128.
//
// Reference math: scores = Q K^T * scale, causal mask, row softmax, P V
// (purego/tensor/attention.go:354-421 GQA, :127-191 MHA, mqa.go:184-243 MQA), K / V read through
// the scheduler's block tables, the cached prefix of a sequence (NumCachedTokens) attended but
// not recomputed.
//
// One CTA = 128 queries of one sequence x one query head; the SM holds two such CTAs, so that
// one's softmax runs under the other's MMAs.  Per key tile j (BN keys):
//   TMA warp      K_j / V_j tiles of the paged pool (one (block, kv head) slab is contiguous;
//                 2-D tensor map over the layer's pool, SWIZZLE_128B) -> shared memory
//   MMA warp      S = Q K_j^T   (128 x BN, K-major x K-major, accumulator in TMEM)
//                 O_j = P_j V_j (128 x hd; P from shared memory, V as the MN-major operand —
//                 the pool keeps V [key][hd], so no transpose anywhere)
//   softmax warps thread r owns query row r (TMEM lane r): the S row is read into registers
//                 once (all tcgen05.ld issued, one wait), row max, exp2 / row sum, bf16 P
//                 written in the swizzled K-major operand layout.  O stays in TMEM across the
//                 tiles (accumulating MMAs); it is rescaled there only when the running
//                 maximum has grown by more than a factor 2^8 ("lazy" flash rescaling), and
//                 read once at the end — nothing but P and the final output leaves the SM
#include <cuda.h>
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"
#include "tc5.cuh"
#include "tma.cuh"

namespace nvllm {

namespace {

constexpr int PT_BM = 128;       // queries per CTA (UMMA M)
constexpr int PT_THREADS = 256;  // warps 0-3 softmax (warpgroup 0), warp 4 TMA, warp 5 MMA, 6-7 idle

template <int HD>
struct PtCfg {
  static constexpr int BN = HD == 64 ? 128 : 64;  // keys per tile
  static constexpr int KST = 2;                    // K stages
  static constexpr int SUB = HD / 64;              // 64-column (128-byte) sub-tiles of a head row
  static constexpr uint32_t Q_BYTES = PT_BM * HD * 2;
  static constexpr uint32_t KV_BYTES = BN * HD * 2;
  static constexpr uint32_t P_BYTES = PT_BM * BN * 2;
  static constexpr uint32_t TILES = Q_BYTES + KST * KV_BYTES + KV_BYTES + P_BYTES;
  static constexpr uint32_t SMEM = 1024 + TILES + 128;
  static constexpr uint32_t TMEM_COLS = 256;  // S: BN columns at 0, O: HD columns at BN
};

__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

}  // namespace

template <int HD>
__global__ void __launch_bounds__(PT_THREADS, 2)
attn_prefill_tc5_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                        const __grid_constant__ CUtensorMap tmV, AttnArgs a, Trace tr) {
  using Cf = PtCfg<HD>;
  constexpr int BN = Cf::BN, SUB = Cf::SUB, KST = Cf::KST;
  TRACE_ENTRY(tr);
  pdl_launch_dependents();
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ = base;
  const uint32_t sK = sQ + Cf::Q_BYTES;
  const uint32_t sV = sK + KST * Cf::KV_BYTES;
  const uint32_t sP = sV + Cf::KV_BYTES;
  const uint32_t bars = sP + Cf::P_BYTES;
  // barriers: q_full, k_full[2], k_empty[2], v_full, v_empty, s_full, p_ready, o_full
  const uint32_t b_qfull = bars, b_kfull = bars + 8, b_kempty = bars + 24, b_vfull = bars + 40,
                 b_vempty = bars + 48, b_sfull = bars + 56, b_pready = bars + 64, b_ofull = bars + 72;
  const uint32_t tmem_slot = bars + 80;
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // heaviest tiles (latest queries of a sequence see the most keys) first
  const int tile = a.n_qtiles - 1 - (int)blockIdx.x, head = blockIdx.y;
  const int seq = a.tile_seq[tile], q0 = a.tile_q0[tile];
  const int row0 = a.cu_q[seq];
  const int n_new = a.cu_q[seq + 1] - row0;
  const int start = a.seq_start[seq];
  const int kvh = head / (a.nh / a.nkv);
  const int rows_valid = min(PT_BM, n_new - q0);
  const int kv_end = start + q0 + rows_valid;  // keys [0, kv_end) are visible to this tile
  const int n_tiles = (kv_end + BN - 1) / BN;
  const int* bt = a.block_tables + (size_t)seq * a.max_blocks;

  if (warp == 4 && lane == 0) {
    mbar_init(b_qfull, 1);
    for (int s = 0; s < KST; s++) {
      mbar_init(b_kfull + 8 * s, 1);
      mbar_init(b_kempty + 8 * s, 1);
    }
    mbar_init(b_vfull, 1);
    mbar_init(b_vempty, 1);
    mbar_init(b_sfull, 1);
    mbar_init(b_pready, 4);  // one arrival per softmax warp
    mbar_init(b_ofull, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmQ) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmK) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmV) : "memory");
  }
  if (warp == 5) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot),
                 "r"(Cf::TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  const uint32_t tS = tmem_base, tO = tmem_base + BN;

  if (warp >= 4) {
    // warpgroup 1 (TMA warp 4, MMA warp 5, two idle warps) hands its registers to the softmax
    // warpgroup, whose threads carry a whole fp32 output row
    asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
    if (warp == 4) {
      // ---------------------------------------------------------------- TMA producer
      pdl_wait();  // Q and this step's K / V rows are the previous kernels' output
      if (elect_one()) {
        mbar_expect_tx(b_qfull, Cf::Q_BYTES);
        for (int s = 0; s < SUB; s++)
          tma_load_2d(sQ + s * (PT_BM * 128), &tmQ, b_qfull, head * HD + 64 * s, row0 + q0);
      }
      __syncwarp();
      for (int j = 0; j < n_tiles; j++) {
        const int k0 = j * BN;
        const int prow = (bt[k0 / KV_BLOCK] * a.nkv + kvh) * KV_BLOCK + (k0 % KV_BLOCK);
        const int ks = j % KST;
        if (j >= KST) mbar_wait(b_kempty + 8 * ks, ((j / KST) & 1) ^ 1);
        if (elect_one()) {
          mbar_expect_tx(b_kfull + 8 * ks, Cf::KV_BYTES);
          for (int s = 0; s < SUB; s++)
            tma_load_2d(sK + ks * Cf::KV_BYTES + s * (BN * 128), &tmK, b_kfull + 8 * ks, 64 * s, prow);
        }
        __syncwarp();
        if (j >= 1) mbar_wait(b_vempty, (j & 1) ^ 1);
        if (elect_one()) {
          mbar_expect_tx(b_vfull, Cf::KV_BYTES);
          for (int s = 0; s < SUB; s++) tma_load_2d(sV + s * (BN * 128), &tmV, b_vfull, 64 * s, prow);
        }
        __syncwarp();
      }
    } else if (warp == 5) {
      // ---------------------------------------------------------------- MMA issuer
      // instruction descriptors (cute/arch/mma_sm100_desc.hpp): fp32 accumulate, bf16 x bf16,
      // M = 128; S: N = BN, both operands K-major; O: N = hd, B (= V) MN-major (bit 16)
      constexpr uint32_t idesc_s = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) |
                                   ((uint32_t)(PT_BM >> 4) << 24);
      constexpr uint32_t idesc_o = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 16) |
                                   ((uint32_t)(HD >> 3) << 17) | ((uint32_t)(PT_BM >> 4) << 24);
      mbar_wait(b_qfull, 0);
      for (int j = 0; j < n_tiles; j++) {
        const int ks = j % KST;
        mbar_wait(b_kfull + 8 * ks, (j / KST) & 1);
        tc_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int s = 0; s < SUB; s++) {
            const uint64_t qd = make_smem_desc(sQ + s * (PT_BM * 128));
            const uint64_t kd = make_smem_desc(sK + ks * Cf::KV_BYTES + s * (BN * 128));
#pragma unroll
            for (int kk = 0; kk < 4; kk++) tc_mma_bf16(tS, qd + 2 * kk, kd + 2 * kk, idesc_s, (s | kk) != 0);
          }
          tc_commit(b_kempty + 8 * ks);
          tc_commit(b_sfull);
        }
        __syncwarp();
        mbar_wait(b_pready, j & 1);  // P_j is in shared memory; S_j and O_{j-1} have been read
        mbar_wait(b_vfull, j & 1);
        tc_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int kk = 0; kk < BN / 16; kk++) {
            const uint64_t pd = make_smem_desc(sP + (kk / 4) * (PT_BM * 128)) + 2 * (kk % 4);
            const uint64_t vd = make_smem_desc_mn(sV + kk * 2048, BN * 128);
            tc_mma_bf16(tO, pd, vd, idesc_o, (j | kk) != 0);
          }
          tc_commit(b_vempty);
          tc_commit(b_ofull);
        }
        __syncwarp();
      }
    }
  } else {
    // ------------------------------------------------------------------ softmax + output
    asm volatile("setmaxnreg.inc.sync.aligned.u32 216;");
    const int q = warp;  // TMEM lane quadrant of this warp
    const int r = q * 32 + lane;
    const int qpos = start + q0 + r;
    const float sl2 = a.scale * 1.4426950408889634f;
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    // O accumulates in TMEM across the key tiles (the P V MMAs run with accumulate = 1).  The
    // exponent reference m_used only follows the running maximum when it has grown by more than
    // 2^RESCALE_LOG2 in probability — then, and only then, O (TMEM) and l are rescaled; in
    // between, p = exp2(s - m_used) may exceed 1 by at most that factor, harmless in fp32 / bf16.
    constexpr float RESCALE_LOG2 = 8.f;
    float m_used = -INFINITY, l_run = 0.f;
    for (int j = 0; j < n_tiles; j++) {
      const int k0 = j * BN;
      // the warp's rows see every key of the tile when its last key <= the first row's position
      const bool need_mask = (k0 + BN - 1) > (start + q0 + q * 32);
      mbar_wait(b_sfull, j & 1);  // also: every earlier P V MMA has retired (in-order pipe)
      tc_fence_after();
      // the whole S row of this tile in registers: all loads issued, one wait
      uint32_t sr[BN];
#pragma unroll
      for (int c0 = 0; c0 < BN; c0 += 32) tmem_ld_32x32_nowait(tS + lane_off + (uint32_t)c0, &sr[c0]);
      tmem_ld_wait();
      float mx = -INFINITY;
      if (need_mask) {
#pragma unroll
        for (int c = 0; c < BN; c++) {
          if (k0 + c > qpos) sr[c] = 0xff800000u;  // -inf: p = 0
          mx = fmaxf(mx, __uint_as_float(sr[c]));
        }
      } else {
#pragma unroll
        for (int c = 0; c < BN; c++) mx = fmaxf(mx, __uint_as_float(sr[c]));
      }
      // rows of the tile that see no key yet (mx = -inf) keep their reference
      const bool grow = (mx - m_used) * sl2 > RESCALE_LOG2 || (m_used == -INFINITY && mx > -INFINITY);
      if (__any_sync(0xffffffffu, grow)) {
        const float m_new = grow ? mx : m_used;
        const float corr = (m_used == -INFINITY) ? 0.f : ex2((m_used - m_new) * sl2);
        l_run *= corr;
        m_used = m_new;
        if (j > 0) {  // tile 0's P V overwrites O: nothing to rescale yet
#pragma unroll 1
          for (int c0 = 0; c0 < HD; c0 += 32) {
            uint32_t rr[32];
            tmem_ld_32x32(tO + lane_off + (uint32_t)c0, rr);
#pragma unroll
            for (int c = 0; c < 32; c++) rr[c] = __float_as_uint(__uint_as_float(rr[c]) * corr);
            tmem_st_32x32(tO + lane_off + (uint32_t)c0, rr);
          }
          tmem_st_wait();
        }
      }
      const float mb = (m_used == -INFINITY) ? 0.f : m_used * sl2;
      float sum = 0.f;
#pragma unroll
      for (int c0 = 0; c0 < BN; c0 += 32) {
        uint32_t pk[16];
#pragma unroll
        for (int c = 0; c < 32; c += 2) {
          const float p0 = ex2(fmaf(__uint_as_float(sr[c0 + c]), sl2, -mb));
          const float p1 = ex2(fmaf(__uint_as_float(sr[c0 + c + 1]), sl2, -mb));
          sum += p0 + p1;
          __nv_bfloat162 b2 = __floats2bfloat162_rn(p0, p1);
          pk[c >> 1] = *reinterpret_cast<uint32_t*>(&b2);
        }
        const uint32_t prow_addr = sP + (uint32_t)(c0 >> 6) * (PT_BM * 128) + (uint32_t)r * 128;
#pragma unroll
        for (int i = 0; i < 4; i++) {
          const uint32_t ch = (uint32_t)(((c0 & 63) >> 3) + i);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(prow_addr + ((ch ^ (uint32_t)(r & 7)) << 4)),
                       "r"(pk[4 * i]), "r"(pk[4 * i + 1]), "r"(pk[4 * i + 2]), "r"(pk[4 * i + 3])
                       : "memory");
        }
      }
      l_run += sum;
      tc_fence_before();
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(b_pready);
    }
    // o_full completes once per tile and is only waited for here.  s_full of the last tile
    // implied that P V of tile n_tiles - 2 had retired, so the barrier is at least in phase
    // n_tiles - 1 now and the parity test below can only pass on the last tile's completion.
    mbar_wait(b_ofull, (n_tiles - 1) & 1);
    tc_fence_after();
    const float inv = 1.0f / l_run;
    bf16* op = (bf16*)a.out + (size_t)(row0 + q0 + r) * a.nh * HD + (size_t)head * HD;
#pragma unroll 1
    for (int c0 = 0; c0 < HD; c0 += 32) {
      uint32_t rr[32];
      tmem_ld_32x32(tO + lane_off + (uint32_t)c0, rr);
      if (r < rows_valid) {
#pragma unroll
        for (int c = 0; c < 32; c += 8) {
          float v[8];
#pragma unroll
          for (int i = 0; i < 8; i++) v[i] = __uint_as_float(rr[c + i]) * inv;
          store8(op + c0 + c, v);
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 5) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                 "r"(Cf::TMEM_COLS)
                 : "memory");
  }
  TRACE_EXIT(tr);
}

// ------------------------------------------------------------------ launcher
bool attn_prefill_tc5_enabled() {
  static const bool on = !(getenv("NVLLM_ATTN_PREFILL_TC5") && atoi(getenv("NVLLM_ATTN_PREFILL_TC5")) == 0);
  return on;
}

template <int HD>
static cudaError_t prefill_tc5_launch(cudaStream_t st, const AttnArgs& a) {
  using Cf = PtCfg<HD>;
  const int pool_rows = a.pool_blocks * a.nkv * KV_BLOCK;
  const TmaDesc* tq = tma_map_cached(a.q, a.total_rows, a.nh * HD, PT_BM);
  const TmaDesc* tk = tma_map_cached(a.kpool, pool_rows, HD, Cf::BN);
  const TmaDesc* tv = tma_map_cached(a.vpool, pool_rows, HD, Cf::BN);
  if (!tq || !tk || !tv) return cudaErrorInvalidValue;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(attn_prefill_tc5_kernel<HD>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cf::SMEM);
    if (e != cudaSuccess) return e;
    attr_set = true;
  }
  (void)launch_k(attn_prefill_tc5_kernel<HD>, dim3(a.n_qtiles, a.nh), dim3(PT_THREADS), Cf::SMEM, st,
                 *reinterpret_cast<const CUtensorMap*>(tq), *reinterpret_cast<const CUtensorMap*>(tk),
                 *reinterpret_cast<const CUtensorMap*>(tv), a, next_trace(13));
  return cudaGetLastError();
}

// needs AttnArgs::pool_blocks / total_rows and a tile list built with 128-query tiles
cudaError_t launch_attn_prefill_tc5(cudaStream_t st, const AttnArgs& a) {
  if (a.n_qtiles <= 0) return cudaSuccess;
  if (a.pool_blocks <= 0 || a.total_rows <= 0) return cudaErrorInvalidValue;
  return a.hd == 64 ? prefill_tc5_launch<64>(st, a) : prefill_tc5_launch<128>(st, a);
}

}  // namespace nvllm
