// Persistent, warp-specialised tcgen05 GEMM for sm_100a.
//
//   acc[m][n] = sum_k A(m,k) * B(n,k)    (bf16 x bf16 or tf32 x tf32 -> fp32 in TMEM; template parameter EB)
//
// Roles inside one 192-thread CTA (one CTA per SM, static round-robin tile schedule):
//   warp 0      TMA producer: cp.async.bulk.tensor tiles of A and B into a STAGES-deep smem ring
//   warp 1      TMEM owner + MMA issuer: one lane issues tcgen05.mma (M=128, N=BN, K=16) x4 per stage,
//               tcgen05.commit releases the smem slot and, after the last k-block, publishes the accumulator
//   warps 2..9  epilogue: tcgen05.ld the 128 x BN fp32 accumulator (one TMEM lane = one output row per
//               thread), fuse bias / activation / activation-gradient / accumulate, store fp32 and/or bf16
// Two TMEM accumulator buffers let the MMA of tile i+1 overlap the epilogue of tile i.
//
// Both operand majors are supported because the three InnerProduct GEMMs need them with the reference's
// {vdim, hdim} weight layout (src/worker/layer.cc:186-213):
//   fwd   Y  = X  W      A = X  (K-major)   B = W  (N-major)
//   dgrad dX = dY W^T    A = dY (K-major)   B = W  (K-major)
//   wgrad dW = X^T dY    A = X  (M-major)   B = dY (N-major)
// smem tiles use the 128-byte swizzle; K-major tiles are [rows][64 k] with 128-byte rows, MN-major tiles
// are stacks of [64 k][64 mn] boxes (8 KB each) — the canonical tcgen05 layouts for SWIZZLE_128B.
#include <cuda.h>
#include <stdlib.h>
#include <mutex>
#include <unordered_map>
#include <string>

#include "../include/singa_b200.h"
#include "common.h"
#include "ptx.cuh"
#include "gemm_common.cuh"

namespace sg {


template <int EB, int BN, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a,
                    const __grid_constant__ CUtensorMap tma_b, const GemmParams p) {
  using Cfg = GemmCfg<BN>;
  using E = Elem<EB>;
  constexpr int BK = E::BK;
  constexpr int UMMA_K = E::UMMA_K;
  constexpr int MN_BOX = E::MN_BOX;
  constexpr int MN_BOX_BYTES = E::MN_BOX_BYTES;
  constexpr int STAGES = Cfg::STAGES;
  constexpr int B_STAGE_BYTES = Cfg::B_STAGE_BYTES;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sA0 = smem_base;
  const uint32_t sB0 = smem_base + STAGES * A_STAGE_BYTES;
  const uint32_t bar0 = smem_base + STAGES * Cfg::STAGE_BYTES;
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (STAGES + s); };
  auto tfull_bar = [&](int a) { return bar0 + 8u * (2 * STAGES + a); };
  auto tempty_bar = [&](int a) { return bar0 + 8u * (2 * STAGES + 2 + a); };
  const uint32_t tmem_slot = bar0 + 8u * (2 * STAGES + 4);
  uint32_t* tmem_slot_ptr =
      reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  if (threadIdx.x == 0) SG_TRACE(0);

  const int m_tiles = (p.M + BM - 1) / BM;
  const int n_tiles = (p.N + BN - 1) / BN;
  const int num_tiles = m_tiles * n_tiles;
  const int num_kb = (p.K + BK - 1) / BK;
  // split-K: work item w = (tile, k-slice); each item reduces its own range of k-blocks and the epilogue adds the
  // partial tile into the zeroed fp32 output with red.global.add (weight gradients of convolutions have a huge
  // K = batch x pixels and only a handful of output tiles)
  const int num_items = num_tiles * p.split_k;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), EPI_WARPS);  // one arrive per epilogue warp
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(tmem_slot, Cfg::TMEM_COLS);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  if (threadIdx.x == 0) SG_TRACE(1);
  // everything above touched only this CTA's shared memory / TMEM: with a programmatic launch it overlapped the tail
  // of the previous kernel.  From here on global memory written by that kernel is read.
  pdl_wait();
  pdl_launch_dependents();

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    // The whole warp walks the loop and one elected lane issues (see elect_one): coordinates stay in uniform registers.
    {
      if (lane == 0 && p.nwait > 0) wait_param_versions(p.wait_flags, p.nwait, p.wait_seq);
      __syncwarp();
      uint32_t stage = 0, phase = 0;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
        const int slice = p.split_k == 1 ? 0 : item / num_tiles;   // no division on the common path
        const int tile = item - slice * num_tiles;
        const int kb_lo = slice * p.kb_per_split;
        const int kb_hi = min(num_kb, kb_lo + p.kb_per_split);
        const int m0 = (tile % m_tiles) * BM;
        const int n0 = (tile / m_tiles) * BN;
        for (int kb = kb_lo; kb < kb_hi; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          const uint32_t fb = full_bar(stage);
          const uint32_t sa = sA0 + stage * A_STAGE_BYTES;
          const uint32_t sb = sB0 + stage * B_STAGE_BYTES;
          const int k0 = kb * BK;
          if (elect_one()) {
            mbar_arrive_expect_tx(fb, Cfg::STAGE_BYTES);
            if (!A_MN) {
              tma_load_2d(sa, &tma_a, fb, k0, m0);
            } else {
#pragma unroll
              for (int b = 0; b < BM / MN_BOX; ++b) tma_load_2d(sa + b * MN_BOX_BYTES, &tma_a, fb, m0 + b * MN_BOX, k0);
            }
            if (!B_MN) {
              tma_load_2d(sb, &tma_b, fb, k0, n0);
            } else {
#pragma unroll
              for (int b = 0; b < BN / MN_BOX; ++b) tma_load_2d(sb + b * MN_BOX_BYTES, &tma_b, fb, n0 + b * MN_BOX, k0);
            }
            if (kb == kb_lo) SG_TRACE(2);
          }
          __syncwarp();
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
      if (lane == 0) SG_TRACE(3);
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    constexpr uint32_t idesc = make_idesc(EB == 2 ? 1u : 2u, BM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
    constexpr uint32_t A_KSTEP = A_MN ? (UMMA_K * 128) : 32;  // bytes per UMMA_K step
    constexpr uint32_t B_KSTEP = B_MN ? (UMMA_K * 128) : 32;
    constexpr uint32_t A_LBO = A_MN ? MN_BOX_BYTES : 16;
    constexpr uint32_t B_LBO = B_MN ? MN_BOX_BYTES : 16;
    // MN-major tf32 tiles use the 32-byte-atom swizzle: groups of 4 k-rows (512 bytes) instead of 8 (1024)
    constexpr uint32_t A_LAYOUT = (A_MN && EB == 4) ? kLayoutSw128Base32 : kLayoutSw128;
    constexpr uint32_t B_LAYOUT = (B_MN && EB == 4) ? kLayoutSw128Base32 : kLayoutSw128;
    constexpr uint32_t A_SBO = (A_MN && EB == 4) ? 512 : 1024;
    constexpr uint32_t B_SBO = (B_MN && EB == 4) ? 512 : 1024;
    uint32_t stage = 0, phase = 0;
    int it = 0;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++it) {
      const int kb_lo = p.split_k == 1 ? 0 : (item / num_tiles) * p.kb_per_split;
      const int kb_hi = min(num_kb, kb_lo + p.kb_per_split);
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      mbar_wait(tempty_bar(acc), acc_phase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * BN;
      for (int kb = kb_lo; kb < kb_hi; ++kb) {
        mbar_wait(full_bar(stage), phase);
        tc_fence_after();
        if (lane == 0 && kb == kb_lo) SG_TRACE(4);
        const uint32_t sa = sA0 + stage * A_STAGE_BYTES;
        const uint32_t sb = sB0 + stage * B_STAGE_BYTES;
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint64_t adesc = make_smem_desc(sa + k * A_KSTEP, A_LBO, A_SBO, A_LAYOUT);
            const uint64_t bdesc = make_smem_desc(sb + k * B_KSTEP, B_LBO, B_SBO, B_LAYOUT);
            umma<EB>(tmem_d, adesc, bdesc, idesc, (kb > kb_lo || k != 0) ? 1u : 0u);
          }
          umma_commit(empty_bar(stage));
          if (kb == kb_hi - 1) { umma_commit(tfull_bar(acc)); SG_TRACE(5); }
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 2..9)
    const int quad = warp & 3;                 // TMEM lane quadrant this warp may read
    // this warp's staging tile sits behind the barrier block
    float* stg = reinterpret_cast<float*>(smem_raw + (bar0 - smem_u32(smem_raw)) + 256) + (warp - 2) * EPI_WARP_FLOATS;
    const bool fast_ok = epilogue_fast_ok(p);
    int it = 0;
    for (int item = blockIdx.x; item < num_items; item += gridDim.x, ++it) {
      const int tile = p.split_k == 1 ? item : item % num_tiles;
      const int m0 = (tile % m_tiles) * BM;
      const int n0 = (tile / m_tiles) * BN;
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      if (warp == 2 && lane == 0) SG_TRACE(6);
      const long long row0 = m0 + quad * 32;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + acc * BN;
#pragma unroll 1
      for (int c = ((warp - 2) >> 2) * 32; c < BN; c += 64) {
        uint32_t r[32];
        tmem_ld_32x32(taddr + c, r);
        tmem_ld_wait();
        if (warp == 2 && lane == 0 && c <= 64) SG_TRACE(10 + 2 * (c >> 6));
        if (c + 64 >= BN) {
          // accumulator fully in registers: hand the TMEM buffer back to the MMA warp early
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(tempty_bar(acc));
        }
        epilogue_chunk(p, r, row0, lane, n0 + c, fast_ok, stg);
        if (warp == 2 && lane == 0 && c <= 64) SG_TRACE(11 + 2 * (c >> 6));
      }
      if (warp == 2 && lane == 0) SG_TRACE(7);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 0) SG_TRACE(8);
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
    if (lane == 0) SG_TRACE(9);
  }
}


// ------------------------------------------------------------------------------------------------
// CTA-pair variant: a cluster of two CTAs (two SMs of one TPC) computes one 256 x BN tile with
// tcgen05.mma.cta_group::2.  Each CTA stages its own 128 rows of A and only HALF of the B tile; the tensor core
// reads the other half from the peer's shared memory, so both the L2->smem traffic and the smem read bandwidth
// per flop drop (per CTA and k-block: 16 KB + BN*64 B instead of 16 KB + BN*128 B).
//   leader (cluster rank 0): warp 1 issues every MMA for the pair; commits are multicast to both CTAs' barriers
//   both CTAs: warp 0 = TMA producer (transaction bytes are counted on the LEADER's full barrier),
//              warps 2..9 = epilogue of the CTA's own 128 accumulator rows (its own TMEM), arriving on the
//              leader's tmem-empty barrier
// ------------------------------------------------------------------------------------------------
template <int EB, int BN, bool A_MN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
gemm_tcgen05_pair_kernel(const __grid_constant__ CUtensorMap tma_a,
                         const __grid_constant__ CUtensorMap tma_b, const GemmParams p) {
  using Cfg = PairCfg<BN>;
  using E = Elem<EB>;
  constexpr int BK = E::BK;
  constexpr int UMMA_K = E::UMMA_K;
  constexpr int MN_BOX = E::MN_BOX;
  constexpr int MN_BOX_BYTES = E::MN_BOX_BYTES;
  constexpr int STAGES = Cfg::STAGES;
  constexpr int B_HALF_BYTES = Cfg::B_HALF_BYTES;
  constexpr int BNH = BN / 2;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sA0 = smem_base;
  const uint32_t sB0 = smem_base + STAGES * A_STAGE_BYTES;
  const uint32_t bar0 = smem_base + STAGES * Cfg::STAGE_BYTES;
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (STAGES + s); };
  auto tfull_bar = [&](int a) { return bar0 + 8u * (2 * STAGES + a); };
  auto tempty_bar = [&](int a) { return bar0 + 8u * (2 * STAGES + 2 + a); };
  const uint32_t tmem_slot = bar0 + 8u * (2 * STAGES + 4);
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int pair_id = blockIdx.x >> 1;
  const int num_pairs = gridDim.x >> 1;

  const int m_tiles = (p.M + 2 * BM - 1) / (2 * BM);
  const int n_tiles = (p.N + BN - 1) / BN;
  const int num_tiles = m_tiles * n_tiles;
  const int num_kb = (p.K + BK - 1) / BK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 2);     // leader's own arrive.expect_tx + the peer's remote arrive
      mbar_init(empty_bar(s), 1);    // multicast commit from the leader's MMA warp
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);    // multicast commit
      mbar_init(tempty_bar(a), 2 * EPI_WARPS);   // epilogue warps x 2 CTAs (only the leader's copy is waited on)
    }
    fence_mbar_init();
  }
  cluster_sync_all();                // barrier inits of both CTAs visible before anyone signals / allocates
  if (warp == 1) {
    tmem_alloc_pair(tmem_slot, Cfg::TMEM_COLS);
    tmem_relinquish_pair();
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (both CTAs)
    {
      if (lane == 0 && p.nwait > 0) wait_param_versions(p.wait_flags, p.nwait, p.wait_seq);
      __syncwarp();
      uint32_t stage = 0, phase = 0;
      for (int tile = pair_id; tile < num_tiles; tile += num_pairs) {
        const int m0 = (tile % m_tiles) * (2 * BM) + static_cast<int>(rank) * BM;   // this CTA's 128 rows
        const int n0 = (tile / m_tiles) * BN + static_cast<int>(rank) * BNH;        // this CTA's half of B
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          const uint32_t fb = full_bar(stage);
          const uint32_t sa = sA0 + stage * A_STAGE_BYTES;
          const uint32_t sb = sB0 + stage * B_HALF_BYTES;
          const int k0 = kb * BK;
          if (elect_one()) {
            if (leader) mbar_arrive_expect_tx(fb, 2 * Cfg::STAGE_BYTES);
            else mbar_arrive_remote(fb, 0);
            if (!A_MN) {
              tma_load_2d_pair(sa, &tma_a, fb, k0, m0);
            } else {
#pragma unroll
              for (int b = 0; b < BM / MN_BOX; ++b) tma_load_2d_pair(sa + b * MN_BOX_BYTES, &tma_a, fb, m0 + b * MN_BOX, k0);
            }
            if (!B_MN) {
              tma_load_2d_pair(sb, &tma_b, fb, k0, n0);
            } else {
#pragma unroll
              for (int b = 0; b < BNH / MN_BOX; ++b) tma_load_2d_pair(sb + b * MN_BOX_BYTES, &tma_b, fb, n0 + b * MN_BOX, k0);
            }
          }
          __syncwarp();
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA only)
    if (leader) {
      constexpr uint32_t idesc = make_idesc(EB == 2 ? 1u : 2u, 2 * BM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
      constexpr uint32_t A_KSTEP = A_MN ? (UMMA_K * 128) : 32;
      constexpr uint32_t B_KSTEP = B_MN ? (UMMA_K * 128) : 32;
      constexpr uint32_t A_LBO = A_MN ? MN_BOX_BYTES : 16;
      constexpr uint32_t B_LBO = B_MN ? MN_BOX_BYTES : 16;
      constexpr uint32_t A_LAYOUT = (A_MN && EB == 4) ? kLayoutSw128Base32 : kLayoutSw128;
      constexpr uint32_t B_LAYOUT = (B_MN && EB == 4) ? kLayoutSw128Base32 : kLayoutSw128;
      constexpr uint32_t A_SBO = (A_MN && EB == 4) ? 512 : 1024;
      constexpr uint32_t B_SBO = (B_MN && EB == 4) ? 512 : 1024;
      uint32_t stage = 0, phase = 0;
      int it = 0;
      for (int tile = pair_id; tile < num_tiles; tile += num_pairs, ++it) {
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        mbar_wait(tempty_bar(acc), acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(full_bar(stage), phase);
          tc_fence_after();
          const uint32_t sa = sA0 + stage * A_STAGE_BYTES;
          const uint32_t sb = sB0 + stage * B_HALF_BYTES;
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) {
              const uint64_t adesc = make_smem_desc(sa + k * A_KSTEP, A_LBO, A_SBO, A_LAYOUT);
              const uint64_t bdesc = make_smem_desc(sb + k * B_KSTEP, B_LBO, B_SBO, B_LAYOUT);
              umma_pair<EB>(tmem_d, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
            }
            umma_commit_pair(empty_bar(stage), 3);                       // frees the slot in BOTH CTAs
            if (kb == num_kb - 1) umma_commit_pair(tfull_bar(acc), 3);   // accumulator ready in BOTH CTAs
          }
          __syncwarp();
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 2..9, both CTAs)
    const int quad = warp & 3;
    // this warp's staging tile sits behind the barrier block
    float* stg = reinterpret_cast<float*>(smem_raw + (bar0 - smem_u32(smem_raw)) + 256) + (warp - 2) * EPI_WARP_FLOATS;
    const bool fast_ok = epilogue_fast_ok(p);
    int it = 0;
    for (int tile = pair_id; tile < num_tiles; tile += num_pairs, ++it) {
      const int m0 = (tile % m_tiles) * (2 * BM) + static_cast<int>(rank) * BM;
      const int n0 = (tile / m_tiles) * BN;
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const long long row0 = m0 + quad * 32;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + acc * BN;
#pragma unroll 1
      for (int c = ((warp - 2) >> 2) * 32; c < BN; c += 64) {
        uint32_t r[32];
        tmem_ld_32x32(taddr + c, r);
        tmem_ld_wait();
        if (c + 64 >= BN) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_remote(tempty_bar(acc), 0);   // the leader's MMA warp waits on ITS barrier
        }
        epilogue_chunk(p, r, row0, lane, n0 + c, fast_ok, stg);
      }
    }
  }

  tc_fence_before();
  cluster_sync_all();      // both CTAs done with both TMEMs and with each other's barriers
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, Cfg::TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------
// Host side: tensor maps and dispatch
// ------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(sym);
  });
  return fn;
}

// 2-D tensor map over bf16 (eb = 2) or fp32 (eb = 4) elements: `inner` contiguous elements per row, `outer` rows,
// row stride ld elements.
static int make_tmap(CUtensorMap* out, int eb, const void* ptr, long long inner, long long outer,
                     long long ld, int box_inner, int box_outer, bool mn_major) {
  EncodeTiledFn enc = get_encode_fn();
  SG_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
  SG_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "gemm operand base must be 16-byte aligned");
  SG_REQUIRE((ld * eb) % 16 == 0, "gemm operand leading dimension (%lld elements of %d bytes) must be a multiple of 16 bytes",
             ld, eb);
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(inner), static_cast<cuuint64_t>(outer)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * eb};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(box_inner), static_cast<cuuint32_t>(box_outer)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, eb == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                   const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   (mn_major && eb == 4) ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  SG_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d) inner=%lld outer=%lld ld=%lld",
             static_cast<int>(r), inner, outer, ld);
  return 0;
}

// Programmatic dependent launch of the single-CTA GEMM (opt-in: SINGA_B200_PDL=1): back-to-back GEMMs of a layer chain
// overlap their launch + set-up with the previous kernel's last tiles.  Measured on the MNIST MLP step (B200, CUDA
// graph): 0.444 ms with it against 0.436 ms without in fp32, 0.341 against 0.342 in bf16 — the early-started CTAs sit
// on the SMs that the concurrent weight-gradient GEMMs of the side streams want — so it is off by default.
static int g_pdl = [] { const char* e = getenv("SINGA_B200_PDL"); return (e != nullptr && e[0] == '1') ? 1 : 0; }();

template <int EB, int BN, bool A_MN, bool B_MN>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p,
                       cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  auto kern = gemm_tcgen05_kernel<EB, BN, A_MN, B_MN>;
  static thread_local int attr_dev = -1;
  int dev = 0;
  cudaGetDevice(&dev);
  if (attr_dev != dev) {
    if (check_cuda(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        Cfg::SMEM_BYTES),
                   "cudaFuncSetAttribute(gemm smem)"))
      return 1;
    attr_dev = dev;
  }
  const int m_tiles = (p.M + BM - 1) / BM;
  const int n_tiles = (p.N + BN - 1) / BN;
  const int tiles = m_tiles * n_tiles * p.split_k;
  const int grid = tiles < num_sms() ? tiles : num_sms();
  if (g_pdl) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(GEMM_THREADS); cfg.dynamicSmemBytes = Cfg::SMEM_BYTES; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    if (check_cuda(cudaLaunchKernelEx(&cfg, kern, ta, tb, p), "gemm_tcgen05 (programmatic launch)")) return 1;
    return 0;
  }
  kern<<<grid, GEMM_THREADS, Cfg::SMEM_BYTES, stream>>>(ta, tb, p);
  SG_CHECK_LAUNCH("gemm_tcgen05");
  return 0;
}

template <int EB, int BN, bool A_MN, bool B_MN>
static int launch_gemm_pair(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p,
                            cudaStream_t stream) {
  using Cfg = PairCfg<BN>;
  auto kern = gemm_tcgen05_pair_kernel<EB, BN, A_MN, B_MN>;
  static thread_local int attr_dev = -1;
  int dev = 0;
  cudaGetDevice(&dev);
  if (attr_dev != dev) {
    if (check_cuda(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES),
                   "cudaFuncSetAttribute(gemm pair smem)"))
      return 1;
    attr_dev = dev;
  }
  const int m_tiles = (p.M + 2 * BM - 1) / (2 * BM);
  const int n_tiles = (p.N + BN - 1) / BN;
  const int tiles = m_tiles * n_tiles;
  const int max_pairs = num_sms() / 2;
  const int pairs = tiles < max_pairs ? tiles : max_pairs;
  kern<<<2 * pairs, GEMM_THREADS, Cfg::SMEM_BYTES, stream>>>(ta, tb, p);
  SG_CHECK_LAUNCH("gemm_tcgen05_pair");
  return 0;
}

template <int EB, int BN>
static int dispatch_major_pair(int a_mn, int b_mn, const CUtensorMap& ta, const CUtensorMap& tb,
                               const GemmParams& p, cudaStream_t s) {
  if (!a_mn && !b_mn) return launch_gemm_pair<EB, BN, false, false>(ta, tb, p, s);
  if (!a_mn && b_mn) return launch_gemm_pair<EB, BN, false, true>(ta, tb, p, s);
  if (a_mn && !b_mn) return launch_gemm_pair<EB, BN, true, false>(ta, tb, p, s);
  return launch_gemm_pair<EB, BN, true, true>(ta, tb, p, s);
}

template <int EB, int BN>
static int dispatch_major(int a_mn, int b_mn, const CUtensorMap& ta, const CUtensorMap& tb,
                          const GemmParams& p, cudaStream_t s) {
  if (!a_mn && !b_mn) return launch_gemm<EB, BN, false, false>(ta, tb, p, s);
  if (!a_mn && b_mn) return launch_gemm<EB, BN, false, true>(ta, tb, p, s);
  if (a_mn && !b_mn) return launch_gemm<EB, BN, true, false>(ta, tb, p, s);
  return launch_gemm<EB, BN, true, true>(ta, tb, p, s);
}

template <int EB>
static int dispatch_tile(bool pair, int bn, int a_mn, int b_mn, const CUtensorMap& ta, const CUtensorMap& tb,
                         const GemmParams& p, cudaStream_t s) {
  if (pair) {
    if (bn == 256) return dispatch_major_pair<EB, 256>(a_mn, b_mn, ta, tb, p, s);
    return dispatch_major_pair<EB, 128>(a_mn, b_mn, ta, tb, p, s);
  }
  switch (bn) {
    case 256: return dispatch_major<EB, 256>(a_mn, b_mn, ta, tb, p, s);
    case 128: return dispatch_major<EB, 128>(a_mn, b_mn, ta, tb, p, s);
    default: return dispatch_major<EB, 64>(a_mn, b_mn, ta, tb, p, s);
  }
}

// Tile choice from a small cost model (cycles), fitted to `tools/gemm_micro.py` sweeps of the 18 MNIST-MLP shapes in
// both precisions (it picks the fastest of the five configurations for all 18 tf32 shapes and 17 of the bf16 ones).
// Measured on B200: a CTA takes in about 40 bytes per cycle through TMA however few CTAs run, and the whole chip about
// 6300; the tensor pipe is rarely the limit.  So the cost of a k-block is its bytes per CTA, and CTA pairs (half of
// the B tile per CTA) win whenever their fewer, larger tiles still cover the SMs:
//   per k-block:  mma = max(2*BN, 128 + BN) cycles   (tensor pipe vs. smem operand reads at 128 B/cycle)
//                 tma = bytes per CTA / min(40, 6300 / active CTAs),  bytes = 16 KB + BN*128 B  (pairs: BN*64 B)
//   total = waves * (num_kb * max(mma, tma) + 1500)
static void pick_tile(int M, int N, int K, int BK, bool allow_single, bool allow_pair, int* bn_out, bool* pair_out) {
  const int sms = num_sms();
  const long long num_kb = (K + BK - 1) / BK;
  const int cands[3] = {256, 128, 64};
  double best = 1e30;
  int best_bn = 64;
  bool best_pair = false;
  for (int pr = 0; pr < 2; ++pr) {
    if (pr == 0 ? !allow_single : !allow_pair) continue;
    for (int i = 0; i < 3; ++i) {
      const int bn = cands[i];
      if (pr == 1 && bn < 128) continue;        // pairs stage BN/2 columns of B per CTA in boxes of >= 64 columns
      if (bn > 64 && N <= bn / 2) continue;     // more than half of the tile would be padding
      const long long rows = pr == 1 ? 2 * BM : BM;
      const long long ctas = ((M + rows - 1) / rows) * ((N + bn - 1) / bn) * (pr == 1 ? 2 : 1);
      const long long active = ctas < sms ? ctas : sms;
      const long long waves = (ctas + sms - 1) / sms;
      const double mma = (2.0 * bn > 128.0 + bn) ? 2.0 * bn : 128.0 + bn;
      double bw = 6300.0 / static_cast<double>(active);
      if (bw > 40.0) bw = 40.0;
      const double tma = (16384.0 + bn * (pr == 1 ? 64.0 : 128.0)) / bw;
      const double total = waves * (num_kb * (mma > tma ? mma : tma) + 1500.0);
      if (total < best) { best = total; best_bn = bn; best_pair = pr == 1; }
    }
  }
  *bn_out = best_bn;
  *pair_out = best_pair;
}

int g_force_bn = 0;  // test hook (sg_gemm_set_tile)
int g_pair_mode = 2; // 0 = single-CTA tiles, 1 = CTA-pair (cta_group::2) tiles, 2 = choose per problem

}  // namespace sg

extern "C" int sg_gemm_set_tile(int bn) {
  if (bn != 0 && bn != 64 && bn != 128 && bn != 256) {
    sg::set_error("sg_gemm_set_tile: bn must be 0, 64, 128 or 256");
    return 2;
  }
  sg::g_force_bn = bn;
  return 0;
}

extern "C" int sg_gemm_set_pair(int mode) {
  if (mode < 0 || mode > 2) {
    sg::set_error("sg_gemm_set_pair: mode must be 0 (off), 1 (on) or 2 (auto)");
    return 2;
  }
  sg::g_pair_mode = mode;
  return 0;
}

extern "C" int sg_gemm(int dtype, const void* A, int a_mn_major, long long lda, const void* B,
                       int b_mn_major, long long ldb, int M, int N, int K,
                       const SgGemmEpilogue* ep, void* stream) {
  using namespace sg;
  SG_REQUIRE(dtype == SG_DTYPE_BF16 || dtype == SG_DTYPE_TF32, "sg_gemm: dtype must be SG_DTYPE_BF16 or SG_DTYPE_TF32");
  const int eb = dtype == SG_DTYPE_BF16 ? 2 : 4;
  const int BK = 128 / eb;          // elements per k-block
  const int MN_BOX = 128 / eb;      // elements per row of an MN-major box
  SG_REQUIRE(M > 0 && N > 0 && K > 0, "sg_gemm: empty problem %dx%dx%d", M, N, K);
  SG_REQUIRE(ep != nullptr && (ep->d32 != nullptr || ep->d16 != nullptr),
             "sg_gemm: epilogue needs an output");
  SG_REQUIRE(ep->grad_act == SG_ACT_NONE || ep->aux != nullptr, "sg_gemm: grad_act needs aux");
  // per-call tile choice (autotuner) > process-wide test hooks > built-in model
  int bn = ep->tile_bn ? ep->tile_bn : g_force_bn;
  SG_REQUIRE(bn == 0 || bn == 64 || bn == 128 || bn == 256, "sg_gemm: tile_bn must be 0, 64, 128 or 256");
  // CTA pairs need BN in {128, 256} (each CTA stages BN/2 columns of B in 64-column boxes)
  bool pair = ep->tile_pair == 2 || (ep->tile_pair == 0 && g_pair_mode == 1);
  const bool pair_free = ep->tile_pair == 0 && g_pair_mode == 2;
  if (bn == 0) {
    pick_tile(M, N, K, BK, !pair, pair || pair_free, &bn, &pair);
  } else if (pair_free) {
    pair = bn >= 128 && static_cast<long long>(M) * N >= 4096LL * 4096LL;
  }
  if (pair && bn < 128) bn = 128;
  CUtensorMap ta, tb;
  int rc;
  if (!a_mn_major) rc = make_tmap(&ta, eb, A, K, M, lda, BK, BM, false);
  else rc = make_tmap(&ta, eb, A, M, K, lda, MN_BOX, BK, true);
  if (rc) return rc;
  if (!b_mn_major) rc = make_tmap(&tb, eb, B, K, N, ldb, BK, pair ? bn / 2 : bn, false);
  else rc = make_tmap(&tb, eb, B, N, K, ldb, MN_BOX, BK, true);
  if (rc) return rc;
  GemmParams p;
  SG_REQUIRE(ep->ldd32 < (1LL << 31) && ep->ldd16 < (1LL << 31) && ep->ldaux < (1LL << 31),
             "sg_gemm: leading dimensions must fit in 31 bits");
  p.M = M; p.N = N; p.K = K;
  p.d32 = ep->d32; p.ldd32 = static_cast<int>(ep->ldd32);
  p.d16 = reinterpret_cast<__nv_bfloat16*>(ep->d16); p.ldd16 = static_cast<int>(ep->ldd16);
  p.bias = ep->bias; p.bias_bf16 = ep->bias_bf16 != 0;
  p.act = static_cast<unsigned char>(ep->act); p.grad_act = static_cast<unsigned char>(ep->grad_act);
  p.aux = ep->aux; p.ldaux = static_cast<int>(ep->ldaux); p.aux_bf16 = ep->aux_bf16 != 0;
  SG_REQUIRE(ep->perm_kk >= 0 && ep->perm_kk < 32768, "sg_gemm: perm_kk out of range");
  p.accumulate = ep->accumulate != 0; p.perm_c = ep->perm_c; p.perm_kk = static_cast<short>(ep->perm_kk); p.perm_valid = 0;
  SG_REQUIRE(ep->nwait >= 0 && ep->nwait <= 64, "sg_gemm: at most 64 wait flags");
  SG_REQUIRE(ep->nwait == 0 || (ep->wait_seq != nullptr && ep->wait_flags != nullptr),
             "sg_gemm: wait flags need the flag list and wait_seq");
  p.wait_flags = ep->wait_flags; p.nwait = static_cast<short>(ep->nwait); p.wait_seq = ep->wait_seq;
  p.split_k = 1; p.kb_per_split = (K + BK - 1) / BK;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  // split-K for plain fp32-output GEMMs (weight gradients) whose tiles cannot fill the machine: take the widest
  // tile (best bytes/flop) and cut the K loop into as many slices as there are idle SMs
  if (!pair && ep->d16 == nullptr && ep->bias == nullptr && ep->act == SG_ACT_NONE &&
      ep->grad_act == SG_ACT_NONE && ep->nwait == 0 && ep->split_k != 1) {
    const int num_kb = (K + BK - 1) / BK;
    int split = ep->split_k > 1 ? ep->split_k : 1;
    if (ep->split_k == 0 && num_kb >= 64 && !ep->tile_bn && !g_force_bn) {
      const int wide = N > 128 ? 256 : (N > 64 ? 128 : 64);
      const long long tiles_w = static_cast<long long>((M + BM - 1) / BM) * ((N + wide - 1) / wide);
      if (tiles_w * 2 <= num_sms()) {
        split = static_cast<int>(num_sms() / tiles_w);
        if (split > num_kb / 8) split = num_kb / 8;
        if (split > 64) split = 64;
        if (split > 1 && wide != bn) {
          bn = wide;   // the B tensor map depends on the tile width: rebuild it
          if (!b_mn_major) { rc = make_tmap(&tb, eb, B, K, N, ldb, BK, bn, false); if (rc) return rc; }
        }
      }
    }
    if (split > 1) {
      p.kb_per_split = (num_kb + split - 1) / split;
      p.split_k = (num_kb + p.kb_per_split - 1) / p.kb_per_split;
      if (!ep->accumulate) {
        if (check_cuda(cudaMemset2DAsync(ep->d32, ep->ldd32 * sizeof(float), 0, N * sizeof(float), M, s),
                       "sg_gemm: zero the split-K output"))
          return 1;
      }
      p.accumulate = 2;   // epilogue: atomic add of the partial tile
    }
  }
  if (eb == 2) return dispatch_tile<2>(pair, bn, a_mn_major, b_mn_major, ta, tb, p, s);
  return dispatch_tile<4>(pair, bn, a_mn_major, b_mn_major, ta, tb, p, s);
}

extern "C" int sg_gemm_bf16(const void* A, int a_mn_major, long long lda, const void* B, int b_mn_major,
                            long long ldb, int M, int N, int K, const SgGemmEpilogue* ep, void* stream) {
  return sg_gemm(SG_DTYPE_BF16, A, a_mn_major, lda, B, b_mn_major, ldb, M, N, K, ep, stream);
}

#ifdef SG_GEMM_TRACE
extern "C" int sg_gemm_trace_read(unsigned long long* out16) {
  return cudaMemcpyFromSymbol(out16, sg::g_gemm_trace, sizeof(unsigned long long) * 32) == cudaSuccess ? 0 : 1;
}
#endif
