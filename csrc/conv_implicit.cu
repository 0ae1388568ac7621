// Implicit-GEMM convolution on tcgen05: the patch matrix is never materialised — the TMA producer builds each
// operand tile straight from the channels-last activation with im2col-mode tensor maps
// (cp.async.bulk.tensor.4d...im2col), which walk the output pixels, apply the convolution stride, add the filter-tap
// offset given with the instruction and zero-fill the padding.  Reference op: src/worker/layer.cc:63-123 (per-image
// unpack_patch2col + sgemm loop, and the pack_col2patch data gradient).
//
// Two operand roles cover the three convolution GEMMs (all: fp32 accumulation in TMEM, bf16 or tf32 operands):
//
//   MODE_ROWS  acc[pixel][n] = sum_k patch(pixel, k) * B(n, k)
//              A = im2col tiles, K-major: 128 pixels x 128 bytes of K per k-block, built from 128/R boxes of
//              [128 pixels][R bytes] — one filter tap (ky,kx) and one channel chunk each; R = 16/32/64/128 bytes with the
//              matching K-major swizzle (none/32B/64B/128B), so any channel count that is a multiple of 16 bytes works
//              (C = 3 or 1 inputs are stored with 16-byte pixels).  B = packed filters [n][K], plain 2-D tiles.
//                forward:        patch over X,  B = W packed as [F][(ky,kx,c)]                      -> Y [pixels][F]
//                data gradient:  patch over dY with the taps flipped (stride 1), B = [C][(jy,jx,f)] -> dX [pixels][C]
//   MODE_COLS  acc[f][n] = sum_pixel dY[pixel][f] * patch(pixel, n)          (weight gradient, split-K over pixels)
//              A = dY [pixels][F], MN-major 2-D tiles;  B = im2col tiles, MN-major: per k-block BK pixels x 128 bytes of
//              channels per box, one box per (tap, 128-byte channel chunk); missing channels are zero-filled by TMA and
//              dropped by the epilogue's column map.
//
// Warp roles, barriers, TMEM double buffering and the fused epilogue are those of gemm_tcgen05.cu.
#include <cuda.h>
#include <stdlib.h>
#include <mutex>

#include "../include/singa_b200.h"
#include "common.h"
#include "ptx.cuh"
#include "gemm_common.cuh"

namespace sg {

struct ConvParams {
  int OW, OHW;             // GEMM-row pixel grid: pixels per row / per image
  int stride, pad_lo;      // base coordinate of pixel (oy, ox) in the source tensor: o * stride - pad_lo
  int ksize;               // filter taps per dimension
  int nchunk;              // channel chunks per tap
  int chunk_elems;         // elements per chunk
  int nsub;                // ksize * ksize * nchunk  (boxes that carry data; the rest of K / N is zero)
  int c_oob;               // a channel coordinate beyond the tensor: boxes loaded there are all zeros
  int R;                   // MODE_ROWS: bytes per box row (16/32/64/128)
  unsigned a_layout, a_sbo, a_lbo;   // MODE_ROWS: smem descriptor fields of the A boxes
};

// cp.async.bulk.tensor.4d im2col: coordinates {c, w, h, n} name the first pixel of the box in the source tensor's
// own coordinates (may be negative: padding), {off_w, off_h} the filter tap.
__device__ __forceinline__ void tma_load_im2col_4d(uint32_t dst, const void* tmap, uint32_t bar, int c, int w, int h,
                                                   int n, int off_w, int off_h) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(bar), "r"(c), "r"(w), "r"(h), "r"(n),
        "h"(static_cast<unsigned short>(off_w)), "h"(static_cast<unsigned short>(off_h))
      : "memory");
}

// The same load issued by either CTA of a pair (cta_group::2): the bytes land in the issuing CTA's shared memory, the
// transaction count goes to the LEADER's barrier (peer bit of the barrier address cleared).
__device__ __forceinline__ void tma_load_im2col_4d_pair(uint32_t dst, const void* tmap, uint32_t bar, int c, int w, int h,
                                                        int n, int off_w, int off_h) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.im2col.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(bar & kPeerBitMask), "r"(c), "r"(w), "r"(h), "r"(n),
        "h"(static_cast<unsigned short>(off_w)), "h"(static_cast<unsigned short>(off_h))
      : "memory");
}

constexpr int MODE_ROWS = 0, MODE_COLS = 1;

template <int EB, int BN, int MODE>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
conv_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                    const GemmParams p, const ConvParams cp) {
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
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const int m_tiles = (p.M + BM - 1) / BM;
  const int n_tiles = (p.N + BN - 1) / BN;
  const int num_tiles = m_tiles * n_tiles;
  const int num_kb = (p.K + BK - 1) / BK;
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
      mbar_init(tempty_bar(a), EPI_WARPS);
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

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    // the whole warp walks the loop (uniform registers for every coordinate), one elected lane issues
    {
      if (lane == 0 && p.nwait > 0) wait_param_versions(p.wait_flags, p.nwait, p.wait_seq);
      __syncwarp();
      uint32_t stage = 0, phase = 0;
      const int conv_oh = cp.OHW / cp.OW;
      for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
        const int slice = p.split_k == 1 ? 0 : item / num_tiles;
        const int tile = item - slice * num_tiles;
        const int kb_lo = slice * p.kb_per_split;
        const int kb_hi = min(num_kb, kb_lo + p.kb_per_split);
        const int m0 = (tile % m_tiles) * BM;
        const int n0 = (tile / m_tiles) * BN;
        // MODE_ROWS: the tile's first pixel, decomposed once per tile
        int img = 0, base_w = 0, base_h = 0;
        // MODE_COLS: this one thread issues 4 + BN/MN_BOX loads per k-block and the kernel was found to be bound by
        // that issue loop (ncu: the MMA warp starves on the full barrier while the producer never waits), so everything
        // that does not change with the k-block is decided here: the (channel, tap) of each of the tile's boxes, and a
        // pixel cursor (image, row, column) that is advanced by BK instead of being divided out again.
        constexpr int NB = BN / MN_BOX;
        int box_c[NB];
        unsigned short box_kx[NB], box_ky[NB];
        int pimg = 0, poy = 0, pox = 0;
        int rq = 0, rch = 0, rkx = 0, rky = 0;      // MODE_ROWS: cursor over the (tap, channel chunk) boxes of K
        if (MODE == MODE_ROWS) {
          rq = kb_lo * (128 / cp.R);
          const int tap = rq / cp.nchunk;
          rch = rq - tap * cp.nchunk;
          rky = tap / cp.ksize;
          rkx = tap - rky * cp.ksize;
          img = m0 / cp.OHW;
          const int rem = m0 - img * cp.OHW;
          const int oy = rem / cp.OW;
          base_h = oy * cp.stride - cp.pad_lo;
          base_w = (rem - oy * cp.OW) * cp.stride - cp.pad_lo;
        } else {
          int q = n0 / MN_BOX;
          int tap = q / cp.nchunk;
          int ch = q - tap * cp.nchunk;
          int ky = tap / cp.ksize;
          int kx = tap - ky * cp.ksize;
#pragma unroll
          for (int b = 0; b < NB; ++b, ++q) {
            const bool live = q < cp.nsub;
            box_c[b] = live ? ch * cp.chunk_elems : cp.c_oob;
            box_kx[b] = static_cast<unsigned short>(live ? kx : 0);
            box_ky[b] = static_cast<unsigned short>(live ? ky : 0);
            if (++ch == cp.nchunk) { ch = 0; if (++kx == cp.ksize) { kx = 0; ++ky; } }
          }
          const int k0 = kb_lo * BK;
          pimg = k0 / cp.OHW;
          const int rem = k0 - pimg * cp.OHW;
          poy = rem / cp.OW;
          pox = rem - poy * cp.OW;
        }
        for (int kb = kb_lo; kb < kb_hi; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          const uint32_t fb = full_bar(stage);
          const bool issuer = elect_one();
          if (issuer) mbar_arrive_expect_tx(fb, Cfg::STAGE_BYTES);
          const uint32_t sa = sA0 + stage * A_STAGE_BYTES;
          const uint32_t sb = sB0 + stage * B_STAGE_BYTES;
          if (MODE == MODE_ROWS) {
            // A: 128/R boxes of [128 pixels][R bytes]; box q of the K dimension = (tap, channel chunk)
            const int per_kb = 128 / cp.R;
            const uint32_t box_bytes = BM * cp.R;
            for (int i = 0; i < per_kb; ++i, ++rq) {
              const bool live = rq < cp.nsub;
              if (issuer)
                tma_load_im2col_4d(sa + i * box_bytes, &tma_a, fb, live ? rch * cp.chunk_elems : cp.c_oob, base_w, base_h,
                                   img, live ? rkx : 0, live ? rky : 0);
              if (++rch == cp.nchunk) { rch = 0; if (++rkx == cp.ksize) { rkx = 0; ++rky; } }
            }
            if (issuer) tma_load_2d(sb, &tma_b, fb, kb * BK, n0);
          } else {
            // A = dY, MN-major 2-D boxes; B: one im2col box [BK pixels][128 bytes] per (tap, channel chunk)
            const int k0 = kb * BK;
            const int bh = poy * cp.stride - cp.pad_lo;
            const int bw = pox * cp.stride - cp.pad_lo;
            if (issuer) {
#pragma unroll
              for (int b = 0; b < BM / MN_BOX; ++b) tma_load_2d(sa + b * MN_BOX_BYTES, &tma_a, fb, m0 + b * MN_BOX, k0);
#pragma unroll
              for (int b = 0; b < NB; ++b)
                tma_load_im2col_4d(sb + b * MN_BOX_BYTES, &tma_b, fb, box_c[b], bw, bh, pimg, box_kx[b], box_ky[b]);
            }
            pox += BK;
            while (pox >= cp.OW) {
              pox -= cp.OW;
              if (++poy == conv_oh) { poy = 0; ++pimg; }
            }
          }
          __syncwarp();
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    constexpr uint32_t idesc = make_idesc(EB == 2 ? 1u : 2u, BM, BN, MODE == MODE_COLS ? 1 : 0, MODE == MODE_COLS ? 1 : 0);
    constexpr uint32_t MN_KSTEP = UMMA_K * 128;
    constexpr uint32_t MN_LAYOUT = EB == 4 ? kLayoutSw128Base32 : kLayoutSw128;
    constexpr uint32_t MN_SBO = EB == 4 ? 512 : 1024;
    // MODE_ROWS: descriptor of the A boxes without the address, and the 16-byte offset of each MMA's 32 bytes of K
    // (inside one box when R >= 32, two 16-byte boxes LBO apart when R == 16) — once, not per k-block
    const uint64_t a_tmpl = make_smem_desc(0, cp.a_lbo, cp.a_sbo, cp.a_layout);
    uint32_t a_off16[BK / UMMA_K];
#pragma unroll
    for (int k = 0; k < BK / UMMA_K; ++k) {
      const uint32_t byte = 32u * k;
      a_off16[k] = (cp.R >= 32 ? (byte / cp.R) * (BM * cp.R) + (byte % cp.R) : (2u * k) * (BM * 16u)) >> 4;
    }
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
        const uint32_t sa = sA0 + stage * A_STAGE_BYTES;
        const uint32_t sb = sB0 + stage * B_STAGE_BYTES;
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            uint64_t adesc, bdesc;
            if (MODE == MODE_ROWS) {
              adesc = a_tmpl + ((sa >> 4) + a_off16[k]);       // the address field (14 bits) cannot carry: smem < 256 KB
              bdesc = make_smem_desc_sw128(sb + 32u * k, 16, 1024);
            } else {
              adesc = make_smem_desc(sa + k * MN_KSTEP, MN_BOX_BYTES, MN_SBO, MN_LAYOUT);
              bdesc = make_smem_desc(sb + k * MN_KSTEP, MN_BOX_BYTES, MN_SBO, MN_LAYOUT);
            }
            umma<EB>(tmem_d, adesc, bdesc, idesc, (kb > kb_lo || k != 0) ? 1u : 0u);
          }
          umma_commit(empty_bar(stage));
          if (kb == kb_hi - 1) umma_commit(tfull_bar(acc));
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 2..9)
    const int quad = warp & 3;
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
          if (lane == 0) mbar_arrive(tempty_bar(acc));
        }
        epilogue_chunk(p, r, row0, lane, n0 + c, fast_ok, stg);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, Cfg::TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------
// Both modes on CTA pairs (cta_group::2): one 256 x BN tile per pair.  Each CTA stages the A tile of ITS 128 rows and
// HALF of the B tile, so a k-block costs 16 KB + BN*64 B per CTA instead of 16 KB + BN*128 B — these kernels sit on
// the L2 -> shared-memory limit (ncu, AlexNet conv2 forward: 11.7 TB/s, tensor pipe 51 %) and the B tile is two
// thirds of the traffic at BN = 256.
//   MODE_ROWS: rows = pixels; this CTA's 128 pixels of patches + half of the filter rows.
//   MODE_COLS: rows = filters (dY^T); this CTA's 128 filters + half of the patch columns; split-K over the pixels.
// Barrier protocol as in gemm_tcgen05_pair_kernel (both CTAs' loads count on the leader's full barrier, the leader's
// MMA warp commits to both CTAs' empty / accumulator-full barriers).
// ------------------------------------------------------------------------------------------------
template <int EB, int BN, int MODE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(GEMM_THREADS, 1)
conv_tcgen05_pair_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                         const GemmParams p, const ConvParams cp) {
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
  const int num_items = num_tiles * p.split_k;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 2);     // leader's own arrive.expect_tx + the peer's remote arrive
      mbar_init(empty_bar(s), 1);    // multicast commit from the leader's MMA warp
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), 2 * EPI_WARPS);
    }
    fence_mbar_init();
  }
  cluster_sync_all();
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
    // the whole warp walks the loop (uniform registers for every coordinate), one elected lane issues
    {
      if (lane == 0 && p.nwait > 0) wait_param_versions(p.wait_flags, p.nwait, p.wait_seq);
      __syncwarp();
      uint32_t stage = 0, phase = 0;
      const int conv_oh = cp.OHW / cp.OW;
      for (int item = pair_id; item < num_items; item += num_pairs) {
        const int slice = p.split_k == 1 ? 0 : item / num_tiles;
        const int tile = item - slice * num_tiles;
        const int kb_lo = slice * p.kb_per_split;
        const int kb_hi = min(num_kb, kb_lo + p.kb_per_split);
        const int m0 = (tile % m_tiles) * (2 * BM) + static_cast<int>(rank) * BM;   // this CTA's 128 rows
        const int n0 = (tile / m_tiles) * BN + static_cast<int>(rank) * BNH;        // this CTA's half of B
        // MODE_ROWS: first pixel of the CTA's rows and a cursor over the (tap, channel chunk) boxes of K
        int img = 0, base_w = 0, base_h = 0, rq = 0, rch = 0, rkx = 0, rky = 0;
        // MODE_COLS: (channel, tap) of each of the CTA's boxes and a pixel cursor advanced by BK per k-block
        constexpr int NB = BNH / MN_BOX > 0 ? BNH / MN_BOX : 1;
        int box_c[NB];
        unsigned short box_kx[NB], box_ky[NB];
        int pimg = 0, poy = 0, pox = 0;
        if (MODE == MODE_ROWS) {
          rq = kb_lo * (128 / cp.R);
          const int tap = rq / cp.nchunk;
          rch = rq - tap * cp.nchunk;
          rky = tap / cp.ksize;
          rkx = tap - rky * cp.ksize;
          img = m0 / cp.OHW;
          const int rem = m0 - img * cp.OHW;
          const int oy = rem / cp.OW;
          base_h = oy * cp.stride - cp.pad_lo;
          base_w = (rem - oy * cp.OW) * cp.stride - cp.pad_lo;
        } else {
          int q = n0 / MN_BOX;
          int tap = q / cp.nchunk;
          int ch = q - tap * cp.nchunk;
          int ky = tap / cp.ksize;
          int kx = tap - ky * cp.ksize;
#pragma unroll
          for (int b = 0; b < NB; ++b, ++q) {
            const bool live = q < cp.nsub;
            box_c[b] = live ? ch * cp.chunk_elems : cp.c_oob;
            box_kx[b] = static_cast<unsigned short>(live ? kx : 0);
            box_ky[b] = static_cast<unsigned short>(live ? ky : 0);
            if (++ch == cp.nchunk) { ch = 0; if (++kx == cp.ksize) { kx = 0; ++ky; } }
          }
          const int k0 = kb_lo * BK;
          pimg = k0 / cp.OHW;
          const int rem = k0 - pimg * cp.OHW;
          poy = rem / cp.OW;
          pox = rem - poy * cp.OW;
        }
        for (int kb = kb_lo; kb < kb_hi; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          const uint32_t fb = full_bar(stage);
          const bool issuer = elect_one();
          if (issuer) {
            if (leader) mbar_arrive_expect_tx(fb, 2 * Cfg::STAGE_BYTES);
            else mbar_arrive_remote(fb, 0);
          }
          const uint32_t sa = sA0 + stage * A_STAGE_BYTES;
          const uint32_t sb = sB0 + stage * B_HALF_BYTES;
          if (MODE == MODE_ROWS) {
            const int per_kb = 128 / cp.R;
            const uint32_t box_bytes = BM * cp.R;
            for (int i = 0; i < per_kb; ++i, ++rq) {
              const bool live = rq < cp.nsub;
              if (issuer)
                tma_load_im2col_4d_pair(sa + i * box_bytes, &tma_a, fb, live ? rch * cp.chunk_elems : cp.c_oob, base_w,
                                        base_h, img, live ? rkx : 0, live ? rky : 0);
              if (++rch == cp.nchunk) { rch = 0; if (++rkx == cp.ksize) { rkx = 0; ++rky; } }
            }
            if (issuer) tma_load_2d_pair(sb, &tma_b, fb, kb * BK, n0);
          } else {
            const int k0 = kb * BK;
            const int bh = poy * cp.stride - cp.pad_lo;
            const int bw = pox * cp.stride - cp.pad_lo;
            if (issuer) {
#pragma unroll
              for (int b = 0; b < BM / MN_BOX; ++b) tma_load_2d_pair(sa + b * MN_BOX_BYTES, &tma_a, fb, m0 + b * MN_BOX, k0);
#pragma unroll
              for (int b = 0; b < NB; ++b)
                tma_load_im2col_4d_pair(sb + b * MN_BOX_BYTES, &tma_b, fb, box_c[b], bw, bh, pimg, box_kx[b], box_ky[b]);
            }
            pox += BK;
            while (pox >= cp.OW) {
              pox -= cp.OW;
              if (++poy == conv_oh) { poy = 0; ++pimg; }
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
      constexpr uint32_t idesc = make_idesc(EB == 2 ? 1u : 2u, 2 * BM, BN, MODE == MODE_COLS ? 1 : 0, MODE == MODE_COLS ? 1 : 0);
      constexpr uint32_t MN_KSTEP = UMMA_K * 128;
      constexpr uint32_t MN_LAYOUT = EB == 4 ? kLayoutSw128Base32 : kLayoutSw128;
      constexpr uint32_t MN_SBO = EB == 4 ? 512 : 1024;
      const uint64_t a_tmpl = make_smem_desc(0, cp.a_lbo, cp.a_sbo, cp.a_layout);
      uint32_t a_off16[BK / UMMA_K];
#pragma unroll
      for (int k = 0; k < BK / UMMA_K; ++k) {
        const uint32_t byte = 32u * k;
        a_off16[k] = (cp.R >= 32 ? (byte / cp.R) * (BM * cp.R) + (byte % cp.R) : (2u * k) * (BM * 16u)) >> 4;
      }
      uint32_t stage = 0, phase = 0;
      int it = 0;
      for (int item = pair_id; item < num_items; item += num_pairs, ++it) {
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
          const uint32_t sa = sA0 + stage * A_STAGE_BYTES;
          const uint32_t sb = sB0 + stage * B_HALF_BYTES;
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) {
              uint64_t adesc, bdesc;
              if (MODE == MODE_ROWS) {
                adesc = a_tmpl + ((sa >> 4) + a_off16[k]);
                bdesc = make_smem_desc_sw128(sb + 32u * k, 16, 1024);
              } else {
                adesc = make_smem_desc(sa + k * MN_KSTEP, MN_BOX_BYTES, MN_SBO, MN_LAYOUT);
                bdesc = make_smem_desc(sb + k * MN_KSTEP, MN_BOX_BYTES, MN_SBO, MN_LAYOUT);
              }
              umma_pair<EB>(tmem_d, adesc, bdesc, idesc, (kb > kb_lo || k != 0) ? 1u : 0u);
            }
            umma_commit_pair(empty_bar(stage), 3);
            if (kb == kb_hi - 1) umma_commit_pair(tfull_bar(acc), 3);
          }
          __syncwarp();
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 2..9, both CTAs)
    const int quad = warp & 3;
    float* stg = reinterpret_cast<float*>(smem_raw + (bar0 - smem_u32(smem_raw)) + 256) + (warp - 2) * EPI_WARP_FLOATS;
    const bool fast_ok = epilogue_fast_ok(p);
    int it = 0;
    for (int item = pair_id; item < num_items; item += num_pairs, ++it) {
      const int tile = p.split_k == 1 ? item : item % num_tiles;
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
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, Cfg::TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------
// Host side
// ------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*,
                                   CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                   CUtensorMapFloatOOBfill);

static void* driver_fn(const char* name) {
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint(name, &sym, cudaEnableDefault, &qres) == cudaSuccess && qres == cudaDriverEntryPointSuccess)
    return sym;
  return nullptr;
}

static int make_tmap_2d(CUtensorMap* out, int eb, const void* ptr, long long inner, long long outer, long long ld,
                        int box_inner, int box_outer, bool mn_major) {
  static EncodeTiledFn enc = reinterpret_cast<EncodeTiledFn>(driver_fn("cuTensorMapEncodeTiled"));
  SG_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
  SG_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "conv operand base must be 16-byte aligned");
  SG_REQUIRE((ld * eb) % 16 == 0, "conv operand leading dimension (%lld elements of %d bytes) must be a multiple of 16 bytes", ld, eb);
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(inner), static_cast<cuuint64_t>(outer)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * eb};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(box_inner), static_cast<cuuint32_t>(box_outer)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, eb == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                   const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   (mn_major && eb == 4) ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  SG_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d) inner=%lld outer=%lld ld=%lld", static_cast<int>(r),
             inner, outer, ld);
  return 0;
}

// im2col map over a channels-last tensor [N][H][W][C] whose pixels are `ldc` elements apart.  The box is
// [pixels][box_c channels]; pixels advance by `stride` over the positions o*stride - pad_lo, o in [0, On).
static int make_tmap_im2col(CUtensorMap* out, int eb, const void* ptr, int N, int H, int W, int C, long long ldc,
                            int OH, int OW, int stride, int pad_lo, int box_c, int box_pixels, CUtensorMapSwizzle swz) {
  static EncodeIm2colFn enc = reinterpret_cast<EncodeIm2colFn>(driver_fn("cuTensorMapEncodeIm2col"));
  SG_REQUIRE(enc != nullptr, "cuTensorMapEncodeIm2col unavailable (no CUDA driver?)");
  SG_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "conv activation base must be 16-byte aligned");
  SG_REQUIRE((ldc * eb) % 16 == 0, "conv activation pixel stride (%lld elements of %d bytes) must be a multiple of 16 bytes",
             ldc, eb);
  cuuint64_t dims[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H),
                        static_cast<cuuint64_t>(N)};
  cuuint64_t strides[3] = {static_cast<cuuint64_t>(ldc) * eb, static_cast<cuuint64_t>(ldc) * eb * W,
                           static_cast<cuuint64_t>(ldc) * eb * W * H};
  // bounding box of the base pixels: first = -pad_lo, last = (On - 1) * stride - pad_lo = dim - 1 + upper
  int lower[2] = {-pad_lo, -pad_lo};
  int upper[2] = {(OW - 1) * stride - pad_lo - (W - 1), (OH - 1) * stride - pad_lo - (H - 1)};
  SG_REQUIRE(lower[0] >= -128 && upper[0] >= -128 && upper[1] >= -128 && upper[0] <= 127 && upper[1] <= 127,
             "convolution geometry outside the im2col tensor-map corner range [-128, 127]");
  SG_REQUIRE(stride >= 1 && stride <= 8, "convolution stride %d outside the im2col tensor-map range [1, 8]", stride);
  cuuint32_t estr[4] = {1, static_cast<cuuint32_t>(stride), static_cast<cuuint32_t>(stride), 1};
  CUresult r = enc(out, eb == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4,
                   const_cast<void*>(ptr), dims, strides, lower, upper, static_cast<cuuint32_t>(box_c),
                   static_cast<cuuint32_t>(box_pixels), estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  SG_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeIm2col failed (%d): N=%d H=%d W=%d C=%d ldc=%lld stride=%d pad=%d box=%dx%d",
             static_cast<int>(r), N, H, W, C, ldc, stride, pad_lo, box_c, box_pixels);
  return 0;
}

template <int EB, int BN, int MODE>
static int launch_conv(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, const ConvParams& cp,
                       cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  auto kern = conv_tcgen05_kernel<EB, BN, MODE>;
  static thread_local int attr_dev = -1;
  int dev = 0;
  cudaGetDevice(&dev);
  if (attr_dev != dev) {
    if (check_cuda(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES),
                   "cudaFuncSetAttribute(conv smem)"))
      return 1;
    attr_dev = dev;
  }
  const int m_tiles = (p.M + BM - 1) / BM;
  const int n_tiles = (p.N + BN - 1) / BN;
  const long long tiles = static_cast<long long>(m_tiles) * n_tiles * p.split_k;
  const int grid = tiles < num_sms() ? static_cast<int>(tiles) : num_sms();
  kern<<<grid, GEMM_THREADS, Cfg::SMEM_BYTES, stream>>>(ta, tb, p, cp);
  SG_CHECK_LAUNCH("conv_tcgen05");
  return 0;
}

template <int EB, int BN, int MODE>
static int launch_conv_pair(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, const ConvParams& cp,
                            cudaStream_t stream) {
  using Cfg = PairCfg<BN>;
  auto kern = conv_tcgen05_pair_kernel<EB, BN, MODE>;
  static thread_local int attr_dev = -1;
  int dev = 0;
  cudaGetDevice(&dev);
  if (attr_dev != dev) {
    if (check_cuda(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES),
                   "cudaFuncSetAttribute(conv pair smem)"))
      return 1;
    attr_dev = dev;
  }
  const long long tiles = static_cast<long long>((p.M + 2 * BM - 1) / (2 * BM)) * ((p.N + BN - 1) / BN) * p.split_k;
  const int max_pairs = num_sms() / 2;
  const int pairs = tiles < max_pairs ? static_cast<int>(tiles) : max_pairs;
  kern<<<2 * pairs, GEMM_THREADS, Cfg::SMEM_BYTES, stream>>>(ta, tb, p, cp);
  SG_CHECK_LAUNCH("conv_tcgen05_pair");
  return 0;
}

template <int EB, int MODE>
static int dispatch_conv(int bn, const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p,
                         const ConvParams& cp, cudaStream_t s) {
  switch (bn) {
    case 256: return launch_conv<EB, 256, MODE>(ta, tb, p, cp, s);
    case 128: return launch_conv<EB, 128, MODE>(ta, tb, p, cp, s);
    default: return launch_conv<EB, 64, MODE>(ta, tb, p, cp, s);
  }
}

static void fill_epilogue(GemmParams& p, const SgGemmEpilogue* ep) {
  p.d32 = ep->d32; p.ldd32 = static_cast<int>(ep->ldd32);
  p.d16 = reinterpret_cast<__nv_bfloat16*>(ep->d16); p.ldd16 = static_cast<int>(ep->ldd16);
  p.bias = ep->bias; p.bias_bf16 = ep->bias_bf16 != 0;
  p.act = static_cast<unsigned char>(ep->act); p.grad_act = static_cast<unsigned char>(ep->grad_act);
  p.aux = ep->aux; p.ldaux = static_cast<int>(ep->ldaux); p.aux_bf16 = ep->aux_bf16 != 0;
  p.accumulate = ep->accumulate != 0; p.perm_c = ep->perm_c; p.perm_kk = static_cast<short>(ep->perm_kk);
  p.perm_valid = 0;
  p.wait_flags = ep->wait_flags; p.nwait = static_cast<short>(ep->nwait); p.wait_seq = ep->wait_seq;
}

}  // namespace sg

using namespace sg;

extern "C" int sg_conv_chunk_bytes(int channels, int elem_bytes) {
  // widest chunk (fewest TMA boxes, widest swizzle) whose zero padding stays under ~15 % of the channels
  const int bytes = channels * elem_bytes;
  const int cands[4] = {128, 64, 32, 16};
  for (int i = 0; i < 4; ++i) {
    const int padded = (bytes + cands[i] - 1) / cands[i] * cands[i];
    if (padded * 100 <= bytes * 115) return cands[i];
  }
  return 16;
}

extern "C" int sg_conv_rows(int dtype, const void* src, const SgConvGeom* g, const void* filt, long long ldf, int ncols,
                            const SgGemmEpilogue* ep, void* stream) {
  SG_REQUIRE(dtype == SG_DTYPE_BF16 || dtype == SG_DTYPE_TF32, "sg_conv_rows: dtype must be SG_DTYPE_BF16 or SG_DTYPE_TF32");
  SG_REQUIRE(g != nullptr && ep != nullptr && (ep->d32 != nullptr || ep->d16 != nullptr), "sg_conv_rows: null arguments");
  const int eb = dtype == SG_DTYPE_BF16 ? 2 : 4;
  const int BK = 128 / eb;
  const int R = g->chunk_bytes;
  SG_REQUIRE(R == 16 || R == 32 || R == 64 || R == 128, "sg_conv_rows: chunk_bytes must be 16, 32, 64 or 128");
  SG_REQUIRE(g->k >= 1 && g->OH >= 1 && g->OW >= 1 && ncols >= 1, "sg_conv_rows: empty problem");
  const int chunk_elems = R / eb;
  const int nchunk = (g->C + chunk_elems - 1) / chunk_elems;
  const int nsub = g->k * g->k * nchunk;
  const long long Kp = (static_cast<long long>(nsub) * chunk_elems + BK - 1) / BK * BK;   // whole k-blocks
  SG_REQUIRE(ldf >= Kp, "sg_conv_rows: packed filter rows hold %lld elements, the K dimension needs %lld", ldf, Kp);
  const long long P = static_cast<long long>(g->N) * g->OH * g->OW;
  SG_REQUIRE(P < (1LL << 31), "sg_conv_rows: too many output pixels");
  CUtensorMap ta, tb;
  const CUtensorMapSwizzle swz = R == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : R == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                               : R == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE;
  if (int rc = make_tmap_im2col(&ta, eb, src, g->N, g->H, g->W, g->C, g->ldc, g->OH, g->OW, g->stride, g->pad_lo,
                                chunk_elems, BM, swz))
    return rc;
  int bn = ep->tile_bn ? ep->tile_bn : (ncols > 128 ? 256 : (ncols > 64 ? 128 : 64));
  SG_REQUIRE(bn == 64 || bn == 128 || bn == 256, "sg_conv_rows: tile_bn must be 0, 64, 128 or 256");
  // CTA pairs: 256-column filter tiles (measured: 1-5 % faster than single CTAs on AlexNet conv2-5 forward / data
  // gradient; with 128-column tiles, e.g. the 96-channel data gradient of conv2, 10 % slower) and enough 256-pixel tiles
  // for every pair of SMs; ep->tile_pair 1 / 2 and SINGA_B200_CONV_PAIR=0 / 1 force the choice
  static const int env_pair = [] { const char* e = getenv("SINGA_B200_CONV_PAIR"); return e == nullptr ? -1 : (e[0] == '1' ? 1 : 0); }();
  bool pair = R >= 32 && bn == 256 && (P + 2 * BM - 1) / (2 * BM) * ((ncols + bn - 1) / bn) >= num_sms() / 2;
  if (env_pair >= 0) pair = pair && env_pair == 1;
  if (ep->tile_pair == 1) pair = false;
  if (ep->tile_pair == 2) pair = bn >= 128;
  if (int rc = make_tmap_2d(&tb, eb, filt, Kp, ncols, ldf, BK, pair ? bn / 2 : bn, false)) return rc;
  GemmParams p;
  fill_epilogue(p, ep);
  p.M = static_cast<int>(P); p.N = ncols; p.K = static_cast<int>(Kp);
  p.split_k = 1; p.kb_per_split = static_cast<int>(Kp / BK);
  SG_REQUIRE(ep->nwait >= 0 && ep->nwait <= 64, "sg_conv_rows: at most 64 wait flags");
  ConvParams cp;
  cp.OW = g->OW; cp.OHW = g->OH * g->OW; cp.stride = g->stride; cp.pad_lo = g->pad_lo; cp.ksize = g->k;
  cp.nchunk = nchunk; cp.chunk_elems = chunk_elems; cp.nsub = nsub; cp.c_oob = nchunk * chunk_elems + chunk_elems;
  cp.R = R;
  cp.a_layout = R == 128 ? 2u : R == 64 ? 4u : R == 32 ? 6u : 0u;
  cp.a_sbo = 8u * R;                       // eight rows of a box
  cp.a_lbo = R == 16 ? BM * 16u : 16u;     // no swizzle: the second 16-byte half of an MMA's K lives in the next box
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (pair) {
    if (eb == 2)
      return bn == 256 ? launch_conv_pair<2, 256, MODE_ROWS>(ta, tb, p, cp, s) : launch_conv_pair<2, 128, MODE_ROWS>(ta, tb, p, cp, s);
    return bn == 256 ? launch_conv_pair<4, 256, MODE_ROWS>(ta, tb, p, cp, s) : launch_conv_pair<4, 128, MODE_ROWS>(ta, tb, p, cp, s);
  }
  if (eb == 2) return dispatch_conv<2, MODE_ROWS>(bn, ta, tb, p, cp, s);
  return dispatch_conv<4, MODE_ROWS>(bn, ta, tb, p, cp, s);
}

extern "C" int sg_conv_wgrad(int dtype, const void* src, const SgConvGeom* g, const void* dy, long long lddy, int F,
                             float* dw, long long lddw, int accumulate, void* stream) {
  SG_REQUIRE(dtype == SG_DTYPE_BF16 || dtype == SG_DTYPE_TF32, "sg_conv_wgrad: dtype must be SG_DTYPE_BF16 or SG_DTYPE_TF32");
  SG_REQUIRE(g != nullptr && dw != nullptr && dy != nullptr, "sg_conv_wgrad: null arguments");
  const int eb = dtype == SG_DTYPE_BF16 ? 2 : 4;
  const int BK = 128 / eb, MN_BOX = 128 / eb;
  const int nchunk = (g->C + MN_BOX - 1) / MN_BOX;       // 128-byte channel chunks
  const int nsub = g->k * g->k * nchunk;
  const long long Np = static_cast<long long>(nsub) * MN_BOX;
  const long long P = static_cast<long long>(g->N) * g->OH * g->OW;
  SG_REQUIRE(P < (1LL << 31) && Np < (1LL << 31), "sg_conv_wgrad: problem too large");
  CUtensorMap ta, tb;
  if (int rc = make_tmap_2d(&ta, eb, dy, F, P, lddy, MN_BOX, BK, true)) return rc;
  if (int rc = make_tmap_im2col(&tb, eb, src, g->N, g->H, g->W, g->C, g->ldc, g->OH, g->OW, g->stride, g->pad_lo, MN_BOX,
                                BK, eb == 4 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B))
    return rc;
  const int bn = Np > 128 ? 256 : (Np > 64 ? 128 : 64);
  GemmParams p;
  SgGemmEpilogue ep = {};
  ep.d32 = dw; ep.ldd32 = lddw;
  fill_epilogue(p, &ep);
  p.M = F; p.N = static_cast<int>(Np); p.K = static_cast<int>(P);
  // column n = (tap * nchunk*MN_BOX + c) -> stored at c * k*k + tap when c < C (the reference's {F, C*k*k} layout)
  SG_REQUIRE(g->C < 32768, "sg_conv_wgrad: too many channels");
  p.perm_c = nchunk * MN_BOX; p.perm_kk = static_cast<short>(g->k * g->k); p.perm_valid = static_cast<short>(g->C);
  const int num_kb = static_cast<int>((P + BK - 1) / BK);
  // CTA pairs when the filters fill whole 256-row tiles (each CTA then stages half of the patch tile): 2-5 % faster on
  // the 256-filter layers of AlexNet; 384 filters would leave a quarter of the second pair tile empty (measured 10-25 %
  // slower than single CTAs)
  static const int env_pair = [] { const char* e = getenv("SINGA_B200_CONV_PAIR"); return e == nullptr ? -1 : (e[0] == '1' ? 1 : 0); }();
  const bool pair = F % (2 * BM) == 0 && bn >= 128 && env_pair != 0;
  const long long tiles = static_cast<long long>((F + (pair ? 2 : 1) * BM - 1) / ((pair ? 2 : 1) * BM)) * ((Np + bn - 1) / bn);
  int split = static_cast<int>((pair ? num_sms() / 2 : num_sms()) / (tiles > 0 ? tiles : 1));
  if (split > num_kb / 8) split = num_kb / 8;
  if (split > 64) split = 64;
  if (split < 1) split = 1;
  p.kb_per_split = (num_kb + split - 1) / split;
  p.split_k = static_cast<short>((num_kb + p.kb_per_split - 1) / p.kb_per_split);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (p.split_k > 1) {
    if (!accumulate &&
        check_cuda(cudaMemset2DAsync(dw, lddw * sizeof(float), 0, static_cast<size_t>(g->C) * g->k * g->k * sizeof(float), F, s),
                   "sg_conv_wgrad: zero the split-K output"))
      return 1;
    p.accumulate = 2;
  } else {
    p.accumulate = accumulate ? 1 : 0;
  }
  ConvParams cp;
  cp.OW = g->OW; cp.OHW = g->OH * g->OW; cp.stride = g->stride; cp.pad_lo = g->pad_lo; cp.ksize = g->k;
  cp.nchunk = nchunk; cp.chunk_elems = MN_BOX; cp.nsub = nsub; cp.c_oob = nchunk * MN_BOX + MN_BOX;
  cp.R = 128; cp.a_layout = 0; cp.a_sbo = 0; cp.a_lbo = 0;
  if (pair) {
    if (eb == 2)
      return bn == 256 ? launch_conv_pair<2, 256, MODE_COLS>(ta, tb, p, cp, s) : launch_conv_pair<2, 128, MODE_COLS>(ta, tb, p, cp, s);
    return bn == 256 ? launch_conv_pair<4, 256, MODE_COLS>(ta, tb, p, cp, s) : launch_conv_pair<4, 128, MODE_COLS>(ta, tb, p, cp, s);
  }
  if (eb == 2) return dispatch_conv<2, MODE_COLS>(bn, ta, tb, p, cp, s);
  return dispatch_conv<4, MODE_COLS>(bn, ta, tb, p, cp, s);
}
