// Persistent, warp-specialised tcgen05 GEMM for sm_100a.
//
//   acc[m][n] = sum_k A(m,k) * B(n,k)    (bf16 x bf16 -> fp32 in TMEM)
//
// Roles inside one 192-thread CTA (one CTA per SM, static round-robin tile schedule):
//   warp 0      TMA producer: cp.async.bulk.tensor tiles of A and B into a STAGES-deep smem ring
//   warp 1      TMEM owner + MMA issuer: one lane issues tcgen05.mma (M=128, N=BN, K=16) x4 per stage,
//               tcgen05.commit releases the smem slot and, after the last k-block, publishes the accumulator
//   warps 2..5  epilogue: tcgen05.ld the 128 x BN fp32 accumulator (one TMEM lane = one output row per
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
#include <mutex>
#include <unordered_map>
#include <string>

#include "../include/singa_b200.h"
#include "common.h"
#include "ptx.cuh"

namespace sg {

constexpr int BM = 128;
constexpr int BK = 64;                       // 64 bf16 = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int A_STAGE_BYTES = BM * BK * 2;   // 16 KB
constexpr int GEMM_THREADS = 192;
constexpr int SMEM_BUDGET = 200 * 1024;

template <int BN> struct GemmCfg {
  static constexpr int B_STAGE_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
  static constexpr int STAGES = (SMEM_BUDGET / STAGE_BYTES) > 8 ? 8 : (SMEM_BUDGET / STAGE_BYTES);
  static constexpr int TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;  // power of two for BN in {64,128,256}
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
};

struct GemmParams {
  int M, N, K;
  float* d32; long long ldd32;
  __nv_bfloat16* d16; long long ldd16;
  const void* bias; int bias_bf16;
  int act, grad_act;
  const void* aux; long long ldaux; int aux_bf16;
  int accumulate;
  int perm_c, perm_kk;
  const unsigned long long* wait_flags[8];
  int nwait;
  const unsigned long long* wait_seq;
};

__device__ __forceinline__ float tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// Scaled tanh of the reference (include/mshadow/cxxnet_op.h:77-87): y = 1.7159047 tanh(0.66666667 x).
constexpr float kStanhA = 1.7159047f;
constexpr float kStanhB = 0.66666667f;

__device__ __forceinline__ float apply_act(float v, int act) {
  if (act == SG_ACT_RELU) return fmaxf(v, 0.f);
  if (act == SG_ACT_STANH) return kStanhA * tanh_fast(kStanhB * v);
  return v;
}
__device__ __forceinline__ float act_grad_from_output(float y, int act) {
  if (act == SG_ACT_RELU) return y > 0.f ? 1.f : 0.f;
  if (act == SG_ACT_STANH) return kStanhA * kStanhB - (kStanhB / kStanhA) * y * y;
  return 1.f;
}

template <int BN, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a,
                         const __grid_constant__ CUtensorMap tma_b, const GemmParams p) {
  using Cfg = GemmCfg<BN>;
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

  const int m_tiles = (p.M + BM - 1) / BM;
  const int n_tiles = (p.N + BN - 1) / BN;
  const int num_tiles = m_tiles * n_tiles;
  const int num_kb = (p.K + BK - 1) / BK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), 4);  // one arrive per epilogue warp
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
    if (lane == 0) {
      if (p.nwait > 0) {
        // fused parameter pull: the weights this GEMM reads are written into this GPU's mirror by the
        // server shards' push kernels (peer stores over NVLink); acquire their version counters here, then
        // order the generic-proxy acquire before the async-proxy (TMA) reads
        const unsigned long long target = *p.wait_seq;
        for (int i = 0; i < p.nwait; ++i) {
          const unsigned long long* f = p.wait_flags[i];
          unsigned long long v;
          const long long t0 = clock64();
          do {
            asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(f) : "memory");
            if (v < target && clock64() - t0 > 60000000000LL) __trap();
          } while (v < target);
        }
        asm volatile("fence.proxy.async;" ::: "memory");
      }
      uint32_t stage = 0, phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m0 = (tile % m_tiles) * BM;
        const int n0 = (tile / m_tiles) * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          const uint32_t fb = full_bar(stage);
          mbar_arrive_expect_tx(fb, Cfg::STAGE_BYTES);
          const uint32_t sa = sA0 + stage * A_STAGE_BYTES;
          const uint32_t sb = sB0 + stage * B_STAGE_BYTES;
          const int k0 = kb * BK;
          if (!A_MN) {
            tma_load_2d(sa, &tma_a, fb, k0, m0);
          } else {
#pragma unroll
            for (int b = 0; b < BM / 64; ++b) tma_load_2d(sa + b * 8192, &tma_a, fb, m0 + b * 64, k0);
          }
          if (!B_MN) {
            tma_load_2d(sb, &tma_b, fb, k0, n0);
          } else {
#pragma unroll
            for (int b = 0; b < BN / 64; ++b) tma_load_2d(sb + b * 8192, &tma_b, fb, n0 + b * 64, k0);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    constexpr uint32_t idesc = make_idesc_bf16(BM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
    constexpr uint32_t A_KSTEP = A_MN ? (UMMA_K * 128) : (UMMA_K * 2);  // bytes per UMMA_K step
    constexpr uint32_t B_KSTEP = B_MN ? (UMMA_K * 128) : (UMMA_K * 2);
    constexpr uint32_t A_LBO = A_MN ? 8192 : 16;
    constexpr uint32_t B_LBO = B_MN ? 8192 : 16;
    uint32_t stage = 0, phase = 0;
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      mbar_wait(tempty_bar(acc), acc_phase ^ 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * BN;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(full_bar(stage), phase);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t sa = sA0 + stage * A_STAGE_BYTES;
          const uint32_t sb = sB0 + stage * B_STAGE_BYTES;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint64_t adesc = make_smem_desc_sw128(sa + k * A_KSTEP, A_LBO, 1024);
            const uint64_t bdesc = make_smem_desc_sw128(sb + k * B_KSTEP, B_LBO, 1024);
            umma_bf16(tmem_d, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(empty_bar(stage));
          if (kb == num_kb - 1) umma_commit(tfull_bar(acc));
        }
        __syncwarp();
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 2..5)
    const int quad = warp & 3;                 // TMEM lane quadrant this warp may read
    const int row_in_tile = quad * 32 + lane;
    const bool vec32_ok = p.d32 == nullptr ||
        ((p.ldd32 & 3) == 0 && (reinterpret_cast<uintptr_t>(p.d32) & 15) == 0);
    const bool vec16_ok = p.d16 == nullptr ||
        ((p.ldd16 & 7) == 0 && (reinterpret_cast<uintptr_t>(p.d16) & 15) == 0);
    const bool vecaux_ok = p.grad_act == SG_ACT_NONE ||
        (p.aux_bf16 ? ((p.ldaux & 7) == 0 && (reinterpret_cast<uintptr_t>(p.aux) & 15) == 0)
                    : ((p.ldaux & 3) == 0 && (reinterpret_cast<uintptr_t>(p.aux) & 15) == 0));
    const bool fast_ok = vec32_ok && vec16_ok && vecaux_ok && p.perm_c == 0;
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int m0 = (tile % m_tiles) * BM;
      const int n0 = (tile / m_tiles) * BN;
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      mbar_wait(tfull_bar(acc), acc_phase);
      tc_fence_after();
      const long long row = m0 + row_in_tile;
      const bool row_ok = row < p.M;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(quad * 32) << 16) + acc * BN;
#pragma unroll 1
      for (int c = 0; c < BN; c += 32) {
        uint32_t r[32];
        tmem_ld_32x32(taddr + c, r);
        tmem_ld_wait();
        if (c + 32 >= BN) {
          // accumulator fully in registers: hand the TMEM buffer back to the MMA warp early
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(tempty_bar(acc));
        }
        const int nb = n0 + c;
        if (!row_ok || nb >= p.N) continue;
        const int nvalid = min(32, p.N - nb);
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
        if (p.bias != nullptr) {
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < nvalid)
              v[j] += p.bias_bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p.bias)[nb + j])
                                  : __ldg(reinterpret_cast<const float*>(p.bias) + nb + j);
        }
        if (p.act != SG_ACT_NONE) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = apply_act(v[j], p.act);
        }
        if (fast_ok && nvalid == 32) {
          if (p.grad_act != SG_ACT_NONE) {
            if (p.aux_bf16) {
              const uint4* ap = reinterpret_cast<const uint4*>(
                  reinterpret_cast<const __nv_bfloat16*>(p.aux) + row * p.ldaux + nb);
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const uint4 u = __ldg(ap + q);
                const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                  const float y0 = __uint_as_float(w[t] << 16);
                  const float y1 = __uint_as_float(w[t] & 0xFFFF0000u);
                  v[q * 8 + t * 2] *= act_grad_from_output(y0, p.grad_act);
                  v[q * 8 + t * 2 + 1] *= act_grad_from_output(y1, p.grad_act);
                }
              }
            } else {
              const float4* ap = reinterpret_cast<const float4*>(
                  reinterpret_cast<const float*>(p.aux) + row * p.ldaux + nb);
#pragma unroll
              for (int q = 0; q < 8; ++q) {
                const float4 y = __ldg(ap + q);
                v[q * 4 + 0] *= act_grad_from_output(y.x, p.grad_act);
                v[q * 4 + 1] *= act_grad_from_output(y.y, p.grad_act);
                v[q * 4 + 2] *= act_grad_from_output(y.z, p.grad_act);
                v[q * 4 + 3] *= act_grad_from_output(y.w, p.grad_act);
              }
            }
          }
          if (p.d32 != nullptr) {
            float4* dp = reinterpret_cast<float4*>(p.d32 + row * p.ldd32 + nb);
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              float4 o = make_float4(v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
              if (p.accumulate) {
                const float4 old = dp[q];
                o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w;
                v[q * 4] = o.x; v[q * 4 + 1] = o.y; v[q * 4 + 2] = o.z; v[q * 4 + 3] = o.w;
              }
              dp[q] = o;
            }
          }
          if (p.d16 != nullptr) {
            uint4* dp = reinterpret_cast<uint4*>(p.d16 + row * p.ldd16 + nb);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              uint32_t w[4];
#pragma unroll
              for (int t = 0; t < 4; ++t) {
                const __nv_bfloat162 h = __floats2bfloat162_rn(v[q * 8 + t * 2], v[q * 8 + t * 2 + 1]);
                w[t] = *reinterpret_cast<const uint32_t*>(&h);
              }
              dp[q] = make_uint4(w[0], w[1], w[2], w[3]);
            }
          }
        } else {
          // generic path: ragged N, unaligned leading dimensions, or permuted columns
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            if (j < nvalid) {
              const int n = nb + j;
              float x = v[j];
              if (p.grad_act != SG_ACT_NONE) {
                const float y = p.aux_bf16
                    ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p.aux)[row * p.ldaux + n])
                    : reinterpret_cast<const float*>(p.aux)[row * p.ldaux + n];
                x *= act_grad_from_output(y, p.grad_act);
              }
              const int nout = p.perm_c > 0 ? (n % p.perm_c) * p.perm_kk + (n / p.perm_c) : n;
              if (p.d32 != nullptr) {
                float* dp = p.d32 + row * p.ldd32 + nout;
                if (p.accumulate) x += *dp;
                *dp = x;
              }
              if (p.d16 != nullptr) p.d16[row * p.ldd16 + nout] = __float2bfloat16_rn(x);
            }
          }
        }
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

// 2-D bf16 tensor map: `inner` contiguous elements per row, `outer` rows, row stride ld elements.
static int make_tmap_bf16(CUtensorMap* out, const void* ptr, long long inner, long long outer,
                          long long ld, int box_inner, int box_outer) {
  EncodeTiledFn enc = get_encode_fn();
  SG_REQUIRE(enc != nullptr, "cuTensorMapEncodeTiled unavailable (no CUDA driver?)");
  SG_REQUIRE((reinterpret_cast<uintptr_t>(ptr) & 15) == 0, "gemm operand base must be 16-byte aligned");
  SG_REQUIRE((ld % 8) == 0, "gemm operand leading dimension (%lld) must be a multiple of 8 bf16", ld);
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(inner), static_cast<cuuint64_t>(outer)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * 2};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(box_inner), static_cast<cuuint32_t>(box_outer)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  SG_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed (%d) inner=%lld outer=%lld ld=%lld",
             static_cast<int>(r), inner, outer, ld);
  return 0;
}

template <int BN, bool A_MN, bool B_MN>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p,
                       cudaStream_t stream) {
  using Cfg = GemmCfg<BN>;
  auto kern = gemm_bf16_tcgen05_kernel<BN, A_MN, B_MN>;
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
  const int tiles = m_tiles * n_tiles;
  const int grid = tiles < num_sms() ? tiles : num_sms();
  kern<<<grid, GEMM_THREADS, Cfg::SMEM_BYTES, stream>>>(ta, tb, p);
  SG_CHECK_LAUNCH("gemm_bf16_tcgen05");
  return 0;
}

template <int BN>
static int dispatch_major(int a_mn, int b_mn, const CUtensorMap& ta, const CUtensorMap& tb,
                          const GemmParams& p, cudaStream_t s) {
  if (!a_mn && !b_mn) return launch_gemm<BN, false, false>(ta, tb, p, s);
  if (!a_mn && b_mn) return launch_gemm<BN, false, true>(ta, tb, p, s);
  if (a_mn && !b_mn) return launch_gemm<BN, true, false>(ta, tb, p, s);
  return launch_gemm<BN, true, true>(ta, tb, p, s);
}

// Tile-width choice from a small cost model (cycles).  Measured on B200: this kernel is bound by the
// L2 -> shared-memory path (about 6300 B/cycle for the whole chip, i.e. ~42 B/cycle/SM with every SM pulling),
// not by the tensor pipe, so wide tiles (fewer bytes per flop) win even when they leave SMs idle:
//   per k-block:  mma = max(2*BN, 128 + BN) cycles   (tensor pipe vs. smem operand reads at 128 B/cycle)
//                 l2  = (16 KB + BN*128 B) / min(64, 6300 / active CTAs)
//   total = waves * (num_kb * max(mma, l2) + fixed per-tile overhead)
static int pick_bn(int M, int N, int K) {
  const int sms = num_sms();
  const long long m_tiles = (M + BM - 1) / BM;
  const long long num_kb = (K + BK - 1) / BK;
  const int cands[3] = {256, 128, 64};
  double best = 1e30;
  int best_bn = 64;
  for (int i = 0; i < 3; ++i) {
    const int bn = cands[i];
    if (bn > 64 && N <= bn / 2) continue;   // more than half of the tile would be padding
    const long long tiles = m_tiles * ((N + bn - 1) / bn);
    const long long active = tiles < sms ? tiles : sms;
    const long long waves = (tiles + sms - 1) / sms;
    const double mma = (2.0 * bn > 128.0 + bn) ? 2.0 * bn : 128.0 + bn;
    double bw = 6300.0 / static_cast<double>(active);
    if (bw > 64.0) bw = 64.0;
    const double l2 = (16384.0 + bn * 128.0) / bw;
    const double per_tile = num_kb * (mma > l2 ? mma : l2) + 2500.0 + 6.0 * bn;
    const double total = waves * per_tile;
    if (total < best) { best = total; best_bn = bn; }
  }
  return best_bn;
}

int g_force_bn = 0;  // test hook (sg_gemm_set_tile)

}  // namespace sg

extern "C" int sg_gemm_set_tile(int bn) {
  if (bn != 0 && bn != 64 && bn != 128 && bn != 256) {
    sg::set_error("sg_gemm_set_tile: bn must be 0, 64, 128 or 256");
    return 2;
  }
  sg::g_force_bn = bn;
  return 0;
}

extern "C" int sg_gemm_bf16(const void* A, int a_mn_major, long long lda, const void* B,
                            int b_mn_major, long long ldb, int M, int N, int K,
                            const SgGemmEpilogue* ep, void* stream) {
  using namespace sg;
  SG_REQUIRE(M > 0 && N > 0 && K > 0, "sg_gemm_bf16: empty problem %dx%dx%d", M, N, K);
  SG_REQUIRE(ep != nullptr && (ep->d32 != nullptr || ep->d16 != nullptr),
             "sg_gemm_bf16: epilogue needs an output");
  SG_REQUIRE(ep->grad_act == SG_ACT_NONE || ep->aux != nullptr, "sg_gemm_bf16: grad_act needs aux");
  const int bn = g_force_bn ? g_force_bn : pick_bn(M, N, K);
  CUtensorMap ta, tb;
  int rc;
  if (!a_mn_major) rc = make_tmap_bf16(&ta, A, K, M, lda, BK, BM);
  else rc = make_tmap_bf16(&ta, A, M, K, lda, 64, BK);
  if (rc) return rc;
  if (!b_mn_major) rc = make_tmap_bf16(&tb, B, K, N, ldb, BK, bn);
  else rc = make_tmap_bf16(&tb, B, N, K, ldb, 64, BK);
  if (rc) return rc;
  GemmParams p;
  p.M = M; p.N = N; p.K = K;
  p.d32 = ep->d32; p.ldd32 = ep->ldd32;
  p.d16 = reinterpret_cast<__nv_bfloat16*>(ep->d16); p.ldd16 = ep->ldd16;
  p.bias = ep->bias; p.bias_bf16 = ep->bias_bf16; p.act = ep->act; p.grad_act = ep->grad_act;
  p.aux = ep->aux; p.ldaux = ep->ldaux; p.aux_bf16 = ep->aux_bf16;
  p.accumulate = ep->accumulate; p.perm_c = ep->perm_c; p.perm_kk = ep->perm_kk;
  SG_REQUIRE(ep->nwait >= 0 && ep->nwait <= 8, "sg_gemm_bf16: at most 8 wait flags");
  SG_REQUIRE(ep->nwait == 0 || ep->wait_seq != nullptr, "sg_gemm_bf16: wait flags need wait_seq");
  for (int i = 0; i < 8; ++i) p.wait_flags[i] = i < ep->nwait ? ep->wait_flags[i] : nullptr;
  p.nwait = ep->nwait; p.wait_seq = ep->wait_seq;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  switch (bn) {
    case 256: return dispatch_major<256>(a_mn_major, b_mn_major, ta, tb, p, s);
    case 128: return dispatch_major<128>(a_mn_major, b_mn_major, ta, tb, p, s);
    default: return dispatch_major<64>(a_mn_major, b_mn_major, ta, tb, p, s);
  }
}
