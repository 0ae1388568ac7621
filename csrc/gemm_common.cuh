// Shared pieces of the tcgen05 GEMM kernels (1-CTA and CTA-pair variants): tile constants, kernel parameters and
// the fused epilogue.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <stdint.h>

#include "../include/singa_b200.h"
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

// Kept at 96 bytes: with the two 128-byte tensor maps the kernel's parameter block stays under 368 bytes.  Growing it
// to ~410 bytes cost ~1.3 us on EVERY launch (measured: +0.026 ms on the 18-GEMM MNIST step), so flags are bytes
// and leading dimensions are 32-bit.
struct GemmParams {
  int M, N, K;
  int kb_per_split;
  float* d32;
  __nv_bfloat16* d16;
  const void* bias;
  const void* aux;
  const unsigned long long* const* wait_flags;   // device array of nwait flag addresses
  const unsigned long long* wait_seq;
  int ldd32, ldd16, ldaux;
  int perm_c, perm_kk;
  short nwait, split_k;
  unsigned char act, grad_act, aux_bf16, accumulate, bias_bf16;
};
static_assert(sizeof(GemmParams) <= 96, "GemmParams grew: see the launch-latency note above");

// Fused parameter pull (src/utils/param_manager.cc:204-234 WaitUpdate): the weights a forward GEMM reads are written
// into this GPU's bf16 mirror by the server shards' push kernels (peer stores over NVLink).  The TMA producer thread
// acquires their version counters here, then orders the generic-proxy acquire before the async-proxy (TMA) reads.
// Kept out of line so the common (no wait) path carries no extra code or registers.
__device__ __noinline__ void wait_param_versions(const unsigned long long* const* flags, int nwait,
                                                 const unsigned long long* seq) {
  const unsigned long long target = *seq;
  for (int i = 0; i < nwait; ++i) {
    const unsigned long long* f = flags[i];
    unsigned long long v;
    const long long t0 = clock64();
    do {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(f) : "memory");
      if (v < target && clock64() - t0 > 60000000000LL) __trap();
    } while (v < target);
  }
  asm volatile("fence.proxy.async;" ::: "memory");
}

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


// Epilogue for one 32-column chunk of one accumulator row: bias, activation, activation gradient, accumulate,
// fp32 / bf16 stores (128-bit when the buffers allow it, element-wise otherwise).
__device__ __forceinline__ void epilogue_chunk(const GemmParams& p, const uint32_t (&r)[32], long long row,
                                               bool row_ok, int nb, bool fast_ok) {
  do {
                if (!row_ok || nb >= p.N) break;
        const int nvalid = min(32, p.N - nb);
        float v[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
        if (p.bias != nullptr) {
          // two separate, fully unrolled loops so the 32 loads of a chunk are independent and issued together
          // (one loop with a per-element fp32/bf16 select serialised them: +5-10 us per forward GEMM)
          if (p.bias_bf16) {
            const __nv_bfloat16* bp = reinterpret_cast<const __nv_bfloat16*>(p.bias) + nb;
            if (nvalid == 32 && (reinterpret_cast<uintptr_t>(bp) & 15) == 0) {
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const uint4 u = __ldg(reinterpret_cast<const uint4*>(bp) + q);
                const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                  v[q * 8 + t * 2] += __uint_as_float(w[t] << 16);
                  v[q * 8 + t * 2 + 1] += __uint_as_float(w[t] & 0xFFFF0000u);
                }
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (j < nvalid) v[j] += __bfloat162float(bp[j]);
            }
          } else {
            const float* bp = reinterpret_cast<const float*>(p.bias) + nb;
            if (nvalid == 32 && (reinterpret_cast<uintptr_t>(bp) & 15) == 0) {
#pragma unroll
              for (int q = 0; q < 8; ++q) {
                const float4 b4 = __ldg(reinterpret_cast<const float4*>(bp) + q);
                v[q * 4] += b4.x; v[q * 4 + 1] += b4.y; v[q * 4 + 2] += b4.z; v[q * 4 + 3] += b4.w;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (j < nvalid) v[j] += __ldg(bp + j);
            }
          }
        }
        if (p.act != SG_ACT_NONE) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = apply_act(v[j], p.act);
        }
        if (fast_ok && nvalid == 32 && p.accumulate != 2) {
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
                if (p.accumulate == 2) {
                  atomicAdd(dp, x);          // split-K partial (result unused: compiles to RED.ADD)
                  continue;
                }
                if (p.accumulate) x += *dp;
                *dp = x;
              }
              if (p.d16 != nullptr) p.d16[row * p.ldd16 + nout] = __float2bfloat16_rn(x);
            }
          }
        }
  } while (false);
}

__device__ __forceinline__ bool epilogue_fast_ok(const GemmParams& p) {
  const bool vec32_ok = p.d32 == nullptr ||
      ((p.ldd32 & 3) == 0 && (reinterpret_cast<uintptr_t>(p.d32) & 15) == 0);
  const bool vec16_ok = p.d16 == nullptr ||
      ((p.ldd16 & 7) == 0 && (reinterpret_cast<uintptr_t>(p.d16) & 15) == 0);
  const bool vecaux_ok = p.grad_act == SG_ACT_NONE ||
      (p.aux_bf16 ? ((p.ldaux & 7) == 0 && (reinterpret_cast<uintptr_t>(p.aux) & 15) == 0)
                  : ((p.ldaux & 3) == 0 && (reinterpret_cast<uintptr_t>(p.aux) & 15) == 0));
  return vec32_ok && vec16_ok && vecaux_ok && p.perm_c == 0;
}

}  // namespace sg
