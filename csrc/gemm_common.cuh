// Shared pieces of the tcgen05 GEMM kernels (1-CTA and CTA-pair variants): tile constants, kernel parameters and
// the fused epilogue.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <stdint.h>

#include "../include/singa_b200.h"
#include "ptx.cuh"

namespace sg {

#ifdef SG_GEMM_TRACE
__device__ unsigned long long g_gemm_trace[32];
#define SG_TRACE(i) do { if (blockIdx.x == 0) g_gemm_trace[i] = clock64(); } while (0)
#else
#define SG_TRACE(i) do { } while (0)
#endif
#define SG_TRACE_W(i) do { if (threadIdx.x == 64 && nb == 0) SG_TRACE(i); } while (0)

constexpr int BM = 128;
// Operand element types.  Everything in the main loop is sized in BYTES: a k-block is one 128-byte swizzle row
// (64 bf16 or 32 tf32/fp32 elements) and one tcgen05.mma consumes 32 bytes of K (UMMA_K = 16 bf16 or 8 tf32), so
// the two precisions share tile sizes, stage counts and descriptors and differ only in the element counts below,
// the tensor-map data type and the MMA kind.  tf32 operands are plain fp32 in memory (the reference computes in
// fp32, src/worker/layer.cc:186-213); the tensor core reads the upper 19 bits.
template <int EB> struct Elem {                // EB = bytes per operand element: 2 (bf16) or 4 (tf32)
  static constexpr int BK = 128 / EB;          // elements per k-block
  static constexpr int UMMA_K = 32 / EB;       // elements per MMA instruction
  static constexpr int MN_BOX = 128 / EB;      // MN-major operand: elements per 128-byte row of a smem box
  static constexpr int MN_BOX_BYTES = BK * 128;  // one [BK k][MN_BOX mn] box
};
constexpr int A_STAGE_BYTES = BM * 128;      // 16 KB
constexpr int GEMM_THREADS = 320;             // warp 0 TMA producer, warp 1 MMA issuer, warps 2..9 epilogue
constexpr int SMEM_BUDGET = 200 * 1024;

template <int BN> struct GemmCfg {
  static constexpr int B_STAGE_BYTES = BN * 128;
  static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
  static constexpr int STAGES = (SMEM_BUDGET / STAGE_BYTES) > 8 ? 8 : (SMEM_BUDGET / STAGE_BYTES);
  static constexpr int TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;  // power of two for BN in {64,128,256}
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/ + 16 * 1024 /*epilogue staging, EPI_SMEM_BYTES*/;
};

// CTA pairs (cta_group::2): 256 x BN tiles, each CTA stages its own 128 rows of A and half of B.
template <int BN> struct PairCfg {
  static constexpr int B_HALF_BYTES = (BN / 2) * 128;
  static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_HALF_BYTES;       // per CTA
  static constexpr int STAGES = (SMEM_BUDGET / STAGE_BYTES) > 8 ? 8 : (SMEM_BUDGET / STAGE_BYTES);
  static constexpr int TMEM_COLS = 2 * BN;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256 + 16 * 1024 /*epilogue staging*/;
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
  int perm_c;                   // > 0: column n = (tap * perm_c + c) is stored at column c * perm_kk + tap ...
  short perm_kk, perm_valid;    // ... and dropped when perm_valid > 0 and c >= perm_valid (zero-padded channel chunks)
  short nwait, split_k;
  unsigned char act, grad_act, aux_bf16, accumulate, bias_bf16;
};
static_assert(sizeof(GemmParams) <= 96, "GemmParams grew: see the launch-latency note above");

// Fused parameter pull (src/utils/param_manager.cc:204-234 WaitUpdate): the weights a forward GEMM reads are written
// into this GPU's bf16 mirror by the server shards' push kernels (peer stores over NVLink).  The TMA producer thread
// acquires their version counters here, then orders the generic-proxy acquire before the async-proxy (TMA) reads.
// Kept out of line so the common (no wait) path carries no extra code or registers.
static __device__ __noinline__ void wait_param_versions(const unsigned long long* const* flags, int nwait,
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

// The bias a forward GEMM adds may be this GPU's weight mirror, which peer GPUs' push kernels overwrite while this
// kernel is already resident (its producer lane is spinning in wait_param_versions).  Non-coherent loads
// (ld.global.nc / __ldg) are only defined for memory that is read-only for the kernel's lifetime, so the bias goes
// through ordinary coherent loads; they are ordered after the producer's acquire by the mbarrier chain
// (acquire -> TMA -> full barrier -> MMA commit -> accumulator-full barrier -> this load).
__device__ __forceinline__ uint4 ld_coherent_u4(const uint4* p) {
  uint4 v;
  asm volatile("ld.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
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


// Epilogue staging.  tcgen05.ld hands each lane one accumulator ROW, so direct stores touch 32 different cache lines
// per instruction (LSU-bound: ~1100 cycles per 32-column fp32 chunk, 14 000 cycles for a 128 x 256 tile), and with
// one epilogue warp per scheduler the code is also issue-latency bound.  Hence: EIGHT epilogue warps (two per TMEM
// lane quadrant, alternating 32-column chunks) and a 32 x 16 fp32 staging tile per warp (2 KB, XOR-swizzled 16-byte
// groups: conflict-free 128-bit accesses both ways).  A chunk is staged in two 16-column halves and read back with
// consecutive lanes on consecutive columns, so every global access of the epilogue (stores, accumulate loads,
// activation-gradient loads) covers whole 32/64-byte row segments.
constexpr int EPI_WARPS = 8;
constexpr int EPI_WARP_FLOATS = 32 * 16;
constexpr int EPI_SMEM_BYTES = EPI_WARPS * EPI_WARP_FLOATS * 4;   // 16 KB

// Coalesced store of one staged 32-column chunk, specialised at compile time so the executed path is short and
// straight (with one or two warps per scheduler the epilogue is issue-latency bound: instruction count is time).
//   D32: fp32 output (else bf16);  AUXK: 0 = none, 1 = bf16, 2 = fp32 activation output for the fused act';
//   ACC: d32 += result.
template <bool D32, int AUXK, bool ACC>
__device__ __forceinline__ void staged_store(const GemmParams& p, const float (&v)[32], long long row0, int lane,
                                             int nb, float* stg) {
  const int sw_w = (lane >> 1) & 3;
  const int grad_act = p.grad_act;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    __syncwarp();                            // the previous half's readers are done with the tile
    if (h == 0) SG_TRACE_W(14);
#pragma unroll
    for (int k = 0; k < 4; ++k)
      *reinterpret_cast<float4*>(stg + lane * 16 + 4 * (k ^ sw_w)) =
          make_float4(v[16 * h + 4 * k], v[16 * h + 4 * k + 1], v[16 * h + 4 * k + 2], v[16 * h + 4 * k + 3]);
    __syncwarp();
    if (h == 0) SG_TRACE_W(15);
    const int cb = nb + 16 * h;
    if (D32) {
      // 4 lanes x 16 B cover one 64-byte row segment; one instruction writes 8 rows
      const int g = lane & 3;
      const int rl = lane >> 2;
      float* dbase = p.d32 + (row0 + rl) * p.ldd32 + cb + 4 * g;
      const int rows_left = static_cast<int>(min(static_cast<long long>(32), p.M - row0)) - rl;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const int rr = q * 8 + rl;
        float4 o = *reinterpret_cast<const float4*>(stg + rr * 16 + 4 * (g ^ ((rr >> 1) & 3)));
        if (q * 8 < rows_left) {
          float4* dp = reinterpret_cast<float4*>(dbase + static_cast<long long>(q * 8) * p.ldd32);
          if (AUXK == 2) {
            const float4 y = __ldg(reinterpret_cast<const float4*>(
                reinterpret_cast<const float*>(p.aux) + (row0 + rr) * p.ldaux + cb + 4 * g));
            o.x *= act_grad_from_output(y.x, grad_act); o.y *= act_grad_from_output(y.y, grad_act);
            o.z *= act_grad_from_output(y.z, grad_act); o.w *= act_grad_from_output(y.w, grad_act);
          }
          if (ACC) {
            const float4 old = *dp;
            o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w;
          }
          *dp = o;
        }
      }
      if (h == 0) SG_TRACE_W(16); else SG_TRACE_W(17);
    } else {
      // 2 lanes x 16 B cover one 32-byte row segment (a full sector); one instruction writes 16 rows
      const int g2 = (lane & 1) * 2;
      const int rl = lane >> 1;
      __nv_bfloat16* dbase = p.d16 + (row0 + rl) * p.ldd16 + cb + 4 * g2;
      const int rows_left = static_cast<int>(min(static_cast<long long>(32), p.M - row0)) - rl;
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int rr = q * 16 + rl;
        const int s = (rr >> 1) & 3;
        const float4 a = *reinterpret_cast<const float4*>(stg + rr * 16 + 4 * (g2 ^ s));
        const float4 b = *reinterpret_cast<const float4*>(stg + rr * 16 + 4 * ((g2 + 1) ^ s));
        float o[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
        if (q * 16 < rows_left) {
          if (AUXK == 1) {
            const uint4 u = __ldg(reinterpret_cast<const uint4*>(
                reinterpret_cast<const __nv_bfloat16*>(p.aux) + (row0 + rr) * p.ldaux + cb + 4 * g2));
            const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
            for (int t = 0; t < 4; ++t) {
              o[t * 2] *= act_grad_from_output(__uint_as_float(w[t] << 16), grad_act);
              o[t * 2 + 1] *= act_grad_from_output(__uint_as_float(w[t] & 0xFFFF0000u), grad_act);
            }
          } else if (AUXK == 2) {
            const float4* ap = reinterpret_cast<const float4*>(
                reinterpret_cast<const float*>(p.aux) + (row0 + rr) * p.ldaux + cb + 4 * g2);
            const float4 y0 = __ldg(ap), y1 = __ldg(ap + 1);
            o[0] *= act_grad_from_output(y0.x, grad_act); o[1] *= act_grad_from_output(y0.y, grad_act);
            o[2] *= act_grad_from_output(y0.z, grad_act); o[3] *= act_grad_from_output(y0.w, grad_act);
            o[4] *= act_grad_from_output(y1.x, grad_act); o[5] *= act_grad_from_output(y1.y, grad_act);
            o[6] *= act_grad_from_output(y1.z, grad_act); o[7] *= act_grad_from_output(y1.w, grad_act);
          }
          uint32_t w[4];
#pragma unroll
          for (int t = 0; t < 4; ++t) {
            const __nv_bfloat162 hh = __floats2bfloat162_rn(o[t * 2], o[t * 2 + 1]);
            w[t] = *reinterpret_cast<const uint32_t*>(&hh);
          }
          *reinterpret_cast<uint4*>(dbase + static_cast<long long>(q * 16) * p.ldd16) = make_uint4(w[0], w[1], w[2], w[3]);
        }
      }
    }
  }
}

// Epilogue for one 32-column chunk of a warp's 32 accumulator rows: bias, activation, activation gradient,
// accumulate, fp32 / bf16 stores.  `row0` is the first row of the warp's slab, `stg` its staging tile.
__device__ __forceinline__ void epilogue_chunk(const GemmParams& p, const uint32_t (&r)[32], long long row0,
                                               int lane, int nb, bool fast_ok, float* stg) {
  if (nb >= p.N) return;                       // warp-uniform
  const int nvalid = min(32, p.N - nb);
  const long long row = row0 + lane;
  const bool row_ok = row < p.M;
  float v[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
  if (p.bias != nullptr) {
    // two separate, fully unrolled loops so the 32 loads of a chunk are independent and issued together
    if (p.bias_bf16) {
      const __nv_bfloat16* bp = reinterpret_cast<const __nv_bfloat16*>(p.bias) + nb;
      if (nvalid == 32 && (reinterpret_cast<uintptr_t>(bp) & 15) == 0) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint4 u = ld_coherent_u4(reinterpret_cast<const uint4*>(bp) + q);
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
          if (j < nvalid)
            v[j] += __uint_as_float(static_cast<uint32_t>(*reinterpret_cast<const volatile unsigned short*>(bp + j)) << 16);
      }
    } else {
      const float* bp = reinterpret_cast<const float*>(p.bias) + nb;
      if (nvalid == 32 && (reinterpret_cast<uintptr_t>(bp) & 15) == 0) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const uint4 u = ld_coherent_u4(reinterpret_cast<const uint4*>(bp) + q);
          const float4 b4 = make_float4(__uint_as_float(u.x), __uint_as_float(u.y), __uint_as_float(u.z),
                                        __uint_as_float(u.w));
          v[q * 4] += b4.x; v[q * 4 + 1] += b4.y; v[q * 4 + 2] += b4.z; v[q * 4 + 3] += b4.w;
        }
      } else {
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (j < nvalid) v[j] += *reinterpret_cast<const volatile float*>(bp + j);
      }
    }
  }
  if (p.act != SG_ACT_NONE) {
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = apply_act(v[j], p.act);
  }
  if (fast_ok && nvalid == 32 && p.accumulate != 2) {           // warp-uniform dispatch to a specialised path
    const bool aux_none = p.grad_act == SG_ACT_NONE;
    if (p.d32 != nullptr && p.d16 == nullptr && aux_none) {
      if (p.accumulate) staged_store<true, 0, true>(p, v, row0, lane, nb, stg);
      else staged_store<true, 0, false>(p, v, row0, lane, nb, stg);
      return;
    }
    if (p.d32 != nullptr && p.d16 == nullptr && !p.aux_bf16 && !p.accumulate) {   // fp32 data gradient with act'
      staged_store<true, 2, false>(p, v, row0, lane, nb, stg);
      return;
    }
    if (p.d16 != nullptr && p.d32 == nullptr && !p.accumulate) {
      if (aux_none) staged_store<false, 0, false>(p, v, row0, lane, nb, stg);
      else if (p.aux_bf16) staged_store<false, 1, false>(p, v, row0, lane, nb, stg);
      else staged_store<false, 2, false>(p, v, row0, lane, nb, stg);
      return;
    }
  }
  // generic path: ragged N, unaligned leading dimensions, permuted columns, split-K partial sums
  if (!row_ok) return;
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
      int nout = n;
      if (p.perm_c > 0) {
        const int tap = n / p.perm_c, c = n - tap * p.perm_c;
        if (p.perm_valid > 0 && c >= p.perm_valid) continue;
        nout = c * p.perm_kk + tap;
      }
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
