// Inner product with at most 16 outputs (the classifier layer: 500 -> 10, 64*4*4 -> 10, ...): CUDA-core kernels.
// A 128-row tcgen05 tile is 6-50 % filled by such a layer and the three GEMM launches of its training step cost
// 35 us on the MNIST MLP (a third of them launch / set-up / exposed-epilogue time) for 30 MFLOP; these kernels take
// the same operands in the same layouts (reference op: InnerProductLayer, src/worker/layer.cc:162-213):
//   forward : y[M][N]  = act(x[M][K] . W[K][N] + b)          one warp per row, lanes over K, W staged in shared memory
//   dgrad   : dx[M][K] = (dy[M][N] . W[K][N]^T) * act'(aux)  eight rows per block, threads over K, W in shared memory
//   wgrad   : dW[K][N] (+)= x^T dy,  db[N] (+)= column sums  (64-wide k strip) x (32-row slice) per block, red.add
// The staged weight matrix must fit 96 KB (K * (N | 1) * 4 bytes); larger layers stay on the GEMM.
// Element types are run-time flags (fp32 / bf16 per operand); accumulation is fp32 FMA.
#include <stdint.h>

#include "../include/singa_b200.h"
#include "common.h"
#include "ptx.cuh"
#include "gemm_common.cuh"

namespace sg {

constexpr int IPS_NMAX = 16;

__device__ __forceinline__ float ld_any(const void* p, long long i, int bf16) {
  return bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i]) : reinterpret_cast<const float*>(p)[i];
}
__device__ __forceinline__ void st_any(void* p, long long i, int bf16, float v) {
  if (bf16) reinterpret_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16(v);
  else reinterpret_cast<float*>(p)[i] = v;
}

// W[K][N] staged once per block as fp32 [K][NP], NP = N | 1 (odd row stride: threads that walk K hit distinct banks).
__device__ __forceinline__ void stage_w(float* ws, const void* w, long long ldw, int w_bf16, int K, int N, int NP) {
  for (int i = threadIdx.x; i < K * N; i += blockDim.x) {
    const int k = i / N, n = i - k * N;
    ws[k * NP + n] = ld_any(w, k * ldw + n, w_bf16);
  }
}

// One warp per row; a lane takes k = lane, lane + 32, ... in chunks of eight loads issued together.
__global__ void __launch_bounds__(256)
ip_small_fwd_kernel(const void* __restrict__ x, long long ldx, int x_bf16, const void* __restrict__ w, long long ldw,
                    int w_bf16, const void* __restrict__ bias, int bias_bf16, void* __restrict__ y, long long ldy,
                    int y_bf16, int M, int K, int N, int act) {
  extern __shared__ float ws[];
  const int NP = N | 1;
  stage_w(ws, w, ldw, w_bf16, K, N, NP);
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int wpb = blockDim.x >> 5;
  for (int row = blockIdx.x * wpb + (threadIdx.x >> 5); row < M; row += gridDim.x * wpb) {
    float acc[IPS_NMAX];
#pragma unroll
    for (int n = 0; n < IPS_NMAX; ++n) acc[n] = 0.f;
    for (int k0 = 0; k0 < K; k0 += 256) {
      float xv[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int k = k0 + lane + 32 * j;
        xv[j] = k < K ? ld_any(x, row * ldx + k, x_bf16) : 0.f;
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int k = k0 + lane + 32 * j;
        if (k < K) {
          const float* wr = ws + k * NP;
#pragma unroll
          for (int n = 0; n < IPS_NMAX; ++n)
            if (n < N) acc[n] = fmaf(xv[j], wr[n], acc[n]);
        }
      }
    }
#pragma unroll
    for (int n = 0; n < IPS_NMAX; ++n) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) acc[n] += __shfl_xor_sync(0xffffffffu, acc[n], o);
    }
    float v = 0.f;     // lane n stores output n (a compile-time select chain keeps acc[] in registers)
#pragma unroll
    for (int n = 0; n < IPS_NMAX; ++n)
      if (lane == n) v = acc[n];
    if (lane < N) {
      if (bias != nullptr) v += ld_any(bias, lane, bias_bf16);
      st_any(y, row * ldy + lane, y_bf16, apply_act(v, act));
    }
  }
}

// A block takes IPS_RB rows; threads walk K (coalesced aux loads and dx stores), dy rows and W come from shared memory.
constexpr int IPS_RB = 8;
__global__ void __launch_bounds__(256)
ip_small_dgrad_kernel(const void* __restrict__ dy, long long lddy, int dy_bf16, const void* __restrict__ w, long long ldw,
                      int w_bf16, const void* __restrict__ aux, long long ldaux, int aux_bf16, int grad_act,
                      void* __restrict__ dx, long long lddx, int dx_bf16, int M, int K, int N) {
  extern __shared__ float ws[];
  __shared__ float g[IPS_RB][IPS_NMAX];
  const int NP = N | 1;
  stage_w(ws, w, ldw, w_bf16, K, N, NP);
  const int row0 = blockIdx.x * IPS_RB;
  if (threadIdx.x < IPS_RB * IPS_NMAX) {
    const int r = threadIdx.x / IPS_NMAX, n = threadIdx.x % IPS_NMAX;
    g[r][n] = (row0 + r < M && n < N) ? ld_any(dy, (row0 + r) * lddy + n, dy_bf16) : 0.f;
  }
  __syncthreads();
  for (int k = threadIdx.x; k < K; k += blockDim.x) {
    float wr[IPS_NMAX];
#pragma unroll
    for (int n = 0; n < IPS_NMAX; ++n) wr[n] = n < N ? ws[k * NP + n] : 0.f;
    float a[IPS_RB];
    if (grad_act != SG_ACT_NONE) {
#pragma unroll
      for (int r = 0; r < IPS_RB; ++r) a[r] = row0 + r < M ? ld_any(aux, (row0 + r) * ldaux + k, aux_bf16) : 0.f;
    }
#pragma unroll
    for (int r = 0; r < IPS_RB; ++r) {
      if (row0 + r >= M) break;
      float sum = 0.f;
#pragma unroll
      for (int n = 0; n < IPS_NMAX; ++n) sum = fmaf(g[r][n], wr[n], sum);
      if (grad_act != SG_ACT_NONE) sum *= act_grad_from_output(a[r], grad_act);
      st_any(dx, (row0 + r) * lddx + k, dx_bf16, sum);
    }
  }
}

// grid (K / 64 strips, row slices of IPS_WM rows); thread = (k of the strip, group of four outputs).  The slice of dy sits
// in shared memory, x is read coalesced along k, eight rows in flight.  Every block stores its partial sums in the
// caller's workspace; the block that arrives last at its strip's counter adds the slices up IN SLICE ORDER (the result
// does not depend on the order of arrival: training runs are reproducible bit for bit) and writes dw / db.
//   ws: float part[slices][K][16], float bpart[slices][16], unsigned counter[strips] (zero before the first launch; the
//   kernel leaves the counters at zero)
constexpr int IPS_WM = 32;
__global__ void __launch_bounds__(256)
ip_small_wgrad_kernel(const void* __restrict__ x, long long ldx, int x_bf16, const void* __restrict__ dy, long long lddy,
                      int dy_bf16, float* __restrict__ dw, long long lddw, float* __restrict__ db, int M, int K, int N,
                      int accumulate, float* __restrict__ ws) {
  __shared__ float g[IPS_WM][IPS_NMAX];
  __shared__ int is_last;
  const int nslices = gridDim.y, slice = blockIdx.y;
  float* part = ws;
  float* bpart = ws + static_cast<long long>(nslices) * K * IPS_NMAX;
  unsigned* counter = reinterpret_cast<unsigned*>(bpart + nslices * IPS_NMAX);
  const int m0 = slice * IPS_WM;
  for (int i = threadIdx.x; i < IPS_WM * IPS_NMAX; i += blockDim.x) {
    const int r = i / IPS_NMAX, n = i % IPS_NMAX;
    g[r][n] = (m0 + r < M && n < N) ? ld_any(dy, (m0 + r) * lddy + n, dy_bf16) : 0.f;
  }
  __syncthreads();
  const int kk = threadIdx.x & 63, ng = threadIdx.x >> 6;
  const int k = blockIdx.x * 64 + kk;
  if (k < K) {
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int r0 = 0; r0 < IPS_WM; r0 += 8) {
      float xv[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) xv[j] = m0 + r0 + j < M ? ld_any(x, (m0 + r0 + j) * ldx + k, x_bf16) : 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j)
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[q] = fmaf(xv[j], g[r0 + j][ng * 4 + q], acc[q]);
    }
    *reinterpret_cast<float4*>(part + (static_cast<long long>(slice) * K + k) * IPS_NMAX + ng * 4) =
        make_float4(acc[0], acc[1], acc[2], acc[3]);
  }
  if (blockIdx.x == 0 && threadIdx.x < IPS_NMAX) {
    float sum = 0.f;
    for (int r = 0; r < IPS_WM; ++r) sum += g[r][threadIdx.x];
    bpart[slice * IPS_NMAX + threadIdx.x] = sum;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(counter + blockIdx.x, 1u) == static_cast<unsigned>(nslices - 1);
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  if (k < K) {
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int y = 0; y < nslices; ++y) {
      const float4 v = __ldcg(reinterpret_cast<const float4*>(part + (static_cast<long long>(y) * K + k) * IPS_NMAX + ng * 4));
      acc[0] += v.x; acc[1] += v.y; acc[2] += v.z; acc[3] += v.w;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int n = ng * 4 + q;
      if (n < N) {
        float* d = dw + k * lddw + n;
        *d = accumulate ? *d + acc[q] : acc[q];
      }
    }
  }
  if (db != nullptr && blockIdx.x == 0 && threadIdx.x < N) {
    float sum = 0.f;
    for (int y = 0; y < nslices; ++y) sum += __ldcg(bpart + y * IPS_NMAX + threadIdx.x);
    db[threadIdx.x] = accumulate ? db[threadIdx.x] + sum : sum;
  }
  if (threadIdx.x == 0) counter[blockIdx.x] = 0u;
}

}  // namespace sg

using namespace sg;

static int ips_smem(int K, int N, const char* who) {
  const long long bytes = static_cast<long long>(K) * (N | 1) * 4;
  if (bytes > 96 * 1024) {
    set_error("%s: the weight matrix (%d x %d) does not fit the 96 KB staging buffer", who, K, N);
    return -1;
  }
  return static_cast<int>(bytes);
}

// 96 KB of dynamic shared memory must be allowed once per kernel and device
static int ips_allow_smem(const void* kern, int which) {
  static thread_local int done_dev[2] = {-1, -1};
  int dev = 0;
  cudaGetDevice(&dev);
  if (done_dev[which] == dev) return 0;
  if (check_cuda(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024),
                 "cudaFuncSetAttribute(ip_small smem)"))
    return 1;
  done_dev[which] = dev;
  return 0;
}

extern "C" int sg_ip_small_fwd(const void* x, long long ldx, int x_bf16, const void* w, long long ldw, int w_bf16,
                               const void* bias, int bias_bf16, void* y, long long ldy, int y_bf16, int M, int K, int N,
                               int act, void* stream) {
  SG_REQUIRE(x != nullptr && w != nullptr && y != nullptr && M > 0 && K > 0 && N > 0 && N <= IPS_NMAX,
             "sg_ip_small_fwd: needs 1..%d outputs", IPS_NMAX);
  const int smem = ips_smem(K, N, "sg_ip_small_fwd");
  if (smem < 0) return 2;
  if (ips_allow_smem(reinterpret_cast<const void*>(ip_small_fwd_kernel), 0)) return 1;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  int grid = (M + 7) / 8;
  if (grid > 2 * num_sms()) grid = 2 * num_sms();
  ip_small_fwd_kernel<<<grid, 256, smem, s>>>(x, ldx, x_bf16, w, ldw, w_bf16, bias, bias_bf16, y, ldy, y_bf16, M, K, N, act);
  SG_CHECK_LAUNCH("ip_small_fwd");
  return 0;
}

extern "C" int sg_ip_small_dgrad(const void* dy, long long lddy, int dy_bf16, const void* w, long long ldw, int w_bf16,
                                 const void* aux, long long ldaux, int aux_bf16, int grad_act, void* dx, long long lddx,
                                 int dx_bf16, int M, int K, int N, void* stream) {
  SG_REQUIRE(dy != nullptr && w != nullptr && dx != nullptr && M > 0 && K > 0 && N > 0 && N <= IPS_NMAX,
             "sg_ip_small_dgrad: needs 1..%d outputs", IPS_NMAX);
  SG_REQUIRE(grad_act == SG_ACT_NONE || aux != nullptr, "sg_ip_small_dgrad: grad_act needs aux");
  const int smem = ips_smem(K, N, "sg_ip_small_dgrad");
  if (smem < 0) return 2;
  if (ips_allow_smem(reinterpret_cast<const void*>(ip_small_dgrad_kernel), 1)) return 1;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  ip_small_dgrad_kernel<<<(M + IPS_RB - 1) / IPS_RB, 256, smem, s>>>(dy, lddy, dy_bf16, w, ldw, w_bf16, aux, ldaux, aux_bf16, grad_act, dx, lddx, dx_bf16, M, K, N);
  SG_CHECK_LAUNCH("ip_small_dgrad");
  return 0;
}

extern "C" long long sg_ip_small_wgrad_ws_bytes(int M, int K) {
  const long long slices = (M + IPS_WM - 1) / IPS_WM, strips = (K + 63) / 64;
  return (slices * K * IPS_NMAX + slices * IPS_NMAX) * 4 + strips * 4;
}

extern "C" int sg_ip_small_wgrad(const void* x, long long ldx, int x_bf16, const void* dy, long long lddy, int dy_bf16,
                                 float* dw, long long lddw, float* db, int M, int K, int N, int accumulate, void* ws,
                                 void* stream) {
  SG_REQUIRE(x != nullptr && dy != nullptr && dw != nullptr && M > 0 && K > 0 && N > 0 && N <= IPS_NMAX,
             "sg_ip_small_wgrad: needs 1..%d outputs", IPS_NMAX);
  SG_REQUIRE(ws != nullptr && (reinterpret_cast<uintptr_t>(ws) & 15) == 0,
             "sg_ip_small_wgrad: needs a 16-byte aligned, zero-initialised workspace of sg_ip_small_wgrad_ws_bytes(M, K)");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  ip_small_wgrad_kernel<<<dim3((K + 63) / 64, (M + IPS_WM - 1) / IPS_WM), 256, 0, s>>>(
      x, ldx, x_bf16, dy, lddy, dy_bf16, dw, lddw, db, M, K, N, accumulate, reinterpret_cast<float*>(ws));
  SG_CHECK_LAUNCH("ip_small_wgrad");
  return 0;
}
