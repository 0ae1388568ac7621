// Inner product with at most 16 outputs (the classifier layer: 500 -> 10, 64*4*4 -> 10, ...): CUDA-core kernels.
// A 128-row tcgen05 tile is 6-50 % filled by such a layer and the three GEMM launches of its training step cost
// 35 us on the MNIST MLP (a third of them launch / set-up / exposed-epilogue time) for 30 MFLOP; these kernels take
// the same operands in the same layouts (reference op: InnerProductLayer, src/worker/layer.cc:162-213):
//   forward : y[M][N]  = act(x[M][K] . W[K][N] + b)                      one warp per row, lanes over K
//   dgrad   : dx[M][K] = (dy[M][N] . W[K][N]^T) * act'(aux)              one block per row, threads over K
//   wgrad   : dW[K][N] (+)= x^T dy,  db[N] (+)= column sums of dy        16 k x 16 n x 4 row groups per block
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

__global__ void __launch_bounds__(256)
ip_small_fwd_kernel(const void* __restrict__ x, long long ldx, int x_bf16, const void* __restrict__ w, long long ldw,
                    int w_bf16, const void* __restrict__ bias, int bias_bf16, void* __restrict__ y, long long ldy,
                    int y_bf16, int M, int K, int N, int act) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= M) return;
  float acc[IPS_NMAX];
#pragma unroll
  for (int n = 0; n < IPS_NMAX; ++n) acc[n] = 0.f;
  for (int k = lane; k < K; k += 32) {
    const float xv = ld_any(x, row * ldx + k, x_bf16);
#pragma unroll
    for (int n = 0; n < IPS_NMAX; ++n)
      if (n < N) acc[n] = fmaf(xv, ld_any(w, k * ldw + n, w_bf16), acc[n]);
  }
#pragma unroll
  for (int n = 0; n < IPS_NMAX; ++n) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc[n] += __shfl_xor_sync(0xffffffffu, acc[n], o);
  }
  // lane n stores output n (a compile-time select chain keeps acc[] in registers)
  float v = 0.f;
#pragma unroll
  for (int n = 0; n < IPS_NMAX; ++n)
    if (lane == n) v = acc[n];
  if (lane < N) {
    if (bias != nullptr) v += ld_any(bias, lane, bias_bf16);
    st_any(y, row * ldy + lane, y_bf16, apply_act(v, act));
  }
}

__global__ void __launch_bounds__(128)
ip_small_dgrad_kernel(const void* __restrict__ dy, long long lddy, int dy_bf16, const void* __restrict__ w, long long ldw,
                      int w_bf16, const void* __restrict__ aux, long long ldaux, int aux_bf16, int grad_act,
                      void* __restrict__ dx, long long lddx, int dx_bf16, int M, int K, int N) {
  const int row = blockIdx.x;
  float g[IPS_NMAX];
#pragma unroll
  for (int n = 0; n < IPS_NMAX; ++n) g[n] = n < N ? ld_any(dy, row * lddy + n, dy_bf16) : 0.f;
  for (int k = threadIdx.x; k < K; k += blockDim.x) {
    float s = 0.f;
#pragma unroll
    for (int n = 0; n < IPS_NMAX; ++n)
      if (n < N) s = fmaf(g[n], ld_any(w, k * ldw + n, w_bf16), s);
    if (grad_act != SG_ACT_NONE) s *= act_grad_from_output(ld_any(aux, row * ldaux + k, aux_bf16), grad_act);
    st_any(dx, row * lddx + k, dx_bf16, s);
  }
}

// threadIdx.x = n (16), threadIdx.y = k within the block's strip (16), threadIdx.z = row group (4)
__global__ void __launch_bounds__(1024)
ip_small_wgrad_kernel(const void* __restrict__ x, long long ldx, int x_bf16, const void* __restrict__ dy, long long lddy,
                      int dy_bf16, float* __restrict__ dw, long long lddw, float* __restrict__ db, int M, int K, int N,
                      int accumulate) {
  __shared__ float part[4][16][IPS_NMAX + 1];
  __shared__ float bpart[4][IPS_NMAX];
  const int n = threadIdx.x, kk = threadIdx.y, grp = threadIdx.z;
  const int k = blockIdx.x * 16 + kk;
  const bool live = n < N && k < K;
  const bool do_bias = db != nullptr && blockIdx.x == 0 && kk == 0 && n < N;
  float acc = 0.f, bacc = 0.f;
  for (int m = grp; m < M; m += 4) {
    const float g = n < N ? ld_any(dy, m * lddy + n, dy_bf16) : 0.f;
    if (live) acc = fmaf(ld_any(x, m * ldx + k, x_bf16), g, acc);
    bacc += g;
  }
  part[grp][kk][n] = acc;
  if (kk == 0) bpart[grp][n] = bacc;
  __syncthreads();
  if (grp == 0) {
    if (live) {
      const float s = part[0][kk][n] + part[1][kk][n] + part[2][kk][n] + part[3][kk][n];
      float* d = dw + k * lddw + n;
      *d = accumulate ? *d + s : s;
    }
    if (do_bias) {
      const float s = bpart[0][n] + bpart[1][n] + bpart[2][n] + bpart[3][n];
      db[n] = accumulate ? db[n] + s : s;
    }
  }
}

}  // namespace sg

using namespace sg;

extern "C" int sg_ip_small_fwd(const void* x, long long ldx, int x_bf16, const void* w, long long ldw, int w_bf16,
                               const void* bias, int bias_bf16, void* y, long long ldy, int y_bf16, int M, int K, int N,
                               int act, void* stream) {
  SG_REQUIRE(x != nullptr && w != nullptr && y != nullptr && M > 0 && K > 0 && N > 0 && N <= IPS_NMAX,
             "sg_ip_small_fwd: needs 1..%d outputs", IPS_NMAX);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  ip_small_fwd_kernel<<<(M + 7) / 8, 256, 0, s>>>(x, ldx, x_bf16, w, ldw, w_bf16, bias, bias_bf16, y, ldy, y_bf16, M, K, N, act);
  SG_CHECK_LAUNCH("ip_small_fwd");
  return 0;
}

extern "C" int sg_ip_small_dgrad(const void* dy, long long lddy, int dy_bf16, const void* w, long long ldw, int w_bf16,
                                 const void* aux, long long ldaux, int aux_bf16, int grad_act, void* dx, long long lddx,
                                 int dx_bf16, int M, int K, int N, void* stream) {
  SG_REQUIRE(dy != nullptr && w != nullptr && dx != nullptr && M > 0 && K > 0 && N > 0 && N <= IPS_NMAX,
             "sg_ip_small_dgrad: needs 1..%d outputs", IPS_NMAX);
  SG_REQUIRE(grad_act == SG_ACT_NONE || aux != nullptr, "sg_ip_small_dgrad: grad_act needs aux");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  ip_small_dgrad_kernel<<<M, 128, 0, s>>>(dy, lddy, dy_bf16, w, ldw, w_bf16, aux, ldaux, aux_bf16, grad_act, dx, lddx, dx_bf16, M, K, N);
  SG_CHECK_LAUNCH("ip_small_dgrad");
  return 0;
}

extern "C" int sg_ip_small_wgrad(const void* x, long long ldx, int x_bf16, const void* dy, long long lddy, int dy_bf16,
                                 float* dw, long long lddw, float* db, int M, int K, int N, int accumulate,
                                 void* stream) {
  SG_REQUIRE(x != nullptr && dy != nullptr && dw != nullptr && M > 0 && K > 0 && N > 0 && N <= IPS_NMAX,
             "sg_ip_small_wgrad: needs 1..%d outputs", IPS_NMAX);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  ip_small_wgrad_kernel<<<(K + 15) / 16, dim3(16, 16, 4), 0, s>>>(x, ldx, x_bf16, dy, lddy, dy_bf16, dw, lddw, db, M, K, N, accumulate);
  SG_CHECK_LAUNCH("ip_small_wgrad");
  return 0;
}
