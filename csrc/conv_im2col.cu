// Convolution lowering for channels-last activations.
//
// The conv layers run as ONE batched GEMM on the tcgen05 kernel (M = N*OH*OW output pixels, N = filters,
// K = k*k*C with the channel innermost) instead of the reference's per-image im2col + SGEMM loop
// (src/worker/layer.cc:63-123).  This file provides the patch gather (im2col) that builds the K-major A
// operand in a caller-supplied scratch buffer, and the gather-form inverse (col2im) for the data gradient.
#include <stdint.h>

#include "../include/singa_b200.h"
#include "common.h"
#include "vec.cuh"

namespace sg {

// col[p][(ky*k + kx)*C + c] = x[n][oy*s - pad + ky][ox*s - pad + kx][c]   (zero outside the image)
// VEC = channels moved per thread (8 -> one 128-bit access; needs C % 8 == 0 and ld % 8 == 0).
template <int VEC>
__global__ void im2col_nhwc_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ col,
                                   long long ld, int N, int H, int W, int C, int OH, int OW, int k,
                                   int s, int pad) {
  const int cv = C / VEC;
  const long long total = static_cast<long long>(N) * OH * OW * k * k * cv;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += stride) {
    const int c = static_cast<int>(i % cv) * VEC;
    long long t = i / cv;
    const int kx = static_cast<int>(t % k); t /= k;
    const int ky = static_cast<int>(t % k); t /= k;
    const long long p = t;
    const int ox = static_cast<int>(t % OW); t /= OW;
    const int oy = static_cast<int>(t % OH);
    const int n = static_cast<int>(t / OH);
    const int yy = oy * s - pad + ky, xx = ox * s - pad + kx;
    const bool in = yy >= 0 && yy < H && xx >= 0 && xx < W;
    __nv_bfloat16* dst = col + p * ld + static_cast<long long>(ky * k + kx) * C + c;
    const __nv_bfloat16* src = x + ((static_cast<long long>(n) * H + yy) * W + xx) * C + c;
    if (VEC == 8) {
      *reinterpret_cast<uint4*>(dst) = in ? __ldg(reinterpret_cast<const uint4*>(src)) : make_uint4(0, 0, 0, 0);
    } else {
      *dst = in ? *src : __float2bfloat16_rn(0.f);
    }
  }
}

// dx[n][y][x][c] = sum over (ky,kx) with (y+pad-ky) % s == 0, (x+pad-kx) % s == 0 and the output pixel in
// range of dcol[p(n,oy,ox)][(ky*k+kx)*C + c]; optionally multiplied by act'(aux) (aux = previous layer's
// activation output, same layout as dx) so the ReLU/tanh backward of the layer below needs no own pass.
__device__ __forceinline__ float actgrad(float y, int act) {
  if (act == SG_ACT_RELU) return y > 0.f ? 1.f : 0.f;
  if (act == SG_ACT_STANH) return 1.7159047f * 0.66666667f - (0.66666667f / 1.7159047f) * y * y;
  return 1.f;
}

__global__ void col2im_nhwc_kernel(const __nv_bfloat16* __restrict__ dcol, long long ld,
                                   __nv_bfloat16* __restrict__ dx, int N, int H, int W, int C, int OH,
                                   int OW, int k, int s, int pad, int grad_act,
                                   const __nv_bfloat16* __restrict__ aux) {
  const long long total = static_cast<long long>(N) * H * W * C;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += stride) {
    const int c = static_cast<int>(i % C);
    long long t = i / C;
    const int xx = static_cast<int>(t % W); t /= W;
    const int yy = static_cast<int>(t % H);
    const int n = static_cast<int>(t / H);
    float g = 0.f;
    for (int ky = 0; ky < k; ++ky) {
      const int ty = yy + pad - ky;
      if (ty < 0 || ty % s != 0) continue;
      const int oy = ty / s;
      if (oy >= OH) continue;
      for (int kx = 0; kx < k; ++kx) {
        const int tx = xx + pad - kx;
        if (tx < 0 || tx % s != 0) continue;
        const int ox = tx / s;
        if (ox >= OW) continue;
        const long long p = (static_cast<long long>(n) * OH + oy) * OW + ox;
        g += __bfloat162float(dcol[p * ld + static_cast<long long>(ky * k + kx) * C + c]);
      }
    }
    if (grad_act != SG_ACT_NONE) g *= actgrad(__bfloat162float(aux[i]), grad_act);
    dx[i] = __float2bfloat16_rn(g);
  }
}

static inline int grid_for(long long n, int threads) {
  long long b = (n + threads - 1) / threads;
  const long long cap = static_cast<long long>(num_sms()) * 16;
  if (b > cap) b = cap;
  return static_cast<int>(b < 1 ? 1 : b);
}

}  // namespace sg

using namespace sg;

extern "C" int sg_im2col_nhwc(const void* x, void* col, long long ld, int N, int H, int W, int C, int OH,
                              int OW, int k, int stride, int pad, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  SG_REQUIRE(k > 0 && stride > 0 && pad >= 0, "sg_im2col_nhwc: bad geometry");
  SG_REQUIRE(ld >= static_cast<long long>(k) * k * C, "sg_im2col_nhwc: ld smaller than k*k*C");
  const bool vec = (C % 8 == 0) && (ld % 8 == 0) && (((uintptr_t)x & 15) == 0) && (((uintptr_t)col & 15) == 0);
  const long long total = static_cast<long long>(N) * OH * OW * k * k * (vec ? C / 8 : C);
  if (total <= 0) return 0;
  if (vec)
    im2col_nhwc_kernel<8><<<grid_for(total, 256), 256, 0, s>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)col, ld, N, H, W, C, OH, OW, k, stride, pad);
  else
    im2col_nhwc_kernel<1><<<grid_for(total, 256), 256, 0, s>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)col, ld, N, H, W, C, OH, OW, k, stride, pad);
  SG_CHECK_LAUNCH("im2col_nhwc");
  return 0;
}

extern "C" int sg_col2im_nhwc(const void* dcol, long long ld, void* dx, int N, int H, int W, int C,
                              int OH, int OW, int k, int stride, int pad, int grad_act, const void* aux,
                              void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  SG_REQUIRE(k > 0 && stride > 0 && pad >= 0, "sg_col2im_nhwc: bad geometry");
  SG_REQUIRE(grad_act == SG_ACT_NONE || aux != nullptr, "sg_col2im_nhwc: grad_act needs aux");
  const long long total = static_cast<long long>(N) * H * W * C;
  if (total <= 0) return 0;
  col2im_nhwc_kernel<<<grid_for(total, 256), 256, 0, s>>>((const __nv_bfloat16*)dcol, ld, (__nv_bfloat16*)dx, N, H, W, C, OH, OW, k,
                                                         stride, pad, grad_act, (const __nv_bfloat16*)aux);
  SG_CHECK_LAUNCH("col2im_nhwc");
  return 0;
}
