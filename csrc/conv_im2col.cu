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
//
// With channels-last input, one (pixel, ky) "row" of the column matrix — k*C elements — is a CONTIGUOUS run of the
// input (the k horizontally adjacent pixels with all their channels).  A warp-row (32 lanes, blockDim.y rows per
// block) copies one such run; the pixel decomposition and the border clipping are computed once per run, not per
// element.  VEC = elements per lane access (8 -> 128-bit; needs C % 8 == 0 and ld % 8 == 0).
template <typename T, int VEC>
__global__ void __launch_bounds__(256)
im2col_nhwc_kernel(const T* __restrict__ x, T* __restrict__ col, long long ld, int N,
                   int H, int W, int C, int OH, int OW, int k, int s, int pad) {
  const long long rows = static_cast<long long>(N) * OH * OW * k;
  const int row_len = k * C;
  const long long row_stride = static_cast<long long>(gridDim.x) * blockDim.y;
  for (long long row = static_cast<long long>(blockIdx.x) * blockDim.y + threadIdx.y; row < rows; row += row_stride) {
    const int ky = static_cast<int>(row % k);
    const long long p = row / k;
    long long t = p;
    const int ox = static_cast<int>(t % OW); t /= OW;
    const int oy = static_cast<int>(t % OH);
    const int n = static_cast<int>(t / OH);
    const int yy = oy * s - pad + ky;
    const int x0 = ox * s - pad;
    // elements [e_lo, e_hi) of the run fall inside the image
    int e_lo = 0, e_hi = 0;
    if (yy >= 0 && yy < H) {
      e_lo = max(0, -x0) * C;
      e_hi = min(k, W - x0) * C;
    }
    T* dst = col + p * ld + static_cast<long long>(ky) * row_len;
    const T* src = x + ((static_cast<long long>(n) * H + yy) * W + x0) * C;
    if (VEC == 8) {
      for (int e = threadIdx.x * 8; e < row_len; e += 32 * 8) {
        const bool in = e >= e_lo && e < e_hi;   // C % 8 == 0: a vector never straddles the border
        *reinterpret_cast<uint4*>(dst + e) = in ? __ldg(reinterpret_cast<const uint4*>(src + e)) : make_uint4(0, 0, 0, 0);
      }
    } else {
      for (int e = threadIdx.x; e < row_len; e += 32)
        dst[e] = (e >= e_lo && e < e_hi) ? src[e] : T(0.f);
    }
  }
}

// Channel counts that are not a multiple of 8 (the RGB input of a first convolution): the run-per-warp kernel above
// degenerates to 2-byte accesses with half of the lanes idle (measured 1.7 ms for AlexNet conv1, 27 % of the training
// step).  Here every thread builds ONE 16-byte vector of the column matrix — 8 consecutive k-indices of one output
// pixel — from scalar (L1-resident) loads, walking (ky, kx, c) incrementally, and stores it with a single coalesced
// 128-bit access.  Needs ld % 8 == 0 and a 16-byte aligned col (true for the padded column buffers).
template <typename T>
__global__ void __launch_bounds__(256)
im2col_nhwc_gather8_kernel(const T* __restrict__ x, T* __restrict__ col, int ld, int N, int H,
                           int W, int C, int OH, int OW, int k, int s, int pad, unsigned total_vecs) {
  constexpr int EPV = 16 / sizeof(T);            // elements per 16-byte vector (8 bf16, 4 fp32)
  const int vpr = ld / EPV;                      // vectors per column-matrix row
  const int K = k * k * C;
  const unsigned stride = gridDim.x * blockDim.x;
  for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < total_vecs; t += stride) {
    const unsigned p = t / vpr;
    const int j0 = static_cast<int>(t - p * vpr) * EPV;
    const int ox = static_cast<int>(p % OW);
    const unsigned q = p / OW;
    const int oy = static_cast<int>(q % OH);
    const int n = static_cast<int>(q / OH);
    int kk = j0 / C;
    int c = j0 - kk * C;
    int ky = kk / k;
    int kx = kk - ky * k;
    const int y0 = oy * s - pad, x0 = ox * s - pad;
    const T* img = x + static_cast<long long>(n) * H * W * C;
    __align__(16) T v[EPV];
#pragma unroll
    for (int e = 0; e < EPV; ++e) {
      const int yy = y0 + ky, xx = x0 + kx;
      const bool in = (j0 + e < K) && yy >= 0 && yy < H && xx >= 0 && xx < W;
      v[e] = in ? img[(static_cast<long long>(yy) * W + xx) * C + c] : T(0.f);
      if (++c == C) { c = 0; if (++kx == k) { kx = 0; ++ky; } }
    }
    *reinterpret_cast<uint4*>(col + static_cast<long long>(p) * ld + j0) = *reinterpret_cast<const uint4*>(v);
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

template <typename T>
__global__ void col2im_nhwc_kernel(const T* __restrict__ dcol, long long ld,
                                   T* __restrict__ dx, int N, int H, int W, int C, int OH,
                                   int OW, int k, int s, int pad, int grad_act,
                                   const T* __restrict__ aux) {
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
        g += static_cast<float>(dcol[p * ld + static_cast<long long>(ky * k + kx) * C + c]);
      }
    }
    if (grad_act != SG_ACT_NONE) g *= actgrad(static_cast<float>(aux[i]), grad_act);
    dx[i] = T(g);
  }
}

// 8 channels per thread (128-bit loads of dcol, fp32 accumulation): needs C % 8 == 0 and ld % 8 == 0.
__global__ void __launch_bounds__(256)
col2im_nhwc_vec8_kernel(const __nv_bfloat16* __restrict__ dcol, long long ld, __nv_bfloat16* __restrict__ dx,
                        int N, int H, int W, int C, int OH, int OW, int k, int s, int pad, int grad_act,
                        const __nv_bfloat16* __restrict__ aux) {
  const int cv = C / 8;
  const long long total = static_cast<long long>(N) * H * W * cv;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += stride) {
    const int c = static_cast<int>(i % cv) * 8;
    long long t = i / cv;
    const int xx = static_cast<int>(t % W); t /= W;
    const int yy = static_cast<int>(t % H);
    const int n = static_cast<int>(t / H);
    float g[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    // taps (ky,kx) with (yy + pad - ky) = oy*s and (xx + pad - kx) = ox*s, oy/ox in range
    for (int ky = (yy + pad) % s; ky < k; ky += s) {
      const int ty = yy + pad - ky;
      if (ty < 0) break;
      const int oy = ty / s;
      if (oy >= OH) continue;
      for (int kx = (xx + pad) % s; kx < k; kx += s) {
        const int tx = xx + pad - kx;
        if (tx < 0) break;
        const int ox = tx / s;
        if (ox >= OW) continue;
        const long long p = (static_cast<long long>(n) * OH + oy) * OW + ox;
        float v[8];
        Vec<__nv_bfloat16>::load(dcol + p * ld + static_cast<long long>(ky * k + kx) * C + c, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] += v[j];
      }
    }
    const long long o = ((static_cast<long long>(n) * H + yy) * W + xx) * C + c;
    if (grad_act != SG_ACT_NONE) {
      float a[8];
      Vec<__nv_bfloat16>::load(aux + o, a);
#pragma unroll
      for (int j = 0; j < 8; ++j) g[j] *= actgrad(a[j], grad_act);
    }
    Vec<__nv_bfloat16>::store(dx + o, g);
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

extern "C" int sg_im2col(const void* x, void* col, long long ld, int N, int H, int W, int C, int OH,
                         int OW, int k, int stride, int pad, int is_bf16, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  SG_REQUIRE(k > 0 && stride > 0 && pad >= 0, "sg_im2col: bad geometry");
  SG_REQUIRE(ld >= static_cast<long long>(k) * k * C, "sg_im2col: ld smaller than k*k*C");
  const int epv = is_bf16 ? 8 : 4;
  const bool vec = is_bf16 && (C % 8 == 0) && (ld % 8 == 0) && (((uintptr_t)x & 15) == 0) && (((uintptr_t)col & 15) == 0);
  const long long rows = static_cast<long long>(N) * OH * OW * k;
  if (rows <= 0) return 0;
  const dim3 block(32, 8);
  long long blocks = (rows + 7) / 8;
  const long long cap = static_cast<long long>(num_sms()) * 16;
  if (blocks > cap) blocks = cap;
  const long long total_vecs = static_cast<long long>(N) * OH * OW * (ld / epv);
  const bool gather_ok = (ld % epv == 0) && (((uintptr_t)col & 15) == 0) && ld < (1LL << 30) && total_vecs < (1LL << 32);
  const long long want = (total_vecs + 255) / 256;
  const long long capv = static_cast<long long>(num_sms()) * 32;
  const unsigned ggrid = static_cast<unsigned>(want < capv ? want : capv);
  if (vec) {
    im2col_nhwc_kernel<__nv_bfloat16, 8><<<static_cast<unsigned>(blocks), block, 0, s>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)col, ld, N, H, W, C, OH, OW, k, stride, pad);
  } else if (gather_ok && is_bf16) {
    im2col_nhwc_gather8_kernel<__nv_bfloat16><<<ggrid, 256, 0, s>>>(
        (const __nv_bfloat16*)x, (__nv_bfloat16*)col, static_cast<int>(ld), N, H, W, C, OH, OW, k, stride, pad,
        static_cast<unsigned>(total_vecs));
  } else if (gather_ok) {
    im2col_nhwc_gather8_kernel<float><<<ggrid, 256, 0, s>>>(
        (const float*)x, (float*)col, static_cast<int>(ld), N, H, W, C, OH, OW, k, stride, pad,
        static_cast<unsigned>(total_vecs));
  } else if (is_bf16) {
    im2col_nhwc_kernel<__nv_bfloat16, 1><<<static_cast<unsigned>(blocks), block, 0, s>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)col, ld, N, H, W, C, OH, OW, k, stride, pad);
  } else {
    im2col_nhwc_kernel<float, 1><<<static_cast<unsigned>(blocks), block, 0, s>>>((const float*)x, (float*)col, ld, N, H, W, C, OH, OW, k, stride, pad);
  }
  SG_CHECK_LAUNCH("im2col");
  return 0;
}

extern "C" int sg_im2col_nhwc(const void* x, void* col, long long ld, int N, int H, int W, int C, int OH,
                              int OW, int k, int stride, int pad, void* stream) {
  return sg_im2col(x, col, ld, N, H, W, C, OH, OW, k, stride, pad, 1, stream);
}

extern "C" int sg_col2im(const void* dcol, long long ld, void* dx, int N, int H, int W, int C,
                         int OH, int OW, int k, int stride, int pad, int grad_act, const void* aux,
                         int is_bf16, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  SG_REQUIRE(k > 0 && stride > 0 && pad >= 0, "sg_col2im: bad geometry");
  SG_REQUIRE(grad_act == SG_ACT_NONE || aux != nullptr, "sg_col2im: grad_act needs aux");
  const long long total = static_cast<long long>(N) * H * W * C;
  if (total <= 0) return 0;
  const bool vec = is_bf16 && (C % 8 == 0) && (ld % 8 == 0) && (((uintptr_t)dcol & 15) == 0) && (((uintptr_t)dx & 15) == 0) &&
                   (aux == nullptr || ((uintptr_t)aux & 15) == 0);
  if (vec)
    col2im_nhwc_vec8_kernel<<<grid_for(total / 8, 256), 256, 0, s>>>((const __nv_bfloat16*)dcol, ld, (__nv_bfloat16*)dx, N, H, W, C,
                                                                     OH, OW, k, stride, pad, grad_act, (const __nv_bfloat16*)aux);
  else if (is_bf16)
    col2im_nhwc_kernel<__nv_bfloat16><<<grid_for(total, 256), 256, 0, s>>>((const __nv_bfloat16*)dcol, ld, (__nv_bfloat16*)dx, N, H, W, C, OH, OW, k,
                                                                          stride, pad, grad_act, (const __nv_bfloat16*)aux);
  else
    col2im_nhwc_kernel<float><<<grid_for(total, 256), 256, 0, s>>>((const float*)dcol, ld, (float*)dx, N, H, W, C, OH, OW, k,
                                                                  stride, pad, grad_act, (const float*)aux);
  SG_CHECK_LAUNCH("col2im");
  return 0;
}

extern "C" int sg_col2im_nhwc(const void* dcol, long long ld, void* dx, int N, int H, int W, int C,
                              int OH, int OW, int k, int stride, int pad, int grad_act, const void* aux,
                              void* stream) {
  return sg_col2im(dcol, ld, dx, N, H, W, C, OH, OW, k, stride, pad, grad_act, aux, 1, stream);
}
