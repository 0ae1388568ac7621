// Softmax-loss (fwd+bwd fused), pooling and LRN for channels-last (NHWC) bf16/fp32 activations.
#include <cuda_bf16.h>
#include <float.h>
#include <stdint.h>

#include "../include/singa_b200.h"
#include "common.h"
#include "vec.cuh"

namespace sg {

// ------------------------------------------------------------------------------------------------
// Softmax loss.  One warp per row: max, exp-sum, NLL, rank of the label (top-k), gradient.
//   metric[0] += scale/N * -log(max(p_label, FLT_MIN));  metric[1] += scale/N * [rank(label) < topk]
//   grad[r][j] = (p_j - [j == label]) * scale / N
// Reference: SoftmaxLossLayer, src/worker/layer.cc:718-764 (its `label < 10` limit is not reproduced).
// ------------------------------------------------------------------------------------------------
template <typename G>
__global__ void softmax_loss_kernel(const float* __restrict__ logits, long long ldl,
                                    const int* __restrict__ labels, int rows, int dim, int topk,
                                    float scale, float* __restrict__ prob, long long ldp,
                                    G* __restrict__ grad, long long ldg, float* __restrict__ metric) {
  const int warps_per_block = blockDim.x >> 5;
  const int lane = threadIdx.x & 31;
  float loss_acc = 0.f, hit_acc = 0.f;
  for (int r = blockIdx.x * warps_per_block + (threadIdx.x >> 5); r < rows;
       r += gridDim.x * warps_per_block) {
    const float* x = logits + static_cast<long long>(r) * ldl;
    float mx = -FLT_MAX;
    for (int j = lane; j < dim; j += 32) mx = fmaxf(mx, x[j]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int j = lane; j < dim; j += 32) sum += __expf(x[j] - mx);
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
    int label = labels[r];
    label = label < 0 ? 0 : (label >= dim ? dim - 1 : label);
    const float xl = x[label];
    float rank = 0.f;
    for (int j = lane; j < dim; j += 32) rank += (x[j] > xl) ? 1.f : 0.f;
    rank = warp_sum(rank);
    const float pl = __expf(xl - mx) * inv;
    if (lane == 0) {
      loss_acc += -logf(fmaxf(pl, FLT_MIN));
      hit_acc += rank < static_cast<float>(topk) ? 1.f : 0.f;
    }
    const float gs = scale / rows;
    for (int j = lane; j < dim; j += 32) {
      const float pj = __expf(x[j] - mx) * inv;
      if (prob) prob[static_cast<long long>(r) * ldp + j] = pj;
      if (grad) grad[static_cast<long long>(r) * ldg + j] = from_float<G>((pj - (j == label ? 1.f : 0.f)) * gs);
    }
  }
  // block-level reduction of the two metrics, one atomic pair per block
  __shared__ float sl[32], sh[32];
  if (lane == 0) { sl[threadIdx.x >> 5] = loss_acc; sh[threadIdx.x >> 5] = hit_acc; }
  __syncthreads();
  if (threadIdx.x < 32) {
    float a = threadIdx.x < warps_per_block ? sl[threadIdx.x] : 0.f;
    float b = threadIdx.x < warps_per_block ? sh[threadIdx.x] : 0.f;
    a = warp_sum(a); b = warp_sum(b);
    if (threadIdx.x == 0 && metric) {
      atomicAdd(metric, a * scale / rows);
      atomicAdd(metric + 1, b * scale / rows);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Pooling, NHWC.  Window clipped at the bottom/right edge; AVE always divides by k*k
// (reference: mshadow pool/unpool, include/mshadow/tensor_expr_ext.h:790-848; layer.cc:512-540).
// Output size is computed by the caller (ceil form, see ops/pool.py).
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void pool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int N, int H, int W, int C,
                                int OH, int OW, int k, int s, int pad, int is_max) {
  const long long total = static_cast<long long>(N) * OH * OW * C;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += stride) {
    const int c = static_cast<int>(i % C);
    long long t = i / C;
    const int ox = static_cast<int>(t % OW); t /= OW;
    const int oy = static_cast<int>(t % OH);
    const int n = static_cast<int>(t / OH);
    const int y0 = oy * s - pad, x0 = ox * s - pad;
    const int y1 = min(y0 + k, H), x1 = min(x0 + k, W);
    float acc = is_max ? -FLT_MAX : 0.f;
    for (int yy = max(y0, 0); yy < y1; ++yy)
      for (int xx = max(x0, 0); xx < x1; ++xx) {
        const float v = to_float(x[((static_cast<long long>(n) * H + yy) * W + xx) * C + c]);
        acc = is_max ? fmaxf(acc, v) : acc + v;
      }
    if (!is_max) acc *= 1.f / (k * k);
    y[i] = from_float<T>(acc);
  }
}

// Gather-form backward (no atomics): every input element sums over the windows that contain it.
// MAX routes the gradient to every element equal to the window max (the reference's tie rule).
template <typename T>
__global__ void pool_bwd_kernel(const T* __restrict__ x, const T* __restrict__ y,
                                const T* __restrict__ dy, T* __restrict__ dx, int N, int H, int W,
                                int C, int OH, int OW, int k, int s, int pad, int is_max) {
  const long long total = static_cast<long long>(N) * H * W * C;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += stride) {
    const int c = static_cast<int>(i % C);
    long long t = i / C;
    const int xx = static_cast<int>(t % W); t /= W;
    const int yy = static_cast<int>(t % H);
    const int n = static_cast<int>(t / H);
    const float xv = to_float(x[i]);
    // windows oy with oy*s - pad <= yy < oy*s - pad + k
    const int oy_lo = max(0, (yy + pad - k + s) / s), oy_hi = min(OH - 1, (yy + pad) / s);
    const int ox_lo = max(0, (xx + pad - k + s) / s), ox_hi = min(OW - 1, (xx + pad) / s);
    float g = 0.f;
    for (int oy = oy_lo; oy <= oy_hi; ++oy)
      for (int ox = ox_lo; ox <= ox_hi; ++ox) {
        const long long o = ((static_cast<long long>(n) * OH + oy) * OW + ox) * C + c;
        if (is_max) {
          if (to_float(y[o]) == xv) g += to_float(dy[o]);
        } else {
          g += to_float(dy[o]);
        }
      }
    if (!is_max) g *= 1.f / (k * k);
    dx[i] = from_float<T>(g);
  }
}

// Vector forms: one block per (image, row), threads over (column, 16-byte channel vector); C % Vec<T>::N == 0.
// Only 32-bit index arithmetic, and the window is a compile-time bound so that every load of a pixel is issued
// before the first use (the loops above wait for each window element in turn: 5-13x slower on AlexNet's fp32 pools;
// several rows per block were tried for the short bf16 rows and lost to the extra index arithmetic).
//   forward : K = window size (2, 3) -> K*K predicated loads of x per output vector
//   backward: MW = ceil(k / stride) = most windows per dimension that contain one input pixel (1, 2, 3)
template <typename T, int K, bool MAX>
__global__ void __launch_bounds__(256)
pool_fwd_row_kernel(const T* __restrict__ x, T* __restrict__ y, int H, int W, int C, int OH, int OW, int s, int pad) {
  constexpr int V = Vec<T>::N;
  const int cv = C / V;
  const int n = blockIdx.x / OH, oy = blockIdx.x - n * OH;
  const int y0 = oy * s - pad;
  const T* xn = x + static_cast<long long>(n) * H * W * C;
  T* yrow = y + static_cast<long long>(blockIdx.x) * OW * C;
  for (int j = threadIdx.x; j < OW * cv; j += blockDim.x) {
    const int ox = j / cv, c = (j - ox * cv) * V;
    const int x0 = ox * s - pad;
    float v[K * K][V];
#pragma unroll
    for (int a = 0; a < K; ++a)
#pragma unroll
      for (int b = 0; b < K; ++b) {
        const int yy = y0 + a, xx = x0 + b;
        if (yy >= 0 && yy < H && xx >= 0 && xx < W) {
          Vec<T>::load(xn + (static_cast<long long>(yy) * W + xx) * C + c, v[a * K + b]);
        } else {
#pragma unroll
          for (int q = 0; q < V; ++q) v[a * K + b][q] = MAX ? -FLT_MAX : 0.f;
        }
      }
    float acc[V];
#pragma unroll
    for (int q = 0; q < V; ++q) acc[q] = v[0][q];
#pragma unroll
    for (int w = 1; w < K * K; ++w)
#pragma unroll
      for (int q = 0; q < V; ++q) acc[q] = MAX ? fmaxf(acc[q], v[w][q]) : acc[q] + v[w][q];
    if (!MAX) {
#pragma unroll
      for (int q = 0; q < V; ++q) acc[q] *= 1.f / (K * K);
    }
    Vec<T>::store(yrow + static_cast<long long>(ox) * C + c, acc);
  }
}

template <typename T, int MW, bool MAX>
__global__ void __launch_bounds__(256)
pool_bwd_row_kernel(const T* __restrict__ x, const T* __restrict__ y, const T* __restrict__ dy, T* __restrict__ dx,
                    int H, int W, int C, int OH, int OW, int k, int s, int pad) {
  constexpr int V = Vec<T>::N;
  const int cv = C / V;
  const int n = blockIdx.x / H, yy = blockIdx.x - n * H;
  const int oy_lo = max(0, (yy + pad - k + s) / s), oy_hi = min(OH - 1, (yy + pad) / s);
  const long long obase = static_cast<long long>(n) * OH * OW * C;
  const T* xrow = x + static_cast<long long>(blockIdx.x) * W * C;
  T* dxrow = dx + static_cast<long long>(blockIdx.x) * W * C;
  const float inv = 1.f / (k * k);
  for (int j = threadIdx.x; j < W * cv; j += blockDim.x) {
    const int xx = j / cv, c = (j - xx * cv) * V;
    const int ox_lo = max(0, (xx + pad - k + s) / s), ox_hi = min(OW - 1, (xx + pad) / s);
    float xv[V], gv[MW * MW][V], yv[MAX ? MW * MW : 1][V];
    if (MAX) Vec<T>::load(xrow + static_cast<long long>(xx) * C + c, xv);
#pragma unroll
    for (int a = 0; a < MW; ++a)
#pragma unroll
      for (int b = 0; b < MW; ++b) {
        const int oy = oy_lo + a, ox = ox_lo + b;
        if (oy <= oy_hi && ox <= ox_hi) {
          const long long o = obase + (static_cast<long long>(oy) * OW + ox) * C + c;
          Vec<T>::load(dy + o, gv[a * MW + b]);
          if (MAX) Vec<T>::load(y + o, yv[a * MW + b]);
        } else {
#pragma unroll
          for (int q = 0; q < V; ++q) {
            gv[a * MW + b][q] = 0.f;
            if (MAX) yv[a * MW + b][q] = FLT_MAX;     // never equal to a finite input
          }
        }
      }
    float g[V];
#pragma unroll
    for (int q = 0; q < V; ++q) g[q] = 0.f;
#pragma unroll
    for (int w = 0; w < MW * MW; ++w)
#pragma unroll
      for (int q = 0; q < V; ++q) g[q] += MAX ? (yv[w][q] == xv[q] ? gv[w][q] : 0.f) : gv[w][q];
    if (!MAX) {
#pragma unroll
      for (int q = 0; q < V; ++q) g[q] *= inv;
    }
    Vec<T>::store(dxrow + static_cast<long long>(xx) * C + c, g);
  }
}

// bf16 max pooling on packed pairs (HMNMX2 / HSET2 / HFMA2): the generic row kernels unpack every element to fp32 and
// were instruction bound at 23-26 % of the copy bandwidth (fp32: 63-76 %).  Same blocks and windows as above.  The
// backward sum of the (at most MW*MW) matching windows is accumulated in bf16 — one rounding per extra term, and only
// pixels shared by several windows that all have their maximum there have more than one term.
struct Bf16x8 {
  uint4 u;
  __device__ __forceinline__ __nv_bfloat162& h(int q) { return reinterpret_cast<__nv_bfloat162*>(&u)[q]; }
};

template <int K>
__global__ void __launch_bounds__(256)
pool_max_fwd_row_bf16_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, int H, int W, int C,
                             int OH, int OW, int s, int pad) {
  const int cv = C / 8;
  const int n = blockIdx.x / OH, oy = blockIdx.x - n * OH;
  const int y0 = oy * s - pad;
  const __nv_bfloat16* xn = x + static_cast<long long>(n) * H * W * C;
  __nv_bfloat16* yrow = y + static_cast<long long>(blockIdx.x) * OW * C;
  Bf16x8 lowest;
  lowest.u = make_uint4(0xFF7FFF7Fu, 0xFF7FFF7Fu, 0xFF7FFF7Fu, 0xFF7FFF7Fu);     // -3.39e38 in both halves
  for (int j = threadIdx.x; j < OW * cv; j += blockDim.x) {
    const int ox = j / cv, c = (j - ox * cv) * 8;
    const int x0 = ox * s - pad;
    Bf16x8 v[K * K];
#pragma unroll
    for (int a = 0; a < K; ++a)
#pragma unroll
      for (int b = 0; b < K; ++b) {
        const int yy = y0 + a, xx = x0 + b;
        if (yy >= 0 && yy < H && xx >= 0 && xx < W)
          v[a * K + b].u = *reinterpret_cast<const uint4*>(xn + (static_cast<long long>(yy) * W + xx) * C + c);
        else
          v[a * K + b] = lowest;
      }
    Bf16x8 acc = v[0];
#pragma unroll
    for (int w = 1; w < K * K; ++w)
#pragma unroll
      for (int q = 0; q < 4; ++q) acc.h(q) = __hmax2(acc.h(q), v[w].h(q));
    *reinterpret_cast<uint4*>(yrow + static_cast<long long>(ox) * C + c) = acc.u;
  }
}

template <int MW>
__global__ void __launch_bounds__(256)
pool_max_bwd_row_bf16_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ y,
                             const __nv_bfloat16* __restrict__ dy, __nv_bfloat16* __restrict__ dx, int H, int W, int C,
                             int OH, int OW, int k, int s, int pad) {
  const int cv = C / 8;
  const int n = blockIdx.x / H, yy = blockIdx.x - n * H;
  const int oy_lo = max(0, (yy + pad - k + s) / s), oy_hi = min(OH - 1, (yy + pad) / s);
  const long long obase = static_cast<long long>(n) * OH * OW * C;
  const __nv_bfloat16* xrow = x + static_cast<long long>(blockIdx.x) * W * C;
  __nv_bfloat16* dxrow = dx + static_cast<long long>(blockIdx.x) * W * C;
  for (int j = threadIdx.x; j < W * cv; j += blockDim.x) {
    const int xx = j / cv, c = (j - xx * cv) * 8;
    const int ox_lo = max(0, (xx + pad - k + s) / s), ox_hi = min(OW - 1, (xx + pad) / s);
    Bf16x8 xv, gv[MW * MW], yv[MW * MW];
    bool live[MW * MW];
    xv.u = *reinterpret_cast<const uint4*>(xrow + static_cast<long long>(xx) * C + c);
#pragma unroll
    for (int a = 0; a < MW; ++a)
#pragma unroll
      for (int b = 0; b < MW; ++b) {
        const int oy = oy_lo + a, ox = ox_lo + b;
        live[a * MW + b] = oy <= oy_hi && ox <= ox_hi;
        if (live[a * MW + b]) {
          const long long o = obase + (static_cast<long long>(oy) * OW + ox) * C + c;
          gv[a * MW + b].u = *reinterpret_cast<const uint4*>(dy + o);
          yv[a * MW + b].u = *reinterpret_cast<const uint4*>(y + o);
        }
      }
    Bf16x8 g;
    g.u = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
    for (int w = 0; w < MW * MW; ++w)
      if (live[w]) {
#pragma unroll
        for (int q = 0; q < 4; ++q) g.h(q) = __hfma2(__heq2(yv[w].h(q), xv.h(q)), gv[w].h(q), g.h(q));
      }
    *reinterpret_cast<uint4*>(dxrow + static_cast<long long>(xx) * C + c) = g.u;
  }
}

template <typename T>
static bool launch_pool_fwd_rows(const T* x, T* y, int N, int H, int W, int C, int OH, int OW, int k, int st, int pad,
                                 int is_max, cudaStream_t s) {
  const int grid = N * OH;
#define SG_POOL_F(K_, M_) pool_fwd_row_kernel<T, K_, M_><<<grid, 256, 0, s>>>(x, y, H, W, C, OH, OW, st, pad)
  if (sizeof(T) == 2 && is_max && (k == 2 || k == 3)) {
    const __nv_bfloat16* xb = reinterpret_cast<const __nv_bfloat16*>(x);
    __nv_bfloat16* yb = reinterpret_cast<__nv_bfloat16*>(y);
    if (k == 2) pool_max_fwd_row_bf16_kernel<2><<<grid, 256, 0, s>>>(xb, yb, H, W, C, OH, OW, st, pad);
    else pool_max_fwd_row_bf16_kernel<3><<<grid, 256, 0, s>>>(xb, yb, H, W, C, OH, OW, st, pad);
    return true;
  }
  if (k == 2) { if (is_max) SG_POOL_F(2, true); else SG_POOL_F(2, false); return true; }
  if (k == 3) { if (is_max) SG_POOL_F(3, true); else SG_POOL_F(3, false); return true; }
#undef SG_POOL_F
  return false;
}

template <typename T>
static bool launch_pool_bwd_rows(const T* x, const T* y, const T* dy, T* dx, int N, int H, int W, int C, int OH, int OW,
                                 int k, int st, int pad, int is_max, cudaStream_t s) {
  const int grid = N * H, mw = (k + st - 1) / st;
#define SG_POOL_B(W_, M_) pool_bwd_row_kernel<T, W_, M_><<<grid, 256, 0, s>>>(x, y, dy, dx, H, W, C, OH, OW, k, st, pad)
  if (sizeof(T) == 2 && is_max && mw >= 1 && mw <= 3) {
    const __nv_bfloat16* xb = reinterpret_cast<const __nv_bfloat16*>(x);
    const __nv_bfloat16* yb = reinterpret_cast<const __nv_bfloat16*>(y);
    const __nv_bfloat16* gb = reinterpret_cast<const __nv_bfloat16*>(dy);
    __nv_bfloat16* db = reinterpret_cast<__nv_bfloat16*>(dx);
    if (mw == 1) pool_max_bwd_row_bf16_kernel<1><<<grid, 256, 0, s>>>(xb, yb, gb, db, H, W, C, OH, OW, k, st, pad);
    else if (mw == 2) pool_max_bwd_row_bf16_kernel<2><<<grid, 256, 0, s>>>(xb, yb, gb, db, H, W, C, OH, OW, k, st, pad);
    else pool_max_bwd_row_bf16_kernel<3><<<grid, 256, 0, s>>>(xb, yb, gb, db, H, W, C, OH, OW, k, st, pad);
    return true;
  }
  if (mw == 1) { if (is_max) SG_POOL_B(1, true); else SG_POOL_B(1, false); return true; }
  if (mw == 2) { if (is_max) SG_POOL_B(2, true); else SG_POOL_B(2, false); return true; }
  if (mw == 3) { if (is_max) SG_POOL_B(3, true); else SG_POOL_B(3, false); return true; }
#undef SG_POOL_B
  return false;
}

// ------------------------------------------------------------------------------------------------
// LRN across channels, NHWC.  norm = knorm + (alpha/n) * sum_{c' in [c-n/2, c+n/2]} x^2; y = x * norm^-beta.
// Backward: dx = dy*norm^-beta - (2*beta*alpha/n) * x * sum_{c'} dy*x*norm^(-beta-1)... evaluated at c'.
// Reference: LRNLayer, src/worker/layer.cc:356-378; window clipping tensor_expr_ext.h:929-934.
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void lrn_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, float* __restrict__ norm,
                               long long pixels, int C, int lsize, float alpha, float beta, float knorm) {
  const long long total = pixels * C;
  const int half = lsize / 2;
  const float salpha = alpha / lsize;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += stride) {
    const int c = static_cast<int>(i % C);
    const long long base = i - c;
    float s = 0.f;
    for (int cc = max(0, c - half); cc <= min(C - 1, c + half); ++cc) {
      const float v = to_float(x[base + cc]);
      s += v * v;
    }
    const float nm = knorm + salpha * s;
    norm[i] = nm;
    y[i] = from_float<T>(to_float(x[i]) * __powf(nm, -beta));
  }
}

template <typename T>
__global__ void lrn_bwd_kernel(const T* __restrict__ x, const float* __restrict__ norm,
                               const T* __restrict__ dy, T* __restrict__ dx, long long pixels, int C,
                               int lsize, float alpha, float beta) {
  const long long total = pixels * C;
  const int half = lsize / 2;
  const float coef = 2.f * beta * alpha / lsize;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += stride) {
    const int c = static_cast<int>(i % C);
    const long long base = i - c;
    float s = 0.f;
    for (int cc = max(0, c - half); cc <= min(C - 1, c + half); ++cc) {
      const float nm = norm[base + cc];
      s += to_float(dy[base + cc]) * to_float(x[base + cc]) * __powf(nm, -beta - 1.f);
    }
    const float g = to_float(dy[i]) * __powf(norm[i], -beta) - coef * to_float(x[i]) * s;
    dx[i] = from_float<T>(g);
  }
}

// Tiled forms (C % Vec<T>::N == 0, C <= LRN_TILE): a block stages whole pixels in shared memory with 16-byte loads,
// then every thread sums its channel window from shared memory — each element is read from global memory once and
// raised to a power once (the loops above re-read x / dy / norm and call __powf for every window element).
constexpr int LRN_TILE = 2048;

// w[0..11] = s[e-4 .. e+7] of one pixel's channel row (c = channel of s[e], a multiple of 4); channels outside
// [0, C) read as zero.  Window half-widths up to 4 (local size <= 9) are served from these three vectors.
__device__ __forceinline__ void lrn_window(const float* s, int e, int c, int C, float (&w)[12]) {
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  const float4 lo = c >= 4 ? *reinterpret_cast<const float4*>(&s[e - 4]) : z;
  const float4 mi = *reinterpret_cast<const float4*>(&s[e]);
  const float4 hi = c + 4 < C ? *reinterpret_cast<const float4*>(&s[e + 4]) : z;
  w[0] = lo.x; w[1] = lo.y; w[2] = lo.z; w[3] = lo.w;
  w[4] = mi.x; w[5] = mi.y; w[6] = mi.z; w[7] = mi.w;
  w[8] = hi.x; w[9] = hi.y; w[10] = hi.z; w[11] = hi.w;
}
__device__ __forceinline__ void store4(float* p, const float (&v)[4]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
}
__device__ __forceinline__ void store4(__nv_bfloat16* p, const float (&v)[4]) {
  const __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
  *reinterpret_cast<uint2*>(p) = make_uint2(*reinterpret_cast<const uint32_t*>(&a), *reinterpret_cast<const uint32_t*>(&b));
}

template <typename T>
__global__ void __launch_bounds__(256)
lrn_fwd_tile_kernel(const T* __restrict__ x, T* __restrict__ y, float* __restrict__ norm, long long pixels, int C,
                    int ppt, int half, float salpha, float beta, float knorm) {
  constexpr int V = Vec<T>::N;
  __shared__ __align__(16) float sx[LRN_TILE];
  for (long long tile = blockIdx.x; tile * ppt < pixels; tile += gridDim.x) {
    const long long p0 = tile * ppt, base = p0 * C;
    const int np = static_cast<int>(min(static_cast<long long>(ppt), pixels - p0)), ne = np * C;
    for (int e = threadIdx.x * V; e < ne; e += blockDim.x * V) {
      float xv[V];
      Vec<T>::load(x + base + e, xv);
#pragma unroll
      for (int q = 0; q < V; q += 4)
        *reinterpret_cast<float4*>(&sx[e + q]) = make_float4(xv[q], xv[q + 1], xv[q + 2], xv[q + 3]);
    }
    __syncthreads();
    for (int e = threadIdx.x * 4; e < ne; e += blockDim.x * 4) {
      float w[12];
      lrn_window(sx, e, e % C, C, w);
#pragma unroll
      for (int q = 0; q < 12; ++q) w[q] *= w[q];
      float nm[4], o[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float sum = w[4 + q];
#pragma unroll
        for (int d = 1; d <= 4; ++d)
          if (d <= half) sum += w[4 + q - d] + w[4 + q + d];
        nm[q] = knorm + salpha * sum;
        o[q] = sx[e + q] * __powf(nm[q], -beta);
      }
      *reinterpret_cast<float4*>(norm + base + e) = make_float4(nm[0], nm[1], nm[2], nm[3]);
      store4(y + base + e, o);
    }
    __syncthreads();
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
lrn_bwd_tile_kernel(const T* __restrict__ x, const float* __restrict__ norm, const T* __restrict__ dy,
                    T* __restrict__ dx, long long pixels, int C, int ppt, int half, float coef, float beta) {
  constexpr int V = Vec<T>::N;
  __shared__ __align__(16) float st[LRN_TILE], sa[LRN_TILE], sx[LRN_TILE];
  for (long long tile = blockIdx.x; tile * ppt < pixels; tile += gridDim.x) {
    const long long p0 = tile * ppt, base = p0 * C;
    const int np = static_cast<int>(min(static_cast<long long>(ppt), pixels - p0)), ne = np * C;
    for (int e = threadIdx.x * V; e < ne; e += blockDim.x * V) {
      float xv[V], gv[V], nv[V], t[V], a[V];
      Vec<T>::load(x + base + e, xv);
      Vec<T>::load(dy + base + e, gv);
#pragma unroll
      for (int q = 0; q < V; q += 4) {
        const float4 n4 = *reinterpret_cast<const float4*>(norm + base + e + q);
        nv[q] = n4.x; nv[q + 1] = n4.y; nv[q + 2] = n4.z; nv[q + 3] = n4.w;
      }
#pragma unroll
      for (int q = 0; q < V; ++q) {
        a[q] = gv[q] * __powf(nv[q], -beta);          // dy * norm^-beta
        t[q] = a[q] * xv[q] / nv[q];                  // dy * x * norm^(-beta-1)
      }
#pragma unroll
      for (int q = 0; q < V; q += 4) {
        *reinterpret_cast<float4*>(&st[e + q]) = make_float4(t[q], t[q + 1], t[q + 2], t[q + 3]);
        *reinterpret_cast<float4*>(&sa[e + q]) = make_float4(a[q], a[q + 1], a[q + 2], a[q + 3]);
        *reinterpret_cast<float4*>(&sx[e + q]) = make_float4(xv[q], xv[q + 1], xv[q + 2], xv[q + 3]);
      }
    }
    __syncthreads();
    for (int e = threadIdx.x * 4; e < ne; e += blockDim.x * 4) {
      float w[12], o[4];
      lrn_window(st, e, e % C, C, w);
      const float4 a4 = *reinterpret_cast<const float4*>(&sa[e]);
      const float4 x4 = *reinterpret_cast<const float4*>(&sx[e]);
      const float av[4] = {a4.x, a4.y, a4.z, a4.w}, xv[4] = {x4.x, x4.y, x4.z, x4.w};
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float sum = w[4 + q];
#pragma unroll
        for (int d = 1; d <= 4; ++d)
          if (d <= half) sum += w[4 + q - d] + w[4 + q + d];
        o[q] = av[q] - coef * xv[q] * sum;
      }
      store4(dx + base + e, o);
    }
    __syncthreads();
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

extern "C" int sg_softmax_loss(const float* logits, long long ldl, const int* labels, int rows, int dim,
                               int topk, float scale, float* prob, long long ldp, void* grad,
                               long long ldg, int grad_bf16, float* metric, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (rows <= 0) return 0;
  SG_REQUIRE(dim > 0 && topk > 0, "sg_softmax_loss: dim and topk must be positive");
  const int threads = 256, wpb = threads / 32;
  int blocks = (rows + wpb - 1) / wpb;
  if (blocks > num_sms() * 4) blocks = num_sms() * 4;
  if (grad_bf16)
    softmax_loss_kernel<__nv_bfloat16><<<blocks, threads, 0, s>>>(logits, ldl, labels, rows, dim, topk, scale, prob, ldp,
                                                                   (__nv_bfloat16*)grad, ldg, metric);
  else
    softmax_loss_kernel<float><<<blocks, threads, 0, s>>>(logits, ldl, labels, rows, dim, topk, scale, prob, ldp,
                                                           (float*)grad, ldg, metric);
  SG_CHECK_LAUNCH("softmax_loss");
  return 0;
}

extern "C" int sg_pool_fwd(const void* x, void* y, int N, int H, int W, int C, int OH, int OW, int k,
                           int stride, int pad, int is_max, int is_bf16, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const long long total = static_cast<long long>(N) * OH * OW * C;
  if (total <= 0) return 0;
  SG_REQUIRE(k > 0 && stride > 0 && pad >= 0 && pad < k, "sg_pool_fwd: bad window");
  const bool aligned = (((uintptr_t)x | (uintptr_t)y) & 15) == 0;
  bool done = false;
  if (aligned && is_bf16 && C % 8 == 0)
    done = launch_pool_fwd_rows((const __nv_bfloat16*)x, (__nv_bfloat16*)y, N, H, W, C, OH, OW, k, stride, pad, is_max, s);
  else if (aligned && !is_bf16 && C % 4 == 0)
    done = launch_pool_fwd_rows((const float*)x, (float*)y, N, H, W, C, OH, OW, k, stride, pad, is_max, s);
  if (done) {
  } else if (is_bf16)
    pool_fwd_kernel<__nv_bfloat16><<<grid_for(total, 256), 256, 0, s>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, N, H, W, C, OH, OW, k, stride, pad, is_max);
  else
    pool_fwd_kernel<float><<<grid_for(total, 256), 256, 0, s>>>((const float*)x, (float*)y, N, H, W, C, OH, OW, k, stride, pad, is_max);
  SG_CHECK_LAUNCH("pool_fwd");
  return 0;
}

extern "C" int sg_pool_bwd(const void* x, const void* y, const void* dy, void* dx, int N, int H, int W,
                           int C, int OH, int OW, int k, int stride, int pad, int is_max, int is_bf16,
                           void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const long long total = static_cast<long long>(N) * H * W * C;
  if (total <= 0) return 0;
  SG_REQUIRE(k > 0 && stride > 0 && pad >= 0 && pad < k, "sg_pool_bwd: bad window");
  const bool aligned = (((uintptr_t)x | (uintptr_t)y | (uintptr_t)dy | (uintptr_t)dx) & 15) == 0;
  bool done = false;
  if (aligned && is_bf16 && C % 8 == 0)
    done = launch_pool_bwd_rows((const __nv_bfloat16*)x, (const __nv_bfloat16*)y, (const __nv_bfloat16*)dy, (__nv_bfloat16*)dx, N, H, W, C, OH, OW, k, stride, pad, is_max, s);
  else if (aligned && !is_bf16 && C % 4 == 0)
    done = launch_pool_bwd_rows((const float*)x, (const float*)y, (const float*)dy, (float*)dx, N, H, W, C, OH, OW, k, stride, pad, is_max, s);
  if (done) {
  } else if (is_bf16)
    pool_bwd_kernel<__nv_bfloat16><<<grid_for(total, 256), 256, 0, s>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)y, (const __nv_bfloat16*)dy, (__nv_bfloat16*)dx, N, H, W, C, OH, OW, k, stride, pad, is_max);
  else
    pool_bwd_kernel<float><<<grid_for(total, 256), 256, 0, s>>>((const float*)x, (const float*)y, (const float*)dy, (float*)dx, N, H, W, C, OH, OW, k, stride, pad, is_max);
  SG_CHECK_LAUNCH("pool_bwd");
  return 0;
}

extern "C" int sg_lrn_fwd(const void* x, void* y, float* norm, long long pixels, int C, int lsize,
                          float alpha, float beta, float knorm, int is_bf16, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const long long total = pixels * C;
  if (total <= 0) return 0;
  SG_REQUIRE(lsize > 0 && (lsize & 1), "sg_lrn_fwd: local size must be odd and positive");
  const int vn = is_bf16 ? 8 : 4;
  if (C % vn == 0 && C <= LRN_TILE && lsize <= 9 && ((((uintptr_t)x | (uintptr_t)y | (uintptr_t)norm)) & 15) == 0) {
    const int ppt = LRN_TILE / C;
    const int grid = grid_for((pixels + ppt - 1) / ppt * 256, 256);
    if (is_bf16)
      lrn_fwd_tile_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, norm, pixels, C, ppt, lsize / 2, alpha / lsize, beta, knorm);
    else
      lrn_fwd_tile_kernel<float><<<grid, 256, 0, s>>>((const float*)x, (float*)y, norm, pixels, C, ppt, lsize / 2, alpha / lsize, beta, knorm);
  } else if (is_bf16)
    lrn_fwd_kernel<__nv_bfloat16><<<grid_for(total, 256), 256, 0, s>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)y, norm, pixels, C, lsize, alpha, beta, knorm);
  else
    lrn_fwd_kernel<float><<<grid_for(total, 256), 256, 0, s>>>((const float*)x, (float*)y, norm, pixels, C, lsize, alpha, beta, knorm);
  SG_CHECK_LAUNCH("lrn_fwd");
  return 0;
}

extern "C" int sg_lrn_bwd(const void* x, const float* norm, const void* dy, void* dx, long long pixels,
                          int C, int lsize, float alpha, float beta, int is_bf16, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const long long total = pixels * C;
  if (total <= 0) return 0;
  SG_REQUIRE(lsize > 0 && (lsize & 1), "sg_lrn_bwd: local size must be odd and positive");
  const int vn = is_bf16 ? 8 : 4;
  if (C % vn == 0 && C <= LRN_TILE && lsize <= 9 && ((((uintptr_t)x | (uintptr_t)dy | (uintptr_t)dx | (uintptr_t)norm)) & 15) == 0) {
    const int ppt = LRN_TILE / C;
    const int grid = grid_for((pixels + ppt - 1) / ppt * 256, 256);
    const float coef = 2.f * beta * alpha / lsize;
    if (is_bf16)
      lrn_bwd_tile_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>((const __nv_bfloat16*)x, norm, (const __nv_bfloat16*)dy, (__nv_bfloat16*)dx, pixels, C, ppt, lsize / 2, coef, beta);
    else
      lrn_bwd_tile_kernel<float><<<grid, 256, 0, s>>>((const float*)x, norm, (const float*)dy, (float*)dx, pixels, C, ppt, lsize / 2, coef, beta);
  } else if (is_bf16)
    lrn_bwd_kernel<__nv_bfloat16><<<grid_for(total, 256), 256, 0, s>>>((const __nv_bfloat16*)x, norm, (const __nv_bfloat16*)dy, (__nv_bfloat16*)dx, pixels, C, lsize, alpha, beta);
  else
    lrn_bwd_kernel<float><<<grid_for(total, 256), 256, 0, s>>>((const float*)x, norm, (const float*)dy, (float*)dx, pixels, C, lsize, alpha, beta);
  SG_CHECK_LAUNCH("lrn_bwd");
  return 0;
}
