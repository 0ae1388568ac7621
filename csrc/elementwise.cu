// Memory-bound element-wise and reduction kernels (the reference lowers all of these to mshadow's
// MapPlanKernel / MapRedKeepLowestKernel, include/mshadow/cuda/tensor_gpu-inl.cuh:56-168).
// Activations are bf16 (or fp32) in HBM, math is fp32 in registers, accesses are 128-bit where the
// buffer allows it.  Grids are sized to a multiple of the SM count and grid-stride over the data.
#include <stdint.h>

#include "../include/singa_b200.h"
#include "common.h"
#include "philox.cuh"
#include "vec.cuh"

namespace sg {

constexpr float kStanhA = 1.7159047f;
constexpr float kStanhB = 0.66666667f;

__device__ __forceinline__ float act_fwd(float v, int act) {
  if (act == SG_ACT_RELU) return fmaxf(v, 0.f);
  if (act == SG_ACT_STANH) return kStanhA * tanhf(kStanhB * v);
  return v;
}
__device__ __forceinline__ float act_bwd_from_y(float y, int act) {
  if (act == SG_ACT_RELU) return y > 0.f ? 1.f : 0.f;
  if (act == SG_ACT_STANH) return kStanhA * kStanhB - (kStanhB / kStanhA) * y * y;
  return 1.f;
}

static inline int ew_grid(long long nvec, int threads) {
  long long blocks = (nvec + threads - 1) / threads;
  const long long cap = static_cast<long long>(num_sms()) * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks);
}

// ---------------------------------------------------------------------------------- activation fwd/bwd
template <typename T>
__global__ void act_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, long long n, int act) {
  constexpr int V = Vec<T>::N;
  const long long nv = n / V;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < nv; i += stride) {
    float v[V];
    Vec<T>::load(x + i * V, v);
#pragma unroll
    for (int j = 0; j < V; ++j) v[j] = act_fwd(v[j], act);
    Vec<T>::store(y + i * V, v);
  }
  if (blockIdx.x == 0) {
    for (long long i = nv * V + threadIdx.x; i < n; i += blockDim.x)
      y[i] = from_float<T>(act_fwd(to_float(x[i]), act));
  }
}

template <typename T>
__global__ void act_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ y, T* __restrict__ dx,
                               long long n, int act) {
  constexpr int V = Vec<T>::N;
  const long long nv = n / V;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < nv; i += stride) {
    float g[V], o[V];
    Vec<T>::load(dy + i * V, g);
    Vec<T>::load(y + i * V, o);
#pragma unroll
    for (int j = 0; j < V; ++j) g[j] *= act_bwd_from_y(o[j], act);
    Vec<T>::store(dx + i * V, g);
  }
  if (blockIdx.x == 0) {
    for (long long i = nv * V + threadIdx.x; i < n; i += blockDim.x)
      dx[i] = from_float<T>(to_float(dy[i]) * act_bwd_from_y(to_float(y[i]), act));
  }
}

// ---------------------------------------------------------------------------------- dropout
// mask(i) = (u(seed, i) < pkeep) / pkeep, regenerated in backward from (seed, i): no mask tensor.
// Reference semantics: src/worker/layer.cc:142-160.
__device__ __forceinline__ void load4(const float* p, float (&v)[4]) {
  const float4 t = *reinterpret_cast<const float4*>(p);
  v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
__device__ __forceinline__ void load4(const __nv_bfloat16* p, float (&v)[4]) {
  const uint2 t = *reinterpret_cast<const uint2*>(p);
  v[0] = __uint_as_float(t.x << 16); v[1] = __uint_as_float(t.x & 0xFFFF0000u);
  v[2] = __uint_as_float(t.y << 16); v[3] = __uint_as_float(t.y & 0xFFFF0000u);
}
__device__ __forceinline__ void store4(float* p, const float (&v)[4]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
}
__device__ __forceinline__ void store4(__nv_bfloat16* p, const float (&v)[4]) {
  const __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]), b = __floats2bfloat162_rn(v[2], v[3]);
  *reinterpret_cast<uint2*>(p) = make_uint2(*reinterpret_cast<const uint32_t*>(&a), *reinterpret_cast<const uint32_t*>(&b));
}

template <typename T>
__global__ void dropout_kernel(const T* __restrict__ x, T* __restrict__ y, long long n, float pkeep,
                               unsigned long long seed, const unsigned long long* __restrict__ seed_dev, bool vec) {
  // seed_dev (optional) holds a per-step value in device memory so a captured CUDA graph draws a fresh
  // mask on every replay.
  if (seed_dev != nullptr) seed ^= (*seed_dev + 1ull) * 0x9E3779B97F4A7C15ull;
  const float inv = 1.f / pkeep;
  const long long n4 = (n + 3) / 4;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4; i += stride) {
    float u[4];
    philox_uniform4(seed, static_cast<unsigned long long>(i), u);
    if (vec && i * 4 + 4 <= n) {          // one 16-byte (fp32) / 8-byte (bf16) access per Philox block
      float v[4];
      load4(x + i * 4, v);
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = u[j] < pkeep ? v[j] * inv : 0.f;
      store4(y + i * 4, v);
      continue;
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long long e = i * 4 + j;
      if (e < n) y[e] = from_float<T>(u[j] < pkeep ? to_float(x[e]) * inv : 0.f);
    }
  }
}

// ---------------------------------------------------------------------------------- column sums (bias grad)
// out[c] (+)= sum_r x[r][c].  One warp-wide column strip per block column, rows split across blockIdx.y,
// partial sums merged with one atomicAdd per (block, column): `out` must be zeroed (or hold the value to
// accumulate into) before the launch.  Reference: sum_rows / sumall_except_dim (layer.cc:107,206).
template <typename T>
__global__ void colsum_kernel(const T* __restrict__ x, long long ld, int rows, int cols,
                              float* __restrict__ out) {
  __shared__ float part[8][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  const int rows_per = (rows + gridDim.y - 1) / gridDim.y;
  const int r0 = blockIdx.y * rows_per;
  const int r1 = min(rows, r0 + rows_per);
  float s = 0.f;
  if (c < cols) {
    for (int r = r0 + threadIdx.y; r < r1; r += 8) s += to_float(x[static_cast<long long>(r) * ld + c]);
  }
  part[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && c < cols) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += part[k][threadIdx.x];
    atomicAdd(out + c, t);
  }
}

// Vector form: every thread owns one 16-byte vector of adjacent columns (8 bf16 or 4 fp32; one 128-bit load per row,
// four rows in flight), CVB column vectors x (256 / CVB) row lanes per block; the row lanes are merged through shared
// memory and the block adds its partial sums with one atomicAdd per column.  Needs a 16-byte aligned base and a row
// stride that is a multiple of 16 bytes (true for every activation / gradient buffer of the framework, whose rows are
// padded); a ragged last vector reads the row padding and its extra columns are dropped.  (The 2- or 4-byte-per-lane
// kernel above reaches 1.4 TB/s; tall-skinny bias gradients of the first convolutions — 774 400 x 96 — took 25 % of
// all reduction time, and the 1000 x 2500 bias gradients of the MLP sat on the critical path of the last push.)
template <typename T>
__global__ void __launch_bounds__(256)
colsum_vec_kernel(const T* __restrict__ x, long long ld, int rows, int cols, int cvecs, int cvb,
                  float* __restrict__ out) {
  constexpr int EPV = 16 / sizeof(T);
  __shared__ float part[256 * EPV];
  const int lanes_r = 256 / cvb;                         // row lanes per block
  const int cv_in = threadIdx.x % cvb;
  const int ry = threadIdx.x / cvb;
  const int cv = blockIdx.x * cvb + cv_in;
  const int rows_per = (rows + gridDim.y - 1) / gridDim.y;
  const int r0 = blockIdx.y * rows_per;
  const int r1 = min(rows, r0 + rows_per);
  float acc[EPV];
#pragma unroll
  for (int j = 0; j < EPV; ++j) acc[j] = 0.f;
  auto add = [&](const uint4& u) {
    if (EPV == 8) {
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        acc[(2 * t) % EPV] += __uint_as_float(w[t] << 16);
        acc[(2 * t + 1) % EPV] += __uint_as_float(w[t] & 0xFFFF0000u);
      }
    } else {
      acc[0] += __uint_as_float(u.x); acc[1] += __uint_as_float(u.y);
      acc[2] += __uint_as_float(u.z); acc[3] += __uint_as_float(u.w);
    }
  };
  if (cv < cvecs && ry < lanes_r) {
    const T* base = x + static_cast<long long>(cv) * EPV;
    int r = r0 + ry;
    for (; r + 3 * lanes_r < r1; r += 4 * lanes_r) {
      uint4 u[4];
#pragma unroll
      for (int i = 0; i < 4; ++i)
        u[i] = __ldg(reinterpret_cast<const uint4*>(base + static_cast<long long>(r + i * lanes_r) * ld));
#pragma unroll
      for (int i = 0; i < 4; ++i) add(u[i]);
    }
    for (; r < r1; r += lanes_r) add(__ldg(reinterpret_cast<const uint4*>(base + static_cast<long long>(r) * ld)));
  }
#pragma unroll
  for (int j = 0; j < EPV; ++j) part[threadIdx.x * EPV + j] = acc[j];
  __syncthreads();
  // thread t < cvb * EPV owns column t of this block: sum over the row lanes
  const int ncol = cvb * EPV;
  for (int t = threadIdx.x; t < ncol; t += 256) {
    const int col = blockIdx.x * ncol + t;
    if (col >= cols) continue;
    float sum = 0.f;
    for (int k = 0; k < lanes_r; ++k) sum += part[(k * cvb + t / EPV) * EPV + (t % EPV)];
    atomicAdd(out + col, sum);
  }
}

// ---------------------------------------------------------------------------------- cast / pack
// dst[r][c] (bf16, row stride ldd, padding columns zeroed) = src[r][pc] (fp32, row stride lds) where
// pc = c for a plain cast, or for conv weights (perm_c = C, perm_kk = k*k): destination column
// j = kk_idx*C + c reads source column c*kk + kk_idx.
template <typename S>
__global__ void cast_pack_kernel(const S* __restrict__ src, long long lds,
                                 __nv_bfloat16* __restrict__ dst, long long ldd, int rows, int cols,
                                 int perm_c, int perm_kk) {
  const long long total = static_cast<long long>(rows) * ldd;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += stride) {
    const int r = static_cast<int>(i / ldd);
    const int c = static_cast<int>(i - static_cast<long long>(r) * ldd);
    float v = 0.f;
    if (c < cols) {
      const int sc = perm_c > 0 ? (c % perm_c) * perm_kk + (c / perm_c) : c;
      v = to_float(src[static_cast<long long>(r) * lds + sc]);
    }
    dst[i] = __float2bfloat16_rn(v);
  }
}

// Filter packing for the implicit-GEMM convolution (see sg_conv_pack in the header): one launch writes the forward
// operand [F][(tap, c)] and the data-gradient operand [C][(flipped tap, f)], both zero padded.
template <typename S, typename D>
__global__ void conv_pack_kernel(const S* __restrict__ w, long long ldw, int F, int C, int k, D* __restrict__ fwd,
                                 long long ldfwd, int cp_f, D* __restrict__ dgrad, long long lddgrad, int fp_d) {
  const int kk = k * k;
  const long long n_fwd = static_cast<long long>(F) * ldfwd;
  const long long n_dg = dgrad != nullptr ? static_cast<long long>(C) * lddgrad : 0;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n_fwd + n_dg; i += stride) {
    float v = 0.f;
    if (i < n_fwd) {
      const int f = static_cast<int>(i / ldfwd);
      const int col = static_cast<int>(i - static_cast<long long>(f) * ldfwd);
      const int tap = col / cp_f, c = col - tap * cp_f;
      if (tap < kk && c < C) v = to_float(w[static_cast<long long>(f) * ldw + c * kk + tap]);
      fwd[i] = from_float<D>(v);
    } else {
      const long long j = i - n_fwd;
      const int c = static_cast<int>(j / lddgrad);
      const int col = static_cast<int>(j - static_cast<long long>(c) * lddgrad);
      const int tap = col / fp_d, f = col - tap * fp_d;
      if (tap < kk && f < F) v = to_float(w[static_cast<long long>(f) * ldw + c * kk + (kk - 1 - tap)]);
      dgrad[j] = from_float<D>(v);
    }
  }
}

// Generic dtype conversion of a flat buffer (fp32 <-> bf16).
template <typename S, typename D>
__global__ void convert_kernel(const S* __restrict__ src, D* __restrict__ dst, long long n) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n; i += stride)
    dst[i] = from_float<D>(to_float(src[i]));
}

// ---------------------------------------------------------------------------------- NCHW <-> NHWC
// 32x32 smem tile transpose of the (C, HW) plane of every image.
template <typename S, typename D>
__global__ void transpose_planes_kernel(const S* __restrict__ src, D* __restrict__ dst, int rows,
                                        int cols) {
  // per image: src[rows][cols] -> dst[cols][rows]
  __shared__ float tile[32][33];
  const long long img = blockIdx.z;
  const S* s = src + img * rows * cols;
  D* d = dst + img * rows * cols;
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int j = threadIdx.y; j < 32; j += 8) {
    const int r = r0 + j, c = c0 + threadIdx.x;
    if (r < rows && c < cols) tile[j][threadIdx.x] = to_float(s[static_cast<long long>(r) * cols + c]);
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += 8) {
    const int c = c0 + j, r = r0 + threadIdx.x;
    if (r < rows && c < cols) d[static_cast<long long>(c) * rows + r] = from_float<D>(tile[threadIdx.x][j]);
  }
}

// ---------------------------------------------------------------------------------- L1 norm (debug norms)
// out[0] += sum |x|  (reference Blob::asum_data = cblas_sasum / count, src/utils/blob.cc:251-255)
template <typename T>
__global__ void asum_kernel(const T* __restrict__ x, long long n, float* __restrict__ out) {
  float s = 0.f;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n; i += stride)
    s += fabsf(to_float(x[i]));
  s = warp_sum(s);
  __shared__ float ws[32];
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = threadIdx.x < (blockDim.x >> 5) ? ws[threadIdx.x] : 0.f;
    t = warp_sum(t);
    if (threadIdx.x == 0) atomicAdd(out, t);
  }
}

// ---------------------------------------------------------------------------------- Philox fills (param init)
// kind 0: uniform in [a, b) * scale; kind 1: gaussian(mean=a, std=b) * scale.  Reference: Param::Init,
// src/utils/param.cc:89-127 (distributions match; streams do not — the reference seeds from wall clock).
__global__ void fill_random_kernel(float* __restrict__ dst, long long n, int kind, float a, float b,
                                   float scale, unsigned long long seed) {
  const long long n4 = (n + 3) / 4;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4; i += stride) {
    float u[4];
    philox_uniform4(seed, static_cast<unsigned long long>(i), u);
    float v[4];
    if (kind == 0) {
#pragma unroll
      for (int j = 0; j < 4; ++j) v[j] = (a + (b - a) * u[j]) * scale;
    } else {
      // Box-Muller on two pairs
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const float r = sqrtf(-2.f * logf(fmaxf(u[2 * j], 1e-12f)));
        float sn, cs;
        sincosf(6.283185307179586f * u[2 * j + 1], &sn, &cs);
        v[2 * j] = (a + b * r * cs) * scale;
        v[2 * j + 1] = (a + b * r * sn) * scale;
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long long e = i * 4 + j;
      if (e < n) dst[e] = v[j];
    }
  }
}

// ---------------------------------------------------------------------------------- parsers
// MNIST: uint8 pixel -> x / norm_a - norm_b  (src/worker/layer.cc:440-445)
template <typename S, typename D>
__global__ void mnist_parse_kernel(const S* __restrict__ src, D* __restrict__ dst, long long n,
                                   float norm_a, float norm_b) {
  const float inv = 1.f / norm_a;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n; i += stride)
    dst[i] = from_float<D>(static_cast<float>(src[i]) * inv - norm_b);
}

// RGB: per image random crop offset (train) + mirror flag, uint8 CHW record -> NHWC output * scale, minus
// optional mean image (fp32 CHW, full size).  Intent of src/worker/layer.cc:573-624 (whose non-mirror path
// never writes the output — defect §2.8 of the survey; the intended copy is implemented here).
// One thread per OUTPUT PIXEL: consecutive threads read consecutive bytes of each colour plane (coalesced) and write
// C adjacent channel values; the index arithmetic is done once per pixel instead of once per element (the
// element-per-thread version took 213 us for a 256 x 3 x 227 x 227 batch, ~10x its memory time).
template <typename S, typename D>
__global__ void rgb_parse_kernel(const S* __restrict__ src, const float* __restrict__ mean,
                                 D* __restrict__ dst, int batch, int C, int H, int W, int crop,
                                 const int* __restrict__ off_h, const int* __restrict__ off_w,
                                 const int* __restrict__ mirror, float scale) {
  const int oh = crop > 0 ? crop : H, ow = crop > 0 ? crop : W;
  const long long pixels = static_cast<long long>(batch) * oh * ow;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const long long plane = static_cast<long long>(H) * W;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < pixels; i += stride) {
    const int x = static_cast<int>(i % ow);
    const long long t = i / ow;
    const int y = static_cast<int>(t % oh);
    const int n = static_cast<int>(t / oh);
    const int sy = y + (off_h ? off_h[n] : 0);
    const int sx = (mirror && mirror[n] ? (ow - 1 - x) : x) + (off_w ? off_w[n] : 0);
    const long long s0 = static_cast<long long>(sy) * W + sx;
    const S* sp = src + static_cast<long long>(n) * C * plane + s0;
    D* dp = dst + i * C;
    for (int c = 0; c < C; ++c) {
      float v = static_cast<float>(sp[c * plane]);
      if (mean) v -= mean[c * plane + s0];
      dp[c] = from_float<D>(v * scale);
    }
  }
}

// MNIST distortion (intent of the reference's disabled block, src/worker/layer.cc:406-438: random scaling by
// +-gamma %, then a random rotation by up to beta degrees or a horizontal shear of beta/90, plus the elastic
// displacement field of MnistProto.kernel / sigma / alpha): every output pixel is bilinearly sampled from the source
// image at  A_n * (x, y, 1) + displacement_n(x, y)  and normalised (x / norm_a - norm_b).  A_n (2x3, per image) and
// the smoothed displacement field are produced by the caller; pixels sampled outside the image read 0.
template <typename S, typename D>
__global__ void mnist_distort_kernel(const S* __restrict__ src, D* __restrict__ dst, int n_img, int in_size, int out_size,
                                     const float* __restrict__ affine, const float* __restrict__ disp, float norm_a,
                                     float norm_b) {
  const long long total = static_cast<long long>(n_img) * out_size * out_size;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const float inv = 1.f / norm_a;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += stride) {
    const int x = static_cast<int>(i % out_size);
    const long long t = i / out_size;
    const int y = static_cast<int>(t % out_size);
    const int n = static_cast<int>(t / out_size);
    const float* a = affine + static_cast<long long>(n) * 6;
    float sx = a[0] * x + a[1] * y + a[2];
    float sy = a[3] * x + a[4] * y + a[5];
    if (disp != nullptr) {
      const long long plane = static_cast<long long>(out_size) * out_size;
      sx += disp[(static_cast<long long>(n) * 2) * plane + static_cast<long long>(y) * out_size + x];
      sy += disp[(static_cast<long long>(n) * 2 + 1) * plane + static_cast<long long>(y) * out_size + x];
    }
    const float fx = floorf(sx), fy = floorf(sy);
    const int x0 = static_cast<int>(fx), y0 = static_cast<int>(fy);
    const float wx = sx - fx, wy = sy - fy;
    const S* img = src + static_cast<long long>(n) * in_size * in_size;
    auto px = [&](int yy, int xx) -> float {
      return (yy >= 0 && yy < in_size && xx >= 0 && xx < in_size) ? static_cast<float>(img[yy * in_size + xx]) : 0.f;
    };
    const float v = (1.f - wy) * ((1.f - wx) * px(y0, x0) + wx * px(y0, x0 + 1)) +
                    wy * ((1.f - wx) * px(y0 + 1, x0) + wx * px(y0 + 1, x0 + 1));
    dst[i] = from_float<D>(v * inv - norm_b);
  }
}

// 2-D copy that changes the row stride (used to pad a [rows][cols] matrix to a TMA-legal ld); T = 2- or 4-byte words.
template <typename T>
__global__ void copy2d_kernel(const T* __restrict__ src, long long lds, T* __restrict__ dst, long long ldd,
                              int rows, int cols) {
  const long long total = static_cast<long long>(rows) * ldd;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += stride) {
    const int r = static_cast<int>(i / ldd);
    const int c = static_cast<int>(i - static_cast<long long>(r) * ldd);
    dst[i] = c < cols ? src[static_cast<long long>(r) * lds + c] : static_cast<T>(0);
  }
}


// ------------------------------------------------------------------------------------------------
// Space-to-depth for strided convolutions over thin inputs (RGB, stride 4: AlexNet conv1).  A k x k, stride-s
// convolution over C channels equals a ceil(k/s) x ceil(k/s), stride-1 convolution over s*s*C channels of
//   xs[n][Y][X][(dy*s + dx)*C + c] = x[n][Y*s + dy - pad][X*s + dx - pad][c]      (0 outside the image),
// whose pixels are wide enough for the implicit-GEMM tile loads (ops/conv.py, S2dConvPlan).  T is a bit
// container (bf16 as 16-bit, fp32 as 32-bit words); one block per (image, row of xs), four channels per thread.
// ------------------------------------------------------------------------------------------------
constexpr int S2D_MAX_CS = 1024;
template <typename T>
__global__ void __launch_bounds__(256)
space_to_depth_kernel(const T* __restrict__ x, T* __restrict__ xs, int H, int W, int C, int Hs, int Ws, int Cs,
                      int s, int pad) {
  // channel of xs -> (dy, dx, offset dx*C + c) of the source, looked up instead of divided out per element
  __shared__ unsigned tab[S2D_MAX_CS];
  const int sc = s * C, real = s * sc;
  for (int ch = threadIdx.x; ch < Cs; ch += blockDim.x) {
    unsigned e = 0xFFFFFFFFu;
    if (ch < real) {
      const int dy = ch / sc, r = ch - dy * sc;
      e = (static_cast<unsigned>(dy) << 24) | (static_cast<unsigned>(r / C) << 16) | static_cast<unsigned>(r);
    }
    tab[ch] = e;
  }
  __syncthreads();
  const int n = blockIdx.x / Hs, Y = blockIdx.x - n * Hs;
  const T* xn = x + static_cast<long long>(n) * H * W * C;
  T* row = xs + static_cast<long long>(blockIdx.x) * Ws * Cs;
  const int groups = Cs / 4;
  for (int j = threadIdx.x; j < Ws * groups; j += blockDim.x) {
    const int X = j / groups, ch0 = (j - X * groups) * 4;
    const int x0 = X * s - pad, y0 = Y * s - pad;
    T v[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const unsigned e = tab[ch0 + q];
      const int yy = y0 + static_cast<int>(e >> 24), xx = x0 + static_cast<int>((e >> 16) & 0xFFu);
      v[q] = 0;
      if (e != 0xFFFFFFFFu && yy >= 0 && yy < H && xx >= 0 && xx < W)
        v[q] = xn[(static_cast<long long>(yy) * W + x0) * C + static_cast<int>(e & 0xFFFFu)];
    }
    T* d = row + X * Cs + ch0;
    if (sizeof(T) == 4) {
      *reinterpret_cast<uint4*>(d) = make_uint4(v[0], v[1], v[2], v[3]);
    } else {
      *reinterpret_cast<uint2*>(d) = make_uint2(static_cast<unsigned>(v[0]) | (static_cast<unsigned>(v[1]) << 16),
                                                static_cast<unsigned>(v[2]) | (static_cast<unsigned>(v[3]) << 16));
    }
  }
}

// dst[r][j] = idx[j] >= 0 ? src[r][idx[j]] : 0   (filter re-layouts)
template <typename S>
__global__ void gather_cols_kernel(const S* __restrict__ src, long long lds, float* __restrict__ dst, long long ldd,
                                   int rows, int ncols, const int* __restrict__ idx) {
  const long long total = static_cast<long long>(rows) * ncols;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += stride) {
    const int r = static_cast<int>(i / ncols), j = static_cast<int>(i - static_cast<long long>(r) * ncols);
    const int k = idx[j];
    dst[r * ldd + j] = k >= 0 ? to_float(src[r * lds + k]) : 0.f;
  }
}

// dst[r][idx[j]] (+)= src[r][j] for idx[j] >= 0   (idx injective on its non-negative entries)
__global__ void scatter_cols_kernel(const float* __restrict__ src, long long lds, float* __restrict__ dst,
                                    long long ldd, int rows, int ncols, const int* __restrict__ idx, int accumulate) {
  const long long total = static_cast<long long>(rows) * ncols;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total; i += stride) {
    const int r = static_cast<int>(i / ncols), j = static_cast<int>(i - static_cast<long long>(r) * ncols);
    const int k = idx[j];
    if (k < 0) continue;
    float* d = dst + r * ldd + k;
    *d = accumulate ? *d + src[r * lds + j] : src[r * lds + j];
  }
}

}  // namespace sg

using namespace sg;

#define DTYPE_SWITCH(is_bf16, T, ...)            \
  if (is_bf16) { using T = __nv_bfloat16; __VA_ARGS__ } \
  else { using T = float; __VA_ARGS__ }

extern "C" int sg_act_fwd(const void* x, void* y, long long n, int act, int is_bf16, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (n <= 0) return 0;
  DTYPE_SWITCH(is_bf16, T, {
    SG_REQUIRE(((uintptr_t)x & 15) == 0 && ((uintptr_t)y & 15) == 0, "sg_act_fwd: 16-byte alignment");
    act_fwd_kernel<T><<<ew_grid(n / Vec<T>::N + 1, 256), 256, 0, s>>>((const T*)x, (T*)y, n, act);
  })
  SG_CHECK_LAUNCH("act_fwd");
  return 0;
}

extern "C" int sg_act_bwd(const void* dy, const void* y, void* dx, long long n, int act, int is_bf16,
                          void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (n <= 0) return 0;
  DTYPE_SWITCH(is_bf16, T, {
    SG_REQUIRE(((uintptr_t)dy & 15) == 0 && ((uintptr_t)y & 15) == 0 && ((uintptr_t)dx & 15) == 0,
               "sg_act_bwd: 16-byte alignment");
    act_bwd_kernel<T><<<ew_grid(n / Vec<T>::N + 1, 256), 256, 0, s>>>((const T*)dy, (const T*)y, (T*)dx, n, act);
  })
  SG_CHECK_LAUNCH("act_bwd");
  return 0;
}

extern "C" int sg_dropout(const void* x, void* y, long long n, float pkeep, unsigned long long seed,
                          const unsigned long long* seed_dev, int is_bf16, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (n <= 0) return 0;
  SG_REQUIRE(pkeep > 0.f && pkeep <= 1.f, "sg_dropout: pkeep must be in (0,1]");
  DTYPE_SWITCH(is_bf16, T, {
    const bool vec = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y)) & 15) == 0;
    dropout_kernel<T><<<ew_grid((n + 3) / 4, 256), 256, 0, s>>>((const T*)x, (T*)y, n, pkeep, seed, seed_dev, vec);
  })
  SG_CHECK_LAUNCH("dropout");
  return 0;
}

extern "C" int sg_colsum(const void* x, long long ld, int rows, int cols, float* out, int accumulate,
                         int is_bf16, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (rows <= 0 || cols <= 0) return 0;
  if (!accumulate) {
    if (check_cuda(cudaMemsetAsync(out, 0, sizeof(float) * cols, s), "colsum memset")) return 1;
  }
  const int epv = is_bf16 ? 8 : 4;
  if (ld % epv == 0 && ((reinterpret_cast<uintptr_t>(x) & 15) == 0)) {
    const int cvecs = (cols + epv - 1) / epv;
    int cvb = 1;
    while (cvb < 32 && cvb < cvecs) cvb <<= 1;           // power of two: 256 % cvb == 0
    const int lanes_r = 256 / cvb;
    const int gxv = (cvecs + cvb - 1) / cvb;
    int gyv = (num_sms() * 8 + gxv - 1) / gxv;
    const int max_gyv = (rows + lanes_r * 8 - 1) / (lanes_r * 8);
    if (gyv > max_gyv) gyv = max_gyv;
    if (gyv < 1) gyv = 1;
    if (is_bf16)
      colsum_vec_kernel<__nv_bfloat16><<<dim3(gxv, gyv), 256, 0, s>>>((const __nv_bfloat16*)x, ld, rows, cols, cvecs, cvb, out);
    else
      colsum_vec_kernel<float><<<dim3(gxv, gyv), 256, 0, s>>>((const float*)x, ld, rows, cols, cvecs, cvb, out);
    SG_CHECK_LAUNCH("colsum_vec");
    return 0;
  }
  const int gx = (cols + 31) / 32;
  int gy = (num_sms() * 4 + gx - 1) / gx;
  const int max_gy = (rows + 63) / 64;
  if (gy > max_gy) gy = max_gy;
  if (gy < 1) gy = 1;
  DTYPE_SWITCH(is_bf16, T, {
    colsum_kernel<T><<<dim3(gx, gy), dim3(32, 8), 0, s>>>((const T*)x, ld, rows, cols, out);
  })
  SG_CHECK_LAUNCH("colsum");
  return 0;
}

extern "C" int sg_cast_pack(const float* src, long long lds, void* dst, long long ldd, int rows,
                            int cols, int perm_c, int perm_kk, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (rows <= 0 || cols <= 0) return 0;
  SG_REQUIRE(ldd >= cols && lds >= cols, "sg_cast_pack: leading dimensions too small");
  SG_REQUIRE(perm_c == 0 || perm_c * perm_kk == cols, "sg_cast_pack: perm_c*perm_kk must equal cols");
  const long long total = static_cast<long long>(rows) * ldd;
  cast_pack_kernel<float><<<ew_grid(total, 256), 256, 0, s>>>(src, lds, (__nv_bfloat16*)dst, ldd, rows, cols,
                                                              perm_c, perm_kk);
  SG_CHECK_LAUNCH("cast_pack");
  return 0;
}

// Same re-layout from a bf16 source (the parameter server's weight mirror) — used by ranks that do not hold the
// fp32 master copy of a convolution weight.
extern "C" int sg_repack_bf16(const void* src, long long lds, void* dst, long long ldd, int rows, int cols,
                              int perm_c, int perm_kk, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (rows <= 0 || cols <= 0) return 0;
  SG_REQUIRE(ldd >= cols && lds >= cols, "sg_repack_bf16: leading dimensions too small");
  SG_REQUIRE(perm_c == 0 || perm_c * perm_kk == cols, "sg_repack_bf16: perm_c*perm_kk must equal cols");
  const long long total = static_cast<long long>(rows) * ldd;
  cast_pack_kernel<__nv_bfloat16><<<ew_grid(total, 256), 256, 0, s>>>((const __nv_bfloat16*)src, lds, (__nv_bfloat16*)dst,
                                                                      ldd, rows, cols, perm_c, perm_kk);
  SG_CHECK_LAUNCH("repack_bf16");
  return 0;
}

extern "C" int sg_convert(const void* src, int src_bf16, void* dst, int dst_bf16, long long n,
                          void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (n <= 0) return 0;
  const int g = ew_grid(n, 256);
  if (src_bf16 && !dst_bf16)
    convert_kernel<__nv_bfloat16, float><<<g, 256, 0, s>>>((const __nv_bfloat16*)src, (float*)dst, n);
  else if (!src_bf16 && dst_bf16)
    convert_kernel<float, __nv_bfloat16><<<g, 256, 0, s>>>((const float*)src, (__nv_bfloat16*)dst, n);
  else if (src_bf16)
    convert_kernel<__nv_bfloat16, __nv_bfloat16><<<g, 256, 0, s>>>((const __nv_bfloat16*)src, (__nv_bfloat16*)dst, n);
  else
    convert_kernel<float, float><<<g, 256, 0, s>>>((const float*)src, (float*)dst, n);
  SG_CHECK_LAUNCH("convert");
  return 0;
}

// per image: [rows][cols] -> [cols][rows]; NCHW->NHWC is rows=C, cols=H*W; NHWC->NCHW is rows=H*W, cols=C.
extern "C" int sg_transpose_planes(const void* src, int src_bf16, void* dst, int dst_bf16, int batch,
                                   int rows, int cols, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (batch <= 0 || rows <= 0 || cols <= 0) return 0;
  SG_REQUIRE(batch <= 65535, "sg_transpose_planes: batch too large for grid.z");
  dim3 grid((cols + 31) / 32, (rows + 31) / 32, batch), block(32, 8);
  if (src_bf16 && dst_bf16)
    transpose_planes_kernel<__nv_bfloat16, __nv_bfloat16><<<grid, block, 0, s>>>((const __nv_bfloat16*)src, (__nv_bfloat16*)dst, rows, cols);
  else if (src_bf16)
    transpose_planes_kernel<__nv_bfloat16, float><<<grid, block, 0, s>>>((const __nv_bfloat16*)src, (float*)dst, rows, cols);
  else if (dst_bf16)
    transpose_planes_kernel<float, __nv_bfloat16><<<grid, block, 0, s>>>((const float*)src, (__nv_bfloat16*)dst, rows, cols);
  else
    transpose_planes_kernel<float, float><<<grid, block, 0, s>>>((const float*)src, (float*)dst, rows, cols);
  SG_CHECK_LAUNCH("transpose_planes");
  return 0;
}

extern "C" int sg_asum(const void* x, long long n, float* out, int is_bf16, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (check_cuda(cudaMemsetAsync(out, 0, sizeof(float), s), "asum memset")) return 1;
  if (n <= 0) return 0;
  DTYPE_SWITCH(is_bf16, T, { asum_kernel<T><<<ew_grid(n, 256), 256, 0, s>>>((const T*)x, n, out); })
  SG_CHECK_LAUNCH("asum");
  return 0;
}

extern "C" int sg_fill_random(float* dst, long long n, int kind, float a, float b, float scale,
                              unsigned long long seed, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (n <= 0) return 0;
  SG_REQUIRE(kind == 0 || kind == 1, "sg_fill_random: kind must be 0 (uniform) or 1 (gaussian)");
  fill_random_kernel<<<ew_grid((n + 3) / 4, 256), 256, 0, s>>>(dst, n, kind, a, b, scale, seed);
  SG_CHECK_LAUNCH("fill_random");
  return 0;
}

// src_f32: the staged records are fp32 (float-valued shards / Datums, e.g. mean-subtracted images) instead of uint8
extern "C" int sg_mnist_parse_src(const void* src, int src_f32, void* dst, long long n, float norm_a, float norm_b,
                                  int dst_bf16, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (n <= 0) return 0;
  SG_REQUIRE(norm_a != 0.f, "sg_mnist_parse: norm_a must be non-zero");
  DTYPE_SWITCH(dst_bf16, T, {
    if (src_f32) mnist_parse_kernel<float, T><<<ew_grid(n, 256), 256, 0, s>>>((const float*)src, (T*)dst, n, norm_a, norm_b);
    else mnist_parse_kernel<uint8_t, T><<<ew_grid(n, 256), 256, 0, s>>>((const uint8_t*)src, (T*)dst, n, norm_a, norm_b);
  })
  SG_CHECK_LAUNCH("mnist_parse");
  return 0;
}

extern "C" int sg_mnist_distort(const void* src, int src_f32, void* dst, int n_img, int in_size, int out_size,
                                const float* affine, const float* disp, float norm_a, float norm_b, int dst_bf16,
                                void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (n_img <= 0) return 0;
  SG_REQUIRE(norm_a != 0.f && affine != nullptr && in_size > 0 && out_size > 0, "sg_mnist_distort: bad arguments");
  const long long total = static_cast<long long>(n_img) * out_size * out_size;
  DTYPE_SWITCH(dst_bf16, T, {
    if (src_f32)
      mnist_distort_kernel<float, T><<<ew_grid(total, 256), 256, 0, s>>>((const float*)src, (T*)dst, n_img, in_size, out_size,
                                                                        affine, disp, norm_a, norm_b);
    else
      mnist_distort_kernel<uint8_t, T><<<ew_grid(total, 256), 256, 0, s>>>((const uint8_t*)src, (T*)dst, n_img, in_size,
                                                                          out_size, affine, disp, norm_a, norm_b);
  })
  SG_CHECK_LAUNCH("mnist_distort");
  return 0;
}

extern "C" int sg_mnist_parse(const void* src_u8, void* dst, long long n, float norm_a, float norm_b,
                              int dst_bf16, void* stream) {
  return sg_mnist_parse_src(src_u8, 0, dst, n, norm_a, norm_b, dst_bf16, stream);
}

extern "C" int sg_rgb_parse_src(const void* src, int src_f32, const float* mean, void* dst, int batch, int C, int H,
                                int W, int crop, const int* off_h, const int* off_w, const int* mirror,
                                float scale, int dst_bf16, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (batch <= 0) return 0;
  SG_REQUIRE(crop >= 0 && crop <= H && crop <= W, "sg_rgb_parse: crop larger than the image");
  const int oh = crop > 0 ? crop : H, ow = crop > 0 ? crop : W;
  const long long total = static_cast<long long>(batch) * oh * ow * C;
  DTYPE_SWITCH(dst_bf16, T, {
    if (src_f32)
      rgb_parse_kernel<float, T><<<ew_grid(total / C, 256), 256, 0, s>>>((const float*)src, mean, (T*)dst, batch, C, H, W,
                                                                       crop, off_h, off_w, mirror, scale);
    else
      rgb_parse_kernel<uint8_t, T><<<ew_grid(total / C, 256), 256, 0, s>>>((const uint8_t*)src, mean, (T*)dst, batch, C, H, W,
                                                                         crop, off_h, off_w, mirror, scale);
  })
  SG_CHECK_LAUNCH("rgb_parse");
  return 0;
}

extern "C" int sg_rgb_parse(const void* src_u8, const float* mean, void* dst, int batch, int C, int H,
                            int W, int crop, const int* off_h, const int* off_w, const int* mirror,
                            float scale, int dst_bf16, void* stream) {
  return sg_rgb_parse_src(src_u8, 0, mean, dst, batch, C, H, W, crop, off_h, off_w, mirror, scale, dst_bf16, stream);
}

extern "C" int sg_copy2d(const void* src, long long lds, void* dst, long long ldd, int rows, int cols,
                         int is_bf16, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (rows <= 0 || cols <= 0) return 0;
  SG_REQUIRE(ldd >= cols && lds >= cols, "sg_copy2d: leading dimensions too small");
  const int grid = ew_grid(static_cast<long long>(rows) * ldd, 256);
  if (is_bf16)   // bit copies: bf16 moves as 16-bit words, fp32 as 32-bit words
    copy2d_kernel<unsigned short><<<grid, 256, 0, s>>>((const unsigned short*)src, lds, (unsigned short*)dst, ldd, rows, cols);
  else
    copy2d_kernel<unsigned int><<<grid, 256, 0, s>>>((const unsigned int*)src, lds, (unsigned int*)dst, ldd, rows, cols);
  SG_CHECK_LAUNCH("copy2d");
  return 0;
}

extern "C" int sg_copy2d_bf16(const void* src, long long lds, void* dst, long long ldd, int rows, int cols,
                              void* stream) {
  return sg_copy2d(src, lds, dst, ldd, rows, cols, 1, stream);
}

extern "C" int sg_conv_pack(const void* w, int w_bf16, long long ldw, int F, int C, int k, void* fwd, long long ldfwd,
                            int cp_f, void* dgrad, long long lddgrad, int fp_d, int out_bf16, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  SG_REQUIRE(w != nullptr && fwd != nullptr && F > 0 && C > 0 && k > 0 && cp_f >= C, "sg_conv_pack: bad arguments");
  SG_REQUIRE(dgrad == nullptr || fp_d >= F, "sg_conv_pack: data-gradient pack narrower than the filter count");
  const long long total = static_cast<long long>(F) * ldfwd + (dgrad ? static_cast<long long>(C) * lddgrad : 0);
  const int grid = ew_grid(total, 256);
#define SG_PACK(S, D) conv_pack_kernel<S, D><<<grid, 256, 0, s>>>((const S*)w, ldw, F, C, k, (D*)fwd, ldfwd, cp_f, (D*)dgrad, lddgrad, fp_d)
  if (w_bf16 && out_bf16) SG_PACK(__nv_bfloat16, __nv_bfloat16);
  else if (w_bf16) SG_PACK(__nv_bfloat16, float);
  else if (out_bf16) SG_PACK(float, __nv_bfloat16);
  else SG_PACK(float, float);
#undef SG_PACK
  SG_CHECK_LAUNCH("conv_pack");
  return 0;
}

extern "C" int sg_space_to_depth(const void* x, void* xs, int N, int H, int W, int C, int Hs, int Ws, int Cs, int s_,
                                 int pad, int is_bf16, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (N <= 0) return 0;
  SG_REQUIRE(s_ > 0 && s_ < 256 && pad >= 0 && Cs >= s_ * s_ * C && Hs > 0 && Ws > 0, "sg_space_to_depth: bad geometry");
  SG_REQUIRE(Cs % 4 == 0 && Cs <= S2D_MAX_CS && s_ * C < 65536 && ((uintptr_t)xs & 15) == 0,
             "sg_space_to_depth: Cs must be a multiple of 4, at most %d, and xs 16-byte aligned", S2D_MAX_CS);
  const int grid = N * Hs;
  if (is_bf16)
    space_to_depth_kernel<unsigned short><<<grid, 256, 0, s>>>((const unsigned short*)x, (unsigned short*)xs, H, W, C, Hs, Ws, Cs, s_, pad);
  else
    space_to_depth_kernel<unsigned int><<<grid, 256, 0, s>>>((const unsigned int*)x, (unsigned int*)xs, H, W, C, Hs, Ws, Cs, s_, pad);
  SG_CHECK_LAUNCH("space_to_depth");
  return 0;
}

extern "C" int sg_gather_cols(const void* src, int src_bf16, long long lds, float* dst, long long ldd, int rows,
                              int ncols, const int* idx, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (rows <= 0 || ncols <= 0) return 0;
  SG_REQUIRE(src != nullptr && dst != nullptr && idx != nullptr && ldd >= ncols, "sg_gather_cols: bad arguments");
  const int grid = ew_grid(static_cast<long long>(rows) * ncols, 256);
  if (src_bf16)
    gather_cols_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>((const __nv_bfloat16*)src, lds, dst, ldd, rows, ncols, idx);
  else
    gather_cols_kernel<float><<<grid, 256, 0, s>>>((const float*)src, lds, dst, ldd, rows, ncols, idx);
  SG_CHECK_LAUNCH("gather_cols");
  return 0;
}

extern "C" int sg_scatter_cols(const float* src, long long lds, float* dst, long long ldd, int rows, int ncols,
                               const int* idx, int accumulate, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (rows <= 0 || ncols <= 0) return 0;
  SG_REQUIRE(src != nullptr && dst != nullptr && idx != nullptr && lds >= ncols, "sg_scatter_cols: bad arguments");
  scatter_cols_kernel<<<ew_grid(static_cast<long long>(rows) * ncols, 256), 256, 0, s>>>(src, lds, dst, ldd, rows, ncols, idx, accumulate);
  SG_CHECK_LAUNCH("scatter_cols");
  return 0;
}
