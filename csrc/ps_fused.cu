// Parameter-server data plane over NVSwitch peer memory: the gradient push (reduce-scatter into the owner
// shard + the server Updater + broadcast of the fresh weights) as ONE kernel, plus the device-side flag
// protocol that replaces the reference's ZeroMQ kSync / WaitUpdate messages
// (src/utils/param_manager.cc:163-234, src/server/server.cc:18-43, src/server/pm_server.cc:262-294).
//
// Layout: every rank holds a full-size gradient arena, bf16 weight mirror and flag block, all allocated with
// cudaMalloc and exported through CUDA IPC, so a kernel on rank s can ld/st any rank's copy.  The arena is cut
// into `world` contiguous chunk ranges; rank s is the SERVER for range s (fp32 master weights + optimizer
// state live only there).
//
// One push launch on the owner of chunks [first_chunk, first_chunk + nchunks):
//   1. acquire-spin until every rank's `ready` flag for this layer reaches the step value (its gradients are
//      complete and it no longer reads the bf16 weights of this layer),
//   2. g = (sum over ranks of peer-loaded gradients) * grad_scale        [reduce-scatter, NVLink reads]
//   3. apply the Updater to the local fp32 master weights / history      [server update, in registers]
//   4. store the new weights as bf16 into EVERY rank's mirror            [all-gather by broadcast, NVLink writes]
//   5. the last CTA to finish (global arrival counter = scratch) release-stores the new version into every
//      rank's `version` flag, which is what the consumer GEMM's producer warp / the wait kernel acquires.
// Memory ordering: data stores -> __threadfence_system() -> counter; last CTA: fence -> st.release.sys flags.
#include <cuda_bf16.h>
#include <stdint.h>
#include <string.h>

#include <mutex>

#include "../include/singa_b200.h"
#include "common.h"

namespace sg {

constexpr int CHUNK = 1024;
constexpr int MAX_WORLD = SG_PS_MAX_WORLD;

constexpr int PUSH_UNROLL = 4;   // chunks per CTA iteration: bytes in flight per thread = U x world x 16 B

struct PushParams {
  const float* grad[MAX_WORLD];
  void* w16[MAX_WORLD];                              // every rank's operand mirror (bf16, or fp32 when mirror_fp32)
  const unsigned long long* ready[MAX_WORLD];
  unsigned long long* version[MAX_WORLD];
  int world;
  float* w; float* h; float* u;
  const int* chunk_param; const float* lr_mult; const float* wd_mult;
  long long first_chunk, nchunks;
  int kind; float lr, momentum, wd, delta, rho, grad_scale; const float* lr_dev;
  const long long* step_dev; long long value_add;   // explicit flag value = *step_dev + value_add (seq == nullptr)
  unsigned long long* scratch;                       // [0] CTA arrival counter (zero between launches)
                                                     // [1] launch sequence number (self-counting flag value)
  int self_count;
  const float* grad_mc;                              // NVLS: multicast address of the gradient arenas (or null)
  void* w16_mc;                                      // NVLS: multicast address of the mirrors (or null)
  int mirror_fp32;                                   // mirrors hold fp32 (tf32 compute path) instead of bf16
  unsigned long long* trace;                         // optional: phase time stamps (ns) of this launch (profiling tool)
  unsigned long long* bump_ready;                    // optional: this rank's own "gradients ready" flag, bumped by
  unsigned long long* bump_seq;                      //   the kernel itself (+ the consumer-side wait sequence)
};

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define PUSH_STAMP(i) do { if (p.trace != nullptr && blockIdx.x == 0 && threadIdx.x == 0) p.trace[i] = globaltimer_ns(); } while (0)


// NVLS (NVLink SHARP) accessors: one instruction reduces / broadcasts across every GPU bound to the multicast
// object, inside the NVSwitch — the reduce-scatter costs each GPU 1/world of the bytes the peer-load version pulls.
__device__ __forceinline__ float4 multimem_ld_reduce_add(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_f4(void* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
               ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void multimem_st_b64(void* mc, uint2 v) {
  asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};"
               ::"l"(mc), "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)) : "memory");
}

__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// Bounded spin: a peer that never arrives becomes a trap (host-visible error), not a hung GPU.
__device__ __forceinline__ void wait_flag_ge(const unsigned long long* p, unsigned long long target) {
  if (ld_acquire_sys(p) >= target) return;
  const long long t0 = clock64();
  while (ld_acquire_sys(p) < target) {
    __nanosleep(200);
    if (clock64() - t0 > 60000000000LL) __trap();   // ~30 s
  }
}

__device__ __forceinline__ void upd(int kind, float lr, float wd, float momentum, float delta, float rho,
                                    float& w, float g, float& h, float& u) {
  switch (kind) {
    case SG_UPD_SGD: {
      if (wd > 0.f) g += wd * w;
      if (momentum > 0.f) { h = momentum * h + lr * g; w -= h; } else w -= lr * g;
      break;
    }
    case SG_UPD_NESTEROV: {
      if (wd > 0.f) g += wd * w;
      const float hold = h;
      h = momentum * h + lr * g;
      w -= h * (1.f + momentum) - hold * momentum;
      break;
    }
    case SG_UPD_ADAGRAD: {
      h += g * g;
      if (wd > 0.f) g += wd * w;
      w -= lr * g * rsqrtf(h + delta);
      break;
    }
    case SG_UPD_RMSPROP: {
      h = rho * h + (1.f - rho) * g * g;
      if (wd > 0.f) g += wd * w;
      w -= lr * g * rsqrtf(h + delta);
      break;
    }
    case SG_UPD_ADADELTA: {
      if (wd > 0.f) g += wd * w;
      h = rho * h + (1.f - rho) * g * g;
      const float t = g * sqrtf(u + delta) * rsqrtf(h + delta);
      u = rho * u + (1.f - rho) * t * t;
      w -= t;
      break;
    }
    default: break;
  }
}

__global__ void __launch_bounds__(256) ps_push_update_kernel(const PushParams p) {
  // Flag value of this launch.  Self-counting mode (the default of the framework): the n-th launch for a
  // layer waits for ready >= n and publishes version = n, with n kept in scratch[1] — no dependence on a step
  // scalar that another stream may already have advanced.
  PUSH_STAMP(0);
  const unsigned long long target = p.self_count
      ? p.scratch[1] + 1ull
      : static_cast<unsigned long long>((p.step_dev ? *p.step_dev : 0) + p.value_add);
  // 0. announce this rank's own gradients (folded in: saves a flag kernel + launch gap per push).  Stream order
  //    guarantees the local gradients are complete when this kernel starts; the flag is set BEFORE any wait, so
  //    two ranks' push kernels can never wait on each other.
  if (p.bump_ready != nullptr && blockIdx.x == 0 && threadIdx.x == 0) {
    if (p.bump_seq != nullptr) *p.bump_seq += 1ull;
    __threadfence_system();
    st_release_sys(p.bump_ready, *p.bump_ready + 1ull);
  }
  PUSH_STAMP(1);
  // 1. wait for every rank's gradients of this layer
  if (threadIdx.x < p.world && p.ready[threadIdx.x] != nullptr) wait_flag_ge(p.ready[threadIdx.x], target);
  __syncthreads();
  PUSH_STAMP(2);

  const float base_lr = p.lr_dev ? __ldg(p.lr_dev) : p.lr;
  const bool need_h = p.kind != SG_UPD_SGD || p.momentum > 0.f;
  const bool need_u = p.kind == SG_UPD_ADADELTA;
  // PUSH_UNROLL chunks per iteration with every peer load issued before the first use: NVLink reads have ~2 us
  // latency, so bandwidth comes from bytes in flight (U x world x 16 B per thread).
  constexpr int U = PUSH_UNROLL;
  for (long long c0 = static_cast<long long>(blockIdx.x) * U; c0 < p.nchunks;
       c0 += static_cast<long long>(gridDim.x) * U) {
    float4 g[U];
    int pid[U];
    long long idx[U];
#pragma unroll
    for (int k = 0; k < U; ++k) {
      const long long c = c0 + k;
      pid[k] = c < p.nchunks ? p.chunk_param[p.first_chunk + c] : -1;
      idx[k] = (p.first_chunk + c) * CHUNK + threadIdx.x * 4;
      g[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    // 2. reduce the peers' gradients: in the switch (NVLS) when a multicast mapping exists, else with
    //    128-bit peer loads (remote ranks go over NVLink)
    if (p.grad_mc != nullptr) {
#pragma unroll
      for (int k = 0; k < U; ++k)
        if (pid[k] >= 0) g[k] = multimem_ld_reduce_add(p.grad_mc + idx[k]);
    } else
#pragma unroll
    for (int r = 0; r < MAX_WORLD; ++r) {
      if (r < p.world) {
        float4 t[U];
#pragma unroll
        for (int k = 0; k < U; ++k)
          t[k] = pid[k] >= 0 ? *reinterpret_cast<const float4*>(p.grad[r] + idx[k]) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int k = 0; k < U; ++k) { g[k].x += t[k].x; g[k].y += t[k].y; g[k].z += t[k].z; g[k].w += t[k].w; }
      }
    }
#pragma unroll
    for (int k = 0; k < U; ++k) {
      if (pid[k] < 0) continue;
      const long long i = idx[k];
      const float lr = base_lr * p.lr_mult[pid[k]];
      const float wd = p.wd * p.wd_mult[pid[k]];
      float4 gv = g[k];
      gv.x *= p.grad_scale; gv.y *= p.grad_scale; gv.z *= p.grad_scale; gv.w *= p.grad_scale;
      // 3. server update on the local master copy
      float4 wv = *reinterpret_cast<const float4*>(p.w + i);
      float4 hv = make_float4(0.f, 0.f, 0.f, 0.f), uv = hv;
      if (need_h) hv = *reinterpret_cast<const float4*>(p.h + i);
      if (need_u) uv = *reinterpret_cast<const float4*>(p.u + i);
      upd(p.kind, lr, wd, p.momentum, p.delta, p.rho, wv.x, gv.x, hv.x, uv.x);
      upd(p.kind, lr, wd, p.momentum, p.delta, p.rho, wv.y, gv.y, hv.y, uv.y);
      upd(p.kind, lr, wd, p.momentum, p.delta, p.rho, wv.z, gv.z, hv.z, uv.z);
      upd(p.kind, lr, wd, p.momentum, p.delta, p.rho, wv.w, gv.w, hv.w, uv.w);
      *reinterpret_cast<float4*>(p.w + i) = wv;
      if (need_h) *reinterpret_cast<float4*>(p.h + i) = hv;
      if (need_u) *reinterpret_cast<float4*>(p.u + i) = uv;
      // 4. broadcast the fresh weights into every rank's mirror (fp32 on the tf32 compute path, else bf16)
      if (p.mirror_fp32) {
        if (p.w16_mc != nullptr) {
          multimem_st_f4(reinterpret_cast<float*>(p.w16_mc) + i, wv);
        } else {
#pragma unroll
          for (int r = 0; r < MAX_WORLD; ++r)
            if (r < p.world && p.w16[r] != nullptr) *reinterpret_cast<float4*>(reinterpret_cast<float*>(p.w16[r]) + i) = wv;
        }
        continue;
      }
      const __nv_bfloat162 a = __floats2bfloat162_rn(wv.x, wv.y), b = __floats2bfloat162_rn(wv.z, wv.w);
      uint2 o;
      o.x = *reinterpret_cast<const uint32_t*>(&a);
      o.y = *reinterpret_cast<const uint32_t*>(&b);
      if (p.w16_mc != nullptr) {
        multimem_st_b64(reinterpret_cast<__nv_bfloat16*>(p.w16_mc) + i, o);
      } else {
#pragma unroll
        for (int r = 0; r < MAX_WORLD; ++r)
          if (r < p.world && p.w16[r] != nullptr) *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(p.w16[r]) + i) = o;
      }
    }
  }
  // 5. publish: last CTA to arrive bumps the version flag on every rank
  PUSH_STAMP(3);
  __threadfence_system();
  __syncthreads();
  PUSH_STAMP(4);
  if (threadIdx.x == 0) {
    const unsigned long long done = atomicAdd(p.scratch, 1ull) + 1ull;
    if (done == gridDim.x) {
      p.scratch[0] = 0ull;   // arrival counter must read zero at the next launch
      p.scratch[1] = target; // every CTA has read the old sequence number by now (all have arrived)
      if (p.trace != nullptr) p.trace[5] = globaltimer_ns();
      __threadfence_system();
      for (int r = 0; r < p.world; ++r)
        if (p.version[r] != nullptr) st_release_sys(p.version[r], target);
      if (p.trace != nullptr) p.trace[6] = globaltimer_ns();
    }
  }
}

__global__ void ps_flag_set_kernel(unsigned long long* flag, const long long* step_dev, long long value_add,
                                   int bump, unsigned long long* local_seq) {
  if (local_seq != nullptr) *local_seq += 1ull;   // consumer-side wait sequence of the same layer (see gemm wait)
  __threadfence_system();
  // bump: flag += 1 (single writer: the owning GPU's stream-ordered launches), else explicit value
  const unsigned long long v = bump ? (*flag + 1ull)
                                    : static_cast<unsigned long long>((step_dev ? *step_dev : 0) + value_add);
  st_release_sys(flag, v);
}

struct WaitParams {
  const unsigned long long* flags[SG_PS_MAX_WAIT];
  int n;
  const long long* step_dev;
  long long value_add;
  const unsigned long long* seq;   // self-counting: wait for flags >= *seq (sg_ps_flag_bump advances it)
};
__global__ void ps_flag_wait_kernel(const WaitParams p) {
  long long explicit_target = (p.step_dev ? *p.step_dev : 0) + p.value_add;
  if (explicit_target < 0) explicit_target = 0;   // e.g. "step - staleness" during the first steps
  const unsigned long long target = p.seq ? *p.seq : static_cast<unsigned long long>(explicit_target);
  if (threadIdx.x < p.n) wait_flag_ge(p.flags[threadIdx.x], target);
}

}  // namespace sg

using namespace sg;

extern "C" int sg_ps_push_update(const SgPsPush* a, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  SG_REQUIRE(a != nullptr, "sg_ps_push_update: null args");
  if (a->nchunks <= 0) return 0;
  SG_REQUIRE(a->world >= 1 && a->world <= MAX_WORLD, "sg_ps_push_update: world %d out of range [1,%d]", a->world, MAX_WORLD);
  SG_REQUIRE(a->scratch != nullptr, "sg_ps_push_update: needs 16 bytes of zero-initialised scratch");
  SG_REQUIRE(a->up.kind >= SG_UPD_SGD && a->up.kind <= SG_UPD_ADADELTA, "sg_ps_push_update: unknown updater");
  const bool need_h = a->up.kind != SG_UPD_SGD || a->up.momentum > 0.f;
  SG_REQUIRE(!need_h || a->h != nullptr, "sg_ps_push_update: updater needs a history buffer");
  SG_REQUIRE(a->up.kind != SG_UPD_ADADELTA || a->u != nullptr, "sg_ps_push_update: AdaDelta needs an update buffer");
  PushParams p;
  for (int r = 0; r < MAX_WORLD; ++r) {
    p.grad[r] = r < a->world ? a->grad[r] : nullptr;
    p.w16[r] = r < a->world ? a->w16[r] : nullptr;
    p.ready[r] = r < a->world ? a->ready[r] : nullptr;
    p.version[r] = r < a->world ? a->version[r] : nullptr;
    // a null mirror pointer is allowed (that rank reads the master copy directly: single-rank fp32 jobs)
    SG_REQUIRE(r >= a->world || p.grad[r] != nullptr, "sg_ps_push_update: null gradient pointer for rank %d", r);
  }
  p.world = a->world;
  p.w = a->w; p.h = a->h; p.u = a->u;
  p.chunk_param = a->chunk_param; p.lr_mult = a->lr_mult; p.wd_mult = a->wd_mult;
  p.first_chunk = a->first_chunk; p.nchunks = a->nchunks;
  p.kind = a->up.kind; p.lr = a->up.lr; p.momentum = a->up.momentum; p.wd = a->up.weight_decay;
  p.delta = a->up.delta; p.rho = a->up.rho; p.grad_scale = a->up.grad_scale; p.lr_dev = a->up.lr_dev;
  p.step_dev = a->step_dev; p.value_add = a->value_add;
  p.scratch = a->scratch; p.self_count = a->self_count;
  p.grad_mc = a->grad_mc; p.w16_mc = a->w16_mc; p.mirror_fp32 = a->mirror_fp32; p.trace = a->trace;
  p.bump_ready = a->bump_ready; p.bump_seq = a->bump_seq;
  long long grid = (a->nchunks + PUSH_UNROLL - 1) / PUSH_UNROLL;
  const long long cap = a->max_ctas > 0 ? a->max_ctas : 64;   // a small CTA budget: leave the SMs to backward
  if (grid > cap) grid = cap;
  ps_push_update_kernel<<<static_cast<unsigned>(grid), 256, 0, s>>>(p);
  SG_CHECK_LAUNCH("ps_push_update");
  return 0;
}

extern "C" int sg_ps_flag_bump(unsigned long long* flag, unsigned long long* local_seq, void* stream) {
  ps_flag_set_kernel<<<1, 1, 0, reinterpret_cast<cudaStream_t>(stream)>>>(flag, nullptr, 0, 1, local_seq);
  SG_CHECK_LAUNCH("ps_flag_bump");
  return 0;
}

extern "C" int sg_ps_flag_set(unsigned long long* flag, const long long* step_dev, long long value_add,
                              void* stream) {
  ps_flag_set_kernel<<<1, 1, 0, reinterpret_cast<cudaStream_t>(stream)>>>(flag, step_dev, value_add, 0, nullptr);
  SG_CHECK_LAUNCH("ps_flag_set");
  return 0;
}

extern "C" int sg_ps_flag_wait(const unsigned long long* const* flags, int n, const long long* step_dev,
                               long long value_add, const unsigned long long* seq, void* stream) {
  SG_REQUIRE(n >= 0 && n <= SG_PS_MAX_WAIT, "sg_ps_flag_wait: at most %d flags per call", SG_PS_MAX_WAIT);
  if (n == 0) return 0;
  WaitParams p;
  for (int i = 0; i < SG_PS_MAX_WAIT; ++i) p.flags[i] = i < n ? flags[i] : nullptr;
  p.n = n; p.step_dev = step_dev; p.value_add = value_add; p.seq = seq;
  ps_flag_wait_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(p);
  SG_CHECK_LAUNCH("ps_flag_wait");
  return 0;
}

// ------------------------------------------------------------------------------------------------
// CUDA IPC helpers: arenas that peers can map (cudaMalloc + cudaIpc*; 64-byte opaque handles travel over the
// control plane).
// ------------------------------------------------------------------------------------------------
extern "C" int sg_ipc_alloc(long long bytes, void** out) {
  SG_REQUIRE(bytes > 0 && out != nullptr, "sg_ipc_alloc: bad arguments");
  if (check_cuda(cudaMalloc(out, static_cast<size_t>(bytes)), "cudaMalloc(ipc arena)")) return 1;
  if (check_cuda(cudaMemset(*out, 0, static_cast<size_t>(bytes)), "cudaMemset(ipc arena)")) return 1;
  return 0;
}
extern "C" int sg_ipc_free(void* p) {
  if (p == nullptr) return 0;
  return check_cuda(cudaFree(p), "cudaFree(ipc arena)");
}
extern "C" int sg_ipc_get_handle(void* p, void* handle64) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  return check_cuda(cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(handle64), p), "cudaIpcGetMemHandle");
}
extern "C" int sg_ipc_open_handle(const void* handle64, void** out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  return check_cuda(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess), "cudaIpcOpenMemHandle");
}
extern "C" int sg_ipc_close_handle(void* p) {
  if (p == nullptr) return 0;
  return check_cuda(cudaIpcCloseMemHandle(p), "cudaIpcCloseMemHandle");
}

// ------------------------------------------------------------------------------------------------
// Allocation-free variant of the push.
//
// Who needs it: code that drives this library from a place where it may not create device memory — a host
// application whose memory manager owns every byte of HBM and hands out nothing at run time, a plug-in invoked
// from a callback in which cudaMalloc (which synchronises the device and takes driver-wide locks) is forbidden,
// or a latency-critical loop that must never block on the allocator.  Such a caller can pass its inputs and
// outputs but has nowhere to get the push kernel's scratch words from, so this entry point keeps them in the
// library image: a file-scope device array, one 256-byte line per slot, SG_PS_NOALLOC_SLOTS slots (16 KB that
// every GPU which loads the library holds until the process exits).
//
// Rules that follow from sharing one built-in array:
//  * In host code the array's name is only a stand-in; the device address is obtained once per device with
//    cudaGetSymbolAddress while that device is current.  The FIRST call on a device performs this set-up, so
//    make that first call before starting any stream capture on that device.
//  * The scratch keeps whatever the previous call left; the words that must start at zero are cleared on the
//    caller's stream before each launch.
//  * All calls with the same slot share the same words: two launches that can overlap on the GPU (different
//    streams, or the same slot used for two layers in flight) race.  Use one slot per concurrently pushed layer.
//  * The looked-up address is cached per device ordinal for the life of the process; it does not survive
//    cudaDeviceReset (or any teardown of the primary context): do not call this entry point again after one.
//  * Only the explicit flag-value mode is available (self_count must be 0): the self-counted sequence number
//    would be part of the cleared scratch.
// Nothing here calls cudaMalloc, cudaFree or any allocating library routine.
// ------------------------------------------------------------------------------------------------
namespace sg {
constexpr int NOALLOC_WORDS_PER_SLOT = 32;   // 256 bytes
static __device__ __align__(256) unsigned long long g_noalloc_scratch[SG_PS_NOALLOC_SLOTS * NOALLOC_WORDS_PER_SLOT];

constexpr int NOALLOC_MAX_DEVICES = 64;
static std::mutex g_noalloc_mu;
static unsigned long long* g_noalloc_dev_ptr[NOALLOC_MAX_DEVICES] = {nullptr};

// Device address of the built-in scratch on the CURRENT device (looked up once per device).
static int noalloc_scratch_base(unsigned long long** out) {
  int dev = -1;
  if (check_cuda(cudaGetDevice(&dev), "sg_ps_push_update_noalloc: cudaGetDevice")) return 1;
  SG_REQUIRE(dev >= 0 && dev < NOALLOC_MAX_DEVICES, "sg_ps_push_update_noalloc: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (g_noalloc_dev_ptr[dev] == nullptr) {
    void* p = nullptr;
    if (check_cuda(cudaGetSymbolAddress(&p, g_noalloc_scratch), "sg_ps_push_update_noalloc: cudaGetSymbolAddress"))
      return 1;
    SG_REQUIRE(p != nullptr, "sg_ps_push_update_noalloc: built-in scratch has no device address");
    g_noalloc_dev_ptr[dev] = static_cast<unsigned long long*>(p);
  }
  *out = g_noalloc_dev_ptr[dev];
  return 0;
}
}  // namespace sg

extern "C" int sg_ps_push_update_noalloc(const SgPsPush* a, int slot, void* stream) {
  SG_REQUIRE(a != nullptr, "sg_ps_push_update_noalloc: null args");
  SG_REQUIRE(slot >= 0 && slot < SG_PS_NOALLOC_SLOTS,
             "sg_ps_push_update_noalloc: slot %d is outside the built-in scratch (%d slots of 256 bytes); "
             "use sg_ps_push_update with caller-provided scratch for more layers in flight",
             slot, SG_PS_NOALLOC_SLOTS);
  SG_REQUIRE(a->self_count == 0,
             "sg_ps_push_update_noalloc: self-counted flag values are not available (the scratch is cleared per call); "
             "pass the flag value through step_dev / value_add");
  if (a->nchunks <= 0) return 0;
  unsigned long long* base = nullptr;
  if (int rc = noalloc_scratch_base(&base)) return rc;
  unsigned long long* mine = base + static_cast<long long>(slot) * NOALLOC_WORDS_PER_SLOT;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  // arrival counter (and the unused sequence word) must read zero when the kernel starts
  if (check_cuda(cudaMemsetAsync(mine, 0, 2 * sizeof(unsigned long long), s), "sg_ps_push_update_noalloc: clear scratch"))
    return 1;
  SgPsPush b = *a;
  b.scratch = mine;
  b.self_count = 0;
  return sg_ps_push_update(&b, stream);
}
