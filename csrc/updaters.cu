// Fused multi-tensor updaters over the flat parameter arena, plus the EASGD / RandomSync exchange rules.
//
// The arena (singa_b200/arena.py) lays every parameter out at an offset that is a multiple of
// SG_ARENA_CHUNK (1024) elements, so one CTA handles one chunk and looks its parameter up in a chunk ->
// parameter table to fetch the per-parameter learning-rate and weight-decay multipliers.  One launch
// updates the whole model (or one layer's slice of it) and also refreshes the bf16 mirror that the GEMMs
// read.  Math restates src/utils/updater.cc:54-182 (with the survey's §2.8 fixes: grad_scale is honoured
// by SGD, Nesterov uses the configured momentum).
// Deliberate deviation, shared by all three statements of the rules (this file, ps_fused.cu::upd and
// ops/reference.py::update_ref): grad_scale multiplies the gradient ONCE, up front, for every updater.  The reference
// applies it only inside the squared-history term of AdaGrad / RMSProp (updater.cc:126-131, 152-157) and squares
// (g + wd*w)*scale for AdaDelta; with grad_scale = 1/world (how the sharded server averages worker gradients) that
// would make the step size depend on the number of workers.
#include <cuda_bf16.h>
#include <stdint.h>

#include "../include/singa_b200.h"
#include "common.h"
#include "philox.cuh"

namespace sg {

constexpr int CHUNK = 1024;

struct UpdHyper {
  int kind;
  float lr, momentum, wd, delta, rho, grad_scale;
  const float* lr_dev;
};

__device__ __forceinline__ void update_one(const UpdHyper& hp, float lr, float wd, float& w, float g,
                                           float& h, float& u) {
  const float sg_ = g * hp.grad_scale;
  switch (hp.kind) {
    case SG_UPD_SGD: {
      float gg = sg_;
      if (wd > 0.f) gg += wd * w;
      if (hp.momentum > 0.f) { h = hp.momentum * h + lr * gg; w -= h; }
      else w -= lr * gg;
      break;
    }
    case SG_UPD_NESTEROV: {
      float gg = sg_;
      if (wd > 0.f) gg += wd * w;
      const float hold = h;
      h = hp.momentum * h + lr * gg;
      w -= h * (1.f + hp.momentum) - hold * hp.momentum;
      break;
    }
    case SG_UPD_ADAGRAD: {
      h += sg_ * sg_;
      float gg = sg_;
      if (wd > 0.f) gg += wd * w;
      w -= lr * gg * rsqrtf(h + hp.delta);
      break;
    }
    case SG_UPD_RMSPROP: {
      h = hp.rho * h + (1.f - hp.rho) * sg_ * sg_;
      float gg = sg_;
      if (wd > 0.f) gg += wd * w;
      w -= lr * gg * rsqrtf(h + hp.delta);
      break;
    }
    case SG_UPD_ADADELTA: {
      float gg = sg_;
      if (wd > 0.f) gg += wd * w;
      h = hp.rho * h + (1.f - hp.rho) * gg * gg;
      const float t = gg * sqrtf(u + hp.delta) * rsqrtf(h + hp.delta);
      u = hp.rho * u + (1.f - hp.rho) * t * t;
      w -= t;
      break;
    }
    default: break;
  }
}

// One CTA (256 threads x float4) per 1024-element chunk.
__global__ void __launch_bounds__(256)
update_arena_kernel(float* __restrict__ w, const float* __restrict__ g, float* __restrict__ h,
                    float* __restrict__ u, __nv_bfloat16* __restrict__ w16,
                    const int* __restrict__ chunk_param, const float* __restrict__ lr_mult,
                    const float* __restrict__ wd_mult, long long first_chunk, UpdHyper hp) {
  const long long chunk = first_chunk + blockIdx.x;
  const int pid = chunk_param[chunk];
  if (pid < 0) return;  // padding chunk
  const float lr = (hp.lr_dev != nullptr ? __ldg(hp.lr_dev) : hp.lr) * lr_mult[pid];
  const float wd = hp.wd * wd_mult[pid];
  const long long i = chunk * CHUNK + threadIdx.x * 4;
  float4 wv = *reinterpret_cast<const float4*>(w + i);
  const float4 gv = *reinterpret_cast<const float4*>(g + i);
  float4 hv = make_float4(0.f, 0.f, 0.f, 0.f), uv = hv;
  const bool need_h = hp.kind != SG_UPD_SGD || hp.momentum > 0.f;
  const bool need_u = hp.kind == SG_UPD_ADADELTA;
  if (need_h) hv = *reinterpret_cast<const float4*>(h + i);
  if (need_u) uv = *reinterpret_cast<const float4*>(u + i);
  update_one(hp, lr, wd, wv.x, gv.x, hv.x, uv.x);
  update_one(hp, lr, wd, wv.y, gv.y, hv.y, uv.y);
  update_one(hp, lr, wd, wv.z, gv.z, hv.z, uv.z);
  update_one(hp, lr, wd, wv.w, gv.w, hv.w, uv.w);
  *reinterpret_cast<float4*>(w + i) = wv;
  if (need_h) *reinterpret_cast<float4*>(h + i) = hv;
  if (need_u) *reinterpret_cast<float4*>(u + i) = uv;
  if (w16 != nullptr) {
    const __nv_bfloat162 a = __floats2bfloat162_rn(wv.x, wv.y), b = __floats2bfloat162_rn(wv.z, wv.w);
    uint2 o;
    o.x = *reinterpret_cast<const uint32_t*>(&a);
    o.y = *reinterpret_cast<const uint32_t*>(&b);
    *reinterpret_cast<uint2*>(w16 + i) = o;
  }
}

// EASGD exchange (src/utils/param.cc:244-284): d = (w - s) * alpha; s += d; w -= d.
// `s` may be a peer-mapped pointer to the server shard on another GPU; the add is atomic so several worker
// groups can hit the same shard concurrently.
__global__ void easgd_kernel(float* __restrict__ w, float* __restrict__ s, long long n, float alpha,
                             __nv_bfloat16* __restrict__ w16) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n; i += stride) {
    const float wi = w[i];
    const float d = (wi - s[i]) * alpha;
    atomicAdd(s + i, d);
    const float nw = wi - d;
    w[i] = nw;
    if (w16) w16[i] = __float2bfloat16_rn(nw);
  }
}

// RandomSync exchange (src/utils/param.cc:130-228).  Index i takes part iff philox(seed, i) < ratio — the
// shared-seed subset both sides can regenerate (a parallel restatement of the reference's sequential
// selection sampling).  delta = w - snap; old = s; s += delta; w += old - snap; snap = w.
__global__ void randomsync_kernel(float* __restrict__ w, float* __restrict__ snap, float* __restrict__ s,
                                  long long n, float ratio, unsigned long long seed,
                                  __nv_bfloat16* __restrict__ w16, unsigned long long* __restrict__ nsel) {
  const long long n4 = (n + 3) / 4;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  unsigned int local = 0;
  for (long long q = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; q < n4; q += stride) {
    float uu[4];
    philox_uniform4(seed, static_cast<unsigned long long>(q), uu);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long long i = q * 4 + j;
      if (i < n && (ratio >= 1.f || uu[j] < ratio)) {
        const float sn = snap[i];
        const float delta = w[i] - sn;
        const float old = atomicAdd(s + i, delta);
        const float nw = w[i] + old - sn;
        w[i] = nw;
        snap[i] = nw;
        if (w16) w16[i] = __float2bfloat16_rn(nw);
        ++local;
      }
    }
  }
  if (nsel != nullptr && local) atomicAdd(nsel, static_cast<unsigned long long>(local));
}

// ---------------------------------------------------------------------------------------------------------
// Sharded forms: ONE launch exchanges the whole arena with all server shards.  Shard k = elements
// [k * per, (k+1) * per) and lives behind table.server[k] (a peer-mapped arena base: the same element index is used
// on every GPU).  `seed_dev` (optional) is mixed into the RandomSync seed on the device so a captured graph draws a
// new subset per replay.
// ---------------------------------------------------------------------------------------------------------
struct ShardTable {
  float* server[SG_PS_MAX_WORLD];
  int nshards;
  long long per;
};

__global__ void easgd_sharded_kernel(float* __restrict__ w, const ShardTable t, long long n, float alpha,
                                     __nv_bfloat16* __restrict__ w16) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n; i += stride) {
    const int k = static_cast<int>(i / t.per);
    float* s = t.server[k < t.nshards ? k : t.nshards - 1];
    const float wi = w[i];
    const float d = (wi - s[i]) * alpha;
    atomicAdd(s + i, d);
    const float nw = wi - d;
    w[i] = nw;
    if (w16) w16[i] = __float2bfloat16_rn(nw);
  }
}

__global__ void randomsync_sharded_kernel(float* __restrict__ w, float* __restrict__ snap, const ShardTable t,
                                          long long n, float ratio, unsigned long long seed,
                                          const unsigned long long* __restrict__ seed_dev,
                                          __nv_bfloat16* __restrict__ w16, unsigned long long* __restrict__ nsel) {
  if (seed_dev != nullptr) seed ^= (*seed_dev + 1ull) * 0x9E3779B97F4A7C15ull;
  const long long n4 = (n + 3) / 4;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  unsigned int local = 0;
  for (long long q = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; q < n4; q += stride) {
    float uu[4];
    philox_uniform4(seed, static_cast<unsigned long long>(q), uu);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long long i = q * 4 + j;
      if (i < n && (ratio >= 1.f || uu[j] < ratio)) {
        const int k = static_cast<int>(i / t.per);
        float* s = t.server[k < t.nshards ? k : t.nshards - 1];
        const float sn = snap[i];
        const float delta = w[i] - sn;
        const float old = atomicAdd(s + i, delta);
        const float nw = w[i] + old - sn;
        w[i] = nw;
        snap[i] = nw;
        if (w16) w16[i] = __float2bfloat16_rn(nw);
        ++local;
      }
    }
  }
  if (nsel != nullptr && local) atomicAdd(nsel, static_cast<unsigned long long>(local));
}

// Server-group neighbour synchronisation (src/server/pm_server.cc:105-127,296-354, src/utils/param_shard.cc:52-58):
// the owner of a shard pulls the same shard of every neighbour group over peer memory and moves its own copy to the
// mean of the group copies.  The move is applied as an atomic add of (mean - mine) so that worker groups exchanging
// with this shard at the same time are not overwritten.
__global__ void group_average_kernel(float* __restrict__ mine, const ShardTable peers, long long lo, long long n) {
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const float inv = 1.f / static_cast<float>(peers.nshards + 1);
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n; i += stride) {
    const long long e = lo + i;
    const float m = mine[e];
    float sum = m;
    for (int k = 0; k < peers.nshards; ++k) sum += __ldcg(peers.server[k] + e);
    atomicAdd(mine + e, sum * inv - m);
  }
}

}  // namespace sg

using namespace sg;

static int fill_table(ShardTable* t, const SgShardTable* a, const char* who) {
  if (a == nullptr || a->nshards < 1 || a->nshards > SG_PS_MAX_WORLD || a->per <= 0) {
    set_error("%s: bad shard table", who);
    return 2;
  }
  for (int k = 0; k < SG_PS_MAX_WORLD; ++k) t->server[k] = k < a->nshards ? a->server[k] : nullptr;
  for (int k = 0; k < a->nshards; ++k)
    if (t->server[k] == nullptr) { set_error("%s: null server pointer for shard %d", who, k); return 2; }
  t->nshards = a->nshards;
  t->per = a->per;
  return 0;
}

static unsigned sweep_grid(long long work) {
  long long blocks = (work + 255) / 256;
  const long long cap = static_cast<long long>(num_sms()) * 8;
  if (blocks > cap) blocks = cap;
  return static_cast<unsigned>(blocks < 1 ? 1 : blocks);
}

extern "C" int sg_easgd_sync_sharded(float* w, const SgShardTable* table, long long n, float alpha, void* w16,
                                     void* stream) {
  if (n <= 0) return 0;
  ShardTable t;
  if (int rc = fill_table(&t, table, "sg_easgd_sync_sharded")) return rc;
  easgd_sharded_kernel<<<sweep_grid(n), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(w, t, n, alpha, (__nv_bfloat16*)w16);
  SG_CHECK_LAUNCH("easgd_sync_sharded");
  return 0;
}

extern "C" int sg_randomsync_sharded(float* w, float* snap, const SgShardTable* table, long long n, float ratio,
                                     unsigned long long seed, const unsigned long long* seed_dev, void* w16,
                                     unsigned long long* nselected, void* stream) {
  if (n <= 0) return 0;
  ShardTable t;
  if (int rc = fill_table(&t, table, "sg_randomsync_sharded")) return rc;
  randomsync_sharded_kernel<<<sweep_grid((n + 3) / 4), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      w, snap, t, n, ratio, seed, seed_dev, (__nv_bfloat16*)w16, nselected);
  SG_CHECK_LAUNCH("randomsync_sharded");
  return 0;
}

extern "C" int sg_group_average(float* mine, const SgShardTable* peers, long long lo, long long n, void* stream) {
  if (n <= 0 || peers == nullptr || peers->nshards == 0) return 0;
  SgShardTable a = *peers;
  if (a.per <= 0) a.per = 1;
  ShardTable t;
  if (int rc = fill_table(&t, &a, "sg_group_average")) return rc;
  group_average_kernel<<<sweep_grid(n), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(mine, t, lo, n);
  SG_CHECK_LAUNCH("group_average");
  return 0;
}

extern "C" int sg_update_arena(float* w, const float* g, float* h, float* u, void* w16,
                               const int* chunk_param, const float* lr_mult, const float* wd_mult,
                               long long first_chunk, long long nchunks, const SgUpdater* up,
                               void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (nchunks <= 0) return 0;
  SG_REQUIRE(up != nullptr, "sg_update_arena: null updater");
  SG_REQUIRE(up->kind >= SG_UPD_SGD && up->kind <= SG_UPD_ADADELTA, "sg_update_arena: unknown updater kind %d", up->kind);
  const bool need_h = up->kind != SG_UPD_SGD || up->momentum > 0.f;
  SG_REQUIRE(!need_h || h != nullptr, "sg_update_arena: updater needs a history buffer");
  SG_REQUIRE(up->kind != SG_UPD_ADADELTA || u != nullptr, "sg_update_arena: AdaDelta needs an update buffer");
  SG_REQUIRE(nchunks < 2147483647LL, "sg_update_arena: too many chunks");
  UpdHyper hp{up->kind, up->lr, up->momentum, up->weight_decay, up->delta, up->rho, up->grad_scale, up->lr_dev};
  update_arena_kernel<<<static_cast<unsigned>(nchunks), 256, 0, s>>>(w, g, h, u, (__nv_bfloat16*)w16, chunk_param, lr_mult,
                                                                    wd_mult, first_chunk, hp);
  SG_CHECK_LAUNCH("update_arena");
  return 0;
}

extern "C" int sg_easgd_sync(float* w, float* server, long long n, float alpha, void* w16, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (n <= 0) return 0;
  long long blocks = (n + 255) / 256;
  const long long cap = static_cast<long long>(num_sms()) * 8;
  if (blocks > cap) blocks = cap;
  easgd_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>(w, server, n, alpha, (__nv_bfloat16*)w16);
  SG_CHECK_LAUNCH("easgd_sync");
  return 0;
}

extern "C" int sg_randomsync(float* w, float* snap, float* server, long long n, float ratio,
                             unsigned long long seed, void* w16, unsigned long long* nselected,
                             void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (n <= 0) return 0;
  long long blocks = ((n + 3) / 4 + 255) / 256;
  const long long cap = static_cast<long long>(num_sms()) * 8;
  if (blocks > cap) blocks = cap;
  randomsync_kernel<<<static_cast<unsigned>(blocks), 256, 0, s>>>(w, snap, server, n, ratio, seed, (__nv_bfloat16*)w16, nselected);
  SG_CHECK_LAUNCH("randomsync");
  return 0;
}
