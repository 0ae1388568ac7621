// Philox-4x32-10 counter RNG: stateless, (seed, index) -> 4 uniforms, so dropout masks can be regenerated
// in backward and parameter init is reproducible for any grid shape.
#pragma once
#include <stdint.h>

namespace sg {

__host__ __device__ inline void philox4x32_10(uint32_t (&ctr)[4], uint32_t k0, uint32_t k1) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = static_cast<uint64_t>(M0) * ctr[0];
    const uint64_t p1 = static_cast<uint64_t>(M1) * ctr[2];
    const uint32_t hi0 = static_cast<uint32_t>(p0 >> 32), lo0 = static_cast<uint32_t>(p0);
    const uint32_t hi1 = static_cast<uint32_t>(p1 >> 32), lo1 = static_cast<uint32_t>(p1);
    const uint32_t n0 = hi1 ^ ctr[1] ^ k0, n1 = lo1, n2 = hi0 ^ ctr[3] ^ k1, n3 = lo0;
    ctr[0] = n0; ctr[1] = n1; ctr[2] = n2; ctr[3] = n3;
    k0 += W0; k1 += W1;
  }
}

// 4 uniforms in [0,1) for counter `idx` under `seed`.
__host__ __device__ inline void philox_uniform4(unsigned long long seed, unsigned long long idx,
                                                float (&u)[4]) {
  uint32_t c[4] = {static_cast<uint32_t>(idx), static_cast<uint32_t>(idx >> 32), 0u, 0u};
  philox4x32_10(c, static_cast<uint32_t>(seed), static_cast<uint32_t>(seed >> 32));
#pragma unroll
  for (int i = 0; i < 4; ++i) u[i] = (c[i] >> 8) * (1.0f / 16777216.0f);
}

}  // namespace sg
