// Counter-based dropout RNG (Philox4x32, 7 rounds).  Stateless: the keep-mask of an element is a pure
// function of (seed, offset, stream, element-group index), so backward and activation recompute
// regenerate exactly the forward mask without storing it, and {seed, offset} live in device memory
// so a captured CUDA graph sees fresh randomness every replay.
//
// One call yields 128 random bits = eight 16-bit lanes -> eight consecutive elements.  An element is
// dropped when its lane < round(p * 65536); p = 0.1 is realised as 6554/65536 = 0.100006.
#pragma once
#include <stdint.h>

namespace b200 {

__host__ __device__ __forceinline__ uint32_t dropout_threshold16(float p) {
  float t = p * 65536.0f + 0.5f;
  return t <= 0.f ? 0u : (t >= 65535.f ? 65535u : (uint32_t)t);
}

__device__ __forceinline__ uint4 philox4x32_7(uint4 c, uint2 k) {
  constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 7; ++r) {
    const uint32_t hi0 = __umulhi(M0, c.x), lo0 = M0 * c.x;
    const uint32_t hi1 = __umulhi(M1, c.z), lo1 = M1 * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += W0;
    k.y += W1;
  }
  return c;
}

// 128 random bits for element group `group` of op `stream`.
__device__ __forceinline__ uint4 dropout_bits(unsigned long long seed, unsigned long long offset, uint32_t stream,
                                              unsigned long long group) {
  uint4 ctr = make_uint4((uint32_t)group, (uint32_t)(group >> 32), (uint32_t)offset,
                         (uint32_t)(offset >> 32) ^ (stream * 0x9E3779B1u));
  uint2 key = make_uint2((uint32_t)seed, (uint32_t)(seed >> 32) ^ stream);
  return philox4x32_7(ctr, key);
}

// bit e of the result = 1 when element 8*group+e is KEPT.
__device__ __forceinline__ uint32_t dropout_keep8(unsigned long long seed, unsigned long long offset, uint32_t stream,
                                                  unsigned long long group, uint32_t thresh) {
  const uint4 r = dropout_bits(seed, offset, stream, group);
  const uint32_t w[4] = {r.x, r.y, r.z, r.w};
  uint32_t keep = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    keep |= ((w[i] & 0xFFFFu) >= thresh ? 1u : 0u) << (2 * i);
    keep |= ((w[i] >> 16) >= thresh ? 1u : 0u) << (2 * i + 1);
  }
  return keep;
}

// Select form of the same mask for the attention kernels: zeroes the dropped ones among eight values
// without materialising the bit mask.  thresh_hi = threshold << 16; a 16-bit lane is kept when
// lane >= threshold, which for the high lane is  w >= thresh_hi  and for the low lane
// (w << 16) >= thresh_hi  -- one compare (plus a shift for the low lane) and one select per element.
__device__ __forceinline__ void dropout_zero8(const uint4 r, uint32_t thresh_hi, float (&v)[8]) {
  const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    v[2 * i] = ((w[i] << 16) >= thresh_hi) ? v[2 * i] : 0.f;
    v[2 * i + 1] = (w[i] >= thresh_hi) ? v[2 * i + 1] : 0.f;
  }
}

}  // namespace b200
