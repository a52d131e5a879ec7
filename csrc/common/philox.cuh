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

// ---- 8-bit lanes: sixteen elements per Philox call (attention probabilities) ---------------------------------
// The attention kernels draw one byte per score: a call covers 16 consecutive kv positions of a query row, which
// halves the generator's share of the softmax instruction stream (the 16-bit form above costs ~9 thread
// instructions per element there, RNG + compare + select).  An element is dropped when its byte < round(p * 256);
// p = 0.1 is realised as 26/256 = 0.1016 and the kernels rescale by the REALISED keep probability
// (dropout_keep_prob8), so the expectation of the output is exact.
__host__ __device__ __forceinline__ uint32_t dropout_threshold8(float p) {
  float t = p * 256.0f + 0.5f;
  return t <= 0.f ? 0u : (t >= 255.f ? 255u : (uint32_t)t);
}
__host__ __device__ __forceinline__ float dropout_keep_prob8(float p) {
  return 1.0f - float(dropout_threshold8(p)) * (1.0f / 256.0f);
}
// SWAR unsigned compare of the four bytes of w against a threshold replicated into every byte of t4:
// bit 7 of byte b of the result = (w.b >= t.b).  (w | H) - (t & ~H) compares the low 7 bits without borrowing
// across bytes; the top bits are merged with one three-input logic op.
__device__ __forceinline__ uint32_t bytes_ge_bit7(uint32_t w, uint32_t t4) {
  const uint32_t z = (w | 0x80808080u) - (t4 & 0x7F7F7F7Fu);
  return (w & ~t4) | (~(w ^ t4) & z);
}
// keep-masks for bf16x2 words: lo = elements (4i, 4i+1), hi = elements (4i+2, 4i+3) of word w
// (0xFFFF per kept element); prmt with the sign-replicate selector expands bit 7 of a byte.
__device__ __forceinline__ void keep_masks_bf16x2(uint32_t w, uint32_t t4, uint32_t& lo, uint32_t& hi) {
  const uint32_t c = bytes_ge_bit7(w, t4);
  // (inline PTX: __byte_perm() documents only the low 3 bits of a selector nibble, the replicate bit is the 4th)
  asm("prmt.b32 %0, %1, 0, 0x9988;" : "=r"(lo) : "r"(c));
  asm("prmt.b32 %0, %1, 0, 0xBBAA;" : "=r"(hi) : "r"(c));
}

// 32-bit all-ones / all-zeros mask from bit 7 of byte B of c (for selecting fp32 values with logic ops)
template <int B>
__device__ __forceinline__ uint32_t byte_sign_mask(uint32_t c) {
  uint32_t m;
  asm("prmt.b32 %0, %1, 0, %2;" : "=r"(m) : "r"(c), "n"(0x8888 + 0x1111 * B));
  return m;
}

}  // namespace b200
