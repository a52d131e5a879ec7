// Bandwidth-bound kernels: embedding gather/scatter, RMSNorm fwd/bwd, dropout+residual, rotary,
// SwiGLU.  Each is a single pass with 128-bit accesses; the reference issues 6-12 ATen launches for
// the same math.  Where a GEMM epilogue already fuses the op
// (B200_EPI_ROPE_BF16 / RESID_DROPOUT / SWIGLU*) the kernel here is the unfused companion used for
// validation and for the layouts the epilogues do not cover.
#include <cuda_bf16.h>
#include <stdlib.h>

#include "../include/b200_trainer.h"
#include "common/host.h"
#include "common/philox.cuh"
#include "common/ptx.cuh"

namespace b200 {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ void red_add_v4(float* addr, float4 v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

// ------------------------------------------------------------------------------------------------ embedding
// one warp per token row; hidden % 8 == 0
__global__ void embed_fwd_kernel(const int64_t* __restrict__ ids, const __nv_bfloat16* __restrict__ w,
                                 float* __restrict__ out, int rows, int hidden, int vocab) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  long long id = ids[row];
  id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
  const uint4* src = reinterpret_cast<const uint4*>(w + size_t(id) * hidden);
  float4* dst = reinterpret_cast<float4*>(out + size_t(row) * hidden);
  for (int i = lane_id(); i < hidden / 8; i += 32) {
    const uint4 q = src[i];
    const float2 a = unpack_bf16x2(q.x), b = unpack_bf16x2(q.y), c = unpack_bf16x2(q.z), d = unpack_bf16x2(q.w);
    dst[2 * i] = make_float4(a.x, a.y, b.x, b.y);
    dst[2 * i + 1] = make_float4(c.x, c.y, d.x, d.y);
  }
}

__global__ void embed_bwd_kernel(const int64_t* __restrict__ ids, const float* __restrict__ dx,
                                 float* __restrict__ dw, int rows, int hidden, int vocab) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  long long id = ids[row];
  if (id < 0 || id >= vocab) return;
  const float4* src = reinterpret_cast<const float4*>(dx + size_t(row) * hidden);
  float* dst = dw + size_t(id) * hidden;
  for (int i = lane_id(); i < hidden / 4; i += 32) red_add_v4(dst + 4 * i, src[i]);
}

// ------------------------------------------------------------------------------------------------ RMSNorm
// One warp per row.  NV = ceil(hidden / 256) > 0: the row is requested up front (2 NV float4 loads in flight per lane)
// and stays in registers for the sum of squares and the scaling -- one read of x.  NV = 0: generic two-pass fallback
// for rows wider than 2048 (the second read hits L1).  hidden % 8 == 0.
template <int NV>
__global__ void __launch_bounds__(256)
rmsnorm_fwd_kernel(const float* __restrict__ x, const __nv_bfloat16* __restrict__ w, __nv_bfloat16* __restrict__ y,
                   float* __restrict__ rstd, int rows, int hidden, float eps) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const float4* xr = reinterpret_cast<const float4*>(x + size_t(row) * hidden);
  const uint4* wr = reinterpret_cast<const uint4*>(w);
  uint4* yr = reinterpret_cast<uint4*>(y + size_t(row) * hidden);
  const int nvec = hidden / 8, lane = lane_id();
  auto scaled = [](const float4 a, const float4 b, const uint4 wq, float r) {
    const float2 w0 = unpack_bf16x2(wq.x), w1 = unpack_bf16x2(wq.y), w2 = unpack_bf16x2(wq.z), w3 = unpack_bf16x2(wq.w);
    return make_uint4(pack_bf16x2(a.x * r * w0.x, a.y * r * w0.y), pack_bf16x2(a.z * r * w1.x, a.w * r * w1.y),
                      pack_bf16x2(b.x * r * w2.x, b.y * r * w2.y), pack_bf16x2(b.z * r * w3.x, b.w * r * w3.y));
  };
  if constexpr (NV > 0) {
    float4 xa[NV], xb[NV];
    uint4 wq[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int i = lane + 32 * k;
      if (i < nvec) {
        xa[k] = xr[2 * i];
        xb[k] = xr[2 * i + 1];
        wq[k] = wr[i];
      } else {
        xa[k] = xb[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        wq[k] = make_uint4(0, 0, 0, 0);
      }
    }
    float ss = 0.f;
#pragma unroll
    for (int k = 0; k < NV; ++k)
      ss += (xa[k].x * xa[k].x + xa[k].y * xa[k].y) + (xa[k].z * xa[k].z + xa[k].w * xa[k].w) +
            (xb[k].x * xb[k].x + xb[k].y * xb[k].y) + (xb[k].z * xb[k].z + xb[k].w * xb[k].w);
    ss = warp_sum(ss);
    const float r = rsqrtf(ss / float(hidden) + eps);
    if (lane == 0) rstd[row] = r;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int i = lane + 32 * k;
      if (i < nvec) yr[i] = scaled(xa[k], xb[k], wq[k], r);
    }
  } else {
    float ss = 0.f;
    for (int i = lane; i < hidden / 4; i += 32) {
      const float4 v = xr[i];
      ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    ss = warp_sum(ss);
    const float r = rsqrtf(ss / float(hidden) + eps);
    if (lane == 0) rstd[row] = r;
    for (int i = lane; i < nvec; i += 32) yr[i] = scaled(xr[2 * i], xr[2 * i + 1], wr[i], r);
  }
}

// Backward.  With g = dy*w, r = rstd:  dx += r*g - x * r^3 * mean(g*x) ;  dw += sum_rows dy * x * r.
// Persistent warps stride over rows and keep their dw partials in registers (columns owned by a lane
// are fixed), then one smem reduction + one red.add per column per CTA.
constexpr int kNormMaxVec = 8;  // 8 x (8 columns) x 32 lanes = 2048 columns max
// NV = ceil(hidden / 256): vectors of 8 columns per lane.  A row's dy / x / dx are all requested up
// front (NV x 80 bytes per lane in flight) and stay in registers for both the dot product and the
// update, so the row is read exactly once; w is loaded once per warp.
template <int NV>
__global__ void __launch_bounds__(256, NV <= 3 ? 2 : 1)
rmsnorm_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const float* __restrict__ x,
                   const __nv_bfloat16* __restrict__ w, const float* __restrict__ rstd, float* __restrict__ dx,
                   float* __restrict__ dw, int rows, int hidden, __nv_bfloat16* __restrict__ dy_next, float p_next,
                   const unsigned long long* __restrict__ rng, uint32_t rng_stream) {
  // dy_next (optional): the updated dx, dropout-masked with stream `rng_stream` and cast to bf16 -- the
  // dY of the residual branch that ends at this point of the stream (what dropout_bwd_cast would produce
  // from dx in a second pass over it).
  extern __shared__ float s_dw[];  // [hidden]
  for (int i = threadIdx.x; i < hidden; i += blockDim.x) s_dw[i] = 0.f;
  __syncthreads();
  const int warps = blockDim.x >> 5, wid = threadIdx.x >> 5, lane = lane_id();
  const int nvec = hidden / 8;
  float acc[NV][8];
  uint4 wq[NV];   // this lane's weight columns, kept packed (unpacking is a shift / mask)
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    const int i = lane + 32 * k;
    wq[k] = i < nvec ? reinterpret_cast<const uint4*>(w)[i] : make_uint4(0, 0, 0, 0);
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[k][e] = 0.f;
  }
  const float inv_h = 1.f / float(hidden);
  for (int row = blockIdx.x * warps + wid; row < rows; row += gridDim.x * warps) {
    const uint4* dyr = reinterpret_cast<const uint4*>(dy + size_t(row) * hidden);
    const float4* xr = reinterpret_cast<const float4*>(x + size_t(row) * hidden);
    float4* dxr = reinterpret_cast<float4*>(dx + size_t(row) * hidden);
    uint4 dq[NV];
    float4 xa[NV], xb[NV], oa[NV], ob[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int i = lane + 32 * k;
      if (i < nvec) {
        dq[k] = dyr[i];
        xa[k] = xr[2 * i];
        xb[k] = xr[2 * i + 1];
        oa[k] = dxr[2 * i];
        ob[k] = dxr[2 * i + 1];
      } else {
        dq[k] = make_uint4(0, 0, 0, 0);
        xa[k] = xb[k] = oa[k] = ob[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
    const float r = rstd[row];
    float dot = 0.f;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const float xs[8] = {xa[k].x, xa[k].y, xa[k].z, xa[k].w, xb[k].x, xb[k].y, xb[k].z, xb[k].w};
      const uint32_t dqw[4] = {dq[k].x, dq[k].y, dq[k].z, dq[k].w};
      const uint32_t wqw[4] = {wq[k].x, wq[k].y, wq[k].z, wq[k].w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 d2 = unpack_bf16x2(dqw[e]), w2 = unpack_bf16x2(wqw[e]);
        dot += d2.x * w2.x * xs[2 * e] + d2.y * w2.y * xs[2 * e + 1];
        acc[k][2 * e] += d2.x * xs[2 * e] * r;
        acc[k][2 * e + 1] += d2.y * xs[2 * e + 1] * r;
      }
    }
    dot = warp_sum(dot);
    const float coef = dot * r * r * r * inv_h;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int i = lane + 32 * k;
      if (i < nvec) {
        const float2 d0 = unpack_bf16x2(dq[k].x), d1 = unpack_bf16x2(dq[k].y), d2 = unpack_bf16x2(dq[k].z), d3 = unpack_bf16x2(dq[k].w);
        const float2 w0 = unpack_bf16x2(wq[k].x), w1 = unpack_bf16x2(wq[k].y), w2 = unpack_bf16x2(wq[k].z), w3 = unpack_bf16x2(wq[k].w);
        oa[k].x += r * d0.x * w0.x - xa[k].x * coef;
        oa[k].y += r * d0.y * w0.y - xa[k].y * coef;
        oa[k].z += r * d1.x * w1.x - xa[k].z * coef;
        oa[k].w += r * d1.y * w1.y - xa[k].w * coef;
        ob[k].x += r * d2.x * w2.x - xb[k].x * coef;
        ob[k].y += r * d2.y * w2.y - xb[k].y * coef;
        ob[k].z += r * d3.x * w3.x - xb[k].z * coef;
        ob[k].w += r * d3.y * w3.y - xb[k].w * coef;
        dxr[2 * i] = oa[k];
        dxr[2 * i + 1] = ob[k];
        if (dy_next != nullptr) {
          float o[8] = {oa[k].x, oa[k].y, oa[k].z, oa[k].w, ob[k].x, ob[k].y, ob[k].z, ob[k].w};
          if (p_next > 0.f) {
            const float inv = 1.f / (1.f - p_next);
#pragma unroll
            for (int e = 0; e < 8; ++e) o[e] *= inv;
            dropout_zero8(dropout_bits(rng[0], rng[1], rng_stream, (unsigned long long)row * nvec + i),
                          dropout_threshold16(p_next) << 16, o);
          }
          reinterpret_cast<uint4*>(dy_next + size_t(row) * hidden)[i] =
              make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7]));
        }
      }
    }
  }
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    const int i = lane + 32 * k;
    if (i < nvec) {
#pragma unroll
      for (int e = 0; e < 8; ++e) atomicAdd(&s_dw[8 * i + e], acc[k][e]);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < hidden; i += blockDim.x) atomicAdd(&dw[i], s_dw[i]);
}

// Same math with R = hidden / 256 warps per row, ONE vector of 8 columns per lane (used when hidden is a multiple of
// 256 and R <= 4): 20 registers of row data per thread instead of 20 x NV, so ~64 registers and four times the
// warps per SM of the kernel above (128 registers, 16 warps per SM: 24 % of the warp slots, 4.1 TB/s) -- while one
// row's warps wait for their loads, others compute or store.  The R partial dot products meet in shared memory behind a
// 32 R-thread named barrier per row (two slots, alternating, so one barrier is enough).
constexpr int kSplitGroups = 4;   // rows in flight per CTA
template <int R>
__global__ void __launch_bounds__(32 * R * kSplitGroups, R <= 3 ? 3 : 2)
rmsnorm_bwd_split_kernel(const __nv_bfloat16* __restrict__ dy, const float* __restrict__ x,
                         const __nv_bfloat16* __restrict__ w, const float* __restrict__ rstd, float* __restrict__ dx,
                         float* __restrict__ dw, int rows, int hidden, __nv_bfloat16* __restrict__ dy_next, float p_next,
                         const unsigned long long* __restrict__ rng, uint32_t rng_stream) {
  extern __shared__ float s_dw[];  // [hidden]
  __shared__ float s_dot[2][kSplitGroups][R];
  for (int i = threadIdx.x; i < hidden; i += blockDim.x) s_dw[i] = 0.f;
  __syncthreads();
  const int wid = threadIdx.x >> 5, lane = lane_id();
  const int grp = wid / R, wr = wid - grp * R;      // row group inside the CTA, warp inside the row
  const int i = wr * 32 + lane;                     // this lane's vector (8 columns) of every row
  const int nvec = hidden / 8;
  const uint4 wq = reinterpret_cast<const uint4*>(w)[i];
  const uint32_t wqw[4] = {wq.x, wq.y, wq.z, wq.w};
  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = 0.f;
  const float inv_h = 1.f / float(hidden);
  const float inv_keep = p_next > 0.f ? 1.f / (1.f - p_next) : 1.f;
  int it = 0;
  for (int row = blockIdx.x * kSplitGroups + grp; row < rows; row += gridDim.x * kSplitGroups, ++it) {
    const uint4 dq = reinterpret_cast<const uint4*>(dy + size_t(row) * hidden)[i];
    const float4* xr = reinterpret_cast<const float4*>(x + size_t(row) * hidden);
    float4* dxr = reinterpret_cast<float4*>(dx + size_t(row) * hidden);
    const float4 xa = xr[2 * i], xb = xr[2 * i + 1];
    float4 oa = dxr[2 * i], ob = dxr[2 * i + 1];
    const float r = rstd[row];
    const float xs[8] = {xa.x, xa.y, xa.z, xa.w, xb.x, xb.y, xb.z, xb.w};
    const uint32_t dqw[4] = {dq.x, dq.y, dq.z, dq.w};
    float g[8];   // dy * w
    float dot = 0.f;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 d2 = unpack_bf16x2(dqw[e]), w2 = unpack_bf16x2(wqw[e]);
      g[2 * e] = d2.x * w2.x;
      g[2 * e + 1] = d2.y * w2.y;
      dot += g[2 * e] * xs[2 * e] + g[2 * e + 1] * xs[2 * e + 1];
      acc[2 * e] += d2.x * xs[2 * e] * r;
      acc[2 * e + 1] += d2.y * xs[2 * e + 1] * r;
    }
    dot = warp_sum(dot);
    if (R > 1) {
      if (lane == 0) s_dot[it & 1][grp][wr] = dot;
      asm volatile("bar.sync %0, %1;" ::"r"(grp + 1), "n"(32 * R) : "memory");
      dot = 0.f;
#pragma unroll
      for (int t = 0; t < R; ++t) dot += s_dot[it & 1][grp][t];
    }
    const float coef = dot * r * r * r * inv_h;
    float o[8] = {oa.x, oa.y, oa.z, oa.w, ob.x, ob.y, ob.z, ob.w};
#pragma unroll
    for (int e = 0; e < 8; ++e) o[e] += r * g[e] - xs[e] * coef;
    dxr[2 * i] = make_float4(o[0], o[1], o[2], o[3]);
    dxr[2 * i + 1] = make_float4(o[4], o[5], o[6], o[7]);
    if (dy_next != nullptr) {
      if (p_next > 0.f) {
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] *= inv_keep;
        dropout_zero8(dropout_bits(rng[0], rng[1], rng_stream, (unsigned long long)row * nvec + i),
                      dropout_threshold16(p_next) << 16, o);
      }
      reinterpret_cast<uint4*>(dy_next + size_t(row) * hidden)[i] =
          make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7]));
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) atomicAdd(&s_dw[8 * i + e], acc[e]);
  __syncthreads();
  for (int c = threadIdx.x; c < hidden; c += blockDim.x) atomicAdd(&dw[c], s_dw[c]);
}

// ------------------------------------------------------------------------------------------------ dropout family
// thread per 8-column group; group index = row * (cols/8) + col/8 (same convention as the GEMM epilogue)
__global__ void resid_dropout_fwd_kernel(const __nv_bfloat16* __restrict__ y, const float* __restrict__ res,
                                         float* __restrict__ out, long long groups, float p,
                                         const unsigned long long* __restrict__ rng, uint32_t stream) {
  const long long g = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (g >= groups) return;
  uint32_t keep = 0xFF;
  float inv = 1.f;
  if (p > 0.f) {
    keep = dropout_keep8(rng[0], rng[1], stream, (unsigned long long)g, dropout_threshold16(p));
    inv = 1.f / (1.f - p);
  }
  const uint4 q = reinterpret_cast<const uint4*>(y)[g];
  const float4 r0 = reinterpret_cast<const float4*>(res)[2 * g], r1 = reinterpret_cast<const float4*>(res)[2 * g + 1];
  const float2 a = unpack_bf16x2(q.x), b = unpack_bf16x2(q.y), c = unpack_bf16x2(q.z), d = unpack_bf16x2(q.w);
  const float v[8] = {a.x, a.y, b.x, b.y, c.x, c.y, d.x, d.y};
  float o[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) o[e] = ((keep >> e) & 1) ? v[e] * inv : 0.f;
  reinterpret_cast<float4*>(out)[2 * g] = make_float4(r0.x + o[0], r0.y + o[1], r0.z + o[2], r0.w + o[3]);
  reinterpret_cast<float4*>(out)[2 * g + 1] = make_float4(r1.x + o[4], r1.y + o[5], r1.z + o[6], r1.w + o[7]);
}

__global__ void dropout_bwd_cast_kernel(const float* __restrict__ dx, __nv_bfloat16* __restrict__ dy, long long groups,
                                        float p, const unsigned long long* __restrict__ rng, uint32_t stream) {
  const long long g = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (g >= groups) return;
  uint32_t keep = 0xFF;
  float inv = 1.f;
  if (p > 0.f) {
    keep = dropout_keep8(rng[0], rng[1], stream, (unsigned long long)g, dropout_threshold16(p));
    inv = 1.f / (1.f - p);
  }
  const float4 a = reinterpret_cast<const float4*>(dx)[2 * g], b = reinterpret_cast<const float4*>(dx)[2 * g + 1];
  const float v[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
  float o[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) o[e] = ((keep >> e) & 1) ? v[e] * inv : 0.f;
  reinterpret_cast<uint4*>(dy)[g] =
      make_uint4(pack_bf16x2(o[0], o[1]), pack_bf16x2(o[2], o[3]), pack_bf16x2(o[4], o[5]), pack_bf16x2(o[6], o[7]));
}

__global__ void dropout_mask_kernel(uint8_t* __restrict__ mask, long long groups, float p,
                                    const unsigned long long* __restrict__ rng, uint32_t stream) {
  const long long g = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (g >= groups) return;
  uint32_t keep = 0xFF;
  if (p > 0.f) keep = dropout_keep8(rng[0], rng[1], stream, (unsigned long long)g, dropout_threshold16(p));
#pragma unroll
  for (int e = 0; e < 8; ++e) mask[8 * g + e] = (keep >> e) & 1;
}

// attention flavour: groups of 16 along the row (ceil(cols/16) groups per row), 8-bit lanes
__global__ void attn_dropout_mask_kernel(uint8_t* __restrict__ mask, long long rows, int cols, float p,
                                         const unsigned long long* __restrict__ rng, uint32_t stream) {
  const int g16 = (cols + 15) / 16;
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= rows * g16) return;
  const long long row = t / g16;
  const int c0 = int(t % g16) * 16;
  uint4 r = make_uint4(~0u, ~0u, ~0u, ~0u);
  const uint32_t thresh = dropout_threshold8(p);
  if (p > 0.f) r = dropout_bits(rng[0], rng[1], stream, (unsigned long long)t);
  const uint32_t w[4] = {r.x, r.y, r.z, r.w};
  for (int e = 0; e < 16 && c0 + e < cols; ++e)
    mask[row * cols + c0 + e] = (((w[e >> 2] >> ((e & 3) * 8)) & 0xFFu) >= thresh) ? 1 : 0;
}

// ------------------------------------------------------------------------------------------------ cast
__global__ void cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long n8) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= n8) return;
  const float4 a = reinterpret_cast<const float4*>(src)[2 * i], b = reinterpret_cast<const float4*>(src)[2 * i + 1];
  reinterpret_cast<uint4*>(dst)[i] =
      make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w));
}

// ------------------------------------------------------------------------------------------------ rotary
// thread per (row, head, 8-wide group of the first half); pairs column j with j+32 (head_dim 64)
__global__ void rope_inplace_kernel(__nv_bfloat16* __restrict__ qkv, const float* __restrict__ cs,
                                    const float* __restrict__ sn, long long total, int ld, int heads2, int seq,
                                    int inverse) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= total) return;
  const int g = int(t & 3);
  const long long rh = t >> 2;
  const int head = int(rh % heads2);
  const long long row = rh / heads2;
  const int pos = int(row % seq);
  __nv_bfloat16* base = qkv + row * ld + head * 64 + g * 8;
  const uint4 lo = *reinterpret_cast<const uint4*>(base), hi = *reinterpret_cast<const uint4*>(base + 32);
  const float4 c0 = reinterpret_cast<const float4*>(cs + size_t(pos) * 32)[2 * g], c1 = reinterpret_cast<const float4*>(cs + size_t(pos) * 32)[2 * g + 1];
  float4 s0 = reinterpret_cast<const float4*>(sn + size_t(pos) * 32)[2 * g], s1 = reinterpret_cast<const float4*>(sn + size_t(pos) * 32)[2 * g + 1];
  const float cc[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
  float ss[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
  if (inverse) {
#pragma unroll
    for (int e = 0; e < 8; ++e) ss[e] = -ss[e];
  }
  const uint32_t low[4] = {lo.x, lo.y, lo.z, lo.w}, hiw[4] = {hi.x, hi.y, hi.z, hi.w};
  uint32_t ol[4], oh[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const float2 a = unpack_bf16x2(low[e]), b = unpack_bf16x2(hiw[e]);
    ol[e] = pack_bf16x2(a.x * cc[2 * e] - b.x * ss[2 * e], a.y * cc[2 * e + 1] - b.y * ss[2 * e + 1]);
    oh[e] = pack_bf16x2(b.x * cc[2 * e] + a.x * ss[2 * e], b.y * cc[2 * e + 1] + a.y * ss[2 * e + 1]);
  }
  *reinterpret_cast<uint4*>(base) = make_uint4(ol[0], ol[1], ol[2], ol[3]);
  *reinterpret_cast<uint4*>(base + 32) = make_uint4(oh[0], oh[1], oh[2], oh[3]);
}

// ------------------------------------------------------------------------------------------------ SwiGLU
__global__ void swiglu_fwd_kernel(const __nv_bfloat16* __restrict__ gu, __nv_bfloat16* __restrict__ act,
                                  long long groups, int inter) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= groups) return;
  const int gpr = inter / 8;
  const long long row = t / gpr;
  const int c = int(t % gpr) * 8;
  const uint4 g = *reinterpret_cast<const uint4*>(gu + row * 2 * inter + c);
  const uint4 u = *reinterpret_cast<const uint4*>(gu + row * 2 * inter + inter + c);
  const uint32_t gw[4] = {g.x, g.y, g.z, g.w}, uw[4] = {u.x, u.y, u.z, u.w};
  uint32_t o[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const float2 a = unpack_bf16x2(gw[e]), b = unpack_bf16x2(uw[e]);
    o[e] = pack_bf16x2(a.x / (1.f + __expf(-a.x)) * b.x, a.y / (1.f + __expf(-a.y)) * b.y);
  }
  *reinterpret_cast<uint4*>(act + row * inter + c) = make_uint4(o[0], o[1], o[2], o[3]);
}

__global__ void swiglu_bwd_kernel(const __nv_bfloat16* __restrict__ dact, const __nv_bfloat16* __restrict__ gu,
                                  __nv_bfloat16* __restrict__ dgu, long long groups, int inter) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= groups) return;
  const int gpr = inter / 8;
  const long long row = t / gpr;
  const int c = int(t % gpr) * 8;
  const uint4 g = *reinterpret_cast<const uint4*>(gu + row * 2 * inter + c);
  const uint4 u = *reinterpret_cast<const uint4*>(gu + row * 2 * inter + inter + c);
  const uint4 d = *reinterpret_cast<const uint4*>(dact + row * inter + c);
  const uint32_t gw[4] = {g.x, g.y, g.z, g.w}, uw[4] = {u.x, u.y, u.z, u.w}, dw[4] = {d.x, d.y, d.z, d.w};
  uint32_t og[4], ou[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const float2 a = unpack_bf16x2(gw[e]), b = unpack_bf16x2(uw[e]), dd = unpack_bf16x2(dw[e]);
    const float s0 = 1.f / (1.f + __expf(-a.x)), s1 = 1.f / (1.f + __expf(-a.y));
    og[e] = pack_bf16x2(dd.x * b.x * s0 * (1.f + a.x * (1.f - s0)), dd.y * b.y * s1 * (1.f + a.y * (1.f - s1)));
    ou[e] = pack_bf16x2(dd.x * a.x * s0, dd.y * a.y * s1);
  }
  *reinterpret_cast<uint4*>(dgu + row * 2 * inter + c) = make_uint4(og[0], og[1], og[2], og[3]);
  *reinterpret_cast<uint4*>(dgu + row * 2 * inter + inter + c) = make_uint4(ou[0], ou[1], ou[2], ou[3]);
}

}  // namespace b200

using namespace b200;
#define STREAM(s) reinterpret_cast<cudaStream_t>(s)

extern "C" {

int b200_embed_fwd(const int64_t* ids, const void* w, float* out, int rows, int hidden, int vocab, void* stream) {
  B200_REQUIRE(hidden % 8 == 0, "embed: hidden %% 8 != 0");
  embed_fwd_kernel<<<(rows + 7) / 8, 256, 0, STREAM(stream)>>>(ids, (const __nv_bfloat16*)w, out, rows, hidden, vocab);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_embed_bwd(const int64_t* ids, const float* dx, float* dw, int rows, int hidden, int vocab, void* stream) {
  B200_REQUIRE(hidden % 4 == 0, "embed: hidden %% 4 != 0");
  embed_bwd_kernel<<<(rows + 7) / 8, 256, 0, STREAM(stream)>>>(ids, dx, dw, rows, hidden, vocab);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_rmsnorm_fwd(const float* x, const void* w, void* y, float* rstd, int rows, int hidden, float eps,
                     void* stream) {
  B200_REQUIRE(hidden % 8 == 0, "rmsnorm: hidden %% 8 != 0");
#define B200_NORM_FWD(NV)                                                                                          \
  rmsnorm_fwd_kernel<NV><<<(rows + 7) / 8, 256, 0, STREAM(stream)>>>(x, (const __nv_bfloat16*)w, (__nv_bfloat16*)y, rstd, \
                                                                      rows, hidden, eps)
  switch ((hidden + 255) / 256) {
    case 1: B200_NORM_FWD(1); break;
    case 2: B200_NORM_FWD(2); break;
    case 3: B200_NORM_FWD(3); break;
    case 4: B200_NORM_FWD(4); break;
    case 5: B200_NORM_FWD(5); break;
    case 6: B200_NORM_FWD(6); break;
    case 7: B200_NORM_FWD(7); break;
    case 8: B200_NORM_FWD(8); break;
    default: B200_NORM_FWD(0);
  }
#undef B200_NORM_FWD
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_rmsnorm_bwd(const void* dy, const float* x, const void* w, const float* rstd, float* dx, float* dw, int rows,
                     int hidden, void* stream) {
  return b200_rmsnorm_bwd_dropout(dy, x, w, rstd, dx, dw, rows, hidden, nullptr, 0.f, nullptr, 0, stream);
}

int b200_rmsnorm_bwd_dropout(const void* dy, const float* x, const void* w, const float* rstd, float* dx, float* dw,
                             int rows, int hidden, void* dy_next_, float p_next, const unsigned long long* rng,
                             unsigned int rng_stream, void* stream) {
  B200_REQUIRE(hidden % 8 == 0 && hidden <= kNormMaxVec * 256, "rmsnorm_bwd: hidden must be a multiple of 8 and <= %d",
               kNormMaxVec * 256);
  B200_REQUIRE(dy_next_ == nullptr || p_next == 0.f || rng != nullptr, "rmsnorm_bwd: rng_state required when p > 0");
  __nv_bfloat16* dy_next = (__nv_bfloat16*)dy_next_;
  {
    // hidden = 256 R, R <= 4: R warps per row, one vector per lane (four times the warps per SM)
    static const int env_split = [] { const char* e = getenv("B200_NORM_SPLIT"); return e ? atoi(e) : 1; }();
    const int R = hidden / 256;
    if (env_split && hidden % 256 == 0 && R >= 1 && R <= 4) {
      static const int env_grid2 = [] { const char* e = getenv("B200_NORM_GRID2"); return e ? atoi(e) : 0; }();
      int grid2 = env_grid2 > 0 ? env_grid2 : sm_count() * (R <= 3 ? 3 : 2);   // every CTA resident
      const int max_grid2 = (rows + kSplitGroups - 1) / kSplitGroups;
      if (grid2 > max_grid2) grid2 = max_grid2;
#define B200_NORM_SPLIT(RR)                                                                                         \
  case RR:                                                                                                          \
    rmsnorm_bwd_split_kernel<RR><<<grid2, 32 * RR * kSplitGroups, hidden * sizeof(float), STREAM(stream)>>>(          \
        (const __nv_bfloat16*)dy, x, (const __nv_bfloat16*)w, rstd, dx, dw, rows, hidden, dy_next, p_next, rng, rng_stream); \
    break;
      switch (R) { B200_NORM_SPLIT(1) B200_NORM_SPLIT(2) B200_NORM_SPLIT(3) B200_NORM_SPLIT(4) }
#undef B200_NORM_SPLIT
      B200_CHECK_CUDA(cudaGetLastError());
      return 0;
    }
  }
  int grid = sm_count() * 2;
  const int max_grid = (rows + 7) / 8;
  if (grid > max_grid) grid = max_grid;
  const int nv = (hidden / 8 + 31) / 32;
#define B200_NORM_BWD(NV)                                                                         \
  case NV:                                                                                        \
    rmsnorm_bwd_kernel<NV><<<grid, 256, hidden * sizeof(float), STREAM(stream)>>>(                \
        (const __nv_bfloat16*)dy, x, (const __nv_bfloat16*)w, rstd, dx, dw, rows, hidden,         \
        dy_next, p_next, rng, rng_stream);                                                        \
    break;
  switch (nv) {
    B200_NORM_BWD(1) B200_NORM_BWD(2) B200_NORM_BWD(3) B200_NORM_BWD(4)
    B200_NORM_BWD(5) B200_NORM_BWD(6) B200_NORM_BWD(7) B200_NORM_BWD(8)
    default: B200_REQUIRE(false, "rmsnorm_bwd: hidden %d too large", hidden);
  }
#undef B200_NORM_BWD
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_resid_dropout_fwd(const void* y, const float* res, float* out, int rows, int cols, float p,
                           const unsigned long long* rng, unsigned int rs, void* stream) {
  B200_REQUIRE(cols % 8 == 0, "resid_dropout: cols %% 8 != 0");
  B200_REQUIRE(p == 0.f || rng != nullptr, "resid_dropout: rng_state required when p > 0");
  const long long groups = (long long)rows * (cols / 8);
  resid_dropout_fwd_kernel<<<(unsigned)((groups + 255) / 256), 256, 0, STREAM(stream)>>>(
      (const __nv_bfloat16*)y, res, out, groups, p, rng, rs);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_dropout_bwd_cast(const float* dx, void* dy, int rows, int cols, float p, const unsigned long long* rng,
                          unsigned int rs, void* stream) {
  B200_REQUIRE(cols % 8 == 0, "dropout_bwd: cols %% 8 != 0");
  B200_REQUIRE(p == 0.f || rng != nullptr, "dropout_bwd: rng_state required when p > 0");
  const long long groups = (long long)rows * (cols / 8);
  dropout_bwd_cast_kernel<<<(unsigned)((groups + 255) / 256), 256, 0, STREAM(stream)>>>(dx, (__nv_bfloat16*)dy, groups,
                                                                                         p, rng, rs);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_dropout_mask(uint8_t* mask, int rows, int cols, float p, const unsigned long long* rng, unsigned int rs,
                      void* stream) {
  B200_REQUIRE(cols % 8 == 0, "dropout_mask: cols %% 8 != 0");
  const long long groups = (long long)rows * (cols / 8);
  dropout_mask_kernel<<<(unsigned)((groups + 255) / 256), 256, 0, STREAM(stream)>>>(mask, groups, p, rng, rs);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_attn_dropout_mask(uint8_t* mask, int rows, int cols, float p, const unsigned long long* rng, unsigned int rs,
                           void* stream) {
  const long long groups = (long long)rows * ((cols + 15) / 16);
  attn_dropout_mask_kernel<<<(unsigned)((groups + 255) / 256), 256, 0, STREAM(stream)>>>(mask, rows, cols, p, rng, rs);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_cast_f32_bf16(const float* src, void* dst, long long n, void* stream) {
  B200_REQUIRE(n % 8 == 0, "cast: n %% 8 != 0");
  if (n == 0) return 0;
  cast_f32_bf16_kernel<<<(unsigned)((n / 8 + 255) / 256), 256, 0, STREAM(stream)>>>(src, (__nv_bfloat16*)dst, n / 8);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_rope_inplace(void* qkv, const float* cs, const float* sn, int rows, int ld, int rope_cols, int seq_len,
                      int inverse, void* stream) {
  B200_REQUIRE(rope_cols % 64 == 0 && ld % 8 == 0, "rope: rope_cols %% 64 / ld %% 8");
  const int heads2 = rope_cols / 64;
  const long long total = (long long)rows * heads2 * 4;
  rope_inplace_kernel<<<(unsigned)((total + 255) / 256), 256, 0, STREAM(stream)>>>((__nv_bfloat16*)qkv, cs, sn, total,
                                                                                   ld, heads2, seq_len, inverse);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_swiglu_fwd(const void* gu, void* act, int rows, int inter, void* stream) {
  B200_REQUIRE(inter % 8 == 0, "swiglu: inter %% 8 != 0");
  const long long groups = (long long)rows * (inter / 8);
  swiglu_fwd_kernel<<<(unsigned)((groups + 255) / 256), 256, 0, STREAM(stream)>>>((const __nv_bfloat16*)gu,
                                                                                  (__nv_bfloat16*)act, groups, inter);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_swiglu_bwd(const void* dact, const void* gu, void* dgu, int rows, int inter, void* stream) {
  B200_REQUIRE(inter % 8 == 0, "swiglu: inter %% 8 != 0");
  const long long groups = (long long)rows * (inter / 8);
  swiglu_bwd_kernel<<<(unsigned)((groups + 255) / 256), 256, 0, STREAM(stream)>>>(
      (const __nv_bfloat16*)dact, (const __nv_bfloat16*)gu, (__nv_bfloat16*)dgu, groups, inter);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"
