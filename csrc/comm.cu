// NVLink 5 / NVSwitch communication kernels over peer-mapped memory (CUDA IPC or symmetric memory):
// no NCCL on these paths.  Every rank passes a device table of N peer pointers per symmetric buffer.
//
//   b200_comm_barrier        signal-pad barrier: st.release.sys of an epoch into every peer's pad, then
//                            ld.acquire.sys spin on the own pad (bounded: a dead peer traps, never hangs)
//   b200_reduce_scatter_f32  this rank's slice of the gradient = sum over peers (P2P loads straight from
//                            the peers' flat fp32 gradient buffers), scaled, written locally, with the
//                            squared-norm partial folded in         (reference K4/K9 + G13)
//   b200_adamw_bcast         AdamW on the owned slice (clip coefficient from device memory) whose bf16
//                            result is stored into EVERY rank's shadow buffer: the all-gather half of the
//                            all-reduce carries updated parameters (2 B/param) instead of gradients
//                            (4 B/param) and the optimizer runs on 1/N of the data  (reference K4 + G14)
//   b200_allreduce_scalars   tiny all-reduce of a few floats through the signal pads (norm^2, loss)
//   b200_all_gather_bf16     FSDP parameter all-gather by P2P pull               (reference K7/K8)
//   b200_reduce_scatter_bf16out  FSDP gradient reduce-scatter                    (reference K9)
//
// Reference call sites being replaced: torch DDP's bucketed NCCL all-reduce (ddp_trainer.py:167-172,
// :342), FSDP1's all_gather_into_tensor / reduce_scatter_tensor (fsdp_trainer.py:379-383),
// clip_grad_norm_ (ddp_trainer.py:350, fsdp_trainer.py:388) and the optimizer steps (:356, :391).
#include <cuda_bf16.h>
#include <math.h>

#include "../include/b200_trainer.h"
#include "common/host.h"
#include "common/ptx.cuh"

namespace b200 {

constexpr int kMaxRanks = 16;
constexpr unsigned long long kCommTimeoutNs = 20000000000ull;  // 20 s

struct PeerTable {
  void* p[kMaxRanks];
};

__device__ __forceinline__ void st_release_sys_u32(uint32_t* addr, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys_u32(const uint32_t* addr) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer_f4(const float* p) {
  float4 v;
  asm volatile("ld.global.relaxed.sys.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ uint4 ld_peer_u4(const void* p) {
  uint4 v;
  asm volatile("ld.global.relaxed.sys.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

// pads.p[r] -> uint32 slots[kMaxRanks * kPadRows] on rank r.  Row `row` keeps barrier epochs separate
// from scalar-exchange traffic.
__device__ __forceinline__ void pad_barrier(const PeerTable& pads, int rank, int world, uint32_t epoch, int row) {
  const int t = threadIdx.x;
  if (t < world) {
    __threadfence_system();
    st_release_sys_u32(reinterpret_cast<uint32_t*>(pads.p[t]) + row * kMaxRanks + rank, epoch);
    const uint32_t* mine = reinterpret_cast<const uint32_t*>(pads.p[rank]) + row * kMaxRanks + t;
    const unsigned long long t0 = global_timer_ns();
    uint32_t spins = 0;
    // epochs are monotonic; signed distance handles wrap-around
    while (int32_t(ld_acquire_sys_u32(mine) - epoch) < 0) {
      if ((++spins & 0xFF) == 0 && global_timer_ns() - t0 > kCommTimeoutNs) {
        printf("b200: comm barrier timed out waiting for rank %d (rank %d, epoch %u)\n", t, rank, epoch);
        __trap();
      }
    }
  }
  __syncthreads();
}

// The epoch of a pad operation is "how many pad operations this rank has executed", kept in device memory and
// bumped by the kernel itself: nothing host-side is baked into the launch, so the kernels can be captured in
// a CUDA graph and replayed.  Every rank runs the same stream-ordered sequence of pad operations, so the
// counters agree.  (Invariant for callers: never have pad operations in flight on two streams at once.)
__device__ __forceinline__ uint32_t next_epoch(unsigned int* counter) {
  __shared__ uint32_t s_epoch;
  if (threadIdx.x == 0) s_epoch = atomicAdd(counter, 1u) + 1u;
  __syncthreads();
  return s_epoch;
}

__global__ void comm_barrier_kernel(PeerTable pads, int rank, int world, unsigned int* counter) {
  const uint32_t epoch = next_epoch(counter);
  pad_barrier(pads, rank, world, epoch, 0);
}

// dst[i] = scale * sum_r src_r[lo + i]  for i in the owned slice; sumsq += sum(dst^2).
// `srcs` are the peers' flat gradient buffers (same global indexing); dst is this rank's own buffer.
template <int kWorld>
__global__ void __launch_bounds__(512)
reduce_scatter_f32_kernel(PeerTable srcs, float* __restrict__ dst, long long lo, long long n4, float scale,
                          float* __restrict__ sumsq, int rank) {
  float acc = 0.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 v[kWorld];
#pragma unroll
    for (int r = 0; r < kWorld; ++r) {
      const int src = (rank + r) % kWorld;  // rotate so the peers are not all hammering rank 0 first
      v[r] = ld_peer_f4(reinterpret_cast<const float*>(srcs.p[src]) + lo + 4 * i);
    }
    float4 s = v[0];
#pragma unroll
    for (int r = 1; r < kWorld; ++r) {
      s.x += v[r].x; s.y += v[r].y; s.z += v[r].z; s.w += v[r].w;
    }
    s.x *= scale; s.y *= scale; s.z *= scale; s.w *= scale;
    *reinterpret_cast<float4*>(dst + lo + 4 * i) = s;
    acc += s.x * s.x + s.y * s.y + s.z * s.z + s.w * s.w;
  }
  if (sumsq != nullptr) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    __shared__ float sh[16];
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
      float t = 0.f;
      for (int i = 0; i < (blockDim.x >> 5); ++i) t += sh[i];
      atomicAdd(sumsq, t);
    }
  }
}

// Generic-world fallback (runtime loop)
__global__ void __launch_bounds__(512)
reduce_scatter_f32_generic(PeerTable srcs, float* __restrict__ dst, long long lo, long long n4, float scale,
                           float* __restrict__ sumsq, int rank, int world) {
  float acc = 0.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = 0; r < world; ++r) {
      const float4 v = ld_peer_f4(reinterpret_cast<const float*>(srcs.p[(rank + r) % world]) + lo + 4 * i);
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    s.x *= scale; s.y *= scale; s.z *= scale; s.w *= scale;
    *reinterpret_cast<float4*>(dst + lo + 4 * i) = s;
    acc += s.x * s.x + s.y * s.y + s.z * s.z + s.w * s.w;
  }
  if (sumsq != nullptr) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    __shared__ float sh[16];
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
      float t = 0.f;
      for (int i = 0; i < (blockDim.x >> 5); ++i) t += sh[i];
      atomicAdd(sumsq, t);
    }
  }
}

// AdamW on [lo, lo + 4*n4) of the flat space; bf16 result stored to every peer's shadow buffer.
__global__ void __launch_bounds__(256)
adamw_bcast_kernel(float* __restrict__ master, const float* __restrict__ grad, float* __restrict__ m,
                   float* __restrict__ v, PeerTable shadows, int world, int rank,
                   const uint8_t* __restrict__ decay_chunks, long long lo, long long n4, float lr, float beta1,
                   float beta2, float eps, float wd, float step_size, float inv_sqrt_bc2,
                   const float* __restrict__ grad_scale) {
  const float gs = grad_scale ? *grad_scale : 1.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const long long gi = lo + 4 * i;
    float4 p = *reinterpret_cast<const float4*>(master + gi);
    const float4 g = *reinterpret_cast<const float4*>(grad + gi);
    float4 mm = reinterpret_cast<const float4*>(m)[i];
    float4 vv = reinterpret_cast<const float4*>(v)[i];
    const float decay = decay_chunks[gi >> 7] ? 1.f - lr * wd : 1.f;
    float* pp = reinterpret_cast<float*>(&p);
    const float* gg = reinterpret_cast<const float*>(&g);
    float* pm = reinterpret_cast<float*>(&mm);
    float* pv = reinterpret_cast<float*>(&vv);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float ge = gg[e] * gs;
      pm[e] = beta1 * pm[e] + (1.f - beta1) * ge;
      pv[e] = beta2 * pv[e] + (1.f - beta2) * ge * ge;
      pp[e] = pp[e] * decay - step_size * (pm[e] / (sqrtf(pv[e]) * inv_sqrt_bc2 + eps));
    }
    *reinterpret_cast<float4*>(master + gi) = p;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
    const uint2 s = make_uint2(pack_bf16x2(pp[0], pp[1]), pack_bf16x2(pp[2], pp[3]));
    for (int r = 0; r < world; ++r) {
      __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(shadows.p[(rank + r) % world]) + gi;
      *reinterpret_cast<uint2*>(dst) = s;
    }
  }
}

// Each rank contributes `n` floats; result (sum over ranks, identical order on every rank) in out[].
// Exchange area: row 1.. of the signal pads, float slots [rank][n] (n <= 8).
__global__ void allreduce_scalars_kernel(PeerTable pads, PeerTable xchg, int rank, int world, unsigned int* counter,
                                         const float* __restrict__ in, float* __restrict__ out, int n) {
  const int t = threadIdx.x;
  const uint32_t epoch = next_epoch(counter);
  // xchg.p[r] -> float area [2][kMaxRanks][8] on rank r, double buffered by epoch parity so a fast
  // rank's next exchange cannot overwrite values a slow rank is still summing
  const int par = (epoch & 1) * kMaxRanks * 8;
  if (t < world) {
    float* dst = reinterpret_cast<float*>(xchg.p[t]) + par + rank * 8;
    for (int i = 0; i < n; ++i) dst[i] = in[i];
  }
  __syncthreads();
  pad_barrier(pads, rank, world, epoch, 0);
  if (t < n) {
    const float* mine = reinterpret_cast<const float*>(xchg.p[rank]) + par;
    float s = 0.f;
    for (int r = 0; r < world; ++r) s += *reinterpret_cast<const volatile float*>(mine + r * 8 + t);
    out[t] = s;
  }
}

// dst[r * shard + i] = src_r[i]   (pull every peer's shard into the local full buffer), 16 B units
__global__ void __launch_bounds__(512)
all_gather_pull_kernel(PeerTable srcs, uint4* __restrict__ dst, long long shard16, int rank, int world,
                       long long excl_lo16, long long excl_hi16) {
  // [excl_lo16, excl_hi16): 16-byte units of the gathered buffer that somebody else fills (the fused
  // all-gather GEMM gathers the rows of its own weight) and this kernel must leave alone
  const long long total = shard16 * world;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int rr = int(i / shard16);
    const int r = (rank + rr) % world;
    const long long off = i - rr * shard16;
    const long long pos = r * shard16 + off;
    if (pos >= excl_lo16 && pos < excl_hi16) continue;
    dst[pos] = ld_peer_u4(reinterpret_cast<const uint4*>(srcs.p[r]) + off);
  }
}

// FSDP gradient reduce-scatter: peers hold full-unit bf16 gradients; this rank sums its shard in fp32
// and ACCUMULATES into its fp32 sharded gradient (gradient accumulation across micro-steps).
__global__ void __launch_bounds__(512)
reduce_scatter_bf16_accum_kernel(PeerTable srcs, float* __restrict__ dst, long long src_off8, long long n8, float scale,
                                 int rank, int world) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int r = 0; r < world; ++r) {
      const uint4 q = ld_peer_u4(reinterpret_cast<const uint4*>(srcs.p[(rank + r) % world]) + src_off8 + i);
      const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 f = unpack_bf16x2(w[e]);
        acc[2 * e] += f.x;
        acc[2 * e + 1] += f.y;
      }
    }
    float4* d = reinterpret_cast<float4*>(dst + 8 * i);
    float4 a = d[0], b = d[1];
    a.x += acc[0] * scale; a.y += acc[1] * scale; a.z += acc[2] * scale; a.w += acc[3] * scale;
    b.x += acc[4] * scale; b.y += acc[5] * scale; b.z += acc[6] * scale; b.w += acc[7] * scale;
    d[0] = a;
    d[1] = b;
  }
}

// FSDP gradient reduce-scatter, fp32 wire: peers hold full-unit fp32 gradient staging buffers; this rank
// sums elements [src_off, src_off + 4*n4) of every peer and ACCUMULATES into its sharded gradient.
__global__ void __launch_bounds__(512)
reduce_scatter_f32_accum_kernel(PeerTable srcs, float* __restrict__ dst, long long src_off, long long n4, float scale,
                                int rank, int world) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
    for (int r = 0; r < world; ++r) {
      const float4 v = ld_peer_f4(reinterpret_cast<const float*>(srcs.p[(rank + r) % world]) + src_off + 4 * i);
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    float4 d = reinterpret_cast<float4*>(dst)[i];
    d.x += s.x * scale; d.y += s.y * scale; d.z += s.z * scale; d.w += s.w * scale;
    reinterpret_cast<float4*>(dst)[i] = d;
  }
}


// ------------------------------------------------------------------------------------------------------------
// Chunk-table kernels: the optimizer side of the fused reduce-scatter.  A rank's owned parameters are a list of
// short contiguous runs (128-row slabs of the weight matrices, whole norm vectors); the host flattens them into
// chunks of <= 4096 elements so that ONE launch covers all of them (one launch per run would be ~100 tiny
// kernels per step).  Offsets are in elements.
//   g   global flat index of the chunk's first element (weight-decay table lookup)
//   pg  offset into the local fp32 master / gradient arrays
//   mv  offset into the local AdamW moment arrays
//   sh  offset into the destination bf16 parameter buffer(s)
//   dst rank whose bf16 buffer receives the updated parameters, or -1 = every rank (DDP replicas)
// ------------------------------------------------------------------------------------------------------------
struct Chunk {
  long long g, pg, mv, sh;
  int n, dst;
};
static_assert(sizeof(Chunk) == sizeof(b200_chunk), "b200_chunk layout");

__device__ __forceinline__ float block_sum_256(float acc) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  __shared__ float sh[8];
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  float t = 0.f;
  if (threadIdx.x == 0)
    for (int i = 0; i < 8; ++i) t += sh[i];
  return t;
}

// out[0] += sq_scale * sum over the chunks of grad^2
__global__ void __launch_bounds__(256)
sumsq_chunks_kernel(const float* __restrict__ grad, const Chunk* __restrict__ ch, int nch, float* __restrict__ out,
                    float sq_scale) {
  float acc = 0.f;
  for (int c = blockIdx.x; c < nch; c += gridDim.x) {
    const Chunk k = ch[c];
    const float4* src = reinterpret_cast<const float4*>(grad + k.pg);
    for (int i = threadIdx.x; i < (k.n >> 2); i += 256) {
      const float4 a = src[i];
      acc += a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w;
    }
  }
  const float t = block_sum_256(acc);
  if (threadIdx.x == 0) atomicAdd(out, t * sq_scale);
}

// AdamW over the chunks; the bf16 result goes to rank `dst`'s parameter buffer (or to every rank's), and the
// gradient is cleared in the same pass (zero_grad) so the next step's reduce-adds start from zero.
// one store, replicated by the NVSwitch into every rank's copy of the buffer (NVLS multicast mapping)
__device__ __forceinline__ void multimem_st_bf16x4(void* mc_addr, uint2 v) {
  asm volatile("multimem.st.relaxed.sys.global.v2.bf16x2 [%0], {%1, %2};" ::"l"(mc_addr), "r"(v.x), "r"(v.y) : "memory");
}

__global__ void __launch_bounds__(256)
adamw_chunks_kernel(float* __restrict__ master, float* __restrict__ grad, float* __restrict__ m, float* __restrict__ v,
                    PeerTable shadows, __nv_bfloat16* __restrict__ mc_base, int world, int rank,
                    const uint8_t* __restrict__ decay_chunks,
                    const Chunk* __restrict__ ch, int nch, float lr, float beta1, float beta2, float eps, float wd,
                    float step_size, float inv_sqrt_bc2, const float* __restrict__ grad_scale, int zero_grad) {
  const float gs = grad_scale ? *grad_scale : 1.f;
  for (int c = blockIdx.x; c < nch; c += gridDim.x) {
    const Chunk k = ch[c];
    for (int i = threadIdx.x; i < (k.n >> 2); i += 256) {
      const long long e = 4LL * i;
      float4 p = *reinterpret_cast<const float4*>(master + k.pg + e);
      const float4 g = *reinterpret_cast<const float4*>(grad + k.pg + e);
      float4 mm = *reinterpret_cast<const float4*>(m + k.mv + e);
      float4 vv = *reinterpret_cast<const float4*>(v + k.mv + e);
      const float decay = decay_chunks[(k.g + e) >> 7] ? 1.f - lr * wd : 1.f;
      float* pp = reinterpret_cast<float*>(&p);
      const float* gg = reinterpret_cast<const float*>(&g);
      float* pm = reinterpret_cast<float*>(&mm);
      float* pv = reinterpret_cast<float*>(&vv);
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float ge = gg[q] * gs;
        pm[q] = beta1 * pm[q] + (1.f - beta1) * ge;
        pv[q] = beta2 * pv[q] + (1.f - beta2) * ge * ge;
        pp[q] = pp[q] * decay - step_size * (pm[q] / (sqrtf(pv[q]) * inv_sqrt_bc2 + eps));
      }
      *reinterpret_cast<float4*>(master + k.pg + e) = p;
      *reinterpret_cast<float4*>(m + k.mv + e) = mm;
      *reinterpret_cast<float4*>(v + k.mv + e) = vv;
      if (zero_grad) *reinterpret_cast<float4*>(grad + k.pg + e) = make_float4(0.f, 0.f, 0.f, 0.f);
      const uint2 s = make_uint2(pack_bf16x2(pp[0], pp[1]), pack_bf16x2(pp[2], pp[3]));
      if (k.dst >= 0) {
        *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(shadows.p[k.dst]) + k.sh + e) = s;
      } else if (mc_base != nullptr) {
        multimem_st_bf16x4(mc_base + k.sh + e, s);   // NVLS: 2 B/param leave this GPU once, not world - 1 times
      } else {
        for (int r = 0; r < world; ++r)
          *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(shadows.p[(rank + r) % world]) + k.sh + e) = s;
      }
    }
  }
}

// bf16 copy of the chunks' fp32 master values into the parameter buffer(s) of their holders, gradient cleared:
// the device half of a step whose AdamW ran somewhere else (--cpu_offload: on the host).
__global__ void __launch_bounds__(256)
cast_scatter_chunks_kernel(const float* __restrict__ master, float* __restrict__ grad, PeerTable shadows, int world,
                           int rank, const Chunk* __restrict__ ch, int nch) {
  for (int c = blockIdx.x; c < nch; c += gridDim.x) {
    const Chunk k = ch[c];
    for (int i = threadIdx.x; i < (k.n >> 2); i += 256) {
      const long long e = 4LL * i;
      const float4 p = *reinterpret_cast<const float4*>(master + k.pg + e);
      if (grad != nullptr) *reinterpret_cast<float4*>(grad + k.pg + e) = make_float4(0.f, 0.f, 0.f, 0.f);
      const uint2 s = make_uint2(pack_bf16x2(p.x, p.y), pack_bf16x2(p.z, p.w));
      if (k.dst >= 0) {
        *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(shadows.p[k.dst]) + k.sh + e) = s;
      } else {
        for (int r = 0; r < world; ++r)
          *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(shadows.p[(rank + r) % world]) + k.sh + e) = s;
      }
    }
  }
}

// Small tensors (norm weights) accumulate in every rank's own buffer; their owner pulls the peers' copies:
// dst[pg ..] = sum_r src_r[g ..], and every rank's copy is cleared for the next step (zero_remote).
__device__ __forceinline__ float4 multimem_ld_reduce_add_f32x4(const float* mc_addr) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc_addr) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_f32x4(float* mc_addr, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w) : "memory");
}

// mc_src != NULL (NVLS multicast address of the same symmetric buffer): the sum over ranks is ONE
// multimem.ld_reduce -- the NVSwitch adds the ranks' copies and returns the result -- and one multimem.st of zeros
// clears every copy, instead of world loads and world stores per element.
__global__ void __launch_bounds__(256)
pull_reduce_chunks_kernel(PeerTable srcs, float* __restrict__ mc_src, float* __restrict__ dst,
                          const Chunk* __restrict__ ch, int nch, int rank, int world, int zero_remote) {
  for (int c = blockIdx.x; c < nch; c += gridDim.x) {
    const Chunk k = ch[c];
    for (int i = threadIdx.x; i < (k.n >> 2); i += 256) {
      const long long e = 4LL * i;
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
      if (mc_src != nullptr) {
        s = multimem_ld_reduce_add_f32x4(mc_src + k.g + e);
        if (zero_remote) {
          __threadfence_system();   // the reduction has read every copy before the zeros go out (tiny tensors: cheap)
          multimem_st_f32x4(mc_src + k.g + e, make_float4(0.f, 0.f, 0.f, 0.f));
        }
      } else {
        for (int r = 0; r < world; ++r) {
          float* src = reinterpret_cast<float*>(srcs.p[(rank + r) % world]) + k.g + e;
          const float4 a = ld_peer_f4(src);
          s.x += a.x; s.y += a.y; s.z += a.z; s.w += a.w;
          if (zero_remote) *reinterpret_cast<float4*>(src) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
      *reinterpret_cast<float4*>(dst + k.pg + e) = s;   // last: dst may be this rank's own source
    }
  }
}

// Embedding-gradient scatter with owner-directed rows: token row t adds dx[t] into row ids[t] of the embedding
// gradient held by the rank that owns that row's 128-row slab (red.global.add over NVLink for remote owners).
__global__ void embed_bwd_rs_kernel(const int64_t* __restrict__ ids, const float* __restrict__ dx, PeerTable owner_base,
                                    int rows, int hidden, int vocab, int world, int slabs, int rot) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const long long id = ids[row];
  if (id < 0 || id >= vocab) return;
  const int s = int(id >> 7);
  const int op = int(((long long)s * world) / slabs);
  int o = op + rot;
  if (o >= world) o -= world;
  const int first = (op * slabs + world - 1) / world;
  const float4* src = reinterpret_cast<const float4*>(dx + size_t(row) * hidden);
  float* dstp = reinterpret_cast<float*>(owner_base.p[o]) + (size_t(id) - (size_t(first) << 7)) * hidden;
  for (int i = threadIdx.x & 31; i < hidden / 4; i += 32) {
    const float4 a = src[i];
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dstp + 4 * i), "f"(a.x), "f"(a.y), "f"(a.z), "f"(a.w)
                 : "memory");
  }
}

static int fill_table(PeerTable* t, const void* const* host_ptrs, int world) {
  B200_REQUIRE(world >= 1 && world <= kMaxRanks, "comm: world size %d out of range (max %d)", world, kMaxRanks);
  for (int i = 0; i < kMaxRanks; ++i) t->p[i] = i < world ? const_cast<void*>(host_ptrs[i]) : nullptr;
  return 0;
}

}  // namespace b200

using namespace b200;
#define STREAM(s) reinterpret_cast<cudaStream_t>(s)

extern "C" {

int b200_comm_pad_bytes(void) { return kMaxRanks * 4 * 4 + 2 * kMaxRanks * 8 * 4; }

int b200_comm_barrier(const void* const* pad_ptrs, int rank, int world, unsigned int* epoch_counter, void* stream) {
  PeerTable pads;
  if (int rc = fill_table(&pads, pad_ptrs, world)) return rc;
  B200_REQUIRE(epoch_counter != nullptr, "comm_barrier: epoch counter (device uint32) required");
  comm_barrier_kernel<<<1, 32, 0, STREAM(stream)>>>(pads, rank, world, epoch_counter);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_reduce_scatter_f32(const void* const* grad_ptrs, float* dst, long long lo, long long hi, float scale,
                            float* sumsq, int rank, int world, int max_ctas, void* stream) {
  B200_REQUIRE(lo % 4 == 0 && hi % 4 == 0 && hi >= lo, "reduce_scatter: range must be 4-element aligned");
  if (hi == lo) return 0;
  PeerTable srcs;
  if (int rc = fill_table(&srcs, grad_ptrs, world)) return rc;
  const long long n4 = (hi - lo) / 4;
  long long blocks = (n4 + 511) / 512;
  const long long cap = max_ctas > 0 ? max_ctas : (long long)sm_count() * 2;
  if (blocks > cap) blocks = cap;
  cudaStream_t s = STREAM(stream);
  switch (world) {
    case 2: reduce_scatter_f32_kernel<2><<<(unsigned)blocks, 512, 0, s>>>(srcs, dst, lo, n4, scale, sumsq, rank); break;
    case 4: reduce_scatter_f32_kernel<4><<<(unsigned)blocks, 512, 0, s>>>(srcs, dst, lo, n4, scale, sumsq, rank); break;
    case 8: reduce_scatter_f32_kernel<8><<<(unsigned)blocks, 512, 0, s>>>(srcs, dst, lo, n4, scale, sumsq, rank); break;
    default: reduce_scatter_f32_generic<<<(unsigned)blocks, 512, 0, s>>>(srcs, dst, lo, n4, scale, sumsq, rank, world);
  }
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_adamw_bcast(float* master, const float* grad, float* exp_avg, float* exp_avg_sq,
                     const void* const* shadow_ptrs, int rank, int world, const uint8_t* decay_chunks, long long lo,
                     long long hi, int step, float lr, float beta1, float beta2, float eps, float weight_decay,
                     const float* grad_scale, int max_ctas, void* stream) {
  B200_REQUIRE(lo % 4 == 0 && hi % 4 == 0 && hi >= lo, "adamw_bcast: range must be 4-element aligned");
  if (hi == lo) return 0;
  PeerTable sh;
  if (int rc = fill_table(&sh, shadow_ptrs, world)) return rc;
  const long long n4 = (hi - lo) / 4;
  const double bc1 = 1.0 - pow((double)beta1, (double)step);
  const double bc2 = 1.0 - pow((double)beta2, (double)step);
  long long blocks = (n4 + 255) / 256;
  const long long cap = max_ctas > 0 ? max_ctas : (long long)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  adamw_bcast_kernel<<<(unsigned)blocks, 256, 0, STREAM(stream)>>>(
      master, grad, exp_avg, exp_avg_sq, sh, world, rank, decay_chunks, lo, n4, lr, beta1, beta2, eps, weight_decay,
      (float)(lr / bc1), (float)(1.0 / sqrt(bc2)), grad_scale);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_allreduce_scalars(const void* const* pad_ptrs, int rank, int world, unsigned int* epoch_counter,
                           const float* in, float* out, int n, void* stream) {
  B200_REQUIRE(n >= 1 && n <= 8, "allreduce_scalars: n must be in [1,8]");
  B200_REQUIRE(epoch_counter != nullptr, "allreduce_scalars: epoch counter (device uint32) required");
  PeerTable pads, xchg;
  if (int rc = fill_table(&pads, pad_ptrs, world)) return rc;
  for (int i = 0; i < kMaxRanks; ++i)
    xchg.p[i] = pads.p[i] ? reinterpret_cast<char*>(pads.p[i]) + kMaxRanks * 4 * 4 : nullptr;
  allreduce_scalars_kernel<<<1, 32, 0, STREAM(stream)>>>(pads, xchg, rank, world, epoch_counter, in, out, n);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_all_gather_bf16(const void* const* shard_ptrs, void* dst, long long shard_elems, int rank, int world,
                         int max_ctas, void* stream) {
  return b200_all_gather_bf16_excl(shard_ptrs, dst, shard_elems, rank, world, 0, 0, max_ctas, stream);
}

int b200_all_gather_bf16_excl(const void* const* shard_ptrs, void* dst, long long shard_elems, int rank, int world,
                              long long excl_lo_elems, long long excl_hi_elems, int max_ctas, void* stream) {
  B200_REQUIRE(shard_elems % 8 == 0, "all_gather: shard must be a multiple of 8 bf16 elements");
  B200_REQUIRE(excl_lo_elems % 8 == 0 && excl_hi_elems % 8 == 0 && excl_hi_elems >= excl_lo_elems,
               "all_gather: excluded window must be 8-element aligned");
  PeerTable srcs;
  if (int rc = fill_table(&srcs, shard_ptrs, world)) return rc;
  const long long shard16 = shard_elems / 8;
  long long blocks = (shard16 * world + 511) / 512;
  const long long cap = max_ctas > 0 ? max_ctas : (long long)sm_count() * 2;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  all_gather_pull_kernel<<<(unsigned)blocks, 512, 0, STREAM(stream)>>>(srcs, (uint4*)dst, shard16, rank, world,
                                                                      excl_lo_elems / 8, excl_hi_elems / 8);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_reduce_scatter_f32_accum(const void* const* full_grad_ptrs, float* dst_shard, long long src_offset,
                                  long long n, float scale, int rank, int world, int max_ctas, void* stream) {
  B200_REQUIRE(src_offset % 4 == 0 && n % 4 == 0, "reduce_scatter_f32_accum: offset/size must be multiples of 4");
  if (n == 0) return 0;
  PeerTable srcs;
  if (int rc = fill_table(&srcs, full_grad_ptrs, world)) return rc;
  long long blocks = (n / 4 + 511) / 512;
  const long long cap = max_ctas > 0 ? max_ctas : (long long)sm_count() * 2;
  if (blocks > cap) blocks = cap;
  reduce_scatter_f32_accum_kernel<<<(unsigned)blocks, 512, 0, STREAM(stream)>>>(srcs, dst_shard, src_offset, n / 4,
                                                                                 scale, rank, world);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_reduce_scatter_bf16_accum(const void* const* full_grad_ptrs, float* dst_shard, long long src_offset,
                                   long long n, float scale, int rank, int world, int max_ctas, void* stream) {
  B200_REQUIRE(src_offset % 8 == 0 && n % 8 == 0, "reduce_scatter_bf16: offset/size must be multiples of 8");
  if (n == 0) return 0;
  PeerTable srcs;
  if (int rc = fill_table(&srcs, full_grad_ptrs, world)) return rc;
  long long blocks = (n / 8 + 511) / 512;
  const long long cap = max_ctas > 0 ? max_ctas : (long long)sm_count() * 2;
  if (blocks > cap) blocks = cap;
  reduce_scatter_bf16_accum_kernel<<<(unsigned)blocks, 512, 0, STREAM(stream)>>>(srcs, dst_shard, src_offset / 8, n / 8,
                                                                                  scale, rank, world);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_sumsq_chunks(const float* grad, const b200_chunk* chunks_dev, int n_chunks, float* out, float sq_scale,
                      void* stream) {
  if (n_chunks <= 0) return 0;
  int blocks = n_chunks < sm_count() * 8 ? n_chunks : sm_count() * 8;
  sumsq_chunks_kernel<<<blocks, 256, 0, STREAM(stream)>>>(grad, reinterpret_cast<const Chunk*>(chunks_dev), n_chunks, out,
                                                         sq_scale);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_adamw_chunks(float* master, float* grad, float* exp_avg, float* exp_avg_sq, const void* const* param_ptrs,
                      void* param_multicast, int rank, int world, const uint8_t* decay_chunks,
                      const b200_chunk* chunks_dev, int n_chunks,
                      int step, float lr, float beta1, float beta2, float eps, float weight_decay,
                      const float* grad_scale, int zero_grad, int max_ctas, void* stream) {
  if (n_chunks <= 0) return 0;
  PeerTable sh;
  if (int rc = fill_table(&sh, param_ptrs, world)) return rc;
  const double bc1 = 1.0 - pow((double)beta1, (double)step);
  const double bc2 = 1.0 - pow((double)beta2, (double)step);
  int cap = max_ctas > 0 ? max_ctas : sm_count() * 8;
  int blocks = n_chunks < cap ? n_chunks : cap;
  adamw_chunks_kernel<<<blocks, 256, 0, STREAM(stream)>>>(master, grad, exp_avg, exp_avg_sq, sh,
                                                         reinterpret_cast<__nv_bfloat16*>(param_multicast), world, rank,
                                                         decay_chunks,
                                                         reinterpret_cast<const Chunk*>(chunks_dev), n_chunks, lr, beta1,
                                                         beta2, eps, weight_decay, (float)(lr / bc1),
                                                         (float)(1.0 / sqrt(bc2)), grad_scale, zero_grad);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_cast_scatter_chunks(const float* master, float* grad_to_zero, const void* const* param_ptrs, int rank, int world,
                              const b200_chunk* chunks_dev, int n_chunks, void* stream) {
  if (n_chunks <= 0) return 0;
  PeerTable sh;
  if (int rc = fill_table(&sh, param_ptrs, world)) return rc;
  int blocks = n_chunks < sm_count() * 8 ? n_chunks : sm_count() * 8;
  cast_scatter_chunks_kernel<<<blocks, 256, 0, STREAM(stream)>>>(master, grad_to_zero, sh, world, rank,
                                                                reinterpret_cast<const Chunk*>(chunks_dev), n_chunks);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_pull_reduce_chunks(const void* const* src_ptrs, void* src_multicast, float* dst, const b200_chunk* chunks_dev,
                            int n_chunks, int rank, int world, int zero_remote, void* stream) {
  if (n_chunks <= 0) return 0;
  PeerTable srcs;
  if (int rc = fill_table(&srcs, src_ptrs, world)) return rc;
  int blocks = n_chunks < sm_count() * 2 ? n_chunks : sm_count() * 2;
  pull_reduce_chunks_kernel<<<blocks, 256, 0, STREAM(stream)>>>(srcs, reinterpret_cast<float*>(src_multicast), dst,
                                                               reinterpret_cast<const Chunk*>(chunks_dev),
                                                               n_chunks, rank, world, zero_remote);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_embed_bwd_rs(const int64_t* ids, const float* dx, const void* const* owner_base, int rows, int hidden,
                      int vocab, int world, int rot, void* stream) {
  B200_REQUIRE(hidden % 4 == 0, "embed_bwd_rs: hidden must be a multiple of 4");
  PeerTable ob;
  if (int rc = fill_table(&ob, owner_base, world)) return rc;
  const int slabs = (vocab + 127) / 128;
  const int r = ((rot % world) + world) % world;
  embed_bwd_rs_kernel<<<(rows + 7) / 8, 256, 0, STREAM(stream)>>>(ids, dx, ob, rows, hidden, vocab, world, slabs, r);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"
