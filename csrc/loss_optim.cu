// Fused softmax cross-entropy (forward + backward in one pass over a logits chunk held in shared
// memory), fused AdamW with bf16 shadow write, gradient-norm partials, clip coefficient, RNG bump.
#include <cuda_bf16.h>
#include <math.h>
#include <stdlib.h>

#include "../include/b200_trainer.h"
#include "common/host.h"
#include "common/ptx.cuh"

namespace b200 {

// ------------------------------------------------------------------------------------------------ cross entropy
// One CTA per row.  The row (<= ~57k bf16) is staged once in shared memory: global traffic is one read
// and one write of the chunk, which the chunked LM-head driver keeps L2-resident.
// Reference: F.cross_entropy(shift_logits, shift_labels) gpt.py:451-453 and its autograd backward.
constexpr int kCeThreads = 512;

__device__ __forceinline__ float ex2_approx_ftz(float x) {   // a single MUFU.EX2
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ float block_reduce(float v, float* scratch, bool is_max) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float t = __shfl_xor_sync(0xffffffffu, v, o);
    v = is_max ? fmaxf(v, t) : v + t;
  }
  __syncthreads();
  if (lane == 0) scratch[wid] = v;
  __syncthreads();
  if (wid == 0) {
    v = lane < (blockDim.x >> 5) ? scratch[lane] : (is_max ? -INFINITY : 0.f);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float t = __shfl_xor_sync(0xffffffffu, v, o);
      v = is_max ? fmaxf(v, t) : v + t;
    }
    if (lane == 0) scratch[0] = v;
  }
  __syncthreads();
  return scratch[0];
}

__global__ void __launch_bounds__(kCeThreads)
ce_fwd_bwd_kernel(__nv_bfloat16* __restrict__ logits, const int64_t* __restrict__ labels, float* __restrict__ loss_sum,
                  int vocab, int ld, float loss_scale, float grad_scale) {
  extern __shared__ __align__(16) uint8_t ce_smem[];
  uint4* srow = reinterpret_cast<uint4*>(ce_smem);
  __shared__ float scratch[32];
  const int row = blockIdx.x;
  uint4* grow = reinterpret_cast<uint4*>(logits + size_t(row) * ld);
  const int nvec = ld / 8;
  const long long label = labels[row];
  if (label < 0 || label >= vocab) {  // ignored position: zero gradient, no loss
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) grow[i] = make_uint4(0, 0, 0, 0);
    return;
  }
  // Pass 1: global -> smem (four independent 16-byte loads in flight per thread) + row max.  Padding
  // columns [vocab, ld) are replaced by -inf while staging, so the later passes carry no bounds checks:
  // they contribute exp2(-inf) = 0 to the sum and get a zero gradient.
  float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
  for (int base = threadIdx.x; base < nvec; base += 4 * kCeThreads) {
    uint4 q[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = base + u * kCeThreads;
      q[u] = i < nvec ? grow[i] : make_uint4(0xFF80FF80u, 0xFF80FF80u, 0xFF80FF80u, 0xFF80FF80u);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = base + u * kCeThreads;
      uint32_t w[4] = {q[u].x, q[u].y, q[u].z, q[u].w};
      if (i * 8 + 8 > vocab) {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int c = i * 8 + 2 * e;
          if (c >= vocab) w[e] = (w[e] & 0xFFFF0000u) | 0xFF80u;
          if (c + 1 >= vocab) w[e] = (w[e] & 0x0000FFFFu) | 0xFF800000u;
        }
      }
      if (i < nvec) srow[i] = make_uint4(w[0], w[1], w[2], w[3]);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 f = unpack_bf16x2(w[e]);
        mx4[e] = fmaxf(mx4[e], fmaxf(f.x, f.y));
      }
    }
  }
  float mx = block_reduce(fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])), scratch, true);
  constexpr float kLog2e = 1.4426950408889634f;
  const float neg_mxs = -mx * kLog2e;
  // Pass 2: exponentials (one FFMA + one MUFU per element), summed in fp32 and written back over the staged row as
  // bf16 -- the gradient pass then needs no second MUFU per element (2 x 4224 x 50k exps per chunk is ~100 us of
  // the SFU pipes, as much as the kernel's HBM time).  The label's logit is picked up on the way.
  __shared__ float s_xl;
  const int lab_vec = int(label >> 3), lab_e = int(label & 7);
  float s4[4] = {0.f, 0.f, 0.f, 0.f};
  for (int i = threadIdx.x; i < nvec; i += kCeThreads) {
    const uint4 q = srow[i];
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
    if (i == lab_vec) {
      const float2 f = unpack_bf16x2(w[lab_e >> 1]);
      s_xl = (lab_e & 1) ? f.y : f.x;
    }
    uint32_t o[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 f = unpack_bf16x2(w[e]);
      const float e0 = ex2_approx_ftz(fmaf(f.x, kLog2e, neg_mxs)), e1 = ex2_approx_ftz(fmaf(f.y, kLog2e, neg_mxs));
      s4[e] += e0 + e1;
      o[e] = pack_bf16x2(e0, e1);
    }
    srow[i] = make_uint4(o[0], o[1], o[2], o[3]);
  }
  const float sum = block_reduce((s4[0] + s4[1]) + (s4[2] + s4[3]), scratch, false);
  const float inv = grad_scale / sum;
  if (threadIdx.x == 0) atomicAdd(loss_sum, (logf(sum) + mx - s_xl) * loss_scale);
  // Pass 3: gradient (softmax - onehot) * grad_scale, written over the logits.
  for (int i = threadIdx.x; i < nvec; i += kCeThreads) {
    const uint4 q = srow[i];
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
    float g[8];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 f = unpack_bf16x2(w[e]);
      g[2 * e] = f.x * inv;
      g[2 * e + 1] = f.y * inv;
    }
    if (i == lab_vec) {
#pragma unroll
      for (int e = 0; e < 8; ++e)
        if (e == lab_e) g[e] -= grad_scale;
    }
    grow[i] = make_uint4(pack_bf16x2(g[0], g[1]), pack_bf16x2(g[2], g[3]), pack_bf16x2(g[4], g[5]), pack_bf16x2(g[6], g[7]));
  }
}

// Persistent, pipelined variant (used when two rows fit in shared memory: ld <= ~57k).  One CTA per SM walks rows
// blockIdx.x, + gridDim.x, ...; a producer warp fetches whole rows with 1-D bulk copies (global -> smem behind an
// mbarrier) into two row buffers, always one row ahead, while 31 consumer warps run the three passes on the row that
// has landed and write the gradient straight to global memory (coalesced 16-byte stores).  The one-CTA-per-row
// kernel above keeps HBM idle while a CTA computes (66 % of copy bandwidth at the benchmark shape); here the next
// row's load is in flight during every pass.  (Also storing by bulk copy from the row buffer was measured: the next
// load then waits for the store to drain the buffer, 172 us per chunk against 186 for the kernel above.)
constexpr int kCeConsumers = 992;   // 31 warps + the producer warp = 1024 threads
constexpr int kCePipeThreads = kCeConsumers + 32;

// reduction over the consumer threads only (named barrier 1); `scratch` alternates between two arrays so one barrier
// per reduction is enough
__device__ __forceinline__ float ce_consumer_reduce(float v, float* scratch, bool is_max) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float t = __shfl_xor_sync(0xffffffffu, v, o);
    v = is_max ? fmaxf(v, t) : v + t;
  }
  if (lane == 0) scratch[wid] = v;
  named_bar_sync(1, kCeConsumers);
  v = lane < kCeConsumers / 32 ? scratch[lane] : (is_max ? -INFINITY : 0.f);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float t = __shfl_xor_sync(0xffffffffu, v, o);
    v = is_max ? fmaxf(v, t) : v + t;
  }
  return v;
}

__global__ void __launch_bounds__(kCePipeThreads, 1)
ce_fwd_bwd_pipe_kernel(__nv_bfloat16* __restrict__ logits, const int64_t* __restrict__ labels, float* __restrict__ loss_sum,
                       int vocab, int ld, int rows, float loss_scale, float grad_scale) {
  extern __shared__ __align__(128) uint8_t ce_smem[];
  __shared__ __align__(8) uint64_t full[2], done[2];
  __shared__ float red_a[32], red_b[32];
  __shared__ float s_xl;
  const uint32_t row_bytes = uint32_t(ld) * 2u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 2; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&done[i], kCeConsumers / 32);
    }
    fence_barrier_init();
  }
  __syncthreads();
  const int mine = (rows - int(blockIdx.x) + int(gridDim.x) - 1) / int(gridDim.x);   // rows this CTA processes

  if (warp == kCeConsumers / 32) {
    // ------------------------------------------------------------------ producer warp: one thread fetches the rows
    if (lane == 0) {
      for (int k = 0; k < mine; ++k) {
        const int b = k & 1;
        const size_t row = size_t(blockIdx.x) + size_t(k) * gridDim.x;
        if (k >= 2) mbar_wait(&done[b], ((k - 2) >> 1) & 1);   // the consumers have finished with row k-2 in this buffer
        mbar_expect_tx(&full[b], row_bytes);
        // a few copies per row rather than one
        constexpr uint32_t kPiece = 16384;
        for (uint32_t off = 0; off < row_bytes; off += kPiece) {
          const uint32_t n = row_bytes - off < kPiece ? row_bytes - off : kPiece;
          bulk_load_g2s(smem_u32(ce_smem + size_t(b) * row_bytes) + off, reinterpret_cast<const uint8_t*>(logits + row * ld) + off, n, &full[b]);
        }
      }
    }
    return;
  }

  // -------------------------------------------------------------------- consumer warps
  // Instruction-issue bound before HBM (19 instructions per element with scalar fp32 math: 172 us per chunk), so the
  // passes use packed operations: bf16x2 max on the raw words, f32x2 FFMA / FADD around the one MUFU per element, and a
  // bf16x2 product for the gradient with 1/sum split into a bf16 head and tail (one rounding, like the scalar path).
  const int nvec = ld / 8;
  constexpr float kLog2e = 1.4426950408889634f;
  auto as_bf162 = [](uint32_t u) { return *reinterpret_cast<const __nv_bfloat162*>(&u); };
  auto as_u32 = [](__nv_bfloat162 v) { return *reinterpret_cast<const uint32_t*>(&v); };
  for (int k = 0; k < mine; ++k) {
    const int b = k & 1;
    const size_t row = size_t(blockIdx.x) + size_t(k) * gridDim.x;
    const long long label = labels[row];
    uint4* srow = reinterpret_cast<uint4*>(ce_smem + size_t(b) * row_bytes);
    const __nv_bfloat16* srow_e = reinterpret_cast<const __nv_bfloat16*>(srow);
    uint4* grow = reinterpret_cast<uint4*>(logits + row * ld);
    mbar_wait(&full[b], (k >> 1) & 1);
    if (label < 0 || label >= vocab) {   // ignored position: zero gradient, no loss
      for (int i = threadIdx.x; i < nvec; i += kCeConsumers) grow[i] = make_uint4(0, 0, 0, 0);
    } else {
      const int lab_vec = int(label >> 3);
      const bool lab_mine = lab_vec % kCeConsumers == int(threadIdx.x);   // this thread's vectors include the label's
      // Pass 1: row max; padding columns [vocab, ld) become -inf in place so the later passes carry no bounds checks
      __nv_bfloat162 m2[2] = {as_bf162(0xFF80FF80u), as_bf162(0xFF80FF80u)};
      for (int i = threadIdx.x; i < nvec; i += kCeConsumers) {
        uint4 q = srow[i];
        if (i * 8 + 8 > vocab) {
          uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int c = i * 8 + 2 * e;
            if (c >= vocab) w[e] = (w[e] & 0xFFFF0000u) | 0xFF80u;
            if (c + 1 >= vocab) w[e] = (w[e] & 0x0000FFFFu) | 0xFF800000u;
          }
          q = make_uint4(w[0], w[1], w[2], w[3]);
          srow[i] = q;
        }
        m2[0] = __hmax2(m2[0], as_bf162(q.x));
        m2[1] = __hmax2(m2[1], as_bf162(q.y));
        m2[0] = __hmax2(m2[0], as_bf162(q.z));
        m2[1] = __hmax2(m2[1], as_bf162(q.w));
      }
      if (lab_mine) s_xl = __bfloat162float(srow_e[label]);   // before pass 2 overwrites the logits with exponentials
      const float2 mf = unpack_bf16x2(as_u32(__hmax2(m2[0], m2[1])));
      const float mx = ce_consumer_reduce(fmaxf(mf.x, mf.y), red_a, true);
      const float neg_mxs = -mx * kLog2e;
      // Pass 2: exponentials (one FFMA2 per pair + one MUFU per element), summed in fp32, written back as bf16
      float2 s2[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
      const float2 l2 = make_float2(kLog2e, kLog2e), n2 = make_float2(neg_mxs, neg_mxs);
      for (int i = threadIdx.x; i < nvec; i += kCeConsumers) {
        const uint4 q = srow[i];
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
        uint32_t o[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 arg = ffma2(unpack_bf16x2(w[e]), l2, n2);
          const float2 ex = make_float2(ex2_approx_ftz(arg.x), ex2_approx_ftz(arg.y));
          s2[e & 1] = fadd2(s2[e & 1], ex);
          o[e] = pack_bf16x2(ex.x, ex.y);
        }
        srow[i] = make_uint4(o[0], o[1], o[2], o[3]);
      }
      const float sum = ce_consumer_reduce((s2[0].x + s2[0].y) + (s2[1].x + s2[1].y), red_b, false);   // (its barrier also publishes s_xl)
      const float inv = grad_scale / sum;
      if (threadIdx.x == 0) atomicAdd(loss_sum, (logf(sum) + mx - s_xl) * loss_scale);
      // Pass 3: gradient softmax * grad_scale written over the logits in global memory; the label's element (minus
      // grad_scale) is fixed up in fp32 by the thread that wrote its vector
      const __nv_bfloat16 ih = __float2bfloat16_rn(inv), il = __float2bfloat16_rn(inv - __bfloat162float(ih));
      const __nv_bfloat162 ih2 = __halves2bfloat162(ih, ih), il2 = __halves2bfloat162(il, il);
      for (int i = threadIdx.x; i < nvec; i += kCeConsumers) {
        const uint4 q = srow[i];
        grow[i] = make_uint4(as_u32(__hfma2(as_bf162(q.x), ih2, __hmul2(as_bf162(q.x), il2))),
                             as_u32(__hfma2(as_bf162(q.y), ih2, __hmul2(as_bf162(q.y), il2))),
                             as_u32(__hfma2(as_bf162(q.z), ih2, __hmul2(as_bf162(q.z), il2))),
                             as_u32(__hfma2(as_bf162(q.w), ih2, __hmul2(as_bf162(q.w), il2))));
      }
      if (lab_mine)
        reinterpret_cast<__nv_bfloat16*>(grow)[label] = __float2bfloat16_rn(__bfloat162float(srow_e[label]) * inv - grad_scale);
    }
    __syncwarp();   // every lane has read its part of the row: the buffer may take row k+2
    if (lane == 0) mbar_arrive(&done[b]);
  }
}

// ------------------------------------------------------------------------------------------------ AdamW
// torch.optim.AdamW math (decoupled decay; denom = sqrt(v)/sqrt(bc2) + eps), 4 elements per thread.
__global__ void __launch_bounds__(256)
adamw_kernel(float* __restrict__ master, const float* __restrict__ grad, float* __restrict__ m, float* __restrict__ v,
             __nv_bfloat16* __restrict__ shadow, const uint8_t* __restrict__ decay_chunks, long long lo, long long n4,
             float lr, float beta1, float beta2, float eps, float wd, float step_size, float inv_sqrt_bc2,
             const float* __restrict__ grad_scale) {
  const float gs = grad_scale ? *grad_scale : 1.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const long long gi = lo + 4 * i;  // global flat element index
    float4 p = *reinterpret_cast<const float4*>(master + gi);
    float4 g = *reinterpret_cast<const float4*>(grad + gi);
    float4 mm = reinterpret_cast<const float4*>(m)[i];
    float4 vv = reinterpret_cast<const float4*>(v)[i];
    const float decay = decay_chunks[gi >> 7] ? 1.f - lr * wd : 1.f;
    float* pp = reinterpret_cast<float*>(&p);
    float* gg = reinterpret_cast<float*>(&g);
    float* pm = reinterpret_cast<float*>(&mm);
    float* pv = reinterpret_cast<float*>(&vv);
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float ge = gg[e] * gs;
      pm[e] = beta1 * pm[e] + (1.f - beta1) * ge;
      pv[e] = beta2 * pv[e] + (1.f - beta2) * ge * ge;
      const float denom = sqrtf(pv[e]) * inv_sqrt_bc2 + eps;
      pp[e] = pp[e] * decay - step_size * (pm[e] / denom);
    }
    *reinterpret_cast<float4*>(master + gi) = p;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
    if (shadow) {
      uint2 s = make_uint2(pack_bf16x2(pp[0], pp[1]), pack_bf16x2(pp[2], pp[3]));
      *reinterpret_cast<uint2*>(shadow + gi) = s;
    }
  }
}

__global__ void __launch_bounds__(256) sumsq_kernel(const float* __restrict__ x, long long n4, float* __restrict__ out) {
  float acc = 0.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const float4 a = reinterpret_cast<const float4*>(x)[i];
    acc += a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  __shared__ float s[8];
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < 8; ++i) t += s[i];
    atomicAdd(out, t);
  }
}

__global__ void clip_coef_kernel(const float* sumsq, float max_norm, float pre_scale, float* scale, float* norm_out) {
  // sumsq is of UNSCALED gradients; the norm of the scaled gradient is pre_scale * sqrt(sumsq)
  const float norm = sqrtf(*sumsq) * fabsf(pre_scale);
  float c = 1.f;
  if (max_norm > 0.f) c = fminf(1.f, max_norm / (norm + 1e-6f));
  *scale = pre_scale * c;
  if (norm_out) *norm_out = norm;
}

__global__ void rng_advance_kernel(unsigned long long* st, unsigned long long delta) { st[1] += delta; }

}  // namespace b200

using namespace b200;
#define STREAM(s) reinterpret_cast<cudaStream_t>(s)

extern "C" {

int b200_ce_fwd_bwd(void* logits, const int64_t* labels, float* loss_sum, int rows, int vocab, int ld,
                    float loss_scale, float grad_scale, void* stream) {
  B200_REQUIRE(ld % 8 == 0 && ld >= vocab, "ce: ld must be a multiple of 8 and >= vocab");
  const size_t smem = size_t(ld) * 2;
  B200_REQUIRE(smem <= 200 * 1024, "ce: a row of %d logits does not fit in shared memory", ld);
  // two rows fit (and there are rows to pipeline): the persistent kernel with bulk-copy row movement
  static const int env_pipe = [] { const char* e = getenv("B200_CE_PIPE"); return e ? atoi(e) : 1; }();
  if (env_pipe && 2 * smem + 1024 <= 227 * 1024 && rows >= 2 * sm_count()) {
    static size_t configured_pipe[kMaxDevices] = {};
    if (2 * smem > configured_pipe[current_device()]) {
      B200_CHECK_CUDA(cudaFuncSetAttribute(ce_fwd_bwd_pipe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(2 * smem)));
      configured_pipe[current_device()] = 2 * smem;
    }
    ce_fwd_bwd_pipe_kernel<<<sm_count(), kCePipeThreads, 2 * smem, STREAM(stream)>>>((__nv_bfloat16*)logits, labels, loss_sum,
                                                                                      vocab, ld, rows, loss_scale, grad_scale);
    B200_CHECK_CUDA(cudaGetLastError());
    return 0;
  }
  static size_t configured[kMaxDevices] = {};   // the attribute is per device
  if (smem > configured[current_device()]) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(ce_fwd_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured[current_device()] = smem;
  }
  ce_fwd_bwd_kernel<<<rows, kCeThreads, smem, STREAM(stream)>>>((__nv_bfloat16*)logits, labels, loss_sum, vocab, ld,
                                                                loss_scale, grad_scale);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_adamw(float* master, const float* grad, float* exp_avg, float* exp_avg_sq, void* shadow,
               const uint8_t* decay_chunks, long long lo, long long hi, int step, float lr, float beta1, float beta2,
               float eps, float weight_decay, const float* grad_scale, void* stream) {
  B200_REQUIRE(lo % 4 == 0 && hi % 4 == 0 && hi >= lo, "adamw: range [%lld,%lld) must be 4-element aligned", lo, hi);
  if (hi == lo) return 0;
  const long long n4 = (hi - lo) / 4;
  const double bc1 = 1.0 - pow((double)beta1, (double)step);
  const double bc2 = 1.0 - pow((double)beta2, (double)step);
  const float step_size = (float)(lr / bc1);
  const float inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2));
  long long blocks = (n4 + 255) / 256;
  const long long cap = (long long)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  adamw_kernel<<<(unsigned)blocks, 256, 0, STREAM(stream)>>>(master, grad, exp_avg, exp_avg_sq, (__nv_bfloat16*)shadow,
                                                             decay_chunks, lo, n4, lr, beta1, beta2, eps, weight_decay,
                                                             step_size, inv_sqrt_bc2, grad_scale);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_sumsq(const float* x, long long lo, long long hi, float* out, void* stream) {
  B200_REQUIRE(lo % 4 == 0 && hi % 4 == 0 && hi >= lo, "sumsq: range must be 4-element aligned");
  if (hi == lo) return 0;
  const long long n4 = (hi - lo) / 4;
  long long blocks = (n4 + 255) / 256;
  const long long cap = (long long)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  sumsq_kernel<<<(unsigned)blocks, 256, 0, STREAM(stream)>>>(x + lo, n4, out);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_clip_coef(const float* sumsq, float max_norm, float pre_scale, float* scale, float* norm_out, void* stream) {
  clip_coef_kernel<<<1, 1, 0, STREAM(stream)>>>(sumsq, max_norm, pre_scale, scale, norm_out);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

int b200_rng_advance(unsigned long long* st, unsigned long long delta, void* stream) {
  rng_advance_kernel<<<1, 1, 0, STREAM(stream)>>>(st, delta);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"
