// Causal flash attention forward, small-footprint variant: same math, layouts and Philox stream as
// attn_fwd_kernel (attention.cu), but sized so that TWO CTAs are resident per SM -- 97 KB of shared memory,
// 256 TMEM columns, 320 threads at <= 102 registers.  The profile of the one-CTA-per-SM kernel showed all of
// its softmax warps marching through the same phase (TMEM load -> max -> barrier -> exp -> smem store) in
// lock-step, so the tensor pipe, the MUFU and the issue slots were never busy at the same time; two independent
// CTAs per SM drift out of phase and fill each other's bubbles.
//
//   warp 0     TMA producer: Q once, K through a 2-stage ring, V single-buffered
//   warp 1     MMA issuer:   S_j = Q K_j^T (TMEM cols 0..127), O_j = P_j V_j (TMEM cols 128..191)
//   warps 2-9  softmax: two threads per query row, 64 score columns each
//
// Reference: gpt.py:199-234 (manual attention / SDPA).  Entry point: b200_attn_fwd2 (same arguments as
// b200_attn_fwd; outputs are interchangeable, the backward kernel consumes either).
#include <cuda_bf16.h>
#include <math.h>

#include "../include/b200_trainer.h"
#include "common/host.h"
#include "common/philox.cuh"
#include "common/ptx.cuh"

namespace b200 {
namespace f2 {

constexpr int kBlk = 128;
constexpr int kD = 64;
constexpr int kTile = kBlk * kD * 2;   // 16 KB
constexpr int kPTile = 2 * kTile;      // 32 KB
constexpr float kScale2 = 0.125f * 1.4426950408889634f;
constexpr int kSmWarps = 8;
constexpr int kThreads = 64 + kSmWarps * 32;

struct Params {
  int B, S, H, S8;
  float p_drop;
  const unsigned long long* rng;
  unsigned int rng_stream;
  __nv_bfloat16* out;
  float* lse;
};

struct Smem {
  static constexpr int kQ = 0;
  static constexpr int kK = kQ + kTile;          // 2 stages
  static constexpr int kV = kK + 2 * kTile;      // 1 stage
  static constexpr int kP = kV + kTile;          // 1 buffer
  static constexpr int kBars = kP + kPTile;
  static constexpr int kXchg = kBars + 256;      // float [2][2][128]
  static constexpr int kTotal = kXchg + 2 * 2 * 128 * 4 + 1024;
};
static_assert(2 * (Smem::kTotal + 1024) <= 228 * 1024, "two CTAs per SM must fit in shared memory");

__device__ __forceinline__ uint32_t swz(uint32_t tile_base, int row, int cc) {
  return tile_base + row * 128 + ((cc ^ (row & 7)) << 4);
}

__global__ void __launch_bounds__(kThreads, 2)
attn_fwd2_kernel(const __grid_constant__ CUtensorMap tm_qkv, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Smem::kBars);
  uint64_t* q_full = bars;          // 1
  uint64_t* k_full = bars + 1;      // 2
  uint64_t* k_empty = bars + 3;     // 2
  uint64_t* v_full = bars + 5;      // 1
  uint64_t* v_empty = bars + 6;     // 1
  uint64_t* s_full = bars + 7;      // 1
  uint64_t* s_free = bars + 8;      // 1 (count 8)
  uint64_t* p_ready = bars + 9;     // 1 (count 8)
  uint64_t* o_full = bars + 10;     // 1
  uint64_t* o_free = bars + 11;     // 1 (count 8)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 12);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = (p.S + kBlk - 1) / kBlk;
  // 1-D grid, longest rows first ACROSS the whole grid (blocks are issued in index order): every
  // (batch, head) pair's 8-block tile starts before any 7-block tile, ... so the tail of the kernel is
  // made of the one-block tiles instead of whichever heavy tiles happened to come last.
  const int bh_count = p.H * p.B;
  const int qi = nq - 1 - int(blockIdx.x) / bh_count;
  const int hd = (int(blockIdx.x) % bh_count) % p.H, b = (int(blockIdx.x) % bh_count) / p.H;
  const int h = p.H * kD;
  const int nkv = qi + 1;
  const int row_base = b * p.S;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tm_qkv);
    mbar_init(q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], 1);
    }
    mbar_init(v_full, 1);
    mbar_init(v_empty, 1);
    mbar_init(s_full, 1);
    mbar_init(s_free, kSmWarps);
    mbar_init(p_ready, kSmWarps);
    mbar_init(o_full, 1);
    mbar_init(o_free, kSmWarps);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<256>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_ptr;
  const uint32_t tS = tmem, tO = tmem + 128;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(q_full, kTile);
      tma_load_2d(smem + Smem::kQ, &tm_qkv, q_full, hd * kD, row_base + qi * kBlk);
      for (int j = 0; j < nkv; ++j) {
        const int st = j & 1;
        mbar_wait(&k_empty[st], ((j >> 1) & 1) ^ 1);
        mbar_expect_tx(&k_full[st], kTile);
        tma_load_2d(smem + Smem::kK + st * kTile, &tm_qkv, &k_full[st], h + hd * kD, row_base + j * kBlk);
        mbar_wait(v_empty, (j & 1) ^ 1);
        mbar_expect_tx(v_full, kTile);
        tma_load_2d(smem + Smem::kV, &tm_qkv, v_full, 2 * h + hd * kD, row_base + j * kBlk);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc_bf16(128, 128, false, false);
      constexpr uint32_t idesc_o = make_idesc_bf16(128, 64, false, true);
      const uint32_t q_addr = smem_u32(smem + Smem::kQ);
      const uint32_t p_addr = smem_u32(smem + Smem::kP);
      const uint32_t v_addr = smem_u32(smem + Smem::kV);
      mbar_wait(q_full, 0);
      auto issue_s = [&](int j) {
        const int st = j & 1;
        mbar_wait(&k_full[st], (j >> 1) & 1);
        mbar_wait(s_free, (j & 1) ^ 1);
        tc_fence_after();
        const uint32_t k_addr = smem_u32(smem + Smem::kK + st * kTile);
#pragma unroll
        for (int t = 0; t < kD / 16; ++t)
          umma_bf16_ss(tS, make_smem_desc_sw128(q_addr + t * 32, 16, 1024), make_smem_desc_sw128(k_addr + t * 32, 16, 1024),
                       idesc_s, t > 0);
        tc_commit(&k_empty[st]);
        tc_commit(s_full);
      };
      issue_s(0);
      for (int j = 0; j < nkv; ++j) {
        mbar_wait(p_ready, j & 1);                 // softmax(j) has written P_j (and is done reading S_j)
        if (j + 1 < nkv) issue_s(j + 1);           // next scores first: the softmax warps need them soonest
        mbar_wait(v_full, j & 1);
        mbar_wait(o_free, (j & 1) ^ 1);
        tc_fence_after();
#pragma unroll
        for (int t = 0; t < kBlk / 16; ++t)
          umma_bf16_ss(tO, make_smem_desc_sw128(p_addr + (t >> 2) * kTile + (t & 3) * 32, 16, 1024),
                       make_smem_desc_sw128(v_addr + t * 2048, kTile, 1024), idesc_o, t > 0);
        tc_commit(v_empty);
        tc_commit(o_full);
      }
    }
  } else {
    const int qd = warp & 3;
    const int hf = (warp - 2) >> 2;
    const int r = qd * 32 + lane;
    const int qrow = qi * kBlk + r;
    const uint32_t lane_off = uint32_t(qd * 32) << 16;
    const bool drop = p.p_drop > 0.f;
    const uint32_t thresh = dropout_threshold16(p.p_drop);
    unsigned long long seed = 0, offs = 0;
    if (drop) { seed = p.rng[0]; offs = p.rng[1]; }
    const unsigned long long grp_row = ((unsigned long long)(b * p.H + hd) * p.S + qrow) * p.S8;
    float* xchg = reinterpret_cast<float*>(smem + Smem::kXchg);
    const uint32_t p_sub = smem_u32(smem + Smem::kP) + hf * kTile;

    float m = -INFINITY, l = 0.f, alpha_prev = 1.f;
    float o[32];
#pragma unroll
    for (int e = 0; e < 32; ++e) o[e] = 0.f;

    auto fold_o = [&](int jdone, float alpha) {    // o = o * alpha + O_jdone
      mbar_wait(o_full, jdone & 1);
      tc_fence_after();
      uint32_t v[32];
      __syncwarp();
      tmem_ld_32x32b_x32(tO + lane_off + hf * 32, v);
      tmem_ld_wait();
#pragma unroll
      for (int e = 0; e < 32; ++e) o[e] = o[e] * alpha + __uint_as_float(v[e]);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(o_free);
    };

    for (int j = 0; j < nkv; ++j) {
      const bool diag = (j == qi);
      mbar_wait(s_full, j & 1);
      tc_fence_after();
      // pass 1: row maximum over this thread's 64 columns (four independent chains)
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        uint32_t v[32];
        __syncwarp();
        tmem_ld_32x32b_x32(tS + lane_off + hf * 64 + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          const float sc = __uint_as_float(v[e]);
          if (!diag || (hf * 64 + c * 32 + e) <= r) mx4[e & 3] = fmaxf(mx4[e & 3], sc);
        }
      }
      float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])) * kScale2;
      const int par = j & 1;
      xchg[(par * 2 + hf) * 128 + r] = mx;
      asm volatile("bar.sync 1, 256;" ::: "memory");
      mx = fmaxf(fmaxf(mx, xchg[(par * 2 + (hf ^ 1)) * 128 + r]), m);
      const float m_use = (mx == -INFINITY) ? 0.f : mx;
      const float alpha = exp2f(m - m_use);
      // P is single-buffered: P V of the previous tile must be finished before it is overwritten; folding its
      // result into the register accumulator here is exactly that wait
      if (j > 0) fold_o(j - 1, alpha_prev);
      float l4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        uint32_t v[32];
        __syncwarp();
        tmem_ld_32x32b_x32(tS + lane_off + hf * 64 + c * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint32_t keep = 0xFF;
          if (drop) keep = dropout_keep8(seed, offs, p.rng_stream, grp_row + ((j * kBlk + hf * 64 + c * 32 + g * 8) >> 3), thresh);
          float pv[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const int col = hf * 64 + c * 32 + g * 8 + e;
            float pe = exp2f(__uint_as_float(v[g * 8 + e]) * kScale2 - m_use);
            if (diag && col > r) pe = 0.f;
            l4[e & 3] += pe;
            pv[e] = ((keep >> e) & 1) ? pe : 0.f;
          }
          st_shared_v4_u32(swz(p_sub, r, c * 4 + g), pack_bf16x2(pv[0], pv[1]), pack_bf16x2(pv[2], pv[3]),
                           pack_bf16x2(pv[4], pv[5]), pack_bf16x2(pv[6], pv[7]));
        }
      }
      l = l * alpha + ((l4[0] + l4[1]) + (l4[2] + l4[3]));
      tc_fence_before();
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(p_ready);
        mbar_arrive(s_free);
      }
      alpha_prev = alpha;
      m = m_use;
    }
    fold_o(nkv - 1, alpha_prev);
    asm volatile("bar.sync 1, 256;" ::: "memory");
    xchg[hf * 128 + r] = l;
    asm volatile("bar.sync 1, 256;" ::: "memory");
    l += xchg[(hf ^ 1) * 128 + r];
    if (qrow < p.S) {
      const float inv = (drop ? 1.f / (1.f - p.p_drop) : 1.f) / l;
      __nv_bfloat16* dst = p.out + size_t(row_base + qrow) * h + hd * kD + hf * 32;
#pragma unroll
      for (int g = 0; g < 4; ++g)
        *reinterpret_cast<uint4*>(dst + g * 8) =
            make_uint4(pack_bf16x2(o[g * 8 + 0] * inv, o[g * 8 + 1] * inv), pack_bf16x2(o[g * 8 + 2] * inv, o[g * 8 + 3] * inv),
                       pack_bf16x2(o[g * 8 + 4] * inv, o[g * 8 + 5] * inv), pack_bf16x2(o[g * 8 + 6] * inv, o[g * 8 + 7] * inv));
      if (hf == 0) p.lse[size_t(b * p.H + hd) * p.S + qrow] = m + log2f(l);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<256>(tmem);
  }
}

}  // namespace f2
}  // namespace b200

using namespace b200;

extern "C" int b200_attn_fwd2(const void* qkv, void* out, float* lse, int batch, int seq, int heads, float dropout_p,
                              const unsigned long long* rng, unsigned int rng_stream, void* stream) {
  using namespace b200::f2;
  B200_REQUIRE(batch > 0 && seq > 0 && heads > 0, "attn_fwd2: empty problem");
  B200_REQUIRE(dropout_p == 0.f || rng != nullptr, "attn_fwd2: rng_state required when dropout_p > 0");
  const int h = heads * kD;
  CUtensorMap tm;
  int rc = make_tmap_2d_bf16(&tm, qkv, 3 * (uint64_t)h, (uint64_t)batch * seq, 3 * (uint64_t)h, 64, kBlk);
  if (rc) return rc;
  Params p{};
  p.B = batch; p.S = seq; p.H = heads; p.S8 = (seq + 7) / 8; p.p_drop = dropout_p; p.rng = rng; p.rng_stream = rng_stream;
  p.out = (__nv_bfloat16*)out; p.lse = lse;
  static bool configured[kMaxDevices] = {};   // the attribute is per device
  if (!configured[current_device()]) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(attn_fwd2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem::kTotal));
    configured[current_device()] = true;
  }
  dim3 grid(((seq + kBlk - 1) / kBlk) * heads * batch);
  attn_fwd2_kernel<<<grid, kThreads, Smem::kTotal, reinterpret_cast<cudaStream_t>(stream)>>>(tm, p);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}
