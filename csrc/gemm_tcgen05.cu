// bf16 GEMM on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM, operands staged by
// TMA into 128B-swizzled shared memory), persistent and warp-specialised:
//
//   warp 0      TMA producer     (one elected lane)   global -> smem ring (kStages deep)
//   warp 1      MMA issuer       (one elected lane)   smem -> TMEM accumulator (double buffered)
//   warps 2..5  epilogue         (128 threads)        TMEM -> registers -> fused epilogue -> global
//
// C[M,N] = A * B with fp32 accumulation.  Both operands may be K-major (contraction dim contiguous:
// the forward y = x W^T case) or MN-major (contraction dim strided: dgrad reads W, wgrad reads dY^T
// and X), selected per operand in the instruction descriptor -- no transposed copies are ever made.
//
// This kernel is the engine behind reference ops G3/G6/G8/G9 (q/k/v/o projections, SwiGLU MLP, LM
// head; /root/reference/src/models/gpt.py:185-187,239,278-281,447) and their backward passes, which
// the reference runs through cuBLASLt.
//
// Epilogues (EPI_*), all fused on the TMEM->register path:
//   BF16          C(bf16)  = acc
//   F32           C(fp32)  = acc                      (beta = 0)
//   F32_ATOMIC    C(fp32) += acc via red.global.add   (wgrad accumulation, split-K friendly)
//   ROPE_BF16     C(bf16)  = rotary(acc) on the q/k column range, plain copy on v   (gpt.py:116-147)
//   RESID_DROPOUT out(fp32) = residual(fp32) + dropout(acc)                         (gpt.py:240,308)
//   SWIGLU        gu(bf16) = acc ; act(bf16) = silu(gate) * up                      (gpt.py:278-281)
//   SWIGLU_BWD    dgu(bf16) = d(silu(g)*u) given acc = d_act
#include <cuda_bf16.h>

#include "../include/b200_trainer.h"
#include "common/host.h"
#include "common/philox.cuh"
#include "common/ptx.cuh"

namespace b200 {

constexpr int BM = 128;
constexpr int BK = 64;  // 64 bf16 = one 128-byte swizzle row
constexpr int UMMA_K = 16;

enum Epi : int {
  EPI_BF16 = 0,
  EPI_F32 = 1,
  EPI_F32_ATOMIC = 2,
  EPI_ROPE_BF16 = 3,
  EPI_RESID_DROPOUT = 4,
  EPI_SWIGLU = 5,
  EPI_SWIGLU_BWD = 6,
};

struct GemmParams {
  int M, N, K;
  int ldc;              // elements
  void* C;              // primary output
  int split_k;          // >= 1
  int k_per_split;      // in BK blocks
  // epilogue extras
  const float* rope_cos;   // [S, 32]
  const float* rope_sin;   // [S, 32]
  int seq_len;
  int rope_cols;           // columns [0, rope_cols) are rotated (q and k), head_dim = 64
  const float* residual;   // [M, ldc] fp32
  float dropout_p;
  const unsigned long long* rng;  // {seed, offset} in device memory (graph-capture safe)
  unsigned int rng_stream;        // per-op stream id mixed into the Philox key
  void* aux;               // SWIGLU: act output [M, N/2] ; SWIGLU_BWD: gu input [M, 2N]
  int ld_aux;
  int half_n;              // SWIGLU: I (up columns live at +half_n in B rows / gu columns)
};

template <int BN>
struct SmemLayout {
  static constexpr int kAStage = BM * BK * 2;
  static constexpr int kBStage = BN * BK * 2;
  static constexpr int kStage = kAStage + kBStage;
  static constexpr int kStages = (BN == 256) ? 4 : 6;
  static constexpr int kBarrierBytes = 256;
  static constexpr int kTotal = kStages * kStage + kBarrierBytes + 1024;  // +1024 for manual alignment
};

__device__ __forceinline__ float silu_f(float x) { return x / (1.0f + __expf(-x)); }

template <int BN, bool A_MN, bool B_MN, int EPI>
__global__ void __launch_bounds__(192, 1)
gemm_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b, const GemmParams p) {
  using L = SmemLayout<BN>;
  constexpr int kStages = L::kStages;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages * L::kAStage;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * L::kStage);
  uint64_t* full_bar = bars;                      // [kStages]
  uint64_t* empty_bar = bars + kStages;           // [kStages]
  uint64_t* tmem_full = bars + 2 * kStages;       // [2]
  uint64_t* tmem_empty = bars + 2 * kStages + 2;  // [2]
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const int tiles_m = (p.M + BM - 1) / BM;
  const int tiles_n = (p.N + BN - 1) / BN;
  const int num_tiles = tiles_m * tiles_n * p.split_k;
  const int k_blocks_total = (p.K + BK - 1) / BK;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 4);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<2 * BN>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        const int ks = t / (tiles_m * tiles_n);
        const int r = t - ks * tiles_m * tiles_n;
        const int m0 = (r % tiles_m) * BM;
        const int n0 = (r / tiles_m) * BN;
        const int kb0 = ks * p.k_per_split;
        const int kb1 = min(k_blocks_total, kb0 + p.k_per_split);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_expect_tx(&full_bar[stage], L::kStage);
          uint8_t* sa = smem_a + stage * L::kAStage;
          uint8_t* sb = smem_b + stage * L::kBStage;
          const int k0 = kb * BK;
          if constexpr (!A_MN) {
            tma_load_2d(sa, &tma_a, &full_bar[stage], k0, m0);
          } else {
#pragma unroll
            for (int c = 0; c < BM / 64; ++c) tma_load_2d(sa + c * (BK * 128), &tma_a, &full_bar[stage], m0 + c * 64, k0);
          }
          if constexpr (EPI == EPI_SWIGLU) {
            // B rows [n0/2, n0/2 + BN/2) of the gate matrix and the same rows of the up matrix
            static_assert(!B_MN || EPI != EPI_SWIGLU, "SwiGLU epilogue expects a K-major weight");
            tma_load_2d(sb, &tma_b, &full_bar[stage], k0, n0 / 2);
            tma_load_2d(sb + (BN / 2) * 128, &tma_b, &full_bar[stage], k0, p.half_n + n0 / 2);
          } else if constexpr (!B_MN) {
            tma_load_2d(sb, &tma_b, &full_bar[stage], k0, n0);
          } else {
#pragma unroll
            for (int c = 0; c < BN / 64; ++c) tma_load_2d(sb + c * (BK * 128), &tma_b, &full_bar[stage], n0 + c * 64, k0);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BM, BN, A_MN, B_MN);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++it) {
        const int ks = t / (tiles_m * tiles_n);
        const int kb0 = ks * p.k_per_split;
        const int kb1 = min(k_blocks_total, kb0 + p.k_per_split);
        const int buf = it & 1;
        const uint32_t buf_phase = (it >> 1) & 1;
        mbar_wait(&tmem_empty[buf], buf_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem_a + stage * L::kAStage);
          const uint32_t b_addr = smem_u32(smem_b + stage * L::kBStage);
#pragma unroll
          for (int j = 0; j < BK / UMMA_K; ++j) {
            const uint64_t da = A_MN ? make_smem_desc_sw128(a_addr + j * (UMMA_K * 128), BK * 128, 1024)
                                     : make_smem_desc_sw128(a_addr + j * (UMMA_K * 2), 16, 1024);
            const uint64_t db = B_MN ? make_smem_desc_sw128(b_addr + j * (UMMA_K * 128), BK * 128, 1024)
                                     : make_smem_desc_sw128(b_addr + j * (UMMA_K * 2), 16, 1024);
            umma_bf16_ss(d_tmem, da, db, idesc, (kb > kb0 || j > 0) ? 1u : 0u);
          }
          tc_commit(&empty_bar[stage]);  // smem slot reusable once these MMAs retire
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        tc_commit(&tmem_full[buf]);  // accumulator complete
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    int it = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++it) {
      const int ks = t / (tiles_m * tiles_n);
      const int r = t - ks * tiles_m * tiles_n;
      const int m0 = (r % tiles_m) * BM;
      const int n0 = (r / tiles_m) * BN;
      const int buf = it & 1;
      const uint32_t buf_phase = (it >> 1) & 1;
      mbar_wait(&tmem_full[buf], buf_phase);
      tc_fence_after();
      const int row = m0 + q * 32 + lane;
      const bool row_ok = row < p.M;
      const uint32_t t_row = tmem_base + (uint32_t(q * 32) << 16) + buf * BN;
      (void)ks;

      if constexpr (EPI == EPI_BF16 || EPI == EPI_F32 || EPI == EPI_F32_ATOMIC || EPI == EPI_RESID_DROPOUT ||
                    EPI == EPI_SWIGLU_BWD) {
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
          uint32_t v[32];
          __syncwarp();
          tmem_ld_32x32b_x32(t_row + c * 32, v);
          tmem_ld_wait();
          const int col0 = n0 + c * 32;
          if (row_ok && col0 < p.N) {
          if constexpr (EPI == EPI_BF16) {
            __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(p.C) + size_t(row) * p.ldc + col0;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              if (col0 + g * 8 < p.N) {
                uint4 o;
                o.x = pack_bf16x2(__uint_as_float(v[g * 8 + 0]), __uint_as_float(v[g * 8 + 1]));
                o.y = pack_bf16x2(__uint_as_float(v[g * 8 + 2]), __uint_as_float(v[g * 8 + 3]));
                o.z = pack_bf16x2(__uint_as_float(v[g * 8 + 4]), __uint_as_float(v[g * 8 + 5]));
                o.w = pack_bf16x2(__uint_as_float(v[g * 8 + 6]), __uint_as_float(v[g * 8 + 7]));
                *reinterpret_cast<uint4*>(out + g * 8) = o;
              }
            }
          } else if constexpr (EPI == EPI_F32) {
            float* out = reinterpret_cast<float*>(p.C) + size_t(row) * p.ldc + col0;
#pragma unroll
            for (int g = 0; g < 8; ++g) {
              if (col0 + g * 4 < p.N) {
                float4 o = make_float4(__uint_as_float(v[g * 4 + 0]), __uint_as_float(v[g * 4 + 1]),
                                       __uint_as_float(v[g * 4 + 2]), __uint_as_float(v[g * 4 + 3]));
                *reinterpret_cast<float4*>(out + g * 4) = o;
              }
            }
          } else if constexpr (EPI == EPI_F32_ATOMIC) {
            float* out = reinterpret_cast<float*>(p.C) + size_t(row) * p.ldc + col0;
#pragma unroll
            for (int g = 0; g < 8; ++g) {
              if (col0 + g * 4 < p.N) {
                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(out + g * 4),
                             "f"(__uint_as_float(v[g * 4 + 0])), "f"(__uint_as_float(v[g * 4 + 1])),
                             "f"(__uint_as_float(v[g * 4 + 2])), "f"(__uint_as_float(v[g * 4 + 3]))
                             : "memory");
              }
            }
          } else if constexpr (EPI == EPI_RESID_DROPOUT) {
            // out = residual + dropout(acc); one Philox call covers 8 consecutive columns (16 bits each)
            const float* res = p.residual + size_t(row) * p.ldc + col0;
            float* out = reinterpret_cast<float*>(p.C) + size_t(row) * p.ldc + col0;
            const bool drop = p.dropout_p > 0.f;
            const uint32_t thresh = dropout_threshold16(p.dropout_p);
            const float inv_keep = drop ? 1.0f / (1.0f - p.dropout_p) : 1.0f;
            unsigned long long seed = 0, offs = 0;
            if (drop) { seed = p.rng[0]; offs = p.rng[1]; }
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              if (col0 + g * 8 < p.N) {
                uint32_t keep = 0xFF;
                if (drop) keep = dropout_keep8(seed, offs, p.rng_stream, (unsigned long long)row * (p.N / 8) + (col0 / 8 + g), thresh);
                const float4 r0 = *reinterpret_cast<const float4*>(res + g * 8);
                const float4 r1 = *reinterpret_cast<const float4*>(res + g * 8 + 4);
                float a[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) a[e] = ((keep >> e) & 1) ? __uint_as_float(v[g * 8 + e]) * inv_keep : 0.f;
                *reinterpret_cast<float4*>(out + g * 8) = make_float4(r0.x + a[0], r0.y + a[1], r0.z + a[2], r0.w + a[3]);
                *reinterpret_cast<float4*>(out + g * 8 + 4) = make_float4(r1.x + a[4], r1.y + a[5], r1.z + a[6], r1.w + a[7]);
              }
            }
          } else {  // EPI_SWIGLU_BWD: acc = d_act[row, col]; gu = [gate | up] with up at +half_n
            const __nv_bfloat16* gu = reinterpret_cast<const __nv_bfloat16*>(p.aux) + size_t(row) * p.ld_aux + col0;
            __nv_bfloat16* dgu = reinterpret_cast<__nv_bfloat16*>(p.C) + size_t(row) * p.ldc + col0;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              if (col0 + g * 8 < p.N) {
                const uint4 gq = *reinterpret_cast<const uint4*>(gu + g * 8);
                const uint4 uq = *reinterpret_cast<const uint4*>(gu + p.half_n + g * 8);
                const uint32_t gw[4] = {gq.x, gq.y, gq.z, gq.w};
                const uint32_t uw[4] = {uq.x, uq.y, uq.z, uq.w};
                uint32_t dgw[4], duw[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  const float2 gg = unpack_bf16x2(gw[e]);
                  const float2 uu = unpack_bf16x2(uw[e]);
                  const float d0 = __uint_as_float(v[g * 8 + 2 * e]);
                  const float d1 = __uint_as_float(v[g * 8 + 2 * e + 1]);
                  const float s0 = 1.0f / (1.0f + __expf(-gg.x));
                  const float s1 = 1.0f / (1.0f + __expf(-gg.y));
                  // d/dg [g*sig(g)] = sig(g) * (1 + g*(1-sig(g)))
                  dgw[e] = pack_bf16x2(d0 * uu.x * s0 * (1.0f + gg.x * (1.0f - s0)),
                                       d1 * uu.y * s1 * (1.0f + gg.y * (1.0f - s1)));
                  duw[e] = pack_bf16x2(d0 * gg.x * s0, d1 * gg.y * s1);
                }
                *reinterpret_cast<uint4*>(dgu + g * 8) = make_uint4(dgw[0], dgw[1], dgw[2], dgw[3]);
                *reinterpret_cast<uint4*>(dgu + p.half_n + g * 8) = make_uint4(duw[0], duw[1], duw[2], duw[3]);
              }
            }
          }
          }  // row_ok && col0 < N
        }
      } else if constexpr (EPI == EPI_ROPE_BF16) {
        // 64-column groups = one head; column j pairs with j+32 (half-split rotation).
        const int pos = row % p.seq_len;
#pragma unroll 1
        for (int hgrp = 0; hgrp < BN / 64; ++hgrp) {
          uint32_t lo[32], hi[32];
          __syncwarp();
          tmem_ld_32x32b_x32(t_row + hgrp * 64, lo);
          tmem_ld_32x32b_x32(t_row + hgrp * 64 + 32, hi);
          tmem_ld_wait();
          const int col0 = n0 + hgrp * 64;
          if (row_ok && col0 < p.N) {
          __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(p.C) + size_t(row) * p.ldc + col0;
          const bool rot = col0 < p.rope_cols;
          const float4* cs = reinterpret_cast<const float4*>(p.rope_cos + size_t(pos) * 32);
          const float4* sn = reinterpret_cast<const float4*>(p.rope_sin + size_t(pos) * 32);
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            float a[8], b[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              a[e] = __uint_as_float(lo[g * 8 + e]);
              b[e] = __uint_as_float(hi[g * 8 + e]);
            }
            if (rot) {
              const float4 c0 = cs[g * 2], c1 = cs[g * 2 + 1], s0 = sn[g * 2], s1 = sn[g * 2 + 1];
              const float cc[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
              const float ss[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
#pragma unroll
              for (int e = 0; e < 8; ++e) {
                const float x1 = a[e], x2 = b[e];
                a[e] = x1 * cc[e] - x2 * ss[e];  // first half:  x1*cos - x2*sin
                b[e] = x2 * cc[e] + x1 * ss[e];  // second half: x2*cos + x1*sin
              }
            }
            *reinterpret_cast<uint4*>(out + g * 8) =
                make_uint4(pack_bf16x2(a[0], a[1]), pack_bf16x2(a[2], a[3]), pack_bf16x2(a[4], a[5]), pack_bf16x2(a[6], a[7]));
            *reinterpret_cast<uint4*>(out + 32 + g * 8) =
                make_uint4(pack_bf16x2(b[0], b[1]), pack_bf16x2(b[2], b[3]), pack_bf16x2(b[4], b[5]), pack_bf16x2(b[6], b[7]));
          }
          }  // row_ok && col0 < N
        }
      } else if constexpr (EPI == EPI_SWIGLU) {
        // accumulator columns [0, BN/2) = gate, [BN/2, BN) = up, for intermediate columns n0/2 ...
        __nv_bfloat16* gu = reinterpret_cast<__nv_bfloat16*>(p.C) + size_t(row) * p.ldc;
        __nv_bfloat16* act = reinterpret_cast<__nv_bfloat16*>(p.aux) + size_t(row) * p.ld_aux;
#pragma unroll 1
        for (int c = 0; c < BN / 64; ++c) {
          uint32_t gv[32], uv[32];
          __syncwarp();
          tmem_ld_32x32b_x32(t_row + c * 32, gv);
          tmem_ld_32x32b_x32(t_row + BN / 2 + c * 32, uv);
          tmem_ld_wait();
          const int col0 = n0 / 2 + c * 32;  // intermediate column
          if (row_ok && col0 < p.half_n) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            uint32_t gw[4], uw[4], aw[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float g0 = __uint_as_float(gv[g * 8 + 2 * e]), g1 = __uint_as_float(gv[g * 8 + 2 * e + 1]);
              const float u0 = __uint_as_float(uv[g * 8 + 2 * e]), u1 = __uint_as_float(uv[g * 8 + 2 * e + 1]);
              gw[e] = pack_bf16x2(g0, g1);
              uw[e] = pack_bf16x2(u0, u1);
              // activation from the bf16-rounded gate/up so recomputation in backward is bit-identical
              const float2 gr = unpack_bf16x2(gw[e]);
              const float2 ur = unpack_bf16x2(uw[e]);
              aw[e] = pack_bf16x2(silu_f(gr.x) * ur.x, silu_f(gr.y) * ur.y);
            }
            *reinterpret_cast<uint4*>(gu + col0 + g * 8) = make_uint4(gw[0], gw[1], gw[2], gw[3]);
            *reinterpret_cast<uint4*>(gu + p.half_n + col0 + g * 8) = make_uint4(uw[0], uw[1], uw[2], uw[3]);
            *reinterpret_cast<uint4*>(act + col0 + g * 8) = make_uint4(aw[0], aw[1], aw[2], aw[3]);
          }
          }  // row_ok && col0 < half_n
        }
      }
      // release the accumulator buffer back to the MMA warp
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[buf]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<2 * BN>(tmem_base);
  }
}

// ------------------------------------------------------------------------------------------------
// host launch
// ------------------------------------------------------------------------------------------------
template <int BN, bool A_MN, bool B_MN, int EPI>
static int launch(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int grid, cudaStream_t stream) {
  auto kfn = gemm_kernel<BN, A_MN, B_MN, EPI>;
  static bool configured[kMaxDevices] = {};   // the attribute is per device
  if (!configured[current_device()]) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, SmemLayout<BN>::kTotal));
    configured[current_device()] = true;
  }
  kfn<<<grid, 192, SmemLayout<BN>::kTotal, stream>>>(ta, tb, p);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

template <int BN, int EPI>
static int dispatch_major(bool a_mn, bool b_mn, const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p,
                          int grid, cudaStream_t s) {
  if (!a_mn && !b_mn) return launch<BN, false, false, EPI>(ta, tb, p, grid, s);
  if constexpr (EPI == EPI_BF16 || EPI == EPI_F32 || EPI == EPI_F32_ATOMIC || EPI == EPI_SWIGLU_BWD) {
    if (!a_mn && b_mn) return launch<BN, false, true, EPI>(ta, tb, p, grid, s);
    if (a_mn && b_mn) return launch<BN, true, true, EPI>(ta, tb, p, grid, s);
    if (a_mn && !b_mn) return launch<BN, true, false, EPI>(ta, tb, p, grid, s);
  }
  set_error("gemm: operand layout (a_mn=%d, b_mn=%d) not instantiated for epilogue %d", (int)a_mn, (int)b_mn, EPI);
  return -2;
}

template <int BN>
static int dispatch_epi(int epi, bool a_mn, bool b_mn, const CUtensorMap& ta, const CUtensorMap& tb,
                        const GemmParams& p, int grid, cudaStream_t s) {
  switch (epi) {
    case EPI_BF16: return dispatch_major<BN, EPI_BF16>(a_mn, b_mn, ta, tb, p, grid, s);
    case EPI_F32: return dispatch_major<BN, EPI_F32>(a_mn, b_mn, ta, tb, p, grid, s);
    case EPI_F32_ATOMIC: return dispatch_major<BN, EPI_F32_ATOMIC>(a_mn, b_mn, ta, tb, p, grid, s);
    case EPI_ROPE_BF16: return dispatch_major<BN, EPI_ROPE_BF16>(a_mn, b_mn, ta, tb, p, grid, s);
    case EPI_RESID_DROPOUT: return dispatch_major<BN, EPI_RESID_DROPOUT>(a_mn, b_mn, ta, tb, p, grid, s);
    case EPI_SWIGLU: return dispatch_major<BN, EPI_SWIGLU>(a_mn, b_mn, ta, tb, p, grid, s);
    case EPI_SWIGLU_BWD: return dispatch_major<BN, EPI_SWIGLU_BWD>(a_mn, b_mn, ta, tb, p, grid, s);
  }
  set_error("gemm: unknown epilogue %d", epi);
  return -2;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_gemm_bf16(const b200_gemm_args* a, void* stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  B200_REQUIRE(a != nullptr, "gemm: null args");
  B200_REQUIRE(a->M > 0 && a->N > 0 && a->K > 0, "gemm: empty problem %dx%dx%d", a->M, a->N, a->K);
  B200_REQUIRE(a->lda % 8 == 0 && a->ldb % 8 == 0, "gemm: lda/ldb must be multiples of 8 elements");
  B200_REQUIRE(a->N % 8 == 0, "gemm: N must be a multiple of 8 (pad the output)");
  const bool a_mn = a->a_mn_major != 0, b_mn = a->b_mn_major != 0;
  // MN-major operands need no M/N alignment: TMA zero-fills the columns of a 64-wide box that fall
  // outside the tensor, and the epilogue predicates rows/columns.

  int bn = a->tile_n;
  if (bn != 128 && bn != 256) {
    // pick the tile width with the better wave efficiency on this device
    const int sms = sm_count();
    auto eff = [&](int w) {
      long tiles = long((a->M + BM - 1) / BM) * ((a->N + w - 1) / w) * (a->split_k > 0 ? a->split_k : 1);
      long waves = (tiles + sms - 1) / sms;
      double useful = double(a->M) * a->N;
      return useful / (double(waves) * sms * BM * w);
    };
    // Measured on B200 (benchmarks/bench_gemm.py): the 256-wide tile wins even at 65% wave efficiency
    // (N = 768) because it halves A re-reads from L2; only fall back when it wastes most of a wave.
    bn = (a->N >= 256 && eff(256) >= eff(128) * 0.6) ? 256 : 128;
  }
  if (a->epilogue == EPI_ROPE_BF16) B200_REQUIRE(a->rope_cols % 64 == 0, "gemm: rope_cols must be a multiple of 64");
  if (a->epilogue == EPI_SWIGLU)
    B200_REQUIRE(a->half_n % (bn / 2) == 0 && a->N == 2 * a->half_n, "gemm: SwiGLU needs N == 2*half_n, half_n %% %d == 0", bn / 2);

  CUtensorMap ta, tb;
  int rc;
  if (!a_mn) rc = make_tmap_2d_bf16(&ta, a->A, a->K, a->M, a->lda, 64, BM);
  else       rc = make_tmap_2d_bf16(&ta, a->A, a->M, a->K, a->lda, 64, BK);
  if (rc) return rc;
  const int b_rows_box = (a->epilogue == EPI_SWIGLU) ? bn / 2 : bn;
  if (!b_mn) rc = make_tmap_2d_bf16(&tb, a->B, a->K, a->N, a->ldb, 64, b_rows_box);
  else       rc = make_tmap_2d_bf16(&tb, a->B, a->N, a->K, a->ldb, 64, BK);
  if (rc) return rc;

  GemmParams p{};
  p.M = a->M; p.N = a->N; p.K = a->K; p.ldc = a->ldc; p.C = a->C;
  const int kblocks = (a->K + BK - 1) / BK;
  int split = a->split_k > 0 ? a->split_k : 1;
  if (split > kblocks) split = kblocks;
  if (split > 1) B200_REQUIRE(a->epilogue == EPI_F32_ATOMIC, "gemm: split-K needs the atomic fp32 epilogue");
  p.k_per_split = (kblocks + split - 1) / split;
  p.split_k = (kblocks + p.k_per_split - 1) / p.k_per_split;
  p.rope_cos = a->rope_cos; p.rope_sin = a->rope_sin; p.seq_len = a->seq_len > 0 ? a->seq_len : 1; p.rope_cols = a->rope_cols;
  p.residual = a->residual; p.dropout_p = a->dropout_p; p.rng = a->rng_state; p.rng_stream = a->rng_stream;
  p.aux = a->aux; p.ld_aux = a->ld_aux; p.half_n = a->half_n;

  const long tiles = long((a->M + BM - 1) / BM) * ((a->N + bn - 1) / bn) * p.split_k;
  int grid = int(tiles < sm_count() ? tiles : sm_count());
  if (a->max_ctas > 0 && grid > a->max_ctas) grid = a->max_ctas;
  if (bn == 256) return dispatch_epi<256>(a->epilogue, a_mn, b_mn, ta, tb, p, grid, stream);
  return dispatch_epi<128>(a->epilogue, a_mn, b_mn, ta, tb, p, grid, stream);
}
