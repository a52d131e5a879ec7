// bf16 GEMM, CTA-pair edition: tcgen05.mma.cta_group::2 over a 256 x 256 output tile shared by the two
// CTAs of a 2-CTA cluster (one TPC).  Each CTA stages its own 128 rows of A and *half* of the B tile, so a
// 128 x 256 slice of output costs 32 KB of operand fill per K-block instead of the 48 KB of the single-CTA
// kernel (gemm_tcgen05.cu) -- the fix for the L2->SM fill bandwidth that bounds the small-K GEMMs of the
// transformer block (K = 768..1024: measured ~11 TB/s of operand fill with the 1-CTA 128x256 tile).
//
//   warp 0      TMA producer (both CTAs): own A rows + own half of B; completion bytes are credited to the
//               leader CTA's full barrier (cp.async.bulk.tensor ... .cta_group::2)
//   warp 1      TMEM allocation (both CTAs, cta_group::2); MMA issue by one thread of the LEADER CTA only;
//               tcgen05.commit multicasts "slot free" / "accumulator ready" to both CTAs
//   warps 2..5  epilogue (both CTAs): TMEM -> registers -> 128B-swizzled smem staging -> TMA store
//               (cp.async.bulk.tensor), or TMA reduce-add (cp.reduce.async.bulk.tensor .add.f32) for the
//               fp32 gradient accumulation -- coalesced full-line writes, no per-thread atomics, rows/columns
//               beyond M/N clipped by the tensor map.
//
// Same operand-layout freedom as the 1-CTA kernel (K-major / MN-major A and B), same b200_gemm_args.
// Reference ops: G3/G6/G8/G9 and their dgrad/wgrad (gpt.py:185-187,239,278-281,447), cuBLASLt there.
#include <cuda_bf16.h>

#include "../include/b200_trainer.h"
#include "common/host.h"
#include "common/ptx.cuh"

namespace b200 {
namespace g2 {

constexpr int BM = 128;        // rows per CTA (256 per pair)
constexpr int BNP = 256;       // tile width of the pair
constexpr int BNH = 128;       // B rows staged per CTA
constexpr int BK = 64;
constexpr int UMMA_K = 16;
constexpr int kAStage = BM * BK * 2;    // 16 KB
constexpr int kBStage = BNH * BK * 2;   // 16 KB
constexpr int kStage = kAStage + kBStage;
constexpr int kStages = 5;
constexpr int kEpiBufs = 4;
constexpr int kEpiBuf = 128 * 128;      // 128 rows x 128 B
constexpr int kBarOff = kStages * kStage + kEpiBufs * kEpiBuf;
constexpr int kSmemTotal = kBarOff + 256 + 1024;

enum { EPI_BF16 = 0, EPI_F32 = 1, EPI_F32_ATOMIC = 2 };

struct Params {
  int M, N, K;
  int split_k, k_per_split;
};

template <bool A_MN, bool B_MN, int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(192, 1)
gemm2_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
             const __grid_constant__ CUtensorMap tma_c, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages * kAStage;
  uint8_t* smem_epi = smem + kStages * kStage;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kBarOff);
  uint64_t* full_bar = bars;                      // [kStages]  (leader's copy is the one that is waited on)
  uint64_t* empty_bar = bars + kStages;           // [kStages]  (per CTA, signalled by multicast commit)
  uint64_t* tmem_full = bars + 2 * kStages;       // [2]        (per CTA, multicast commit)
  uint64_t* tmem_empty = bars + 2 * kStages + 2;  // [2]        (leader's copy: 8 epilogue warps arrive)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta_rank = cluster_ctarank();
  const bool leader = cta_rank == 0;
  const int pair = blockIdx.x >> 1, npairs = gridDim.x >> 1;

  const int tiles_m = (p.M + 2 * BM - 1) / (2 * BM);
  const int tiles_n = (p.N + BNP - 1) / BNP;
  const int num_tiles = tiles_m * tiles_n * p.split_k;
  const int k_blocks_total = (p.K + BK - 1) / BK;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    tma_prefetch_desc(&tma_c);
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 8);
    }
    fence_barrier_init();
  }
  cluster_sync_all();   // barrier inits visible to the peer before any remote arrive / multicast commit
  if (warp == 1) tmem_alloc_cg2<512>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (both CTAs)
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = pair; t < num_tiles; t += npairs) {
        const int ks = t / (tiles_m * tiles_n);
        const int r = t - ks * tiles_m * tiles_n;
        const int m0 = (r % tiles_m) * 2 * BM + int(cta_rank) * BM;   // this CTA's rows
        const int n0 = (r / tiles_m) * BNP + int(cta_rank) * BNH;     // this CTA's half of the B tile
        const int kb0 = ks * p.k_per_split;
        const int kb1 = min(k_blocks_total, kb0 + p.k_per_split);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if (leader) mbar_expect_tx(&full_bar[stage], 2 * kStage);   // bytes of both CTAs land on this barrier
          uint8_t* sa = smem_a + stage * kAStage;
          uint8_t* sb = smem_b + stage * kBStage;
          const int k0 = kb * BK;
          if constexpr (!A_MN) {
            tma_load_2d_cg2(sa, &tma_a, &full_bar[stage], k0, m0);
          } else {
#pragma unroll
            for (int c = 0; c < BM / 64; ++c) tma_load_2d_cg2(sa + c * (BK * 128), &tma_a, &full_bar[stage], m0 + c * 64, k0);
          }
          if constexpr (!B_MN) {
            tma_load_2d_cg2(sb, &tma_b, &full_bar[stage], k0, n0);
          } else {
#pragma unroll
            for (int c = 0; c < BNH / 64; ++c) tma_load_2d_cg2(sb + c * (BK * 128), &tma_b, &full_bar[stage], n0 + c * 64, k0);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA only)
    if (leader && lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(2 * BM, BNP, A_MN, B_MN);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int t = pair; t < num_tiles; t += npairs, ++it) {
        const int ks = t / (tiles_m * tiles_n);
        const int kb0 = ks * p.k_per_split;
        const int kb1 = min(k_blocks_total, kb0 + p.k_per_split);
        const int buf = it & 1;
        const uint32_t buf_phase = (it >> 1) & 1;
        mbar_wait(&tmem_empty[buf], buf_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * BNP;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem_a + stage * kAStage);
          const uint32_t b_addr = smem_u32(smem_b + stage * kBStage);
#pragma unroll
          for (int j = 0; j < BK / UMMA_K; ++j) {
            const uint64_t da = A_MN ? make_smem_desc_sw128(a_addr + j * (UMMA_K * 128), BK * 128, 1024)
                                     : make_smem_desc_sw128(a_addr + j * (UMMA_K * 2), 16, 1024);
            const uint64_t db = B_MN ? make_smem_desc_sw128(b_addr + j * (UMMA_K * 128), BK * 128, 1024)
                                     : make_smem_desc_sw128(b_addr + j * (UMMA_K * 2), 16, 1024);
            umma_bf16_ss_cg2(d_tmem, da, db, idesc, (kb > kb0 || j > 0) ? 1u : 0u);
          }
          tc_commit_cg2_mc(&empty_bar[stage], 0b11);   // both CTAs may refill this slot
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        tc_commit_cg2_mc(&tmem_full[buf], 0b11);       // both CTAs' epilogues may drain their half
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps (both CTAs)
    const int q = warp & 3;
    const int r_in = q * 32 + lane;                       // row within this CTA's 128 x 256 accumulator
    const bool epi_leader = threadIdx.x == 64;
    constexpr int kChunkCols = (EPI == EPI_BF16) ? 64 : 32;   // 128 bytes per row per staging buffer
    constexpr int kChunks = BNP / kChunkCols;
    uint32_t g = 0;                                       // running chunk counter -> staging ring slot
    int it = 0;
    for (int t = pair; t < num_tiles; t += npairs, ++it) {
      const int ks = t / (tiles_m * tiles_n);
      const int r = t - ks * tiles_m * tiles_n;
      const int m0 = (r % tiles_m) * 2 * BM + int(cta_rank) * BM;
      const int n0 = (r / tiles_m) * BNP;
      const int buf = it & 1;
      const uint32_t buf_phase = (it >> 1) & 1;
      (void)ks;
      mbar_wait(&tmem_full[buf], buf_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + (uint32_t(q * 32) << 16) + buf * BNP;
#pragma unroll 1
      for (int c = 0; c < kChunks; ++c, ++g) {
        const uint32_t slot = smem_u32(smem_epi + (g % kEpiBufs) * kEpiBuf);
        // the TMA store that last read this slot (kEpiBufs chunks ago) must be done with it
        if (epi_leader) tma_store_wait_read<kEpiBufs - 1>();
        named_bar_sync(1, 128);
        const uint32_t row_addr = slot + r_in * 128;
        if constexpr (EPI == EPI_BF16) {
          uint32_t v0[32], v1[32];
          tmem_ld_32x32b_x32(t_row + c * 64, v0);
          tmem_ld_32x32b_x32(t_row + c * 64 + 32, v1);
          tmem_ld_wait();
#pragma unroll
          for (int cc = 0; cc < 4; ++cc)
            st_shared_v4_u32(row_addr + ((cc ^ (r_in & 7)) << 4),
                             pack_bf16x2(__uint_as_float(v0[cc * 8 + 0]), __uint_as_float(v0[cc * 8 + 1])),
                             pack_bf16x2(__uint_as_float(v0[cc * 8 + 2]), __uint_as_float(v0[cc * 8 + 3])),
                             pack_bf16x2(__uint_as_float(v0[cc * 8 + 4]), __uint_as_float(v0[cc * 8 + 5])),
                             pack_bf16x2(__uint_as_float(v0[cc * 8 + 6]), __uint_as_float(v0[cc * 8 + 7])));
#pragma unroll
          for (int cc = 0; cc < 4; ++cc)
            st_shared_v4_u32(row_addr + (((cc + 4) ^ (r_in & 7)) << 4),
                             pack_bf16x2(__uint_as_float(v1[cc * 8 + 0]), __uint_as_float(v1[cc * 8 + 1])),
                             pack_bf16x2(__uint_as_float(v1[cc * 8 + 2]), __uint_as_float(v1[cc * 8 + 3])),
                             pack_bf16x2(__uint_as_float(v1[cc * 8 + 4]), __uint_as_float(v1[cc * 8 + 5])),
                             pack_bf16x2(__uint_as_float(v1[cc * 8 + 6]), __uint_as_float(v1[cc * 8 + 7])));
        } else {
          uint32_t v[32];
          tmem_ld_32x32b_x32(t_row + c * 32, v);
          tmem_ld_wait();
#pragma unroll
          for (int cc = 0; cc < 8; ++cc)
            st_shared_v4_u32(row_addr + ((cc ^ (r_in & 7)) << 4), v[cc * 4 + 0], v[cc * 4 + 1], v[cc * 4 + 2], v[cc * 4 + 3]);
        }
        if (c == kChunks - 1) {
          // accumulator fully read: hand the TMEM buffer back to the MMA thread of the leader CTA
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_leader(&tmem_empty[buf]);
        }
        fence_proxy_async_smem();
        named_bar_sync(1, 128);
        if (epi_leader) {
          if constexpr (EPI == EPI_F32_ATOMIC) tma_reduce_add_2d(&tma_c, smem_epi + (g % kEpiBufs) * kEpiBuf, n0 + c * kChunkCols, m0);
          else tma_store_2d(&tma_c, smem_epi + (g % kEpiBufs) * kEpiBuf, n0 + c * kChunkCols, m0);
          tma_store_commit();
        }
      }
    }
    if (epi_leader) tma_store_wait<0>();   // all bulk stores complete before the CTA (and its smem) goes away
  }

  tc_fence_before();
  cluster_sync_all();   // nobody exits (or frees TMEM) while the peer may still signal it
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_cg2<512>(tmem_base);
  }
}

template <bool A_MN, bool B_MN, int EPI>
static int launch(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const Params& p, int grid,
                  cudaStream_t stream) {
  auto kfn = gemm2_kernel<A_MN, B_MN, EPI>;
  static bool configured = false;
  if (!configured) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemTotal));
    configured = true;
  }
  kfn<<<grid, 192, kSmemTotal, stream>>>(ta, tb, tc, p);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

template <int EPI>
static int dispatch(bool a_mn, bool b_mn, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc,
                    const Params& p, int grid, cudaStream_t s) {
  if (!a_mn && !b_mn) return launch<false, false, EPI>(ta, tb, tc, p, grid, s);
  if (!a_mn && b_mn) return launch<false, true, EPI>(ta, tb, tc, p, grid, s);
  if (a_mn && b_mn) return launch<true, true, EPI>(ta, tb, tc, p, grid, s);
  return launch<true, false, EPI>(ta, tb, tc, p, grid, s);
}

}  // namespace g2
}  // namespace b200

using namespace b200;

extern "C" int b200_gemm2_bf16(const b200_gemm_args* a, void* stream_) {
  using namespace b200::g2;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  B200_REQUIRE(a != nullptr, "gemm2: null args");
  B200_REQUIRE(a->M > 0 && a->N > 0 && a->K > 0, "gemm2: empty problem %dx%dx%d", a->M, a->N, a->K);
  B200_REQUIRE(a->lda % 8 == 0 && a->ldb % 8 == 0, "gemm2: lda/ldb must be multiples of 8 elements");
  B200_REQUIRE(a->epilogue >= EPI_BF16 && a->epilogue <= EPI_F32_ATOMIC,
               "gemm2: epilogue %d is only available in the single-CTA kernel (b200_gemm_bf16)", a->epilogue);
  const bool a_mn = a->a_mn_major != 0, b_mn = a->b_mn_major != 0;
  const int c_elem = a->epilogue == EPI_BF16 ? 2 : 4;
  B200_REQUIRE((size_t(a->ldc) * c_elem) % 16 == 0, "gemm2: C row pitch must be a multiple of 16 bytes");

  CUtensorMap ta, tb, tc;
  int rc;
  if (!a_mn) rc = make_tmap_2d_bf16(&ta, a->A, a->K, a->M, a->lda, 64, BM);
  else       rc = make_tmap_2d_bf16(&ta, a->A, a->M, a->K, a->lda, 64, BK);
  if (rc) return rc;
  if (!b_mn) rc = make_tmap_2d_bf16(&tb, a->B, a->K, a->N, a->ldb, 64, BNH);
  else       rc = make_tmap_2d_bf16(&tb, a->B, a->N, a->K, a->ldb, 64, BK);
  if (rc) return rc;
  rc = make_tmap_2d_sw128(&tc, a->C, c_elem, a->N, a->M, a->ldc, 128);
  if (rc) return rc;

  Params p{};
  p.M = a->M; p.N = a->N; p.K = a->K;
  const int kblocks = (a->K + BK - 1) / BK;
  int split = a->split_k > 0 ? a->split_k : 1;
  if (split > kblocks) split = kblocks;
  if (split > 1) B200_REQUIRE(a->epilogue == EPI_F32_ATOMIC, "gemm2: split-K needs the fp32 reduce-add epilogue");
  p.k_per_split = (kblocks + split - 1) / split;
  p.split_k = (kblocks + p.k_per_split - 1) / p.k_per_split;

  const long tiles = long((a->M + 2 * BM - 1) / (2 * BM)) * ((a->N + BNP - 1) / BNP) * p.split_k;
  int sms = sm_count();
  if (a->max_ctas > 0 && a->max_ctas < sms) sms = a->max_ctas;
  long pairs = sms / 2;
  if (pairs < 1) pairs = 1;
  if (tiles < pairs) pairs = tiles;
  const int grid = int(pairs * 2);
  switch (a->epilogue) {
    case EPI_BF16: return dispatch<EPI_BF16>(a_mn, b_mn, ta, tb, tc, p, grid, stream);
    case EPI_F32: return dispatch<EPI_F32>(a_mn, b_mn, ta, tb, tc, p, grid, stream);
    default: return dispatch<EPI_F32_ATOMIC>(a_mn, b_mn, ta, tb, tc, p, grid, stream);
  }
}
