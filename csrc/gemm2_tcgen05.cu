// bf16 GEMM, CTA-pair edition: tcgen05.mma.cta_group::2 over a 256 x BNP output tile shared by the two
// CTAs of a 2-CTA cluster (one TPC).  Each CTA stages its own 128 rows of A and *half* of the B tile, so a
// 128 x 256 slice of output costs 32 KB of operand fill per K-block instead of the 48 KB of the single-CTA
// kernel (gemm_tcgen05.cu) -- the fix for the L2->SM fill bandwidth that bounds the small-K GEMMs of the
// transformer block (K = 768..1024: measured ~11 TB/s of operand fill with the 1-CTA 128x256 tile).
//
//   warp 0      TMA producer (both CTAs): own A rows + own half of B; completion bytes are credited to the
//               leader CTA's full barrier (cp.async.bulk.tensor ... .cta_group::2)
//   warp 1      TMEM allocation (both CTAs, cta_group::2); MMA issue by one thread of the LEADER CTA only;
//               tcgen05.commit multicasts "slot free" / "accumulator ready" to both CTAs
//   warps 2..5  epilogue (both CTAs): TMEM -> registers -> fused math -> 128B-swizzled smem staging ring ->
//               TMA store (or TMA reduce-add for fp32 gradient accumulation).  Epilogue *inputs* (fp32
//               residual, saved gate|up) are prefetched into the same ring by TMA loads, so every global
//               access of the epilogue is a full-line bulk transfer and rows/columns beyond M/N are clipped by
//               the tensor maps.
//
// Epilogues: BF16, F32, F32_ATOMIC (TMA reduce-add), ROPE_BF16, RESID_DROPOUT, SWIGLU, SWIGLU_BWD -- the same
// set and the same b200_gemm_args as the single-CTA kernel; see that file / the header for the math.
// Reference ops: G3/G4/G6/G7/G8/G9 and their dgrad/wgrad (gpt.py:116-147,185-187,239-240,278-282,447).
#include <cuda_bf16.h>

#include "../include/b200_trainer.h"
#include "common/host.h"
#include "common/philox.cuh"
#include "common/ptx.cuh"

namespace b200 {
namespace g2 {

constexpr int BM = 128;   // rows per CTA (256 per pair)
constexpr int BK = 64;
constexpr int UMMA_K = 16;
constexpr int kAStage = BM * BK * 2;   // 16 KB
constexpr int kEpiBuf = 128 * 128;     // staging slot: 128 rows x 128 B
#ifndef B200_EPI_WARPS
#define B200_EPI_WARPS 8
#endif
// Epilogue warps: two per scheduler; the warps of a lane quarter split a unit's columns kParts ways.  Going from one
// to two warps per scheduler halved the fused epilogues' time in round 1; -DB200_EPI_WARPS=16 (four per scheduler)
// measured no further gain in round 2 (SwiGLU 70.6 vs 69.6 us, SwiGLU' 62.3 vs 62.4, RoPE 37.9 vs 34.8), and an L2
// prefetch of a tile's epilogue inputs made them slower (SwiGLU' 70.5 us): these epilogues are bound by the staging
// ring's load -> compute -> store round trip per unit, not by issue slots.  Trading one operand stage for a deeper
// input prefetch (residual 4 units ahead, SwiGLU' 2) was slower as well (small down+resid 53.3 -> 60.4 us, SwiGLU'
// 63.5 -> 68.5 us): the mainloop needs its 4-5 stages more than the epilogue needs the slots.  An ablation of the
// SwiGLU' epilogue (8192 x 3072 x 768: everything 62.5 us; no stores 56.3; no loads 52.2; no math 60.4; loads + stores
// only 60.4; only the accumulator read 37.9) says what the two compete for: shared-memory bandwidth.  The main loop
// moves ~80 KB per k-block through smem (TMA in, UMMA operand reads), a 64-column epilogue unit another 128 KB (TMA
// in, LDS, STS, TMA out), together close to the 128 B/clk an SM delivers.
constexpr int kEpiWarps = B200_EPI_WARPS;
static_assert(kEpiWarps == 8 || kEpiWarps == 16, "8 or 16 epilogue warps");
constexpr int kParts = kEpiWarps / 4;  // column parts of a unit (one per warp of a lane quarter)
constexpr int kCB = 64 / kParts;       // bf16 columns per thread in a 64-column unit
constexpr int kCF = 32 / kParts;       // fp32 columns per thread in a 32-column unit
constexpr int kCh = 8 / kParts;        // 16-byte chunks of a staging row per thread
constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kThreads = 64 + kEpiThreads;

enum {
  EPI_BF16 = 0, EPI_F32 = 1, EPI_F32_ATOMIC = 2, EPI_ROPE_BF16 = 3, EPI_RESID_DROPOUT = 4, EPI_SWIGLU = 5,
  EPI_SWIGLU_BWD = 6
};

template <int BNP, int EPI>
struct Cfg {
  static constexpr int BNH = BNP / 2;                       // B rows staged per CTA
  static constexpr int kBStage = BNH * BK * 2;
  static constexpr int kStage = kAStage + kBStage;
  static constexpr int kEpiBufs = (EPI == EPI_SWIGLU_BWD) ? 6 : 4;
  static constexpr int kStages = (EPI == EPI_SWIGLU_BWD || BNP == 384) ? 4 : (BNP == 256 ? 5 : 6);
  static constexpr int kEpiOff = kStages * kStage;
  static constexpr int kBarOff = kEpiOff + kEpiBufs * kEpiBuf;
  static constexpr int kSmemTotal = kBarOff + 256 + 1024;
  // Accumulator: double buffered, except the 384-wide tile (N = 768 problems: 64 tiles = one wave on 74
  // pairs instead of two 65%-full waves of 256-wide tiles), which owns 384 of the 512 columns once.
  static constexpr int kAccBufs = BNP == 384 ? 1 : 2;
  static constexpr uint32_t kTmemCols = BNP == 384 ? 512 : 2 * BNP;
  static_assert(BNP != 384 || EPI == EPI_BF16 || EPI == EPI_RESID_DROPOUT, "384-wide tile: plain / residual epilogues only");
  static_assert(kSmemTotal <= 232448, "shared memory budget exceeded");
};

struct Params {
  int M, N, K;
  int split_k, k_per_split;
  const float* rope_cos;
  const float* rope_sin;
  int seq_len, rope_cols;
  float dropout_p;
  const unsigned long long* rng;
  unsigned int rng_stream;
  int half_n;
  // ---- all-gather fused into the GEMM (AG instantiations): B is a weight sharded across ranks -------------
  const void* ag_src[16];      // every rank's bf16 shard of the unit that contains B (peer-mapped)
  void* ag_dst;                // local gathered copy of B ([N, K] bf16, contiguous) = what the B tensor map reads
  long long ag_part;           // elements of the unit owned by each rank
  long long ag_b_off;          // element offset of B inside the unit
  unsigned int* ag_flags;      // local, one per 128-row block of B: == ag_epoch once the block has landed
  unsigned int ag_epoch;
  const unsigned int* ag_epoch_ptr;   // when set, the epoch is read from device memory (CUDA-graph replays see a fresh one)
  int ag_pairs;                // CTA pairs that gather instead of multiplying
  int ag_rot;                  // n-tile rotation so tiles over locally owned rows are visited first
  int ag_rank, ag_world;
  // ---- reduce-scatter fused into the epilogue (OWN instantiations): C's 128-row slabs belong to ranks ---------
  int own_world, own_slabs, own_rot;   // owner of slab s = ((s * world) / slabs + rot) % world
  int n_fast;                          // tile order: n-tiles of a row block consecutive (see mtile/ntile)
  int own_prev;                        // tma_x = this rank's own accumulator: owner += acc + prev
};

// One output tensor map per owner (OWN instantiations): map o covers the rows of C that rank o owns, at their
// address in rank o's memory (peer-mapped over NVLink, or local for o == this rank).
constexpr int kMaxOwners = 8;
template <bool OWN>
struct OwnerMaps {};
template <>
struct OwnerMaps<true> {
  CUtensorMap m[kMaxOwners];
};

__device__ __forceinline__ uint32_t ld_acquire_gpu_u32(const unsigned int* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu_u32(unsigned int* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ uint4 ld_shared_v4(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
// sigmoid(x) = 0.5 tanh(x / 2) + 0.5: ONE MUFU op (tanh.approx, max relative error 2^-11) and two FMA-pipe ops.  The
// SwiGLU epilogues are bound by their own instruction stream (two epilogue warps per scheduler), and the SFU pipe
// issues 16 results per clock per SM: ex2 + rcp cost twice the MUFU slots.
__device__ __forceinline__ float sigmoid_f(float x) {
  float t;
  asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(0.5f * x));
  return fmaf(t, 0.5f, 0.5f);
}

// tma_c: primary output.  tma_x: second tensor -- RESID_DROPOUT: residual (fp32, input); SWIGLU: act (bf16,
// output); SWIGLU_BWD: saved [gate|up] (bf16, input).
template <int BNP, bool A_MN, bool B_MN, int EPI, bool AG = false, bool OWN = false>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
gemm2_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
             const __grid_constant__ CUtensorMap tma_c, const __grid_constant__ CUtensorMap tma_x, const Params p,
             const __grid_constant__ OwnerMaps<OWN> om) {
  static_assert(!OWN || EPI == EPI_F32_ATOMIC, "owner-directed output needs the reduce-add epilogue");
  using C = Cfg<BNP, EPI>;
  constexpr int kStages = C::kStages;
  constexpr int NB = C::kEpiBufs;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages * kAStage;
  uint8_t* smem_epi = smem + C::kEpiOff;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + C::kBarOff);
  uint64_t* full_bar = bars;                      // [kStages]  leader's copy is waited on
  uint64_t* empty_bar = bars + kStages;           // [kStages]  per CTA, multicast commit
  uint64_t* tmem_full = bars + 2 * kStages;       // [2]        per CTA, multicast commit
  uint64_t* tmem_empty = bars + 2 * kStages + 2;  // [2]        leader's copy: 8 epilogue warps arrive
  uint64_t* ld_bar = bars + 2 * kStages + 4;      // [6]        epilogue input loads (per CTA)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 10);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta_rank = cluster_ctarank();
  const bool leader = cta_rank == 0;
  // AG: the first ag_pairs pairs of the grid are gather CTAs (pair < 0 below); the rest multiply
  const int ag_pairs = AG ? p.ag_pairs : 0;
  // epoch published into the per-block flags: a launch argument, or -- so that a captured launch can be replayed --
  // a word in device memory that the caller bumps (stream-ordered) before every launch
  uint32_t ag_epoch = 0;
  if constexpr (AG) ag_epoch = p.ag_epoch_ptr != nullptr ? *reinterpret_cast<const volatile unsigned int*>(p.ag_epoch_ptr) : p.ag_epoch;
  const int pair = int(blockIdx.x >> 1) - ag_pairs, npairs = int(gridDim.x >> 1) - ag_pairs;

  const int tiles_m = (p.M + 2 * BM - 1) / (2 * BM);
  const int tiles_n = (p.N + BNP - 1) / BNP;
  // n-tile of linear tile index r (m fastest).  AG rotates the n order so locally owned weight rows come first.
  // Rasterisation: m fastest by default (consecutive tiles share the B tile).  n_fast: the n-tiles of one row block
  // are consecutive instead, so they run at the same time on neighbouring pairs and A -- when A is the operand
  // that does not fit in L2 (LM-head dW: 416 MB of dlogits, three n-tiles) -- comes from HBM once, not per n-tile.
  auto mtile = [&](int r) { return p.n_fast ? r / tiles_n : r % tiles_m; };
  auto ntile = [&](int r) {
    int qn = p.n_fast ? r % tiles_n : r / tiles_m;
    if constexpr (AG) {
      qn += p.ag_rot;
      if (qn >= tiles_n) qn -= tiles_n;
    }
    return qn;
  };
  const int tiles_mn = tiles_m * tiles_n;
  const int num_tiles = tiles_mn * p.split_k;
  const int k_blocks_total = (p.K + BK - 1) / BK;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    tma_prefetch_desc(&tma_c);
    if constexpr (EPI >= EPI_RESID_DROPOUT) tma_prefetch_desc(&tma_x);
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], 2 * kEpiWarps);
    }
    for (int i = 0; i < 6; ++i) mbar_init(&ld_bar[i], 1);
    fence_barrier_init();
  }
  cluster_sync_all();   // barrier inits visible to the peer before any remote arrive / multicast commit
  if (warp == 1) tmem_alloc_cg2<C::kTmemCols>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (AG && pair < 0) {
    // ------------------------------------------------------------------ gather CTAs (fused all-gather)
    // One thread per CTA streams 128-row blocks of the weight from their owners' shards (peer memory over
    // NVLink, or the local shard) through a ring of 32 KB smem slots with bulk async copies
    // (peer -> smem -> local gathered buffer), then publishes the block's flag.  Blocks are taken round-robin
    // by the gather CTAs, starting at the locally owned rows -- the order the GEMM pairs consume them in.
    if constexpr (AG) if (threadIdx.x == 0) {
      constexpr int NS = 5;                       // ring depth: 5 x 32 KB (the GEMM's operand stages, unused here)
      constexpr long long CH = 16384;             // elements per chunk (32 KB)
      static_assert(NS * CH * 2 <= kStages * C::kStage, "gather ring does not fit in the operand stages");
      const int gi = blockIdx.x, GC = 2 * ag_pairs;
      const int nblocks = (p.N + 127) / 128;
      const int first = p.ag_rot * (BNP / 128);   // first row block the GEMM pairs will ask for
      const long long blk_elems = 128LL * p.K;
      const long long b_total = (long long)p.N * p.K;
      const __nv_bfloat16* const* srcs = reinterpret_cast<const __nv_bfloat16* const*>(p.ag_src);
      __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.ag_dst);
      uint32_t G = 0;                             // running chunk counter: ring slot and mbarrier phase
      for (int k = gi; k < nblocks; k += GC) {
        int nb = first + k;
        if (nb >= nblocks) nb -= nblocks;
        const long long b0 = nb * blk_elems;                               // element offset inside B
        const long long total = min(blk_elems, b_total - b0);
        const int n = int((total + CH - 1) / CH);
        const uint32_t Gbase = G;
        for (int i = 0; i < n + (NS - 2); ++i) {
          if (i < n) {
            if (i >= NS) tma_store_wait_read<1>();   // the store that last read this slot has finished with it
            const uint32_t slot = (Gbase + i) % NS;
            const long long off = b0 + i * CH, len = min(CH, total - i * CH);
            const long long u0 = p.ag_b_off + off;                           // element offset inside the unit
            const int owner = int(u0 / p.ag_part);
            const long long in_owner = min(len, (owner + 1) * p.ag_part - u0);
            uint64_t* bar = &full_bar[slot];
            mbar_expect_tx(bar, uint32_t(len * 2));
            const uint32_t sdst = smem_u32(smem_a + slot * (CH * 2));
            bulk_load_g2s(sdst, srcs[owner] + (u0 - owner * p.ag_part), uint32_t(in_owner * 2), bar);
            if (in_owner < len)   // chunk straddles two owners
              bulk_load_g2s(sdst + uint32_t(in_owner * 2), srcs[owner + 1], uint32_t((len - in_owner) * 2), bar);
          }
          const int d = i - (NS - 2);
          if (d >= 0 && d < n) {
            const uint32_t gd = Gbase + d, slot = gd % NS;
            mbar_wait(&full_bar[slot], (gd / NS) & 1);
            const long long off = b0 + d * CH, len = min(CH, total - d * CH);
            bulk_store_s2g(dst + off, smem_u32(smem_a + slot * (CH * 2)), uint32_t(len * 2));
            tma_store_commit();
          }
        }
        G = Gbase + n;
        tma_store_wait<0>();        // every byte of this block has been written
        __threadfence();
        st_release_gpu_u32(p.ag_flags + nb, ag_epoch);
      }
    }
  } else if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (both CTAs)
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = pair; t < num_tiles; t += npairs) {
        const int ks = t / tiles_mn;
        const int r = t - ks * tiles_mn;
        const int m0 = mtile(r) * 2 * BM + int(cta_rank) * BM;   // this CTA's rows
        const int nt = ntile(r) * BNP;
        // this CTA's half of the B tile; SwiGLU: leader stages gate rows, peer the matching up rows
        const int n0 = (EPI == EPI_SWIGLU) ? (cta_rank == 0 ? nt / 2 : p.half_n + nt / 2) : nt + int(cta_rank) * C::BNH;
        const int kb0 = ks * p.k_per_split;
        const int kb1 = min(k_blocks_total, kb0 + p.k_per_split);
        if constexpr (AG) {
          // the 128 weight rows this CTA is about to stage must have arrived from their owners
          if (n0 < p.N) {
            const unsigned int* flag = p.ag_flags + n0 / 128;
            const unsigned long long t0 = global_timer_ns();
            uint32_t spins = 0;
            while (ld_acquire_gpu_u32(flag) != ag_epoch) {
              if ((++spins & 0xFF) == 0 && global_timer_ns() - t0 > B200_WAIT_TIMEOUT_NS) {
                printf("b200: fused all-gather GEMM timed out waiting for weight rows %d (block %d)\n", n0, blockIdx.x);
                __trap();
              }
            }
            asm volatile("fence.proxy.async;" ::: "memory");   // bulk-store writes -> visible to this TMA read
          }
        }
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          if (leader) mbar_expect_tx(&full_bar[stage], 2 * C::kStage);   // bytes of both CTAs land on this barrier
          uint8_t* sa = smem_a + stage * kAStage;
          uint8_t* sb = smem_b + stage * C::kBStage;
          const int k0 = kb * BK;
          if constexpr (!A_MN) {
            tma_load_2d_cg2(sa, &tma_a, &full_bar[stage], k0, m0);
          } else {
#pragma unroll
            for (int c = 0; c < BM / 64; ++c) tma_load_2d_cg2(sa + c * (BK * 128), &tma_a, &full_bar[stage], m0 + c * 64, k0);
          }
          if constexpr (BNP == 384) {
            // Two MMAs cover the tile: N = 256 (accumulator columns 0..255) then N = 128 (256..383); each
            // takes half of its B rows from each CTA.  Staging chunks 0,1 from tile columns
            // [128 rank, +128) and chunk 2 from [256 + 64 rank, +64) makes accumulator column j == tile
            // column j.  A 64-row chunk is 8 KB in both operand layouts.
#pragma unroll
            for (int c = 0; c < 3; ++c) {
              const int nc = nt + (c < 2 ? int(cta_rank) * 128 + c * 64 : 256 + int(cta_rank) * 64);
              if constexpr (!B_MN) tma_load_2d_cg2(sb + c * 8192, &tma_b, &full_bar[stage], k0, nc);
              else                 tma_load_2d_cg2(sb + c * 8192, &tma_b, &full_bar[stage], nc, k0);
            }
          } else if constexpr (!B_MN) {
            tma_load_2d_cg2(sb, &tma_b, &full_bar[stage], k0, n0);
          } else {
#pragma unroll
            for (int c = 0; c < C::BNH / 64; ++c) tma_load_2d_cg2(sb + c * (BK * 128), &tma_b, &full_bar[stage], n0 + c * 64, k0);
          }
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA only)
    if (leader && lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(2 * BM, BNP == 384 ? 256 : BNP, A_MN, B_MN);
      constexpr uint32_t idesc_tail = make_idesc_bf16(2 * BM, 128, A_MN, B_MN);   // 384-wide tile: columns 256..383
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int t = pair; t < num_tiles; t += npairs, ++it) {
        const int ks = t / tiles_mn;
        const int kb0 = ks * p.k_per_split;
        const int kb1 = min(k_blocks_total, kb0 + p.k_per_split);
        const int buf = C::kAccBufs == 2 ? (it & 1) : 0;
        const uint32_t buf_phase = C::kAccBufs == 2 ? ((it >> 1) & 1) : (it & 1);
        mbar_wait(&tmem_empty[buf], buf_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * BNP;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem_a + stage * kAStage);
          const uint32_t b_addr = smem_u32(smem_b + stage * C::kBStage);
#pragma unroll
          for (int j = 0; j < BK / UMMA_K; ++j) {
            const uint64_t da = A_MN ? make_smem_desc_sw128(a_addr + j * (UMMA_K * 128), BK * 128, 1024)
                                     : make_smem_desc_sw128(a_addr + j * (UMMA_K * 2), 16, 1024);
            const uint64_t db = B_MN ? make_smem_desc_sw128(b_addr + j * (UMMA_K * 128), BK * 128, 1024)
                                     : make_smem_desc_sw128(b_addr + j * (UMMA_K * 2), 16, 1024);
            umma_bf16_ss_cg2(d_tmem, da, db, idesc, (kb > kb0 || j > 0) ? 1u : 0u);
            if constexpr (BNP == 384) {
              // third 64-row chunk of this CTA's B stage (16 KB in, either layout)
              const uint64_t db2 = B_MN ? make_smem_desc_sw128(b_addr + 16384 + j * (UMMA_K * 128), BK * 128, 1024)
                                        : make_smem_desc_sw128(b_addr + 16384 + j * (UMMA_K * 2), 16, 1024);
              umma_bf16_ss_cg2(d_tmem + 256, da, db2, idesc_tail, (kb > kb0 || j > 0) ? 1u : 0u);
            }
          }
          tc_commit_cg2_mc(&empty_bar[stage], 0b11);   // both CTAs may refill this slot
          if (++stage == kStages) { stage = 0; phase ^= 1; }
        }
        tc_commit_cg2_mc(&tmem_full[buf], 0b11);       // both CTAs' epilogues may drain their half
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps (both CTAs)
    // Eight warps: warps e and e+4 share a TMEM lane quarter (row r_in) and split every unit's columns in
    // halves, so each scheduler has two epilogue warps to overlap tcgen05.ld / smem / MUFU latencies.
    const int q = warp & 3;
    const int pt = (warp - 2) >> 2;                       // column part owned inside a unit (0 .. kParts-1)
    const int r_in = q * 32 + lane;                       // row within this CTA's 128 x BNP accumulator
    const int sw = r_in & 7;
    const bool epi_leader = threadIdx.x == 64;
    const uint32_t lane_off = uint32_t(q * 32) << 16;
    const uint32_t epi_base = smem_u32(smem_epi);
    const int my_tiles = (num_tiles - pair + npairs - 1) / npairs;   // tiles this pair processes

    // "unit" = the amount of output handled per staging step (128 bytes per row); units are numbered across
    // tiles so that the staging ring and the input prefetch run ahead over tile boundaries.
    constexpr int kUnitCols = (EPI == EPI_F32 || EPI == EPI_F32_ATOMIC || EPI == EPI_RESID_DROPOUT) ? 32 : 64;
    constexpr int kUnits = (EPI == EPI_SWIGLU) ? BNP / 128 : BNP / kUnitCols;   // per tile
    const long total_units = (long)my_tiles * kUnits;
    auto unit_coords = [&](long u, int& m0, int& n0, int& c) {
      const int it = int(u / kUnits);
      c = int(u - (long)it * kUnits);
      const int t = pair + it * npairs;
      const int r = t % tiles_mn;
      m0 = mtile(r) * 2 * BM + int(cta_rank) * BM;
      n0 = ntile(r) * BNP;
    };
    uint32_t g = 0;   // store-only staging ring position
    auto stage_begin = [&]() -> uint32_t {
      if (epi_leader) tma_store_wait_read<NB - 1>();   // the store that last read this slot is done with it
      named_bar_sync(1, kEpiThreads);
      return epi_base + (g % NB) * kEpiBuf;
    };
    auto stage_end = [&](const CUtensorMap* map, int col, int row, bool reduce) {
      fence_proxy_async_smem();
      named_bar_sync(1, kEpiThreads);
      if (epi_leader) {
        if (reduce) tma_reduce_add_2d(map, smem_epi + (g % NB) * kEpiBuf, col, row);
        else tma_store_2d(map, smem_epi + (g % NB) * kEpiBuf, col, row);
        tma_store_commit();
      }
      ++g;
    };
    // this thread's kCB bf16 columns (fp32 bits in v) -> chunks pt*kCh .. of its staging row
    auto put_bf16 = [&](uint32_t slot, const uint32_t* v) {
      const uint32_t row_addr = slot + r_in * 128;
#pragma unroll
      for (int cc = 0; cc < kCh; ++cc)
        st_shared_v4_u32(row_addr + (((pt * kCh + cc) ^ sw) << 4),
                         pack_bf16x2(__uint_as_float(v[cc * 8 + 0]), __uint_as_float(v[cc * 8 + 1])),
                         pack_bf16x2(__uint_as_float(v[cc * 8 + 2]), __uint_as_float(v[cc * 8 + 3])),
                         pack_bf16x2(__uint_as_float(v[cc * 8 + 4]), __uint_as_float(v[cc * 8 + 5])),
                         pack_bf16x2(__uint_as_float(v[cc * 8 + 6]), __uint_as_float(v[cc * 8 + 7])));
    };
    auto release_tmem = [&](int buf) {
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_leader(&tmem_empty[buf]);
    };

    if constexpr (EPI == EPI_RESID_DROPOUT) {
      if (epi_leader) {   // residual prefetch runs two units ahead
        for (long u = 0; u < 2 && u < total_units; ++u) {
          int m0, n0, c;
          unit_coords(u, m0, n0, c);
          mbar_expect_tx(&ld_bar[u % NB], kEpiBuf);
          tma_load_2d(smem_epi + (u % NB) * kEpiBuf, &tma_x, &ld_bar[u % NB], n0 + c * 32, m0);
        }
      }
    }
    if constexpr (OWN) {
      if (p.own_prev && epi_leader) {   // locally accumulated gradient: prefetch runs two units ahead
        for (long u = 0; u < 2 && u < total_units; ++u) {
          int m0, n0, c;
          unit_coords(u, m0, n0, c);
          mbar_expect_tx(&ld_bar[u % NB], kEpiBuf);
          tma_load_2d(smem_epi + (u % NB) * kEpiBuf, &tma_x, &ld_bar[u % NB], n0 + c * 32, m0);
        }
      }
    }
    if constexpr (EPI == EPI_SWIGLU_BWD) {
      if (epi_leader && total_units > 0) {   // [gate|up] prefetch runs one unit (a pair of slots) ahead
        int m0, n0, c;
        unit_coords(0, m0, n0, c);
        mbar_expect_tx(&ld_bar[0], 2 * kEpiBuf);
        tma_load_2d(smem_epi, &tma_x, &ld_bar[0], n0 + c * 64, m0);
        tma_load_2d(smem_epi + kEpiBuf, &tma_x, &ld_bar[0], p.half_n + n0 + c * 64, m0);
      }
    }

    const bool drop = p.dropout_p > 0.f;
    const uint32_t thresh = dropout_threshold16(p.dropout_p);
    const float inv_keep = drop ? 1.0f / (1.0f - p.dropout_p) : 1.0f;
    unsigned long long seed = 0, offs = 0;
    if (EPI == EPI_RESID_DROPOUT && drop) { seed = p.rng[0]; offs = p.rng[1]; }

    long u_glob = 0;
    for (int it = 0; it < my_tiles; ++it) {
      const int t = pair + it * npairs;
      const int r = t % tiles_mn;
      const int m0 = mtile(r) * 2 * BM + int(cta_rank) * BM;
      const int n0 = ntile(r) * BNP;
      const int row = m0 + r_in;
      const int buf = C::kAccBufs == 2 ? (it & 1) : 0;
      const uint32_t buf_phase = C::kAccBufs == 2 ? ((it >> 1) & 1) : (it & 1);
      mbar_wait(&tmem_full[buf], buf_phase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + lane_off + buf * BNP;

      if constexpr (EPI == EPI_BF16) {
#pragma unroll 1
        for (int c = 0; c < kUnits; ++c) {
          const uint32_t slot = stage_begin();
          uint32_t v[kCB];
          tmem_ld_cols<kCB>(t_row + c * 64 + pt * kCB, v);
          tmem_ld_wait();
          if (c == kUnits - 1) release_tmem(buf);
          put_bf16(slot, v);
          stage_end(&tma_c, n0 + c * 64, m0, false);
        }
      } else if constexpr (EPI == EPI_ROPE_BF16) {
        // one unit = one head; column j (< 32) pairs with j + 32, so both halves are read and each thread
        // produces its own half.  cos/sin depend on the row only: cached in registers for the whole tile.
        constexpr int kHalfParts = kParts / 2;
        const int hcol = (pt % kHalfParts) * kCB;     // this thread's first column inside its half of the head
        float cs[kCB], sn[kCB];
        {
          const int pos = (row < p.M ? row : 0) % p.seq_len;
          const float4* c4 = reinterpret_cast<const float4*>(p.rope_cos + size_t(pos) * 32 + hcol);
          const float4* s4 = reinterpret_cast<const float4*>(p.rope_sin + size_t(pos) * 32 + hcol);
#pragma unroll
          for (int i = 0; i < kCB / 4; ++i) {
            const float4 a = c4[i], b = s4[i];
            cs[4 * i] = a.x; cs[4 * i + 1] = a.y; cs[4 * i + 2] = a.z; cs[4 * i + 3] = a.w;
            sn[4 * i] = b.x; sn[4 * i + 1] = b.y; sn[4 * i + 2] = b.z; sn[4 * i + 3] = b.w;
          }
        }
#pragma unroll 1
        for (int c = 0; c < kUnits; ++c) {
          const uint32_t slot = stage_begin();
          uint32_t v[kCB];
          if (n0 + c * 64 < p.rope_cols) {
            uint32_t w[kCB];
            tmem_ld_cols<kCB>(t_row + c * 64 + pt * kCB, v);                  // own columns
            tmem_ld_cols<kCB>(t_row + c * 64 + (pt ^ kHalfParts) * kCB, w);   // partner columns (j <-> j + 32)
            tmem_ld_wait();
            const float sgn = pt < kHalfParts ? -1.0f : 1.0f;   // first half: x1*c - x2*s ; second half: x2*c + x1*s
#pragma unroll
            for (int e = 0; e < kCB; ++e)
              v[e] = __float_as_uint(__uint_as_float(v[e]) * cs[e] + sgn * __uint_as_float(w[e]) * sn[e]);
          } else {
            tmem_ld_cols<kCB>(t_row + c * 64 + pt * kCB, v);
            tmem_ld_wait();
          }
          if (c == kUnits - 1) release_tmem(buf);
          put_bf16(slot, v);
          stage_end(&tma_c, n0 + c * 64, m0, false);
        }
      } else if constexpr (EPI == EPI_F32 || EPI == EPI_F32_ATOMIC) {
        if constexpr (OWN) {
          // GEMM -> reduce-scatter in one kernel: this CTA's 128 rows are one slab of C; its owner's map (peer
          // memory over NVLink unless the owner is this rank) takes the reduce-add.  Rows beyond M belong to
          // nobody (the per-owner maps clip the last slab).  With own_prev the locally accumulated gradient of the
          // earlier micro-steps (tma_x) is prefetched into the staging slot and travels in the same reduce-add:
          // owner += acc + prev (added by the first K split only).
          const bool own_valid = m0 < p.M;
          const int own_s = m0 >> 7;
          const int own_op = own_valid ? (own_s * p.own_world) / p.own_slabs : 0;   // owner before rotation
          int own_o = own_op + p.own_rot;
          if (own_o >= p.own_world) own_o -= p.own_world;
          const int own_row = m0 - (((own_op * p.own_slabs + p.own_world - 1) / p.own_world) << 7);
          const bool add_prev = p.own_prev != 0 && t < tiles_mn;                    // K split 0
#pragma unroll 1
          for (int c = 0; c < kUnits; ++c, ++u_glob) {
            uint32_t slot;
            if (p.own_prev) {
              if (epi_leader) {   // prev prefetch two units ahead; that slot was last read by the reduce of unit u-2
                tma_store_wait_read<1>();
                const long un = u_glob + 2;
                if (un < total_units) {
                  int pm0, pn0, pc;
                  unit_coords(un, pm0, pn0, pc);
                  mbar_expect_tx(&ld_bar[un % NB], kEpiBuf);
                  tma_load_2d(smem_epi + (un % NB) * kEpiBuf, &tma_x, &ld_bar[un % NB], pn0 + pc * 32, pm0);
                }
              }
              slot = epi_base + uint32_t(u_glob % NB) * kEpiBuf;
            } else {
              slot = stage_begin();
            }
            uint32_t v[kCF];
            tmem_ld_cols<kCF>(t_row + c * 32 + pt * kCF, v);
            tmem_ld_wait();
            if (c == kUnits - 1) release_tmem(buf);
            const uint32_t row_addr = slot + r_in * 128;
            if (p.own_prev) mbar_wait(&ld_bar[u_glob % NB], uint32_t(u_glob / NB) & 1);
#pragma unroll
            for (int cc = 0; cc < kCh; ++cc) {
              const uint32_t addr = row_addr + (((pt * kCh + cc) ^ sw) << 4);
              if (add_prev) {
                const uint4 q4 = ld_shared_v4(addr);
                st_shared_v4_u32(addr, __float_as_uint(__uint_as_float(q4.x) + __uint_as_float(v[cc * 4 + 0])),
                                 __float_as_uint(__uint_as_float(q4.y) + __uint_as_float(v[cc * 4 + 1])),
                                 __float_as_uint(__uint_as_float(q4.z) + __uint_as_float(v[cc * 4 + 2])),
                                 __float_as_uint(__uint_as_float(q4.w) + __uint_as_float(v[cc * 4 + 3])));
              } else {
                st_shared_v4_u32(addr, v[cc * 4 + 0], v[cc * 4 + 1], v[cc * 4 + 2], v[cc * 4 + 3]);
              }
            }
            fence_proxy_async_smem();
            named_bar_sync(1, kEpiThreads);
            if (epi_leader) {
              if (own_valid)
                tma_reduce_add_2d(&om.m[own_o], reinterpret_cast<void*>(smem_epi + (p.own_prev ? (u_glob % NB) : (g % NB)) * kEpiBuf),
                                  n0 + c * 32, own_row);
              tma_store_commit();   // one (possibly empty) group per unit: the slot-reuse waits count groups
            }
            ++g;
          }
        } else {
#pragma unroll 1
          for (int c = 0; c < kUnits; ++c) {
            const uint32_t slot = stage_begin();
            uint32_t v[kCF];
            tmem_ld_cols<kCF>(t_row + c * 32 + pt * kCF, v);
            tmem_ld_wait();
            if (c == kUnits - 1) release_tmem(buf);
            const uint32_t row_addr = slot + r_in * 128;
#pragma unroll
            for (int cc = 0; cc < kCh; ++cc)
              st_shared_v4_u32(row_addr + (((pt * kCh + cc) ^ sw) << 4), v[cc * 4 + 0], v[cc * 4 + 1], v[cc * 4 + 2], v[cc * 4 + 3]);
            stage_end(&tma_c, n0 + c * 32, m0, EPI == EPI_F32_ATOMIC);
          }
        }
      } else if constexpr (EPI == EPI_RESID_DROPOUT) {
#pragma unroll 1
        for (int c = 0; c < kUnits; ++c, ++u_glob) {
          // keep the residual prefetch two units ahead; its slot was last read by the store of unit u-2
          if (epi_leader) {
            tma_store_wait_read<1>();
            const long un = u_glob + 2;
            if (un < total_units) {
              int pm0, pn0, pc;
              unit_coords(un, pm0, pn0, pc);
              mbar_expect_tx(&ld_bar[un % NB], kEpiBuf);
              tma_load_2d(smem_epi + (un % NB) * kEpiBuf, &tma_x, &ld_bar[un % NB], pn0 + pc * 32, pm0);
            }
          }
          const uint32_t slot = epi_base + uint32_t(u_glob % NB) * kEpiBuf;
          uint32_t v[kCF];
          tmem_ld_cols<kCF>(t_row + c * 32 + pt * kCF, v);
          tmem_ld_wait();
          if (c == kUnits - 1) release_tmem(buf);
          mbar_wait(&ld_bar[u_glob % NB], uint32_t(u_glob / NB) & 1);
          const uint32_t row_addr = slot + r_in * 128;
          const int col0 = n0 + c * 32;
#pragma unroll
          for (int g8 = 0; g8 < kCF / 8; ++g8) {
            float a8[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) a8[e] = __uint_as_float(v[g8 * 8 + e]) * inv_keep;
            if (drop)   // compare + select per element straight off the Philox words (no bit-mask round trip)
              dropout_zero8(dropout_bits(seed, offs, p.rng_stream, (unsigned long long)row * (p.N / 8) + (col0 / 8 + pt * (kCF / 8) + g8)),
                            thresh << 16, a8);
#pragma unroll
            for (int hlf = 0; hlf < 2; ++hlf) {
              const int cl = g8 * 2 + hlf;                 // local 16-byte chunk (4 floats) of this thread's columns
              const uint32_t addr = row_addr + (((pt * kCh + cl) ^ sw) << 4);
              const uint4 rq = ld_shared_v4(addr);
              st_shared_v4_u32(addr, __float_as_uint(__uint_as_float(rq.x) + a8[hlf * 4 + 0]),
                               __float_as_uint(__uint_as_float(rq.y) + a8[hlf * 4 + 1]),
                               __float_as_uint(__uint_as_float(rq.z) + a8[hlf * 4 + 2]),
                               __float_as_uint(__uint_as_float(rq.w) + a8[hlf * 4 + 3]));
            }
          }
          fence_proxy_async_smem();
          named_bar_sync(1, kEpiThreads);
          if (epi_leader) {
            tma_store_2d(&tma_c, smem_epi + (u_glob % NB) * kEpiBuf, col0, m0);
            tma_store_commit();
          }
        }
      } else if constexpr (EPI == EPI_SWIGLU) {
        // accumulator columns [0, BNP/2) = gate, [BNP/2, BNP) = up, for intermediate columns n0/2 ...
#pragma unroll 1
        for (int c = 0; c < kUnits; ++c) {
          uint32_t gv[kCB], uv[kCB];
          tmem_ld_cols<kCB>(t_row + c * 64 + pt * kCB, gv);
          tmem_ld_cols<kCB>(t_row + BNP / 2 + c * 64 + pt * kCB, uv);
          tmem_ld_wait();
          if (c == kUnits - 1) release_tmem(buf);
          const int icol = n0 / 2 + c * 64;   // intermediate column of this unit
          uint32_t slot = stage_begin();
          put_bf16(slot, gv);
          stage_end(&tma_c, icol, m0, false);
          slot = stage_begin();
          put_bf16(slot, uv);
          stage_end(&tma_c, p.half_n + icol, m0, false);
          // activation from the bf16-rounded gate / up so that backward recomputation is bit-identical
#pragma unroll
          for (int e = 0; e < kCB; ++e) {
            const float ga = __bfloat162float(__float2bfloat16_rn(__uint_as_float(gv[e])));
            const float ua = __bfloat162float(__float2bfloat16_rn(__uint_as_float(uv[e])));
            gv[e] = __float_as_uint(ga * sigmoid_f(ga) * ua);
          }
          slot = stage_begin();
          put_bf16(slot, gv);
          stage_end(&tma_x, icol, m0, false);
        }
      } else {  // EPI_SWIGLU_BWD: acc = d_act; inputs gate/up via TMA; outputs [dgate | dup]
#pragma unroll 1
        for (int c = 0; c < kUnits; ++c, ++u_glob) {
          const int pr = int(u_glob % 3);
          if (epi_leader) {
            // the pair of slots unit u+1 loads into was last read by the stores of unit u-2
            tma_store_wait_read<1>();
            const long un = u_glob + 1;
            if (un < total_units) {
              int pm0, pn0, pc;
              unit_coords(un, pm0, pn0, pc);
              const int npr = int(un % 3);
              mbar_expect_tx(&ld_bar[npr], 2 * kEpiBuf);
              tma_load_2d(smem_epi + (2 * npr) * kEpiBuf, &tma_x, &ld_bar[npr], pn0 + pc * 64, pm0);
              tma_load_2d(smem_epi + (2 * npr + 1) * kEpiBuf, &tma_x, &ld_bar[npr], p.half_n + pn0 + pc * 64, pm0);
            }
          }
          uint32_t d[kCB];
          tmem_ld_cols<kCB>(t_row + c * 64 + pt * kCB, d);
          tmem_ld_wait();
          if (c == kUnits - 1) release_tmem(buf);
          mbar_wait(&ld_bar[pr], uint32_t(u_glob / 3) & 1);
          const uint32_t g_row = epi_base + (2 * pr) * kEpiBuf + r_in * 128;
          const uint32_t u_row = g_row + kEpiBuf;
#pragma unroll
          for (int cc = 0; cc < kCh; ++cc) {
            const uint32_t off = (((pt * kCh + cc) ^ sw) << 4);
            const uint4 gq = ld_shared_v4(g_row + off), uq = ld_shared_v4(u_row + off);
            const uint32_t gw[4] = {gq.x, gq.y, gq.z, gq.w}, uw[4] = {uq.x, uq.y, uq.z, uq.w};
            uint32_t og[4], ou[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float2 gg = unpack_bf16x2(gw[e]), uu = unpack_bf16x2(uw[e]);
              const float da = __uint_as_float(d[cc * 8 + 2 * e]), db = __uint_as_float(d[cc * 8 + 2 * e + 1]);
              const float s0 = sigmoid_f(gg.x), s1 = sigmoid_f(gg.y);
              const float t0 = da * s0, t1 = db * s1;                       // d_act * sigmoid(gate)
              // d silu(g)/dg = s (1 + g (1 - s));  1 + g (1 - s) = fma(g, 1 - s, 1)
              og[e] = pack_bf16x2(t0 * uu.x * fmaf(gg.x, 1.0f - s0, 1.0f), t1 * uu.y * fmaf(gg.y, 1.0f - s1, 1.0f));
              ou[e] = pack_bf16x2(t0 * gg.x, t1 * gg.y);
            }
            st_shared_v4_u32(g_row + off, og[0], og[1], og[2], og[3]);
            st_shared_v4_u32(u_row + off, ou[0], ou[1], ou[2], ou[3]);
          }
          fence_proxy_async_smem();
          named_bar_sync(1, kEpiThreads);
          // C = [dgate | dup] is 2 * half_n wide, so a unit past column half_n (last tile of an intermediate size
          // that is not a multiple of the tile width) is NOT clipped by the tensor map: its dgate store would
          // land on the first dup columns.  Such units are dropped here (their loads were harmless).
          if (epi_leader) {
            if (n0 + c * 64 < p.N) {
              tma_store_2d(&tma_c, smem_epi + (2 * pr) * kEpiBuf, n0 + c * 64, m0);
              tma_store_2d(&tma_c, smem_epi + (2 * pr + 1) * kEpiBuf, p.half_n + n0 + c * 64, m0);
            }
            tma_store_commit();   // one (possibly empty) group per unit: the slot-reuse waits count groups
          }
        }
      }
    }
    if (epi_leader) tma_store_wait<0>();   // all bulk stores complete before the CTA (and its smem) goes away
  }

  tc_fence_before();
  cluster_sync_all();   // nobody exits (or frees TMEM) while the peer may still signal it
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_cg2<C::kTmemCols>(tmem_base);
  }
}

// cudaFuncSetAttribute is per device: remember which devices have this instantiation configured.
template <int BNP, bool A_MN, bool B_MN, int EPI, bool AG = false, bool OWN = false>
static int launch(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const CUtensorMap& tx,
                  const Params& p, int grid, cudaStream_t stream, const OwnerMaps<OWN>& om = OwnerMaps<OWN>{}) {
  auto kfn = gemm2_kernel<BNP, A_MN, B_MN, EPI, AG, OWN>;
  static bool configured[kMaxDevices] = {};
  const int dev = current_device();
  if (!configured[dev]) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BNP, EPI>::kSmemTotal));
    configured[dev] = true;
  }
  kfn<<<grid, kThreads, Cfg<BNP, EPI>::kSmemTotal, stream>>>(ta, tb, tc, tx, p, om);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

template <int BNP, int EPI>
static int dispatch(bool a_mn, bool b_mn, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc,
                    const CUtensorMap& tx, const Params& p, int grid, cudaStream_t s) {
  if (!a_mn && !b_mn) return launch<BNP, false, false, EPI>(ta, tb, tc, tx, p, grid, s);
  if (!a_mn && b_mn) return launch<BNP, false, true, EPI>(ta, tb, tc, tx, p, grid, s);
  if constexpr (EPI <= EPI_F32_ATOMIC) {
    if (a_mn && b_mn) return launch<BNP, true, true, EPI>(ta, tb, tc, tx, p, grid, s);
    return launch<BNP, true, false, EPI>(ta, tb, tc, tx, p, grid, s);
  }
  set_error("gemm2: MN-major A is not instantiated for epilogue %d", EPI);
  return -2;
}

template <int BNP>
static int dispatch_epi(int epi, bool a_mn, bool b_mn, const CUtensorMap& ta, const CUtensorMap& tb,
                        const CUtensorMap& tc, const CUtensorMap& tx, const Params& p, int grid, cudaStream_t s) {
  switch (epi) {
    case EPI_BF16: return dispatch<BNP, EPI_BF16>(a_mn, b_mn, ta, tb, tc, tx, p, grid, s);
    case EPI_F32: return dispatch<BNP, EPI_F32>(a_mn, b_mn, ta, tb, tc, tx, p, grid, s);
    case EPI_F32_ATOMIC: return dispatch<BNP, EPI_F32_ATOMIC>(a_mn, b_mn, ta, tb, tc, tx, p, grid, s);
    case EPI_ROPE_BF16: return dispatch<BNP, EPI_ROPE_BF16>(a_mn, b_mn, ta, tb, tc, tx, p, grid, s);
    case EPI_RESID_DROPOUT: return dispatch<BNP, EPI_RESID_DROPOUT>(a_mn, b_mn, ta, tb, tc, tx, p, grid, s);
    case EPI_SWIGLU_BWD: return dispatch<BNP, EPI_SWIGLU_BWD>(a_mn, b_mn, ta, tb, tc, tx, p, grid, s);
  }
  set_error("gemm2: unknown epilogue %d", epi);
  return -2;
}

}  // namespace g2
}  // namespace b200

using namespace b200;

static int run_gemm2(const b200_gemm_args* a, const b200_ag_args* ag, const b200_rs_args* rs, void* stream_);

extern "C" int b200_gemm2_bf16(const b200_gemm_args* a, void* stream_) { return run_gemm2(a, nullptr, nullptr, stream_); }

extern "C" int b200_gemm2_ag_bf16(const b200_gemm_args* a, const b200_ag_args* ag, void* stream_) {
  B200_REQUIRE(ag != nullptr, "gemm2_ag: null all-gather arguments");
  return run_gemm2(a, ag, nullptr, stream_);
}

extern "C" int b200_gemm2_rs_bf16(const b200_gemm_args* a, const b200_rs_args* rs, void* stream_) {
  B200_REQUIRE(rs != nullptr, "gemm2_rs: null reduce-scatter arguments");
  return run_gemm2(a, nullptr, rs, stream_);
}

extern "C" int b200_slab_owner(int slab, int slabs, int world, int rot) {
  if (slabs <= 0 || world <= 0 || slab < 0 || slab >= slabs) return -1;
  return int(((long long)slab * world / slabs + rot) % world);
}

extern "C" int b200_owner_slab_range(int owner, int slabs, int world, int rot, int* first, int* last) {
  if (slabs <= 0 || world <= 0 || owner < 0 || owner >= world) return -1;
  const long long op = ((owner - rot) % world + world) % world;
  *first = int((op * slabs + world - 1) / world);
  *last = int(((op + 1) * slabs + world - 1) / world);
  return 0;
}

static int run_gemm2(const b200_gemm_args* a, const b200_ag_args* ag, const b200_rs_args* rs, void* stream_) {
  using namespace b200::g2;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  B200_REQUIRE(a != nullptr, "gemm2: null args");
  B200_REQUIRE(a->M > 0 && a->N > 0 && a->K > 0, "gemm2: empty problem %dx%dx%d", a->M, a->N, a->K);
  B200_REQUIRE(a->lda % 8 == 0 && a->ldb % 8 == 0, "gemm2: lda/ldb must be multiples of 8 elements");
  const int epi = a->epilogue;
  B200_REQUIRE(epi >= EPI_BF16 && epi <= EPI_SWIGLU_BWD, "gemm2: unknown epilogue %d", epi);
  const bool a_mn = a->a_mn_major != 0, b_mn = a->b_mn_major != 0;
  const bool c_f32 = epi == EPI_F32 || epi == EPI_F32_ATOMIC || epi == EPI_RESID_DROPOUT;
  const int c_elem = c_f32 ? 4 : 2;
  B200_REQUIRE((size_t(a->ldc) * c_elem) % 16 == 0, "gemm2: C row pitch must be a multiple of 16 bytes");
  if (epi == EPI_ROPE_BF16) B200_REQUIRE(a->rope_cols % 64 == 0 && a->rope_cos && a->rope_sin && a->seq_len > 0, "gemm2: bad rotary arguments");
  if (epi == EPI_RESID_DROPOUT) {
    B200_REQUIRE(a->residual != nullptr, "gemm2: residual required");
    B200_REQUIRE(a->dropout_p == 0.f || a->rng_state != nullptr, "gemm2: rng_state required when dropout_p > 0");
    B200_REQUIRE(a->N % 8 == 0, "gemm2: N %% 8 != 0");
  }
  if (epi == EPI_SWIGLU) B200_REQUIRE(a->N == 2 * a->half_n && a->half_n % 128 == 0 && a->aux != nullptr && !b_mn, "gemm2: SwiGLU needs N == 2*half_n, half_n %% 128 == 0, K-major B and an act output");
  if (epi == EPI_SWIGLU_BWD) B200_REQUIRE(a->N == a->half_n && a->aux != nullptr, "gemm2: SwiGLU' needs N == half_n and the saved [gate|up]");

  // tile width: 256 unless the narrower pair tile fills the machine better (N = 768/1024 with few row tiles)
  // 256-wide pair tile unless the caller forces 128 (measured slower on every model shape, kept for narrow N)
  int bnp = (a->tile_n == 128 || a->N < 256) ? 128 : 256;
  if (epi == EPI_SWIGLU) bnp = 256;
  // 384-wide tile when it needs fewer column-waves than the 256-wide one (N = 768: 64 tiles in one wave
  // instead of 96 tiles in two), for the epilogues it is instantiated with
  if (bnp == 256 && ag == nullptr && !a_mn && a->N % 384 == 0 && a->split_k <= 1 &&
      (epi == EPI_BF16 || (epi == EPI_RESID_DROPOUT && !b_mn)) && a->tile_n != 256) {
    int hw = sm_count();
    if (a->max_ctas > 0 && a->max_ctas < hw) hw = a->max_ctas;
    const long hw_pairs = hw / 2 > 0 ? hw / 2 : 1;
    const long tm = (a->M + 2 * BM - 1) / (2 * BM);
    const long cost256 = ((tm * ((a->N + 255) / 256) + hw_pairs - 1) / hw_pairs) * 256;
    const long cost384 = ((tm * (a->N / 384) + hw_pairs - 1) / hw_pairs) * 384;
    if (a->tile_n == 384 || cost384 < cost256) bnp = 384;
  }

  CUtensorMap ta, tb, tc, tx;
  int rc;
  if (!a_mn) rc = make_tmap_2d_bf16(&ta, a->A, a->K, a->M, a->lda, 64, BM);
  else       rc = make_tmap_2d_bf16(&ta, a->A, a->M, a->K, a->lda, 64, BK);
  if (rc) return rc;
  if (!b_mn) rc = make_tmap_2d_bf16(&tb, a->B, a->K, a->N, a->ldb, 64, bnp == 384 ? 64 : bnp / 2);   // 384: three 64-row chunks
  else       rc = make_tmap_2d_bf16(&tb, a->B, a->N, a->K, a->ldb, 64, BK);
  if (rc) return rc;
  const int c_cols = epi == EPI_SWIGLU_BWD ? 2 * a->half_n : a->N;   // SwiGLU': C = [dgate | dup] is 2*half_n wide
  const void* c_base = a->C;
  if (rs != nullptr && c_base == nullptr && rs->owner_base != nullptr)   // owner-directed output: C itself is optional
    for (int o = 0; o < rs->world && c_base == nullptr; ++o) c_base = rs->owner_base[o];
  rc = make_tmap_2d_sw128(&tc, c_base, c_elem, c_cols, a->M, a->ldc, 128);
  if (rc) return rc;
  tx = tc;
  if (epi == EPI_RESID_DROPOUT) rc = make_tmap_2d_sw128(&tx, a->residual, 4, a->N, a->M, a->ldc, 128);
  else if (epi == EPI_SWIGLU) rc = make_tmap_2d_sw128(&tx, a->aux, 2, a->half_n, a->M, a->ld_aux, 128);
  else if (epi == EPI_SWIGLU_BWD) rc = make_tmap_2d_sw128(&tx, a->aux, 2, 2 * a->half_n, a->M, a->ld_aux, 128);
  if (rc) return rc;

  Params p{};
  p.M = a->M; p.N = a->N; p.K = a->K;
  const int kblocks = (a->K + BK - 1) / BK;
  int split = a->split_k > 0 ? a->split_k : 1;
  if (split > kblocks) split = kblocks;
  if (split > 1) B200_REQUIRE(epi == EPI_F32_ATOMIC, "gemm2: split-K needs the fp32 reduce-add epilogue");
  p.k_per_split = (kblocks + split - 1) / split;
  p.split_k = (kblocks + p.k_per_split - 1) / p.k_per_split;
  p.rope_cos = a->rope_cos; p.rope_sin = a->rope_sin; p.seq_len = a->seq_len > 0 ? a->seq_len : 1; p.rope_cols = a->rope_cols;
  p.dropout_p = a->dropout_p; p.rng = a->rng_state; p.rng_stream = a->rng_stream; p.half_n = a->half_n;

  {
    // n-fastest order when A is too big for L2 (126 MB) while B fits easily and there are few n-tiles per row block
    const double a_bytes = 2.0 * a->M * a->K, b_bytes = 2.0 * a->N * a->K;
    const int tn = (a->N + bnp - 1) / bnp;
    p.n_fast = (ag == nullptr && a_bytes > 96e6 && b_bytes < 48e6 && tn > 1 && tn <= 8) ? 1 : 0;
    if (a->raster == 1) p.n_fast = 0;
    if (a->raster == 2 && ag == nullptr) p.n_fast = 1;
  }
  const long tiles = long((a->M + 2 * BM - 1) / (2 * BM)) * ((a->N + bnp - 1) / bnp) * p.split_k;
  int sms = sm_count();
  if (a->max_ctas > 0 && a->max_ctas < sms) sms = a->max_ctas;
  long pairs = sms / 2;
  if (pairs < 1) pairs = 1;
  if (tiles < pairs) pairs = tiles;
  if (ag != nullptr) {
    // all-gather fused into the GEMM: some pairs gather the weight rows over NVLink, the others multiply
    B200_REQUIRE(epi == EPI_BF16 || epi == EPI_ROPE_BF16, "gemm2_ag: epilogue %d is not instantiated with the fused gather", epi);
    B200_REQUIRE(!a_mn && !b_mn && bnp == 256 && p.split_k == 1, "gemm2_ag: needs K-major operands, the 256-wide tile, no split-K");
    B200_REQUIRE(a->ldb == a->K, "gemm2_ag: the gathered B must be contiguous (ldb == K)");
    B200_REQUIRE(ag->world >= 1 && ag->world <= 16 && ag->rank >= 0 && ag->rank < ag->world, "gemm2_ag: bad rank/world");
    B200_REQUIRE(ag->part_elems >= 16384 && ag->part_elems % 8 == 0 && ag->b_offset_elems % 8 == 0,
                 "gemm2_ag: shard size must be >= 16384 elements and 16-byte aligned");
    B200_REQUIRE(ag->b_offset_elems + (long long)a->N * a->K <= ag->part_elems * ag->world, "gemm2_ag: B lies outside the unit");
    B200_REQUIRE(ag->flags != nullptr && ag->shard_ptrs != nullptr, "gemm2_ag: flags / shard pointers missing");
    int gp = ag->gather_pairs > 0 ? ag->gather_pairs : 4;
    const int hw_pairs = sms / 2;
    if (gp > hw_pairs - 1) gp = hw_pairs - 1;
    B200_REQUIRE(gp >= 1, "gemm2_ag: not enough SMs for a gather pair");
    long gemm_pairs = hw_pairs - gp;
    if (tiles < gemm_pairs) gemm_pairs = tiles;
    for (int i = 0; i < 16; ++i) p.ag_src[i] = i < ag->world ? ag->shard_ptrs[i] : nullptr;
    p.ag_dst = const_cast<void*>(a->B);
    p.ag_part = ag->part_elems; p.ag_b_off = ag->b_offset_elems; p.ag_flags = ag->flags; p.ag_epoch = ag->epoch; p.ag_epoch_ptr = ag->epoch_dev;
    p.ag_pairs = gp; p.ag_rank = ag->rank; p.ag_world = ag->world;
    // first n-tile that contains a locally owned weight row
    long long first_local = ag->rank * ag->part_elems - ag->b_offset_elems;
    if (first_local < 0) first_local = 0;
    const int tiles_n = (a->N + 255) / 256;
    int rot = int((first_local / a->K) / 256);
    if (rot >= tiles_n) rot = 0;
    p.ag_rot = rot;
    const int ag_grid = int((gp + gemm_pairs) * 2);
    if (epi == EPI_BF16) return launch<256, false, false, EPI_BF16, true>(ta, tb, tc, tx, p, ag_grid, stream);
    return launch<256, false, false, EPI_ROPE_BF16, true>(ta, tb, tc, tx, p, ag_grid, stream);
  }
  const int grid = int(pairs * 2);
  if (rs != nullptr) {
    // GEMM -> reduce-scatter: every 128-row slab of C is reduce-added into its owner's memory
    B200_REQUIRE(epi == EPI_F32_ATOMIC && a_mn && b_mn && (bnp == 256 || bnp == 128),
                 "gemm2_rs: instantiated for the wgrad form only (MN-major A and B, fp32 reduce-add epilogue)");
    B200_REQUIRE(rs->world >= 1 && rs->world <= kMaxOwners && rs->owner_base != nullptr, "gemm2_rs: world must be in [1,%d]", kMaxOwners);
    const int slabs = (a->M + 127) / 128;
    const int rot = ((rs->rot % rs->world) + rs->world) % rs->world;
    p.own_world = rs->world; p.own_slabs = slabs; p.own_rot = rot;
    p.own_prev = rs->prev != nullptr;
    if (rs->prev != nullptr) {
      rc = make_tmap_2d_sw128(&tx, rs->prev, 4, a->N, a->M, a->ldc, 128);
      if (rc) return rc;
    }
    OwnerMaps<true> om;
    for (int o = 0; o < kMaxOwners; ++o) {
      om.m[o] = tc;   // owners without rows keep a valid (never used) descriptor
      if (o >= rs->world) continue;
      int first = 0, last = 0;
      b200_owner_slab_range(o, slabs, rs->world, rot, &first, &last);
      if (last <= first) continue;
      const long long rows = (long long)(last * 128 < a->M ? last * 128 : a->M) - (long long)first * 128;
      B200_REQUIRE(rs->owner_base[o] != nullptr, "gemm2_rs: owner %d owns rows but has no base pointer", o);
      rc = make_tmap_2d_sw128(&om.m[o], rs->owner_base[o], 4, a->N, (uint64_t)rows, a->ldc, 128);
      if (rc) return rc;
    }
    if (bnp == 256) return launch<256, true, true, EPI_F32_ATOMIC, false, true>(ta, tb, tc, tx, p, grid, stream, om);
    return launch<128, true, true, EPI_F32_ATOMIC, false, true>(ta, tb, tc, tx, p, grid, stream, om);
  }
  if (epi == EPI_SWIGLU) return dispatch<256, EPI_SWIGLU>(a_mn, b_mn, ta, tb, tc, tx, p, grid, stream);
  if (bnp == 384) {
    if (epi == EPI_RESID_DROPOUT) return launch<384, false, false, EPI_RESID_DROPOUT>(ta, tb, tc, tx, p, grid, stream);
    if (b_mn) return launch<384, false, true, EPI_BF16>(ta, tb, tc, tx, p, grid, stream);
    return launch<384, false, false, EPI_BF16>(ta, tb, tc, tx, p, grid, stream);
  }
  if (bnp == 256) return dispatch_epi<256>(epi, a_mn, b_mn, ta, tb, tc, tx, p, grid, stream);
  return dispatch_epi<128>(epi, a_mn, b_mn, ta, tb, tc, tx, p, grid, stream);
}
