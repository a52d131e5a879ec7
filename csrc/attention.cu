// Causal flash attention forward + backward on tcgen05 (head_dim 64), with Philox dropout on the
// probabilities.  Replaces the reference's manual O(S^2) path (gpt.py:230-234: QK^T bmm -> triu mask
// -> fp32 softmax -> dropout -> @V, which materialises [B,H,S,S]) and its SDPA path (gpt.py:199-206).
//
// Data layout: packed qkv [B*S, 3h] bf16 straight out of the fused QKV GEMM; one TMA tensor map over
// it serves Q, K and V tiles (column offsets hd*64, h + hd*64, 2h + hd*64).  Output is written in
// [B, S, H, D] order, i.e. the [B*S, h] matrix o_proj consumes.
//
// Both kernels are persistent: one CTA per SM walks a boustrophedon list of work items ordered heaviest
// first (snake_tile), so barriers / TMEM / tensor maps are set up once and every ring index and mbarrier
// phase derives from running counters.
//
// Forward work item = (q block of 128 rows, head, batch); CTA = 18 warps:
//   warp 0     TMA producer: Q (double buffered across items), K/V tiles through 2-stage rings
//   warp 1     MMA issuer:   S_j = Q K_j^T  (TMEM, double buffered)  and  O_j = P_j V_j (TMEM, double buffered)
//   warps 2-17 softmax: four threads per query row, each owning 32 of the 128 score columns (four warps
//              per scheduler: measured, one or two cannot hide the fixed-latency dependency stalls).  tcgen05.ld S ->
//              online softmax in the exp2 domain on RAW scores (scale folded into the exponent FFMA2, one
//              MUFU.EX2 per element, row max exchanged once per tile through smem) -> dropout as compare +
//              select on the Philox words -> bf16 P tile written to 128B-swizzled smem (A operand of P V)
//              -> running O kept in registers: o = o * alpha + O_j (no TMEM rescale pass).
// Backward work item = (kv block, head, batch) looping over q blocks i >= j; five MMAs per pair:
//   S = Q K^T, dP = dO V^T, dV += P~^T dO, dK += dS^T Q, dQ_i = dS K
// S / dP of pair i+1 are issued as soon as pair i's are in registers; dS (smem) and dQ (TMEM) are double buffered
// so the other three MMAs of pair i run under the element-wise work of pair i+1 (P~ has one buffer, released by
// the dV MMA, which runs first; Q / dO have three stages because a stage is held until the pair's LAST MMAs).
// dQ tiles leave through TMA reduce-add into the fp32 accumulator, staged in the pair's own dS buffer: the
// red.global.add.v4 it replaces touched 32 lines per warp instruction and held the loop on its source registers.
// P~ / dS tiles are written once to smem and consumed both K-major (dS K) and MN-major (P~^T dO,
// dS^T Q) through different UMMA descriptors -- no transposes.
//
// LSE is stored in the base-2 domain *including* the 1/sqrt(D) scale:  lse2 = max2 + log2(sum).
#include <cuda_bf16.h>
#include <math.h>

#include <mutex>

#include "../include/b200_trainer.h"
#include "common/host.h"
#include "common/philox.cuh"
#include "common/ptx.cuh"

namespace b200 {

constexpr int kBlk = 128;      // q rows per tile == kv rows per tile
constexpr int kD = 64;         // head_dim
constexpr int kTile = kBlk * kD * 2;          // 16 KB: [128 rows][128 B]
constexpr int kPTile = 2 * kTile;             // 32 KB: two 64-column sub-tiles of a 128x128 bf16 matrix
constexpr float kScale = 0.125f;              // 1/sqrt(64)
constexpr float kScale2 = 0.125f * 1.4426950408889634f;

struct AttnParams {
  int B, S, H;
  int S16;               // ceil(S/16): dropout groups (one Philox call, 8-bit lanes) per score row
  float p_drop;
  const unsigned long long* rng;
  unsigned int rng_stream;
  __nv_bfloat16* out;    // fwd: [B*S, h]
  float* lse;            // [B, H, S] base-2
  // backward
  const float* delta;    // [B, H, S]
  float* dq_accum;       // [B*S, h] fp32
  __nv_bfloat16* dqkv;   // [B*S, 3h]
  int tma_out;           // S % 128 == 0: output tiles (fwd O, bwd dK / dV) leave through shared memory + TMA stores
};

// 16-byte chunk `cc` (0..7) of row `row` inside a [rows][128 B] 128B-swizzled tile
__device__ __forceinline__ uint32_t swz(uint32_t tile_base, int row, int cc) {
  return tile_base + row * 128 + ((cc ^ (row & 7)) << 4);
}
// single MUFU.EX2 (exp2f() adds a denormal-range compare and two predicated multiplies per element)
__device__ __forceinline__ float ex2_ftz(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// ================================================================================================ forward
struct FwdSmem {
  static constexpr int kQ = 0;                     // 2 buffers (next tile's Q lands while this one finishes)
  static constexpr int kK = kQ + 2 * kTile;        // 2 stages
  static constexpr int kV = kK + 2 * kTile;        // 2 stages
  static constexpr int kO = kV + 2 * kTile;        // output tile staging for the TMA store
  static constexpr int kBars = kO + kTile;         // (the probabilities never touch shared memory: they live in TMEM)
  static constexpr int kXchg = kBars + 256;                 // float [2][4][128] row-max exchange
  static constexpr int kXchgL = kXchg + 2 * 4 * 128 * 4;    // float [4][128] row-sum exchange (end of a tile)
  static constexpr int kTotal = kXchgL + 4 * 128 * 4 + 1024;
};

constexpr int kSoftmaxWarps = 16;   // four per scheduler: each thread owns one row x 32 score columns
constexpr int kAttnThreads = 64 + kSoftmaxWarps * 32;

// The four threads that share a score row sit in the four warps with the same lane quarter `qd`
// (one per 32-column quarter), so the row-max exchange only needs those 128 threads: one named
// barrier per quarter instead of one over all 512 softmax threads.
__device__ __forceinline__ void softmax_bar_sync(int qd) { asm volatile("bar.sync %0, 128;" ::"r"(qd + 1) : "memory"); }

// Work list of a persistent attention CTA.  Tiles are numbered heaviest first (tile / bh_count = size
// rank) and dealt to the CTAs in boustrophedon order -- round 0 left to right, round 1 right to left, ... --
// which for causal tile sizes 8,7,..,1 x 96 on 148 CTAs gives every CTA 21..24 kv blocks (mean 23.4), as
// good as a greedy scheduler and with no atomics, barrier re-initialisation or TMEM re-allocation
// between tiles.  Returns the tile index of this CTA's round r, or -1 when that slot is past the end.
__device__ __forceinline__ int snake_tile(int r, int total) {
  const int P = int(gridDim.x), c = int(blockIdx.x);
  const int idx = r * P + ((r & 1) ? P - 1 - c : c);
  return idx < total ? idx : -1;
}

__global__ void __launch_bounds__(kAttnThreads, 1)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tm_qkv, const __grid_constant__ CUtensorMap tm_out, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + FwdSmem::kBars);
  uint64_t* q_full = bars;            // 2
  uint64_t* q_empty = bars + 2;       // 2
  uint64_t* k_full = bars + 4;        // 2
  uint64_t* k_empty = bars + 6;       // 2
  uint64_t* v_full = bars + 8;        // 2
  uint64_t* v_empty = bars + 10;      // 2
  uint64_t* s_full = bars + 12;       // 2
  uint64_t* s_free = bars + 14;       // 2 (count = softmax warps)
  uint64_t* p_ready = bars + 16;      // 2 (count = softmax warps)
  uint64_t* o_full = bars + 18;       // 2
  uint64_t* o_free = bars + 20;       // 2 (count = softmax warps)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 22);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = (p.S + kBlk - 1) / kBlk;
  const int bh_count = p.H * p.B;
  const int total = nq * bh_count;
  const int rounds = (total + int(gridDim.x) - 1) / int(gridDim.x);
  const int h = p.H * kD;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tm_qkv);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&q_full[i], 1);
      mbar_init(&q_empty[i], 1);
      mbar_init(&k_full[i], 1);
      mbar_init(&k_empty[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&v_empty[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&s_free[i], kSoftmaxWarps);
      mbar_init(&p_ready[i], kSoftmaxWarps);
      mbar_init(&o_full[i], 1);
      mbar_init(&o_free[i], kSoftmaxWarps);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<512>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_ptr;
  // TMEM columns: S [0, 256) and O [256, 384) double buffered; P~ [384, 512) (bf16, two per 32-bit column: 64
  // columns per 128 kv positions) is written by the softmax threads with tcgen05.st and read by the P V MMA as its A
  // operand straight from tensor memory.  (Computed, not tabulated: an indexed array lands in local memory.)
  auto tS = [&](int st) { return tmem + uint32_t(st) * 128u; };
  auto tO = [&](int st) { return tmem + 256u + uint32_t(st) * 64u; };
  auto tP = [&](int st) { return tmem + 384u + uint32_t(st) * 64u; };

  // Every role walks the same tile list; `ti` counts this CTA's tiles (Q buffer / phase) and `G` its kv
  // blocks across tiles (K/V stages, S/P/O buffers and their phases), so no barrier is ever re-armed.
  if (warp == 0) {
    if (lane == 0) {
      uint32_t G = 0;
      int ti = 0;
      for (int r = 0; r < rounds; ++r) {
        const int tile = snake_tile(r, total);
        if (tile < 0) continue;
        const int qi = nq - 1 - tile / bh_count, bh = tile % bh_count;
        const int hd = bh % p.H, row_base = (bh / p.H) * p.S;
        const int qb = ti & 1;
        mbar_wait(&q_empty[qb], ((ti >> 1) & 1) ^ 1);
        mbar_expect_tx(&q_full[qb], kTile);
        tma_load_2d(smem + FwdSmem::kQ + qb * kTile, &tm_qkv, &q_full[qb], hd * kD, row_base + qi * kBlk);
        for (int j = 0; j <= qi; ++j, ++G) {
          const int st = G & 1;
          const uint32_t ph = (G >> 1) & 1;
          mbar_wait(&k_empty[st], ph ^ 1);
          mbar_expect_tx(&k_full[st], kTile);
          tma_load_2d(smem + FwdSmem::kK + st * kTile, &tm_qkv, &k_full[st], h + hd * kD, row_base + j * kBlk);
          mbar_wait(&v_empty[st], ph ^ 1);
          mbar_expect_tx(&v_full[st], kTile);
          tma_load_2d(smem + FwdSmem::kV + st * kTile, &tm_qkv, &v_full[st], 2 * h + hd * kD, row_base + j * kBlk);
        }
        ++ti;
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc_bf16(128, 128, false, false);
      constexpr uint32_t idesc_o = make_idesc_bf16(128, 64, false, true);
      auto issue_s = [&](uint32_t g, uint32_t q_addr) {
        const int st = g & 1;
        const uint32_t ph = (g >> 1) & 1;
        mbar_wait(&k_full[st], ph);
        mbar_wait(&s_free[st], ph ^ 1);
        tc_fence_after();
        const uint32_t k_addr = smem_u32(smem + FwdSmem::kK + st * kTile);
#pragma unroll
        for (int t = 0; t < kD / 16; ++t)
          umma_bf16_ss(tS(st), make_smem_desc_sw128(q_addr + t * 32, 16, 1024),
                       make_smem_desc_sw128(k_addr + t * 32, 16, 1024), idesc_s, t > 0);
        tc_commit(&k_empty[st]);
        tc_commit(&s_full[st]);
      };
      auto issue_pv = [&](uint32_t g) {
        const int st = g & 1;
        const uint32_t ph = (g >> 1) & 1;
        mbar_wait(&p_ready[st], ph);
        mbar_wait(&v_full[st], ph);
        mbar_wait(&o_free[st], ph ^ 1);
        tc_fence_after();
        const uint32_t v_addr = smem_u32(smem + FwdSmem::kV + st * kTile);
#pragma unroll
        for (int t = 0; t < kBlk / 16; ++t)   // A = P~ from TMEM: 16 kv positions = 8 packed columns per K step
          umma_bf16_ts(tO(st), tP(st) + t * 8, make_smem_desc_sw128(v_addr + t * 2048, kTile, 1024), idesc_o, t > 0);
        tc_commit(&v_empty[st]);
        tc_commit(&o_full[st]);
      };
      uint32_t G = 0;
      int ti = 0;
      for (int r = 0; r < rounds; ++r) {
        const int tile = snake_tile(r, total);
        if (tile < 0) continue;
        const int nkv = nq - tile / bh_count;   // = qi + 1
        const int qb = ti & 1;
        const uint32_t q_addr = smem_u32(smem + FwdSmem::kQ + qb * kTile);
        mbar_wait(&q_full[qb], (ti >> 1) & 1);
        issue_s(G, q_addr);
        for (int j = 0; j < nkv; ++j) {
          if (j + 1 < nkv) issue_s(G + j + 1, q_addr);
          else tc_commit(&q_empty[qb]);          // every S MMA of this tile has been issued: Q may be replaced
          issue_pv(G + j);
        }
        G += nkv;
        ++ti;
      }
    }
  } else {
    const int qd = warp & 3;
    const int qt = (warp - 2) >> 2;               // which 32-column quarter of the score tile this thread owns
    const int r = qd * 32 + lane;                 // row within the q tile
    const uint32_t lane_off = uint32_t(qd * 32) << 16;
    const bool drop = p.p_drop > 0.f;
    const uint32_t t4 = dropout_threshold8(p.p_drop) * 0x01010101u;
    const float inv_keep = drop ? 1.f / dropout_keep_prob8(p.p_drop) : 1.f;
    unsigned long long seed = 0, offs = 0;
    if (drop) { seed = p.rng[0]; offs = p.rng[1]; }
    float* xchg = reinterpret_cast<float*>(smem + FwdSmem::kXchg);     // [2][4][128]
    float* xchg_l = reinterpret_cast<float*>(smem + FwdSmem::kXchgL);  // [4][128]

    uint32_t G = 0;
    for (int rr = 0; rr < rounds; ++rr) {
      const int tile = snake_tile(rr, total);
      if (tile < 0) continue;
      const int qi = nq - 1 - tile / bh_count, bh = tile % bh_count;
      const int hd = bh % p.H, b = bh / p.H;
      const int nkv = qi + 1;
      const int row_base = b * p.S;
      const int qrow = qi * kBlk + r;               // position in the sequence
      const unsigned long long grp_row = ((unsigned long long)bh * p.S + qrow) * p.S16;

      float m = -INFINITY, l = 0.f, alpha_prev = 1.f;
      float o[16];
#pragma unroll
      for (int e = 0; e < 16; ++e) o[e] = 0.f;

      auto fold_o = [&](uint32_t g, float alpha) {
        const int pb = g & 1;
        mbar_wait(&o_full[pb], (g >> 1) & 1);
        tc_fence_after();
        uint32_t v[16];
        __syncwarp();
        tmem_ld_32x32b_x16(tO(pb) + lane_off + qt * 16, v);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 16; e += 2) {   // FFMA2: two accumulator lanes per issue slot
          const float2 t = ffma2(make_float2(o[e], o[e + 1]), make_float2(alpha, alpha),
                                 make_float2(__uint_as_float(v[e]), __uint_as_float(v[e + 1])));
          o[e] = t.x;
          o[e + 1] = t.y;
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&o_free[pb]);
      };

      for (int j = 0; j < nkv; ++j) {
        const uint32_t g = G + j;
        const int st = g & 1;
        const uint32_t ph = (g >> 1) & 1;
        const bool diag = (j == qi);
        mbar_wait(&s_full[st], ph);
        tc_fence_after();
        // this thread's 32 scores stay in registers for both the max and the exp pass
        uint32_t v[32];
        __syncwarp();
        tmem_ld_32x32b_x32(tS(st) + lane_off + qt * 32, v);
        tmem_ld_wait();
        // The causal mask only touches the diagonal tile: a warp-uniform branch keeps the compare +
        // select out of every other tile.  Scores stay raw (unscaled) -- the 1/sqrt(d) * log2(e) factor
        // is folded into the exponent's FFMA below, and max commutes with a positive scale.
        if (diag) {
#pragma unroll
          for (int e = 0; e < 32; ++e)
            if ((qt * 32 + e) > r) v[e] = 0xFF800000u;   // -inf
        }
        float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};   // four chains instead of one serial max
#pragma unroll
        for (int e = 0; e < 32; ++e) mx4[e & 3] = fmaxf(mx4[e & 3], __uint_as_float(v[e]));
        float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
        xchg[(st * 4 + qt) * 128 + r] = mx;
        softmax_bar_sync(qd);
#pragma unroll
        for (int t = 0; t < 4; ++t) mx = fmaxf(mx, xchg[(st * 4 + t) * 128 + r]);
        mx = fmaxf(mx, m);
        const float m_use = (mx == -INFINITY) ? 0.f : mx;   // m, m_use: running max of the RAW scores
        const float alpha = ex2_ftz((m - m_use) * kScale2);
        const float neg_msc = -m_use * kScale2;
        // probabilities -> TMEM (bf16 pairs, dropout applied, unnormalised)
        float2 l2[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
        const float2 sc2 = make_float2(kScale2, kScale2), nm2 = make_float2(neg_msc, neg_msc);
#pragma unroll
        for (int g16 = 0; g16 < 2; ++g16) {
          // one Philox call = sixteen 8-bit lanes = this thread's next 16 kv positions
          uint4 rb = make_uint4(0u, 0u, 0u, 0u);
          if (drop) rb = dropout_bits(seed, offs, p.rng_stream, grp_row + ((j * kBlk + qt * 32 + g16 * 16) >> 4));
          const uint32_t w[4] = {rb.x, rb.y, rb.z, rb.w};
          uint32_t pk8[8];
#pragma unroll
          for (int hf = 0; hf < 2; ++hf) {
            const int gq = g16 * 2 + hf;
            float pv[8];
#pragma unroll
            for (int e = 0; e < 8; e += 2) {
              // per PAIR of elements: one FFMA2 (exponent), two MUFU, one FADD2 (row sum); masked entries are ex2(-inf) = 0
              const float2 arg = ffma2(make_float2(__uint_as_float(v[gq * 8 + e]), __uint_as_float(v[gq * 8 + e + 1])), sc2, nm2);
              pv[e] = ex2_ftz(arg.x);
              pv[e + 1] = ex2_ftz(arg.y);
              l2[(e >> 1) & 1] = fadd2(l2[(e >> 1) & 1], make_float2(pv[e], pv[e + 1]));
            }
            uint32_t pk[4] = {pack_bf16x2(pv[0], pv[1]), pack_bf16x2(pv[2], pv[3]), pack_bf16x2(pv[4], pv[5]),
                              pack_bf16x2(pv[6], pv[7])};
            if (drop) {   // dropped probabilities are cleared on the packed bf16 pairs: compare 4 bytes at once, AND
              uint32_t lo, hi;
              keep_masks_bf16x2(w[hf * 2], t4, lo, hi);
              pk[0] &= lo; pk[1] &= hi;
              keep_masks_bf16x2(w[hf * 2 + 1], t4, lo, hi);
              pk[2] &= lo; pk[3] &= hi;
            }
#pragma unroll
            for (int e = 0; e < 4; ++e) pk8[hf * 4 + e] = pk[e];
          }
          tmem_st_32x32b_x8(tP(st) + lane_off + qt * 16 + g16 * 8, pk8);
        }
        l = l * alpha + ((l2[0].x + l2[0].y) + (l2[1].x + l2[1].y));
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(&p_ready[st]);
          mbar_arrive(&s_free[st]);
        }
        if (j > 0) fold_o(g - 1, alpha_prev);   // fold in O_{j-1}
        alpha_prev = alpha;
        m = m_use;
      }
      // total row sum = the four column-quarters' partial sums (same running max on all four threads); exchanged
      // while the last P V product is still running, then that product is folded in
      xchg_l[qt * 128 + r] = l;
      const bool out_leader = qt == 0 && lane == 0;
      if (p.tma_out && out_leader) tma_store_wait_read<0>();   // the previous tile's store has left the staging rows
      softmax_bar_sync(qd);
      l = (xchg_l[r] + xchg_l[128 + r]) + (xchg_l[256 + r] + xchg_l[384 + r]);
      fold_o(G + nkv - 1, alpha_prev);
      G += nkv;
      if (p.tma_out) {
        // A thread's 16 output columns are 32 B of a row whose neighbours sit 2 h bytes apart: stored directly, every
        // warp instruction touches 32 lines.  The quarter's 32 rows x 128 B go through shared memory and one TMA store.
        const float inv = inv_keep / l;
        const uint32_t o_tile = smem_u32(smem + FwdSmem::kO);
#pragma unroll
        for (int gq = 0; gq < 2; ++gq)
          st_shared_v4(swz(o_tile, r, qt * 2 + gq),
                       pack_bf16x2(o[gq * 8 + 0] * inv, o[gq * 8 + 1] * inv), pack_bf16x2(o[gq * 8 + 2] * inv, o[gq * 8 + 3] * inv),
                       pack_bf16x2(o[gq * 8 + 4] * inv, o[gq * 8 + 5] * inv), pack_bf16x2(o[gq * 8 + 6] * inv, o[gq * 8 + 7] * inv));
        if (qt == 0) p.lse[size_t(bh) * p.S + qrow] = m * kScale2 + log2f(l);
        fence_proxy_async_smem();
        softmax_bar_sync(qd);
        if (out_leader) {
          tma_store_2d(&tm_out, smem + FwdSmem::kO + qd * 4096, hd * kD, row_base + qi * kBlk + qd * 32);
          tma_store_commit();
        }
      } else if (qrow < p.S) {
        const float inv = inv_keep / l;
        __nv_bfloat16* dst = p.out + size_t(row_base + qrow) * h + hd * kD + qt * 16;
#pragma unroll
        for (int gq = 0; gq < 2; ++gq)
          *reinterpret_cast<uint4*>(dst + gq * 8) =
              make_uint4(pack_bf16x2(o[gq * 8 + 0] * inv, o[gq * 8 + 1] * inv), pack_bf16x2(o[gq * 8 + 2] * inv, o[gq * 8 + 3] * inv),
                         pack_bf16x2(o[gq * 8 + 4] * inv, o[gq * 8 + 5] * inv), pack_bf16x2(o[gq * 8 + 6] * inv, o[gq * 8 + 7] * inv));
        if (qt == 0) p.lse[size_t(bh) * p.S + qrow] = m * kScale2 + log2f(l);
      }
      // the row-sum slots are rewritten by the next tile only after its own kv loop, which contains
      // at least one quarter-wide barrier: no thread can still be reading this tile's sums by then
    }
  }

  if (warp >= 2 && warp < 6 && lane == 0) tma_store_wait<0>();   // output stores have left shared memory
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<512>(tmem);
  }
}

// ================================================================================================ backward
// delta[b,h,s] = sum_d dO * O   (one thread per (row, head))
__global__ void attn_delta_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ d_o,
                                  float* __restrict__ delta, int B, int S, int H) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= (long long)B * S * H) return;
  const int hd = int(t % H);
  const long long row = t / H;
  const int b = int(row / S), s = int(row % S);
  const uint4* po = reinterpret_cast<const uint4*>(o + row * (H * kD) + hd * kD);
  const uint4* pd = reinterpret_cast<const uint4*>(d_o + row * (H * kD) + hd * kD);
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const uint4 a = po[i], c = pd[i];
    const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, cw[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 x = unpack_bf16x2(aw[e]), y = unpack_bf16x2(cw[e]);
      acc += x.x * y.x + x.y * y.y;
    }
  }
  delta[size_t(b * H + hd) * S + s] = acc;
}

// dq (fp32 accumulator) -> bf16 q-columns of dqkv
__global__ void attn_dq_cast_kernel(const float* __restrict__ dq, __nv_bfloat16* __restrict__ dqkv, long long rows,
                                    int h) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const int gpr = h / 8;
  if (t >= rows * gpr) return;
  const long long row = t / gpr;
  const int c = int(t % gpr) * 8;
  const float4 a = *reinterpret_cast<const float4*>(dq + row * h + c), bq = *reinterpret_cast<const float4*>(dq + row * h + c + 4);
  *reinterpret_cast<uint4*>(dqkv + row * 3 * h + c) =
      make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(bq.x, bq.y), pack_bf16x2(bq.z, bq.w));
}

// The two wrappers above plus the memset of the dQ accumulator and the rotary backward are four
// memory-bound launches around the main kernel; the *_rope entry point runs them as two:
//   attn_delta_zero_kernel : delta AND the zero-fill of dq_accum (eight lanes per (row, head), coalesced)
//   attn_dq_rope_kernel    : dQ fp32 -> inverse rotary -> bf16 q columns, and the inverse rotary of the dK
//                            columns in place (the rotation is applied to the fp32 dQ: one rounding, not two)
__global__ void attn_delta_zero_kernel(const __nv_bfloat16* __restrict__ o, const __nv_bfloat16* __restrict__ d_o,
                                       float* __restrict__ delta, float* __restrict__ dq, long long pairs, int S, int H) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  const long long rh = t >> 3;          // (row, head) pair; consecutive pairs are consecutive 64-column groups
  const int sub = int(t & 7);
  float acc = 0.f;
  if (rh < pairs) {
    const uint4 a = reinterpret_cast<const uint4*>(o + rh * kD)[sub], c = reinterpret_cast<const uint4*>(d_o + rh * kD)[sub];
    const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, cw[4] = {c.x, c.y, c.z, c.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 x = unpack_bf16x2(aw[e]), y = unpack_bf16x2(cw[e]);
      acc += x.x * y.x + x.y * y.y;
    }
    float4* z = reinterpret_cast<float4*>(dq + rh * kD);
    z[sub] = make_float4(0.f, 0.f, 0.f, 0.f);
    z[sub + 8] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
#pragma unroll
  for (int off = 4; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
  if (rh < pairs && sub == 0) {
    const long long row = rh / H;
    const int hd = int(rh % H);
    delta[((row / S) * H + hd) * S + (row % S)] = acc;
  }
}

// thread per (row, which in {dQ, dK}, head, 8-wide group of the first half); pairs column j with j + 32
__global__ void attn_dq_rope_kernel(const float* __restrict__ dq, __nv_bfloat16* __restrict__ dqkv,
                                    const float* __restrict__ cs, const float* __restrict__ sn, long long total, int H,
                                    int S) {
  const long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (t >= total) return;
  const int g = int(t & 3);
  const long long rh = t >> 2;
  const int head2 = int(rh % (2 * H));   // 0..H-1: dQ heads, H..2H-1: dK heads
  const long long row = rh / (2 * H);
  const int pos = int(row % S), h = H * kD;
  const float4* c4 = reinterpret_cast<const float4*>(cs + size_t(pos) * 32) + 2 * g;
  const float4* s4 = reinterpret_cast<const float4*>(sn + size_t(pos) * 32) + 2 * g;
  const float4 c0 = c4[0], c1 = c4[1], s0 = s4[0], s1 = s4[1];
  const float cc[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
  const float ss[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
  __nv_bfloat16* dst = dqkv + row * 3 * h + head2 * kD + g * 8;   // dK columns start at h = H * 64
  float a[8], b[8];
  if (head2 < H) {
    const float* src = dq + row * h + head2 * kD + g * 8;
    const float4 a0 = *reinterpret_cast<const float4*>(src), a1 = *reinterpret_cast<const float4*>(src + 4);
    const float4 b0 = *reinterpret_cast<const float4*>(src + 32), b1 = *reinterpret_cast<const float4*>(src + 36);
    a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
    b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w; b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
  } else {
    const uint4 lo = *reinterpret_cast<const uint4*>(dst), hi = *reinterpret_cast<const uint4*>(dst + 32);
    const uint32_t low[4] = {lo.x, lo.y, lo.z, lo.w}, hiw[4] = {hi.x, hi.y, hi.z, hi.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 x = unpack_bf16x2(low[e]), y = unpack_bf16x2(hiw[e]);
      a[2 * e] = x.x; a[2 * e + 1] = x.y; b[2 * e] = y.x; b[2 * e + 1] = y.y;
    }
  }
  uint32_t ol[4], oh[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {   // transpose of the forward rotation
    ol[e] = pack_bf16x2(a[2 * e] * cc[2 * e] + b[2 * e] * ss[2 * e], a[2 * e + 1] * cc[2 * e + 1] + b[2 * e + 1] * ss[2 * e + 1]);
    oh[e] = pack_bf16x2(b[2 * e] * cc[2 * e] - a[2 * e] * ss[2 * e], b[2 * e + 1] * cc[2 * e + 1] - a[2 * e + 1] * ss[2 * e + 1]);
  }
  *reinterpret_cast<uint4*>(dst) = make_uint4(ol[0], ol[1], ol[2], ol[3]);
  *reinterpret_cast<uint4*>(dst + 32) = make_uint4(oh[0], oh[1], oh[2], oh[3]);
}

struct BwdSmem {
  static constexpr int kK = 0;
  static constexpr int kV = kK + kTile;
  static constexpr int kQ = kV + kTile;            // 3 stages: a stage is held until the pair's dK MMA has read it, and
  static constexpr int kdO = kQ + 3 * kTile;       // 3 stages  with two the next-but-one load started a pair too late
  static constexpr int kP = kdO + 3 * kTile;       // 1 buffer of 32 KB (freed by the dV MMA: p_free)
  static constexpr int kdS = kP + kPTile;          // 2 buffers x 32 KB (also the staging of the dQ reduce-add)
  static constexpr int kBars = kdS + 2 * kPTile;
  static constexpr int kTotal = kBars + 256 + 1024;
};
static_assert(BwdSmem::kTotal <= 227 * 1024, "backward smem exceeds the 227 KB per-CTA limit");

// Persistent: one CTA per SM walks its boustrophedon share of the (kv block, head, batch) list (see
// snake_tile), heaviest kv blocks (small j = most q blocks) first.  `ti` counts this CTA's tiles and `I`
// its (kv block, q block) pairs across tiles; every ring / buffer index and mbarrier phase derives from
// those two counters, so nothing is re-initialised between tiles.
__global__ void __launch_bounds__(kAttnThreads, 1)
attn_bwd_kernel(const __grid_constant__ CUtensorMap tm_qkv, const __grid_constant__ CUtensorMap tm_do,
                const __grid_constant__ CUtensorMap tm_dq, const __grid_constant__ CUtensorMap tm_dkv, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + BwdSmem::kBars);
  uint64_t* kv_full = bars;            // 1
  uint64_t* kv_empty = bars + 1;       // 1  every MMA of the tile has completed: K / V may be replaced
  uint64_t* qdo_full = bars + 2;       // 3
  uint64_t* qdo_empty = bars + 5;      // 3
  uint64_t* sdp_full = bars + 8;       // 1
  uint64_t* s_free = bars + 9;         // 1 (count = softmax warps)
  uint64_t* pds_ready = bars + 10;     // 2 (count = softmax warps)
  uint64_t* dq_full = bars + 12;       // 2
  uint64_t* dq_free = bars + 14;       // 2 (count = softmax warps)
  uint64_t* acc_full = bars + 16;      // 1
  uint64_t* acc_free = bars + 17;      // 1 (count = softmax warps): dK / dV have been read out of TMEM
  uint64_t* p_free = bars + 18;        // 1  the dV MMA of the pair has read P~
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 19);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nq = (p.S + kBlk - 1) / kBlk;
  const int bh_count = p.H * p.B;
  const int total = nq * bh_count;
  const int rounds = (total + int(gridDim.x) - 1) / int(gridDim.x);
  const int h = p.H * kD;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tm_qkv);
    tma_prefetch_desc(&tm_do);
    tma_prefetch_desc(&tm_dq);
    tma_prefetch_desc(&tm_dkv);
    mbar_init(kv_full, 1);
    mbar_init(kv_empty, 1);
    for (int i = 0; i < 3; ++i) {
      mbar_init(&qdo_full[i], 1);
      mbar_init(&qdo_empty[i], 1);
    }
    mbar_init(p_free, 1);
    mbar_init(sdp_full, 1);
    mbar_init(s_free, kSoftmaxWarps);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&pds_ready[i], kSoftmaxWarps);
      mbar_init(&dq_full[i], 1);
      mbar_init(&dq_free[i], kSoftmaxWarps);
    }
    mbar_init(acc_full, 1);
    mbar_init(acc_free, kSoftmaxWarps);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<512>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_ptr;
  const uint32_t tS = tmem, tdP = tmem + 128, tdV = tmem + 256, tdK = tmem + 320, tdQ = tmem + 384;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t I = 0;
      int ti = 0;
      for (int r = 0; r < rounds; ++r) {
        const int tile = snake_tile(r, total);
        if (tile < 0) continue;
        const int j = tile / bh_count, bh = tile % bh_count;
        const int hd = bh % p.H, row_base = (bh / p.H) * p.S;
        mbar_wait(kv_empty, (ti & 1) ^ 1);
        mbar_expect_tx(kv_full, 2 * kTile);
        tma_load_2d(smem + BwdSmem::kK, &tm_qkv, kv_full, h + hd * kD, row_base + j * kBlk);
        tma_load_2d(smem + BwdSmem::kV, &tm_qkv, kv_full, 2 * h + hd * kD, row_base + j * kBlk);
        for (int it = 0; it < nq - j; ++it, ++I) {
          const int st = I % 3;
          mbar_wait(&qdo_empty[st], ((I / 3) & 1) ^ 1);
          mbar_expect_tx(&qdo_full[st], 2 * kTile);
          tma_load_2d(smem + BwdSmem::kQ + st * kTile, &tm_qkv, &qdo_full[st], hd * kD, row_base + (j + it) * kBlk);
          tma_load_2d(smem + BwdSmem::kdO + st * kTile, &tm_do, &qdo_full[st], hd * kD, row_base + (j + it) * kBlk);
        }
        ++ti;
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc_bf16(128, 128, false, false);    // Q K^T, dO V^T
      constexpr uint32_t idesc_tt = make_idesc_bf16(128, 64, true, true);      // P^T dO, dS^T Q
      constexpr uint32_t idesc_dq = make_idesc_bf16(128, 64, false, true);     // dS K
      const uint32_t k_addr = smem_u32(smem + BwdSmem::kK), v_addr = smem_u32(smem + BwdSmem::kV);
      const uint32_t p_addr = smem_u32(smem + BwdSmem::kP), ds_addr = smem_u32(smem + BwdSmem::kdS);
      // S = Q K^T and dP = dO V^T of pair I (one TMEM buffer each; the element-wise warps free it as soon
      // as both are in registers, so the next pair's S / dP run under this pair's softmax).
      auto issue_sdp = [&](uint32_t I) {
        const int st = I % 3;
        const uint32_t q_addr = smem_u32(smem + BwdSmem::kQ + st * kTile);
        const uint32_t do_addr = smem_u32(smem + BwdSmem::kdO + st * kTile);
        mbar_wait(&qdo_full[st], (I / 3) & 1);
        mbar_wait(s_free, (I & 1) ^ 1);
        tc_fence_after();
#pragma unroll
        for (int t = 0; t < 4; ++t)
          umma_bf16_ss(tS, make_smem_desc_sw128(q_addr + t * 32, 16, 1024), make_smem_desc_sw128(k_addr + t * 32, 16, 1024),
                       idesc_s, t > 0);
#pragma unroll
        for (int t = 0; t < 4; ++t)
          umma_bf16_ss(tdP, make_smem_desc_sw128(do_addr + t * 32, 16, 1024), make_smem_desc_sw128(v_addr + t * 32, 16, 1024),
                       idesc_s, t > 0);
        tc_commit(sdp_full);
      };
      uint32_t I0 = 0;
      int ti = 0;
      for (int r = 0; r < rounds; ++r) {
        const int tile = snake_tile(r, total);
        if (tile < 0) continue;
        const int niter = nq - tile / bh_count;
        mbar_wait(kv_full, ti & 1);
        issue_sdp(I0);
        mbar_wait(acc_free, (ti & 1) ^ 1);     // previous tile's dK / dV have left TMEM
        for (int it = 0; it < niter; ++it) {
          const uint32_t I = I0 + it;
          const int st = I & 1;                 // dS buffer and dQ buffer of this pair
          const int qst = I % 3;                // its Q / dO stage
          const uint32_t use = (I >> 1) & 1;
          const uint32_t q_addr = smem_u32(smem + BwdSmem::kQ + qst * kTile);
          const uint32_t do_addr = smem_u32(smem + BwdSmem::kdO + qst * kTile);
          const uint32_t pb = p_addr, dsb = ds_addr + st * kPTile;
          if (it + 1 < niter) issue_sdp(I + 1);
          mbar_wait(&pds_ready[st], use);
          mbar_wait(&dq_free[st], use ^ 1);
          tc_fence_after();
#pragma unroll
          for (int t = 0; t < 8; ++t)  // dV += P~^T dO   (K = 128 q rows, 16 per step)
            umma_bf16_ss(tdV, make_smem_desc_sw128(pb + t * 2048, kTile, 1024),
                         make_smem_desc_sw128(do_addr + t * 2048, kTile, 1024), idesc_tt, (it > 0 || t > 0));
          tc_commit(p_free);           // the single P~ buffer may be rewritten by the next pair
#pragma unroll
          for (int t = 0; t < 8; ++t)  // dK += dS^T Q
            umma_bf16_ss(tdK, make_smem_desc_sw128(dsb + t * 2048, kTile, 1024),
                         make_smem_desc_sw128(q_addr + t * 2048, kTile, 1024), idesc_tt, (it > 0 || t > 0));
#pragma unroll
          for (int t = 0; t < 8; ++t)  // dQ_i = dS K     (K = 128 kv, 16 per step)
            umma_bf16_ss(tdQ + st * 64, make_smem_desc_sw128(dsb + (t >> 2) * kTile + (t & 3) * 32, 16, 1024),
                         make_smem_desc_sw128(k_addr + t * 2048, kTile, 1024), idesc_dq, t > 0);
          tc_commit(&qdo_empty[qst]);
          tc_commit(&dq_full[st]);
        }
        tc_commit(acc_full);
        tc_commit(kv_empty);
        I0 += niter;
        ++ti;
      }
    }
  } else {
    const int qd = warp & 3;
    const int qt = (warp - 2) >> 2;   // four threads per row: each owns 32 kv columns / 16 head-dim columns
    const int r = qd * 32 + lane;
    const uint32_t lane_off = uint32_t(qd * 32) << 16;
    const bool drop = p.p_drop > 0.f;
    const uint32_t t4 = dropout_threshold8(p.p_drop) * 0x01010101u;
    const float keep_prob = drop ? dropout_keep_prob8(p.p_drop) : 1.f;
    const float log2_inv_keep = -log2f(keep_prob);
    const float keep_scale = keep_prob * kScale;
    unsigned long long seed = 0, offs = 0;
    if (drop) { seed = p.rng[0]; offs = p.rng[1]; }
    const uint32_t p_sub0 = smem_u32(smem + BwdSmem::kP) + (qt >> 1) * kTile;
    const uint32_t ds_sub0 = smem_u32(smem + BwdSmem::kdS) + (qt >> 1) * kTile;
    // dQ leaves through TMA reduce-add.  The two warps (qd, qt) and (qd, qt ^ 1) hold rows [32 qd, +32) x head-dim
    // columns [32 (qt >> 1), +32) of it: as fp32 that is a 32 x 128 B swizzled box = exactly the 4 KB of the dS
    // buffer that only these two warps ever write (each thread stages into the bytes of its own dS chunks), so the
    // staging needs no shared memory of its own and only a 64-thread named barrier.
    const int pair_bar = 1 + qd * 2 + (qt >> 1);
    const bool dq_leader = (qt & 1) == 0 && lane == 0;
    uint8_t* const dq_stage0 = smem + BwdSmem::kdS + (qt >> 1) * kTile + qd * 4096;

    uint32_t I0 = 0;
    int ti = 0;
    for (int rr = 0; rr < rounds; ++rr) {
      const int tile = snake_tile(rr, total);
      if (tile < 0) continue;
      const int j = tile / bh_count, bh = tile % bh_count;
      const int hd = bh % p.H, row_base = (bh / p.H) * p.S;
      const int niter = nq - j;

      // dQ of pair I (local q block `it`) -> fp32 accumulator in global memory.  dq_full(I) also says the MMAs that
      // read P~ / dS buffer I&1 are done, so that dS buffer is free to stage the tile until pair I+2 writes it.
      // Rows past the end of the sequence carry dS = 0 (their probabilities are 0), so the whole box may be added.
      auto drain_dq = [&](uint32_t I, int it) {
        mbar_wait(&dq_full[I & 1], (I >> 1) & 1);
        tc_fence_after();
        uint32_t v[16];
        __syncwarp();
        tmem_ld_32x32b_x16(tdQ + (I & 1) * 64 + lane_off + qt * 16, v);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&dq_free[I & 1]);
        const uint32_t ds_sub = ds_sub0 + (I & 1) * kPTile;
#pragma unroll
        for (int g = 0; g < 4; ++g)
          st_shared_v4(swz(ds_sub, r, (qt & 1) * 4 + g), v[g * 4 + 0], v[g * 4 + 1], v[g * 4 + 2], v[g * 4 + 3]);
        fence_proxy_async_smem();
        named_bar_sync(pair_bar, 64);
        if (dq_leader) {
          tma_reduce_add_2d(&tm_dq, dq_stage0 + (I & 1) * kPTile, hd * kD + (qt >> 1) * 32, row_base + (j + it) * kBlk + qd * 32);
          tma_store_commit();
        }
      };

      // per-row statistics are fetched one block ahead so their global-load latency hides under the math
      const float* lse_bh = p.lse + size_t(bh) * p.S;
      const float* delta_bh = p.delta + size_t(bh) * p.S;
      float lse_nx = 0.f, dl_nx = 0.f;
      if (j * kBlk + r < p.S) { lse_nx = lse_bh[j * kBlk + r]; dl_nx = delta_bh[j * kBlk + r]; }
      // Pair I writes dS buffer I&1; the MMAs of pair I-2 that read it have completed because dQ(I-2)
      // was drained (dq_full covers every earlier MMA) during pair I-1.  The drain of dQ(I-1) comes AFTER
      // this pair's element-wise work so its MMAs run underneath that work.
      for (int it = 0; it < niter; ++it) {
        const uint32_t I = I0 + it;
        const int i = j + it;
        const int qrow = i * kBlk + r;
        const bool diag = (it == 0);
        const bool row_ok = qrow < p.S;
        const float neg_lse = row_ok ? log2_inv_keep - lse_nx : -INFINITY;
        const float dl = row_ok ? dl_nx : 0.f;
        if (it + 1 < niter && qrow + kBlk < p.S) { lse_nx = lse_bh[qrow + kBlk]; dl_nx = delta_bh[qrow + kBlk]; }
        const unsigned long long grp_row = ((unsigned long long)bh * p.S + qrow) * p.S16;
        mbar_wait(sdp_full, I & 1);
        tc_fence_after();
        const uint32_t p_sub = p_sub0, ds_sub = ds_sub0 + (I & 1) * kPTile;
        // With q = 1/(1-p):  P~ = keep ? q P : 0  and  dS = P (dP~ - delta) / sqrt(d),  dP~ = keep ? q dP : 0.
        // Folding q into the exponent gives qP from one FFMA + MUFU, and
        //   dS = qP * (keep ? dP/sqrt(d) - c : -c),  c = delta (1-p) / sqrt(d)
        // so an element costs FFMA, MUFU, compare, 2 selects, FFMA, FMUL and its share of two packs.
        const float neg_c = -dl * keep_scale;
        // Two halves of 16 columns: 32 live accumulator registers instead of 64 (the kernel is capped at 96
        // registers per thread by its 18 warps), and the second half's TMEM loads fly under the first's math.
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          uint32_t sv[16], dv[16];
          __syncwarp();
          tmem_ld_32x32b_x16(tS + lane_off + qt * 32 + hh * 16, sv);
          tmem_ld_32x32b_x16(tdP + lane_off + qt * 32 + hh * 16, dv);
          tmem_ld_wait();
          if (hh == 1) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(s_free);   // S / dP are in registers: the next pair's may overwrite them
          }
          if (diag) {   // causal mask: only the diagonal tile pays for it (warp-uniform branch)
#pragma unroll
            for (int e = 0; e < 16; ++e)
              if ((qt * 32 + hh * 16 + e) > r) sv[e] = 0xFF800000u;   // -inf -> probability 0 -> dS 0
          }
          // one Philox call (sixteen 8-bit lanes) covers this half's 16 kv positions
          uint4 rb = make_uint4(0u, 0u, 0u, 0u);
          if (drop) rb = dropout_bits(seed, offs, p.rng_stream, grp_row + ((j * kBlk + qt * 32 + hh * 16) >> 4));
          // keep flags: bit 7 of every byte (four scores per word compared at once); all kept without dropout
          const uint32_t cm[4] = {drop ? bytes_ge_bit7(rb.x, t4) : 0x80808080u, drop ? bytes_ge_bit7(rb.y, t4) : 0x80808080u,
                                  drop ? bytes_ge_bit7(rb.z, t4) : 0x80808080u, drop ? bytes_ge_bit7(rb.w, t4) : 0x80808080u};
          const uint32_t negc_bits = __float_as_uint(neg_c);
#pragma unroll
          for (int g2 = 0; g2 < 2; ++g2) {
            const int g = hh * 2 + g2;
            float pd[8], ds[8];
#pragma unroll
            for (int e = 0; e < 8; e += 2) {   // pairs: FFMA2 (exponent), FFMA2 (dP scale - c), FMUL2 (dS)
              const float2 arg = ffma2(make_float2(__uint_as_float(sv[g2 * 8 + e]), __uint_as_float(sv[g2 * 8 + e + 1])),
                                       make_float2(kScale2, kScale2), make_float2(neg_lse, neg_lse));
              const float2 t = ffma2(make_float2(__uint_as_float(dv[g2 * 8 + e]), __uint_as_float(dv[g2 * 8 + e + 1])),
                                     make_float2(kScale, kScale), make_float2(neg_c, neg_c));
              const float qp0 = ex2_ftz(arg.x), qp1 = ex2_ftz(arg.y);
              // selects as logic ops on 32-bit masks expanded from the keep bits (prmt sign-replicate)
              const uint32_t cw = cm[g2 * 2 + (e >> 2)];
              const uint32_t m0 = (e & 2) ? byte_sign_mask<2>(cw) : byte_sign_mask<0>(cw);
              const uint32_t m1 = (e & 2) ? byte_sign_mask<3>(cw) : byte_sign_mask<1>(cw);
              const float t0 = __uint_as_float((__float_as_uint(t.x) & m0) | (negc_bits & ~m0));
              const float t1 = __uint_as_float((__float_as_uint(t.y) & m1) | (negc_bits & ~m1));
              const float2 d2 = fmul2(make_float2(qp0, qp1), make_float2(t0, t1));
              ds[e] = d2.x;
              ds[e + 1] = d2.y;
              pd[e] = __uint_as_float(__float_as_uint(qp0) & m0);
              pd[e + 1] = __uint_as_float(__float_as_uint(qp1) & m1);
            }
            const int cc = (qt & 1) * 4 + g;
            if (g == 0) {   // the reduce-add of dQ(I-2) has finished reading this dS buffer, dV(I-1) has read P~
              if (dq_leader) tma_store_wait_read<0>();
              named_bar_sync(pair_bar, 64);
              mbar_wait(p_free, (I & 1) ^ 1);
            }
            st_shared_v4(swz(p_sub, r, cc), pack_bf16x2(pd[0], pd[1]), pack_bf16x2(pd[2], pd[3]),
                         pack_bf16x2(pd[4], pd[5]), pack_bf16x2(pd[6], pd[7]));
            st_shared_v4(swz(ds_sub, r, cc), pack_bf16x2(ds[0], ds[1]), pack_bf16x2(ds[2], ds[3]),
                         pack_bf16x2(ds[4], ds[5]), pack_bf16x2(ds[6], ds[7]));
          }
        }
        tc_fence_before();
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&pds_ready[I & 1]);
        if (it > 0) drain_dq(I - 1, it - 1);
      }
      drain_dq(I0 + niter - 1, niter - 1);

      // dK / dV of this kv block: thread (r, qt) owns kv row r, head-dim columns [16 qt, 16 qt + 16)
      mbar_wait(acc_full, ti & 1);
      tc_fence_after();
      const int kvrow = j * kBlk + r;
      __nv_bfloat16* dk_dst = p.dqkv + size_t(row_base + kvrow) * 3 * h + h + hd * kD + qt * 16;
      __nv_bfloat16* dv_dst = dk_dst + h;
      {
        uint32_t kv_[16], vv_[16];
        __syncwarp();
        tmem_ld_32x32b_x16(tdK + lane_off + qt * 16, kv_);
        tmem_ld_32x32b_x16(tdV + lane_off + qt * 16, vv_);
        tmem_ld_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(acc_free);   // accumulators are in registers: the next tile may start
        if (p.tma_out) {
          // dK -> first half of the (now idle) P~ buffer, dV -> second half, as 128B-swizzled [128][64] bf16 tiles; the
          // four warps of a row quarter meet, then one thread stores the quarter's dK rows and another its dV rows
          // (the same threads that own those rows' later P~ writes wait for the store's read: see the pair barrier)
          const uint32_t dk_tile = smem_u32(smem + BwdSmem::kP), dv_tile = dk_tile + kTile;
#pragma unroll
          for (int g = 0; g < 2; ++g) {
            st_shared_v4(swz(dk_tile, r, qt * 2 + g),
                         pack_bf16x2(__uint_as_float(kv_[g * 8 + 0]), __uint_as_float(kv_[g * 8 + 1])),
                         pack_bf16x2(__uint_as_float(kv_[g * 8 + 2]), __uint_as_float(kv_[g * 8 + 3])),
                         pack_bf16x2(__uint_as_float(kv_[g * 8 + 4]), __uint_as_float(kv_[g * 8 + 5])),
                         pack_bf16x2(__uint_as_float(kv_[g * 8 + 6]), __uint_as_float(kv_[g * 8 + 7])));
            st_shared_v4(swz(dv_tile, r, qt * 2 + g),
                         pack_bf16x2(__uint_as_float(vv_[g * 8 + 0]), __uint_as_float(vv_[g * 8 + 1])),
                         pack_bf16x2(__uint_as_float(vv_[g * 8 + 2]), __uint_as_float(vv_[g * 8 + 3])),
                         pack_bf16x2(__uint_as_float(vv_[g * 8 + 4]), __uint_as_float(vv_[g * 8 + 5])),
                         pack_bf16x2(__uint_as_float(vv_[g * 8 + 6]), __uint_as_float(vv_[g * 8 + 7])));
          }
          fence_proxy_async_smem();
          named_bar_sync(9 + qd, 128);
          if (dq_leader) {   // qt = 0 stores dK (P~ sub-tile 0), qt = 2 stores dV (sub-tile 1)
            tma_store_2d(&tm_dkv, smem + BwdSmem::kP + (qt >> 1) * kTile + qd * 4096, (1 + (qt >> 1)) * h + hd * kD,
                         row_base + j * kBlk + qd * 32);
            tma_store_commit();
          }
        } else if (kvrow < p.S) {
#pragma unroll
          for (int g = 0; g < 2; ++g) {
            *reinterpret_cast<uint4*>(dk_dst + g * 8) = make_uint4(
                pack_bf16x2(__uint_as_float(kv_[g * 8 + 0]), __uint_as_float(kv_[g * 8 + 1])),
                pack_bf16x2(__uint_as_float(kv_[g * 8 + 2]), __uint_as_float(kv_[g * 8 + 3])),
                pack_bf16x2(__uint_as_float(kv_[g * 8 + 4]), __uint_as_float(kv_[g * 8 + 5])),
                pack_bf16x2(__uint_as_float(kv_[g * 8 + 6]), __uint_as_float(kv_[g * 8 + 7])));
            *reinterpret_cast<uint4*>(dv_dst + g * 8) = make_uint4(
                pack_bf16x2(__uint_as_float(vv_[g * 8 + 0]), __uint_as_float(vv_[g * 8 + 1])),
                pack_bf16x2(__uint_as_float(vv_[g * 8 + 2]), __uint_as_float(vv_[g * 8 + 3])),
                pack_bf16x2(__uint_as_float(vv_[g * 8 + 4]), __uint_as_float(vv_[g * 8 + 5])),
                pack_bf16x2(__uint_as_float(vv_[g * 8 + 6]), __uint_as_float(vv_[g * 8 + 7])));
          }
        }
      }
      I0 += niter;
      ++ti;
    }
  }

  if (warp >= 2 && (((warp - 2) >> 2) & 1) == 0 && lane == 0) tma_store_wait<0>();   // dQ reduce-adds have left shared memory
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<512>(tmem);
  }
}

}  // namespace b200

using namespace b200;
#define STREAM(s) reinterpret_cast<cudaStream_t>(s)

extern "C" {

int b200_attn_fwd(const void* qkv, void* out, float* lse, int batch, int seq, int heads, float dropout_p,
                  const unsigned long long* rng, unsigned int rng_stream, void* stream) {
  B200_REQUIRE(batch > 0 && seq > 0 && heads > 0, "attn_fwd: empty problem");
  B200_REQUIRE(dropout_p == 0.f || rng != nullptr, "attn_fwd: rng_state required when dropout_p > 0");
  const int h = heads * kD;
  CUtensorMap tm;
  int rc = make_tmap_2d_bf16(&tm, qkv, 3 * (uint64_t)h, (uint64_t)batch * seq, 3 * (uint64_t)h, 64, kBlk);
  if (rc) return rc;
  AttnParams p{};
  p.B = batch; p.S = seq; p.H = heads; p.S16 = (seq + 15) / 16; p.p_drop = dropout_p; p.rng = rng; p.rng_stream = rng_stream;
  p.out = (__nv_bfloat16*)out; p.lse = lse;
  p.tma_out = seq % kBlk == 0;
  CUtensorMap tm_out;   // 32-row boxes of the output (only used when whole tiles are in range)
  rc = make_tmap_2d_bf16(&tm_out, out, (uint64_t)h, (uint64_t)batch * seq, (uint64_t)h, 64, 32);
  if (rc) return rc;
  static bool configured[kMaxDevices] = {};   // the attribute is per device
  if (!configured[current_device()]) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, FwdSmem::kTotal));
    configured[current_device()] = true;
  }
  // persistent: one CTA per SM walks its share of the (q tile, head, batch) list
  const int tiles = ((seq + kBlk - 1) / kBlk) * heads * batch;
  dim3 grid(tiles < sm_count() ? tiles : sm_count());
  attn_fwd_kernel<<<grid, kAttnThreads, FwdSmem::kTotal, STREAM(stream)>>>(tm, tm_out, p);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

static int attn_bwd_impl(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, float* delta,
                         float* dq_accum, int batch, int seq, int heads, float dropout_p, const unsigned long long* rng,
                         unsigned int rng_stream, const float* rope_cos, const float* rope_sin, void* stream);

int b200_attn_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, float* delta,
                  float* dq_accum, int batch, int seq, int heads, float dropout_p, const unsigned long long* rng,
                  unsigned int rng_stream, void* stream) {
  return attn_bwd_impl(qkv, out, dout, lse, dqkv, delta, dq_accum, batch, seq, heads, dropout_p, rng, rng_stream, nullptr,
                       nullptr, stream);
}

int b200_attn_bwd_rope(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, float* delta,
                       float* dq_accum, int batch, int seq, int heads, float dropout_p, const unsigned long long* rng,
                       unsigned int rng_stream, const float* rope_cos, const float* rope_sin, void* stream) {
  B200_REQUIRE(rope_cos != nullptr && rope_sin != nullptr, "attn_bwd_rope: rotary tables required");
  return attn_bwd_impl(qkv, out, dout, lse, dqkv, delta, dq_accum, batch, seq, heads, dropout_p, rng, rng_stream, rope_cos,
                       rope_sin, stream);
}

static int attn_bwd_impl(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, float* delta,
                         float* dq_accum, int batch, int seq, int heads, float dropout_p, const unsigned long long* rng,
                         unsigned int rng_stream, const float* rope_cos, const float* rope_sin, void* stream) {
  B200_REQUIRE(batch > 0 && seq > 0 && heads > 0, "attn_bwd: empty problem");
  B200_REQUIRE(dropout_p == 0.f || rng != nullptr, "attn_bwd: rng_state required when dropout_p > 0");
  const int h = heads * kD;
  const long long rows = (long long)batch * seq;
  CUtensorMap tm_qkv, tm_do;
  int rc = make_tmap_2d_bf16(&tm_qkv, qkv, 3 * (uint64_t)h, (uint64_t)rows, 3 * (uint64_t)h, 64, kBlk);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&tm_do, dout, (uint64_t)h, (uint64_t)rows, (uint64_t)h, 64, kBlk);
  if (rc) return rc;
  CUtensorMap tm_dq;   // fp32 dQ accumulator, 32 x 32 boxes for the reduce-add
  rc = make_tmap_2d_sw128(&tm_dq, dq_accum, 4, (uint64_t)h, (uint64_t)rows, (uint64_t)h, 32);
  if (rc) return rc;
  CUtensorMap tm_dkv;  // 32-row boxes of the dK / dV columns of dqkv
  rc = make_tmap_2d_bf16(&tm_dkv, dqkv, 3 * (uint64_t)h, (uint64_t)rows, 3 * (uint64_t)h, 64, 32);
  if (rc) return rc;
  cudaStream_t s = STREAM(stream);
  const long long nd = rows * heads;
  if (rope_cos != nullptr) {
    attn_delta_zero_kernel<<<(unsigned)((nd * 8 + 255) / 256), 256, 0, s>>>(
        (const __nv_bfloat16*)out, (const __nv_bfloat16*)dout, delta, dq_accum, nd, seq, heads);
    B200_CHECK_CUDA(cudaGetLastError());
  } else {
    attn_delta_kernel<<<(unsigned)((nd + 255) / 256), 256, 0, s>>>((const __nv_bfloat16*)out, (const __nv_bfloat16*)dout,
                                                                   delta, batch, seq, heads);
    B200_CHECK_CUDA(cudaGetLastError());
    B200_CHECK_CUDA(cudaMemsetAsync(dq_accum, 0, size_t(rows) * h * sizeof(float), s));
  }
  AttnParams p{};
  p.B = batch; p.S = seq; p.H = heads; p.S16 = (seq + 15) / 16; p.p_drop = dropout_p; p.rng = rng; p.rng_stream = rng_stream;
  p.lse = const_cast<float*>(lse); p.delta = delta; p.dq_accum = dq_accum; p.dqkv = (__nv_bfloat16*)dqkv;
  p.tma_out = seq % kBlk == 0;
  static bool configured[kMaxDevices] = {};   // the attribute is per device
  if (!configured[current_device()]) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, BwdSmem::kTotal));
    configured[current_device()] = true;
  }
  const int tiles = ((seq + kBlk - 1) / kBlk) * heads * batch;   // persistent: one CTA per SM
  dim3 grid(tiles < sm_count() ? tiles : sm_count());
  attn_bwd_kernel<<<grid, kAttnThreads, BwdSmem::kTotal, s>>>(tm_qkv, tm_do, tm_dq, tm_dkv, p);
  B200_CHECK_CUDA(cudaGetLastError());
  if (rope_cos != nullptr) {
    const long long nt = rows * 2 * heads * 4;
    attn_dq_rope_kernel<<<(unsigned)((nt + 255) / 256), 256, 0, s>>>(dq_accum, (__nv_bfloat16*)dqkv, rope_cos, rope_sin, nt,
                                                                     heads, seq);
  } else {
    const long long ng = rows * (h / 8);
    attn_dq_cast_kernel<<<(unsigned)((ng + 255) / 256), 256, 0, s>>>(dq_accum, (__nv_bfloat16*)dqkv, rows, h);
  }
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"

// ================================================================================================
// Attention backward for callers that must not allocate device memory.
//
// b200_attn_bwd() above needs two scratch buffers from its caller (the softmax-gradient row sums and an
// fp32 dQ accumulator).  Not every caller can provide them: a host whose framework allocator already
// owns all of HBM, a serving loop where cudaMalloc is banned because it stalls the whole device, or a
// plain C program linked against this library with no allocator of its own.  For those callers the
// scratch lives inside the library image, in the fixed-size array below.  It is part of the module, so
// every GPU that loads libb200trainer.so carries these 32 MiB from load until the process exits, whether
// or not this entry point is ever called.
// ================================================================================================
#define B200_ATTN_BWD_NOALLOC_SCRATCH_BYTES (32u << 20)
static __device__ __align__(256) unsigned char g_attn_bwd_noalloc_scratch[B200_ATTN_BWD_NOALLOC_SCRATCH_BYTES];

namespace {

constexpr int kNoAllocMaxDevices = 64;
void* g_noalloc_scratch_addr[kNoAllocMaxDevices] = {nullptr};  // device address of the array, per device ordinal
bool g_noalloc_func_configured[kNoAllocMaxDevices] = {false};
std::mutex g_noalloc_mutex;

// In host code the identifier g_attn_bwd_noalloc_scratch is only a handle for the symbol, not a pointer a
// kernel may dereference; the real address differs per device and has to be asked for with that device
// current.  Done once per device, on the first call made while that device is current.
int noalloc_scratch_for_current_device(void** out) {
  int dev = -1;
  B200_CHECK_CUDA(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < kNoAllocMaxDevices, "attn_bwd_noalloc: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lock(g_noalloc_mutex);
  if (g_noalloc_scratch_addr[dev] == nullptr) {
    void* ptr = nullptr;
    B200_CHECK_CUDA(cudaGetSymbolAddress((void**)&ptr, g_attn_bwd_noalloc_scratch));
    B200_REQUIRE(ptr != nullptr, "attn_bwd_noalloc: cudaGetSymbolAddress returned a null address");
    g_noalloc_scratch_addr[dev] = ptr;
  }
  if (!g_noalloc_func_configured[dev]) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(b200::attn_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         b200::BwdSmem::kTotal));
    g_noalloc_func_configured[dev] = true;
  }
  *out = g_noalloc_scratch_addr[dev];
  return 0;
}

}  // namespace

extern "C" {

unsigned long long b200_attn_bwd_noalloc_max_scratch_bytes(void) { return B200_ATTN_BWD_NOALLOC_SCRATCH_BYTES; }

// Same result as b200_attn_bwd(); the caller passes inputs and outputs only.
//
//  * Size limit: batch*seq*heads*64*4 bytes (dQ accumulator) plus batch*heads*seq*4 bytes (row sums), each
//    rounded up to 256, must fit in 32 MiB, i.e. batch*seq*heads <= 129,055 -- 8 x 1024 tokens at 12 heads need 24.4 MiB and fit
//    (the cap at 12 heads is 10,754 tokens), 16 x 1024 do not.  Larger problems are refused with an error; nothing is launched.
//  * The first call on each device looks up the array's device address and raises the kernel's shared-memory
//    limit.  Make that first call during start-up, before any stream capture begins.
//  * All calls on a device share the one array and it still holds whatever the previous call left there.
//    The part that must start from zero (the dQ accumulator) is cleared on `stream` at the start of every
//    call; the row sums are fully rewritten.  Two calls whose GPU work can overlap -- different streams, or
//    several host threads -- race on the array and corrupt each other: keep them on one stream or
//    serialise them yourself.
//  * No cudaMalloc, cudaFree or other allocating call is made, here or in anything this function calls.
int b200_attn_bwd_noalloc(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv, int batch,
                          int seq, int heads, float dropout_p, const unsigned long long* rng, unsigned int rng_stream,
                          void* stream) {
  B200_REQUIRE(batch > 0 && seq > 0 && heads > 0, "attn_bwd_noalloc: empty problem");
  B200_REQUIRE(dropout_p == 0.f || rng != nullptr, "attn_bwd_noalloc: rng_state required when dropout_p > 0");
  const int h = heads * kD;
  const long long rows = (long long)batch * seq;
  const unsigned long long dq_bytes = ((unsigned long long)rows * h * sizeof(float) + 255ull) & ~255ull;
  const unsigned long long delta_bytes = ((unsigned long long)rows * heads * sizeof(float) + 255ull) & ~255ull;
  B200_REQUIRE(dq_bytes + delta_bytes <= B200_ATTN_BWD_NOALLOC_SCRATCH_BYTES,
               "attn_bwd_noalloc: batch=%d seq=%d heads=%d needs %llu bytes of scratch, the built-in array holds %u "
               "(use b200_attn_bwd with caller-provided scratch for larger problems)",
               batch, seq, heads, dq_bytes + delta_bytes, B200_ATTN_BWD_NOALLOC_SCRATCH_BYTES);
  void* base = nullptr;
  if (int rc = noalloc_scratch_for_current_device(&base)) return rc;
  float* dq_accum = reinterpret_cast<float*>(base);
  float* delta = reinterpret_cast<float*>(static_cast<unsigned char*>(base) + dq_bytes);

  CUtensorMap tm_qkv, tm_do;
  int rc = make_tmap_2d_bf16(&tm_qkv, qkv, 3 * (uint64_t)h, (uint64_t)rows, 3 * (uint64_t)h, 64, kBlk);
  if (rc) return rc;
  rc = make_tmap_2d_bf16(&tm_do, dout, (uint64_t)h, (uint64_t)rows, (uint64_t)h, 64, kBlk);
  if (rc) return rc;
  CUtensorMap tm_dq;   // the kernel reduce-adds dQ into the scratch through a tensor map (encoded on the host: no allocation)
  rc = make_tmap_2d_sw128(&tm_dq, dq_accum, 4, (uint64_t)h, (uint64_t)rows, (uint64_t)h, 32);
  if (rc) return rc;
  CUtensorMap tm_dkv;  // 32-row boxes of the dK / dV columns of dqkv
  rc = make_tmap_2d_bf16(&tm_dkv, dqkv, 3 * (uint64_t)h, (uint64_t)rows, 3 * (uint64_t)h, 64, 32);
  if (rc) return rc;
  cudaStream_t s = STREAM(stream);
  const long long nd = rows * heads;
  attn_delta_kernel<<<(unsigned)((nd + 255) / 256), 256, 0, s>>>((const __nv_bfloat16*)out, (const __nv_bfloat16*)dout,
                                                                 delta, batch, seq, heads);
  B200_CHECK_CUDA(cudaGetLastError());
  B200_CHECK_CUDA(cudaMemsetAsync(dq_accum, 0, size_t(rows) * h * sizeof(float), s));  // stale from the last call
  AttnParams p{};
  p.B = batch; p.S = seq; p.H = heads; p.S16 = (seq + 15) / 16; p.p_drop = dropout_p; p.rng = rng; p.rng_stream = rng_stream;
  p.lse = const_cast<float*>(lse); p.delta = delta; p.dq_accum = dq_accum; p.dqkv = (__nv_bfloat16*)dqkv;
  p.tma_out = seq % kBlk == 0;
  const int tiles = ((seq + kBlk - 1) / kBlk) * heads * batch;   // persistent: one CTA per SM
  dim3 grid(tiles < sm_count() ? tiles : sm_count());
  attn_bwd_kernel<<<grid, kAttnThreads, BwdSmem::kTotal, s>>>(tm_qkv, tm_do, tm_dq, tm_dkv, p);
  B200_CHECK_CUDA(cudaGetLastError());
  const long long ng = rows * (h / 8);
  attn_dq_cast_kernel<<<(unsigned)((ng + 255) / 256), 256, 0, s>>>(dq_accum, (__nv_bfloat16*)dqkv, rows, h);
  B200_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // extern "C"
