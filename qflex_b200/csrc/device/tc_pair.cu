// CTA-pair tensor-core contraction (tcgen05 cta_group::2) for the classes whose
// whole K fits one TMEM accumulation (k <= 8 bits) and whose N fits the 512 TMEM
// columns (n <= 8 bits): bristlecone_70 1-32-1's dominant 2^20 x 256 x 256
// expands and the d = 16 expands.
//
// Same math and operand handling as tc_fused.cu (reference: A.reorder +
// cblas_cgemm of src/tensor.cpp:1045-1194, 960-980; split-bf16 emulation of
// fp32, A gathered by TMA and converted in the kernel), different data flow:
//   * two CTAs of a cluster (one TPC) own a 256-row tile; every tcgen05.mma is
//     M = 256 across the pair, so each CTA stages only HALF of every B' tile
//     (bn/2 rows) -- the L2 -> SM traffic of the small operand halves;
//   * the n tiles are the INNER loop: a k block of A is fetched from HBM and
//     converted once and multiplied against every n tile (tc_fused re-fetches and
//     re-converts A per n tile); the accumulators of all n tiles of a tile live
//     in TMEM side by side (tiles_n * bn <= 512 columns);
//   * one accumulation per tile (the whole K, <= 512 real k: same error as the
//     chunked fold, tools/tc_error.py), drained per n tile so that the epilogue
//     of n tile 0 overlaps the MMAs of n tile 1 and the next tile's MMAs start
//     as soon as their accumulator is free.
// Only the leader CTA (cluster rank 0) issues MMAs.  Barriers the MMA warp
// waits on live in the leader's shared memory and are arrived on remotely by
// the peer (converters: A ready; B producer: expect_tx + TMA complete_tx;
// epilogue: accumulator drained); barriers that release buffers are signalled
// in both CTAs by a multicast tcgen05.commit.
#include <cstdio>

#include "tc_common.cuh"

namespace {

constexpr uint32_t PEER_MASK = 0xFEFFFFFFu;  // shared::cluster address of the pair's even CTA

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on a barrier of the leader CTA (works from either CTA of the pair)
__device__ __forceinline__ void mbar_arrive_leader(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];"
               ::"r"(bar & PEER_MASK)
               : "memory");
}
__device__ __forceinline__ void mbar_expect_tx_leader(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cluster.b64 _, [%0], %1;"
               ::"r"(bar & PEER_MASK), "r"(bytes)
               : "memory");
}
// Wait on a leader barrier the peer arrives on.  Default (CTA-scope) semantics on
// purpose: .acquire.cluster / .release.cluster compile to MEMBAR.ALL.GPU + CCTL.IVALL
// per poll (measured: 1.85 ms instead of ~1 ms for 2^20 x 256 x 256).  What crosses
// the CTA boundary is read by the async proxy (tcgen05.mma operands, made visible by
// fence.proxy.async before the arrive) or is TMEM, ordered by the tcgen05 fences.
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {
  uint32_t ok = 0, spins = 0;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    if (!ok && ++spins > (1u << 28)) __trap();
  } while (!ok);
}
// this CTA's half of a B' tile; completion is counted on the leader's barrier
__device__ __forceinline__ void tma_load_2d_pair(const CUtensorMap* map, uint32_t bar,
                                                 uint32_t dst, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar & PEER_MASK), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma2_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                           uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive (once the MMAs issued so far have completed) on the barrier at this
// offset in BOTH CTAs of the pair
__device__ __forceinline__ void umma2_commit(uint32_t bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64"
      " [%0], %1;"
      ::"r"(bar), "h"((uint16_t)3)
      : "memory");
}

// timing experiment (QFLEX_FUSED_DBG & 256): clocks the MMA issuer of pair 0 spends in
// each of its waits (lagged schedule only), written to g.part
#define MMA_TIMED(slot, stmt)                 \
  do {                                        \
    if (prof) {                               \
      const long long t0_ = clock64();        \
      stmt;                                   \
      wt[slot] += clock64() - t0_;            \
    } else {                                  \
      stmt;                                   \
    }                                         \
  } while (0)

// FusedArgs fields as in tc_fused.cu, except: tiles = number of 256-row pair
// tiles, b_slot_bytes = bn/2 * 128 (this CTA's half), acc_stages = number of
// accumulator SETS (one set = tiles_n accumulators of bn columns).
// STORE: 0 row-per-thread stores (any C layout), 1 quad-transposed (natural C, wide
// rows), 2 transposed C.  A template parameter so that an instantiation carries the
// registers of one store path only (the d = 16 classes lost 35 % when all three
// paths lived in one epilogue).
template <int STORE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(FUSED_THREADS, 1)
    tc_pair_bf16x3(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmBh,
                   const __grid_constant__ CUtensorMap tmBm, const __grid_constant__ FusedArgs g) {
  extern __shared__ uint8_t smem_raw[];
  // the dynamic window starts at the same offset in both CTAs, so this rounding
  // gives identical layouts (the MMA applies one descriptor to both CTAs)
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~(uintptr_t)1023);
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t a16_base = smem_base;                                   // sc x (hi, mid)
  const uint32_t b_base = a16_base + (uint32_t)g.sc * 2 * A16_TILE;      // sb slots
  const uint32_t a32_base = b_base + (uint32_t)g.sb * g.b_slot_bytes;    // sa stages
  const uint32_t bar_base = a32_base + (uint32_t)g.sa * g.a32_bytes;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + (bar_base - smem_base) + 8 * NUM_BARS);
  auto a_full = [&](int s) { return bar_base + 8u * s; };                       // local
  auto a_empty = [&](int s) { return bar_base + 8u * (MAX_SA + s); };           // local
  auto c_full = [&](int s) { return bar_base + 8u * (2 * MAX_SA + s); };        // leader's
  auto c_empty = [&](int s) { return bar_base + 8u * (2 * MAX_SA + MAX_SC + s); };  // both
  auto b_full = [&](int s) { return bar_base + 8u * (2 * MAX_SA + 2 * MAX_SC + s); };  // leader's
  auto b_empty = [&](int s) { return bar_base + 8u * (2 * MAX_SA + 2 * MAX_SC + MAX_SB + s); };
  auto tfull_bar = [&](int s) {                                                 // both
    return bar_base + 8u * (2 * MAX_SA + 2 * MAX_SC + 2 * MAX_SB + s);
  };
  auto tempty_bar = [&](int s) {                                                // leader's
    return bar_base + 8u * (2 * MAX_SA + 2 * MAX_SC + 2 * MAX_SB + MAX_ACC + s);
  };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int bn = g.bn;
  const int subs = g.dual ? 2 : g.tiles_n;  // accumulators per tile
  const int n_acc = g.acc_stages * subs;
  const uint32_t tmem_cols = (uint32_t)max(32, n_acc * bn);
  const int pair0 = (int)(blockIdx.x >> 1), pair_step = (int)(gridDim.x >> 1);
  // tile -> pair assignment: interleaved (tile t to pair t mod P) or blocked (each pair
  // a contiguous range of tiles)
  int t_begin = pair0, t_end = g.tiles, t_step = pair_step;
  if (g.blocked) {
    const int per = (g.tiles + pair_step - 1) / pair_step;
    t_begin = pair0 * per;
    t_end = min(g.tiles, t_begin + per);
    t_step = 1;
  }

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < MAX_SA; ++s) {
      mbar_init(a_full(s), 1);
      mbar_init(a_empty(s), g.one_arrive ? 1 : CONV_WARPS);
    }
    for (int s = 0; s < MAX_SC; ++s) {
      mbar_init(c_full(s), g.one_arrive ? 2 : 2 * CONV_WARPS);
      mbar_init(c_empty(s), 1);
    }
    for (int s = 0; s < MAX_SB; ++s) {
      mbar_init(b_full(s), 2);
      mbar_init(b_empty(s), 1);
    }
    for (int s = 0; s < MAX_ACC; ++s) {
      mbar_init(tfull_bar(s), 1);
      mbar_init(tempty_bar(s), g.one_arrive ? 2 : (g.dual_epi ? 2 * 4 : 2 * 8));
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    // the same warp of both CTAs allocates (and later frees) the pair's columns
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_u32(tmem_slot)), "r"(tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  cluster_sync_all();  // barriers of both CTAs initialised before any remote arrive
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < CONV_WARP0) {
    reg_dec<40>();
    if (warp == 0) {
      // ---------------- A producer (own 128 rows): as in tc_fused.cu
      const int m_sh = lane < g.n_m_hi ? g.a_m_hi[lane] - 1 : -1;
      const int k_sh = lane < g.n_k_hi ? g.a_k_hi[lane] - 1 : -1;
      int stage = 0;
      uint32_t phase = 0;
      // optional start stagger (ns per pair): spreads the pairs' load / store bursts
      if (g.stagger_ns > 0) {
        const long long t0 = clock64();
        const long long wait = (long long)g.stagger_ns * pair0 * 2;  // ~2 clocks per ns
        while (clock64() - t0 < wait) {}
      }
      for (int t = t_begin; t < t_end; t += t_step) {
        const unsigned mt = 2u * (unsigned)t + rank;
        const unsigned row_base =
            __reduce_or_sync(0xffffffffu, (m_sh >= 0 && ((mt >> lane) & 1u)) ? (1u << m_sh) : 0u);
        for (int kb = 0; kb < g.kblocks; ++kb) {
          const unsigned off16 =
              row_base | __reduce_or_sync(0xffffffffu, (k_sh >= 0 && (((unsigned)kb >> lane) & 1u))
                                                           ? (1u << k_sh)
                                                           : 0u);
          if (lane == 0) {
            mbar_wait_spin(a_empty(stage), phase ^ 1u);
            if (g.dbg & 64) {  // timing experiment: handshakes only, no A traffic
              mbar_arrive(a_full(stage));
            } else {
              mbar_expect_tx(a_full(stage), g.a32_bytes);
              tma_load_a(&tmA, g.a_rank, a_full(stage), a32_base + (uint32_t)stage * g.a32_bytes,
                         (int)off16);
            }
          }
          if (++stage == g.sa) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    } else if (warp == 2 && lane == 0) {
      // ---------------- B producer: this CTA's bn/2 rows of B'_mid then B'_hi of
      // every (k block, n tile); both CTAs' bytes are counted on the leader's barrier
      int slot = 0;
      uint32_t phase = 0;
      const int half_rows = bn >> 1;
      // B' is written by the split kernel launched just before this one; with a
      // programmatic dependent launch this kernel may have started before that one
      // finished (returns at once otherwise)
      asm volatile("griddepcontrol.wait;" ::: "memory");
      auto load_b = [&](int kb, int nt) {
#pragma unroll
        for (int part = 0; part < 2; ++part) {
          if (!g.b_resident) mbar_wait_spin(b_empty(slot), phase ^ 1u);
          mbar_expect_tx_leader(b_full(slot), g.b_slot_bytes);
          tma_load_2d_pair(part == 0 ? &tmBm : &tmBh, b_full(slot),
                           b_base + (uint32_t)slot * g.b_slot_bytes, kb * BK,
                           nt * bn + (int)rank * half_rows);
          if (++slot == g.sb) {
            slot = 0;
            phase ^= 1u;
          }
        }
      };
      for (int t = t_begin; t < t_end; t += t_step) {
        if (g.b_resident && t != t_begin) break;
        if (g.lag) {  // same order as the MMA issuer's lagged schedule (below)
          for (int kb = 0; kb < g.kblocks; ++kb) {
            load_b(kb, 0);
            if (kb >= 1) load_b(kb - 1, 1);
          }
          load_b(g.kblocks - 1, 1);
        } else {
          for (int kb = 0; kb < g.kblocks; ++kb)
            for (int nt = 0; nt < g.tiles_n; ++nt) load_b(kb, nt);
        }
      }
    } else if (warp == 1 && rank == 0) {
      // ---------------- MMA issuer (leader CTA): warp-uniform control flow, one
      // elected lane issues.  M = 256 over the pair, N = bn, K = 16 per instruction.
      const bool leader = elect_one();
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(bn >> 3) << 17) |
                             ((uint32_t)((2 * BM) >> 4) << 24);
      int cs = 0, bs = 0, set = 0;
      uint32_t cph = 0, bph = 0, set_phase = 0;
      const int nkk = (g.dbg & 4) ? 0 : g.nkk;
      if (g.dual) {
        // K = 16: one stage = two m tiles side by side in the bf16 rows (columns 0..31
        // and 32..63); B' (resident, slots 0 = mid, 1 = hi) is the same for both
        const uint64_t dBm = smem_desc_sw128(b_base);
        const uint64_t dBh = smem_desc_sw128(b_base + g.b_slot_bytes);
        mbar_wait_cluster(b_full(0), 0u);
        mbar_wait_cluster(b_full(1), 0u);
        const bool prof = (g.dbg & 256) && blockIdx.x == 0;
        long long wt[4] = {0, 0, 0, 0};
        const long long role_t0 = clock64();
        for (int t = t_begin; t < t_end; t += t_step) {
          const uint64_t dAh = smem_desc_sw128(a16_base + (uint32_t)cs * 2 * A16_TILE);
          const uint64_t dAm = dAh + (uint64_t)(A16_TILE >> 4);
          MMA_TIMED(3, mbar_wait_cluster(c_full(cs), cph));
          for (int h = 0; h < 2; ++h) {
            const int acc = set * 2 + h;
            MMA_TIMED(h, mbar_wait_cluster(tempty_bar(acc), set_phase ^ 1u));
            const uint32_t d_tmem = tmem_base + (uint32_t)(acc * bn);
            tc_fence_after();
            if (leader) {
              for (int kk = 0; kk < nkk; ++kk)
                umma2_bf16(d_tmem, dAh + 2 * (2 * h + kk), dBm + 2 * kk, idesc, kk == 0 ? 0u : 1u);
              for (int kk = 0; kk < nkk; ++kk)
                umma2_bf16(d_tmem, dAm + 2 * (2 * h + kk), dBh + 2 * kk, idesc, 1u);
              for (int kk = 0; kk < nkk; ++kk)
                umma2_bf16(d_tmem, dAh + 2 * (2 * h + kk), dBh + 2 * kk, idesc, 1u);
              if (h == 1) umma2_commit(c_empty(cs));
              // one "accumulators full" commit per stage: both m tiles are drained after
              // the second one's MMAs anyway (2^24 x 16 x 16: 0.910 -> 0.884 ms)
              if (h == 1) umma2_commit(tfull_bar(acc));
            }
          }
          if (++cs == g.sc) {
            cs = 0;
            cph ^= 1u;
          }
          if (++set == g.acc_stages) {
            set = 0;
            set_phase ^= 1u;
          }
        }
        if (prof && lane == 0) {
          long long* out = reinterpret_cast<long long*>(g.part);
          out[0] = clock64() - role_t0;
          for (int i = 0; i < 4; ++i) out[1 + i] = wt[i];
        }
      } else if (g.lag) {
        // Two n tiles (N' = 512: the whole TMEM): both accumulators used to complete
        // within one sub-step (12 MMAs, ~1500 clocks) of each other at the end of a tile,
        // and the next tile's first MMAs then waited for the drains -- ~3300 clocks of
        // TMEM reads per accumulator plus the stores of the first before the second is
        // touched: ~8100 of a tile's ~32900 clocks (tools: QFLEX_FUSED_DBG=256 runs,
        // profiles/pair_skeleton_r2z2.log).  Here n tile 1 runs ONE k block behind n
        // tile 0 and catches up at the end of the tile, so that accumulator 0 is
        // drained behind the last two sub-steps of n tile 1, and accumulator 1 behind
        // the first two of the next tile's n tile 0.  An A' stage lives one step longer
        // (released after n tile 1's use): sc >= 3.
        int csb = 0;  // stage of n tile 1's k block (already waited for by n tile 0)
        const bool prof = (g.dbg & 256) && blockIdx.x == 0;
        long long wt[4] = {0, 0, 0, 0};
        const long long role_t0 = clock64();
        auto sub_step = [&](int kb, int nt, int stage, bool last_use) {
          const uint64_t dAh = smem_desc_sw128(a16_base + (uint32_t)stage * 2 * A16_TILE);
          const uint64_t dAm = dAh + (uint64_t)(A16_TILE >> 4);
          const int acc = set * 2 + nt;
          if (kb == 0) MMA_TIMED(nt, mbar_wait_cluster(tempty_bar(acc), set_phase ^ 1u));
          const uint32_t d_tmem = tmem_base + (uint32_t)(acc * bn);
          const uint32_t first_acc = kb == 0 ? 0u : 1u;
          const uint64_t dBm = smem_desc_sw128(b_base + (uint32_t)bs * g.b_slot_bytes);
          MMA_TIMED(2, mbar_wait_cluster(b_full(bs), bph));
          tc_fence_after();
          if (leader) {
            for (int kk = 0; kk < nkk; ++kk)
              umma2_bf16(d_tmem, dAh + 2 * kk, dBm + 2 * kk, idesc, kk == 0 ? first_acc : 1u);
            if (!g.b_resident) umma2_commit(b_empty(bs));
          }
          if (++bs == g.sb) {
            bs = 0;
            bph ^= g.b_resident ? 0u : 1u;
          }
          const uint64_t dBh = smem_desc_sw128(b_base + (uint32_t)bs * g.b_slot_bytes);
          MMA_TIMED(2, mbar_wait_cluster(b_full(bs), bph));
          tc_fence_after();
          if (leader) {
            for (int kk = 0; kk < nkk; ++kk)
              umma2_bf16(d_tmem, dAm + 2 * kk, dBh + 2 * kk, idesc, 1u);
            for (int kk = 0; kk < nkk; ++kk)
              umma2_bf16(d_tmem, dAh + 2 * kk, dBh + 2 * kk, idesc, 1u);
            if (!g.b_resident) umma2_commit(b_empty(bs));
            if (last_use) umma2_commit(c_empty(stage));
            if (kb == g.kblocks - 1) umma2_commit(tfull_bar(acc));
          }
          if (++bs == g.sb) {
            bs = 0;
            bph ^= g.b_resident ? 0u : 1u;
          }
        };
        for (int t = t_begin; t < t_end; t += t_step) {
          for (int kb = 0; kb < g.kblocks; ++kb) {
            MMA_TIMED(3, mbar_wait_cluster(c_full(cs), cph));
            sub_step(kb, 0, cs, false);
            if (++cs == g.sc) {
              cs = 0;
              cph ^= 1u;
            }
            if (kb >= 1) {
              sub_step(kb - 1, 1, csb, true);
              if (++csb == g.sc) csb = 0;
            }
          }
          sub_step(g.kblocks - 1, 1, csb, true);
          if (++csb == g.sc) csb = 0;
          if (++set == g.acc_stages) {
            set = 0;
            set_phase ^= 1u;
          }
        }
        if (prof && lane == 0) {
          long long* out = reinterpret_cast<long long*>(g.part);
          out[0] = clock64() - role_t0;
          for (int i = 0; i < 4; ++i) out[1 + i] = wt[i];
        }
      } else
      for (int t = t_begin; t < t_end; t += t_step) {
        for (int kb = 0; kb < g.kblocks; ++kb) {
          const uint64_t dAh = smem_desc_sw128(a16_base + (uint32_t)cs * 2 * A16_TILE);
          const uint64_t dAm = dAh + (uint64_t)(A16_TILE >> 4);
          mbar_wait_cluster(c_full(cs), cph);
          for (int nt = 0; nt < g.tiles_n; ++nt) {
            const int acc = set * g.tiles_n + nt;
            if (kb == 0) mbar_wait_cluster(tempty_bar(acc), set_phase ^ 1u);
            const uint32_t d_tmem = tmem_base + (uint32_t)(acc * bn);
            const uint32_t first_acc = kb == 0 ? 0u : 1u;
            const uint64_t dBm = smem_desc_sw128(b_base + (uint32_t)bs * g.b_slot_bytes);
            mbar_wait_cluster(b_full(bs), bph);
            tc_fence_after();
            if (leader) {
              if (nkk == 4) {
#pragma unroll
                for (int kk = 0; kk < 4; ++kk)
                  umma2_bf16(d_tmem, dAh + 2 * kk, dBm + 2 * kk, idesc, kk == 0 ? first_acc : 1u);
              } else {
                for (int kk = 0; kk < nkk; ++kk)
                  umma2_bf16(d_tmem, dAh + 2 * kk, dBm + 2 * kk, idesc, kk == 0 ? first_acc : 1u);
              }
              if (!g.b_resident) umma2_commit(b_empty(bs));
            }
            if (++bs == g.sb) {
              bs = 0;
              bph ^= g.b_resident ? 0u : 1u;
            }
            const uint64_t dBh = smem_desc_sw128(b_base + (uint32_t)bs * g.b_slot_bytes);
            mbar_wait_cluster(b_full(bs), bph);
            tc_fence_after();
            if (leader) {
              if (nkk == 4) {
#pragma unroll
                for (int kk = 0; kk < 4; ++kk)
                  umma2_bf16(d_tmem, dAm + 2 * kk, dBh + 2 * kk, idesc, 1u);
#pragma unroll
                for (int kk = 0; kk < 4; ++kk)
                  umma2_bf16(d_tmem, dAh + 2 * kk, dBh + 2 * kk, idesc, 1u);
              } else {
                for (int kk = 0; kk < nkk; ++kk)
                  umma2_bf16(d_tmem, dAm + 2 * kk, dBh + 2 * kk, idesc, 1u);
                for (int kk = 0; kk < nkk; ++kk)
                  umma2_bf16(d_tmem, dAh + 2 * kk, dBh + 2 * kk, idesc, 1u);
              }
              if (!g.b_resident) umma2_commit(b_empty(bs));
              if (nt == g.tiles_n - 1) umma2_commit(c_empty(cs));
              if (kb == g.kblocks - 1) umma2_commit(tfull_bar(acc));  // this n tile is complete
            }
            if (++bs == g.sb) {
              bs = 0;
              bph ^= g.b_resident ? 0u : 1u;
            }
          }
          if (++cs == g.sc) {
            cs = 0;
            cph ^= 1u;
          }
        }
        if (++set == g.acc_stages) {
          set = 0;
          set_phase ^= 1u;
        }
      }
    }
  } else if (warp < EPI_WARP0) {
    reg_dec<64>();
    // ---------------- converters (own 128 rows): as in tc_fused.cu, but "A ready"
    // is signalled on the leader's barrier
    const unsigned ctid = threadIdx.x - CONV_WARP0 * 32;
    uint32_t src_t = 0, dst_t = 0;
#pragma unroll
    for (int b = 0; b < 8; ++b)
      if ((ctid >> b) & 1u) {
        src_t ^= g.cv_src[b];
        dst_t ^= g.cv_dst[b];
      }
    const uint32_t ok0 = g.cv_ok0, ok1 = g.cv_ok1;
    const bool vec = g.cv_vec != 0;
    const uint32_t flip = (vec && ((ctid >> 2) & 1u)) ? ok1 : 0u;
    int as = 0, cs = 0;
    uint32_t aph = 0, cph = 0;
    for (int t = t_begin; t < t_end; t += t_step) {
      for (int kb = 0; kb < g.kblocks; ++kb) {
        mbar_wait_spin(a_full(as), aph);
        mbar_wait_spin(c_empty(cs), cph ^ 1u);
        const uint32_t src = a32_base + (uint32_t)as * g.a32_bytes;
        const uint32_t hi_tile = a16_base + (uint32_t)cs * 2 * A16_TILE;
#pragma unroll
        for (int it = 0; it < 4; ++it) {
          if (it < g.cv_its && !(g.dbg & 1)) {
            const uint32_t s = src_t ^ g.cv_src_it[it];
            float2 e00, e10, e01, e11;  // e<k0><k1>
            if (vec) {
              const float4 p = lds128(src + (s ^ flip)), q = lds128(src + (s ^ flip ^ ok1));
              const float4 lo = flip ? q : p, hi = flip ? p : q;
              e00 = make_float2(lo.x, lo.y);
              e10 = make_float2(lo.z, lo.w);
              e01 = make_float2(hi.x, hi.y);
              e11 = make_float2(hi.z, hi.w);
            } else {
              e00 = lds64(src + s);
              e10 = lds64(src + (s ^ ok0));
              e01 = lds64(src + (s ^ ok1));
              e11 = lds64(src + (s ^ ok0 ^ ok1));
            }
            uint32_t h0, h1, h2, h3, m0, m1, m2, m3;
            split_pair(e00.x, e00.y, h0, m0);
            split_pair(e10.x, e10.y, h1, m1);
            split_pair(e01.x, e01.y, h2, m2);
            split_pair(e11.x, e11.y, h3, m3);
            const uint32_t d = hi_tile + (dst_t ^ g.cv_dst_it[it]);
            sts128(d, h0, h1, h2, h3);
            sts128(d + A16_TILE, m0, m1, m2, m3);
          }
        }
        fence_proxy_async();
        if (g.one_arrive) {
          // ONE (remote) arrive per CTA and stage instead of one per converter warp:
          // arrivals from the peer CTA are serialised at the leader's barrier, and 8 of
          // them per 16-KB stage bounded the K = 16 classes (tools/dbg_sweep.sh)
          asm volatile("bar.sync 1, %0;" ::"n"(CONV_WARPS * 32) : "memory");
          if (ctid == 0) {
            mbar_arrive_leader(c_full(cs));
            mbar_arrive(a_empty(as));
          }
        } else {
          __syncwarp();
          if (lane == 0) {
            mbar_arrive_leader(c_full(cs));
            mbar_arrive(a_empty(as));
          }
        }
        if (++as == g.sa) {
          as = 0;
          aph ^= 1u;
        }
        if (++cs == g.sc) {
          cs = 0;
          cph ^= 1u;
        }
      }
    }
  } else {
    reg_inc<152>();
    // ---------------- epilogue (own 128 rows of the pair tile): per n tile, TMEM ->
    // registers, release the accumulator, then C rows with 32-byte stores
    const int lane_group = warp & 3;
    const int half = (warp - EPI_WARP0) >> 2;
    const int row = lane_group * 32 + lane;
    unsigned m_local = 0;
#pragma unroll
    for (int i = 0; i < 7; ++i) m_local |= (unsigned)((row >> i) & 1) << g.row_perm[i];
    const int half_cols = bn >> 1;
    const int npieces = half_cols >> 4;
    int set = 0;
    uint32_t set_phase = 0;
    const unsigned long long N = 1ull << g.n_bits;
    // quad-transposed stores: the tile rows of the four lanes of this lane's quad
    unsigned mq[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = lane_group * 32 + (lane & ~3) + i;
      mq[i] = 0;
#pragma unroll
      for (int b = 0; b < 7; ++b) mq[i] |= (unsigned)((r >> b) & 1) << g.row_perm[b];
    }
    const bool quad_stores = STORE == 1;
    TStore ts{};
    const bool t_stores = STORE == 2 && tstore_setup(g, g.C, row, lane, &ts);
    unsigned m_local_b = 0;  // first of the 4 memory-adjacent rows of this lane's group
    if (STORE == 2 && t_stores) {
#pragma unroll
      for (int i = 0; i < 7; ++i) m_local_b |= (unsigned)((ts.rowb >> i) & 1) << g.row_perm[i];
    }
    if (STORE == 0 && g.dual_epi) {
      // K = 16 (dual stages): the two accumulators of a stage are two m tiles that become
      // ready together.  The kernel is bound by the epilogue's per-accumulator round trip
      // (wait, TMEM load, release, ~1300 clocks of mostly fixed cost for 16 floats per
      // thread; profiles/pair_skeleton_r2z*.log: with every load, MMA and store disabled
      // the handshakes alone take 0.70 of the 0.88 ms).  So warps 0-3 drain accumulator
      // 0 and warps 4-7 accumulator 1, each thread a WHOLE row (bn floats): one round
      // trip per warp and stage instead of two.
      const int h = half;
      const int np = bn >> 4;
      const float scale = g.scale_full;
      for (int t = t_begin; t < t_end; t += t_step) {
        const unsigned long long mtile = (2ull * (unsigned long long)t + rank) * 2 + (unsigned)h;
        const unsigned long long m = mtile * BM + m_local;
        unsigned long long row_off = 0;
        if (g.c_natural) {
          row_off = m * N;
        } else {
          for (int q = 0; q < g.m_bits; ++q) row_off |= ((m >> q) & 1ull) << g.c_m[q];
        }
        const int acc = set * 2 + h;
        float sum[8][16];
        mbar_wait_spin(tfull_bar(set * 2 + 1), set_phase);
        tc_fence_after();
        const uint32_t t_row = tmem_base + ((uint32_t)(lane_group * 32) << 16) + (uint32_t)(acc * bn);
#pragma unroll
        for (int p = 0; p < 8; ++p)
          if (p < np && !(g.dbg & 8)) tmem_ld16_async(t_row + (uint32_t)(p * 16), sum[p]);
        tmem_wait_ld();
#pragma unroll
        for (int p = 0; p < 8; ++p)
          if (p < np && !(g.dbg & 8)) tmem_reg_fence(sum[p]);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_leader(tempty_bar(acc));
#pragma unroll
        for (int p = 0; p < 8; ++p)
          if (p < np) {
#pragma unroll
            for (int j = 0; j < 16; ++j) sum[p][j] *= scale;
          }
        if (g.c_natural && np >= 2 && !(g.dbg & 128)) {
          // natural C: transpose inside every quad so that each store instruction
          // writes full 128-byte lines (8 per warp) instead of 32 sectors in 32 lines
          float* tile_c = reinterpret_cast<float*>(g.C + mtile * BM * N);
#pragma unroll
          for (int pp = 0; pp < 4; ++pp) {
            if (2 * pp < np && !(g.dbg & 2)) {
              quad_transpose(sum[2 * pp], sum[2 * pp + 1], lane);
              float* dst = tile_c + 8 * (lane & 3) + pp * 32;
              stg256(dst + 2ull * mq[0] * N, &sum[2 * pp][0]);
              stg256(dst + 2ull * mq[1] * N, &sum[2 * pp][8]);
              stg256(dst + 2ull * mq[2] * N, &sum[2 * pp + 1][0]);
              stg256(dst + 2ull * mq[3] * N, &sum[2 * pp + 1][8]);
            }
          }
        } else
#pragma unroll
        for (int p = 0; p < 8; ++p) {
          if (p < np && !(g.dbg & 2)) {
            const unsigned long long n0 = (unsigned long long)(p * 8);
            if (g.c_natural) {
              float* dst = reinterpret_cast<float*>(g.C + row_off + n0);
              stg256(dst, &sum[p][0]);
              stg256(dst + 8, &sum[p][8]);
            } else {
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const unsigned long long n = n0 + j;
                unsigned long long co = 0;
                for (int q = 0; q < g.n_bits; ++q) co |= ((n >> q) & 1ull) << g.c_n[q];
                g.C[row_off + co] = make_float2(sum[p][2 * j], sum[p][2 * j + 1]);
              }
            }
          }
        }
        if (++set == g.acc_stages) {
          set = 0;
          set_phase ^= 1u;
        }
      }
    } else
    for (int t = t_begin; t < t_end; t += t_step) {
      for (int sub = 0; sub < subs; ++sub) {
        // dual: the two accumulators are two m tiles; else the n tiles of one m tile
        const int nt = g.dual ? 0 : sub;
        const unsigned long long mtile =
            g.dual ? (2ull * (unsigned long long)t + rank) * 2 + sub
                   : 2ull * (unsigned long long)t + rank;
        const unsigned long long m = mtile * BM + m_local;
        unsigned long long row_off = 0;
        if (g.c_natural) {
          row_off = m * N;
        } else {
          for (int q = 0; q < g.m_bits; ++q) row_off |= ((m >> q) & 1ull) << g.c_m[q];
        }
        const int acc = set * subs + sub;
        float sum[8][16];
        mbar_wait_spin(tfull_bar(g.dual ? set * 2 + 1 : acc), set_phase);
        tc_fence_after();
        const uint32_t t_row = tmem_base + ((uint32_t)(lane_group * 32) << 16) +
                               (uint32_t)(acc * bn + half * half_cols);
#pragma unroll
        for (int p = 0; p < 8; ++p)
          if (p < npieces && !(g.dbg & 8)) tmem_ld16_async(t_row + (uint32_t)(p * 16), sum[p]);
        tmem_wait_ld();
#pragma unroll
        for (int p = 0; p < 8; ++p)
          if (p < npieces && !(g.dbg & 8)) tmem_reg_fence(sum[p]);
        tc_fence_before();
        if (g.one_arrive) {
          asm volatile("bar.sync 2, %0;" ::"n"(8 * 32) : "memory");
          if (warp == EPI_WARP0 && lane == 0) mbar_arrive_leader(tempty_bar(acc));
        } else {
          __syncwarp();
          if (lane == 0) mbar_arrive_leader(tempty_bar(acc));
        }
        {
          // accumulation-bias compensation: registers only, so AFTER the accumulator is
          // released (the MMA issuer waits for that release at the start of a tile)
          const float scale = g.scale_full;
#pragma unroll
          for (int p = 0; p < 8; ++p)
            if (p < npieces) {
#pragma unroll
              for (int j = 0; j < 16; ++j) sum[p][j] *= scale;
            }
        }
        if (STORE == 2 && t_stores) {
          const unsigned long long mb = mtile * BM + m_local_b;
          unsigned long long rb = 0;
          for (int q = 0; q < g.m_bits; ++q) rb |= ((mb >> q) & 1ull) << g.c_m[q];
          const unsigned long long n_base =
              ((unsigned long long)nt * bn + half * half_cols) >> 1;
#pragma unroll
          for (int p = 0; p < 8; ++p) {
            if (p < npieces && !(g.dbg & 2)) {
              unsigned long long co[2];
#pragma unroll
              for (int g4 = 0; g4 < 2; ++g4) {
                const unsigned long long n = n_base + p * 8 + g4 * 4;
                co[g4] = ts.coj;
                for (int q = 2; q < g.n_bits; ++q) co[g4] |= ((n >> q) & 1ull) << g.c_n[q];
              }
              tstore_piece(sum[p], g.C + rb + co[0], g.C + rb + co[1], ts.xm0, ts.xm1, lane);
            }
          }
          continue;
        }
        if (STORE == 1 && quad_stores) {
          float* tile_c = reinterpret_cast<float*>(g.C + mtile * BM * N);
          const int col0 = nt * bn + half * half_cols + 8 * (lane & 3);  // float column
#pragma unroll
          for (int pp = 0; pp < 4; ++pp) {
            if (2 * pp < npieces && !(g.dbg & 2)) {
              quad_transpose(sum[2 * pp], sum[2 * pp + 1], lane);
              float* dst = tile_c + col0 + pp * 32;
              stg256(dst + 2ull * mq[0] * N, &sum[2 * pp][0]);
              stg256(dst + 2ull * mq[1] * N, &sum[2 * pp][8]);
              stg256(dst + 2ull * mq[2] * N, &sum[2 * pp + 1][0]);
              stg256(dst + 2ull * mq[3] * N, &sum[2 * pp + 1][8]);
            }
          }
          continue;
        }
#pragma unroll
        for (int p = 0; p < 8; ++p) {
          if (p < npieces && !(g.dbg & 2)) {
            const unsigned long long n0 =
                ((unsigned long long)nt * bn + half * half_cols + p * 16) >> 1;
            if (g.dbg & 16) {
              // timing experiment: same bytes into the same tile, but every warp store
              // covers 1 KB of contiguous memory (results are wrong)
              const unsigned long long tile0 =
                  ((2ull * (unsigned long long)t + rank) * BM + lane_group * 32) * N;
              float* dst = reinterpret_cast<float*>(g.C + tile0) +
                           (size_t)((nt * 2 + half) * 16 + p * 2) * 256 + lane * 8;
              stg256(dst, &sum[p][0]);
              stg256(dst + 256, &sum[p][8]);
            } else if (g.c_natural) {
              float* dst = reinterpret_cast<float*>(g.C + row_off + n0);
              stg256(dst, &sum[p][0]);
              stg256(dst + 8, &sum[p][8]);
            } else {
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const unsigned long long n = n0 + j;
                unsigned long long co = 0;
                for (int q = 0; q < g.n_bits; ++q) co |= ((n >> q) & 1ull) << g.c_n[q];
                g.C[row_off + co] = make_float2(sum[p][2 * j], sum[p][2 * j + 1]);
              }
            }
          }
        }
      }
      if (++set == g.acc_stages) {
        set = 0;
        set_phase ^= 1u;
      }
    }
  }

  // both CTAs are done with each other's shared memory and with TMEM
  __syncwarp();
  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                 "r"(tmem_cols)
                 : "memory");
  }
}

}  // namespace

// Roles already swapped (m_bits >= n_bits).  *done = false: not applicable (the
// caller tries tc_fused next).  `scratch` = where B' hi/mid go: nullptr = the context's
// workspace (grown on demand), else a caller-owned buffer of >= 2 * N * K * 8 bytes
// (qfx_contract_pair_static below).
static int launch_pair(qfx_ctx* ctx, const qfx_contract_desc& din, const float2* A,
                       const float2* B, float2* C, bool* done, uint8_t* scratch) {
  *done = false;
  qfx_contract_desc d = din;
  if (d.m_bits < 8 || d.n_bits < 4 || d.k_bits < 4 || d.n_bits > 8 || d.k_bits > 8) return QFX_OK;
  const uint64_t M = 1ull << d.m_bits, N = 1ull << d.n_bits, K = 1ull << d.k_bits;
  const int n_pairs = ctx->sm_count / 2;
  // enough 256-row tiles to occupy every pair (smaller problems: tc_fused's 128-row tiles)
  if ((int)(M / (2 * BM)) < n_pairs) return QFX_OK;
  {
    std::vector<int> idx(d.k_bits);
    for (int i = 0; i < d.k_bits; ++i) idx[i] = i;
    std::sort(idx.begin(), idx.end(), [&](int x, int y) { return d.a_k[x] < d.a_k[y]; });
    int8_t ak[QFX_MAX_BITS], bk[QFX_MAX_BITS];
    for (int i = 0; i < d.k_bits; ++i) {
      ak[i] = d.a_k[idx[i]];
      bk[i] = d.b_k[idx[i]];
    }
    memcpy(d.a_k, ak, d.k_bits);
    memcpy(d.b_k, bk, d.k_bits);
  }
  FusedArgs g{};
  // K = 16 (d = 16 bonds, HBM-bound): a 128-row stage is only 16 KB and the per-stage
  // handshakes, not HBM, bound the kernel (~1200 clocks per stage whatever its size).
  // Dual mode doubles the stage: m bit 7 rides along as a fifth k bit of the A tile,
  // so the bf16 rows hold two m tiles side by side, multiplied into two accumulators.
  static const bool dual_off = [] {
    const char* e = getenv("QFLEX_PAIR_DUAL");
    return e != nullptr && atoi(e) == 0;
  }();
  g.dual = (!dual_off && d.k_bits == 4 && d.n_bits <= 6 && (int)(M / (4 * BM)) >= n_pairs) ? 1 : 0;
  g.kb_bits = g.dual ? 5 : std::min(5, (int)d.k_bits);
  g.tile_bits = 7 + g.kb_bits;
  uint64_t span = 0;
  for (int i = 0; i < d.m_bits; ++i) span |= 1ull << d.a_m[i];
  for (int i = 0; i < d.k_bits; ++i) span |= 1ull << d.a_k[i];
  span += 1;
  CUtensorMap mA, mBh, mBm;
  {
    qfx_contract_desc dm = d;
    if (g.dual) dm.a_k[4] = d.a_m[7];  // the tile's fifth "k" bit
    if (!make_map_a(&mA, A, dm, g.kb_bits, span, &g)) return QFX_OK;
  }

  g.C = C;
  g.m_bits = d.m_bits;
  g.n_bits = d.n_bits;
  g.kblocks = (int)std::max<uint64_t>(1, (2 * K) / BK);
  g.nkk = (int)std::min<uint64_t>(4, (2 * K) / 16);
  g.chunk = g.kblocks;
  g.nchunks = 1;
  g.scale_full = g.scale_last = tc_debias_scale(3 * g.nkk * g.kblocks, ctx->tc_debias);
  // UMMA N per instruction: 256, except N' = 256 where two n tiles of 128 (two
  // accumulators draining alternately) beat one of 256 (2^20 x 128 x 256: 0.61 ->
  // 0.51 ms).  For N' = 512, 128-wide tiles (8-KB B' slots, twice as many) measured
  // 2 % slower than 256.
  int bn_max = (2 * N == 256) ? 128 : MAX_BN;
  if (const char* e = getenv("QFLEX_PAIR_BN")) bn_max = std::max(32, std::min(MAX_BN, atoi(e)));
  g.bn = (int)std::min<uint64_t>(2 * N, (uint64_t)bn_max);
  g.tiles_n = (int)((2 * N) / g.bn);
  g.tiles = (int)(M / (2 * BM));
  if (g.dual) {
    g.tiles = (int)(M / (4 * BM));
    g.kblocks = 1;
    g.chunk = 1;
  }
  g.c_natural = 1;
  for (int i = 0; i < d.n_bits; ++i)
    if (d.c_n[i] != i) g.c_natural = 0;
  for (int i = 0; i < d.m_bits; ++i)
    if (d.c_m[i] != d.n_bits + i) g.c_natural = 0;
  if ((reinterpret_cast<uintptr_t>(C) & 31) != 0) g.c_natural = 0;
  memcpy(g.c_m, d.c_m, QFX_MAX_BITS);
  memcpy(g.c_n, d.c_n, QFX_MAX_BITS);
  const int m_lo = g.dual ? 8 : 7;  // m bits inside a stage
  g.n_m_hi = d.m_bits - m_lo;
  for (int i = 0; i < g.n_m_hi; ++i) g.a_m_hi[i] = d.a_m[m_lo + i];
  g.n_k_hi = g.dual ? 0 : d.k_bits - g.kb_bits;
  for (int i = 0; i < g.n_k_hi; ++i) g.a_k_hi[i] = d.a_k[g.kb_bits + i];

  // shared memory plan per CTA: bf16 A stages, half-tile B slots, fp32 A stages
  g.a32_bytes = 8u << g.tile_bits;
  g.b_slot_bytes = (uint32_t)(g.bn / 2) * 128u;
  const int subs = g.dual ? 2 : g.tiles_n;  // accumulators per tile
  g.acc_stages = std::max(1, std::min(MAX_ACC / subs, 512 / (g.bn * subs)));
  const int fixed = SMEM_FIXED;
  const int slot = (int)g.b_slot_bytes, a32 = (int)g.a32_bytes, a16 = 2 * A16_TILE;
  const int slots_per_kb = 2 * g.tiles_n;
  static const int lag_env = [] {
    const char* e = getenv("QFLEX_PAIR_LAG");
    return e ? atoi(e) : 1;
  }();
  // (only when the two accumulators fill TMEM: with N' = 256 there are two accumulator
  // sets and the lag only costs B' slots -- 2^20 x 128 x 256: 0.51 -> 0.60 ms)
  g.lag = (lag_env != 0 && !g.dual && g.tiles_n == 2 && g.acc_stages == 1 && g.kblocks >= 2) ? 1 : 0;
  g.sc = g.lag ? 3 : 2;
  g.sa = 2;
  g.sb = slots_per_kb;
  g.b_resident = (slots_per_kb * g.kblocks <= MAX_SB) ? 1 : 0;
  if (g.b_resident) g.sb = slots_per_kb * g.kblocks;
  auto total = [&]() { return fixed + g.sc * a16 + g.sb * slot + g.sa * a32; };
  if (g.b_resident && total() > SMEM_LIMIT) {  // all of B' does not fit: stream it
    g.b_resident = 0;
    g.sb = slots_per_kb;
  }
  if (g.dual && !g.b_resident) return QFX_OK;  // (cannot happen: 2 slots of <= 8 KB)
  if (g.lag && (g.sb > MAX_SB || total() > SMEM_LIMIT)) {  // no room for the third A' stage
    g.lag = 0;
    g.sc = 2;
  }
  if (g.sb > MAX_SB || total() > SMEM_LIMIT) return QFX_OK;
  for (bool grew = true; grew;) {
    grew = false;
    if (!g.b_resident && g.sb < MAX_SB && total() + slot <= SMEM_LIMIT) { ++g.sb; grew = true; }
    if (g.sa < MAX_SA && total() + a32 <= SMEM_LIMIT) { ++g.sa; grew = true; }
    if (g.sc < MAX_SC && g.sc < g.sa && total() + a16 <= SMEM_LIMIT) { ++g.sc; grew = true; }
  }
  if (const char* e = getenv("QFLEX_PAIR_BLOCKED")) g.blocked = atoi(e);
  // (experiment, off: one barrier arrive per CTA instead of one per warp -- fewer remote
  // arrivals, but the extra named barrier costs more than they do: 0.892 -> 0.911 ms)
  g.one_arrive = 0;
  if (const char* e = getenv("QFLEX_PAIR_ONE_ARRIVE")) g.one_arrive = atoi(e) != 0;
  // start stagger: the pairs otherwise run in lockstep and hit HBM with synchronized
  // load / store bursts (2^20 x 256 x 256: 0.947 -> 0.926 ms); only when a pair has
  // enough tiles for the start delay (<= one tile time for the last pair) to pay off
  g.stagger_ns = (g.kblocks >= 4 && g.tiles >= 16 * n_pairs) ? 120 : 0;
  if (const char* e = getenv("QFLEX_PAIR_STAGGER")) g.stagger_ns = atoi(e);
  // tuning overrides (exact stage counts; rejected when they do not fit)
  if (getenv("QFLEX_PAIR_SA") || getenv("QFLEX_PAIR_SB") || getenv("QFLEX_PAIR_SC")) {
    const int sa0 = g.sa, sb0 = g.sb, sc0 = g.sc;
    if (const char* e = getenv("QFLEX_PAIR_SA")) g.sa = std::max(1, std::min(MAX_SA, atoi(e)));
    if (const char* e = getenv("QFLEX_PAIR_SC")) g.sc = std::max(1, std::min(MAX_SC, atoi(e)));
    if (const char* e = getenv("QFLEX_PAIR_SB"))
      if (!g.b_resident) g.sb = std::max(slots_per_kb, std::min(MAX_SB, atoi(e)));
    if (total() > SMEM_LIMIT) { g.sa = sa0; g.sb = sb0; g.sc = sc0; }
  }
  if (const char* e = getenv("QFLEX_PAIR_ACC")) g.acc_stages = std::max(1, std::min(g.acc_stages, atoi(e)));
  const int smem_bytes = total();
  if (const char* e = getenv("QFLEX_FUSED_DBG")) g.dbg = atoi(e);

  // workspace: B' hi/mid (2N x 2K bf16 each)
  const size_t b_bytes = (size_t)N * K * 8;
  int rc = QFX_OK;
  uint8_t* ws = scratch;
  if (ws == nullptr) {
    rc = qfx_ensure_ws(ctx, 2 * b_bytes + 4096);
    if (rc) return rc;
    ws = reinterpret_cast<uint8_t*>(ctx->ws);
  }
  uint32_t* b_hi = reinterpret_cast<uint32_t*>(ws);
  uint32_t* b_mid = reinterpret_cast<uint32_t*>(ws + b_bytes);
  rc = qfx_launch_split(ctx, B, 1, d.b_n, d.n_bits, d.b_k, d.k_bits, b_hi, b_mid);
  if (rc) return rc;
  if (!make_map_b(&mBh, b_hi, 2 * N, 2 * K, (uint32_t)(g.bn / 2)) ||
      !make_map_b(&mBm, b_mid, 2 * N, 2 * K, (uint32_t)(g.bn / 2))) {
    qfx_set_error("cuTensorMapEncodeTiled failed");
    return QFX_ERR_CUDA;
  }
  static qfx_per_device_once attr;
  if (attr.need(ctx->device)) {
    QFX_CUDA(cudaFuncSetAttribute(tc_pair_bf16x3<0>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  SMEM_LIMIT));
    QFX_CUDA(cudaFuncSetAttribute(tc_pair_bf16x3<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  SMEM_LIMIT));
    QFX_CUDA(cudaFuncSetAttribute(tc_pair_bf16x3<2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  SMEM_LIMIT));
    attr.mark(ctx->device);
  }
  const int grid = 2 * std::min(g.tiles, n_pairs);
  static const bool dual_epi_off = [] {
    const char* e = getenv("QFLEX_PAIR_DUAL_EPI");
    return e != nullptr && atoi(e) == 0;
  }();
  g.dual_epi = (g.dual && !g.one_arrive && !dual_epi_off && pick_store_mode(g, C) == 0) ? 1 : 0;
  static long long* prof_buf = nullptr;  // timing experiment only (QFLEX_FUSED_DBG & 256)
  if (scratch != nullptr) g.dbg &= ~256;  // (the no-allocation entry point never allocates)
  if (g.dbg & 256) {
    if (prof_buf == nullptr) QFX_CUDA(cudaMalloc(&prof_buf, 8 * sizeof(long long)));
    QFX_CUDA(cudaMemsetAsync(prof_buf, 0, 8 * sizeof(long long), ctx->stream));
    g.part = reinterpret_cast<float2*>(prof_buf);
  }
  {
    // launched as a programmatic dependent of the split kernel: the prologue and the
    // first A loads overlap the split's tail and the launch latency (QFLEX_PDL=0: plain)
    static const bool pdl = [] {
      const char* e = getenv("QFLEX_PDL");
      return e == nullptr || atoi(e) != 0;
    }();
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)grid);
    cfg.blockDim = dim3(FUSED_THREADS);
    cfg.dynamicSmemBytes = (size_t)smem_bytes;
    cfg.stream = ctx->stream;
    cudaLaunchAttribute attrs[1];
    attrs[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attrs[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attrs;
    cfg.numAttrs = pdl ? 1 : 0;
    switch (pick_store_mode(g, C)) {
      case 1: QFX_CUDA(cudaLaunchKernelEx(&cfg, tc_pair_bf16x3<1>, mA, mBh, mBm, g)); break;
      case 2: QFX_CUDA(cudaLaunchKernelEx(&cfg, tc_pair_bf16x3<2>, mA, mBh, mBm, g)); break;
      default: QFX_CUDA(cudaLaunchKernelEx(&cfg, tc_pair_bf16x3<0>, mA, mBh, mBm, g));
    }
  }
  QFX_CUDA(cudaGetLastError());
  if ((g.dbg & 256) && (g.lag || g.dual)) {
    long long h[8];
    QFX_CUDA(cudaStreamSynchronize(ctx->stream));
    QFX_CUDA(cudaMemcpy(h, prof_buf, sizeof(h), cudaMemcpyDeviceToHost));
    fprintf(stderr, "pair-prof MMA issuer: %lld clk; waits: accumulator 0 %lld, accumulator 1 %lld, "
            "B slots %lld, A stage %lld\n", h[0], h[1], h[2], h[3], h[4]);
  }
  ctx->stats.launches_gemm_tc++;
  ctx->stats.launches_gemm_tc_fused++;
  ctx->stats.launches_gemm_tc_pair++;
  ctx->stats.launches_total++;
  *done = true;
  return QFX_OK;
}

int qfx_launch_contract_tc_pair(qfx_ctx* ctx, const qfx_contract_desc& din, const float2* A,
                                const float2* B, float2* C, bool* done) {
  *done = false;
  static const bool disabled = [] {
    const char* e = getenv("QFLEX_TC_PAIR");
    return e != nullptr && atoi(e) == 0;
  }();
  if (disabled) return QFX_OK;
  return launch_pair(ctx, din, A, B, C, done, nullptr);
}

// ---------------------------------------------------------------- no-allocation entry
// qfx_contract_pair_static: the CTA-pair kernel above with its scratch (the small
// operand split into bf16 hi / mid, 2 x N x K x 8 bytes) in a file-scope device array
// instead of the context's cudaMalloc'ed workspace, for callers that may not allocate
// (CUDA graph capture, serving loops).  No counterpart in the reference.
// Every GPU that loads the library holds the array
// (1 MiB) until exit.  The split kernel rewrites every byte the pair kernel reads, so
// nothing has to start at zero; all calls share the array, so calls that overlap on
// the GPU (different contexts / streams on one device) race.
static __device__ __align__(256) unsigned char
    qfx_pair_static_scratch[2ull * 8ull << QFX_STATIC_MAX_B_BITS];

extern "C" int qfx_contract_pair_static(qfx_ctx* ctx, const qfx_contract_desc* din) {
  if (ctx == nullptr || din == nullptr) {
    qfx_set_error("qfx_contract_pair_static: null argument");
    return QFX_ERR_ARG;
  }
  qfx_contract_desc d = *din;
  if (d.m_bits < 0 || d.n_bits < 0 || d.k_bits < 0 || d.m_bits + d.k_bits > 32 ||
      d.n_bits + d.k_bits > 32 || d.m_bits + d.n_bits > 32) {
    qfx_set_error("qfx_contract_pair_static: operand rank out of range (max 32 bits each)");
    return QFX_ERR_ARG;
  }
  auto span_of = [](const int8_t* p, int np, const int8_t* q, int nq, uint64_t* span) {
    uint64_t seen = 0;
    for (int pass = 0; pass < 2; ++pass) {
      const int8_t* v = pass ? q : p;
      const int n = pass ? nq : np;
      for (int i = 0; i < n; ++i) {
        if (v[i] < 0 || v[i] >= 32 || ((seen >> v[i]) & 1ull)) return false;
        seen |= 1ull << v[i];
      }
    }
    *span = seen + 1;
    return true;
  };
  uint64_t sa, sb, sc;
  if (!span_of(d.a_m, d.m_bits, d.a_k, d.k_bits, &sa) ||
      !span_of(d.b_k, d.k_bits, d.b_n, d.n_bits, &sb) ||
      !span_of(d.c_m, d.m_bits, d.c_n, d.n_bits, &sc)) {
    qfx_set_error("qfx_contract_pair_static: bit positions of an operand must be distinct and < 32");
    return QFX_ERR_ARG;
  }
  if (d.off_a + sa > ctx->arena_elems || d.off_b + sb > ctx->arena_elems ||
      d.off_c + sc > ctx->arena_elems) {
    qfx_set_error("qfx_contract_pair_static: operand outside the arena");
    return QFX_ERR_ARG;
  }
  if ((d.off_a < d.off_c + sc && d.off_c < d.off_a + sa) ||
      (d.off_b < d.off_c + sc && d.off_c < d.off_b + sb)) {
    qfx_set_error("qfx_contract_pair_static: C overlaps an input operand");
    return QFX_ERR_ARG;
  }
  if (d.n_bits > d.m_bits) {  // the larger free side plays the rows (as qfx_contract does)
    std::swap(d.off_a, d.off_b);
    std::swap(d.m_bits, d.n_bits);
    for (int i = 0; i < QFX_MAX_BITS; ++i) {
      std::swap(d.a_m[i], d.b_n[i]);
      std::swap(d.c_m[i], d.c_n[i]);
      std::swap(d.a_k[i], d.b_k[i]);
    }
  }
  if (d.n_bits + d.k_bits > QFX_STATIC_MAX_B_BITS || d.n_bits > 8 || d.k_bits > 8) {
    qfx_set_error("qfx_contract_pair_static: the smaller operand (2^" +
                  std::to_string(d.n_bits + d.k_bits) + " elements, n_bits " +
                  std::to_string(d.n_bits) + ", k_bits " + std::to_string(d.k_bits) +
                  ") exceeds the static scratch: at most 2^" +
                  std::to_string(QFX_STATIC_MAX_B_BITS) +
                  " elements with n_bits <= 8 and k_bits <= 8");
    return QFX_ERR_ARG;
  }
  QFX_CUDA(cudaSetDevice(ctx->device));
  // In host code the array's name is a host-side stand-in: its device address is
  // looked up once per device, with that device current.
  static qfx_per_device_once looked_up;
  static uint8_t* scratch_of[64] = {};
  if (ctx->device < 0 || ctx->device >= 64) {
    qfx_set_error("qfx_contract_pair_static: device index out of range");
    return QFX_ERR_ARG;
  }
  if (looked_up.need(ctx->device)) {
    void* p = nullptr;
    QFX_CUDA(cudaGetSymbolAddress(&p, qfx_pair_static_scratch));
    scratch_of[ctx->device] = static_cast<uint8_t*>(p);
    looked_up.mark(ctx->device);
  }
  bool done = false;
  int rc = launch_pair(ctx, d, ctx->arena + d.off_a, ctx->arena + d.off_b, ctx->arena + d.off_c,
                       &done, scratch_of[ctx->device]);
  if (rc) return rc;
  if (!done) {
    qfx_set_error("qfx_contract_pair_static: shape outside the CTA-pair kernel's range "
                  "(needs m_bits >= 8 with 2^m_bits / 256 >= SM pairs, 4 <= n_bits <= 8, "
                  "4 <= k_bits <= 8, TMA-addressable A)");
    return QFX_ERR_ARG;
  }
  const double M = (double)(1ull << d.m_bits), N = (double)(1ull << d.n_bits),
               K = (double)(1ull << d.k_bits);
  ctx->stats.flops += 8.0 * M * N * K;
  ctx->stats.bytes += 8.0 * (M * K + K * N + M * N);
  return QFX_OK;
}
