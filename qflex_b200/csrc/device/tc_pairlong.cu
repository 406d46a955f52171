// CTA-pair tensor-core contraction for the long-K / wide-N classes (k > 8 bits or
// n > 8 bits: the merges of the 7x7 and Sycamore schedules, 2^15 x 2^15 x 2^15,
// 2^18 x 2^12 x 2^10).
//
// Same math and operand handling as tc_fused.cu (reference: A.reorder + cblas_cgemm of
// src/tensor.cpp:1045-1194, 960-980; split-bf16 emulation of fp32; A gathered by TMA
// and converted in the kernel; K accumulated in TMEM in chunks that are folded in fp32
// registers), with the data flow of tc_pair.cu: the two CTAs of a cluster own a 256-row
// tile, every tcgen05.mma is M = 256 across the pair (cta_group::2), and each CTA stages
// only HALF of every B' tile.  tc_fused streams a full 256-row B' tile (64 KB of hi +
// mid per k block) into every SM for 12 MMAs: ~42 B/clk/SM, 11.8 TB/s of L2 -> SM
// traffic over the chip, which is the L2's limit and the reason those classes sat at
// ~0.6 of the tensor roof.  Here it is half of that.
//
// Work item = (256-row m tile, n tile of bn <= 256 columns); n tiles are rasterised in
// groups whose B' slab stays in L2 (as tc_fused).  Barriers as in tc_pair.cu: what the
// MMA warp waits on lives in the leader CTA and is arrived on remotely by the peer;
// buffer releases are multicast tcgen05.commit.
#include <cmath>

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
__device__ __forceinline__ void mbar_arrive_leader(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(bar & PEER_MASK) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx_leader(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cluster.b64 _, [%0], %1;"
               ::"r"(bar & PEER_MASK), "r"(bytes)
               : "memory");
}
// (CTA-scope semantics on purpose, see tc_pair.cu)
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
__device__ __forceinline__ void umma2_commit(uint32_t bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64"
      " [%0], %1;"
      ::"r"(bar), "h"((uint16_t)3)
      : "memory");
}

// work item -> (pair m tile, n tile): groups of n_group n tiles, m-major inside a group
__device__ __forceinline__ void decode_item(const FusedArgs& g, int w, int* mt, int* nt) {
  const int per_group = (g.tiles / g.tiles_n) * g.n_group;
  const int grp = w / per_group, r = w - grp * per_group;
  *mt = r / g.n_group;
  *nt = grp * g.n_group + (r - *mt * g.n_group);
}

// FusedArgs as in tc_fused.cu, except: tiles = (M / 256) * tiles_n, b_slot_bytes =
// bn/2 * 128 (this CTA's half of a B' tile).
template <int STORE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(FUSED_THREADS, 1)
    tc_pairlong_bf16x3(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmBh,
                       const __grid_constant__ CUtensorMap tmBm, const __grid_constant__ FusedArgs g) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~(uintptr_t)1023);
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t a16_base = smem_base;
  const uint32_t b_base = a16_base + (uint32_t)g.sc * 2 * A16_TILE;
  const uint32_t a32_base = b_base + (uint32_t)g.sb * g.b_slot_bytes;
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
  const uint32_t tmem_cols = (uint32_t)max(32, g.acc_stages * bn);
  const int pair0 = (int)(blockIdx.x >> 1), pair_step = (int)(gridDim.x >> 1);

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < MAX_SA; ++s) {
      mbar_init(a_full(s), 1);
      mbar_init(a_empty(s), CONV_WARPS);
    }
    for (int s = 0; s < MAX_SC; ++s) {
      mbar_init(c_full(s), 2 * CONV_WARPS);
      mbar_init(c_empty(s), 1);
    }
    for (int s = 0; s < MAX_SB; ++s) {
      mbar_init(b_full(s), 2);
      mbar_init(b_empty(s), 1);
    }
    for (int s = 0; s < MAX_ACC; ++s) {
      mbar_init(tfull_bar(s), 1);
      mbar_init(tempty_bar(s), 2 * 8);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_u32(tmem_slot)), "r"(tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < CONV_WARP0) {
    reg_dec<40>();
    if (warp == 0) {
      // ---------------- A producer (own 128 rows of the pair tile)
      const int m_sh = lane < g.n_m_hi ? g.a_m_hi[lane] - 1 : -1;
      const int k_sh = lane < g.n_k_hi ? g.a_k_hi[lane] - 1 : -1;
      int stage = 0;
      uint32_t phase = 0;
      for (int w = pair0; w < g.tiles * g.ksplit; w += pair_step) {
        const int kb0 = (w / g.tiles) * g.kblocks;  // split-K: this item's share of K
        int mt_, nt_;
        decode_item(g, w % g.tiles, &mt_, &nt_);
        const unsigned mt = 2u * (unsigned)mt_ + rank;
        const unsigned row_base =
            __reduce_or_sync(0xffffffffu, (m_sh >= 0 && ((mt >> lane) & 1u)) ? (1u << m_sh) : 0u);
        for (int kb = 0; kb < g.kblocks; ++kb) {
          const unsigned off16 =
              row_base | __reduce_or_sync(0xffffffffu, (k_sh >= 0 && (((unsigned)(kb0 + kb) >> lane) & 1u))
                                                           ? (1u << k_sh)
                                                           : 0u);
          if (lane == 0) {
            mbar_wait_spin(a_empty(stage), phase ^ 1u);
            mbar_expect_tx(a_full(stage), g.a32_bytes);
            tma_load_a(&tmA, g.a_rank, a_full(stage), a32_base + (uint32_t)stage * g.a32_bytes,
                       (int)off16);
          }
          if (++stage == g.sa) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    } else if (warp == 2 && lane == 0) {
      // ---------------- B producer: this CTA's bn/2 rows of B'_mid then B'_hi of every
      // k block; both CTAs' bytes are counted on the leader's barrier
      int slot = 0;
      uint32_t phase = 0;
      const int half_rows = bn >> 1;
      // (programmatic dependent launch: B' comes from the split kernel launched before)
      asm volatile("griddepcontrol.wait;" ::: "memory");
      for (int w = pair0; w < g.tiles * g.ksplit; w += pair_step) {
        const int kb0 = (w / g.tiles) * g.kblocks;
        int mt, nt;
        decode_item(g, w % g.tiles, &mt, &nt);
        for (int kb = kb0; kb < kb0 + g.kblocks; ++kb) {
#pragma unroll
          for (int part = 0; part < 2; ++part) {
            mbar_wait_spin(b_empty(slot), phase ^ 1u);
            mbar_expect_tx_leader(b_full(slot), g.b_slot_bytes);
            tma_load_2d_pair(part == 0 ? &tmBm : &tmBh, b_full(slot),
                             b_base + (uint32_t)slot * g.b_slot_bytes, kb * BK,
                             nt * bn + (int)rank * half_rows);
            if (++slot == g.sb) {
              slot = 0;
              phase ^= 1u;
            }
          }
        }
      }
    } else if (warp == 1 && rank == 0) {
      // ---------------- MMA issuer (leader CTA): M = 256 over the pair, N = bn, K = 16
      const bool leader = elect_one();
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(bn >> 3) << 17) |
                             ((uint32_t)((2 * BM) >> 4) << 24);
      int cs = 0, bs = 0, acc = 0;
      uint32_t cph = 0, bph = 0, acc_phase = 0;
      const int nkk = (g.dbg & 4) ? 0 : g.nkk;
      for (int w = pair0; w < g.tiles * g.ksplit; w += pair_step) {
        for (int ch = 0; ch < g.nchunks; ++ch) {
          mbar_wait_cluster(tempty_bar(acc), acc_phase ^ 1u);
          const uint32_t d_tmem = tmem_base + (uint32_t)(acc * bn);
          const int kb_end = min(g.kblocks, (ch + 1) * g.chunk);
          for (int kb = ch * g.chunk; kb < kb_end; ++kb) {
            const uint32_t first_acc = (kb == ch * g.chunk) ? 0u : 1u;
            const uint64_t dAh = smem_desc_sw128(a16_base + (uint32_t)cs * 2 * A16_TILE);
            const uint64_t dAm = dAh + (uint64_t)(A16_TILE >> 4);
            const uint64_t dBm = smem_desc_sw128(b_base + (uint32_t)bs * g.b_slot_bytes);
            mbar_wait_cluster(c_full(cs), cph);
            mbar_wait_cluster(b_full(bs), bph);
            tc_fence_after();
            if (leader) {
              for (int kk = 0; kk < nkk; ++kk)
                umma2_bf16(d_tmem, dAh + 2 * kk, dBm + 2 * kk, idesc, kk == 0 ? first_acc : 1u);
              umma2_commit(b_empty(bs));
            }
            if (++bs == g.sb) {
              bs = 0;
              bph ^= 1u;
            }
            const uint64_t dBh = smem_desc_sw128(b_base + (uint32_t)bs * g.b_slot_bytes);
            mbar_wait_cluster(b_full(bs), bph);
            tc_fence_after();
            if (leader) {
              for (int kk = 0; kk < nkk; ++kk)
                umma2_bf16(d_tmem, dAm + 2 * kk, dBh + 2 * kk, idesc, 1u);
              for (int kk = 0; kk < nkk; ++kk)
                umma2_bf16(d_tmem, dAh + 2 * kk, dBh + 2 * kk, idesc, 1u);
              umma2_commit(b_empty(bs));
              umma2_commit(c_empty(cs));
            }
            if (++bs == g.sb) {
              bs = 0;
              bph ^= 1u;
            }
            if (++cs == g.sc) {
              cs = 0;
              cph ^= 1u;
            }
          }
          if (leader) umma2_commit(tfull_bar(acc));  // chunk accumulator complete
          if (++acc == g.acc_stages) {
            acc = 0;
            acc_phase ^= 1u;
          }
        }
      }
    }
  } else if (warp < EPI_WARP0) {
    reg_dec<64>();
    // ---------------- converters (own 128 rows): as in tc_pair.cu
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
    for (int w = pair0; w < g.tiles * g.ksplit; w += pair_step) {
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
        __syncwarp();
        if (lane == 0) {
          mbar_arrive_leader(c_full(cs));
          mbar_arrive(a_empty(as));
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
    // ---------------- epilogue (own 128 rows): fp32 fold of the k chunks, then C
    const int lane_group = warp & 3;
    const int half = (warp - EPI_WARP0) >> 2;
    const int row = lane_group * 32 + lane;
    unsigned m_local = 0;
#pragma unroll
    for (int i = 0; i < 7; ++i) m_local |= (unsigned)((row >> i) & 1) << g.row_perm[i];
    const int half_cols = bn >> 1;
    const int npieces = half_cols >> 4;
    int acc = 0;
    uint32_t acc_phase = 0;
    const unsigned long long N = 1ull << g.n_bits;
    unsigned mq[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = lane_group * 32 + (lane & ~3) + i;
      mq[i] = 0;
#pragma unroll
      for (int b = 0; b < 7; ++b) mq[i] |= (unsigned)((r >> b) & 1) << g.row_perm[b];
    }
    TStore ts{};
    const bool t_stores = STORE == 2 && tstore_setup(g, g.C, row, lane, &ts);
    unsigned m_local_b = 0;
    if (STORE == 2 && t_stores) {
#pragma unroll
      for (int i = 0; i < 7; ++i) m_local_b |= (unsigned)((ts.rowb >> i) & 1) << g.row_perm[i];
    }
    for (int w = pair0; w < g.tiles * g.ksplit; w += pair_step) {
      int mt_, nt;
      decode_item(g, w % g.tiles, &mt_, &nt);
      const unsigned long long mtile = 2ull * (unsigned long long)mt_ + rank;
      // split-K: this item's partial sums go to its own [M][N] slab (added in a fixed
      // order by tc_splitk_finish, which also applies C's layout)
      float2* const Cout =
          g.ksplit > 1 ? g.part + ((unsigned long long)(w / g.tiles) << (g.m_bits + g.n_bits)) : g.C;
      float sum[8][16];
      for (int ch = 0; ch < g.nchunks; ++ch) {
        mbar_wait_spin(tfull_bar(acc), acc_phase);
        tc_fence_after();
        const uint32_t t_row = tmem_base + ((uint32_t)(lane_group * 32) << 16) +
                               (uint32_t)(acc * bn + half * half_cols);
#pragma unroll
        for (int p = 0; p < 8; ++p) {
          if (p < npieces && !(g.dbg & 8)) {
            uint32_t r[16];
            tmem_ld16(t_row + (uint32_t)(p * 16), r);
            const float scale = (ch + 1 == g.nchunks) ? g.scale_last : g.scale_full;
            if (ch == 0) {
#pragma unroll
              for (int j = 0; j < 16; ++j) sum[p][j] = __uint_as_float(r[j]) * scale;
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j) sum[p][j] = fmaf(__uint_as_float(r[j]), scale, sum[p][j]);
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_leader(tempty_bar(acc));
        if (++acc == g.acc_stages) {
          acc = 0;
          acc_phase ^= 1u;
        }
      }
      const unsigned long long m = mtile * BM + m_local;
      unsigned long long row_off = 0;
      if (g.c_natural) {
        row_off = m * N;
      } else {
        for (int q = 0; q < g.m_bits; ++q) row_off |= ((m >> q) & 1ull) << g.c_m[q];
      }
      if (STORE == 2 && t_stores) {
        const unsigned long long mb = mtile * BM + m_local_b;
        unsigned long long rb = 0;
        for (int q = 0; q < g.m_bits; ++q) rb |= ((mb >> q) & 1ull) << g.c_m[q];
        const unsigned long long n_base = ((unsigned long long)nt * bn + half * half_cols) >> 1;
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
            tstore_piece(sum[p], Cout + rb + co[0], Cout + rb + co[1], ts.xm0, ts.xm1, lane);
          }
        }
        continue;
      }
      if (STORE == 1) {
        float* tile_c = reinterpret_cast<float*>(Cout + mtile * BM * N);
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
          if (g.c_natural) {
            float* dst = reinterpret_cast<float*>(Cout + row_off + n0);
            stg256(dst, &sum[p][0]);
            stg256(dst + 8, &sum[p][8]);
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const unsigned long long n = n0 + j;
              unsigned long long co = 0;
              for (int q = 0; q < g.n_bits; ++q) co |= ((n >> q) & 1ull) << g.c_n[q];
              Cout[row_off + co] = make_float2(sum[p][2 * j], sum[p][2 * j + 1]);
            }
          }
        }
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

// tc_pair.cu's tile is 256 rows x ALL of N' (up to 512 columns).  When the tiles are few
// waves over the pairs and the last wave is mostly empty (2^16 x 256 x 256: 256 tiles over
// 74 pairs = 3.46 waves, 4 are paid), the same contraction as (256 rows x 256 columns)
// work items of this kernel wastes less: 512 items = 6.92 waves.
static bool din_takes_pair_tail(const qfx_contract_desc& d, int n_pairs) {
  static const int mode = [] {
    const char* e = getenv("QFLEX_TC_PAIRLONG_TAIL");
    return e ? atoi(e) : 1;
  }();
  if (mode == 0 || d.n_bits != 8 || d.k_bits < 5) return false;  // two n tiles of 256 columns
  const double tiles = (double)(1ull << d.m_bits) / 256.0;
  if (tiles < n_pairs || tiles > 8.0 * n_pairs) return false;
  const double eff_pair = tiles / n_pairs / std::ceil(tiles / n_pairs);
  const double eff_long = 2.0 * tiles / n_pairs / std::ceil(2.0 * tiles / n_pairs);
  return mode == 2 || eff_long > eff_pair + 0.08;
}

// Roles already swapped (m_bits >= n_bits).  *done = false: not applicable (the caller
// tries tc_fused next): shapes the pair kernel of tc_pair.cu takes (k, n <= 8 bits),
// fewer than two waves of 256-row work items (tc_fused splits K for those), or an A
// layout the TMA map cannot express.
int qfx_launch_contract_tc_pairlong(qfx_ctx* ctx, const qfx_contract_desc& din, const float2* A,
                                    const float2* B, float2* C, bool* done) {
  *done = false;
  static const bool enabled = [] {
    const char* e = getenv("QFLEX_TC_PAIRLONG");
    return e == nullptr || atoi(e) != 0;
  }();
  if (!enabled) return QFX_OK;
  qfx_contract_desc d = din;
  static const int min_k_bits = [] {
    const char* e = getenv("QFLEX_TC_PAIRLONG_MINK");
    return e ? atoi(e) : 5;  // K = 16: 2^16 x 2^12 x 16 gains 6 %, 2^18 x 512 x 16 and
  }();                       // 2^20 x 1024 x 16 lose 2-4 % against tc_fused: left there
  if (d.m_bits < 8 || d.n_bits < 4 || d.k_bits < std::max(4, min_k_bits)) return QFX_OK;
  const bool pair_shape = d.n_bits <= 8 && d.k_bits <= 8;  // what tc_pair.cu takes
  if (pair_shape && !din_takes_pair_tail(d, ctx->sm_count / 2)) return QFX_OK;
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
  const uint64_t M = 1ull << d.m_bits, N = 1ull << d.n_bits, K = 1ull << d.k_bits;
  const int n_pairs = ctx->sm_count / 2;
  FusedArgs g{};
  g.bn = (int)std::min<uint64_t>(2 * N, MAX_BN);
  g.tiles_n = (int)((2 * N) / g.bn);
  g.tiles = (int)((M / (2 * BM)) * g.tiles_n);
  // few work items and a long K (the final merges, 2^12 x 256 x 2^16 = 32 items for 74
  // pairs): split K over the items as tc_fused.cu does; partial [M][N] slabs, summed by
  // tc_splitk_finish in a fixed order
  const int kblocks_all = (int)std::max<uint64_t>(1, (2 * K) / BK);
  int ksplit = 1;
  if (g.tiles < 2 * n_pairs) {
    static const bool splitk_on = [] {
      const char* e = getenv("QFLEX_TC_PAIRLONG_SPLITK");
      return e == nullptr || atoi(e) != 0;
    }();
    if (!splitk_on || kblocks_all < 64) return QFX_OK;
    while (g.tiles * ksplit * 2 <= 8 * n_pairs && kblocks_all / (ksplit * 2) >= 8) ksplit *= 2;
    if (g.tiles * ksplit < 2 * n_pairs) return QFX_OK;
  }
  g.kb_bits = std::min(5, (int)d.k_bits);
  g.tile_bits = 7 + g.kb_bits;
  uint64_t span = 0;
  for (int i = 0; i < d.m_bits; ++i) span |= 1ull << d.a_m[i];
  for (int i = 0; i < d.k_bits; ++i) span |= 1ull << d.a_k[i];
  span += 1;
  CUtensorMap mA, mBh, mBm;
  if (!make_map_a(&mA, A, d, g.kb_bits, span, &g)) return QFX_OK;

  g.C = C;
  g.ksplit = ksplit;
  g.m_bits = d.m_bits;
  g.n_bits = d.n_bits;
  g.kblocks = kblocks_all / ksplit;
  g.nkk = (int)std::min<uint64_t>(4, (2 * K) / 16);
  g.chunk = ctx->tc_chunk > 0 ? ctx->tc_chunk : 4;
  g.nchunks = (g.kblocks + g.chunk - 1) / g.chunk;
  {
    const int last = g.kblocks - (g.nchunks - 1) * g.chunk;
    g.scale_full = tc_debias_scale(3 * g.nkk * std::min(g.chunk, g.kblocks), ctx->tc_debias);
    g.scale_last = tc_debias_scale(3 * g.nkk * last, ctx->tc_debias);
  }
  {
    // n tiles per rasterisation group: the group's B' slab stays in L2 while the m
    // tiles sweep past it (as tc_fused.cu)
    const double slab_per_ntile = (double)g.bn * (double)(2 * K) * 2.0 * 2.0;
    int G = 1;
    while (G * 2 <= g.tiles_n && (G * 2) * slab_per_ntile <= 16.0 * 1048576.0) G *= 2;
    if (const char* e = getenv("QFLEX_TC_NGROUP")) {
      G = std::max(1, atoi(e));
      while (g.tiles_n % G) --G;
    }
    g.n_group = G;
  }
  g.c_natural = 1;
  for (int i = 0; i < d.n_bits; ++i)
    if (d.c_n[i] != i) g.c_natural = 0;
  for (int i = 0; i < d.m_bits; ++i)
    if (d.c_m[i] != d.n_bits + i) g.c_natural = 0;
  if ((reinterpret_cast<uintptr_t>(C) & 31) != 0) g.c_natural = 0;
  const int c_is_natural = g.c_natural;
  if (ksplit > 1) g.c_natural = 1;  // the slabs are [M][N]; C's layout is applied by the finish
  memcpy(g.c_m, d.c_m, QFX_MAX_BITS);
  memcpy(g.c_n, d.c_n, QFX_MAX_BITS);
  g.n_m_hi = d.m_bits - 7;
  for (int i = 0; i < g.n_m_hi; ++i) g.a_m_hi[i] = d.a_m[7 + i];
  g.n_k_hi = d.k_bits - g.kb_bits;
  for (int i = 0; i < g.n_k_hi; ++i) g.a_k_hi[i] = d.a_k[g.kb_bits + i];

  // shared memory plan per CTA: bf16 A stages, half-tile B slots, fp32 A stages
  g.a32_bytes = 8u << g.tile_bits;
  g.b_slot_bytes = (uint32_t)(g.bn / 2) * 128u;
  g.acc_stages = std::min(MAX_ACC, 512 / g.bn);
  const int fixed = SMEM_FIXED;
  const int slot = (int)g.b_slot_bytes, a32 = (int)g.a32_bytes, a16 = 2 * A16_TILE;
  g.sc = 2;
  g.sb = 4;
  g.sa = 2;
  g.b_resident = 0;
  auto total = [&]() { return fixed + g.sc * a16 + g.sb * slot + g.sa * a32; };
  if (total() > SMEM_LIMIT) return QFX_OK;
  for (bool grew = true; grew;) {
    grew = false;
    if (g.sa < MAX_SA && total() + a32 <= SMEM_LIMIT) { ++g.sa; grew = true; }
    if (g.sb < MAX_SB && total() + slot <= SMEM_LIMIT) { ++g.sb; grew = true; }
    if (g.sc < MAX_SC && g.sc < g.sa && total() + a16 <= SMEM_LIMIT) { ++g.sc; grew = true; }
  }
  const int smem_bytes = total();
  if (const char* e = getenv("QFLEX_FUSED_DBG")) g.dbg = atoi(e);

  // workspace: B' hi/mid (2N x 2K bf16 each), then the split-K slabs
  const size_t b_bytes = (size_t)N * K * 8;
  const size_t part_bytes = ksplit > 1 ? (size_t)ksplit * M * N * 8 : 0;
  int rc = qfx_ensure_ws(ctx, 2 * b_bytes + 4096 + part_bytes);
  if (rc) return rc;
  uint8_t* ws = reinterpret_cast<uint8_t*>(ctx->ws);
  g.part = reinterpret_cast<float2*>(ws + ((2 * b_bytes + 4095) & ~(size_t)4095));
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
    QFX_CUDA(cudaFuncSetAttribute(tc_pairlong_bf16x3<0>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  SMEM_LIMIT));
    QFX_CUDA(cudaFuncSetAttribute(tc_pairlong_bf16x3<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  SMEM_LIMIT));
    QFX_CUDA(cudaFuncSetAttribute(tc_pairlong_bf16x3<2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  SMEM_LIMIT));
    attr.mark(ctx->device);
  }
  const int grid = 2 * std::min(g.tiles * g.ksplit, n_pairs);
  {
    // programmatic dependent of the split kernel (see tc_pair.cu); QFLEX_PDL=0: plain launch
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
      case 1: QFX_CUDA(cudaLaunchKernelEx(&cfg, tc_pairlong_bf16x3<1>, mA, mBh, mBm, g)); break;
      case 2: QFX_CUDA(cudaLaunchKernelEx(&cfg, tc_pairlong_bf16x3<2>, mA, mBh, mBm, g)); break;
      default: QFX_CUDA(cudaLaunchKernelEx(&cfg, tc_pairlong_bf16x3<0>, mA, mBh, mBm, g));
    }
  }
  QFX_CUDA(cudaGetLastError());
  if (ksplit > 1) {
    rc = qfx_launch_tc_splitk_finish(ctx, g.part, C, d, ksplit, c_is_natural);
    if (rc) return rc;
  }
  ctx->stats.launches_gemm_tc++;
  ctx->stats.launches_gemm_tc_fused++;
  ctx->stats.launches_gemm_tc_pair++;
  ctx->stats.launches_total++;
  *done = true;
  return QFX_OK;
}
