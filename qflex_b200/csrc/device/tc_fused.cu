// Fused tensor-core contraction: the fp32 -> split-bf16 conversion of the big
// (streamed) operand happens INSIDE the GEMM kernel, so a contraction reads A
// once from HBM as complex64, wherever its index bits live, and writes C once.
//
// Replaces, for the classes SURVEY 8d marks tensor-core candidates (and the
// d=16 HBM-bound expands), the reference's A.reorder + cblas_cgemm
// (src/tensor.cpp:1045-1194, 960-980): the reorder of A is folded into a
// multi-dimensional TMA tensor map over A's address bits, the product is
// tcgen05.mma kind::f16 on bf16 hi/mid pairs (x ~= hi + mid):
//     a*b ~= a_mid*b_hi + a_hi*b_mid + a_hi*b_hi        (rel. error ~ 2^-17)
// run as ONE real GEMM on the raw float views (see tc_split.cu for B'):
//     C'[m][2n+c] = sum_k' A'[m][k'] * B'[2n+c][k'],  k' = 2k + {re,im}
//
// Persistent kernel, one CTA per SM, 512 threads in four warpgroups:
//   warp 0      TMA producer for A: one box of 128 rows x 32 complex k per
//               stage, gathered through <= 4 address-bit runs + one "offset"
//               dimension (stride 16 B) that carries the tile base;
//   warp 1      MMA issuer (one elected lane, tcgen05.mma cta_group::1);
//   warp 2      TMA producer for B' (pre-split bf16 hi/mid, small, L2-resident);
//   warps 4-7   converters: fp32 tile (smem) -> bf16 hi + mid tiles in the
//               128B-swizzled K-major layout the MMA descriptors expect;
//   warps 8-15  epilogue: tcgen05.ld -> fp32 fold of the k chunks -> C.
// Register budget is rebalanced with setmaxnreg (40 / 72 / 200 / 200).
#include "tc_common.cuh"

namespace {

// tile index -> (m tile, n tile): groups of n_group n tiles, m-major inside a group
__device__ __forceinline__ void decode_tile(const FusedArgs& g, int t, int* mt, int* nt) {
  const int per_group = (g.tiles / g.tiles_n) * g.n_group;
  const int grp = t / per_group, r = t - grp * per_group;
  *mt = r / g.n_group;
  *nt = grp * g.n_group + (r - *mt * g.n_group);
}

// STORE: 0 row-per-thread stores (any C layout), 1 quad-transposed (natural C, wide
// rows), 2 transposed C; see tc_pair.cu.
template <int STORE>
__global__ void __launch_bounds__(FUSED_THREADS, 1)
    tc_fused_bf16x3(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmBh,
                    const __grid_constant__ CUtensorMap tmBm, const __grid_constant__ FusedArgs g) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~(uintptr_t)1023);
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t a16_base = smem_base;                                   // sc x (hi, mid)
  const uint32_t b_base = a16_base + (uint32_t)g.sc * 2 * A16_TILE;      // sb slots
  const uint32_t a32_base = b_base + (uint32_t)g.sb * g.b_slot_bytes;    // sa stages
  const uint32_t bar_base = a32_base + (uint32_t)g.sa * g.a32_bytes;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + (bar_base - smem_base) + 8 * NUM_BARS);
  auto a_full = [&](int s) { return bar_base + 8u * s; };
  auto a_empty = [&](int s) { return bar_base + 8u * (MAX_SA + s); };
  auto c_full = [&](int s) { return bar_base + 8u * (2 * MAX_SA + s); };
  auto c_empty = [&](int s) { return bar_base + 8u * (2 * MAX_SA + MAX_SC + s); };
  auto b_full = [&](int s) { return bar_base + 8u * (2 * MAX_SA + 2 * MAX_SC + s); };
  auto b_empty = [&](int s) { return bar_base + 8u * (2 * MAX_SA + 2 * MAX_SC + MAX_SB + s); };
  auto tfull_bar = [&](int s) {
    return bar_base + 8u * (2 * MAX_SA + 2 * MAX_SC + 2 * MAX_SB + s);
  };
  auto tempty_bar = [&](int s) {
    return bar_base + 8u * (2 * MAX_SA + 2 * MAX_SC + 2 * MAX_SB + MAX_ACC + s);
  };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bn = g.bn;
  const uint32_t tmem_cols = (uint32_t)max(32, g.acc_stages * bn);

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < MAX_SA; ++s) {
      mbar_init(a_full(s), 1);
      mbar_init(a_empty(s), CONV_WARPS);
    }
    for (int s = 0; s < MAX_SC; ++s) {
      mbar_init(c_full(s), CONV_WARPS);
      mbar_init(c_empty(s), 1);
    }
    for (int s = 0; s < MAX_SB; ++s) {
      mbar_init(b_full(s), 1);
      mbar_init(b_empty(s), 1);
    }
    for (int s = 0; s < MAX_ACC; ++s) {
      mbar_init(tfull_bar(s), 1);
      mbar_init(tempty_bar(s), 8);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                 ::"r"(smem_u32(tmem_slot)), "r"(tmem_cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < CONV_WARP0) {
    reg_dec<40>();
    if (warp == 0) {
      // ---------------- A producer: one gathered fp32 box per k block.  Lane i owns
      // bit i of the tile / k-block index; the tile base (in 16-byte units) is a
      // warp-wide OR, so the per-tile cost does not grow with the operand rank.
      const int m_sh = lane < g.n_m_hi ? g.a_m_hi[lane] - 1 : -1;
      const int k_sh = lane < g.n_k_hi ? g.a_k_hi[lane] - 1 : -1;
      int stage = 0;
      uint32_t phase = 0;
      for (int w = blockIdx.x; w < g.tiles * g.ksplit; w += gridDim.x) {
        const int t = w % g.tiles, kb0 = (w / g.tiles) * g.kblocks;
        int mt_, nt_;
        decode_tile(g, t, &mt_, &nt_);
        const unsigned mt = (unsigned)mt_;
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
      // ---------------- B producer: B'_mid then B'_hi of every k block
      int slot = 0;
      uint32_t phase = 0;
      // (programmatic dependent launch: B' comes from the split kernel launched before)
      asm volatile("griddepcontrol.wait;" ::: "memory");
      for (int w = blockIdx.x; w < g.tiles * g.ksplit; w += gridDim.x) {
        const int t = w % g.tiles, kb0 = (w / g.tiles) * g.kblocks;
        int mt, nt;
        decode_tile(g, t, &mt, &nt);
        if (g.b_resident && w != (int)blockIdx.x) break;
        for (int kb = kb0; kb < kb0 + g.kblocks; ++kb) {
#pragma unroll
          for (int part = 0; part < 2; ++part) {
            if (!g.b_resident) mbar_wait_spin(b_empty(slot), phase ^ 1u);
            mbar_expect_tx(b_full(slot), g.b_slot_bytes);
            tma_load_2d(part == 0 ? &tmBm : &tmBh, b_full(slot),
                        b_base + (uint32_t)slot * g.b_slot_bytes, kb * BK, nt * bn);
            if (++slot == g.sb) {
              slot = 0;
              phase ^= 1u;
            }
          }
        }
      }
    } else if (warp == 1) {
      // ---------------- MMA issuer.  The whole warp runs the (warp-uniform) control
      // flow and the barrier waits, so that descriptors live in uniform registers
      // and there are only a handful of instructions between two tcgen05.mma; one
      // elected lane issues the MMAs and commits.  The issue rate of this warp is
      // on the critical path of the kernel.
      const bool leader = elect_one();
      // instruction descriptor: D=f32, A=B=bf16, both K-major, N=bn, M=128
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(bn >> 3) << 17) |
                             ((uint32_t)(BM >> 4) << 24);
      int cs = 0, bs = 0, acc = 0;
      uint32_t cph = 0, bph = 0, acc_phase = 0;
      const int nkk = (g.dbg & 4) ? 0 : g.nkk;
      for (int w = blockIdx.x; w < g.tiles * g.ksplit; w += gridDim.x) {
        for (int ch = 0; ch < g.nchunks; ++ch) {
          mbar_wait_spin(tempty_bar(acc), acc_phase ^ 1u);
          const uint32_t d_tmem = tmem_base + (uint32_t)(acc * bn);
          const int kb_end = min(g.kblocks, (ch + 1) * g.chunk);
          for (int kb = ch * g.chunk; kb < kb_end; ++kb) {
            const uint32_t first_acc = (kb == ch * g.chunk) ? 0u : 1u;
            const uint64_t dAh = smem_desc_sw128(a16_base + (uint32_t)cs * 2 * A16_TILE);
            const uint64_t dAm = dAh + (uint64_t)(A16_TILE >> 4);
            const uint64_t dBm = smem_desc_sw128(b_base + (uint32_t)bs * g.b_slot_bytes);
            mbar_wait_spin(c_full(cs), cph);
            mbar_wait_spin(b_full(bs), bph);
            tc_fence_after();
            if (leader) {
              if (nkk == 4) {
#pragma unroll
                for (int kk = 0; kk < 4; ++kk)
                  umma_bf16(d_tmem, dAh + 2 * kk, dBm + 2 * kk, idesc, kk == 0 ? first_acc : 1u);
              } else {
                for (int kk = 0; kk < nkk; ++kk)
                  umma_bf16(d_tmem, dAh + 2 * kk, dBm + 2 * kk, idesc, kk == 0 ? first_acc : 1u);
              }
              if (!g.b_resident) umma_commit(b_empty(bs));
            }
            if (++bs == g.sb) {
              bs = 0;
              bph ^= g.b_resident ? 0u : 1u;
            }
            const uint64_t dBh = smem_desc_sw128(b_base + (uint32_t)bs * g.b_slot_bytes);
            mbar_wait_spin(b_full(bs), bph);
            tc_fence_after();
            if (leader) {
              if (nkk == 4) {
#pragma unroll
                for (int kk = 0; kk < 4; ++kk)
                  umma_bf16(d_tmem, dAm + 2 * kk, dBh + 2 * kk, idesc, 1u);
#pragma unroll
                for (int kk = 0; kk < 4; ++kk)
                  umma_bf16(d_tmem, dAh + 2 * kk, dBh + 2 * kk, idesc, 1u);
              } else {
                for (int kk = 0; kk < nkk; ++kk)
                  umma_bf16(d_tmem, dAm + 2 * kk, dBh + 2 * kk, idesc, 1u);
                for (int kk = 0; kk < nkk; ++kk)
                  umma_bf16(d_tmem, dAh + 2 * kk, dBh + 2 * kk, idesc, 1u);
              }
              if (!g.b_resident) umma_commit(b_empty(bs));
              umma_commit(c_empty(cs));
            }
            if (++bs == g.sb) {
              bs = 0;
              bph ^= g.b_resident ? 0u : 1u;
            }
            if (++cs == g.sc) {
              cs = 0;
              cph ^= 1u;
            }
          }
          if (leader) umma_commit(tfull_bar(acc));  // chunk accumulator complete
          if (++acc == g.acc_stages) {
            acc = 0;
            acc_phase ^= 1u;
          }
        }
      }
    }
  } else if (warp < EPI_WARP0) {
    reg_dec<64>();
    // ---------------- converters: fp32 stage -> bf16 hi / mid (SW128 K-major).
    // One item = the 4 complex (k bits 0,1) of one row = one 16-byte chunk of the
    // hi tile and one of the mid tile.
    const unsigned ctid = threadIdx.x - CONV_WARP0 * 32;
    uint32_t src_t = 0, dst_t = 0;
#pragma unroll
    for (int b = 0; b < 8; ++b)
      if ((ctid >> b) & 1u) {
        src_t ^= g.cv_src[b];
        dst_t ^= g.cv_dst[b];
      }
    const uint32_t ok0 = g.cv_ok0, ok1 = g.cv_ok1;
    // 16-byte loads (k0 = lowest smem bit): lanes with bit 2 set read the k1 = 1
    // half first, so that a quarter warp touches 8 distinct 16-byte bank groups
    const bool vec = g.cv_vec != 0;
    const uint32_t flip = (vec && ((ctid >> 2) & 1u)) ? ok1 : 0u;
    int as = 0, cs = 0;
    uint32_t aph = 0, cph = 0;
    for (int w = blockIdx.x; w < g.tiles * g.ksplit; w += gridDim.x) {
      for (int kb = 0; kb < g.kblocks; ++kb) {
        mbar_wait_spin(a_full(as), aph);
        mbar_wait_spin(c_empty(cs), cph ^ 1u);
        // offsets are XOR-combined first, bases (only 1 KB aligned) added last
        const uint32_t src = a32_base + (uint32_t)as * g.a32_bytes;
        const uint32_t hi_tile = a16_base + (uint32_t)cs * 2 * A16_TILE;
#pragma unroll
        for (int it = 0; it < 4; ++it) {
          if (it < g.cv_its && !(g.dbg & 1)) {
            const uint32_t s = src_t ^ g.cv_src_it[it];
            float2 e00, e10, e01, e11;  // e<k0><k1>
            if (vec) {
              const float4 p = lds128(src + (s ^ flip)), q = lds128(src + (s ^ flip ^ ok1));
              const float4 lo = flip ? q : p, hi = flip ? p : q;  // lo: k1 = 0, hi: k1 = 1
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
          mbar_arrive(c_full(cs));
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
    // ---------------- epilogue: warp w may only touch TMEM lanes 32*(w%4) .. +31; two
    // warps share a lane group and split the tile's columns.  The tensor core
    // accumulates with truncation, which biases long sums towards zero, so an
    // accumulator only integrates `chunk` k blocks; chunks are folded here in
    // fp32 with round-to-nearest adds, in registers, and C is written once.
    const int lane_group = warp & 3;
    const int half = (warp - EPI_WARP0) >> 2;
    const int row = lane_group * 32 + lane;
    unsigned m_local = 0;  // tile rows are a bit permutation of the tile's m values
#pragma unroll
    for (int i = 0; i < 7; ++i) m_local |= (unsigned)((row >> i) & 1) << g.row_perm[i];
    const int half_cols = bn >> 1;
    const int npieces = half_cols >> 4;
    int acc = 0;
    uint32_t acc_phase = 0;
    const unsigned long long N = 1ull << g.n_bits;
    // quad-transposed stores (tc_common.cuh): tile rows of the four lanes of this quad
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
    for (int w = blockIdx.x; w < g.tiles * g.ksplit; w += gridDim.x) {
      const int t = w % g.tiles;
      int mt, nt;
      decode_tile(g, t, &mt, &nt);
      // split-K: this item's partial sums go to its own [M][N] slab (summed in a
      // fixed order by tc_splitk_finish)
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
            // (scale: accumulation-bias compensation of this chunk, tc_common.cuh)
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
        if (lane == 0) mbar_arrive(tempty_bar(acc));
        if (++acc == g.acc_stages) {
          acc = 0;
          acc_phase ^= 1u;
        }
      }
      const unsigned long long m = (unsigned long long)mt * BM + m_local;
      unsigned long long row_off = 0;
      if (g.c_natural) {
        row_off = m * N;
      } else {
        for (int q = 0; q < g.m_bits; ++q) row_off |= ((m >> q) & 1ull) << g.c_m[q];
      }
      if (STORE == 2 && t_stores) {
        const unsigned long long mb = (unsigned long long)mt * BM + m_local_b;
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
            tstore_piece(sum[p], Cout + rb + co[0], Cout + rb + co[1], ts.xm0, ts.xm1, lane);
          }
        }
        continue;
      }
      if (STORE == 1 && quad_stores) {
        float* tile_c = reinterpret_cast<float*>(Cout + (unsigned long long)mt * BM * N);
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
          // complex column of the first element of this piece
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

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                 "r"(tmem_cols)
                 : "memory");
  }
}

struct FinishArgs {
  const float2* part;
  float2* C;
  int m_bits, n_bits, ksplit, natural;
  int8_t c_m[QFX_MAX_BITS], c_n[QFX_MAX_BITS];
};

// C = sum over the split-K slabs, in slab order; two complex per thread (n >= 4 bits)
__global__ void tc_splitk_finish(const __grid_constant__ FinishArgs g) {
  const unsigned long long total = 1ull << (g.m_bits + g.n_bits - 1);
  const float4* p = reinterpret_cast<const float4*>(g.part);
  for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
       i < total; i += (unsigned long long)gridDim.x * blockDim.x) {
    float4 s = p[i];
    for (int k = 1; k < g.ksplit; ++k) {
      const float4 v = p[((unsigned long long)k << (g.m_bits + g.n_bits - 1)) + i];
      s.x += v.x;
      s.y += v.y;
      s.z += v.z;
      s.w += v.w;
    }
    if (g.natural) {
      reinterpret_cast<float4*>(g.C)[i] = s;
    } else {
      const unsigned long long e = 2 * i;
      const unsigned long long n = e & ((1ull << g.n_bits) - 1ull), m = e >> g.n_bits;
      unsigned long long o = 0;
      for (int q = 1; q < g.n_bits; ++q) o |= ((n >> q) & 1ull) << g.c_n[q];
      for (int q = 0; q < g.m_bits; ++q) o |= ((m >> q) & 1ull) << g.c_m[q];
      g.C[o] = make_float2(s.x, s.y);
      g.C[o | (1ull << g.c_n[0])] = make_float2(s.z, s.w);
    }
  }
}

}  // namespace

// C = sum of the ksplit [M][N] slabs at `part` (also used by tc_pairlong.cu)
int qfx_launch_tc_splitk_finish(qfx_ctx* ctx, const float2* part, float2* C,
                                const qfx_contract_desc& d, int ksplit, int c_is_natural) {
  FinishArgs f{};
  f.part = part;
  f.C = C;
  f.m_bits = d.m_bits;
  f.n_bits = d.n_bits;
  f.ksplit = ksplit;
  f.natural = c_is_natural;
  memcpy(f.c_m, d.c_m, QFX_MAX_BITS);
  memcpy(f.c_n, d.c_n, QFX_MAX_BITS);
  const unsigned long long pairs = 1ull << (d.m_bits + d.n_bits - 1);
  const unsigned fb = (unsigned)std::min<unsigned long long>((pairs + 255) / 256,
                                                             16ull * ctx->sm_count);
  tc_splitk_finish<<<fb, 256, 0, ctx->stream>>>(f);
  QFX_CUDA(cudaGetLastError());
  ctx->stats.launches_reduce++;
  ctx->stats.launches_total++;
  return QFX_OK;
}

// Roles already swapped (m_bits >= n_bits).  Returns QFX_OK and sets *done = true
// when the fused kernel was launched; *done = false means "not applicable, use the
// pre-split path" (not an error).
int qfx_launch_contract_tc_fused(qfx_ctx* ctx, const qfx_contract_desc& din, const float2* A,
                                 const float2* B, float2* C, bool* done) {
  *done = false;
  static const bool disabled = [] {
    const char* e = getenv("QFLEX_TC_FUSED");
    return e != nullptr && atoi(e) == 0;
  }();
  if (disabled) return QFX_OK;
  qfx_contract_desc d = din;
  if (d.m_bits < 7 || d.n_bits < 4 || d.k_bits < 4) return QFX_OK;
  // K enumeration is free: lowest A address bits first, so that a k block is
  // made of A's most contiguous k bits (B' is gathered by the split pre-pass).
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
  FusedArgs g{};
  g.kb_bits = std::min(5, (int)d.k_bits);
  g.tile_bits = 7 + g.kb_bits;
  uint64_t span = 0;
  for (int i = 0; i < d.m_bits; ++i) span |= 1ull << d.a_m[i];
  for (int i = 0; i < d.k_bits; ++i) span |= 1ull << d.a_k[i];
  span += 1;
  CUtensorMap mA, mBh, mBm;
  if (!make_map_a(&mA, A, d, g.kb_bits, span, &g)) return QFX_OK;

  // split-K: few output tiles and a long K (the final merges, e.g. 2^12 x 256 x 2^16
  // = 64 tiles for 148 SMs).  Work items = tiles x ksplit, each item accumulates a
  // power-of-two share of the k blocks into its own [M][N] slab; tc_splitk_finish
  // adds the slabs in a fixed order (deterministic) and applies C's layout.
  const int kblocks_all = (int)std::max<uint64_t>(1, (2 * K) / BK);
  const int tiles_all = (int)((M / BM) * ((2 * N) / std::min<uint64_t>(2 * N, MAX_BN)));
  int ksplit = 1;
  if (tiles_all < 2 * ctx->sm_count && kblocks_all >= 64) {
    while (tiles_all * ksplit * 2 <= 8 * ctx->sm_count && kblocks_all / (ksplit * 2) >= 8) ksplit *= 2;
  }
  if (const char* e = getenv("QFLEX_TC_KSPLIT")) ksplit = std::max(1, atoi(e));
  const size_t part_bytes = ksplit > 1 ? (size_t)ksplit * M * N * 8 : 0;

  // workspace: B' hi/mid (2N x 2K bf16 each), then the split-K slabs
  const size_t b_bytes = (size_t)N * K * 8;
  int rc = qfx_ensure_ws(ctx, 2 * b_bytes + 4096 + part_bytes);
  if (rc) return rc;
  uint8_t* ws = reinterpret_cast<uint8_t*>(ctx->ws);
  uint32_t* b_hi = reinterpret_cast<uint32_t*>(ws);
  uint32_t* b_mid = reinterpret_cast<uint32_t*>(ws + b_bytes);
  rc = qfx_launch_split(ctx, B, 1, d.b_n, d.n_bits, d.b_k, d.k_bits, b_hi, b_mid);
  if (rc) return rc;

  g.C = C;
  g.ksplit = ksplit;
  g.part = reinterpret_cast<float2*>(ws + ((2 * b_bytes + 4095) & ~(size_t)4095));
  g.m_bits = d.m_bits;
  g.n_bits = d.n_bits;
  g.kblocks = kblocks_all / ksplit;
  g.nkk = (int)std::min<uint64_t>(4, (2 * K) / 16);
  g.chunk = ctx->tc_chunk > 0 ? ctx->tc_chunk : 4;
  g.nchunks = (g.kblocks + g.chunk - 1) / g.chunk;
  {
    // MMA instructions per accumulator: 3 products x nkk k16-steps x k blocks
    const int last = g.kblocks - (g.nchunks - 1) * g.chunk;
    g.scale_full = tc_debias_scale(3 * g.nkk * std::min(g.chunk, g.kblocks), ctx->tc_debias);
    g.scale_last = tc_debias_scale(3 * g.nkk * last, ctx->tc_debias);
  }
  g.bn = (int)std::min<uint64_t>(2 * N, MAX_BN);
  g.tiles_n = (int)((2 * N) / g.bn);
  g.tiles = (int)((M / BM) * g.tiles_n);
  // n tiles per rasterisation group: the group's B' slab (hi + mid) should stay in L2
  // while the m tiles sweep past it (wide N: 2^18 x 2^12 x 2^10 streamed all 67 MB of
  // B' against every few m tiles and ran at 156 TFLOP/s)
  {
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

  // shared memory plan: bf16 A stages, B slots, fp32 A stages
  g.a32_bytes = 8u << g.tile_bits;
  g.b_slot_bytes = (uint32_t)g.bn * 128u;
  g.acc_stages = std::min(MAX_ACC, 512 / g.bn);
  const int fixed = SMEM_FIXED;
  const int slot = (int)g.b_slot_bytes, a32 = (int)g.a32_bytes, a16 = 2 * A16_TILE;
  g.sc = 2;
  g.sb = 3;
  g.sa = 2;
  g.b_resident = (ksplit == 1 && g.tiles_n == 1 && 2 * g.kblocks <= FUSED_MAX_SB) ? 1 : 0;
  if (g.b_resident) g.sb = 2 * g.kblocks;
  auto total = [&]() { return fixed + g.sc * a16 + g.sb * slot + g.sa * a32; };
  if (g.b_resident && total() > SMEM_LIMIT) {  // all of B' does not fit: stream it
    g.b_resident = 0;
    g.sb = 3;
  }
  if (total() > SMEM_LIMIT) return QFX_OK;
  // grow round-robin while it fits: fp32 stages (memory-level parallelism) first
  for (bool grew = true; grew;) {
    grew = false;
    if (g.sa < MAX_SA && total() + a32 <= SMEM_LIMIT) { ++g.sa; grew = true; }
    if (!g.b_resident && g.sb < FUSED_MAX_SB && total() + slot <= SMEM_LIMIT) { ++g.sb; grew = true; }
    if (g.sc < MAX_SC && g.sc < g.sa && total() + a16 <= SMEM_LIMIT) { ++g.sc; grew = true; }
  }
  const int smem_bytes = total();
  if (const char* e = getenv("QFLEX_FUSED_DBG")) g.dbg = atoi(e);
  if (const char* e = getenv("QFLEX_FUSED_ACC")) g.acc_stages = std::min(g.acc_stages, std::max(1, atoi(e)));

  if (!make_map_b(&mBh, b_hi, 2 * N, 2 * K, (uint32_t)g.bn) ||
      !make_map_b(&mBm, b_mid, 2 * N, 2 * K, (uint32_t)g.bn)) {
    qfx_set_error("cuTensorMapEncodeTiled failed");
    return QFX_ERR_CUDA;
  }
  static qfx_per_device_once attr;
  if (attr.need(ctx->device)) {
    QFX_CUDA(cudaFuncSetAttribute(tc_fused_bf16x3<0>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  SMEM_LIMIT));
    QFX_CUDA(cudaFuncSetAttribute(tc_fused_bf16x3<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  SMEM_LIMIT));
    QFX_CUDA(cudaFuncSetAttribute(tc_fused_bf16x3<2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  SMEM_LIMIT));
    attr.mark(ctx->device);
  }
  const int grid = std::min(g.tiles * g.ksplit, ctx->sm_count);
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
      case 1: QFX_CUDA(cudaLaunchKernelEx(&cfg, tc_fused_bf16x3<1>, mA, mBh, mBm, g)); break;
      case 2: QFX_CUDA(cudaLaunchKernelEx(&cfg, tc_fused_bf16x3<2>, mA, mBh, mBm, g)); break;
      default: QFX_CUDA(cudaLaunchKernelEx(&cfg, tc_fused_bf16x3<0>, mA, mBh, mBm, g));
    }
  }
  QFX_CUDA(cudaGetLastError());
  if (ksplit > 1) {
    rc = qfx_launch_tc_splitk_finish(ctx, g.part, C, d, ksplit, c_is_natural);
    if (rc) return rc;
  }
  ctx->stats.launches_gemm_tc++;
  ctx->stats.launches_gemm_tc_fused++;
  ctx->stats.launches_total++;
  *done = true;
  return QFX_OK;
}
