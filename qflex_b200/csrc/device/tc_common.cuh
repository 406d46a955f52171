// Shared pieces of the tensor-core contraction kernels (tc_fused.cu: one CTA per
// SM; tc_pair.cu: CTA pairs, tcgen05 cta_group::2): PTX wrappers, the kernel
// argument block, and the host-side tensor-map builders (the gathered fp32 view
// of the streamed operand and the converter XOR tables).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "qfx_internal.h"

int qfx_launch_split(qfx_ctx* ctx, const float2* src, int mode, const int8_t* row_pos,
                     int row_bits, const int8_t* k_pos, int k_bits, uint32_t* hi,
                     uint32_t* mid);

namespace {

constexpr int BM = 128;
constexpr int BK = 64;                       // bf16 per k block = one 128-byte swizzle row
constexpr int MAX_BN = 256;
constexpr int A16_TILE = BM * BK * 2;        // 16 KB: one bf16 tile (hi or mid)
constexpr int MAX_SC = 4;                    // bf16 A stages (hi + mid each)
constexpr int CONV_WARP0 = 4, CONV_WARPS = 8, EPI_WARP0 = CONV_WARP0 + CONV_WARPS;
constexpr int CONV_THREADS = CONV_WARPS * 32;
constexpr int FUSED_THREADS = (EPI_WARP0 + 8) * 32;  // 640
constexpr int MAX_SA = 6, MAX_SB = 16, MAX_ACC = 8;  // barrier slots (tc_fused uses <= 8 B slots)
constexpr int FUSED_MAX_SB = 8;
constexpr int SMEM_FIXED = 1024 + 640;       // alignment slack + barriers + TMEM slot
constexpr int NUM_BARS = 2 * MAX_SA + 2 * MAX_SC + 2 * MAX_SB + 2 * MAX_ACC;
constexpr int SMEM_LIMIT = 232448;           // 227 KB opt-in maximum per CTA
constexpr int MAX_TILE_BITS = 12;

struct FusedArgs {
  float2* C;
  int m_bits, n_bits;
  int kblocks;        // k blocks per tile
  int nkk;            // 16-wide MMA k steps per k block (<= 4)
  int chunk, nchunks; // k blocks accumulated in TMEM before the fp32 fold
  int bn, tiles_n, tiles;
  int c_natural;
  int a_rank;         // rank of the A tensor map (in-tile runs + offset dim)
  int kb_bits;        // complex k bits inside a k block
  int tile_bits;      // 7 + kb_bits
  int sa, sb;         // fp32 A stages, B slots (each slot = one bn x 64 bf16 tile)
  int sc, acc_stages; // bf16 A stages, TMEM accumulator stages
  int b_resident;     // all of B' fits the slots (one n tile): loaded once, never released
  uint32_t a32_bytes; // bytes of one fp32 A stage
  uint32_t b_slot_bytes;
  // tile base: address bit of every m bit >= 7 and every k bit >= kb_bits
  int8_t a_m_hi[QFX_MAX_BITS], a_k_hi[QFX_MAX_BITS];
  int n_m_hi, n_k_hi;
  // converters: one work item = one 16-byte chunk of an A' row (4 complex k of one
  // row).  Source / destination byte offsets are XOR-linear in the item index:
  // bits 0..7 = converter thread, higher bits = iteration.
  uint32_t cv_src[8], cv_dst[8];
  uint32_t cv_src_it[4], cv_dst_it[4];
  int cv_its;         // items per thread per k block
  uint32_t cv_ok0, cv_ok1;  // source byte offsets of k bit 0 / k bit 1
  int cv_vec;         // k0, k1 are the two lowest smem element bits: 16-byte loads
  int8_t row_perm[7]; // tile row bit i is m bit row_perm[i]
  int n_group;        // tc_fused rasterisation: n tiles per group (power of two, divides tiles_n);
                      // tiles run group by group, m-major inside, so a group's B' slab stays in L2
  int dual;           // tc_pair, K = 16: a stage carries TWO m tiles (m bit 7 rides as a fifth k bit
                      // of the A tile: columns 32..63 of the bf16 row); two accumulators per stage
  int lag;            // tc_pair, two n tiles: n tile 1 runs one k block behind n tile 0 (sc >= 3)
  int dual_epi;       // tc_pair dual: epilogue warps 0-3 / 4-7 drain accumulator 0 / 1, a whole row per thread
  int stagger_ns;     // tc_pair: start delay per pair index (tuning experiment)
  int blocked;        // tc_pair: each pair owns a contiguous range of tiles (else interleaved)
  int one_arrive;     // tc_pair: one barrier arrive per CTA (not per warp) for "A converted" / "accumulator drained"
  int ksplit;         // tc_fused split-K: work items = tiles * ksplit, kblocks = k blocks per item
  float2* part;       // split-K partial results [ksplit][M][N] (natural layout)
  int dbg;            // QFLEX_FUSED_DBG: 1 no convert, 2 no C stores, 4 no MMA, 8 no TMEM loads, 64 no A loads (timing experiments)
  int8_t c_m[QFX_MAX_BITS], c_n[QFX_MAX_BITS];
  // Compensation of the tensor core's accumulation bias (tc_debias_scale below): what
  // is read from a TMEM accumulator is multiplied by scale_full (an accumulator that
  // integrated `chunk` k blocks) or scale_last (the shorter last chunk of a tile).
  float scale_full, scale_last;
};

// The tcgen05 accumulator adds with round-toward-zero.  For data of random sign that
// shrinks every result by the same expected factor, about 1.2e-8 per MMA instruction
// accumulated into the same TMEM tile (measured: tools/tc_bias.py,
// profiles/tc_bias_r2_off.log -- 24 instructions 2.7e-7, 48: 5.5e-7, 96: 1.27e-6; with
// the compensation, profiles/tc_bias_r2_on.log: within +-7e-8).  One contraction
// hides that inside its 4e-6 random error, but a bias adds up LINEARLY over the ~45
// tensor-core contractions behind an amplitude of bristlecone_70 1-32-1 (measured
// -5.6e-5 against the reference, the FFMA path: 1e-6) while the random parts add in
// quadrature.  The epilogue therefore multiplies what it reads from TMEM by
// 1 + beta(instructions); QFLEX_TC_DEBIAS=0 switches it off.
inline float tc_debias_scale(int mma_instructions, bool enabled) {
  if (!enabled || mma_instructions <= 0) return 1.0f;
  static const double steps[] = {0, 6, 12, 24, 48, 96, 192};
  static const double beta[] = {0, 0.6e-7, 1.3e-7, 2.7e-7, 5.5e-7, 1.27e-6, 2.9e-6};
  const double t = (double)mma_instructions;
  int i = 0;
  while (i < 5 && t > steps[i + 1]) ++i;
  const double f = (t - steps[i]) / (steps[i + 1] - steps[i]);
  return (float)(1.0 + beta[i] + f * (beta[i + 1] - beta[i]));
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Bounded spin: a protocol bug traps instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok = 0;
  const long long t0 = clock64();
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    if (!ok && clock64() - t0 > 4000000000ll) __trap();
  } while (!ok);
}

__device__ __forceinline__ void tma_load_2d(const CUtensorMap* map, uint32_t bar, uint32_t dst,
                                            int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}

// A tile: all in-tile coordinates are 0, dimension 1 (stride 16 bytes, box 1)
// carries the tile base offset in units of two complex64.
__device__ __forceinline__ void tma_load_a(const CUtensorMap* map, int rank, uint32_t bar,
                                           uint32_t dst, int off16) {
  const uint64_t m = reinterpret_cast<uint64_t>(map);
  const int z = 0;
  if (rank == 2) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(m), "r"(bar), "r"(z), "r"(off16)
        : "memory");
  } else if (rank == 3) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %3}], [%2];"
        ::"r"(dst), "l"(m), "r"(bar), "r"(z), "r"(off16)
        : "memory");
  } else if (rank == 4) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %3, %3}], [%2];"
        ::"r"(dst), "l"(m), "r"(bar), "r"(z), "r"(off16)
        : "memory");
  } else {
    asm volatile(
        "cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %3, %3, %3}], [%2];"
        ::"r"(dst), "l"(m), "r"(bar), "r"(z), "r"(off16)
        : "memory");
  }
}

__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                          uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
               ::"r"(bar)
               : "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
// generic-proxy smem writes -> visible to the async proxy (UMMA operand reads)
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// K-major, 128-byte swizzle, dense 128-byte rows: SBO = 8 rows * 128 B = 1024.
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t addr) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFFu) >> 4);  // start address
  d |= (uint64_t)1 << 16;                   // leading byte offset (unused for SW128 K-major)
  d |= (uint64_t)(1024 >> 4) << 32;         // stride byte offset
  d |= (uint64_t)1 << 46;                   // descriptor version (sm_100)
  d |= (uint64_t)2 << 61;                   // SWIZZLE_128B
  return d;
}

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// Asynchronous form: issue several loads back to back, then ONE tmem_wait_ld() and a
// tmem_reg_fence() per destination (keeps register uses behind the wait).  With a
// wait after every x16 load, draining a 128-column accumulator cost ~8 load
// latencies on the epilogue's critical path.
__device__ __forceinline__ void tmem_ld16_async(uint32_t taddr, float (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=f"(r[0]), "=f"(r[1]), "=f"(r[2]), "=f"(r[3]), "=f"(r[4]), "=f"(r[5]), "=f"(r[6]),
        "=f"(r[7]), "=f"(r[8]), "=f"(r[9]), "=f"(r[10]), "=f"(r[11]), "=f"(r[12]), "=f"(r[13]),
        "=f"(r[14]), "=f"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_reg_fence(float (&r)[16]) {
  asm volatile("" : "+f"(r[0]), "+f"(r[1]), "+f"(r[2]), "+f"(r[3]), "+f"(r[4]), "+f"(r[5]),
                    "+f"(r[6]), "+f"(r[7]), "+f"(r[8]), "+f"(r[9]), "+f"(r[10]), "+f"(r[11]),
                    "+f"(r[12]), "+f"(r[13]), "+f"(r[14]), "+f"(r[15])
               :
               : "memory");
}

__device__ __forceinline__ float4 lds128(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts32(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

// 32-byte global store (sm_100 256-bit vector store): one full L2 sector per thread
__device__ __forceinline__ void stg256(float* p, const float* v) {
  asm volatile("st.global.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "f"(v[0]),
               "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7])
               : "memory");
}

// Transpose, inside every quad of lanes, the 4 rows x 32 floats the quad holds.
// In: lane j of a quad holds 32 consecutive columns of ITS row (x = columns 0..15,
// y = 16..31).  Out: lane j holds columns 8j..8j+7 of the quad's four rows:
// x[0..7] row 0, x[8..15] row 1, y[0..7] row 2, y[8..15] row 3.  A quad then
// stores one full 128-byte line of a C row per instruction (8 lines per warp store
// instead of 32: the row-per-thread stores cost 0.2 ms of a 1.1 ms launch).
__device__ __forceinline__ void quad_transpose(float (&x)[16], float (&y)[16], int lane) {
  const bool odd = (lane & 1) != 0, hi = (lane & 2) != 0;
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const float sx = odd ? x[e] : x[8 + e];
    const float sy = odd ? y[e] : y[8 + e];
    const float rx = __shfl_xor_sync(0xffffffffu, sx, 1);
    const float ry = __shfl_xor_sync(0xffffffffu, sy, 1);
    if (odd) {
      x[e] = rx;
      y[e] = ry;
    } else {
      x[8 + e] = rx;
      y[8 + e] = ry;
    }
  }
#pragma unroll
  for (int e = 0; e < 16; ++e) {
    const float s = hi ? x[e] : y[e];
    const float r = __shfl_xor_sync(0xffffffffu, s, 2);
    if (hi) x[e] = r;
    else y[e] = r;
  }
}

// Transposed C (the rows of the tile are the FAST dimension of C in memory: what a
// role swap produces).  A thread holds 8 complex columns of its row in v; the four
// lanes that differ in lane bits xm0 / xm1 hold the four rows that are adjacent in
// memory (C address bits 0 and 1).  A 4x4 transpose of complex elements inside that
// lane group leaves every lane with ONE column of the four rows = 32 contiguous
// bytes: one 256-bit store instead of four 8-byte stores to four 2 MB-strided lines.
// dst0 / dst1: address of (first of the 4 rows, this lane's column) for columns 0..3
// and 4..7 of the piece.
__device__ __forceinline__ void tstore_piece(float (&v)[16], float2* dst0, float2* dst1,
                                             unsigned xm0, unsigned xm1, unsigned lane) {
  const bool P = (lane & xm0) != 0, H = (lane & xm1) != 0;
#pragma unroll
  for (int g4 = 0; g4 < 2; ++g4) {
    float* e = &v[8 * g4];  // e[2i], e[2i+1] = complex column i of this group
#pragma unroll
    for (int b = 0; b < 4; b += 2) {
      const float sx = P ? e[2 * b] : e[2 * b + 2], sy = P ? e[2 * b + 1] : e[2 * b + 3];
      const float rx = __shfl_xor_sync(0xffffffffu, sx, xm0);
      const float ry = __shfl_xor_sync(0xffffffffu, sy, xm0);
      if (P) {
        e[2 * b] = rx;
        e[2 * b + 1] = ry;
      } else {
        e[2 * b + 2] = rx;
        e[2 * b + 3] = ry;
      }
    }
#pragma unroll
    for (int s0 = 0; s0 < 2; ++s0) {
      const float sx = H ? e[2 * s0] : e[2 * s0 + 4], sy = H ? e[2 * s0 + 1] : e[2 * s0 + 5];
      const float rx = __shfl_xor_sync(0xffffffffu, sx, xm1);
      const float ry = __shfl_xor_sync(0xffffffffu, sy, xm1);
      if (H) {
        e[2 * s0] = rx;
        e[2 * s0 + 1] = ry;
      } else {
        e[2 * s0 + 4] = rx;
        e[2 * s0 + 5] = ry;
      }
    }
    stg256(reinterpret_cast<float*>(g4 == 0 ? dst0 : dst1), e);
  }
}

// Setup of the transposed-store path: finds the tile-row bits that are C address
// bits 0 and 1.  Returns false when they are not both lane bits (rows = 32 *
// lane_group + lane) or C is not 32-byte aligned.
struct TStore {
  unsigned xm0, xm1;       // lane masks of the two row bits
  unsigned long long coj;  // column offset of this lane's column (n bits 0, 1)
  int rowb;                // this lane's row with the two bits cleared
};
__device__ __forceinline__ bool tstore_setup(const FusedArgs& g, const float2* C, int row, int lane,
                                             TStore* t) {
  if (g.c_natural || g.n_bits < 3) return false;
  if ((reinterpret_cast<uintptr_t>(C) & 31) != 0) return false;
  int r0 = -1, r1 = -1;
  for (int i = 0; i < 7; ++i) {
    if (g.c_m[g.row_perm[i]] == 0) r0 = i;
    if (g.c_m[g.row_perm[i]] == 1) r1 = i;
  }
  if (r0 < 0 || r1 < 0 || r0 > 4 || r1 > 4) return false;
  t->xm0 = 1u << r0;
  t->xm1 = 1u << r1;
  const int j = ((lane >> r0) & 1) | (((lane >> r1) & 1) << 1);
  t->coj = ((j & 1) ? (1ull << g.c_n[0]) : 0ull) | ((j & 2) ? (1ull << g.c_n[1]) : 0ull);
  t->rowb = row & ~(int)(t->xm0 | t->xm1);
  return true;
}

// one lane of a converged warp (the same lane every time)
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}

template <int N>
__device__ __forceinline__ void reg_dec() {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N));
}
template <int N>
__device__ __forceinline__ void reg_inc() {
  asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N));
}

// fp32 pair (re, im) -> packed bf16 hi and packed bf16 mid (re in the low half)
__device__ __forceinline__ void split_pair(float re, float im, uint32_t& hi, uint32_t& mid) {
  const __nv_bfloat162 h = __float22bfloat162_rn(make_float2(re, im));
  hi = *reinterpret_cast<const uint32_t*>(&h);
  const float hre = __uint_as_float(hi << 16), him = __uint_as_float(hi & 0xffff0000u);
  const __nv_bfloat162 m = __float22bfloat162_rn(make_float2(re - hre, im - him));
  mid = *reinterpret_cast<const uint32_t*>(&m);
}

__device__ __forceinline__ float2 lds64(uint32_t addr) {
  float2 v;
  asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c,
                                       uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c),
               "r"(d)
               : "memory");
}

// Spin on an mbarrier phase.  Bounded: a protocol bug traps instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait_spin(uint32_t bar, uint32_t parity) {
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

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// 2-D bf16 tensor [rows][cols] row-major, box = 64 cols x box_rows, 128B swizzle.
// cols may be 32 (K = 16): the box then reaches past the tensor and the
// out-of-bounds half of every 128-byte row is zero-filled (never read by the MMA).
bool make_map_b(CUtensorMap* map, void* base, uint64_t rows, uint64_t cols, uint32_t box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) return false;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {cols * 2};
  cuuint32_t box[2] = {(cuuint32_t)BK, box_rows};
  cuuint32_t estr[2] = {1, 1};
  return fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
            CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

struct TileBit {
  int addr, kind, idx;  // kind 0 = row (m) bit, 1 = k bit
};

// Gathered fp32 view of A: the tile's address bits, grouped into runs of
// consecutive addresses, are the box dimensions; dimension 1 is an "offset"
// dimension (stride 16 bytes, box 1) whose coordinate carries the tile base.
// Fills the converter XOR tables.  Returns false when the tile cannot be
// expressed (caller falls back to the pre-split path).
bool make_map_a(CUtensorMap* map, const float2* A, const qfx_contract_desc& d, int kb_bits,
                uint64_t span_elems, FusedArgs* g) {
  EncodeTiledFn fn = encode_fn();
  if (fn == nullptr) return false;
  if ((reinterpret_cast<uintptr_t>(A) & 15) != 0) return false;
  std::vector<TileBit> bits;
  for (int i = 0; i < 7; ++i) bits.push_back({d.a_m[i], 0, i});
  for (int j = 0; j < kb_bits; ++j) bits.push_back({d.a_k[j], 1, j});
  std::sort(bits.begin(), bits.end(),
            [](const TileBit& x, const TileBit& y) { return x.addr < y.addr; });
  if (bits[0].addr != 0) return false;
  // runs of consecutive addresses (first run <= 7 bits: 2 * 2^7 floats = 256 box limit)
  std::vector<std::pair<int, int>> runs;  // (first addr, length)
  static const int run0_cap = [] {
    const char* e = getenv("QFLEX_MAP_RUN0_CAP");
    return e ? std::max(1, std::min(7, atoi(e))) : 7;
  }();
  for (size_t s = 0; s < bits.size(); ++s) {
    const int cap = runs.size() == 1 ? run0_cap : 8;
    if (!runs.empty() && runs.back().first + runs.back().second == bits[s].addr &&
        runs.back().second < cap)
      runs.back().second++;
    else
      runs.push_back({bits[s].addr, 1});
  }
  if (runs.size() > 4) return false;
  const int rank = (int)runs.size() + 1;
  cuuint64_t dims[5];
  cuuint64_t strides[4];
  cuuint32_t box[5], estr[5] = {1, 1, 1, 1, 1};
  dims[0] = 2ull << runs[0].second;  // floats
  box[0] = (cuuint32_t)dims[0];
  dims[1] = (span_elems + 1) / 2;    // tile base in units of 16 bytes
  box[1] = 1;
  strides[0] = 16;
  for (size_t r = 1; r < runs.size(); ++r) {
    dims[r + 1] = 1ull << runs[r].second;
    box[r + 1] = (cuuint32_t)dims[r + 1];
    strides[r] = 8ull << runs[r].first;
  }
  if (fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)rank, const_cast<float2*>(A), dims,
         strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
    return false;
  g->a_rank = rank;
  // ---- converter tables.  smem element bit s of the fp32 stage is tile bit bits[s].
  const int nb = (int)bits.size();
  int sk0 = -1, sk1 = -1;
  for (int s = 0; s < nb; ++s)
    if (bits[s].kind == 1 && bits[s].idx == 0) sk0 = s;
    else if (bits[s].kind == 1 && bits[s].idx == 1) sk1 = s;
  std::vector<int> item;  // smem bits enumerated by the item index (all but k0, k1)
  for (int s = 0; s < nb; ++s)
    if (s != sk0 && s != sk1) item.push_back(s);
  // tile-row bit order: the three lowest row bits take part in the 128B swizzle
  // (chunk ^= row & 7); pick them so that lanes 0..2 of a converter warp hit 8
  // distinct 16-byte bank groups: a lane that walks k bit 2+j owns slot j, a lane
  // that walks an m bit takes a free slot as row bit.
  int slot_m[3] = {-1, -1, -1};
  bool slot_k[3] = {false, false, false};
  for (int i = 0; i < 3; ++i)
    if (bits[item[i]].kind == 1) slot_k[bits[item[i]].idx - 2] = true;
  for (int i = 0; i < 3; ++i)
    if (bits[item[i]].kind == 0)
      for (int j = 0; j < 3; ++j)
        if (!slot_k[j] && slot_m[j] < 0) {
          slot_m[j] = bits[item[i]].idx;
          break;
        }
  bool used[7] = {false, false, false, false, false, false, false};
  for (int j = 0; j < 3; ++j)
    if (slot_m[j] >= 0) used[slot_m[j]] = true;
  int next = 0;
  int pos_of_m[7];
  for (int q = 0; q < 7; ++q) {
    int mi = q < 3 ? slot_m[q] : -1;
    if (mi < 0) {
      while (used[next]) ++next;
      mi = next;
      used[mi] = true;
    }
    g->row_perm[q] = (int8_t)mi;
    pos_of_m[mi] = q;
  }
  auto dst_mask = [&](const TileBit& tb) -> uint32_t {
    if (tb.kind == 1) return 1u << (4 + (tb.idx - 2));
    const int q = pos_of_m[tb.idx];
    return (1u << (7 + q)) | (q < 3 ? (1u << (4 + q)) : 0u);
  };
  const int n_item = (int)item.size();  // 9 (K = 16) or 10
  for (int b = 0; b < 8; ++b) {
    g->cv_src[b] = 8u << item[b];
    g->cv_dst[b] = dst_mask(bits[item[b]]);
  }
  g->cv_its = 1 << (n_item - 8);
  for (int it = 0; it < 4; ++it) {
    uint32_t xs = 0, xd = 0;
    for (int b = 0; 8 + b < n_item; ++b)
      if ((it >> b) & 1) {
        xs ^= 8u << item[8 + b];
        xd ^= dst_mask(bits[item[8 + b]]);
      }
    g->cv_src_it[it] = xs;
    g->cv_dst_it[it] = xd;
  }
  g->cv_ok0 = 8u << sk0;
  g->cv_ok1 = 8u << sk1;
  g->cv_vec = (sk0 == 0 && sk1 == 1) ? 1 : 0;
  return true;
}

// Host mirror of the epilogue's store-path choice (template parameter STORE).
inline int pick_store_mode(const FusedArgs& g, const float2* C) {
  if (g.dbg & 32) return 0;
  const int npieces = (g.bn >> 1) >> 4;
  if (g.c_natural) return npieces >= 4 ? 1 : 0;
  if (g.n_bits < 3 || (reinterpret_cast<uintptr_t>(C) & 31) != 0) return 0;
  int r0 = -1, r1 = -1;
  for (int i = 0; i < 7; ++i) {
    if (g.c_m[g.row_perm[i]] == 0) r0 = i;
    if (g.c_m[g.row_perm[i]] == 1) r1 = i;
  }
  return (r0 >= 0 && r1 >= 0 && r0 <= 4 && r1 <= 4) ? 2 : 0;
}

}  // namespace
