// Tensor-core contraction path for compute-bound contractions (d >= 32):
// tcgen05.mma (kind::f16, bf16 inputs, fp32 accumulators in TMEM) with operands
// staged in shared memory by TMA, split-bf16 emulation of fp32:
//     a*b ~= a_hi*b_hi + a_hi*b_mid + a_mid*b_hi         (rel. error ~ 2^-17)
// The complex product is run as ONE real GEMM on the raw float views
//     C'[m][2n+c] = sum_k' A'[m][k'] * B'[2n+c][k']        (see tc_split.cu)
// so no flops are wasted on the real/imag split (2*M*2N*2K = 8MNK).
//
// Replaces cblas_cgemm of the reference (src/tensor.cpp:960-980) for the
// contraction classes SURVEY 8d marks "tensor-core candidates".
//
// Kernel structure (persistent, one CTA per SM, 192 threads):
//   warp 0   : TMA producer   (4 tiles per stage: A_hi, A_mid, B_hi, B_mid)
//   warp 1   : MMA issuer     (one elected lane; 12 tcgen05.mma per 64-wide k block)
//   warps 2-5: epilogue       (tcgen05.ld TMEM -> registers -> global C)
//   TMEM accumulators are double-buffered so the epilogue of tile i overlaps
//   the MMAs of tile i+1.
#include <cuda.h>

#include <algorithm>
#include <cstring>

#include "qfx_internal.h"

int qfx_launch_split(qfx_ctx* ctx, const float2* src, int mode, const int8_t* row_pos,
                     int row_bits, const int8_t* k_pos, int k_bits, uint32_t* hi,
                     uint32_t* mid);

namespace {

constexpr int BM = 128;
constexpr int BK = 64;  // bf16 elements per k block = 128 bytes = one swizzle row
constexpr int STAGES = 2;
constexpr int MAX_BN = 256;
constexpr int A_TILE_BYTES = BM * BK * 2;                                  // 16 KB
constexpr int STAGE_BYTES = 2 * A_TILE_BYTES + 2 * MAX_BN * BK * 2;       // 96 KB
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/;
constexpr int TC_THREADS = 320;  // TMA warp + MMA warp + 8 epilogue warps
constexpr int MAX_PIECES = MAX_BN / 2 / 32;  // 32-column pieces per epilogue thread

struct TcArgs {
  float2* C;
  int m_bits, n_bits;  // complex extents
  int kblocks;         // (2K) / 64
  int chunk;           // k blocks accumulated in TMEM before the fp32 fold
  int nchunks;
  int bn;              // real columns per tile: min(2N, 256)
  int tiles_n;         // 2N / bn
  int tiles;           // tiles_m * tiles_n
  int c_natural;       // C rows are contiguous runs of N complex
  int8_t c_m[QFX_MAX_BITS], c_n[QFX_MAX_BITS];
};

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

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]),
        "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]),
        "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__global__ void __launch_bounds__(TC_THREADS, 1)
    tc_gemm_bf16x3(const __grid_constant__ CUtensorMap tmAh, const __grid_constant__ CUtensorMap tmAm,
                   const __grid_constant__ CUtensorMap tmBh, const __grid_constant__ CUtensorMap tmBm,
                   const __grid_constant__ TcArgs g) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~(uintptr_t)1023);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  // bars[0..1] full, [2..3] empty, [4..5] tmem_full, [6..7] tmem_empty, [8] tmem ptr
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t bar_base = smem_u32(bars);
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (2 + s); };
  auto tfull_bar = [&](int s) { return bar_base + 8u * (4 + s); };
  auto tempty_bar = [&](int s) { return bar_base + 8u * (6 + s); };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bn = g.bn;
  const uint32_t tmem_cols = (2 * bn < 32) ? 32u : (uint32_t)(2 * bn);

  if (warp == 0 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
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

  const uint32_t stage_tx = 2u * A_TILE_BYTES + 2u * (uint32_t)bn * BK * 2u;

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < g.tiles; t += gridDim.x) {
        const int nt = t % g.tiles_n, mt = t / g.tiles_n;
        for (int kb = 0; kb < g.kblocks; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1u);
          mbar_expect_tx(full_bar(stage), stage_tx);
          const uint32_t sb = smem_base + (uint32_t)stage * STAGE_BYTES;
          tma_load_2d(&tmAh, full_bar(stage), sb, kb * BK, mt * BM);
          tma_load_2d(&tmAm, full_bar(stage), sb + A_TILE_BYTES, kb * BK, mt * BM);
          tma_load_2d(&tmBh, full_bar(stage), sb + 2 * A_TILE_BYTES, kb * BK, nt * bn);
          tma_load_2d(&tmBm, full_bar(stage), sb + 2 * A_TILE_BYTES + bn * BK * 2, kb * BK,
                      nt * bn);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1u;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // instruction descriptor: D=f32, A=B=bf16, both K-major, N=bn, M=128
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(bn >> 3) << 17) |
                             ((uint32_t)(BM >> 4) << 24);
      int stage = 0, acc = 0;
      uint32_t phase = 0, acc_phase = 0;
      for (int t = blockIdx.x; t < g.tiles; t += gridDim.x) {
        for (int ch = 0; ch < g.nchunks; ++ch) {
          mbar_wait(tempty_bar(acc), acc_phase ^ 1u);
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + (uint32_t)(acc * bn);
          const int kb_end = min(g.kblocks, (ch + 1) * g.chunk);
          for (int kb = ch * g.chunk; kb < kb_end; ++kb) {
            mbar_wait(full_bar(stage), phase);
            tc_fence_after();
            const uint32_t sb = smem_base + (uint32_t)stage * STAGE_BYTES;
            const uint64_t dAh = smem_desc_sw128(sb);
            const uint64_t dAm = smem_desc_sw128(sb + A_TILE_BYTES);
            const uint64_t dBh = smem_desc_sw128(sb + 2 * A_TILE_BYTES);
            const uint64_t dBm = smem_desc_sw128(sb + 2 * A_TILE_BYTES + bn * BK * 2);
            const bool first = (kb == ch * g.chunk);
#pragma unroll
            for (int kk = 0; kk < BK / 16; ++kk) {
              const uint64_t adv = (uint64_t)(kk * 32 >> 4);  // 16 bf16 = 32 bytes along K
              umma_bf16(d_tmem, dAm + adv, dBh + adv, idesc, (first && kk == 0) ? 0u : 1u);
              umma_bf16(d_tmem, dAh + adv, dBm + adv, idesc, 1u);
              umma_bf16(d_tmem, dAh + adv, dBh + adv, idesc, 1u);
            }
            umma_commit(empty_bar(stage));  // frees the smem stage when the MMAs retire
            if (++stage == STAGES) {
              stage = 0;
              phase ^= 1u;
            }
          }
          umma_commit(tfull_bar(acc));  // chunk accumulator complete
          if (++acc == 2) {
            acc = 0;
            acc_phase ^= 1u;
          }
        }
      }
    }
  } else {
    // epilogue: warp w may only touch TMEM lanes 32*(w%4) .. +31; two warps share
    // a lane group and split the tile's columns.  The tensor core accumulates
    // with truncation, which biases long sums towards zero, so the accumulator
    // only integrates `chunk` k blocks; chunks are folded here in fp32 with
    // round-to-nearest adds, in registers, and C is written once per tile.
    const int lane_group = warp & 3;
    const int half = (warp - 2) >> 2;
    const int row = lane_group * 32 + lane;
    const int half_cols = bn >> 1;
    const int npieces = half_cols >> 5;
    int acc = 0;
    uint32_t acc_phase = 0;
    const unsigned long long N = 1ull << g.n_bits;
    for (int t = blockIdx.x; t < g.tiles; t += gridDim.x) {
      const int nt = t % g.tiles_n, mt = t / g.tiles_n;
      float sum[MAX_PIECES][32];
      for (int ch = 0; ch < g.nchunks; ++ch) {
        mbar_wait(tfull_bar(acc), acc_phase);
        tc_fence_after();
        const uint32_t t_row = tmem_base + ((uint32_t)(lane_group * 32) << 16) +
                               (uint32_t)(acc * bn + half * half_cols);
#pragma unroll
        for (int p = 0; p < MAX_PIECES; ++p) {
          if (p < npieces) {
            uint32_t r[32];
            tmem_ld32(t_row + (uint32_t)(p * 32), r);
            if (ch == 0) {
#pragma unroll
              for (int j = 0; j < 32; ++j) sum[p][j] = __uint_as_float(r[j]);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) sum[p][j] += __uint_as_float(r[j]);
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tempty_bar(acc));
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1u;
        }
      }
      const unsigned long long m = (unsigned long long)mt * BM + row;
      unsigned long long row_off = 0;
      if (g.c_natural) {
        row_off = m * N;
      } else {
        for (int q = 0; q < g.m_bits; ++q) row_off |= ((m >> q) & 1ull) << g.c_m[q];
      }
#pragma unroll
      for (int p = 0; p < MAX_PIECES; ++p) {
        if (p < npieces) {
          // complex column of the first element of this piece
          const unsigned long long n0 =
              ((unsigned long long)nt * bn + half * half_cols + p * 32) >> 1;
          if (g.c_natural) {
            float4* dst = reinterpret_cast<float4*>(g.C + row_off + n0);
#pragma unroll
            for (int j = 0; j < 8; ++j)
              dst[j] = make_float4(sum[p][4 * j], sum[p][4 * j + 1], sum[p][4 * j + 2],
                                   sum[p][4 * j + 3]);
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const unsigned long long n = n0 + j;
              unsigned long long co = 0;
              for (int q = 0; q < g.n_bits; ++q) co |= ((n >> q) & 1ull) << g.c_n[q];
              g.C[row_off + co] = make_float2(sum[p][2 * j], sum[p][2 * j + 1]);
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

// 2-D bf16 tensor [rows][cols] row-major, box = 64 cols x box_rows, 128B swizzle
bool make_map(CUtensorMap* map, void* base, uint64_t rows, uint64_t cols, uint32_t box_rows) {
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

}  // namespace

// Eligibility: called with roles already swapped (m_bits >= n_bits).
bool qfx_tc_eligible(const qfx_contract_desc& d) {
  if (d.m_bits < 7) return false;                  // one full 128-row tile
  if (d.n_bits < 4) return false;                  // UMMA N = 2N >= 32 (pre-split path: >= 64)
  if (d.k_bits < 4) return false;                  // 2K >= 32 (pre-split path: one 64-wide k block)
  const double flops = 8.0 * (double)(1ull << d.m_bits) * (double)(1ull << d.n_bits) *
                       (double)(1ull << d.k_bits);
  return flops >= (double)(1ull << 27);
}

int qfx_launch_contract_tc(qfx_ctx* ctx, const qfx_contract_desc& d, const float2* A,
                           const float2* B, float2* C) {
  const uint64_t M = 1ull << d.m_bits, N = 1ull << d.n_bits, K = 1ull << d.k_bits;
  // workspace: A' hi/mid (M x 2K bf16 each) and B' hi/mid (2N x 2K bf16 each)
  const size_t a_bytes = (size_t)M * K * 4, b_bytes = (size_t)N * K * 8;
  const size_t need = 2 * a_bytes + 2 * b_bytes + 4096;
  int rc = qfx_ensure_ws(ctx, need);
  if (rc) return rc;
  uint8_t* ws = reinterpret_cast<uint8_t*>(ctx->ws);
  uint32_t* a_hi = reinterpret_cast<uint32_t*>(ws);
  uint32_t* a_mid = reinterpret_cast<uint32_t*>(ws + a_bytes);
  uint32_t* b_hi = reinterpret_cast<uint32_t*>(ws + 2 * a_bytes);
  uint32_t* b_mid = reinterpret_cast<uint32_t*>(ws + 2 * a_bytes + b_bytes);
  rc = qfx_launch_split(ctx, A, 0, d.a_m, d.m_bits, d.a_k, d.k_bits, a_hi, a_mid);
  if (rc) return rc;
  rc = qfx_launch_split(ctx, B, 1, d.b_n, d.n_bits, d.b_k, d.k_bits, b_hi, b_mid);
  if (rc) return rc;

  TcArgs g{};
  g.C = C;
  g.m_bits = d.m_bits;
  g.n_bits = d.n_bits;
  g.kblocks = (int)((2 * K) / BK);
  g.chunk = ctx->tc_chunk > 0 ? ctx->tc_chunk : 2;
  g.nchunks = (g.kblocks + g.chunk - 1) / g.chunk;
  g.bn = (int)std::min<uint64_t>(2 * N, MAX_BN);
  g.tiles_n = (int)((2 * N) / g.bn);
  g.tiles = (int)((M / BM) * g.tiles_n);
  g.c_natural = 1;
  for (int i = 0; i < d.n_bits; ++i)
    if (d.c_n[i] != i) g.c_natural = 0;
  for (int i = 0; i < d.m_bits; ++i)
    if (d.c_m[i] != d.n_bits + i) g.c_natural = 0;
  if ((reinterpret_cast<uintptr_t>(C) & 15) != 0) g.c_natural = 0;
  memcpy(g.c_m, d.c_m, QFX_MAX_BITS);
  memcpy(g.c_n, d.c_n, QFX_MAX_BITS);

  CUtensorMap mAh, mAm, mBh, mBm;
  if (!make_map(&mAh, a_hi, M, 2 * K, BM) || !make_map(&mAm, a_mid, M, 2 * K, BM) ||
      !make_map(&mBh, b_hi, 2 * N, 2 * K, (uint32_t)g.bn) ||
      !make_map(&mBm, b_mid, 2 * N, 2 * K, (uint32_t)g.bn)) {
    qfx_set_error("cuTensorMapEncodeTiled failed");
    return QFX_ERR_CUDA;
  }
  static qfx_per_device_once attr;
  if (attr.need(ctx->device)) {
    QFX_CUDA(cudaFuncSetAttribute(tc_gemm_bf16x3, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  SMEM_BYTES));
    attr.mark(ctx->device);
  }
  const int grid = std::min(g.tiles, ctx->sm_count);
  tc_gemm_bf16x3<<<grid, TC_THREADS, SMEM_BYTES, ctx->stream>>>(mAh, mAm, mBh, mBm, g);
  QFX_CUDA(cudaGetLastError());
  ctx->stats.launches_gemm_tc++;
  ctx->stats.launches_total++;
  return QFX_OK;
}
