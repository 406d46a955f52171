// Pre-pass of the tensor-core contraction path: gather an operand through its
// bit tables (this IS the reference's Tensor::reorder, src/tensor.cpp:866-911,
// as one pass) and, in the same pass, split every fp32 into bf16 hi + bf16 mid
// (x ~= hi + mid, |x - hi - mid| <= 2^-17 |x|) and lay the result out as the
// K-major real matrices the tcgen05 kernel consumes:
//
//   A  (complex M x K)  ->  A'[m][2k+c]            = (re, im) of A[m,k]      (M  x 2K)
//   B  (complex K x N)  ->  B'[2n  ][2k], [2k+1]   = ( Br, -Bi)              (2N x 2K)
//                            B'[2n+1][2k], [2k+1]   = ( Bi,  Br)
//
// so that C'[m][2n+c] = sum_k' A'[m][k'] B'[2n+c][k'] is the complex product
// with C' the raw float view of C.  Same tiling idea as permute_bits: a CTA owns
// the low source bits plus the pre-images of the low destination bits, so both
// global sides move contiguous runs.
#include <cuda_bf16.h>

#include <algorithm>
#include <vector>

#include "qfx_internal.h"

namespace {

constexpr int kRunBits = 5;
constexpr int kTileBits = 10;
constexpr int kBigRunBits = 6;   // big tiles: >= 512-byte runs on both sides
constexpr int kBigTileBits = 12; // 4096 elements = 32 KB of shared memory
constexpr int kMaxTileBits = 12;
constexpr int kThreads = 256;

struct SplitArgs {
  const float2* src;
  uint32_t* hi;   // 2 x bf16 per complex element
  uint32_t* mid;
  int t_bits, o_bits, k_bits;
  int8_t e_src[kMaxTileBits];
  int8_t f_dst[kMaxTileBits];
  int8_t f_to_e[kMaxTileBits];
  int8_t o_src[QFX_MAX_BITS], o_dst[QFX_MAX_BITS];
};

__device__ __forceinline__ void split2(float x, __nv_bfloat16& h, __nv_bfloat16& m) {
  h = __float2bfloat16_rn(x);
  m = __float2bfloat16_rn(x - __bfloat162float(h));
}

__device__ __forceinline__ uint32_t pack(__nv_bfloat16 lo, __nv_bfloat16 hi) {
  return (uint32_t)__bfloat16_as_ushort(lo) | ((uint32_t)__bfloat16_as_ushort(hi) << 16);
}

// MODE 0: operand A (row = m, natural [M][K]); MODE 1: operand B (expanded rows)
template <int MODE>
__global__ void __launch_bounds__(kThreads) split_gather(const __grid_constant__ SplitArgs g) {
  // Programmatic dependent launch: the contraction kernel that follows may start its
  // prologue (barrier init, TMEM allocation, first A loads) now; it waits
  // (griddepcontrol.wait) before it reads what this kernel writes.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  __shared__ float2 tile[1 << kTileBits];
  __shared__ unsigned src_off[1 << kTileBits];
  __shared__ unsigned dst_off[1 << kTileBits];
  __shared__ unsigned short e_of_f[1 << kTileBits];
  const int T = 1 << g.t_bits;
  for (int x = threadIdx.x; x < T; x += kThreads) {
    unsigned so = 0, dof = 0, e = 0;
    for (int i = 0; i < g.t_bits; ++i) {
      const unsigned b = (x >> i) & 1u;
      so |= b << g.e_src[i];
      dof |= b << g.f_dst[i];
      e |= b << g.f_to_e[i];
    }
    src_off[x] = so;
    dst_off[x] = dof;
    e_of_f[x] = (unsigned short)e;
  }
  __syncthreads();
  const unsigned long long n_outer = 1ull << g.o_bits;
  const unsigned long long kmask = (1ull << g.k_bits) - 1ull;
  for (unsigned long long o = blockIdx.x; o < n_outer; o += gridDim.x) {
    unsigned long long sb = 0, db = 0;
    for (int i = 0; i < g.o_bits; ++i) {
      const unsigned long long b = (o >> i) & 1ull;
      sb |= b << g.o_src[i];
      db |= b << g.o_dst[i];
    }
    const float2* __restrict__ s = g.src + sb;
    for (int x = threadIdx.x; x < T; x += kThreads) tile[x] = __ldg(s + src_off[x]);
    __syncthreads();
    for (int x = threadIdx.x; x < T; x += kThreads) {
      const float2 v = tile[e_of_f[x]];
      const unsigned long long d = db + dst_off[x];  // = row*K + k (complex units)
      __nv_bfloat16 rh, rm, ih, im;
      split2(v.x, rh, rm);
      split2(v.y, ih, im);
      if (MODE == 0) {
        g.hi[d] = pack(rh, ih);
        g.mid[d] = pack(rm, im);
      } else {
        const unsigned long long k = d & kmask, row0 = 2ull * (d - k) + k;
        const unsigned long long row1 = row0 + (kmask + 1ull);
        g.hi[row0] = pack(rh, __hneg(ih));
        g.mid[row0] = pack(rm, __hneg(im));
        g.hi[row1] = pack(ih, rh);
        g.mid[row1] = pack(im, rm);
      }
    }
    __syncthreads();
  }
}

// XOR swizzle of the tile index (bits 1..3 <- higher bits): whatever subset of
// index bits a warp varies, its 8-byte accesses spread over the banks; pairs
// (bit 0) stay adjacent for the 16-byte stores.
__device__ __forceinline__ unsigned swz(unsigned e) {
  return e ^ ((((e >> 4) ^ (e >> 7) ^ (e >> 10)) & 7u) << 1);
}

// Operand A, tensors of >= 2^12 elements: 4096-element tiles, every thread keeps
// 8 independent 16-byte loads in flight and writes 16-byte runs of hi and mid.
// Offsets split into a per-thread part (registers) and a per-iteration part
// (tiny shared tables), so the inner loops are one add per access.
__global__ void __launch_bounds__(kThreads) split_a_big(const __grid_constant__ SplitArgs g) {
  __shared__ __align__(16) float2 tile[1 << kBigTileBits];
  __shared__ unsigned s_hi[8], d_hi[4], e_hi[4];
  const unsigned tid = threadIdx.x;
  // source side: pair u = it*256 + tid, element e0 = 2u (bit 0 clear)
  unsigned s_lo = 0;
  for (int b = 0; b < 8; ++b) s_lo |= ((tid >> b) & 1u) << g.e_src[b + 1];
  // destination side: quad q = it*256 + tid, element f0 = 4q
  unsigned d_lo = 0, e_lo = 0;
  for (int b = 0; b < 8; ++b) {
    d_lo |= ((tid >> b) & 1u) << g.f_dst[b + 2];
    e_lo |= ((tid >> b) & 1u) << g.f_to_e[b + 2];
  }
  if (tid < 8) {
    unsigned v = 0;
    for (int b = 0; b < 3; ++b) v |= ((tid >> b) & 1u) << g.e_src[b + 9];
    s_hi[tid] = v;
  }
  if (tid < 4) {
    unsigned v = 0, w = 0;
    for (int b = 0; b < 2; ++b) {
      v |= ((tid >> b) & 1u) << g.f_dst[b + 10];
      w |= ((tid >> b) & 1u) << g.f_to_e[b + 10];
    }
    d_hi[tid] = v;
    e_hi[tid] = w;
  }
  const unsigned e_j1 = 1u << g.f_to_e[0], e_j2 = 1u << g.f_to_e[1];
  const unsigned s_d1 = 1u << g.e_src[0];
  const bool vec = (g.e_src[0] == 0) && ((reinterpret_cast<uintptr_t>(g.src) & 15) == 0);
  __syncthreads();
  const unsigned long long n_outer = 1ull << g.o_bits;
  for (unsigned long long o = blockIdx.x; o < n_outer; o += gridDim.x) {
    unsigned long long sb = 0, db = 0;
    for (int i = 0; i < g.o_bits; ++i) {
      const unsigned long long b = (o >> i) & 1ull;
      sb |= b << g.o_src[i];
      db |= b << g.o_dst[i];
    }
    const float2* __restrict__ s = g.src + sb + s_lo;
    float4 v[8];
    if (vec) {
#pragma unroll
      for (int it = 0; it < 8; ++it) v[it] = __ldg(reinterpret_cast<const float4*>(s + s_hi[it]));
    } else {
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        const float2 a = __ldg(s + s_hi[it]), b = __ldg(s + s_hi[it] + s_d1);
        v[it] = make_float4(a.x, a.y, b.x, b.y);
      }
    }
#pragma unroll
    for (int it = 0; it < 8; ++it)
      *reinterpret_cast<float4*>(&tile[swz(2u * (it * 256u + tid))]) = v[it];
    __syncthreads();
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const unsigned e0 = e_lo | e_hi[it];
      uint32_t h[4], m[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 x = tile[swz(e0 + ((j & 1) ? e_j1 : 0u) + ((j & 2) ? e_j2 : 0u))];
        __nv_bfloat16 rh, rm, ih, im;
        split2(x.x, rh, rm);
        split2(x.y, ih, im);
        h[j] = pack(rh, ih);
        m[j] = pack(rm, im);
      }
      const unsigned long long d = db + d_lo + d_hi[it];
      *reinterpret_cast<uint4*>(g.hi + d) = make_uint4(h[0], h[1], h[2], h[3]);
      *reinterpret_cast<uint4*>(g.mid + d) = make_uint4(m[0], m[1], m[2], m[3]);
    }
    __syncthreads();
  }
}

}  // namespace

// row_pos/k_pos: operand address bit of every row (m or n) bit and of every k
// bit.  Output row-major [rows][K] complex order (k lowest).
int qfx_launch_split(qfx_ctx* ctx, const float2* src, int mode, const int8_t* row_pos,
                     int row_bits, const int8_t* k_pos, int k_bits, uint32_t* hi,
                     uint32_t* mid) {
  const int n = row_bits + k_bits;
  std::vector<int> sp(n), dp(n);
  for (int i = 0; i < k_bits; ++i) {
    sp[i] = k_pos[i];
    dp[i] = i;
  }
  for (int i = 0; i < row_bits; ++i) {
    sp[k_bits + i] = row_pos[i];
    dp[k_bits + i] = k_bits + i;
  }
  // tile: bits lowest in the source + bits lowest in the destination, padded
  std::vector<int> by_src(n);
  for (int i = 0; i < n; ++i) by_src[i] = i;
  std::sort(by_src.begin(), by_src.end(), [&](int x, int y) { return sp[x] < sp[y]; });
  const bool big = (mode == 0 && n >= kBigTileBits);
  const int run_bits = big ? kBigRunBits : kRunBits;
  const int tile_bits = big ? kBigTileBits : kTileBits;
  std::vector<char> in_tile(n, 0);
  int t = 0;
  for (int r = 0; r < n && r < run_bits; ++r) {
    in_tile[by_src[r]] = 1;
    ++t;
  }
  for (int i = 0; i < n; ++i)
    if (dp[i] < run_bits && !in_tile[i]) {
      in_tile[i] = 1;
      ++t;
    }
  for (int r = 0; r < n && t < tile_bits; ++r)
    if (!in_tile[by_src[r]]) {
      in_tile[by_src[r]] = 1;
      ++t;
    }
  SplitArgs g{};
  g.src = src;
  g.hi = hi;
  g.mid = mid;
  g.t_bits = t;
  g.o_bits = n - t;
  g.k_bits = k_bits;
  std::vector<int> tb;  // tile bits ascending by source position
  int oi = 0;
  for (int r = 0; r < n; ++r) {
    const int i = by_src[r];
    if (in_tile[i]) {
      tb.push_back(i);
    } else {
      g.o_src[oi] = (int8_t)sp[i];
      g.o_dst[oi] = (int8_t)dp[i];
      ++oi;
    }
  }
  for (int i = 0; i < t; ++i) g.e_src[i] = (int8_t)sp[tb[i]];
  std::vector<int> by_dst(t);
  for (int i = 0; i < t; ++i) by_dst[i] = i;
  std::sort(by_dst.begin(), by_dst.end(), [&](int x, int y) { return dp[tb[x]] < dp[tb[y]]; });
  for (int i = 0; i < t; ++i) {
    g.f_dst[i] = (int8_t)dp[tb[by_dst[i]]];
    g.f_to_e[i] = (int8_t)by_dst[i];
  }
  const unsigned long long n_outer = 1ull << g.o_bits;
  const unsigned blocks =
      (unsigned)std::min<unsigned long long>(n_outer, 16ull * ctx->sm_count);
  if (big && t == kBigTileBits && g.f_dst[0] == 0 && g.f_dst[1] == 1)
    split_a_big<<<blocks, kThreads, 0, ctx->stream>>>(g);
  else if (mode == 0)
    split_gather<0><<<blocks, kThreads, 0, ctx->stream>>>(g);
  else
    split_gather<1><<<blocks, kThreads, 0, ctx->stream>>>(g);
  QFX_CUDA(cudaGetLastError());
  ctx->stats.launches_permute++;
  ctx->stats.launches_total++;
  return QFX_OK;
}
