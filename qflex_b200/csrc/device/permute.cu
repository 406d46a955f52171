// Single-pass index permutation (bit permutation of the linear offset).
//
// Replaces Tensor::reorder -> _fast_reorder / _left_reorder / _right_reorder /
// _generate_binary_reordering_map (reference src/tensor.cpp:494-911,1250-1303):
// the reference needs up to three passes (L-R-L, ~10x tensor bytes of traffic)
// plus memoised size_t maps; here every element is read once and written once.
//
// A CTA owns a tile made of the low source bits and the (pre-images of the) low
// destination bits, so both the global reads and the global writes are
// contiguous runs of >= 32 elements (256 B); the transpose happens in shared
// memory.
#include <algorithm>
#include <cstring>
#include <vector>

#include "qfx_internal.h"

namespace {

constexpr int kRunBits = 5;    // 32 elements = 256 B contiguous on each side
constexpr int kTileBits = 10;  // up to 1024 elements (8 KB) per tile
constexpr int kPermThreads = 256;

struct PermArgs {
  const float2* src;
  float2* dst;
  int n_bits, t_bits, o_bits;
  int8_t e_src[kTileBits];   // tile bit i (ascending source position): src pos
  int8_t f_dst[kTileBits];   // tile bit i (ascending destination position): dst pos
  int8_t f_to_e[kTileBits];  // ... and which e-bit it is
  int8_t o_src[QFX_MAX_BITS], o_dst[QFX_MAX_BITS];  // remaining bits
};

__global__ void __launch_bounds__(kPermThreads)
    permute_bits(const __grid_constant__ PermArgs g) {
  __shared__ float2 tile[1 << kTileBits];
  __shared__ unsigned src_off[1 << kTileBits];
  __shared__ unsigned dst_off[1 << kTileBits];
  __shared__ unsigned short e_of_f[1 << kTileBits];
  const int T = 1 << g.t_bits;
  for (int x = threadIdx.x; x < T; x += kPermThreads) {
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
  for (unsigned long long o = blockIdx.x; o < n_outer; o += gridDim.x) {
    unsigned long long sb = 0, db = 0;
    for (int i = 0; i < g.o_bits; ++i) {
      const unsigned long long b = (o >> i) & 1ull;
      sb |= b << g.o_src[i];
      db |= b << g.o_dst[i];
    }
    const float2* __restrict__ s = g.src + sb;
    float2* __restrict__ d = g.dst + db;
    for (int x = threadIdx.x; x < T; x += kPermThreads) tile[x] = __ldg(s + src_off[x]);
    __syncthreads();
    for (int x = threadIdx.x; x < T; x += kPermThreads) d[dst_off[x]] = tile[e_of_f[x]];
    __syncthreads();
  }
}

}  // namespace

int qfx_launch_permute(qfx_ctx* ctx, uint64_t src_off, uint64_t dst_off,
                       int n_bits, const int8_t* dst_pos) {
  if (ctx == nullptr || (n_bits > 0 && dst_pos == nullptr) || n_bits < 0 ||
      n_bits > 32) {
    qfx_set_error("qfx_permute: bad argument");
    return QFX_ERR_ARG;
  }
  uint64_t seen = 0;
  for (int i = 0; i < n_bits; ++i) {
    if (dst_pos[i] < 0 || dst_pos[i] >= n_bits || (seen >> dst_pos[i]) & 1) {
      qfx_set_error("qfx_permute: dst_pos is not a permutation");
      return QFX_ERR_ARG;
    }
    seen |= 1ull << dst_pos[i];
  }
  const uint64_t n = 1ull << n_bits;
  if (src_off + n > ctx->arena_elems || dst_off + n > ctx->arena_elems) {
    qfx_set_error("qfx_permute: span outside the arena");
    return QFX_ERR_ARG;
  }
  if (src_off < dst_off + n && dst_off < src_off + n) {
    qfx_set_error("qfx_permute: source and destination overlap");
    return QFX_ERR_ARG;
  }
  return qfx_launch_permute_ptr(ctx, ctx->arena + src_off, ctx->arena + dst_off, n_bits, dst_pos);
}

// Same on raw device pointers (dst_pos already validated: a permutation of 0..n_bits-1).
int qfx_launch_permute_ptr(qfx_ctx* ctx, const float2* src, float2* dst, int n_bits,
                           const int8_t* dst_pos) {
  // tile = low source bits U pre-images of low destination bits, padded with
  // further low source bits up to kTileBits
  std::vector<char> in_tile(n_bits, 0);
  int t = 0;
  for (int b = 0; b < n_bits && b < kRunBits; ++b) {
    in_tile[b] = 1;
    ++t;
  }
  for (int b = 0; b < n_bits; ++b)
    if (dst_pos[b] < kRunBits && !in_tile[b]) {
      in_tile[b] = 1;
      ++t;
    }
  for (int b = 0; b < n_bits && t < kTileBits; ++b)
    if (!in_tile[b]) {
      in_tile[b] = 1;
      ++t;
    }
  PermArgs g{};
  g.src = src;
  g.dst = dst;
  g.n_bits = n_bits;
  g.t_bits = t;
  g.o_bits = n_bits - t;
  std::vector<int> tb;  // tile bits ascending by source position
  int oi = 0;
  for (int b = 0; b < n_bits; ++b) {
    if (in_tile[b]) {
      tb.push_back(b);
    } else {
      g.o_src[oi] = (int8_t)b;
      g.o_dst[oi] = dst_pos[b];
      ++oi;
    }
  }
  for (int i = 0; i < t; ++i) g.e_src[i] = (int8_t)tb[i];
  std::vector<int> by_dst(t);
  for (int i = 0; i < t; ++i) by_dst[i] = i;
  std::sort(by_dst.begin(), by_dst.end(),
            [&](int x, int y) { return dst_pos[tb[x]] < dst_pos[tb[y]]; });
  for (int i = 0; i < t; ++i) {
    g.f_dst[i] = dst_pos[tb[by_dst[i]]];
    g.f_to_e[i] = (int8_t)by_dst[i];
  }
  const unsigned long long n_outer = 1ull << g.o_bits;
  const unsigned blocks = (unsigned)std::min<unsigned long long>(
      n_outer, 16ull * ctx->sm_count);
  permute_bits<<<blocks, kPermThreads, 0, ctx->stream>>>(g);
  QFX_CUDA(cudaGetLastError());
  ctx->stats.launches_permute++;
  ctx->stats.launches_total++;
  return QFX_OK;
}

// ---------------------------------------------------------------- general dimensions
// Tensor::_naive_reorder (reference src/tensor.cpp:377-492): axes of arbitrary
// (non-power-of-two, unit) dimensions.  The pipeline never produces such tensors --
// every index of a circuit network has dimension 2^k -- so this is the plain form:
// one thread per DESTINATION element (coalesced writes), the source offset rebuilt
// from the mixed-radix digits of the destination offset (gathered reads).
namespace {
struct GeneralReorderArgs {
  const float2* src;
  float2* dst;
  unsigned long long total;
  int rank;
  uint32_t new_dim[QFX_MAX_RANK];            // dimension of new axis j
  unsigned long long src_stride[QFX_MAX_RANK];  // source stride of new axis j
};

__global__ void reorder_general_kernel(const __grid_constant__ GeneralReorderArgs g) {
  const unsigned long long step = (unsigned long long)gridDim.x * blockDim.x;
  for (unsigned long long o = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
       o < g.total; o += step) {
    unsigned long long rest = o, s = 0;
    for (int j = g.rank - 1; j >= 0; --j) {
      const unsigned long long digit = rest % g.new_dim[j];
      rest /= g.new_dim[j];
      s += digit * g.src_stride[j];
    }
    g.dst[o] = g.src[s];
  }
}
}  // namespace

extern "C" int qfx_reorder_general(qfx_ctx* ctx, uint64_t src_off, uint64_t dst_off, int32_t rank,
                                   const uint32_t* dims, const int32_t* new_order) {
  if (ctx == nullptr || (rank > 0 && (dims == nullptr || new_order == nullptr))) {
    qfx_set_error("qfx_reorder_general: null argument");
    return QFX_ERR_ARG;
  }
  if (rank <= 0) {
    // reference: "Number of indices cannot be zero." (src/tensor.cpp:393)
    qfx_set_error("qfx_reorder_general: Number of indices cannot be zero.");
    return QFX_ERR_ARG;
  }
  if (rank > QFX_MAX_RANK) {
    qfx_set_error("qfx_reorder_general: rank above QFX_MAX_RANK");
    return QFX_ERR_ARG;
  }
  unsigned long long total = 1;
  uint32_t seen = 0;
  for (int i = 0; i < rank; ++i) {
    if (dims[i] == 0 || total > (1ull << 40) / dims[i]) {
      qfx_set_error("qfx_reorder_general: bad dimension");
      return QFX_ERR_ARG;
    }
    total *= dims[i];
    if (new_order[i] < 0 || new_order[i] >= rank || ((seen >> new_order[i]) & 1u)) {
      qfx_set_error("qfx_reorder_general: new_order is not a permutation of the axes");
      return QFX_ERR_ARG;
    }
    seen |= 1u << new_order[i];
  }
  if (src_off + total > ctx->arena_elems || dst_off + total > ctx->arena_elems) {
    qfx_set_error("qfx_reorder_general: operand outside the arena");
    return QFX_ERR_ARG;
  }
  if (src_off < dst_off + total && dst_off < src_off + total) {
    qfx_set_error("qfx_reorder_general: source and destination overlap");
    return QFX_ERR_ARG;
  }
  unsigned long long old_stride[QFX_MAX_RANK];
  old_stride[rank - 1] = 1;
  for (int i = rank - 1; i > 0; --i) old_stride[i - 1] = old_stride[i] * dims[i];
  GeneralReorderArgs g{};
  g.src = ctx->arena + src_off;
  g.dst = ctx->arena + dst_off;
  g.total = total;
  g.rank = rank;
  for (int j = 0; j < rank; ++j) {
    g.new_dim[j] = dims[new_order[j]];
    g.src_stride[j] = old_stride[new_order[j]];
  }
  const unsigned blocks = (unsigned)std::min<unsigned long long>((total + 255) / 256,
                                                                 16ull * ctx->sm_count);
  reorder_general_kernel<<<blocks, 256, 0, ctx->stream>>>(g);
  QFX_CUDA(cudaGetLastError());
  ctx->stats.launches_permute++;
  ctx->stats.launches_total++;
  ctx->stats.bytes += 16.0 * (double)total;
  return QFX_OK;
}
