// Pairwise tensor contraction kernels (FFMA path) for sm_100a.
//
// Replaces qflex::multiply = A.reorder + B.reorder + cblas_cgemm/cgemv/cdotu
// (reference src/tensor.cpp:1045-1194, 960-1042).  The two reorder passes are
// folded into the operand gathers: every operand is addressed through per-bit
// position tables (include/qflex_b200.h), so a contraction reads A and B once
// from wherever their index bits live and writes C once.
//
// Kernels
//   gemm_gather<...>  shared-memory tiled complex64 GEMM on FFMA, register
//                     tiles TMxTN per thread, double-buffered smem, 16-byte
//                     gathered global loads ordered by operand address so that a
//                     warp touches contiguous runs; optional split-K partials.
//   splitk_finish     deterministic sum of split-K partials -> C.
//   dotk / dotk_finish  contractions with <= 64 outputs (cgemv tails, cdotu):
//                     block-wide reduction over K in fp64, optional accumulation
//                     into a complex128 amplitude slot
//                     (reference src/contraction_utils.cpp:352).
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <vector>

#include "qfx_internal.h"

namespace {

constexpr int kMaxTileBits = 16;

struct GemmArgs {
  const float2* A;
  const float2* B;
  float2* C;
  float2* ws;  // split-K partial sums, or nullptr
  int m_bits, n_bits, k_bits;
  int tm, tn, tk;  // runtime tile bits (<= template maxima)
  int ks_bits;     // log2(number of K splits)
  int8_t a_m[QFX_MAX_BITS], c_m[QFX_MAX_BITS];
  int8_t b_n[QFX_MAX_BITS], c_n[QFX_MAX_BITS];
  int8_t a_k[QFX_MAX_BITS], b_k[QFX_MAX_BITS];
  // tile element enumeration, sorted by operand address bit
  int8_t ta_addr[kMaxTileBits], ta_smem[kMaxTileBits];
  int8_t tb_addr[kMaxTileBits], tb_smem[kMaxTileBits];
  int ta_n, tb_n;
  int a_vec, b_vec, c_vec;
};

__device__ __forceinline__ float4 ldg_f4(const float2* p) {
  return __ldg(reinterpret_cast<const float4*>(p));
}

template <int TMB, int TNB, int TKB, int TM, int TN, bool FULLK>
__global__ void __launch_bounds__(((1 << TMB) / TM) * ((1 << TNB) / TN))
    gemm_gather(const __grid_constant__ GemmArgs g) {
  constexpr int BM = 1 << TMB, BN = 1 << TNB, BKMAX = 1 << TKB;
  constexpr int TX = BN / TN, TY = BM / TM, NT = TX * TY;
  constexpr int PA = (BM * BKMAX / 2 + NT - 1) / NT;
  constexpr int PB = (BKMAX * BN / 2 + NT - 1) / NT;
  static_assert(TM % 2 == 0, "TM must be even");
  static_assert(TN == 1 || TN % 2 == 0, "TN must be 1 or even");

  extern __shared__ __align__(16) float2 smem[];
  float2* As = smem;                    // [2][BKMAX][BM]
  float2* Bs = smem + 2 * BKMAX * BM;   // [2][BKMAX][BN]

  const int tid = threadIdx.x;
  const int tx = tid % TX, ty = tid / TX;
  const int tk = FULLK ? TKB : g.tk;
  const int BK = 1 << tk;

  // ---- block decode: n tile fastest (CTAs sharing an A tile are adjacent)
  unsigned bid = blockIdx.x;
  const int ntn_bits = g.n_bits - g.tn, ntm_bits = g.m_bits - g.tm;
  const unsigned ntile = bid & ((1u << ntn_bits) - 1u);
  bid >>= ntn_bits;
  const unsigned mtile = bid & ((1u << ntm_bits) - 1u);
  bid >>= ntm_bits;
  const unsigned split = bid;

  unsigned long long baseA = 0, baseB = 0, baseC = 0;
  for (int i = 0; i < ntm_bits; ++i) {
    unsigned long long b = (mtile >> i) & 1u;
    baseA |= b << g.a_m[g.tm + i];
    baseC |= b << g.c_m[g.tm + i];
  }
  for (int i = 0; i < ntn_bits; ++i) {
    unsigned long long b = (ntile >> i) & 1u;
    baseB |= b << g.b_n[g.tn + i];
    baseC |= b << g.c_n[g.tn + i];
  }
  const float2* __restrict__ Ab = g.A + baseA;
  const float2* __restrict__ Bb = g.B + baseB;

  // ---- per-thread gather slots (pairs of tile elements)
  const int EA = 1 << (g.tm + tk), EB = 1 << (tk + g.tn);
  unsigned gA[PA], gB[PB];
  int sA[PA], sB[PB];
  const int ta_n = g.tm + tk, tb_n = tk + g.tn;
  // When the runtime tk is below TKB the host sorted tables for the runtime
  // tile, so ta_n/tb_n bits are exactly the populated entries.
#pragma unroll
  for (int j = 0; j < PA; ++j) {
    const unsigned e0 = 2u * (tid + j * NT);
    unsigned go = 0;
    int so = 0;
    for (int i = 1; i < ta_n; ++i) {
      const unsigned b = (e0 >> i) & 1u;
      go |= b << g.ta_addr[i];
      so |= b << g.ta_smem[i];
    }
    gA[j] = go;
    sA[j] = so;
  }
#pragma unroll
  for (int j = 0; j < PB; ++j) {
    const unsigned e0 = 2u * (tid + j * NT);
    unsigned go = 0;
    int so = 0;
    for (int i = 1; i < tb_n; ++i) {
      const unsigned b = (e0 >> i) & 1u;
      go |= b << g.tb_addr[i];
      so |= b << g.tb_smem[i];
    }
    gB[j] = go;
    sB[j] = so;
  }
  const unsigned dA = ta_n > 0 ? (1u << g.ta_addr[0]) : 0u;
  const unsigned dB = tb_n > 0 ? (1u << g.tb_addr[0]) : 0u;
  const int sdA = ta_n > 0 ? (1 << g.ta_smem[0]) : 0;
  const int sdB = tb_n > 0 ? (1 << g.tb_smem[0]) : 0;
  const bool a_vec = g.a_vec != 0, b_vec = g.b_vec != 0;

  float4 ra[PA], rb[PB];
  const int khi_bits = g.k_bits - tk;

  auto ldg_tiles = [&](unsigned kc) {
    unsigned long long kA = 0, kB = 0;
    for (int i = 0; i < khi_bits; ++i) {
      unsigned long long b = (kc >> i) & 1u;
      kA |= b << g.a_k[tk + i];
      kB |= b << g.b_k[tk + i];
    }
    const float2* __restrict__ Ak = Ab + kA;
    const float2* __restrict__ Bk = Bb + kB;
#pragma unroll
    for (int j = 0; j < PA; ++j) {
      const int e0 = 2 * (tid + j * NT);
      if (e0 < EA) {
        if (a_vec) {
          ra[j] = ldg_f4(Ak + gA[j]);
        } else {
          const float2 lo = __ldg(Ak + gA[j]);
          float2 hi = make_float2(0.f, 0.f);
          if (e0 + 1 < EA) hi = __ldg(Ak + gA[j] + dA);
          ra[j] = make_float4(lo.x, lo.y, hi.x, hi.y);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < PB; ++j) {
      const int e0 = 2 * (tid + j * NT);
      if (e0 < EB) {
        if (b_vec) {
          rb[j] = ldg_f4(Bk + gB[j]);
        } else {
          const float2 lo = __ldg(Bk + gB[j]);
          float2 hi = make_float2(0.f, 0.f);
          if (e0 + 1 < EB) hi = __ldg(Bk + gB[j] + dB);
          rb[j] = make_float4(lo.x, lo.y, hi.x, hi.y);
        }
      }
    }
  };
  auto sts_tiles = [&](int buf) {
    float2* Ad = As + buf * BKMAX * BM;
    float2* Bd = Bs + buf * BKMAX * BN;
#pragma unroll
    for (int j = 0; j < PA; ++j) {
      const int e0 = 2 * (tid + j * NT);
      if (e0 < EA) {
        Ad[sA[j]] = make_float2(ra[j].x, ra[j].y);
        if (e0 + 1 < EA) Ad[sA[j] + sdA] = make_float2(ra[j].z, ra[j].w);
      }
    }
#pragma unroll
    for (int j = 0; j < PB; ++j) {
      const int e0 = 2 * (tid + j * NT);
      if (e0 < EB) {
        Bd[sB[j]] = make_float2(rb[j].x, rb[j].y);
        if (e0 + 1 < EB) Bd[sB[j] + sdB] = make_float2(rb[j].z, rb[j].w);
      }
    }
  };

  float2 acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = make_float2(0.f, 0.f);

  const unsigned nch = 1u << (khi_bits - g.ks_bits);
  const unsigned kc0 = split * nch;

  ldg_tiles(kc0);
  sts_tiles(0);
  __syncthreads();

  for (unsigned it = 0; it < nch; ++it) {
    const int cur = it & 1;
    const bool more = (it + 1 < nch);
    if (more) ldg_tiles(kc0 + it + 1);

    const float2* Ac = As + cur * BKMAX * BM;
    const float2* Bc = Bs + cur * BKMAX * BN;
    auto kstep = [&](int kk) {
      float2 a[TM], b[TN];
      const float2* Ak = Ac + kk * BM;
      const float2* Bk = Bc + kk * BN;
#pragma unroll
      for (int i = 0; i < TM / 2; ++i) {
        const float4 v =
            *reinterpret_cast<const float4*>(Ak + (i * TY + ty) * 2);
        a[2 * i] = make_float2(v.x, v.y);
        a[2 * i + 1] = make_float2(v.z, v.w);
      }
      if constexpr (TN == 1) {
        b[0] = Bk[tx];
      } else {
#pragma unroll
        for (int j = 0; j < TN / 2; ++j) {
          const float4 v =
              *reinterpret_cast<const float4*>(Bk + (j * TX + tx) * 2);
          b[2 * j] = make_float2(v.x, v.y);
          b[2 * j + 1] = make_float2(v.z, v.w);
        }
      }
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) {
          acc[i][j].x = fmaf(a[i].x, b[j].x, acc[i][j].x);
          acc[i][j].x = fmaf(-a[i].y, b[j].y, acc[i][j].x);
          acc[i][j].y = fmaf(a[i].x, b[j].y, acc[i][j].y);
          acc[i][j].y = fmaf(a[i].y, b[j].x, acc[i][j].y);
        }
    };
    if constexpr (FULLK) {
#pragma unroll
      for (int kk = 0; kk < BKMAX; ++kk) kstep(kk);
    } else {
      for (int kk = 0; kk < BK; ++kk) kstep(kk);
    }

    if (more) sts_tiles(cur ^ 1);
    __syncthreads();
  }

  // ---- epilogue
  const int mvalid = 1 << g.tm, nvalid = 1 << g.tn;
  if (g.ws == nullptr) {
    float2* __restrict__ Cb = g.C + baseC;
    unsigned colOff[TN];
#pragma unroll
    for (int j = 0; j < TN; ++j) {
      const int nl = (TN == 1) ? tx : (((j >> 1) * TX + tx) * 2 + (j & 1));
      unsigned o = 0;
      for (int q = 0; q < g.tn; ++q) o |= ((nl >> q) & 1u) << g.c_n[q];
      colOff[j] = o;
    }
#pragma unroll
    for (int i = 0; i < TM; ++i) {
      const int ml = ((i >> 1) * TY + ty) * 2 + (i & 1);
      if (ml >= mvalid) continue;
      unsigned ro = 0;
      for (int q = 0; q < g.tm; ++q) ro |= ((ml >> q) & 1u) << g.c_m[q];
      if constexpr (TN == 1) {
        if (tx < nvalid) Cb[ro + colOff[0]] = acc[i][0];
      } else {
#pragma unroll
        for (int j = 0; j < TN; j += 2) {
          const int nl = ((j >> 1) * TX + tx) * 2;
          if (nl >= nvalid) continue;
          if (g.c_vec) {
            *reinterpret_cast<float4*>(Cb + ro + colOff[j]) = make_float4(
                acc[i][j].x, acc[i][j].y, acc[i][j + 1].x, acc[i][j + 1].y);
          } else {
            Cb[ro + colOff[j]] = acc[i][j];
            if (nl + 1 < nvalid) Cb[ro + colOff[j + 1]] = acc[i][j + 1];
          }
        }
      }
    }
  } else {
    // split-K partials: ws[split][m][n] in natural (m,n) numbering
    float2* __restrict__ W =
        g.ws + ((unsigned long long)split << (g.m_bits + g.n_bits));
#pragma unroll
    for (int i = 0; i < TM; ++i) {
      const int ml = ((i >> 1) * TY + ty) * 2 + (i & 1);
      if (ml >= mvalid) continue;
      const unsigned long long mg = ((unsigned long long)mtile << g.tm) | ml;
#pragma unroll
      for (int j = 0; j < TN; ++j) {
        const int nl = (TN == 1) ? tx : (((j >> 1) * TX + tx) * 2 + (j & 1));
        if (nl >= nvalid) continue;
        const unsigned long long ng = ((unsigned long long)ntile << g.tn) | nl;
        W[(mg << g.n_bits) | ng] = acc[i][j];
      }
    }
  }
}

struct FinishArgs {
  const float2* ws;
  float2* C;
  int m_bits, n_bits, splits;
  int8_t c_m[QFX_MAX_BITS], c_n[QFX_MAX_BITS];
};

__global__ void splitk_finish(const __grid_constant__ FinishArgs g) {
  const unsigned long long total = 1ull << (g.m_bits + g.n_bits);
  for (unsigned long long idx =
           blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
       idx < total; idx += (unsigned long long)gridDim.x * blockDim.x) {
    float2 s = make_float2(0.f, 0.f);
    for (int sp = 0; sp < g.splits; ++sp) {
      const float2 v = g.ws[((unsigned long long)sp << (g.m_bits + g.n_bits)) + idx];
      s.x += v.x;
      s.y += v.y;
    }
    const unsigned long long n = idx & ((1ull << g.n_bits) - 1ull);
    const unsigned long long m = idx >> g.n_bits;
    unsigned long long o = 0;
    for (int q = 0; q < g.n_bits; ++q) o |= ((n >> q) & 1ull) << g.c_n[q];
    for (int q = 0; q < g.m_bits; ++q) o |= ((m >> q) & 1ull) << g.c_m[q];
    g.C[o] = s;
  }
}

// ---------------------------------------------------------------- small outputs
struct DotArgs {
  const float2* A;
  const float2* B;
  float2* C;      // used when amp == nullptr
  double2* amp;   // amplitude slot (accumulated) or nullptr
  double2* ws;    // [outputs][splits] partials (splits > 1)
  int m_bits, n_bits, k_bits, ks_bits;
  int vec;  // k bit 0 is address bit 0 of both operands and both bases are even
  int8_t a_m[QFX_MAX_BITS], c_m[QFX_MAX_BITS];
  int8_t b_n[QFX_MAX_BITS], c_n[QFX_MAX_BITS];
  int8_t a_k[QFX_MAX_BITS], b_k[QFX_MAX_BITS];
};

constexpr int kDotThreads = 256;
constexpr int kDotLoBits = 8;

__device__ __forceinline__ unsigned long long dot_out_offset(const DotArgs& g,
                                                             unsigned out) {
  const unsigned n = out & ((1u << g.n_bits) - 1u), m = out >> g.n_bits;
  unsigned long long o = 0;
  for (int q = 0; q < g.n_bits; ++q) o |= (unsigned long long)((n >> q) & 1u) << g.c_n[q];
  for (int q = 0; q < g.m_bits; ++q) o |= (unsigned long long)((m >> q) & 1u) << g.c_m[q];
  return o;
}

__global__ void __launch_bounds__(kDotThreads)
    dotk(const __grid_constant__ DotArgs g) {
  // K is enumerated lowest-A-address-bit first (host sorts the k tables): the
  // optional vector bit (k bit 0 = address bit 0 of both operands: 16-byte loads),
  // then up to 8 bits across the threads, then the remaining bits walked by every
  // thread in Gray-code order, so that one step flips ONE address bit per operand
  // (the summation order is free; accumulation is fp64).
  const unsigned out = blockIdx.x, split = blockIdx.y;
  const unsigned n = out & ((1u << g.n_bits) - 1u), m = out >> g.n_bits;
  unsigned long long baseA = 0, baseB = 0;
  for (int q = 0; q < g.m_bits; ++q) baseA |= (unsigned long long)((m >> q) & 1u) << g.a_m[q];
  for (int q = 0; q < g.n_bits; ++q) baseB |= (unsigned long long)((n >> q) & 1u) << g.b_n[q];

  const int v = g.vec;
  const int lo_bits = (g.k_bits - v) < kDotLoBits ? (g.k_bits - v) : kDotLoBits;
  const unsigned tid = threadIdx.x;
  unsigned long long offA = baseA, offB = baseB;
  for (int q = 0; q < lo_bits; ++q) {
    unsigned long long b = (tid >> q) & 1u;
    offA |= b << g.a_k[v + q];
    offB |= b << g.b_k[v + q];
  }
  const bool active = tid < (1u << lo_bits);
  const int h0 = v + lo_bits;                 // first walked k bit
  const int hi_bits = g.k_bits - h0;
  const int walk_bits = hi_bits - g.ks_bits;  // per block; the top ks bits select the split
  for (int q = 0; q < g.ks_bits; ++q) {
    unsigned long long b = (split >> q) & 1u;
    offA |= b << g.a_k[h0 + walk_bits + q];
    offB |= b << g.b_k[h0 + walk_bits + q];
  }
  const unsigned nch = 1u << walk_bits;
  const float2* __restrict__ A = g.A;
  const float2* __restrict__ B = g.B;

  double sx = 0.0, sy = 0.0;
  if (active) {
    for (unsigned c = 0; c < nch; c += 4) {
      unsigned long long oa[4], ob[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        oa[u] = offA;
        ob[u] = offB;
        const unsigned nxt = c + u + 1;
        if (nxt < nch) {
          const int bit = __ffs(nxt) - 1;  // Gray code: step i flips bit ctz(i)
          offA ^= 1ull << g.a_k[h0 + bit];
          offB ^= 1ull << g.b_k[h0 + bit];
        }
      }
      if (v) {
        float4 a[4], b[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (c + u < nch) {
            a[u] = __ldg(reinterpret_cast<const float4*>(A + oa[u]));
            b[u] = __ldg(reinterpret_cast<const float4*>(B + ob[u]));
          }
        // fp32 products and a short fp32 sum (<= 8 terms, what cblas_cdotu does for
        // the whole K), then fp64: fp32->fp64 conversions run at a few per clock per
        // SM and bounded this kernel at 0.8 TB/s when every product was converted
        float px = 0.f, py = 0.f;
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (c + u < nch) {
            px = fmaf(a[u].x, b[u].x, px);
            px = fmaf(-a[u].y, b[u].y, px);
            py = fmaf(a[u].x, b[u].y, py);
            py = fmaf(a[u].y, b[u].x, py);
            px = fmaf(a[u].z, b[u].z, px);
            px = fmaf(-a[u].w, b[u].w, px);
            py = fmaf(a[u].z, b[u].w, py);
            py = fmaf(a[u].w, b[u].z, py);
          }
        sx += (double)px;
        sy += (double)py;
      } else {
        float2 a[4], b[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (c + u < nch) {
            a[u] = __ldg(A + oa[u]);
            b[u] = __ldg(B + ob[u]);
          }
        float px = 0.f, py = 0.f;
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (c + u < nch) {
            px = fmaf(a[u].x, b[u].x, px);
            px = fmaf(-a[u].y, b[u].y, px);
            py = fmaf(a[u].x, b[u].y, py);
            py = fmaf(a[u].y, b[u].x, py);
          }
        sx += (double)px;
        sy += (double)py;
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    sx += __shfl_xor_sync(0xffffffffu, sx, o);
    sy += __shfl_xor_sync(0xffffffffu, sy, o);
  }
  __shared__ double red[2][kDotThreads / 32];
  if ((tid & 31) == 0) {
    red[0][tid >> 5] = sx;
    red[1][tid >> 5] = sy;
  }
  __syncthreads();
  if (tid == 0) {
    double tx_ = 0.0, ty_ = 0.0;
    for (int w = 0; w < kDotThreads / 32; ++w) {
      tx_ += red[0][w];
      ty_ += red[1][w];
    }
    if (g.ks_bits > 0) {
      g.ws[((unsigned long long)out << g.ks_bits) + split] = make_double2(tx_, ty_);
    } else if (g.amp != nullptr) {
      g.amp->x += tx_;
      g.amp->y += ty_;
    } else {
      g.C[dot_out_offset(g, out)] = make_float2((float)tx_, (float)ty_);
    }
  }
}

// ---------------------------------------------------------------- stream_small
// C[m, n] = sum_k A[m, k] * B[k, n] with N, K <= 8 and a long M: one world-line step
// of a large grid tensor (cut-aware flatten of Sycamore m >= 14: 2^28..2^30 elements
// times a 1- or 2-qubit gate), or the expansion of a patch by a grid tensor with
// bond dimension <= 8 (bristlecone 1-24-1: 2^24 x 8 x 8, AI 4 flop/B).  Pure
// streaming: the tiled GEMM spent
// its time on per-tile bookkeeping (0.95 TB/s).  One thread per m, m enumerated in
// A-address order; the m bits are grouped into runs that are contiguous in both A and
// C so that an address is a few shift-mask-or steps; B lives in shared memory.
struct StreamSeg {
  int start, len, a_shift, c_shift;
};
struct StreamArgs {
  const float2* A;
  const float2* B;
  float2* C;
  int m_bits, n_bits, k_bits, n_seg;
  int c_vec;  // the n bits are C's lowest address bits and C's base is even: 16-byte stores
  int a_vec;  // the k bits are A's lowest address bits and A's base is even: 16-byte loads
  StreamSeg seg[8];
  int8_t a_k[3], b_k[3], b_n[3], c_n[3];
};

template <int KB, int NB>
__global__ void __launch_bounds__(256) stream_small(const __grid_constant__ StreamArgs g) {
  constexpr int K = 1 << KB, N = 1 << NB;
  __shared__ float2 Bs[K * N];
  __shared__ unsigned long long ka[K], cn[N];
  if ((int)threadIdx.x < K * N) {
    const int k = threadIdx.x / N, n = threadIdx.x % N;
    unsigned long long o = 0;
    for (int q = 0; q < KB; ++q) o |= (unsigned long long)((k >> q) & 1) << g.b_k[q];
    for (int q = 0; q < NB; ++q) o |= (unsigned long long)((n >> q) & 1) << g.b_n[q];
    Bs[threadIdx.x] = g.B[o];
  }
  if (threadIdx.x < 8) {
    unsigned long long oa = 0, oc = 0;
    for (int q = 0; q < KB; ++q) oa |= (unsigned long long)((threadIdx.x >> q) & 1) << g.a_k[q];
    for (int q = 0; q < NB; ++q) oc |= (unsigned long long)((threadIdx.x >> q) & 1) << g.c_n[q];
    if ((int)threadIdx.x < K) ka[threadIdx.x] = oa;
    if ((int)threadIdx.x < N) cn[threadIdx.x] = oc;
  }
  __syncthreads();
  const unsigned long long M = 1ull << g.m_bits;
  const float2* __restrict__ A = g.A;
  float2* __restrict__ C = g.C;
  for (unsigned long long m = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; m < M;
       m += (unsigned long long)gridDim.x * blockDim.x) {
    unsigned long long ao = 0, co = 0;
    for (int s = 0; s < g.n_seg; ++s) {
      const unsigned long long f = (m >> g.seg[s].start) & ((1ull << g.seg[s].len) - 1ull);
      ao |= f << g.seg[s].a_shift;
      co |= f << g.seg[s].c_shift;
    }
    float2 a[K];
    if (g.a_vec && K >= 2) {
      // the k bits are A's lowest address bits: one row = K contiguous elements
#pragma unroll
      for (int k = 0; k < K; k += 2) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(A + ao + k));
        a[k] = make_float2(v.x, v.y);
        a[k + 1] = make_float2(v.z, v.w);
      }
    } else {
#pragma unroll
      for (int k = 0; k < K; ++k) a[k] = __ldg(A + (ao | ka[k]));
    }
    float2 prev = make_float2(0.f, 0.f);
#pragma unroll
    for (int n = 0; n < N; ++n) {
      float sx = 0.f, sy = 0.f;
#pragma unroll
      for (int k = 0; k < K; ++k) {
        const float2 b = Bs[k * N + n];
        sx = fmaf(a[k].x, b.x, sx);
        sx = fmaf(-a[k].y, b.y, sx);
        sy = fmaf(a[k].x, b.y, sy);
        sy = fmaf(a[k].y, b.x, sy);
      }
      if (!g.c_vec) {
        C[co | cn[n]] = make_float2(sx, sy);
      } else if (n & 1) {
        *reinterpret_cast<float4*>(C + co + n - 1) = make_float4(prev.x, prev.y, sx, sy);
      } else {
        prev = make_float2(sx, sy);
      }
    }
  }
}

template <int KB>
void launch_stream_small_n(const StreamArgs& g, unsigned blocks, cudaStream_t stream) {
  switch (g.n_bits) {
    case 0: stream_small<KB, 0><<<blocks, 256, 0, stream>>>(g); break;
    case 1: stream_small<KB, 1><<<blocks, 256, 0, stream>>>(g); break;
    case 2: stream_small<KB, 2><<<blocks, 256, 0, stream>>>(g); break;
    default: stream_small<KB, 3><<<blocks, 256, 0, stream>>>(g); break;
  }
}

void launch_stream_small(const StreamArgs& g, unsigned blocks, cudaStream_t stream) {
  switch (g.k_bits) {
    case 0: launch_stream_small_n<0>(g, blocks, stream); break;
    case 1: launch_stream_small_n<1>(g, blocks, stream); break;
    case 2: launch_stream_small_n<2>(g, blocks, stream); break;
    default: launch_stream_small_n<3>(g, blocks, stream); break;
  }
}

// ---------------------------------------------------------------- outer_small
// C[m, n] = sum_k A[m, k] * B[k, n] with K <= 4 and BOTH free sides long, both operands
// small next to the result (the root of a tree-ordered world line: 2^12 x 2^12 x 2,
// two 64-KB halves, a 128-MB result written in the grid tensor's final layout).  Pure
// write streaming, so the work is enumerated in C-ADDRESS order: thread t owns the two
// elements 2t, 2t+1 (one 16-byte store); a run of C address bits that is a run of m
// bits in A (or of n bits in B) turns into one shift-mask-or; A and B come from L1/L2
// (neighbouring threads read neighbouring or identical addresses).  The tiled GEMM
// reached 1.9 TB/s on these (scattered C tiles, K = 2 per tile pass).
struct OuterSeg {
  int start, len, a_shift, b_shift;  // C address bits [start, start+len): bits of A (b_shift < 0) or of B
};
struct OuterArgs {
  const float2* A;
  const float2* B;
  float2* C;
  int c_bits, k_bits, n_seg;
  int low_is_a;             // C address bit 0 is an m bit (else an n bit)
  unsigned low_delta;       // its offset in A (or B)
  OuterSeg seg[16];         // C address bits >= 1
  int8_t a_k[2], b_k[2];
};

// A block works on tiles of 2^12 consecutive C elements (2048 pairs): the operand
// offsets of C address bits 1..11 come from two 6-bit / 5-bit tables in shared memory
// (built once per block), those of the bits above from the run list, once per tile.
template <int KB>
__global__ void __launch_bounds__(256) outer_small(const __grid_constant__ OuterArgs g) {
  constexpr int K = 1 << KB;
  __shared__ unsigned ka[K], kb[K];
  __shared__ unsigned lo_a[64], lo_b[64], mid_a[32], mid_b[32];  // C bits 1..6, 7..11
  __shared__ unsigned tile_a, tile_b;
  const unsigned tid = threadIdx.x;
  if ((int)tid < K) {
    unsigned oa = 0, ob = 0;
    for (int q = 0; q < KB; ++q) {
      oa |= ((tid >> q) & 1u) << g.a_k[q];
      ob |= ((tid >> q) & 1u) << g.b_k[q];
    }
    ka[tid] = oa;
    kb[tid] = ob;
  }
  if (tid >= 64 && tid < 64 + 96) {
    // offsets of a group of low C address bits, from the run list
    const unsigned e = tid - 64, idx = e < 64 ? e : e - 64;
    const unsigned long long c = e < 64 ? (unsigned long long)idx << 1 : (unsigned long long)idx << 7;
    const unsigned long long mask = e < 64 ? 0x7eull : 0xf80ull;
    unsigned oa = 0, ob = 0;
    for (int s = 0; s < g.n_seg; ++s)
      for (int q = 0; q < g.seg[s].len; ++q) {
        const int p = g.seg[s].start + q;
        if (!((mask >> p) & 1ull) || !((c >> p) & 1ull)) continue;
        if (g.seg[s].b_shift < 0) oa |= 1u << (g.seg[s].a_shift + q);
        else ob |= 1u << (g.seg[s].b_shift + q);
      }
    if (e < 64) lo_a[idx] = oa, lo_b[idx] = ob;
    else mid_a[idx] = oa, mid_b[idx] = ob;
  }
  const unsigned long long tiles = g.c_bits > 12 ? 1ull << (g.c_bits - 12) : 1ull;
  const unsigned pairs_in_tile = g.c_bits > 12 ? 2048u : 1u << (g.c_bits - 1);
  const float2* __restrict__ A = g.A;
  const float2* __restrict__ B = g.B;
  for (unsigned long long tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    __syncthreads();
    if (tid == 0) {
      const unsigned long long c = tile << 12;
      unsigned oa = 0, ob = 0;
      for (int s = 0; s < g.n_seg; ++s) {
        // the part of the run at C address bits >= 12
        int start = g.seg[s].start, len = g.seg[s].len, skip = 0;
        if (start + len <= 12) continue;
        if (start < 12) skip = 12 - start;
        const unsigned f = (unsigned)(c >> (start + skip)) & ((1u << (len - skip)) - 1u);
        if (g.seg[s].b_shift < 0) oa |= f << (g.seg[s].a_shift + skip);
        else ob |= f << (g.seg[s].b_shift + skip);
      }
      tile_a = oa;
      tile_b = ob;
    }
    __syncthreads();
    const unsigned ta = tile_a, tb = tile_b;
#pragma unroll 2
    for (unsigned p = tid; p < pairs_in_tile; p += 256) {
      const unsigned ao = ta | lo_a[p & 63u] | mid_a[p >> 6], bo = tb | lo_b[p & 63u] | mid_b[p >> 6];
      const unsigned ao1 = g.low_is_a ? ao | g.low_delta : ao;
      const unsigned bo1 = g.low_is_a ? bo : bo | g.low_delta;
      float x0 = 0.f, y0 = 0.f, x1 = 0.f, y1 = 0.f;
#pragma unroll
      for (int k = 0; k < K; ++k) {
        const float2 a0 = __ldg(A + (ao | ka[k])), b0 = __ldg(B + (bo | kb[k]));
        const float2 a1 = g.low_is_a ? __ldg(A + (ao1 | ka[k])) : a0;
        const float2 b1 = g.low_is_a ? b0 : __ldg(B + (bo1 | kb[k]));
        x0 = fmaf(a0.x, b0.x, x0);
        x0 = fmaf(-a0.y, b0.y, x0);
        y0 = fmaf(a0.x, b0.y, y0);
        y0 = fmaf(a0.y, b0.x, y0);
        x1 = fmaf(a1.x, b1.x, x1);
        x1 = fmaf(-a1.y, b1.y, x1);
        y1 = fmaf(a1.x, b1.y, y1);
        y1 = fmaf(a1.y, b1.x, y1);
      }
      *reinterpret_cast<float4*>(g.C + (tile << 12) + 2ull * p) = make_float4(x0, y0, x1, y1);
    }
  }
}

// ---------------------------------------------------------------- gemv_rows
// out[m] = sum_k A[m,k] * B[k] with 2..64 outputs and a long K, A dense: the tail of a
// schedule whose output indices stayed open (the last merge yields all 2^k amplitudes).
// dotk runs one block per output and walks A with the output bits fixed -- when those
// bits sit low in A's address every block touches 8 bytes out of each 32-byte sector.
// Here A is read ONCE, linearly, with 16-byte loads: address bit 0 = the two halves of a
// load, bits 1..8 = the thread, higher bits = the block (bits that are output bits, and
// the K split) or walked in Gray-code order; a (thread, half) pair therefore always
// feeds the same output and needs one accumulator.  B (2^k elements) comes from L1/L2.
struct GemvArgs {
  const float2* A;
  const float2* B;
  double2* ws;                   // [outputs][splits]
  int m_bits, ks_bits;
  int n_hm, n_hk;                // output / k bits at A address positions >= kGemvLow
  int8_t kind[16];               // A address bit p < kGemvLow: k index, or -(m index) - 1
  int8_t hm_pos[8], hm_idx[8];   // high output bits: A address position, m index
  int8_t hk_pos[32], hk_bk[32];  // high k bits, ascending A position: A / B position
  int8_t b_k[QFX_MAX_BITS];
};
constexpr int kGemvLow = 9;

__global__ void __launch_bounds__(256) gemv_rows(const __grid_constant__ GemvArgs g) {
  __shared__ double2 part[512];
  __shared__ uint8_t part_m[512];
  const unsigned tid = threadIdx.x;
  unsigned mlow[2] = {0u, 0u};
  unsigned long long offB[2] = {0ull, 0ull};
#pragma unroll
  for (int e = 0; e < 2; ++e) {
    const unsigned j = (tid << 1) | e;
    for (int p = 0; p < kGemvLow; ++p)
      if ((j >> p) & 1u) {
        const int kd = g.kind[p];
        if (kd >= 0) offB[e] |= 1ull << g.b_k[kd];
        else mlow[e] |= 1u << (-kd - 1);
      }
  }
  unsigned long long offA = (unsigned long long)tid << 1, offBc = 0;
  unsigned mhigh = 0;
  for (int q = 0; q < g.n_hm; ++q)
    if ((blockIdx.x >> q) & 1u) {
      offA |= 1ull << g.hm_pos[q];
      mhigh |= 1u << g.hm_idx[q];
    }
  const int walk_bits = g.n_hk - g.ks_bits;
  for (int q = 0; q < g.ks_bits; ++q)
    if ((blockIdx.y >> q) & 1u) {
      offA |= 1ull << g.hk_pos[walk_bits + q];
      offBc |= 1ull << g.hk_bk[walk_bits + q];
    }
  const unsigned nch = 1u << walk_bits;
  const float2* __restrict__ A = g.A;
  const float2* __restrict__ B = g.B;
  double s0x = 0.0, s0y = 0.0, s1x = 0.0, s1y = 0.0;
  for (unsigned c = 0; c < nch; c += 4) {
    unsigned long long oa[4], ob[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      oa[u] = offA;
      ob[u] = offBc;
      const unsigned nxt = c + u + 1;
      if (nxt < nch) {
        const int bit = __ffs(nxt) - 1;  // Gray code: step i flips bit ctz(i)
        offA ^= 1ull << g.hk_pos[bit];
        offBc ^= 1ull << g.hk_bk[bit];
      }
    }
    float4 a[4];
    float2 b0[4], b1[4];
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (c + u < nch) {
        a[u] = __ldg(reinterpret_cast<const float4*>(A + oa[u]));
        b0[u] = __ldg(B + (ob[u] | offB[0]));
        b1[u] = __ldg(B + (ob[u] | offB[1]));
      }
    float p0x = 0.f, p0y = 0.f, p1x = 0.f, p1y = 0.f;  // short fp32 sums, then fp64
#pragma unroll
    for (int u = 0; u < 4; ++u)
      if (c + u < nch) {
        p0x = fmaf(a[u].x, b0[u].x, p0x);
        p0x = fmaf(-a[u].y, b0[u].y, p0x);
        p0y = fmaf(a[u].x, b0[u].y, p0y);
        p0y = fmaf(a[u].y, b0[u].x, p0y);
        p1x = fmaf(a[u].z, b1[u].x, p1x);
        p1x = fmaf(-a[u].w, b1[u].y, p1x);
        p1y = fmaf(a[u].z, b1[u].y, p1y);
        p1y = fmaf(a[u].w, b1[u].x, p1y);
      }
    s0x += (double)p0x;
    s0y += (double)p0y;
    s1x += (double)p1x;
    s1y += (double)p1y;
  }
  part[(tid << 1)] = make_double2(s0x, s0y);
  part[(tid << 1) | 1] = make_double2(s1x, s1y);
  part_m[(tid << 1)] = (uint8_t)mlow[0];
  part_m[(tid << 1) | 1] = (uint8_t)mlow[1];
  __syncthreads();
  // outputs of this block: every value of the low output bits; one warp per output,
  // lanes stride over the 512 partials in a fixed order (deterministic)
  unsigned low_mask = 0;
  for (int p = 0; p < kGemvLow; ++p)
    if (g.kind[p] < 0) low_mask |= 1u << (-g.kind[p] - 1);
  const unsigned warp = tid >> 5, lane = tid & 31;
  for (unsigned m = warp; m < (1u << g.m_bits); m += 8) {
    if ((m & ~low_mask) != 0) continue;  // enumerate the subsets of low_mask only
    double sx = 0.0, sy = 0.0;
    for (unsigned j = lane; j < 512; j += 32)
      if (part_m[j] == m) {
        sx += part[j].x;
        sy += part[j].y;
      }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      sx += __shfl_xor_sync(0xffffffffu, sx, o);
      sy += __shfl_xor_sync(0xffffffffu, sy, o);
    }
    if (lane == 0)
      g.ws[((unsigned long long)(m | mhigh) << g.ks_bits) + blockIdx.y] = make_double2(sx, sy);
  }
}

// One warp per output: lanes stride over the split partials, then a shuffle tree
// (fixed order: deterministic).  A single thread walking 1024 partials took 140 us,
// three times the dot product itself.
__global__ void dotk_finish(const __grid_constant__ DotArgs g) {
  const unsigned outs = 1u << (g.m_bits + g.n_bits);
  const unsigned out = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (out >= outs) return;
  double sx = 0.0, sy = 0.0;
  const unsigned splits = 1u << g.ks_bits;
  for (unsigned s = lane; s < splits; s += 32) {
    const double2 v = g.ws[((unsigned long long)out << g.ks_bits) + s];
    sx += v.x;
    sy += v.y;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    sx += __shfl_xor_sync(0xffffffffu, sx, o);
    sy += __shfl_xor_sync(0xffffffffu, sy, o);
  }
  if (lane != 0) return;
  if (g.amp != nullptr) {
    g.amp->x += sx;
    g.amp->y += sy;
  } else {
    g.C[dot_out_offset(g, out)] = make_float2((float)sx, (float)sy);
  }
}

__global__ void scalar_accumulate_kernel(const float2* src, double2* amp) {
  amp->x += (double)src->x;
  amp->y += (double)src->y;
}

// ---------------------------------------------------------------- host side
template <int TMB, int TNB, int TKB, int TM, int TN>
int launch_gemm_cfg(qfx_ctx* ctx, GemmArgs& g, unsigned blocks) {
  constexpr int BM = 1 << TMB, BN = 1 << TNB, BKMAX = 1 << TKB;
  constexpr int NT = (BM / TM) * (BN / TN);
  constexpr size_t smem = 2ull * BKMAX * (BM + BN) * sizeof(float2);
  static qfx_per_device_once attr_full, attr_part;
  if (g.tk == TKB) {
    auto kern = gemm_gather<TMB, TNB, TKB, TM, TN, true>;
    if (attr_full.need(ctx->device)) {
      QFX_CUDA(cudaFuncSetAttribute(
          kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      attr_full.mark(ctx->device);
    }
    kern<<<blocks, NT, smem, ctx->stream>>>(g);
  } else {
    auto kern = gemm_gather<TMB, TNB, TKB, TM, TN, false>;
    if (attr_part.need(ctx->device)) {
      QFX_CUDA(cudaFuncSetAttribute(
          kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      attr_part.mark(ctx->device);
    }
    kern<<<blocks, NT, smem, ctx->stream>>>(g);
  }
  QFX_CUDA(cudaGetLastError());
  return QFX_OK;
}

struct TileBit {
  int addr, smem;
};

void build_tile_table(const int8_t* row_addr, int trow, const int8_t* k_addr,
                      int tk, int k_smem_shift, int8_t* out_addr,
                      int8_t* out_smem, int* vec) {
  std::vector<TileBit> bits;
  for (int q = 0; q < trow; ++q) bits.push_back({row_addr[q], q});
  for (int q = 0; q < tk; ++q) bits.push_back({k_addr[q], k_smem_shift + q});
  std::sort(bits.begin(), bits.end(),
            [](const TileBit& x, const TileBit& y) { return x.addr < y.addr; });
  for (size_t i = 0; i < bits.size(); ++i) {
    out_addr[i] = (int8_t)bits[i].addr;
    out_smem[i] = (int8_t)bits[i].smem;
  }
  *vec = (!bits.empty() && bits[0].addr == 0) ? 1 : 0;
}

// Positions of one operand must be distinct and < 32.  Returns the OR of all
// position bits (the operand's extent minus one) or ~0 on error.
uint64_t operand_mask(const int8_t* p1, int n1, const int8_t* p2, int n2) {
  uint64_t seen = 0;
  for (int i = 0; i < n1; ++i) {
    if (p1[i] < 0 || p1[i] >= 32 || (seen >> p1[i]) & 1) return ~0ull;
    seen |= 1ull << p1[i];
  }
  for (int i = 0; i < n2; ++i) {
    if (p2[i] < 0 || p2[i] >= 32 || (seen >> p2[i]) & 1) return ~0ull;
    seen |= 1ull << p2[i];
  }
  return seen;
}

// the streamed operand of the GEMV path: its m and k bits fill 0 .. m+k-1
bool mask_a_dense(const qfx_contract_desc& d) {
  uint64_t seen = 0;
  for (int i = 0; i < d.m_bits; ++i) seen |= 1ull << d.a_m[i];
  for (int i = 0; i < d.k_bits; ++i) seen |= 1ull << d.a_k[i];
  const int bits = d.m_bits + d.k_bits;
  return bits <= 32 && seen == (bits == 64 ? ~0ull : (1ull << bits) - 1ull);
}

}  // namespace

int qfx_launch_contract(qfx_ctx* ctx, const qfx_contract_desc* din,
                        bool accumulate_amp, uint64_t slot) {
  if (ctx == nullptr || din == nullptr) {
    qfx_set_error("qfx_contract: null argument");
    return QFX_ERR_ARG;
  }
  qfx_contract_desc d = *din;
  if (d.m_bits < 0 || d.n_bits < 0 || d.k_bits < 0 ||
      d.m_bits + d.k_bits > 32 || d.n_bits + d.k_bits > 32 ||
      d.m_bits + d.n_bits > 32) {
    qfx_set_error("qfx_contract: operand rank out of range (max 32 bits each)");
    return QFX_ERR_ARG;
  }
  const uint64_t mask_a = operand_mask(d.a_m, d.m_bits, d.a_k, d.k_bits);
  const uint64_t mask_b = operand_mask(d.b_k, d.k_bits, d.b_n, d.n_bits);
  const uint64_t mask_c = operand_mask(d.c_m, d.m_bits, d.c_n, d.n_bits);
  if (mask_a == ~0ull || mask_b == ~0ull || mask_c == ~0ull) {
    qfx_set_error(
        "qfx_contract: bit positions of an operand must be distinct and < 32");
    return QFX_ERR_ARG;
  }
  // spans actually touched (positions may have holes: projected/cut bits are
  // folded into the base offset)
  const uint64_t sa = mask_a + 1, sb = mask_b + 1, sc = mask_c + 1;
  if (d.off_a + sa > ctx->arena_elems || d.off_b + sb > ctx->arena_elems ||
      (!accumulate_amp && d.off_c + sc > ctx->arena_elems)) {
    qfx_set_error("qfx_contract: operand outside the arena");
    return QFX_ERR_ARG;
  }
  if (!accumulate_amp) {
    auto overlap = [](uint64_t o1, uint64_t s1, uint64_t o2, uint64_t s2) {
      return o1 < o2 + s2 && o2 < o1 + s1;
    };
    if (overlap(d.off_a, sa, d.off_c, sc) || overlap(d.off_b, sb, d.off_c, sc)) {
      // reference: "A and C cannot be the same tensor" (src/tensor.cpp:1050-1057)
      qfx_set_error("qfx_contract: C overlaps an input operand");
      return QFX_ERR_ARG;
    }
  } else {
    if (d.m_bits != 0 || d.n_bits != 0) {
      // reference: "Contraction did not complete" (src/contraction_utils.cpp:348)
      qfx_set_error("qfx_dot_accumulate: result is not a scalar");
      return QFX_ERR_ARG;
    }
    if (slot >= ctx->n_amps) {
      qfx_set_error("qfx_dot_accumulate: amplitude slot out of range");
      return QFX_ERR_ARG;
    }
  }

  // Larger free side plays "M" (rows); the contraction is symmetric in A/B once
  // C is addressed through tables.
  if (d.n_bits > d.m_bits) {
    std::swap(d.off_a, d.off_b);
    std::swap(d.m_bits, d.n_bits);
    for (int i = 0; i < QFX_MAX_BITS; ++i) {
      std::swap(d.a_m[i], d.b_n[i]);
      std::swap(d.c_m[i], d.c_n[i]);
      std::swap(d.a_k[i], d.b_k[i]);
    }
  }
  // K enumeration order is free (it only permutes the summation): put the K
  // bits that sit lowest in either operand first so that a K chunk covers the
  // operands' contiguous low address bits.
  {
    std::vector<int> idx(d.k_bits);
    for (int i = 0; i < d.k_bits; ++i) idx[i] = i;
    std::stable_sort(idx.begin(), idx.end(), [&](int x, int y) {
      return std::min(d.a_k[x], d.b_k[x]) < std::min(d.a_k[y], d.b_k[y]);
    });
    int8_t ak[QFX_MAX_BITS], bk[QFX_MAX_BITS];
    for (int i = 0; i < d.k_bits; ++i) {
      ak[i] = d.a_k[idx[i]];
      bk[i] = d.b_k[idx[i]];
    }
    memcpy(d.a_k, ak, d.k_bits);
    memcpy(d.b_k, bk, d.k_bits);
  }

  const double M = (double)(1ull << d.m_bits), N = (double)(1ull << d.n_bits),
               K = (double)(1ull << d.k_bits);
  ctx->stats.flops += 8.0 * M * N * K;
  ctx->stats.bytes += 8.0 * (M * K + K * N + M * N);
  qfx_prof_scope prof_scope(ctx, 0, din->m_bits, din->n_bits, din->k_bits, 8.0 * M * N * K,
                            8.0 * (M * K + K * N + M * N));

  const float2* A = ctx->arena + d.off_a;
  const float2* B = ctx->arena + d.off_b;
  float2* C = accumulate_amp ? nullptr : ctx->arena + d.off_c;

  // ---------------- compute-bound classes: tcgen05 split-bf16 path
  if (!accumulate_amp && d.precision != QFX_PREC_FFMA && qfx_tc_eligible(d)) {
    bool done = false;
    // (tc_pairlong first: it declines what tc_pair does better)
    int rc = qfx_launch_contract_tc_pairlong(ctx, d, A, B, C, &done);
    if (rc || done) return rc;
    rc = qfx_launch_contract_tc_pair(ctx, d, A, B, C, &done);
    if (rc || done) return rc;
    rc = qfx_launch_contract_tc_fused(ctx, d, A, B, C, &done);
    if (rc || done) return rc;
    if (d.n_bits >= 5 && d.k_bits >= 5) return qfx_launch_contract_tc(ctx, d, A, B, C);
  }

  // ---------------- K <= 4, both free sides long, small operands: C-ordered outer product
  if (!accumulate_amp && d.k_bits <= 2 && d.n_bits >= 4 && d.m_bits + d.n_bits >= 18 &&
      d.m_bits + d.k_bits <= 21 && d.n_bits + d.k_bits <= 21 && !(d.off_c & 1) &&
      getenv("QFLEX_NO_OUTER") == nullptr) {
    const int c_bits = d.m_bits + d.n_bits;
    // owner of every C address bit: (is_a, position in that operand)
    std::vector<int> own_a(c_bits, -1), own_b(c_bits, -1);
    bool dense = true;
    for (int i = 0; i < d.m_bits; ++i) {
      if (d.c_m[i] >= c_bits) dense = false;
      else own_a[d.c_m[i]] = d.a_m[i];
    }
    for (int i = 0; i < d.n_bits; ++i) {
      if (d.c_n[i] >= c_bits) dense = false;
      else own_b[d.c_n[i]] = d.b_n[i];
    }
    if (dense) {
      OuterArgs g{};
      g.A = A;
      g.B = B;
      g.C = C;
      g.c_bits = c_bits;
      g.k_bits = d.k_bits;
      for (int q = 0; q < d.k_bits; ++q) g.a_k[q] = d.a_k[q], g.b_k[q] = d.b_k[q];
      g.low_is_a = own_a[0] >= 0 ? 1 : 0;
      g.low_delta = 1u << (g.low_is_a ? own_a[0] : own_b[0]);
      bool ok = true;
      for (int p = 1; p < c_bits;) {
        const bool is_a = own_a[p] >= 0;
        const int first = is_a ? own_a[p] : own_b[p];
        int q = p + 1;
        while (q < c_bits && (own_a[q] >= 0) == is_a &&
               (is_a ? own_a[q] : own_b[q]) == first + (q - p))
          ++q;
        if (g.n_seg == 16) {
          ok = false;
          break;
        }
        g.seg[g.n_seg++] = OuterSeg{p, q - p, is_a ? first : 0, is_a ? -1 : first};
        p = q;
      }
      if (ok) {
        const unsigned long long tiles = c_bits > 12 ? 1ull << (c_bits - 12) : 1ull;
        const unsigned blocks = (unsigned)std::min<unsigned long long>(tiles, 8ull * ctx->sm_count);
        switch (d.k_bits) {
          case 0: outer_small<0><<<blocks, 256, 0, ctx->stream>>>(g); break;
          case 1: outer_small<1><<<blocks, 256, 0, ctx->stream>>>(g); break;
          default: outer_small<2><<<blocks, 256, 0, ctx->stream>>>(g); break;
        }
        QFX_CUDA(cudaGetLastError());
        ctx->stats.launches_gemm_ffma++;
        ctx->stats.launches_total++;
        return QFX_OK;
      }
    }
  }

  // ---------------- tiny N and K, long M: streaming kernel
  static const int stream_max_bits = [] {
    const char* e = getenv("QFLEX_STREAM_BITS");
    return e ? std::max(0, std::min(3, atoi(e))) : 3;
  }();
  if (!accumulate_amp && d.m_bits >= 16 && d.n_bits <= stream_max_bits &&
      d.k_bits <= stream_max_bits && getenv("QFLEX_NO_STREAM") == nullptr) {
    StreamArgs g{};
    g.A = A;
    g.B = B;
    g.C = C;
    g.m_bits = d.m_bits;
    g.n_bits = d.n_bits;
    g.k_bits = d.k_bits;
    for (int q = 0; q < d.k_bits; ++q) g.a_k[q] = d.a_k[q], g.b_k[q] = d.b_k[q];
    for (int q = 0; q < d.n_bits; ++q) g.b_n[q] = d.b_n[q], g.c_n[q] = d.c_n[q];
    // m bits in A-address order (which m bit is which is free as long as a_m and c_m
    // move together), then runs contiguous in both operands
    std::vector<int> idx(d.m_bits);
    for (int i = 0; i < d.m_bits; ++i) idx[i] = i;
    std::sort(idx.begin(), idx.end(), [&](int x, int y) { return d.a_m[x] < d.a_m[y]; });
    g.c_vec = (d.n_bits >= 1 && !(d.off_c & 1)) ? 1 : 0;
    for (int q = 0; q < d.n_bits; ++q)
      if (d.c_n[q] != q) g.c_vec = 0;
    g.a_vec = (d.k_bits >= 1 && !(d.off_a & 1)) ? 1 : 0;
    for (int q = 0; q < d.k_bits; ++q)
      if (d.a_k[q] != q) g.a_vec = 0;
    bool ok = true;
    for (int i = 0; i < d.m_bits;) {
      int j = i + 1;
      while (j < d.m_bits && d.a_m[idx[j]] == d.a_m[idx[j - 1]] + 1 &&
             d.c_m[idx[j]] == d.c_m[idx[j - 1]] + 1)
        ++j;
      if (g.n_seg == 8) {
        ok = false;
        break;
      }
      g.seg[g.n_seg++] = StreamSeg{i, j - i, d.a_m[idx[i]], d.c_m[idx[i]]};
      i = j;
    }
    if (ok) {
      const unsigned long long M = 1ull << d.m_bits;
      const unsigned blocks =
          (unsigned)std::min<unsigned long long>(M / 256, 16ull * ctx->sm_count);
      launch_stream_small(g, blocks, ctx->stream);
      QFX_CUDA(cudaGetLastError());
      ctx->stats.launches_gemm_ffma++;
      ctx->stats.launches_total++;
      return QFX_OK;
    }
  }

  // ---------------- a few outputs on one side, long K, dense big operand: linear GEMV
  if (!accumulate_amp && d.n_bits == 0 && d.m_bits >= 1 && d.m_bits <= 6 && d.k_bits >= 10 &&
      mask_a_dense(d) && !(d.off_a & 1) && getenv("QFLEX_NO_GEMV") == nullptr) {
    // The big operand is read linearly; the vector is read at the addresses its own
    // layout gives the same k.  When the two layouts order the k bits differently those
    // reads scatter (7x7 tails: 2^24-element vector, 3.0 of 6.4 TB/s).  For long K the
    // vector is first permuted into the big operand's k order (one extra read + write
    // of the SMALL side, 1/2^m of the big one).
    static const bool gemv_permute = [] {
      const char* e = getenv("QFLEX_GEMV_PERMUTE");
      return e == nullptr || e[0] != '0';
    }();
    size_t b_perm_bytes = 0;
    {
      bool same_order = true;
      std::vector<int> by_a(d.k_bits);
      for (int i = 0; i < d.k_bits; ++i) by_a[i] = i;
      std::sort(by_a.begin(), by_a.end(), [&](int x, int y) { return d.a_k[x] < d.a_k[y]; });
      for (int i = 0; i + 1 < d.k_bits; ++i)
        if (d.b_k[by_a[i]] > d.b_k[by_a[i + 1]]) same_order = false;
      uint64_t mask_b_dense = 0;
      for (int i = 0; i < d.k_bits; ++i) mask_b_dense |= 1ull << d.b_k[i];
      const bool b_dense = mask_b_dense == (1ull << d.k_bits) - 1ull;
      if (gemv_permute && !same_order && b_dense && d.k_bits >= 18) {
        b_perm_bytes = sizeof(float2) << d.k_bits;
        int8_t dst_pos[QFX_MAX_BITS];
        for (int r = 0; r < d.k_bits; ++r) dst_pos[d.b_k[by_a[r]]] = (int8_t)r;
        int ks0 = 0;  // (same split as below: the reduce workspace sits behind the copy)
        int rc = qfx_ensure_ws(ctx, b_perm_bytes + (sizeof(double2) << (d.m_bits + 12)));
        if (rc) return rc;
        (void)ks0;
        rc = qfx_launch_permute_ptr(ctx, B, reinterpret_cast<float2*>(ctx->ws), d.k_bits, dst_pos);
        if (rc) return rc;
        B = reinterpret_cast<const float2*>(ctx->ws);
        for (int r = 0; r < d.k_bits; ++r) d.b_k[by_a[r]] = (int8_t)r;
      }
    }
    GemvArgs g{};
    g.A = A;
    g.B = B;
    g.m_bits = d.m_bits;
    memcpy(g.b_k, d.b_k, QFX_MAX_BITS);
    const int bits = d.m_bits + d.k_bits;
    std::vector<int> owner(bits, 0);  // per A address bit: k index or -(m index) - 1
    for (int i = 0; i < d.k_bits; ++i) owner[d.a_k[i]] = i;
    for (int i = 0; i < d.m_bits; ++i) owner[d.a_m[i]] = -i - 1;
    for (int p = 0; p < kGemvLow; ++p) g.kind[p] = (int8_t)owner[p];
    for (int p = kGemvLow; p < bits; ++p) {
      if (owner[p] >= 0) {
        g.hk_pos[g.n_hk] = (int8_t)p;
        g.hk_bk[g.n_hk] = d.b_k[owner[p]];
        ++g.n_hk;
      } else {
        g.hm_pos[g.n_hm] = (int8_t)p;
        g.hm_idx[g.n_hm] = (int8_t)(-owner[p] - 1);
        ++g.n_hm;
      }
    }
    int ks = 0;
    while (ks < g.n_hk - 4 && (1ll << (g.n_hm + ks)) < 8ll * ctx->sm_count) ++ks;
    g.ks_bits = ks;
    int rc = qfx_ensure_ws(ctx, b_perm_bytes + (sizeof(double2) << (d.m_bits + ks)));
    if (rc) return rc;
    if (b_perm_bytes) g.B = reinterpret_cast<const float2*>(ctx->ws);  // (ws may not move: sized above)
    g.ws = reinterpret_cast<double2*>(reinterpret_cast<uint8_t*>(ctx->ws) + b_perm_bytes);
    dim3 grid(1u << g.n_hm, 1u << ks);
    gemv_rows<<<grid, 256, 0, ctx->stream>>>(g);
    QFX_CUDA(cudaGetLastError());
    DotArgs f{};
    f.C = C;
    f.ws = g.ws;
    f.m_bits = d.m_bits;
    f.ks_bits = ks;
    memcpy(f.c_m, d.c_m, QFX_MAX_BITS);
    const unsigned outs = 1u << d.m_bits;
    dotk_finish<<<(outs * 32 + 127) / 128, 128, 0, ctx->stream>>>(f);
    QFX_CUDA(cudaGetLastError());
    ctx->stats.launches_reduce += 2;
    ctx->stats.launches_total += 2;
    return QFX_OK;
  }

  // ---------------- tiny outputs: block reduction over K
  if (d.m_bits + d.n_bits <= 6) {
    DotArgs g{};
    g.A = A;
    g.B = B;
    g.C = C;
    g.amp = accumulate_amp ? ctx->amps + slot : nullptr;
    g.m_bits = d.m_bits;
    g.n_bits = d.n_bits;
    g.k_bits = d.k_bits;
    memcpy(g.a_m, d.a_m, QFX_MAX_BITS);
    memcpy(g.c_m, d.c_m, QFX_MAX_BITS);
    memcpy(g.b_n, d.b_n, QFX_MAX_BITS);
    memcpy(g.c_n, d.c_n, QFX_MAX_BITS);
    {  // K enumeration is free: lowest A address bit first (contiguous runs per warp)
      std::vector<int> idx(d.k_bits);
      for (int i = 0; i < d.k_bits; ++i) idx[i] = i;
      std::sort(idx.begin(), idx.end(), [&](int x, int y) { return d.a_k[x] < d.a_k[y]; });
      for (int i = 0; i < d.k_bits; ++i) {
        g.a_k[i] = d.a_k[idx[i]];
        g.b_k[i] = d.b_k[idx[i]];
      }
    }
    g.vec = (d.k_bits >= 1 && g.a_k[0] == 0 && g.b_k[0] == 0 && !(d.off_a & 1) && !(d.off_b & 1))
                ? 1 : 0;
    const int outs_bits = d.m_bits + d.n_bits;
    const int lo_bits = std::min(d.k_bits - g.vec, kDotLoBits);
    const int hi_bits = d.k_bits - g.vec - lo_bits;
    // each block should stream at least 16 chunks of 256 elements
    int ks = 0;
    while (ks < hi_bits - 4 && ((1 << (outs_bits + ks)) < 4 * ctx->sm_count)) ++ks;
    g.ks_bits = ks;
    if (ks > 0) {
      const size_t need = sizeof(double2) << (outs_bits + ks);
      int rc = qfx_ensure_ws(ctx, need);
      if (rc) return rc;
      g.ws = reinterpret_cast<double2*>(ctx->ws);
    }
    dim3 grid(1u << outs_bits, 1u << ks);
    dotk<<<grid, kDotThreads, 0, ctx->stream>>>(g);
    QFX_CUDA(cudaGetLastError());
    ctx->stats.launches_reduce++;
    ctx->stats.launches_total++;
    if (ks > 0) {
      const unsigned outs = 1u << outs_bits;
      dotk_finish<<<(outs * 32 + 127) / 128, 128, 0, ctx->stream>>>(g);
      QFX_CUDA(cudaGetLastError());
      ctx->stats.launches_reduce++;
      ctx->stats.launches_total++;
    }
    return QFX_OK;
  }

  // ---------------- tiled GEMM
  GemmArgs g{};
  g.A = A;
  g.B = B;
  g.C = C;
  g.ws = nullptr;
  g.m_bits = d.m_bits;
  g.n_bits = d.n_bits;
  g.k_bits = d.k_bits;
  memcpy(g.a_m, d.a_m, QFX_MAX_BITS);
  memcpy(g.c_m, d.c_m, QFX_MAX_BITS);
  memcpy(g.b_n, d.b_n, QFX_MAX_BITS);
  memcpy(g.c_n, d.c_n, QFX_MAX_BITS);
  memcpy(g.a_k, d.a_k, QFX_MAX_BITS);
  memcpy(g.b_k, d.b_k, QFX_MAX_BITS);

  int TMB, TNB, TKB;
  const int nb = d.n_bits;
  if (nb >= 7) { TMB = 7; TNB = 7; TKB = 3; }
  else if (nb == 6) { TMB = 7; TNB = 6; TKB = 4; }
  else if (nb == 5) { TMB = 8; TNB = 5; TKB = 3; }
  else if (nb == 4) { TMB = 8; TNB = 4; TKB = 3; }
  else if (nb == 3) { TMB = 8; TNB = 3; TKB = 3; }
  else if (nb == 2) { TMB = 8; TNB = 2; TKB = 3; }
  else if (nb == 1) { TMB = 8; TNB = 1; TKB = 3; }
  else { TMB = 9; TNB = 0; TKB = 3; }
  g.tm = std::min(TMB, d.m_bits);
  g.tn = std::min(TNB, d.n_bits);
  g.tk = std::min(TKB, d.k_bits);
  g.ta_n = g.tm + g.tk;
  g.tb_n = g.tk + g.tn;
  build_tile_table(g.a_m, g.tm, g.a_k, g.tk, TMB, g.ta_addr, g.ta_smem, &g.a_vec);
  build_tile_table(g.b_n, g.tn, g.b_k, g.tk, TNB, g.tb_addr, g.tb_smem, &g.b_vec);
  g.c_vec = (g.tn >= 1 && g.c_n[0] == 0) ? 1 : 0;
  // 16-byte accesses need 16-byte aligned bases
  if (d.off_a & 1) g.a_vec = 0;
  if (d.off_b & 1) g.b_vec = 0;
  if (d.off_c & 1) g.c_vec = 0;

  const int tile_bits = (d.m_bits - g.tm) + (d.n_bits - g.tn);
  const int khi_bits = d.k_bits - g.tk;
  int ks = 0;
  // split K when the output tiles cannot fill the GPU
  while (ks < khi_bits - 2 && (1ll << (tile_bits + ks)) < 2ll * ctx->sm_count &&
         (sizeof(float2) << (d.m_bits + d.n_bits + ks + 1)) <= (512ull << 20))
    ++ks;
  g.ks_bits = ks;
  if (ks > 0) {
    const size_t need = sizeof(float2) << (d.m_bits + d.n_bits + ks);
    int rc = qfx_ensure_ws(ctx, need);
    if (rc) return rc;
    g.ws = reinterpret_cast<float2*>(ctx->ws);
  }
  const unsigned blocks = 1u << (tile_bits + ks);
  int rc;
  if (nb >= 7) rc = launch_gemm_cfg<7, 7, 3, 8, 8>(ctx, g, blocks);
  else if (nb == 6) rc = launch_gemm_cfg<7, 6, 4, 8, 4>(ctx, g, blocks);
  else if (nb == 5) rc = launch_gemm_cfg<8, 5, 3, 8, 4>(ctx, g, blocks);
  else if (nb == 4) rc = launch_gemm_cfg<8, 4, 3, 4, 4>(ctx, g, blocks);
  else if (nb == 3) rc = launch_gemm_cfg<8, 3, 3, 4, 2>(ctx, g, blocks);
  else if (nb == 2) rc = launch_gemm_cfg<8, 2, 3, 4, 1>(ctx, g, blocks);
  else if (nb == 1) rc = launch_gemm_cfg<8, 1, 3, 2, 1>(ctx, g, blocks);
  else rc = launch_gemm_cfg<9, 0, 3, 2, 1>(ctx, g, blocks);
  if (rc) return rc;
  ctx->stats.launches_gemm_ffma++;
  ctx->stats.launches_total++;
  if (ks > 0) {
    FinishArgs f{};
    f.ws = g.ws;
    f.C = C;
    f.m_bits = d.m_bits;
    f.n_bits = d.n_bits;
    f.splits = 1 << ks;
    memcpy(f.c_m, d.c_m, QFX_MAX_BITS);
    memcpy(f.c_n, d.c_n, QFX_MAX_BITS);
    const unsigned long long total = 1ull << (d.m_bits + d.n_bits);
    const unsigned fb = (unsigned)std::min<unsigned long long>((total + 255) / 256,
                                                               8ull * ctx->sm_count);
    splitk_finish<<<fb, 256, 0, ctx->stream>>>(f);
    QFX_CUDA(cudaGetLastError());
    ctx->stats.launches_reduce++;
    ctx->stats.launches_total++;
  }
  return QFX_OK;
}

extern "C" int qfx_scalar_accumulate(qfx_ctx* ctx, uint64_t off, uint64_t slot) {
  if (ctx == nullptr || off >= ctx->arena_elems || slot >= ctx->n_amps) {
    qfx_set_error("qfx_scalar_accumulate: bad argument");
    return QFX_ERR_ARG;
  }
  scalar_accumulate_kernel<<<1, 1, 0, ctx->stream>>>(ctx->arena + off,
                                                      ctx->amps + slot);
  QFX_CUDA(cudaGetLastError());
  ctx->stats.launches_reduce++;
  ctx->stats.launches_total++;
  return QFX_OK;
}

namespace {
struct TensorAccArgs {
  const float2* src;
  double2* amps;
  int n_bits;
  uint32_t stride[QFX_MAX_BITS];
};
__global__ void tensor_accumulate_kernel(const __grid_constant__ TensorAccArgs g) {
  const unsigned long long total = 1ull << g.n_bits;
  for (unsigned long long i = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x;
       i < total; i += (unsigned long long)gridDim.x * blockDim.x) {
    unsigned long long slot = 0;
    for (int b = 0; b < g.n_bits; ++b) slot += ((i >> b) & 1ull) * g.stride[b];
    const float2 v = g.src[i];
    double2 a = g.amps[slot];  // slots of distinct i are distinct (host checks)
    a.x += (double)v.x;
    a.y += (double)v.y;
    g.amps[slot] = a;
  }
}
}  // namespace

extern "C" int qfx_tensor_accumulate(qfx_ctx* ctx, uint64_t off, int32_t n_bits,
                                     const uint32_t* slot_stride, uint64_t slot) {
  if (ctx == nullptr || n_bits < 0 || n_bits > QFX_MAX_BITS ||
      (n_bits > 0 && slot_stride == nullptr) || off + (1ull << n_bits) > ctx->arena_elems) {
    qfx_set_error("qfx_tensor_accumulate: bad argument");
    return QFX_ERR_ARG;
  }
  // distinct elements must land in distinct slots: with the strides sorted, each one
  // has to exceed the sum of the smaller ones
  uint64_t reach = 0;
  {
    std::vector<uint64_t> st(slot_stride, slot_stride + n_bits);
    std::sort(st.begin(), st.end());
    for (uint64_t x : st) {
      if (x == 0 || x <= reach) {
        qfx_set_error("qfx_tensor_accumulate: slot strides overlap");
        return QFX_ERR_ARG;
      }
      reach += x;
    }
  }
  if (slot + reach >= ctx->n_amps) {
    qfx_set_error("qfx_tensor_accumulate: slot outside the amplitude table");
    return QFX_ERR_ARG;
  }
  TensorAccArgs g{};
  g.src = ctx->arena + off;
  g.amps = ctx->amps + slot;
  g.n_bits = n_bits;
  for (int b = 0; b < n_bits; ++b) g.stride[b] = slot_stride[b];
  const unsigned long long total = 1ull << n_bits;
  const unsigned blocks =
      (unsigned)std::min<unsigned long long>((total + 255) / 256, 8ull * ctx->sm_count);
  tensor_accumulate_kernel<<<blocks, 256, 0, ctx->stream>>>(g);
  QFX_CUDA(cudaGetLastError());
  ctx->stats.launches_reduce++;
  ctx->stats.launches_total++;
  return QFX_OK;
}

// ---------------------------------------------------------------- chains
// World-line flattening (reference flatten_grid_of_tensors, src/read_circuit.cpp:
// 548-689): every grid site multiplies its gate tensors one after the other, a
// chain of ~30 tiny contractions (K = 2 or 4, results of 2 .. 2^20 elements), and
// the sites are independent.  One thread block runs one whole chain, step after
// step with a block barrier in between, so ~2000 launch-bound kernels (12 ms for
// bristlecone_70 1-32-1) become one launch.
namespace {

__device__ __forceinline__ uint32_t chain_deposit(uint32_t v, int bits, const int8_t* pos) {
  uint32_t o = 0;
  for (int q = 0; q < bits; ++q) o |= ((v >> q) & 1u) << pos[q];
  return o;
}

__device__ __forceinline__ void chain_point(const float2* __restrict__ A,
                                            const float2* __restrict__ B, float2* C, uint32_t ao,
                                            uint32_t bo, uint32_t co, unsigned K,
                                            const uint32_t* ka, const uint32_t* kb) {
  float sx = 0.f, sy = 0.f;
  for (unsigned k = 0; k < K; ++k) {
    const float2 a = A[ao | ka[k]];
    const float2 b = B[bo | kb[k]];
    sx = fmaf(a.x, b.x, sx);
    sx = fmaf(-a.y, b.y, sx);
    sy = fmaf(a.x, b.y, sy);
    sy = fmaf(a.y, b.x, sy);
  }
  C[co] = make_float2(sx, sy);
}

// One block = one chain.  Per step the bit tables of the descriptor become three
// per-output-bit offset words (xa/xb/xc) and two per-k offset tables in shared
// memory; a thread owns the outputs whose low 10 index bits are its thread id and
// walks the remaining bits in Gray-code order (one XOR per operand per output),
// four outputs per trip so that a late step's 2^16..2^20 outputs keep several loads
// in flight per thread -- the block is alone on its SM and latency is all it pays.
struct ChainVars {
  const qfx_var_term* terms;   // nullptr: no slice-dependent bases
  const int32_t* term_first;   // [2 * n_steps + 1]
  uint32_t values[QFX_MAX_CUT_VALUES];
};

__device__ __forceinline__ unsigned long long chain_var_offset(const ChainVars& v, int first,
                                                               int last) {
  unsigned long long off = 0;
  for (int t = first; t < last; ++t) {
    const qfx_var_term term = v.terms[t];
    unsigned long long x = (unsigned long long)v.values[term.cut] >> term.from;
    if (term.width < 32) x &= (1ull << term.width) - 1ull;
    off += x << term.shift;
  }
  return off;
}

__global__ void __launch_bounds__(1024)
    contract_chains(float2* arena, const qfx_contract_desc* descs, const int32_t* chain_first,
                    const __grid_constant__ ChainVars vars) {
  __shared__ qfx_contract_desc d;
  __shared__ uint32_t xa[32], xb[32], xc[32], ka[64], kb[64];
  const int first = chain_first[blockIdx.x], last = chain_first[blockIdx.x + 1];
  const unsigned tid = threadIdx.x;
  for (int s = first; s < last; ++s) {
    __syncthreads();  // previous step's results visible; tables no longer in use
    for (int i = tid; i < (int)(sizeof(qfx_contract_desc) / 4); i += blockDim.x)
      reinterpret_cast<uint32_t*>(&d)[i] = reinterpret_cast<const uint32_t*>(descs + s)[i];
    __syncthreads();
    if (vars.terms != nullptr && tid < 2) {
      // projected gate tensors: this slice's base of operand a (tid 0) / b (tid 1)
      const unsigned long long add = chain_var_offset(vars, vars.term_first[2 * s + tid],
                                                      vars.term_first[2 * s + tid + 1]);
      if (tid == 0) d.off_a += add;
      else d.off_b += add;
    }
    if (vars.terms != nullptr) __syncthreads();
    const int nb = d.n_bits, tb = d.m_bits + d.n_bits;
    const unsigned K = 1u << d.k_bits;
    if ((int)tid < tb) {
      const int q = (int)tid - nb;
      xa[tid] = q < 0 ? 0u : 1u << d.a_m[q];
      xb[tid] = q < 0 ? 1u << d.b_n[tid] : 0u;
      xc[tid] = q < 0 ? 1u << d.c_n[tid] : 1u << d.c_m[q];
    } else if (tid >= 64 && tid < 64 + K) {
      ka[tid - 64] = chain_deposit(tid - 64, d.k_bits, d.a_k);
      kb[tid - 64] = chain_deposit(tid - 64, d.k_bits, d.b_k);
    }
    __syncthreads();
    const float2* A = arena + d.off_a;
    const float2* B = arena + d.off_b;
    float2* C = arena + d.off_c;
    const int low = tb < 10 ? tb : 10;
    if (tid >= (1u << low)) continue;  // fewer outputs than threads
    uint32_t ao = 0, bo = 0, co = 0;
    for (int b = 0; b < low; ++b)
      if ((tid >> b) & 1u) ao |= xa[b], bo |= xb[b], co |= xc[b];
    const int hi = tb - low;  // index bits walked by this thread
    if (hi < 2) {
      chain_point(A, B, C, ao, bo, co, K, ka, kb);
      if (hi == 1) chain_point(A, B, C, ao ^ xa[low], bo ^ xb[low], co ^ xc[low], K, ka, kb);
      continue;
    }
    const uint32_t a0 = xa[low], b0 = xb[low], c0 = xc[low];
    const uint32_t a1 = xa[low + 1], b1 = xb[low + 1], c1 = xc[low + 1];
    const unsigned trips = 1u << (hi - 2);
    for (unsigned j = 0;;) {
      chain_point(A, B, C, ao, bo, co, K, ka, kb);
      chain_point(A, B, C, ao ^ a0, bo ^ b0, co ^ c0, K, ka, kb);
      chain_point(A, B, C, ao ^ a1, bo ^ b1, co ^ c1, K, ka, kb);
      chain_point(A, B, C, ao ^ a0 ^ a1, bo ^ b0 ^ b1, co ^ c0 ^ c1, K, ka, kb);
      if (++j == trips) break;
      const int f = low + 2 + __ffs(j) - 1;  // Gray code: trip j flips bit ctz(j)
      ao ^= xa[f], bo ^= xb[f], co ^= xc[f];
    }
  }
}

}  // namespace

extern "C" int qfx_contract_chains(qfx_ctx* ctx, const qfx_contract_desc* descs,
                                   const int32_t* chain_first, int32_t n_chains) {
  if (ctx == nullptr || descs == nullptr || chain_first == nullptr || n_chains < 0) {
    qfx_set_error("qfx_contract_chains: bad argument");
    return QFX_ERR_ARG;
  }
  if (n_chains == 0) return QFX_OK;
  const int n_steps = chain_first[n_chains];
  double flops = 0, bytes = 0;
  for (int c = 0; c < n_chains; ++c)
    if (chain_first[c] > chain_first[c + 1] || chain_first[c] < 0) {
      qfx_set_error("qfx_contract_chains: chain_first must be non-decreasing");
      return QFX_ERR_ARG;
    }
  for (int s = 0; s < n_steps; ++s) {
    const qfx_contract_desc& d = descs[s];
    if (d.m_bits < 0 || d.n_bits < 0 || d.k_bits < 0 || d.k_bits > 6 ||
        d.m_bits + d.n_bits > 26) {
      qfx_set_error("qfx_contract_chains: a step is not a small contraction (k <= 6 bits, "
                    "result <= 2^26 elements)");
      return QFX_ERR_ARG;
    }
    const uint64_t mask_a = operand_mask(d.a_m, d.m_bits, d.a_k, d.k_bits);
    const uint64_t mask_b = operand_mask(d.b_k, d.k_bits, d.b_n, d.n_bits);
    const uint64_t mask_c = operand_mask(d.c_m, d.m_bits, d.c_n, d.n_bits);
    if (mask_a == ~0ull || mask_b == ~0ull || mask_c == ~0ull) {
      qfx_set_error("qfx_contract_chains: bit positions of an operand must be distinct and < 32");
      return QFX_ERR_ARG;
    }
    if (d.off_a + mask_a + 1 > ctx->arena_elems || d.off_b + mask_b + 1 > ctx->arena_elems ||
        d.off_c + mask_c + 1 > ctx->arena_elems) {
      qfx_set_error("qfx_contract_chains: operand outside the arena");
      return QFX_ERR_ARG;
    }
    const double M = (double)(1ull << d.m_bits), N = (double)(1ull << d.n_bits),
                 K = (double)(1ull << d.k_bits);
    flops += 8.0 * M * N * K;
    bytes += 8.0 * (M * K + K * N + M * N);
  }
  // descriptors + chain table travel through the workspace
  const size_t desc_bytes = (size_t)n_steps * sizeof(qfx_contract_desc);
  const size_t first_bytes = (size_t)(n_chains + 1) * sizeof(int32_t);
  const size_t desc_pad = (desc_bytes + 255) & ~(size_t)255;
  int rc = qfx_ensure_ws(ctx, desc_pad + first_bytes);
  if (rc) return rc;
  uint8_t* ws = reinterpret_cast<uint8_t*>(ctx->ws);
  QFX_CUDA(cudaMemcpyAsync(ws, descs, desc_bytes, cudaMemcpyHostToDevice, ctx->stream));
  QFX_CUDA(cudaMemcpyAsync(ws + desc_pad, chain_first, first_bytes, cudaMemcpyHostToDevice,
                           ctx->stream));
  QFX_CUDA(cudaStreamSynchronize(ctx->stream));  // pageable host buffers may be reused
  ChainVars no_vars{};
  contract_chains<<<n_chains, 1024, 0, ctx->stream>>>(
      ctx->arena, reinterpret_cast<const qfx_contract_desc*>(ws),
      reinterpret_cast<const int32_t*>(ws + desc_pad), no_vars);
  QFX_CUDA(cudaGetLastError());
  ctx->stats.launches_gemm_ffma++;
  ctx->stats.launches_total++;
  ctx->stats.flops += flops;
  ctx->stats.bytes += bytes;
  return QFX_OK;
}


// ---------------------------------------------------------------- resident programs
struct qfx_program {
  uint8_t* dev = nullptr;  // descs | chain_first | terms | term_first
  size_t off_first = 0, off_terms = 0, off_term_first = 0;
  int32_t n_chains = 0, n_steps = 0, n_terms = 0;
  int32_t max_cut = -1;
  double flops = 0, bytes = 0;
};

extern "C" int qfx_program_create(qfx_ctx* ctx, const qfx_contract_desc* descs,
                                  const int32_t* chain_first, int32_t n_chains,
                                  const qfx_var_term* terms, const int32_t* term_first,
                                  qfx_program** program_out) {
  if (ctx == nullptr || descs == nullptr || chain_first == nullptr || n_chains < 0 ||
      term_first == nullptr || program_out == nullptr) {
    qfx_set_error("qfx_program_create: bad argument");
    return QFX_ERR_ARG;
  }
  *program_out = nullptr;
  const int n_steps = n_chains ? chain_first[n_chains] : 0;
  const int n_terms = term_first[2 * n_steps];
  if (n_terms > 0 && terms == nullptr) {
    qfx_set_error("qfx_program_create: terms missing");
    return QFX_ERR_ARG;
  }
  qfx_program* p = new (std::nothrow) qfx_program();
  if (p == nullptr) {
    qfx_set_error("qfx_program_create: host allocation failed");
    return QFX_ERR_OOM;
  }
  p->n_chains = n_chains;
  p->n_steps = n_steps;
  p->n_terms = n_terms;
  for (int c = 0; c < n_chains; ++c)
    if (chain_first[c] > chain_first[c + 1] || chain_first[c] < 0) {
      delete p;
      qfx_set_error("qfx_program_create: chain_first must be non-decreasing");
      return QFX_ERR_ARG;
    }
  for (int s = 0; s < n_steps; ++s) {
    const qfx_contract_desc& d = descs[s];
    bool ok = d.m_bits >= 0 && d.n_bits >= 0 && d.k_bits >= 0 && d.k_bits <= 6 &&
              d.m_bits + d.n_bits <= 26;
    const uint64_t mask_a = ok ? operand_mask(d.a_m, d.m_bits, d.a_k, d.k_bits) : ~0ull;
    const uint64_t mask_b = ok ? operand_mask(d.b_k, d.k_bits, d.b_n, d.n_bits) : ~0ull;
    const uint64_t mask_c = ok ? operand_mask(d.c_m, d.m_bits, d.c_n, d.n_bits) : ~0ull;
    ok = ok && mask_a != ~0ull && mask_b != ~0ull && mask_c != ~0ull &&
         d.off_c + mask_c + 1 <= ctx->arena_elems && term_first[2 * s] <= term_first[2 * s + 1] &&
         term_first[2 * s + 1] <= term_first[2 * s + 2];
    if (!ok) {
      delete p;
      qfx_set_error("qfx_program_create: a step is not a small contraction inside the arena "
                    "(k <= 6 bits, result <= 2^26 elements, positions distinct and < 32)");
      return QFX_ERR_ARG;
    }
    const double M = (double)(1ull << d.m_bits), N = (double)(1ull << d.n_bits),
                 K = (double)(1ull << d.k_bits);
    p->flops += 8.0 * M * N * K;
    p->bytes += 8.0 * (M * K + K * N + M * N);
  }
  for (int t = 0; t < n_terms; ++t) {
    if (terms[t].cut < 0 || terms[t].cut >= QFX_MAX_CUT_VALUES || terms[t].shift < 0 ||
        terms[t].shift > 40 || terms[t].from < 0 || terms[t].from > 31 || terms[t].width < 0) {
      delete p;
      qfx_set_error("qfx_program_create: bad var term");
      return QFX_ERR_ARG;
    }
    p->max_cut = std::max(p->max_cut, terms[t].cut);
  }
  auto pad = [](size_t x) { return (x + 255) & ~(size_t)255; };
  const size_t b_desc = pad((size_t)n_steps * sizeof(qfx_contract_desc));
  const size_t b_first = pad((size_t)(n_chains + 1) * sizeof(int32_t));
  const size_t b_terms = pad((size_t)std::max(n_terms, 1) * sizeof(qfx_var_term));
  const size_t b_tf = pad((size_t)(2 * n_steps + 1) * sizeof(int32_t));
  p->off_first = b_desc;
  p->off_terms = b_desc + b_first;
  p->off_term_first = b_desc + b_first + b_terms;
  cudaSetDevice(ctx->device);
  if (cudaMalloc(&p->dev, b_desc + b_first + b_terms + b_tf) != cudaSuccess) {
    cudaGetLastError();
    delete p;
    qfx_set_error("out of device memory allocating a chain program");
    return QFX_ERR_OOM;
  }
  cudaError_t e = cudaMemcpyAsync(p->dev, descs, (size_t)n_steps * sizeof(qfx_contract_desc),
                                  cudaMemcpyHostToDevice, ctx->stream);
  if (e == cudaSuccess)
    e = cudaMemcpyAsync(p->dev + p->off_first, chain_first, (size_t)(n_chains + 1) * sizeof(int32_t),
                        cudaMemcpyHostToDevice, ctx->stream);
  if (e == cudaSuccess && n_terms > 0)
    e = cudaMemcpyAsync(p->dev + p->off_terms, terms, (size_t)n_terms * sizeof(qfx_var_term),
                        cudaMemcpyHostToDevice, ctx->stream);
  if (e == cudaSuccess)
    e = cudaMemcpyAsync(p->dev + p->off_term_first, term_first,
                        (size_t)(2 * n_steps + 1) * sizeof(int32_t), cudaMemcpyHostToDevice,
                        ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) {
    cudaFree(p->dev);
    delete p;
    return qfx_fail_cuda(e, "qfx_program_create (upload)");
  }
  *program_out = p;
  return QFX_OK;
}

extern "C" int qfx_program_run(qfx_ctx* ctx, qfx_program* p, const uint32_t* cut_values,
                               int32_t n_values) {
  if (ctx == nullptr || p == nullptr || n_values < 0 || n_values > QFX_MAX_CUT_VALUES ||
      (n_values > 0 && cut_values == nullptr) || p->max_cut >= n_values) {
    qfx_set_error("qfx_program_run: bad argument (cut values missing for a var term?)");
    return QFX_ERR_ARG;
  }
  if (p->n_chains == 0) return QFX_OK;
  qfx_prof_scope prof_scope(ctx, 2, p->n_steps, p->n_chains, 0, p->flops, p->bytes);
  ChainVars v{};
  v.terms = reinterpret_cast<const qfx_var_term*>(p->dev + p->off_terms);
  v.term_first = reinterpret_cast<const int32_t*>(p->dev + p->off_term_first);
  for (int i = 0; i < n_values; ++i) v.values[i] = cut_values[i];
  contract_chains<<<p->n_chains, 1024, 0, ctx->stream>>>(
      ctx->arena, reinterpret_cast<const qfx_contract_desc*>(p->dev),
      reinterpret_cast<const int32_t*>(p->dev + p->off_first), v);
  QFX_CUDA(cudaGetLastError());
  ctx->stats.launches_gemm_ffma++;
  ctx->stats.launches_total++;
  ctx->stats.flops += p->flops;
  ctx->stats.bytes += p->bytes;
  return QFX_OK;
}

extern "C" int qfx_program_destroy(qfx_ctx* ctx, qfx_program* p) {
  if (p == nullptr) return QFX_OK;
  if (ctx != nullptr) {
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
  }
  if (p->dev) cudaFree(p->dev);
  delete p;
  return QFX_OK;
}
