// TEST INFRASTRUCTURE ONLY (oracle/).  A plain-C++ CPU emulation of the device
// C ABI in include/qflex_b200.h, used by the `-m "not gpu"` tests to check the
// HOST logic (parsers, planner, executor recursion, slice dealing) in a
// container without a GPU.  It is built into tests/_build/ and linked ONLY into
// a test copy of the host layer; nothing under qflex_b200/ links, loads or
// falls back to it, and the product library libqflex_b200.so fails loudly when
// no sm_100 GPU is present.
//
// Semantics restated from the reference:
//   qfx_contract  = qflex::multiply (src/tensor.cpp:1045-1194): C[l,r] =
//                   sum_c A[l,c]*B[c,r] in complex64 with fp32 accumulation;
//   qfx_permute   = Tensor::reorder (src/tensor.cpp:866-911) as a bit permutation;
//   qfx_dot_accumulate = _multiply_vv + amplitude += (src/contraction_utils.cpp:352).
#include <algorithm>
#include <complex>
#include <cstring>
#include <string>
#include <vector>

#include "qflex_b200.h"

struct qfx_ctx {
  std::vector<std::complex<float>> arena;
  std::vector<std::complex<double>> amps;
  qfx_stats stats{};
};

namespace {
thread_local std::string g_err;
int fail(const std::string& m) {
  g_err = m;
  return QFX_ERR_ARG;
}
inline uint64_t deposit(uint64_t x, const int8_t* pos, int n) {
  uint64_t o = 0;
  for (int i = 0; i < n; ++i) o |= ((x >> i) & 1ull) << pos[i];
  return o;
}
}  // namespace

extern "C" {

const char* qfx_last_error(void) { return g_err.c_str(); }
int qfx_device_count(int* n) {
  *n = 1;
  return QFX_OK;
}
int qfx_device_memory(int, uint64_t* free_b, uint64_t* total_b) {
  *free_b = 0;
  *total_b = 0;
  return QFX_OK;
}
int qfx_create(int, qfx_ctx** out) {
  *out = new qfx_ctx();
  return QFX_OK;
}
int qfx_destroy(qfx_ctx* c) {
  delete c;
  return QFX_OK;
}
int qfx_trim_cache(int) { return QFX_OK; }
int qfx_bind_thread(qfx_ctx*) { return QFX_OK; }
int qfx_sync(qfx_ctx*) { return QFX_OK; }
int qfx_arena_reserve(qfx_ctx* c, uint64_t n) {
  if (n > c->arena.size()) c->arena.assign(n, {0.f, 0.f});
  return QFX_OK;
}
int qfx_arena_size(qfx_ctx* c, uint64_t* n) {
  *n = c->arena.size();
  return QFX_OK;
}
int qfx_upload(qfx_ctx* c, uint64_t off, const float* h, uint64_t n) {
  if (off + n > c->arena.size()) return fail("qfx_upload: span outside the arena");
  std::memcpy(&c->arena[off], h, n * 8);
  return QFX_OK;
}
int qfx_download(qfx_ctx* c, uint64_t off, float* h, uint64_t n) {
  if (off + n > c->arena.size()) return fail("qfx_download: span outside the arena");
  std::memcpy(h, &c->arena[off], n * 8);
  return QFX_OK;
}
int qfx_copy(qfx_ctx* c, uint64_t s, uint64_t d, uint64_t n) {
  if (s + n > c->arena.size() || d + n > c->arena.size())
    return fail("qfx_copy: span outside the arena");
  std::memmove(&c->arena[d], &c->arena[s], n * 8);
  return QFX_OK;
}

static int contract_impl(qfx_ctx* c, const qfx_contract_desc* d, bool acc, uint64_t slot) {
  const uint64_t M = 1ull << d->m_bits, N = 1ull << d->n_bits, K = 1ull << d->k_bits;
  uint64_t mask_a = 0, mask_b = 0, mask_c = 0;
  for (int i = 0; i < d->m_bits; ++i) mask_a |= 1ull << d->a_m[i], mask_c |= 1ull << d->c_m[i];
  for (int i = 0; i < d->n_bits; ++i) mask_b |= 1ull << d->b_n[i], mask_c |= 1ull << d->c_n[i];
  for (int i = 0; i < d->k_bits; ++i) mask_a |= 1ull << d->a_k[i], mask_b |= 1ull << d->b_k[i];
  if (d->off_a + mask_a >= c->arena.size() || d->off_b + mask_b >= c->arena.size() ||
      (!acc && d->off_c + mask_c >= c->arena.size()))
    return fail("qfx_contract: operand outside the arena");
  if (acc && (d->m_bits || d->n_bits)) return fail("qfx_dot_accumulate: result is not a scalar");
  if (acc && slot >= c->amps.size()) return fail("qfx_dot_accumulate: amplitude slot out of range");
  if (!acc) {
    const uint64_t sa = mask_a + 1, sb = mask_b + 1, sc = mask_c + 1;
    if ((d->off_a < d->off_c + sc && d->off_c < d->off_a + sa) ||
        (d->off_b < d->off_c + sc && d->off_c < d->off_b + sb))
      return fail("qfx_contract: C overlaps an input operand");
  }
  std::vector<uint64_t> ka(K), kb(K), nb(N), nc(N);
  for (uint64_t k = 0; k < K; ++k) {
    ka[k] = deposit(k, d->a_k, d->k_bits);
    kb[k] = deposit(k, d->b_k, d->k_bits);
  }
  for (uint64_t n = 0; n < N; ++n) {
    nb[n] = deposit(n, d->b_n, d->n_bits);
    nc[n] = deposit(n, d->c_n, d->n_bits);
  }
  const std::complex<float>* A = &c->arena[d->off_a];
  const std::complex<float>* B = &c->arena[d->off_b];
  c->stats.flops += 8.0 * M * N * K;
  c->stats.bytes += 8.0 * (M * K + K * N + M * N);
  c->stats.launches_total++;
  if (acc) {
    std::complex<double> s = 0;
    for (uint64_t k = 0; k < K; ++k) s += std::complex<double>(A[ka[k]]) * std::complex<double>(B[kb[k]]);
    c->amps[slot] += s;
    c->stats.launches_reduce++;
    return QFX_OK;
  }
  std::complex<float>* C = &c->arena[d->off_c];
#pragma omp parallel for schedule(static)
  for (int64_t m = 0; m < (int64_t)M; ++m) {
    const uint64_t ma = deposit(m, d->a_m, d->m_bits), mc = deposit(m, d->c_m, d->m_bits);
    std::vector<float> re(N, 0.f), im(N, 0.f);
    for (uint64_t k = 0; k < K; ++k) {
      const std::complex<float> a = A[ma + ka[k]];
      const std::complex<float>* Bk = B + kb[k];
      for (uint64_t n = 0; n < N; ++n) {
        const std::complex<float> b = Bk[nb[n]];
        re[n] += a.real() * b.real() - a.imag() * b.imag();
        im[n] += a.real() * b.imag() + a.imag() * b.real();
      }
    }
    for (uint64_t n = 0; n < N; ++n) C[mc + nc[n]] = {re[n], im[n]};
  }
  c->stats.launches_gemm_ffma++;
  return QFX_OK;
}

int qfx_contract(qfx_ctx* c, const qfx_contract_desc* d) { return contract_impl(c, d, false, 0); }
int qfx_contract_chains(qfx_ctx* c, const qfx_contract_desc* d, const int32_t* first, int32_t n) {
  if (c == nullptr || d == nullptr || first == nullptr || n < 0)
    return fail("qfx_contract_chains: bad argument");
  for (int32_t ch = 0; ch < n; ++ch)
    for (int32_t s = first[ch]; s < first[ch + 1]; ++s)
      if (int rc = contract_impl(c, d + s, false, 0)) return rc;
  return QFX_OK;
}
struct qfx_program {
  std::vector<qfx_contract_desc> descs;
  std::vector<int32_t> chain_first, term_first;
  std::vector<qfx_var_term> terms;
};
int qfx_program_create(qfx_ctx* c, const qfx_contract_desc* d, const int32_t* first, int32_t n,
                       const qfx_var_term* terms, const int32_t* term_first, qfx_program** out) {
  if (c == nullptr || d == nullptr || first == nullptr || n < 0 || term_first == nullptr || out == nullptr)
    return fail("qfx_program_create: bad argument");
  auto* p = new qfx_program();
  const int n_steps = n ? first[n] : 0;
  p->descs.assign(d, d + n_steps);
  p->chain_first.assign(first, first + n + 1);
  p->term_first.assign(term_first, term_first + 2 * n_steps + 1);
  if (p->term_first.back() > 0) p->terms.assign(terms, terms + p->term_first.back());
  *out = p;
  return QFX_OK;
}
int qfx_program_run(qfx_ctx* c, qfx_program* p, const uint32_t* values, int32_t n_values) {
  if (c == nullptr || p == nullptr) return fail("qfx_program_run: bad argument");
  auto offset = [&](int first, int last, uint64_t* off) {
    for (int t = first; t < last; ++t) {
      const qfx_var_term& term = p->terms[t];
      if (term.cut < 0 || term.cut >= n_values) return false;
      uint64_t x = (uint64_t)values[term.cut] >> term.from;
      if (term.width < 32) x &= (1ull << term.width) - 1ull;
      *off += x << term.shift;
    }
    return true;
  };
  const uint64_t launches = c->stats.launches_total;
  for (std::size_t s = 0; s < p->descs.size(); ++s) {
    qfx_contract_desc d = p->descs[s];
    if (!offset(p->term_first[2 * s], p->term_first[2 * s + 1], &d.off_a) ||
        !offset(p->term_first[2 * s + 1], p->term_first[2 * s + 2], &d.off_b))
      return fail("qfx_program_run: cut values missing for a var term");
    if (int rc = contract_impl(c, &d, false, 0)) return rc;
  }
  c->stats.launches_total = launches + 1;  // one launch on the device
  return QFX_OK;
}
int qfx_program_destroy(qfx_ctx*, qfx_program* p) {
  delete p;
  return QFX_OK;
}
int qfx_dot_accumulate(qfx_ctx* c, const qfx_contract_desc* d, uint64_t slot) {
  return contract_impl(c, d, true, slot);
}
int qfx_scalar_accumulate(qfx_ctx* c, uint64_t off, uint64_t slot) {
  if (off >= c->arena.size() || slot >= c->amps.size()) return fail("qfx_scalar_accumulate: bad argument");
  c->amps[slot] += std::complex<double>(c->arena[off]);
  c->stats.launches_total++;
  return QFX_OK;
}
int qfx_tensor_accumulate(qfx_ctx* c, uint64_t off, int32_t nb, const uint32_t* stride,
                          uint64_t slot) {
  if (c == nullptr || nb < 0 || nb > QFX_MAX_BITS || (nb > 0 && stride == nullptr) ||
      off + (1ull << nb) > c->arena.size())
    return fail("qfx_tensor_accumulate: bad argument");
  std::vector<uint64_t> st(stride, stride + nb);
  std::sort(st.begin(), st.end());
  uint64_t reach = 0;
  for (uint64_t x : st) {
    if (x == 0 || x <= reach) return fail("qfx_tensor_accumulate: slot strides overlap");
    reach += x;
  }
  if (slot + reach >= c->amps.size())
    return fail("qfx_tensor_accumulate: slot outside the amplitude table");
  for (uint64_t i = 0; i < (1ull << nb); ++i) {
    uint64_t s = slot;
    for (int b = 0; b < nb; ++b) s += ((i >> b) & 1ull) * stride[b];
    c->amps[s] += std::complex<double>(c->arena[off + i]);
  }
  c->stats.launches_total++;
  return QFX_OK;
}
int qfx_amps_reserve(qfx_ctx* c, uint64_t n) {
  if (n == 0) return fail("qfx_amps_reserve: bad argument");
  c->amps.assign(std::max<uint64_t>(n, c->amps.size()), {0.0, 0.0});
  return QFX_OK;
}
int qfx_amps_download(qfx_ctx* c, double* h, uint64_t n) {
  if (n > c->amps.size()) return fail("qfx_amps_download: bad argument");
  std::memcpy(h, c->amps.data(), n * 16);
  return QFX_OK;
}
int qfx_permute(qfx_ctx* c, uint64_t s, uint64_t d, int32_t nb, const int8_t* pos) {
  const uint64_t n = 1ull << nb;
  if (s + n > c->arena.size() || d + n > c->arena.size()) return fail("qfx_permute: span outside the arena");
  if (s < d + n && d < s + n) return fail("qfx_permute: source and destination overlap");
  for (uint64_t i = 0; i < n; ++i) c->arena[d + deposit(i, pos, nb)] = c->arena[s + i];
  c->stats.launches_total++;
  c->stats.launches_permute++;
  return QFX_OK;
}
int qfx_timer_start(qfx_ctx*) { return QFX_OK; }
int qfx_timer_stop_ms(qfx_ctx*, float* ms) {
  *ms = 0.f;
  return QFX_OK;
}
int qfx_profile_begin(qfx_ctx*) { return QFX_OK; }
int qfx_profile_end(qfx_ctx*, qfx_profile_row*, int32_t, int32_t* n) {
  *n = 0;
  return QFX_OK;
}
int qfx_stats_get(qfx_ctx* c, qfx_stats* o) {
  *o = c->stats;
  return QFX_OK;
}
int qfx_stats_reset(qfx_ctx* c) {
  c->stats = qfx_stats{};
  return QFX_OK;
}

}  // extern "C"
