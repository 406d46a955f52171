// TEST INFRASTRUCTURE ONLY (oracle/). Thin extern "C" shim over the UNMODIFIED
// reference sources, compiled where they lie under /root/reference/src by
// oracle/Makefile into oracle/_ref/libqflex_ref.so. It lets the Python tests and
// bench.py's cpu_baseline / --impl reference legs call the reference's own
//   - qflex::EvaluateCircuit      (src/evaluate_circuit.cpp:106-253)
//   - qflex::multiply             (src/tensor.cpp:1045-1194)
//   - qflex::Tensor::reorder      (src/tensor.cpp:866-911)
// through ctypes. Nothing under qflex_b200/ links or loads this library.
#include <chrono>
#include <complex>
#include <cstring>
#include <sstream>
#include <string>
#include <vector>

#include "cblas.h"
#include "evaluate_circuit.h"
#include "global.h"
#include "input.h"
#include "tensor.h"

namespace {

std::vector<std::string> split(const char* s, char sep) {
  std::vector<std::string> out;
  if (s == nullptr || *s == 0) return out;
  std::string cur;
  for (const char* p = s; *p; ++p) {
    if (*p == sep) {
      out.push_back(cur);
      cur.clear();
    } else {
      cur.push_back(*p);
    }
  }
  out.push_back(cur);
  return out;
}

void set_err(char* err, std::size_t cap, const std::string& msg) {
  if (err == nullptr || cap == 0) return;
  std::size_t n = std::min(cap - 1, msg.size());
  std::memcpy(err, msg.data(), n);
  err[n] = 0;
}

std::string join(const std::vector<std::string>& v, char sep) {
  std::string out;
  for (std::size_t i = 0; i < v.size(); ++i) {
    if (i) out.push_back(sep);
    out += v[i];
  }
  return out;
}

}  // namespace

extern "C" {

void ref_set_blas_threads(int n) { scipy_openblas_set_num_threads(n); }
int ref_get_blas_threads() { return scipy_openblas_get_num_threads(); }

// Runs the reference EvaluateCircuit on in-memory texts.
// states_out: n_out strings of length state_len, '\n'-separated, NUL-terminated.
// amps_out:   2*n_out doubles (re, im).
// Returns n_out (>= 0) or -1 on error (message in err).
long ref_evaluate(const char* circuit_text, const char* ordering_text,
                  const char* grid_text, const char* initial_state,
                  const char* final_state, unsigned long memory_limit_bytes,
                  char* states_out, unsigned long states_cap, double* amps_out,
                  unsigned long amps_cap_complex, double* seconds, char* err,
                  unsigned long err_cap) {
  try {
    qflex::global::verbose = 0;
    qflex::global::memory_limit = memory_limit_bytes;
    qflex::QflexInput input;
    input.circuit.load(std::stringstream(circuit_text));
    input.ordering.load(std::stringstream(ordering_text));
    input.grid.load(std::stringstream(grid_text));
    input.initial_state = initial_state ? initial_state : "";
    input.final_state = final_state ? final_state : "";
    auto t0 = std::chrono::steady_clock::now();
    auto result = qflex::EvaluateCircuit(&input);
    auto t1 = std::chrono::steady_clock::now();
    if (seconds) *seconds = std::chrono::duration<double>(t1 - t0).count();
    if (result.size() > amps_cap_complex) {
      set_err(err, err_cap, "amps_out too small");
      return -1;
    }
    std::string states;
    for (std::size_t i = 0; i < result.size(); ++i) {
      if (i) states.push_back('\n');
      states += result[i].first;
      amps_out[2 * i] = result[i].second.real();
      amps_out[2 * i + 1] = result[i].second.imag();
    }
    if (states.size() + 1 > states_cap) {
      set_err(err, err_cap, "states_out too small");
      return -1;
    }
    std::memcpy(states_out, states.c_str(), states.size() + 1);
    return static_cast<long>(result.size());
  } catch (const std::string& msg) {
    set_err(err, err_cap, msg);
    return -1;
  } catch (const std::exception& ex) {
    set_err(err, err_cap, ex.what());
    return -1;
  }
}

// multiply(A, B, C, scratch) on caller-provided complex64 data.
// Index names are ';'-separated. c_indices_out receives ';'-joined names.
// A and B data are updated in place (the reference reorders them).
long ref_multiply(const char* a_indices, const unsigned long* a_dims,
                  float* a_data, const char* b_indices,
                  const unsigned long* b_dims, float* b_data, float* c_data,
                  unsigned long c_capacity, char* c_indices_out,
                  unsigned long c_indices_cap, unsigned long* c_dims_out,
                  char* err, unsigned long err_cap) {
  try {
    using qflex::s_type;
    auto ai = split(a_indices, ';');
    auto bi = split(b_indices, ';');
    std::vector<std::size_t> ad(a_dims, a_dims + ai.size());
    std::vector<std::size_t> bd(b_dims, b_dims + bi.size());
    std::size_t as = 1, bs = 1;
    for (auto d : ad) as *= d;
    for (auto d : bd) bs *= d;
    const s_type* ap = reinterpret_cast<const s_type*>(a_data);
    const s_type* bp = reinterpret_cast<const s_type*>(b_data);
    qflex::Tensor A(ai, ad, std::vector<s_type>(ap, ap + as));
    qflex::Tensor B(bi, bd, std::vector<s_type>(bp, bp + bs));
    qflex::Tensor C({"_c_"}, {static_cast<std::size_t>(c_capacity)});
    std::vector<s_type> scratch(std::max(as, bs));
    qflex::multiply(A, B, C, scratch.data());
    std::memcpy(a_data, A.data(), as * sizeof(s_type));
    std::memcpy(b_data, B.data(), bs * sizeof(s_type));
    std::memcpy(c_data, C.data(), C.size() * sizeof(s_type));
    std::string names = join(C.get_indices(), ';');
    if (names.size() + 1 > c_indices_cap) {
      set_err(err, err_cap, "c_indices_out too small");
      return -1;
    }
    std::memcpy(c_indices_out, names.c_str(), names.size() + 1);
    for (std::size_t i = 0; i < C.get_dimensions().size(); ++i)
      c_dims_out[i] = C.get_dimensions()[i];
    return static_cast<long>(C.get_indices().size());
  } catch (const std::string& msg) {
    set_err(err, err_cap, msg);
    return -1;
  } catch (const std::exception& ex) {
    set_err(err, err_cap, ex.what());
    return -1;
  }
}

// Tensor::reorder(new_ordering, scratch) in place on caller data.
long ref_reorder(const char* indices, const unsigned long* dims, float* data,
                 const char* new_indices, char* err, unsigned long err_cap) {
  try {
    using qflex::s_type;
    auto idx = split(indices, ';');
    auto nidx = split(new_indices, ';');
    std::vector<std::size_t> d(dims, dims + idx.size());
    std::size_t n = 1;
    for (auto x : d) n *= x;
    const s_type* p = reinterpret_cast<const s_type*>(data);
    qflex::Tensor T(idx, d, std::vector<s_type>(p, p + n));
    std::vector<s_type> scratch(n);
    T.reorder(nidx, scratch.data());
    std::memcpy(data, T.data(), n * sizeof(s_type));
    return 0;
  } catch (const std::string& msg) {
    set_err(err, err_cap, msg);
    return -1;
  } catch (const std::exception& ex) {
    set_err(err, err_cap, ex.what());
    return -1;
  }
}

}  // extern "C"
