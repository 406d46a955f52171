// extern "C" surface of libqflex_host.so (include/qflex_b200_host.h).
#include <algorithm>
#include <cstring>
#include <exception>
#include <sstream>
#include <string>
#include <vector>

#include "errors.h"
#include "evaluate_circuit.h"
#include "qflex_b200_host.h"

namespace {

void set_err(char* err, uint64_t cap, const std::string& msg) {
  if (err == nullptr || cap == 0) return;
  const std::size_t n = std::min<std::size_t>(cap - 1, msg.size());
  std::memcpy(err, msg.data(), n);
  err[n] = 0;
}

void load_input(qflex::QflexInput* input, const char* circuit_text,
                const char* ordering_text, const char* grid_text,
                const char* initial_state, const char* final_state) {
  input->circuit.load(std::stringstream(circuit_text ? circuit_text : ""));
  input->ordering.load(std::stringstream(ordering_text ? ordering_text : ""));
  input->grid.load(std::stringstream(grid_text ? grid_text : ""));
  input->initial_state = initial_state ? initial_state : "";
  input->final_state = final_state ? final_state : "";
}

void fill_stats(const qflex::Session& s, double* out) {
  if (out == nullptr) return;
  const qflex::ExecStats& e = s.stats();
  const qflex::Plan& p = s.plan();
  out[0] = e.seconds_upload;
  out[1] = e.seconds_flatten;
  out[2] = e.seconds_contract;
  out[3] = e.device_ms_contract;
  out[4] = (double)e.leaves;
  out[5] = (double)e.device.launches_total;
  out[6] = e.device.flops;
  out[7] = e.device.bytes;
  out[8] = (double)p.arena_elems * 8.0;
  out[9] = p.flops_per_pass;
  out[10] = p.bytes_per_pass;
  out[11] = (double)e.device.launches_gemm_ffma;
  out[12] = (double)e.device.launches_gemm_tc;
  out[13] = (double)e.device.launches_reduce;
  out[14] = (double)e.device.launches_permute;
  out[15] = (double)p.n_amplitudes;
  out[16] = (double)p.fused_bits;
  out[17] = (double)p.open_outputs;
  out[18] = (double)s.lanes();
  // flops of ONE execution of the whole job (every slice, each step counted once per
  // combination of the cuts before it) -- what a single rank alone would execute
  double all = 0.0;
  {
    std::vector<double> fan(p.cuts.size() + 1, 1.0);
    for (std::size_t c = 0; c < p.cuts.size(); ++c)
      fan[c + 1] = fan[c] * (double)std::max<std::size_t>(1, p.cuts[c].values.size());
    for (const auto& st : p.steps) all += st.flops * fan[std::min<std::size_t>(st.level, p.cuts.size())];
  }
  out[19] = all;
}

std::string join_states(const std::vector<std::string>& states) {
  std::string s;
  for (std::size_t i = 0; i < states.size(); ++i) {
    if (i) s.push_back('\n');
    s += states[i];
  }
  return s;
}

const char* kind_name(qflex::PlanStep::Kind k) {
  switch (k) {
    case qflex::PlanStep::CONTRACT: return "contract";
    case qflex::PlanStep::DOT_ACCUMULATE: return "dot";
    case qflex::PlanStep::SCALAR_ACCUMULATE: return "scalar";
    case qflex::PlanStep::TENSOR_ACCUMULATE: return "amps";
    case qflex::PlanStep::PERMUTE: return "permute";
    case qflex::PlanStep::PROGRAM: return "program";
  }
  return "?";
}

}  // namespace

#define QFLEX_CAPI_CATCH(ret)                                   \
  catch (const std::string& msg) {                              \
    set_err(err, err_cap, msg);                                 \
    return ret;                                                 \
  }                                                             \
  catch (const std::exception& ex) {                            \
    set_err(err, err_cap, ex.what());                           \
    return ret;                                                 \
  }

extern "C" {

void qflex_b200_set_verbose(int level) { qflex::global::verbose = level; }

void* qflex_b200_prepare_sharded(const char* circuit_text, const char* ordering_text,
                                 const char* grid_text, const char* initial_state,
                                 const char* final_state, uint64_t memory_limit_bytes,
                                 int device, int precision, int world, char* err,
                                 uint64_t err_cap) {
  try {
    if (memory_limit_bytes) qflex::global::memory_limit = memory_limit_bytes;
    qflex::QflexInput input;
    load_input(&input, circuit_text, ordering_text, grid_text, initial_state, final_state);
    qflex::DevicePlacement where;
    where.device = device;
    where.precision = precision;
    where.world = world < 1 ? 1 : world;
    return qflex::PrepareCircuit(&input, where).release();
  }
  QFLEX_CAPI_CATCH(nullptr)
}

void* qflex_b200_prepare_on(const char* circuit_text, const char* ordering_text,
                            const char* grid_text, const char* initial_state,
                            const char* final_state, uint64_t memory_limit_bytes,
                            const int* devices, int n_devices, int precision, int world,
                            char* err, uint64_t err_cap) {
  try {
    if (memory_limit_bytes) qflex::global::memory_limit = memory_limit_bytes;
    if (devices == nullptr || n_devices < 1) throw std::string("empty device list");
    qflex::QflexInput input;
    load_input(&input, circuit_text, ordering_text, grid_text, initial_state, final_state);
    qflex::DevicePlacement where;
    where.devices.assign(devices, devices + n_devices);
    where.device = devices[0];
    where.precision = precision;
    where.world = world < 1 ? 1 : world;
    return qflex::PrepareCircuit(&input, where).release();
  }
  QFLEX_CAPI_CATCH(nullptr)
}

void* qflex_b200_prepare_batch(const char* circuit_text, const char* ordering_text,
                               const char* grid_text, const char* initial_state,
                               const char* final_state, uint64_t memory_limit_bytes,
                               int device, int precision, int world, char* err,
                               uint64_t err_cap) {
  try {
    if (memory_limit_bytes) qflex::global::memory_limit = memory_limit_bytes;
    qflex::DevicePlacement where;
    where.device = device;
    where.precision = precision;
    where.world = world < 1 ? 1 : world;
    // first with the final state as projections of open output legs (another bitstring
    // costs nothing but the contraction); if that plan does not fit, the plain one
    // (another bitstring = rewritten deltas + flatten)
    for (const bool projected : {true, false}) {
      qflex::QflexInput input;
      load_input(&input, circuit_text, ordering_text, grid_text, initial_state, final_state);
      where.resident_finals = projected;
      try {
        return qflex::PrepareCircuit(&input, where).release();
      } catch (const std::string&) {
        if (!projected) throw;
      }
    }
    return nullptr;
  }
  QFLEX_CAPI_CATCH(nullptr)
}

void* qflex_b200_prepare(const char* circuit_text, const char* ordering_text,
                         const char* grid_text, const char* initial_state,
                         const char* final_state, uint64_t memory_limit_bytes,
                         int device, int precision, char* err, uint64_t err_cap) {
  return qflex_b200_prepare_sharded(circuit_text, ordering_text, grid_text, initial_state,
                                    final_state, memory_limit_bytes, device, precision, 1, err,
                                    err_cap);
}

long qflex_b200_contract(void* prepared, int rank, int world, double* amps_out,
                         uint64_t amps_cap_complex, double* stats_out, char* err,
                         uint64_t err_cap) {
  try {
    if (prepared == nullptr) throw std::string("null handle");
    auto* p = static_cast<qflex::PreparedCircuit*>(prepared);
    std::vector<std::complex<double>> amps;
    p->session->Contract(rank, world, &amps);
    if (amps.size() > amps_cap_complex) throw std::string("amps_out too small");
    for (std::size_t i = 0; i < amps.size(); ++i) {
      amps_out[2 * i] = amps[i].real();
      amps_out[2 * i + 1] = amps[i].imag();
    }
    fill_stats(*p->session, stats_out);
    return (long)amps.size();
  }
  QFLEX_CAPI_CATCH(-1)
}

long qflex_b200_reload(void* prepared, char* err, uint64_t err_cap) {
  try {
    if (prepared == nullptr) throw std::string("null handle");
    auto* p = static_cast<qflex::PreparedCircuit*>(prepared);
    p->session->Upload();
    p->session->Flatten();
    return 0;
  }
  QFLEX_CAPI_CATCH(-1)
}

long qflex_b200_plan_stats(void* prepared, double* stats_out) {
  if (prepared == nullptr || stats_out == nullptr) return -1;
  fill_stats(*static_cast<qflex::PreparedCircuit*>(prepared)->session, stats_out);
  return QFLEX_B200_N_STATS;
}

long qflex_b200_set_final_state(void* prepared, const char* final_state, char* err,
                                uint64_t err_cap) {
  try {
    if (prepared == nullptr) throw std::string("null handle");
    static_cast<qflex::PreparedCircuit*>(prepared)->SetFinalState(final_state ? final_state : "");
    return 0;
  }
  QFLEX_CAPI_CATCH(-1)
}

long qflex_b200_states(void* prepared, char* states_out, uint64_t states_cap) {
  if (prepared == nullptr) return -1;
  auto* p = static_cast<qflex::PreparedCircuit*>(prepared);
  const std::string s = join_states(p->output_states);
  if (states_out != nullptr && s.size() + 1 <= states_cap)
    std::memcpy(states_out, s.c_str(), s.size() + 1);
  return (long)s.size() + 1;
}

long qflex_b200_describe(void* prepared, char* out, uint64_t cap) {
  if (prepared == nullptr) return -1;
  auto* p = static_cast<qflex::PreparedCircuit*>(prepared);
  std::stringstream ss;
  auto dump = [&](const char* phase, const std::vector<qflex::PlanStep>& steps) {
    for (const auto& st : steps) {
      ss << phase << ' ' << st.level << ' ' << kind_name(st.kind) << ' '
         << (st.kind == qflex::PlanStep::PERMUTE || st.kind == qflex::PlanStep::PROGRAM ? st.n_bits : st.desc.m_bits) << ' '
         << st.desc.n_bits << ' ' << st.desc.k_bits << ' ' << st.label;
      if (st.kind != qflex::PlanStep::PERMUTE && st.kind != qflex::PlanStep::PROGRAM) {
        // operand bit tables (debugging aid for kernel work)
        auto tab = [&](const char* name, const int8_t* t, int n) {
          ss << ' ' << name << '=';
          for (int i = 0; i < n; ++i) ss << (i ? "," : "") << (int)t[i];
        };
        ss << " |";
        tab("a_m", st.desc.a_m, st.desc.m_bits);
        tab("a_k", st.desc.a_k, st.desc.k_bits);
        tab("b_n", st.desc.b_n, st.desc.n_bits);
        tab("b_k", st.desc.b_k, st.desc.k_bits);
        tab("c_m", st.desc.c_m, st.desc.m_bits);
        tab("c_n", st.desc.c_n, st.desc.n_bits);
      }
      ss << '\n';
    }
  };
  dump("flatten", p->session->plan().flatten_steps);
  dump("grid", p->session->plan().steps);
  const std::string s = ss.str();
  if (out != nullptr && s.size() + 1 <= cap) std::memcpy(out, s.c_str(), s.size() + 1);
  return (long)s.size() + 1;
}

void qflex_b200_profile(void* prepared, int enable) {
  if (prepared == nullptr) return;
  static_cast<qflex::PreparedCircuit*>(prepared)->session->set_profile(enable != 0);
}

long qflex_b200_profile_rows(void* prepared, double* out, uint64_t cap_rows) {
  if (prepared == nullptr) return -1;
  const auto& rows = static_cast<qflex::PreparedCircuit*>(prepared)->session->profile_rows();
  for (std::size_t i = 0; i < rows.size() && i < cap_rows && out != nullptr; ++i) {
    double* o = out + 8 * i;
    o[0] = rows[i].kind;
    o[1] = rows[i].m_bits;
    o[2] = rows[i].n_bits;
    o[3] = rows[i].k_bits;
    o[4] = (double)rows[i].launches;
    o[5] = rows[i].ms;
    o[6] = rows[i].flops;
    o[7] = rows[i].bytes;
  }
  return (long)rows.size();
}

void qflex_b200_release(void* prepared) {
  delete static_cast<qflex::PreparedCircuit*>(prepared);
}

long qflex_b200_evaluate(const char* circuit_text, const char* ordering_text,
                         const char* grid_text, const char* initial_state,
                         const char* final_state, uint64_t memory_limit_bytes,
                         int device, int rank, int world, int precision,
                         char* states_out, uint64_t states_cap, double* amps_out,
                         uint64_t amps_cap_complex, double* stats_out, char* err,
                         uint64_t err_cap) {
  try {
    if (memory_limit_bytes) qflex::global::memory_limit = memory_limit_bytes;
    qflex::QflexInput input;
    load_input(&input, circuit_text, ordering_text, grid_text, initial_state, final_state);
    qflex::DevicePlacement where;
    where.device = device;
    where.rank = rank;
    where.world = world;
    where.precision = precision;
    qflex::ExecStats stats;
    auto prepared = qflex::PrepareCircuit(&input, where);
    // capacity checks before any device work, so a caller can retry cheaply
    if (prepared->output_states.size() > amps_cap_complex)
      throw std::string("amps_out too small");
    const std::string states = join_states(prepared->output_states);
    if (states.size() + 1 > states_cap) throw std::string("states_out too small");
    std::vector<std::complex<double>> amps;
    prepared->session->Contract(rank, world, &amps);
    if (amps.size() > amps_cap_complex) throw std::string("amps_out too small");
    std::memcpy(states_out, states.c_str(), states.size() + 1);
    for (std::size_t i = 0; i < amps.size(); ++i) {
      amps_out[2 * i] = amps[i].real();
      amps_out[2 * i + 1] = amps[i].imag();
    }
    fill_stats(*prepared->session, stats_out);
    return (long)amps.size();
  }
  QFLEX_CAPI_CATCH(-1)
}

}  // extern "C"
