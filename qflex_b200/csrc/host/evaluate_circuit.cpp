// EvaluateCircuit on the B200 engine.  Control flow follows the reference
// driver (src/evaluate_circuit.cpp:106-253): ordering -> output states ->
// network -> flatten -> contract -> zip(states, amplitudes); the numeric work is
// a device plan (plan.h) instead of Tensor::multiply/reorder calls.
#include "evaluate_circuit.h"

#include <algorithm>
#include <cstdlib>

#include <chrono>
#include <iostream>

#include "errors.h"

namespace qflex {
namespace {

// Position of grid site `pos` in the string of active qubits.
std::size_t find_output_pos(const QflexInput* input, const std::vector<std::size_t>& pos) {
  std::size_t p = pos[0] * input->grid.J + pos[1];
  for (const auto& off : input->grid.qubits_off)
    if (off[0] < pos[0] || (off[0] == pos[0] && off[1] < pos[1])) --p;
  return p;
}

double seconds_since(const std::chrono::steady_clock::time_point& t0) {
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

}  // namespace

std::string get_output_states(
    const QflexInput* input, const std::list<ContractionOperation>& ordering,
    std::vector<std::vector<std::size_t>>* final_qubits,
    std::vector<std::string>* output_states) {
  if (input == nullptr) throw ERROR_MSG("Input must be non-null.");
  if (final_qubits == nullptr) throw ERROR_MSG("Final qubits must be non-null.");
  if (output_states == nullptr) throw ERROR_MSG("Output states must be non-null");

  std::string base_state = input->final_state;
  const bool unspecified = input->final_state.empty();
  if (unspecified) base_state = std::string(input->initial_state.length(), '0');

  std::vector<std::size_t> where;
  std::vector<std::vector<std::size_t>> what;
  for (const auto& op : ordering) {
    if (op.op_type != ContractionOperation::CUT || op.cut.tensors.size() != 1) continue;
    const std::size_t p = find_output_pos(input, op.cut.tensors[0]);
    if (unspecified) base_state[p] = 'x';
    where.push_back(p);
    what.push_back(op.cut.values.empty() ? std::vector<std::size_t>{0, 1} : op.cut.values);
    final_qubits->push_back(op.cut.tensors[0]);
  }
  // each terminal cut multiplies the list; earlier cuts vary slowest
  output_states->push_back(base_state);
  for (std::size_t c = 0; c < where.size(); ++c) {
    std::vector<std::string> grown;
    for (const std::string& state : *output_states)
      for (const std::size_t val : what[c]) {
        std::string s = state;
        s[where[c]] = std::to_string(val).at(0);
        grown.push_back(std::move(s));
      }
    *output_states = std::move(grown);
  }
  for (std::size_t i = 0; i < output_states->at(0).length(); ++i) {
    const char c = output_states->at(0)[i];
    if (c != '0' && c != '1')
      throw ERROR_MSG("Final state has non-binary character ", c, " at index ", i,
                      "despite having no terminal cut there.");
  }
  return base_state;
}

std::unique_ptr<PreparedCircuit> PrepareCircuit(QflexInput* input,
                                                const DevicePlacement& where) {
  if (input == nullptr) throw ERROR_MSG("Input must be non-null.");
  const auto t0 = std::chrono::steady_clock::now();
  double t_prev = 0;
  auto lap = [&](const char* what) {
    const double t = seconds_since(t0);
    if (global::verbose > 0)
      std::cerr << WARN_MSG(what, t - t_prev, "s") << std::endl;
    t_prev = t;
  };

  std::list<ContractionOperation> ordering;
  ordering_data_to_contraction_ordering(*input, &ordering);
  lap("Time spent making contraction ordering: ");

  const std::size_t n_active =
      input->grid.I * input->grid.J - input->grid.qubits_off.size();
  if (input->initial_state.empty()) input->initial_state = std::string(n_active, '0');

  auto prepared = std::make_unique<PreparedCircuit>();
  std::vector<std::vector<std::size_t>> final_qubits;
  try {
    input->final_state =
        get_output_states(input, ordering, &final_qubits, &prepared->output_states);
  } catch (const std::string& err_msg) {
    throw ERROR_MSG("Failed to call get_output_states(). Error:\n\t[", err_msg, "]");
  }

  // As many terminal cuts as fit the memory limit stay open (counted from the end of
  // the ordering; the arena grows monotonically with that count, so bisect).
  // QFLEX_OPEN_OUTPUTS=<n> caps the count, 0 = the reference's projected leaves, whose
  // plan is also the one whose errors are reported.
  // When that plan cannot be built or does not fit (grid tensors next to cut bonds too
  // large), the same once more with the cut-aware flatten (BuildPlan, plan.h);
  // QFLEX_CUT_AWARE_FLATTEN=1 forces it, =0 forbids it.
  Plan plan;
  int max_open = (int)final_qubits.size();
  if (const char* e = getenv("QFLEX_OPEN_OUTPUTS")) max_open = std::min(max_open, std::max(0, atoi(e)));
  const auto fits = [&](const Plan& p) {
    return (double)p.arena_elems * sizeof(s_type) <= (double)global::memory_limit;
  };
  auto plan_with = [&](bool cut_aware, Plan* result) {
    bool planned = false;
    auto try_open = [&](int n) {
      try {
        Plan p = BuildPlan(*input, ordering, final_qubits, n, cut_aware);
        if (p.open_outputs == 0 || !fits(p)) return false;
        *result = std::move(p);
        return true;
      } catch (const std::string&) {
        return false;
      }
    };
    if (max_open > 0) {
      planned = try_open(max_open);
      if (!planned) {
        int lo = 0, hi = max_open;  // lo fits (0 = projected), hi does not
        while (hi - lo > 1) {
          const int mid = (lo + hi) / 2;
          if (try_open(mid)) lo = mid, planned = true;
          else hi = mid;
        }
      }
    }
    if (!planned) *result = BuildPlan(*input, ordering, final_qubits, 0, cut_aware);
    if (!fits(*result)) return;

    // Then undo as much of the bond-cut slicing as the memory limit allows (BuildPlan's
    // fuse_bits: 2^f slices per pass of larger kernels), keeping at least one outer
    // slice per rank of a multi-GPU run.  QFLEX_FUSE_BITS=<n> caps f, 0 = the ordering
    // as written.  The arena grows monotonically with f, so bisect.
    const int n_open = result->open_outputs;
    int max_fuse = 62;
    if (const char* e = getenv("QFLEX_FUSE_BITS")) max_fuse = std::max(0, atoi(e));
    double slices = 1;
    for (const auto& c : result->cuts)
      if (!c.terminal) slices *= (double)std::max<std::size_t>(1, c.values.size());
    const int ranks = std::max(1, where.world);
    if (slices >= ranks) {
      int cap = 0;
      while (slices / (double)(uint64_t(2) << cap) >= ranks) ++cap;
      max_fuse = std::min(max_fuse, cap);
    }
    if (max_fuse <= 0) return;
    auto try_fuse = [&](int f, Plan* out) {
      try {
        Plan p = BuildPlan(*input, ordering, final_qubits, n_open, cut_aware, f);
        if (p.open_outputs != n_open || !fits(p)) return false;
        *out = std::move(p);
        return true;
      } catch (const std::string&) {
        return false;
      }
    };
    Plan best;
    int have = 0;
    {
      // how many bits the ordering offers at all
      Plan probe;
      int offered = 0;
      try {
        probe = BuildPlan(*input, ordering, final_qubits, n_open, cut_aware, max_fuse);
        offered = probe.fused_bits;
        if (offered > 0 && probe.open_outputs == n_open && fits(probe)) {
          best = std::move(probe);
          have = offered;
        }
      } catch (const std::string&) {
        offered = max_fuse;
      }
      if (have == 0 && offered > 0) {
        int lo = 0, hi = std::min(offered, max_fuse);  // lo fits, hi does not
        while (hi - lo > 1) {
          const int mid = (lo + hi) / 2;
          if (try_fuse(mid, &best)) lo = mid, have = mid;
          else hi = mid;
        }
      }
    }
    if (have > 0) *result = std::move(best);
  };
  const char* ca_env = getenv("QFLEX_CUT_AWARE_FLATTEN");
  const bool ca_forced = ca_env && ca_env[0] == '1', ca_forbidden = ca_env && ca_env[0] == '0';
  if (ca_forced) {
    plan_with(true, &plan);
  } else {
    std::string first_error;
    bool ok = false;
    try {
      plan_with(false, &plan);
      ok = fits(plan);
    } catch (const std::string& err_msg) {
      first_error = err_msg;
    }
    if (!ok && !ca_forbidden) {
      try {
        Plan second;
        plan_with(true, &second);
        if (first_error.empty() && !fits(second)) throw std::string("does not fit either");
        plan = std::move(second);
        ok = true;
      } catch (const std::string&) {
        // report the plain plan's outcome
      }
    }
    if (!ok && !first_error.empty()) throw first_error;
  }
  plan.precision = where.precision;
  for (auto& st : plan.steps) st.desc.precision = where.precision;
  lap("Time spent planning the contraction: ");

  const double alloc_size = (double)plan.arena_elems * sizeof(s_type);
  if (alloc_size > (double)global::memory_limit)
    throw ERROR_MSG("Required space (", utils::readable_memory_string(alloc_size),
                    ") exceeds memory limit (",
                    utils::readable_memory_string((double)global::memory_limit),
                    "). Cancelling simulation.");
  if (global::verbose > 0)
    std::cerr << WARN_MSG("Allocating ", utils::readable_memory_string(alloc_size),
                          " of device memory for this simulation.")
              << std::endl;
  if (plan.n_amplitudes != prepared->output_states.size())
    throw ERROR_MSG("Planner produced ", plan.n_amplitudes, " amplitudes for ",
                    prepared->output_states.size(), " output states.");

  for (const auto& q : final_qubits)
    prepared->output_positions.push_back(find_output_pos(input, q));
  prepared->session = std::make_unique<Session>(std::move(plan), where.device);
  lap("Time spent creating 2D grid of tensors on the device: ");
  return prepared;
}

void PreparedCircuit::SetFinalState(const std::string& final_state) {
  if (!session) throw ERROR_MSG("Circuit is not resident.");
  session->SetFinalState(final_state);
  for (std::string& s : output_states) {
    std::string next = final_state;
    for (const std::size_t p : output_positions) next[p] = s[p];
    s = std::move(next);
  }
}

std::vector<std::pair<std::string, std::complex<double>>> EvaluateCircuit(
    QflexInput* input, const DevicePlacement& where, ExecStats* stats) {
  const auto t0 = std::chrono::steady_clock::now();
  auto prepared = PrepareCircuit(input, where);
  std::vector<std::complex<double>> amplitudes;
  try {
    prepared->session->Contract(where.rank, where.world, &amplitudes);
  } catch (const std::string& err_msg) {
    throw ERROR_MSG("Failed to call ContractGrid(). Error:\n\t[", err_msg, "]");
  }
  std::vector<std::pair<std::string, std::complex<double>>> result;
  for (std::size_t c = 0; c < amplitudes.size(); ++c)
    result.emplace_back(prepared->output_states[c], amplitudes[c]);
  if (stats) *stats = prepared->session->stats();
  if (global::verbose > 0) {
    std::cerr << WARN_MSG("Time spent contracting tensors: ",
                          prepared->session->stats().seconds_contract, "s")
              << std::endl;
    std::cerr << WARN_MSG("Total time: ", seconds_since(t0), "s") << std::endl;
  }
  return result;
}

std::vector<std::pair<std::string, std::complex<double>>> EvaluateCircuit(
    QflexInput* input) {
  return EvaluateCircuit(input, DevicePlacement{}, nullptr);
}

}  // namespace qflex
