// EvaluateCircuit on the B200 engine.  Control flow follows the reference
// driver (src/evaluate_circuit.cpp:106-253): ordering -> output states ->
// network -> flatten -> contract -> zip(states, amplitudes); the numeric work is
// a device plan (plan.h) instead of Tensor::multiply/reorder calls.
#include "evaluate_circuit.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>

#include <chrono>
#include <iostream>
#include <mutex>

#include "errors.h"

namespace qflex {
namespace {

// Position of grid site `pos` in the state strings: its rank among the active qubits
// in row-major order.
std::size_t find_output_pos(const QflexInput* input, const std::vector<std::size_t>& pos) {
  const std::size_t J = input->grid.J, target = pos[0] * J + pos[1];
  std::size_t skipped = 0;
  for (const auto& q : input->grid.qubits_off)
    if (q.size() == 2 && q[0] * J + q[1] < target) ++skipped;
  return target - skipped;
}

// ---- plan memory (see PrepareCircuit)
std::string PlanKey(const QflexInput& input, const std::vector<std::vector<std::size_t>>& final_qubits,
                    const DevicePlacement& where) {
  std::string key;
  key.reserve(64 * 1024);
  auto add = [&key](const std::string& s) {
    key += s;
    key += '\n';
  };
  add(utils::concat("grid ", input.grid.I, "x", input.grid.J));
  for (const auto& q : input.grid.qubits_off) add(utils::concat("off ", q[0], ",", q[1]));
  add(utils::concat("circuit ", input.circuit.num_active_qubits, " ", input.circuit.gates.size()));
  for (const auto& g : input.circuit.gates) add(g.raw);
  for (const auto& line : input.ordering.instructions) add(line);
  add("initial " + input.initial_state);
  // final state: only where a terminal cut makes it an output marker; the bits of the
  // other qubits do not change the plan
  std::string fs = input.final_state;
  {
    std::vector<char> is_output(fs.size(), 0);
    const std::size_t J = input.grid.J;
    for (const auto& q : final_qubits) {
      std::size_t p = q[0] * J + q[1];
      for (const auto& o : input.grid.qubits_off)
        if (o.size() == 2 && o[0] * J + o[1] < q[0] * J + q[1]) --p;
      if (p < is_output.size()) is_output[p] = 1;
    }
    for (std::size_t i = 0; i < fs.size(); ++i)
      if (!is_output[i] && (fs[i] == '0' || fs[i] == '1')) fs[i] = 'b';
  }
  add("final " + fs);
  add(utils::concat("memory ", global::memory_limit, " world ", where.world, " devices ",
                    where.devices.size(), " finals ", (int)where.resident_finals));
  for (const char* var : {"QFLEX_OPEN_OUTPUTS", "QFLEX_FUSE_BITS", "QFLEX_FUSE_TARGET_FLOPS",
                          "QFLEX_CUT_AWARE_FLATTEN", "QFLEX_WORLDLINE_TREE", "QFLEX_FOLD_LAYOUT",
                          "QFLEX_LINE_PROGRAMS"}) {
    const char* e = getenv(var);
    add(utils::concat(var, "=", e ? e : ""));
  }
  return key;
}

std::mutex g_plan_mu;
std::list<std::pair<std::string, Plan>> g_plans;  // most recent first
constexpr std::size_t kPlansKept = 4;

bool RecallPlan(const std::string& key, Plan* plan) {
  if (const char* e = getenv("QFLEX_PLAN_MEMORY"); e != nullptr && e[0] == '0') return false;
  std::lock_guard<std::mutex> lock(g_plan_mu);
  for (auto it = g_plans.begin(); it != g_plans.end(); ++it)
    if (it->first == key) {
      g_plans.splice(g_plans.begin(), g_plans, it);
      *plan = g_plans.front().second;
      return true;
    }
  return false;
}

void RememberPlan(const std::string& key, const Plan& plan) {
  if (const char* e = getenv("QFLEX_PLAN_MEMORY"); e != nullptr && e[0] == '0') return;
  std::lock_guard<std::mutex> lock(g_plan_mu);
  g_plans.emplace_front(key, plan);
  while (g_plans.size() > kPlansKept) g_plans.pop_back();
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

  // One digit per terminal cut, in ordering order: where it sits in the state string
  // and which values it takes.
  struct Digit {
    std::size_t where;
    std::vector<std::size_t> values;
  };
  std::vector<Digit> digits;
  for (const auto& op : ordering) {
    if (op.op_type != ContractionOperation::CUT || op.cut.tensors.size() != 1) continue;
    Digit d{find_output_pos(input, op.cut.tensors[0]), op.cut.values};
    if (d.values.empty()) d.values = {0, 1};
    digits.push_back(std::move(d));
    final_qubits->push_back(op.cut.tensors[0]);
  }

  // The template: the requested final state, or all zeros with 'x' at the outputs.
  std::string base = input->final_state;
  if (base.empty()) {
    base.assign(input->initial_state.length(), '0');
    for (const auto& d : digits) base[d.where] = 'x';
  }

  // Mixed-radix count over the digits, first terminal cut most significant: amplitude
  // number n belongs to the state whose digits are n's expansion (the slot formula
  // output_index * |values| + i of src/contraction_utils.cpp:300-314).
  std::size_t total = 1;
  for (const auto& d : digits) total *= d.values.size();
  output_states->reserve(output_states->size() + total);
  for (std::size_t n = 0; n < total; ++n) {
    std::string state = base;
    std::size_t rest = n;
    for (std::size_t k = digits.size(); k--;) {
      const Digit& d = digits[k];
      state[d.where] = (char)('0' + d.values[rest % d.values.size()] % 10);
      rest /= d.values.size();
    }
    output_states->push_back(std::move(state));
  }
  const std::string& first = output_states->at(0);
  for (std::size_t i = 0; i < first.length(); ++i)
    if (first[i] != '0' && first[i] != '1')
      throw ERROR_MSG("Final state has non-binary character ", first[i], " at index ", i,
                      "despite having no terminal cut there.");
  return base;
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
  if (where.resident_finals) {
    // one projected terminal cut per qubit the ordering does not cut itself, in front
    // of everything else (a cut must precede the use of its tensor); value = the
    // qubit's bit of the requested final state
    std::vector<char> has_cut(input->grid.I * input->grid.J, 0);
    for (const auto& op : ordering)
      if (op.op_type == ContractionOperation::CUT && op.cut.tensors.size() == 1)
        has_cut[op.cut.tensors[0][0] * input->grid.J + op.cut.tensors[0][1]] = 1;
    std::vector<char> off(input->grid.I * input->grid.J, 0);
    for (const auto& q : input->grid.qubits_off)
      if (q.size() == 2 && q[0] < input->grid.I && q[1] < input->grid.J)
        off[q[0] * input->grid.J + q[1]] = 1;
    if (!input->final_state.empty() && input->final_state.size() != n_active)
      throw ERROR_MSG("Size of final_conf: ", input->final_state.size(),
                      ", must be equal to size of initial_conf: ", n_active, ".");
    std::list<ContractionOperation> cuts;
    std::size_t a = 0;
    int id = 0;
    for (std::size_t q = 0; q < off.size(); ++q) {
      if (off[q]) continue;
      if (has_cut[q]) {
        prepared->state_cut.push_back(-1);
      } else {
        const char c = input->final_state.empty() ? '0' : input->final_state[a];
        if (c != '0' && c != '1')
          throw ERROR_MSG("Final state has non-binary character ", c, " at index ", a,
                          "despite having no terminal cut there.");
        cuts.emplace_back(CutIndex({get_qubit_coords(q, input->grid.J)}, {std::size_t(c - '0')}));
        prepared->state_cut.push_back(id++);
      }
      ++a;
    }
    ordering.splice(ordering.begin(), cuts);
  }
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
  // Plans are remembered across calls (the reference's Python interface calls
  // simulate() once per output bitstring, qflexcirq/interface/qflex_simulator.py:44-66:
  // same circuit, grid and ordering every time).  The key is everything the plan
  // depends on EXCEPT the bits of the final state on qubits without a terminal cut --
  // those only select the closing delta tensors, which are patched on a hit.
  const std::string plan_key = PlanKey(*input, final_qubits, where);
  if (!RecallPlan(plan_key, &plan)) {
    int max_open = (int)final_qubits.size();
    if (const char* e = getenv("QFLEX_OPEN_OUTPUTS")) max_open = std::min(max_open, std::max(0, atoi(e)));
    const auto fits = [&](const Plan& p) {
      return (double)p.arena_elems * sizeof(s_type) <= (double)global::memory_limit;
    };
    PlanCache cache;  // phase 1 of BuildPlan, shared by the searches below
    auto plan_with = [&](bool cut_aware, Plan* result) {
      bool planned = false;
      auto try_open = [&](int n) {
        try {
          Plan p = BuildPlan(*input, ordering, final_qubits, n, cut_aware, 0, &cache);
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
      if (!planned) *result = BuildPlan(*input, ordering, final_qubits, 0, cut_aware, 0, &cache);
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
      const int ranks = std::max(1, where.world) * std::max<int>(1, (int)where.devices.size());
      if (slices >= ranks) {
        int cap = 0;
        while (slices / (double)(uint64_t(2) << cap) >= ranks) ++cap;
        max_fuse = std::min(max_fuse, cap);
      }
      if (max_fuse <= 0) return;
      auto try_fuse = [&](int f, Plan* out) {
        try {
          Plan p = BuildPlan(*input, ordering, final_qubits, n_open, cut_aware, f, &cache);
          if (p.open_outputs != n_open || !fits(p)) return false;
          *out = std::move(p);
          return true;
        } catch (const std::string&) {
          return false;
        }
      };
      // Fusing pays while the schedule is a stream of small launches; once the average
      // contraction is large (>= ~1e11 flop, hundreds of microseconds on the tensor
      // cores) it only grows the arena and the merges' K.  Every fused bit about doubles
      // the work per launch: aim for the smallest f that reaches the target
      // (QFLEX_FUSE_TARGET_FLOPS), back off by bisection when it does not fit.
      double target = 1e11;
      if (const char* e = getenv("QFLEX_FUSE_TARGET_FLOPS")) target = std::max(1.0, atof(e));
      const double mean0 =
          result->flops_per_pass / std::max<double>(1.0, (double)result->steps.size());
      int want = 0;
      while (want < max_fuse && std::ldexp(mean0, want) < target) ++want;
      if (want <= 0) return;
      Plan best;
      int have = 0;
      if (try_fuse(want, &best)) {
        have = want;
      } else {
        int lo = 0, hi = want;  // lo fits, hi does not
        while (hi - lo > 1) {
          const int mid = (lo + hi) / 2;
          if (try_fuse(mid, &best)) lo = mid, have = mid;
          else hi = mid;
        }
        // `best` holds the last successful try, which is lo (bisection only raises lo)
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
    RememberPlan(plan_key, plan);
  }
  // closing deltas of THIS final state (a recalled plan carries another one's)
  if (!where.resident_finals) {
    const std::string& fs = input->final_state;
    for (std::size_t a = 0; a < plan.final_delta_upload.size() && a < fs.size(); ++a) {
      const int u = plan.final_delta_upload[a];
      if (u < 0 || plan.uploads[u].data.size() != 2) continue;
      plan.uploads[u].data[0] = s_type(fs[a] == '0' ? 1.f : 0.f, 0.f);
      plan.uploads[u].data[1] = s_type(fs[a] == '1' ? 1.f : 0.f, 0.f);
    }
  } else {
    // projected output legs: the values of the state cuts
    std::size_t a = 0;
    for (std::size_t q = 0; q < prepared->state_cut.size(); ++q, ++a) {
      const int c = prepared->state_cut[q];
      if (c >= 0 && c < (int)plan.cuts.size() && plan.cuts[c].values.size() == 1 &&
          a < input->final_state.size())
        plan.cuts[c].values[0] = std::size_t(input->final_state[a] - '0');
    }
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
  prepared->session =
      where.devices.empty()
          ? std::make_unique<Session>(std::move(plan), where.device)
          : std::make_unique<Session>(std::move(plan), where.devices);
  lap("Time spent creating 2D grid of tensors on the device: ");
  return prepared;
}

void PreparedCircuit::SetFinalState(const std::string& final_state) {
  if (!session) throw ERROR_MSG("Circuit is not resident.");
  if (!state_cut.empty()) {
    // every non-output qubit carries a projected terminal cut: move the projections
    if (final_state.size() != state_cut.size())
      throw ERROR_MSG("Size of final_conf: ", final_state.size(),
                      ", must be equal to size of initial_conf: ", state_cut.size(), ".");
    for (std::size_t a = 0; a < state_cut.size(); ++a) {
      if (state_cut[a] < 0) continue;
      const char c = final_state[a];
      if (c != '0' && c != '1')
        throw ERROR_MSG("Final state has non-binary character ", c, " at index ", a,
                        "despite having no terminal cut there.");
      session->SetCutValue(state_cut[a], std::size_t(c - '0'));
    }
    for (std::string& s : output_states) {
      std::string next = final_state;
      for (std::size_t a = 0; a < state_cut.size(); ++a)
        if (state_cut[a] < 0) next[a] = s[a];
      s = std::move(next);
    }
    return;
  }
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

std::vector<int> ParseDeviceList(const std::string& text) {
  std::vector<int> out;
  if (text.empty()) return out;
  if (text == "all") {
    int n = 0;
    if (qfx_device_count(&n) != QFX_OK) throw ERROR_MSG("No CUDA device: ", qfx_last_error());
    for (int d = 0; d < n; ++d) out.push_back(d);
    return out;
  }
  std::size_t pos = 0;
  while (pos <= text.size()) {
    const std::size_t comma = std::min(text.find(',', pos), text.size());
    const std::string item = text.substr(pos, comma - pos);
    if (item.empty() || item.find_first_not_of("0123456789") != std::string::npos)
      throw ERROR_MSG("Device list '", text, "' must be comma-separated device numbers or 'all'.");
    out.push_back(std::atoi(item.c_str()));
    pos = comma + 1;
  }
  return out;
}

std::vector<std::pair<std::string, std::complex<double>>> EvaluateCircuit(
    QflexInput* input) {
  // the reference's signature; the GPUs come from the environment
  DevicePlacement where;
  if (const char* e = getenv("QFLEX_DEVICES")) where.devices = ParseDeviceList(e);
  return EvaluateCircuit(input, where, nullptr);
}

}  // namespace qflex
