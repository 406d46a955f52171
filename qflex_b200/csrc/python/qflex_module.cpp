// Python module `qflex` on the B200 engine: the reference's pybind surface
// (src/pybind_main.h:19-23, src/pybind_main.cpp:75-138) --
//     qflex.simulate(options: dict) -> list[(bitstring, complex)]
// with the same option keys (circuit / circuit_filename, ordering /
// ordering_filename, grid / grid_filename, initial_state(s), final_state(s),
// memory_limit) and the same error behaviour (message on stderr, empty list), so that
// `from qflexcirq import qflex` style callers (qflexcirq/interface/qflex_simulator.py:
// 5,44-66) can import it instead.  Beyond the reference: a LIST of final states is
// served from one resident circuit (the reference answers "Not yet supported",
// src/pybind_main.cpp:111-113), and the additive keys device / rank / world /
// precision place the run.
#include <pybind11/complex.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <complex>
#include <iostream>
#include <memory>
#include <sstream>
#include <string>
#include <utility>
#include <vector>

#include "errors.h"
#include "evaluate_circuit.h"

namespace py = pybind11;

namespace {

using Amplitudes = std::vector<std::pair<std::string, std::complex<double>>>;

std::string text_of(const py::dict& options, const char* key, bool* found) {
  *found = false;
  if (!options.contains(key)) return "";
  const py::object value = options[key];
  if (py::isinstance<py::str>(value) || !py::isinstance<py::iterable>(value))
    throw ERROR_MSG("'", key, "' must be a list of strings.");
  std::string text;
  for (const auto& line : value.cast<py::iterable>()) {
    if (!py::isinstance<py::str>(line)) throw ERROR_MSG(key, " must be a list of strings.");
    text += line.cast<std::string>();
    text += '\n';
  }
  *found = true;
  return text;
}

template <typename Part>
void load_part(const py::dict& options, const std::string& key, Part* part) {
  bool found = false;
  const std::string text = text_of(options, key.c_str(), &found);
  if (found) {
    part->load(std::stringstream(text));
    return;
  }
  const std::string fkey = key + "_filename";
  if (!options.contains(fkey.c_str())) return;
  const py::object name = options[fkey.c_str()];
  if (!py::isinstance<py::str>(name)) throw ERROR_MSG("'", key, "_filename' must be a string.");
  part->load(name.cast<std::string>());
}

std::vector<std::string> states_of(const py::dict& options, const std::string& which) {
  const std::string many = which + "_states", one = which + "_state";
  std::vector<std::string> out;
  if (options.contains(many.c_str())) {
    const py::object v = options[many.c_str()];
    if (py::isinstance<py::str>(v)) {
      out.push_back(v.cast<std::string>());
    } else if (py::isinstance<py::iterable>(v)) {
      for (const auto& x : v.cast<py::iterable>()) {
        if (!py::isinstance<py::str>(x)) throw ERROR_MSG("states must be strings.");
        out.push_back(x.cast<std::string>());
      }
    } else {
      throw ERROR_MSG(which, "_states must be a list of strings.");
    }
  } else if (options.contains(one.c_str())) {
    const py::object v = options[one.c_str()];
    if (!py::isinstance<py::str>(v)) throw ERROR_MSG("states must be strings.");
    out.push_back(v.cast<std::string>());
  } else {
    out.push_back("");
  }
  return out;
}

int int_option(const py::dict& options, const char* key, int fallback) {
  return options.contains(key) ? options[key].cast<int>() : fallback;
}

Amplitudes simulate(const py::dict& options) {
  try {
    qflex::QflexInput input;
    if (options.contains("memory_limit")) {
      const py::object ml = options["memory_limit"];
      if (py::isinstance<py::int_>(ml)) {
        const long v = ml.cast<long>();
        if (v <= 0) throw ERROR_MSG("'memory_limit' must be positive.");
        qflex::global::memory_limit = (std::size_t)v;
      } else if (py::isinstance<py::str>(ml)) {
        qflex::global::memory_limit =
            qflex::utils::from_readable_memory_string(ml.cast<std::string>());
      } else {
        throw ERROR_MSG("'memory_limit' must be valid.");
      }
    } else {
      qflex::global::memory_limit = 1L << 30;
    }
    load_part(options, "grid", &input.grid);
    load_part(options, "circuit", &input.circuit);
    load_part(options, "ordering", &input.ordering);
    const std::vector<std::string> initial = states_of(options, "initial");
    const std::vector<std::string> final_states = states_of(options, "final");
    if (initial.size() != 1) throw ERROR_MSG("Not yet supported");

    qflex::DevicePlacement where;
    where.device = int_option(options, "device", 0);
    where.rank = int_option(options, "rank", 0);
    where.world = int_option(options, "world", 1);
    if (options.contains("precision")) {
      const std::string p = options["precision"].cast<std::string>();
      if (p == "ffma") where.precision = QFX_PREC_FFMA;
      else if (p == "bf16x3") where.precision = QFX_PREC_BF16X3;
      else if (p != "auto") throw ERROR_MSG("'precision' must be auto, ffma or bf16x3.");
    }
    input.initial_state = initial[0];
    input.final_state = final_states[0];

    Amplitudes result;
    py::gil_scoped_release release;
    if (final_states.size() == 1) return qflex::EvaluateCircuit(&input, where, nullptr);
    // a batch of output bitstrings: one resident circuit, one delta rewrite +
    // flatten + contraction per state
    std::unique_ptr<qflex::PreparedCircuit> prepared;
    {
      // final states as projections of open output legs when that fits (another
      // bitstring then costs the contraction alone), else rewritten deltas + flatten
      qflex::QflexInput projected = input;
      where.resident_finals = true;
      try {
        prepared = qflex::PrepareCircuit(&projected, where);
      } catch (const std::string&) {
        where.resident_finals = false;
        prepared = qflex::PrepareCircuit(&input, where);
      }
    }
    std::vector<std::complex<double>> amplitudes;
    for (std::size_t s = 0; s < final_states.size(); ++s) {
      if (s) prepared->SetFinalState(final_states[s]);
      prepared->session->Contract(where.rank, where.world, &amplitudes);
      for (std::size_t c = 0; c < amplitudes.size(); ++c)
        result.emplace_back(prepared->output_states[c], amplitudes[c]);
    }
    return result;
  } catch (const std::string& err_msg) {
    std::cerr << err_msg << std::endl;
  } catch (const char* err_msg) {
    std::cerr << err_msg << std::endl;
  } catch (const py::error_already_set&) {
    throw;
  } catch (const std::exception& e) {
    std::cerr << e.what() << std::endl;
  }
  return {};
}

}  // namespace

PYBIND11_MODULE(qflex, m) {
  m.doc() = "qFlex simulate() on the B200 contraction engine (qflex_b200)";
  m.def("simulate", &simulate, "Call the simulator with the normal parameters");
}
