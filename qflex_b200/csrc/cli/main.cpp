// qflex_b200.x -- command-line front end with the reference's usage, output
// format and exit codes (reference src/main.cpp:14-158): prints one line
// "<initial> --> <final>: <re> <im>" per output bitstring; exit 1 on
// std::exception, 2 on a qFlex error string.
// Additive options: --device, --precision.
#include <complex>
#include <cstring>
#include <fstream>
#include <iostream>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "../host/errors.h"
#include "../host/evaluate_circuit.h"
#include "../host/memory.h"

namespace {

const char* kVersion = "qFlex-B200 0.1";

std::string usage() {
  return qflex::utils::concat(
      R"(Flexible Quantum Circuit Simulator (qFlex), B200 tensor-network contraction engine.

  Usage:
    qflex_b200.x <circuit_filename> <ordering_filename> <grid_filename> [<initial_conf> <final_conf> <verbosity_level> <memory_limit> <track_memory_seconds>]
    qflex_b200.x -c <circuit_filename> -o <ordering_filename> -g <grid_filename> [--initial-conf=<initial_conf> --final-conf=<final_conf> --verbosity=<verbosity_level> --memory=<memory_limit> --track-memory=<seconds>] [--device=<id> | --devices=<list>] [--precision=<auto|ffma|bf16x3>]
    qflex_b200.x (-h | --help)
    qflex_b200.x --version

  Options:
    -h,--help                              Show this help.
    -c,--circuit=<circuit_filename>        Circuit filename.
    -o,--ordering=<ordering_filename>      Ordering filename.
    -g,--grid=<grid_filename>              Grid filename.
    -v,--verbosity=<verbosity_level>       Verbosity level [default: )",
      qflex::global::verbose, R"(].
    -m,--memory=<memory_limit>             Memory limit [default: )",
      qflex::utils::readable_memory_string((double)qflex::global::memory_limit), R"(].
    -t,--track-memory=<seconds>            If <verbosity_level> > 0, track memory usage [default: )",
      qflex::global::track_memory_seconds, R"(].
    --initial-conf=<initial_conf>          Initial configuration.
    --final-conf=<final_conf>              Final configuration.
    --device=<id>                          CUDA device [default: 0].
    --devices=<list>                       Several GPUs, e.g. 0,1,2,3 or all: the cut slices are
                                           dealt over them (default: $QFLEX_DEVICES).
    --precision=<mode>                     auto | ffma | bf16x3 [default: auto].
    --version                              Show version.
)");
}

struct Args {
  std::map<std::string, std::string> opt;
  std::vector<std::string> positional;
};

// "-c X", "-cX", "--circuit=X", "--circuit X"
Args parse(int argc, char** argv) {
  static const std::map<std::string, std::string> shorts = {
      {"-c", "--circuit"}, {"-o", "--ordering"}, {"-g", "--grid"},
      {"-v", "--verbosity"}, {"-m", "--memory"}, {"-t", "--track-memory"}};
  static const std::vector<std::string> longs = {
      "--circuit", "--ordering", "--grid", "--verbosity", "--memory",
      "--track-memory", "--initial-conf", "--final-conf", "--device", "--devices", "--precision"};
  Args a;
  for (int i = 1; i < argc; ++i) {
    std::string tok = argv[i];
    if (tok == "-h" || tok == "--help") {
      a.opt["--help"] = "1";
    } else if (tok == "--version") {
      a.opt["--version"] = "1";
    } else if (tok.rfind("--", 0) == 0) {
      std::string key = tok, val;
      const std::size_t eq = tok.find('=');
      bool has_val = false;
      if (eq != std::string::npos) {
        key = tok.substr(0, eq);
        val = tok.substr(eq + 1);
        has_val = true;
      }
      bool known = false;
      for (const auto& l : longs) known = known || (l == key);
      if (!known) throw std::runtime_error("Unexpected argument: " + tok + "\n" + usage());
      if (!has_val) {
        if (i + 1 >= argc) throw std::runtime_error(key + " requires an argument\n" + usage());
        val = argv[++i];
      }
      a.opt[key] = val;
    } else if (tok.size() >= 2 && tok[0] == '-' && shorts.count(tok.substr(0, 2))) {
      const std::string key = shorts.at(tok.substr(0, 2));
      std::string val = tok.substr(2);
      if (val.empty()) {
        if (i + 1 >= argc) throw std::runtime_error(tok + " requires an argument\n" + usage());
        val = argv[++i];
      }
      a.opt[key] = val;
    } else if (tok.size() >= 2 && tok[0] == '-' && !(tok[1] >= '0' && tok[1] <= '9')) {
      throw std::runtime_error("Unexpected argument: " + tok + "\n" + usage());
    } else {
      a.positional.push_back(tok);
    }
  }
  return a;
}

}  // namespace

int main(int argc, char** argv) {
  try {
    const Args args = parse(argc, argv);
    if (args.opt.count("--help")) {
      std::cout << usage() << std::endl;
      return 0;
    }
    if (args.opt.count("--version")) {
      std::cout << kVersion << std::endl;
      return 0;
    }
    const bool flag_form = args.opt.count("--circuit") || args.opt.count("--ordering") ||
                           args.opt.count("--grid");
    auto pick = [&](const char* key, std::size_t pos) -> std::string {
      auto it = args.opt.find(key);
      if (it != args.opt.end()) return it->second;
      if (!flag_form && pos < args.positional.size()) return args.positional[pos];
      return "";
    };
    if (flag_form) {
      if (!args.positional.empty() || !args.opt.count("--circuit") ||
          !args.opt.count("--ordering") || !args.opt.count("--grid"))
        throw std::runtime_error("Arguments did not match expected patterns\n" + usage());
    } else {
      if (args.positional.size() < 3 || args.positional.size() > 8 ||
          args.opt.count("--initial-conf") || args.opt.count("--final-conf") ||
          args.opt.count("--verbosity") || args.opt.count("--memory") ||
          args.opt.count("--track-memory"))
        throw std::runtime_error("Arguments did not match expected patterns\n" + usage());
    }
    const std::string circuit_filename = pick("--circuit", 0);
    const std::string ordering_filename = pick("--ordering", 1);
    const std::string grid_filename = pick("--grid", 2);

    qflex::QflexInput input;
    input.initial_state = pick("--initial-conf", 3);
    input.final_state = pick("--final-conf", 4);
    if (const std::string v = pick("--verbosity", 5); !v.empty())
      qflex::global::verbose = std::stol(v);
    if (const std::string m = pick("--memory", 6); !m.empty())
      qflex::global::memory_limit = qflex::utils::from_readable_memory_string(m);
    if (const std::string t = pick("--track-memory", 7); !t.empty())
      qflex::global::track_memory_seconds = std::stol(t);

    qflex::DevicePlacement where;
    if (args.opt.count("--device")) where.device = std::stoi(args.opt.at("--device"));
    if (args.opt.count("--devices")) where.devices = qflex::ParseDeviceList(args.opt.at("--devices"));
    else if (const char* e = getenv("QFLEX_DEVICES")) where.devices = qflex::ParseDeviceList(e);
    // --track-memory: the reference's SIGALRM report (src/main.cpp:108-112), host + device
    if (qflex::global::verbose > 0 && qflex::global::track_memory_seconds > 0) {
      signal(SIGALRM, qflex::memory::print_memory_usage);
      alarm(qflex::global::track_memory_seconds);
    }
    if (args.opt.count("--precision")) {
      const std::string p = args.opt.at("--precision");
      if (p == "auto") where.precision = QFX_PREC_AUTO;
      else if (p == "ffma") where.precision = QFX_PREC_FFMA;
      else if (p == "bf16x3") where.precision = QFX_PREC_BF16X3;
      else throw std::runtime_error("Unknown precision: " + p);
    }

    if (qflex::global::verbose > 0)
      std::cerr << WARN_MSG("Maximum allowed memory: ",
                            qflex::utils::readable_memory_string(
                                (double)qflex::global::memory_limit))
                << std::endl;

    input.circuit.load(std::ifstream(circuit_filename));
    input.ordering.load(std::ifstream(ordering_filename));
    input.grid.load(grid_filename);

    std::vector<std::pair<std::string, std::complex<double>>> amplitudes;
    try {
      amplitudes = qflex::EvaluateCircuit(&input, where, nullptr);
    } catch (const std::string& err_msg) {
      throw ERROR_MSG("Failed to call EvaluateCircuit(). Error:\n\t[", err_msg, "]");
    }
    for (const auto& [state, amplitude] : amplitudes)
      std::cout << input.initial_state << " --> " << state << ": "
                << std::real(amplitude) << " " << std::imag(amplitude) << std::endl;
  } catch (const std::exception& ex) {
    std::cerr << ex.what() << std::endl;
    return 1;
  } catch (const std::string& msg) {
    std::cerr << msg << std::endl;
    return 2;
  }
  return 0;
}
