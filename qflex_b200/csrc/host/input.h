// Text front-end: the three qFlex input formats (circuit, grid, ordering).
// Same public structs and .load() overloads as the reference
// (src/circuit.h:12-36, src/grid.h:13-20, src/ordering.h:13-19, src/input.h:11-17)
// so that callers written against qflex::QflexInput compile unchanged.
// Grammar: SURVEY Appendix C / reference docs/input_formats.md.
#ifndef QFLEX_HOST_INPUT_H_
#define QFLEX_HOST_INPUT_H_

#include <cstddef>
#include <istream>
#include <string>
#include <vector>

namespace qflex {

struct QflexGate {
  std::string name;
  std::size_t cycle{0};
  std::vector<std::size_t> qubits;
  std::vector<double> params;
  std::string raw;
};

struct QflexCircuit {
  std::size_t num_active_qubits{0};
  std::size_t depth{0};
  std::vector<QflexGate> gates;

  void clear();
  void load(std::istream& istream);
  void load(std::istream&& istream);
  void load(const std::string& filename);
};

struct QflexGrid {
  std::size_t I{0}, J{0};
  std::vector<std::vector<std::size_t>> qubits_off;

  void clear();
  void load(std::istream& istream);
  void load(std::istream&& istream);
  void load(const std::string& filename);
};

struct QflexOrdering {
  std::vector<std::string> instructions;

  void clear();
  void load(std::istream& istream);
  void load(std::istream&& istream);
  void load(const std::string& filename);
};

struct QflexInput {
  QflexOrdering ordering;
  QflexCircuit circuit;
  QflexGrid grid;
  std::string initial_state;
  std::string final_state;
};

}  // namespace qflex

#endif  // QFLEX_HOST_INPUT_H_
