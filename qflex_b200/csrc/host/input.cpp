// Parsers for the circuit / grid / ordering text formats.
// Behaviour follows the reference loaders (src/circuit.cpp:27-206,
// src/grid.cpp:11-45, src/ordering.cpp:10-53); the line cleaner is a
// hand-written scanner instead of a chain of std::regex replacements.
#include "input.h"

#include <algorithm>
#include <fstream>
#include <unordered_map>
#include <unordered_set>

#include "errors.h"

namespace qflex {
namespace {

bool is_space(char c) {
  return c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '\f' || c == '\v';
}

bool is_kept(char c) {
  if (c == '(' || c == ')' || is_space(c)) return true;
  if ((c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z') || (c >= '0' && c <= '9'))
    return true;
  return c == '_' || c == '.' || c == ',' || c == '-';
}

// Normalises one input line: drops comments and foreign characters, collapses
// blanks, trims the right end, glues "name (" (circuit files only) and removes
// blanks inside parentheses.
std::string clean_line(const std::string& raw, bool glue_open_paren) {
  std::string s;
  for (char c : raw) {
    if (c == '#') break;
    if (!is_kept(c)) continue;
    s.push_back(c == '\t' ? ' ' : c);
  }
  // runs of two or more blanks -> one space
  std::string t;
  for (std::size_t i = 0; i < s.size();) {
    if (!is_space(s[i])) {
      t.push_back(s[i++]);
      continue;
    }
    std::size_t j = i;
    while (j < s.size() && is_space(s[j])) ++j;
    if (j - i >= 2) t.push_back(' ');
    else t.push_back(s[i]);
    i = j;
  }
  while (!t.empty() && is_space(t.back())) t.pop_back();
  // blanks before '(' and blanks whose next parenthesis is ')'
  std::string out;
  for (std::size_t i = 0; i < t.size();) {
    if (!is_space(t[i])) {
      out.push_back(t[i++]);
      continue;
    }
    std::size_t j = i;
    while (j < t.size() && is_space(t[j])) ++j;
    bool drop = false;
    if (glue_open_paren && j < t.size() && t[j] == '(') drop = true;
    if (!drop) {
      for (std::size_t p = j; p < t.size(); ++p) {
        if (t[p] == '(') break;
        if (t[p] == ')') {
          drop = true;
          break;
        }
      }
    }
    if (!drop) out.append(t, i, j - i);
    i = j;
  }
  return out;
}

std::vector<std::string> split_blank(const std::string& line) {
  std::vector<std::string> tokens;
  std::string cur;
  for (char c : line) {
    if (is_space(c)) {
      if (!cur.empty()) tokens.push_back(cur);
      cur.clear();
    } else {
      cur.push_back(c);
    }
  }
  if (!cur.empty()) tokens.push_back(cur);
  return tokens;
}

std::vector<std::string> split_comma(const std::string& s) {
  std::vector<std::string> tokens;
  std::string cur;
  for (char c : s) {
    if (c == ',') {
      if (!cur.empty()) tokens.push_back(cur);
      cur.clear();
    } else {
      cur.push_back(c);
    }
  }
  if (!cur.empty()) tokens.push_back(cur);
  return tokens;
}

bool looks_like_number(const std::string& token) {
  try {
    (void)std::stol(token);
  } catch (...) {
    return false;
  }
  return true;
}

bool looks_like_integer(const std::string& token) {
  return looks_like_number(token) && std::stol(token) == std::stod(token);
}

}  // namespace

// ------------------------------------------------------------------ circuit
void QflexCircuit::clear() {
  num_active_qubits = 0;
  depth = 0;
  gates.clear();
}

void QflexCircuit::load(std::istream& istream) { load(std::move(istream)); }

void QflexCircuit::load(std::istream&& istream) {
  clear();
  std::size_t line_no = 0, last_cycle = 0;
  std::unordered_set<std::size_t> busy;
  std::string line;
  auto fail = [&](const std::string& why) {
    return ERROR_MSG("[", std::to_string(line_no + 1), ": ", line, "] ", why);
  };

  for (; std::getline(istream, line); ++line_no) {
    line = clean_line(line, /*glue_open_paren=*/true);
    if (line.empty()) continue;

    const auto n_open = std::count(line.begin(), line.end(), '(');
    const auto n_close = std::count(line.begin(), line.end(), ')');
    if (n_open != n_close || n_open > 1) throw fail("Not-matching parentheses.");

    const auto tokens = split_blank(line);
    if (line_no == 0) {
      if (tokens.size() != 1 || std::stol(tokens[0]) <= 0)
        throw fail("First line in circuit must be the number of active qubits.");
      num_active_qubits = std::stol(tokens[0]);
      continue;
    }
    if (tokens.size() < 3)
      throw fail(
          "Gate must be specified as: cycle gate_name[(p1[,p2,...])] q1 "
          "[q2, ...]");

    gates.emplace_back();
    QflexGate& gate = gates.back();
    gate.raw = line;
    if (!looks_like_integer(tokens[0]))
      throw fail("First token must be a valid cycle number.");
    gate.cycle = std::stol(tokens[0]);
    if (gate.cycle < last_cycle) throw fail("Cycle number can only increase.");
    if (gate.cycle != last_cycle) {
      busy.clear();
      last_cycle = gate.cycle;
    }

    const std::size_t open = tokens[1].find('(');
    if (open == std::string::npos) {
      gate.name = tokens[1];
    } else {
      gate.name = tokens[1].substr(0, open);
      const std::size_t close = tokens[1].find(')');
      for (const auto& p : split_comma(tokens[1].substr(open + 1, close - open - 1))) {
        if (!looks_like_number(p)) throw ERROR_MSG("Params must be valid numbers.");
        gate.params.push_back(std::stod(p));
      }
    }
    for (std::size_t i = 2; i < tokens.size(); ++i) {
      if (!looks_like_integer(tokens[i])) throw fail("Qubit must be a valid number.");
      gate.qubits.push_back(std::stol(tokens[i]));
    }
    for (const auto q : gate.qubits) {
      if (!busy.insert(q).second) throw fail("Qubits can only used once per cycle.");
    }
  }

  // depth = max over qubit pairs of (1 per cz/cx, 2 per fsim)
  std::unordered_map<std::size_t, std::unordered_map<std::size_t, std::size_t>> per_pair;
  for (const auto& gate : gates) {
    if (gate.qubits.size() > 2)
      throw fail("Depth calculation does not handle k-qubit gates with k > 2.");
    if (gate.qubits.size() != 2) continue;
    const std::size_t hi = std::max(gate.qubits[0], gate.qubits[1]);
    const std::size_t lo = std::min(gate.qubits[0], gate.qubits[1]);
    std::size_t& d = per_pair[hi][lo];
    d += (gate.name == "fsim") ? 2 : 1;
    depth = std::max(depth, d);
  }
}

void QflexCircuit::load(const std::string& filename) {
  std::ifstream in(filename);
  if (!in.good()) throw ERROR_MSG("Cannot open circuit file ", filename, ".");
  load(in);
}

// --------------------------------------------------------------------- grid
void QflexGrid::clear() {
  I = 0;
  J = 0;
  qubits_off.clear();
}

void QflexGrid::load(std::istream& istream) { load(std::move(istream)); }

void QflexGrid::load(std::istream&& istream) {
  clear();
  std::string line;
  std::size_t line_no = 0;
  while (std::getline(istream, line)) {
    if (line.empty() || line[0] == '#') continue;
    ++line_no;
    std::string row;
    for (char c : line)
      if (c == '0' || c == '1') row.push_back(c);
    if (row.empty()) continue;
    if (J != 0 && J != row.size())
      throw ERROR_MSG("Grid size at line ", line_no,
                      " is inconsistent with a width of ", row.size(),
                      " instead of ", J, ".");
    J = row.size();
    for (std::size_t q = 0; q < J; ++q)
      if (row[q] == '0') qubits_off.push_back({I, q});
    ++I;
  }
}

void QflexGrid::load(const std::string& filename) {
  std::ifstream in(filename);
  if (!in.good()) throw ERROR_MSG("Cannot open grid file: ", filename, ".");
  load(in);
}

// ----------------------------------------------------------------- ordering
void QflexOrdering::clear() { instructions.clear(); }

void QflexOrdering::load(std::istream& istream) { load(std::move(istream)); }

void QflexOrdering::load(std::istream&& istream) {
  clear();
  std::string line;
  while (std::getline(istream, line)) {
    line = clean_line(line, /*glue_open_paren=*/false);
    if (!line.empty()) instructions.push_back(line);
  }
}

void QflexOrdering::load(const std::string& filename) {
  std::ifstream in(filename);
  if (!in.good()) throw std::string("Cannot open ordering file: ") + filename;
  load(in);
}

// -------------------------------------------------------------------- utils
namespace utils {

std::size_t from_readable_memory_string(std::string memory) {
  std::string s;
  for (char c : memory) {
    if (!is_kept(c) || is_space(c)) continue;
    s.push_back((char)std::toupper((unsigned char)c));
  }
  const std::size_t cut = std::min(s.find_first_not_of("0123456789"), s.size());
  std::size_t m = std::stoll(s.substr(0, cut));
  const std::string suffix = s.substr(cut);
  if (suffix == "GB") m <<= 30;
  else if (suffix == "MB") m <<= 20;
  else if (suffix == "KB") m <<= 10;
  else if (suffix == "B" || suffix.empty()) m <<= 0;
  else throw ERROR_MSG("Suffix ", suffix, " not recognized.");
  return m;
}

std::string readable_memory_string(double memory) {
  static const char* const suffix[] = {" B", " kB", " MB", " GB"};
  std::size_t scale = 0;
  while (memory >= 1024.0 && scale < 3) {
    ++scale;
    memory /= 1024.0;
  }
  return concat(static_cast<std::size_t>(memory * 100) / 100., suffix[scale]);
}

}  // namespace utils
}  // namespace qflex
