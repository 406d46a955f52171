// Ordering file -> contraction operations, validation, index naming.
// Behaviour: reference src/contraction_utils.cpp:358-610.
#include "schedule.h"

#include <algorithm>
#include <cstdlib>
#include <sstream>
#include <unordered_map>
#include <unordered_set>

#include "errors.h"

namespace qflex {

std::string index_name(const std::vector<std::size_t>& p1,
                       const std::vector<std::size_t>& p2) {
  if (p1.size() == 2 && p2.size() == 2)
    return utils::concat("(", p1[0], ",", p1[1], "),(", p2[0], ",", p2[1], ")");
  if (p1.size() == 3 && p2.size() == 3)
    return utils::concat("(", p1[0], ",", p1[1], ",", p1[2], "),(", p2[0], ",",
                         p2[1], ",", p2[2], ")");
  if (p1.size() == 2 && p2.empty())
    return utils::concat("(", p1[0], ",", p1[1], "),(o)");
  std::stringstream ss;
  ss << "Failed to construct tensor name with the following vectors: p1 = [";
  for (const auto& p : p1) ss << p << ',';
  ss << "] and p2 = [";
  for (const auto& p : p2) ss << p << ',';
  ss << "].";
  throw ERROR_MSG(ss.str());
}

std::string index_name(const std::vector<std::vector<std::size_t>>& tensors) {
  try {
    if (tensors.size() == 2) return index_name(tensors[0], tensors[1]);
    if (tensors.size() == 1) return index_name(tensors[0], {});
  } catch (const std::string& err_msg) {
    throw ERROR_MSG("Failed to call index_name(). Error:\n\t[", err_msg, "]");
  }
  throw ERROR_MSG("Failed to construct tensor name with input tensors size: ",
                  tensors.size(), ".");
}

std::vector<std::size_t> get_qubit_coords(std::size_t q, std::size_t J) {
  return {q / J, q % J};
}

bool find_grid_coord_in_list(
    const std::optional<std::vector<std::vector<std::size_t>>>& coord_list,
    const std::size_t i, const std::size_t j) {
  if (!coord_list.has_value()) return false;
  for (const auto& c : coord_list.value())
    if (c.size() == 2 && c[0] == i && c[1] == j) return true;
  return false;
}

namespace {

// "(0,1,3)" -> {0,1,3}; "()" -> {}.  False when the token is malformed.
bool parse_cut_values(const std::string& token, std::vector<std::size_t>* out) {
  if (token.size() < 2 || token.front() != '(' || token.back() != ')') return false;
  for (std::size_t i = 1; i + 1 < token.size(); ++i)
    if (!(token[i] == ',' || (token[i] >= '0' && token[i] <= '9'))) return false;
  std::string cur;
  const std::string inner = token.substr(1, token.size() - 2);
  if (inner.empty()) return true;
  std::size_t start = 0;
  while (true) {
    const std::size_t comma = inner.find(',', start);
    const std::string piece = inner.substr(
        start, comma == std::string::npos ? std::string::npos : comma - start);
    out->push_back(std::atoi(piece.c_str()));
    if (comma == std::string::npos) break;
    start = comma + 1;
  }
  return true;
}

}  // namespace

namespace {

// One line of the ordering grammar:
//   expand <patch> <qubit> | cut (<v,..>) <qubit> [<qubit>] | merge <patch> <patch>
// tokenised once; qubit ids are resolved against the grid by the caller.
struct OrderingLine {
  std::string verb;
  std::vector<std::string> words;  // everything after the verb
};

OrderingLine tokenize(const std::string& line) {
  OrderingLine out;
  std::stringstream ss(line);
  ss >> out.verb;
  for (std::string w; ss >> w;) out.words.push_back(w);
  return out;
}

}  // namespace

void ordering_data_to_contraction_ordering(
    const QflexInput& input, std::list<ContractionOperation>* ordering) {
  if (ordering == nullptr) throw ERROR_MSG("Ordering must be non-null.");

  const std::size_t J = input.grid.J;
  const std::size_t sites = input.grid.I * J;
  std::vector<char> off(sites, 0);
  for (const auto& q : input.grid.qubits_off)
    if (q.size() == 2 && q[0] < input.grid.I && q[1] < J) off[q[0] * J + q[1]] = 1;

  for (const std::string& line : input.ordering.instructions) {
    if (line.empty() || line[0] == '#') continue;
    auto fail = [&line](const std::string& why) {
      return ERROR_MSG("Parsing failed on line: '", line, "' with error: ", why);
    };
    // qubit id (a word of the line) -> grid position; `which` names it in messages
    auto site_of = [&](const std::string& word, const char* which) {
      char* end = nullptr;
      const long id = std::strtol(word.c_str(), &end, 10);
      if (static_cast<std::size_t>(static_cast<long>(sites)) != sites) throw fail("Grid is too big.");
      if (id < 0 || id >= static_cast<long>(sites))  // (a non-number reads as 0, as >> does)
        throw fail(utils::concat("Index ", id, " must be within grid boundaries."));
      if (off[id]) throw fail(utils::concat(which, " must specify an active qubit."));
      return get_qubit_coords((std::size_t)id, J);
    };
    const OrderingLine t = tokenize(line);
    const auto word = [&](std::size_t i) { return i < t.words.size() ? t.words[i] : std::string("0"); };
    if (t.verb == "expand") {
      ordering->emplace_back(ExpandPatch(word(0), site_of(word(1), "Index")));
    } else if (t.verb == "merge") {
      ordering->emplace_back(MergePatches(t.words.size() > 0 ? t.words[0] : "",
                                          t.words.size() > 1 ? t.words[1] : ""));
    } else if (t.verb == "cut") {
      std::vector<std::size_t> values;
      if (!parse_cut_values(t.words.empty() ? "" : t.words[0], &values))
        throw fail("Cut values must be comma-separated ints, e.g. (0,1,3).");
      const auto first = site_of(word(1), "Index 1");
      if (t.words.size() < 3) {
        ordering->emplace_back(CutIndex({first}, values));  // terminal cut
      } else {
        const auto second = site_of(word(2), "Index 2");
        // a bond is named lower qubit id first
        const bool in_order = first[0] * J + first[1] < second[0] * J + second[1];
        ordering->emplace_back(in_order ? CutIndex({first, second}, values)
                                        : CutIndex({second, first}, values));
      }
    } else {
      throw fail("Received an invalid operation in config.");
    }
  }
  ValidateOrdering(*ordering);
}

// The rules of src/contraction_utils.h:189-204, stated on two clocks: `epoch` counts
// the cuts seen so far; a patch remembers the epoch in which it first held a tensor.
// A patch that held something before the latest cut is frozen -- it was contracted
// inside an outer slice loop and may not grow inside an inner one -- and so is a patch
// that has been merged away.  Tensors are expanded once, indices cut once, patches
// merged away once.  All violations are collected into one message.
void ValidateOrdering(const std::list<ContractionOperation>& ordering) {
  constexpr long kEmpty = -1;
  struct PatchLife {
    long born = kEmpty;   // epoch of the first expand / merge into this patch
    bool gone = false;    // merged into another patch
  };
  std::unordered_map<std::string, PatchLife> life;
  std::unordered_set<std::string> seen_tensor, seen_cut;
  std::vector<std::string> complaints;
  long epoch = 0;
  auto frozen = [&](const PatchLife& p) { return p.born != kEmpty && p.born < epoch; };

  for (const auto& op : ordering) {
    if (op.op_type == ContractionOperation::CUT) {
      std::string index;
      try {
        index = index_name(op.cut.tensors);
      } catch (const std::string& err_msg) {
        throw ERROR_MSG("Failed to during CUT. Error:\n\t[", err_msg, "]");
      }
      if (!seen_cut.insert(index).second)
        complaints.push_back(utils::concat("Index ", index, " is cut multiple times."));
      ++epoch;
    } else if (op.op_type == ContractionOperation::EXPAND) {
      PatchLife& p = life[op.expand.id];
      const auto& at = op.expand.tensor;
      if (frozen(p) || p.gone)
        complaints.push_back(utils::concat("Tensor at (", at[0], ",", at[1],
                                           ") is added to non-empty patch ", op.expand.id,
                                           " after a cut."));
      const std::string name = utils::concat("(", at[0], ",", at[1], ")");
      if (!seen_tensor.insert(name).second)
        complaints.push_back(utils::concat("Tensor ", name, " is contracted multiple times."));
      if (p.born == kEmpty) p.born = epoch;
    } else {
      PatchLife& from = life[op.merge.source_id];
      if (from.gone)
        complaints.push_back(
            utils::concat("Patch ", op.merge.source_id, " is merged multiple times."));
      from.gone = true;
      PatchLife& into = life[op.merge.target_id];  // (may rehash: `from` is not used below)
      if (frozen(into))
        complaints.push_back(utils::concat("Patch ", op.merge.source_id,
                                           " is merged into non-empty patch ",
                                           op.merge.target_id, " after a cut."));
      if (into.born == kEmpty) into.born = epoch;
    }
  }
  if (complaints.empty()) return;
  std::string all;
  for (const auto& c : complaints) all += "\n" + c;
  throw all;
}

}  // namespace qflex
