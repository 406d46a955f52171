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

void ordering_data_to_contraction_ordering(
    const QflexInput& input, std::list<ContractionOperation>* ordering) {
  if (ordering == nullptr) throw ERROR_MSG("Ordering must be non-null.");

  const std::size_t grid_size = input.grid.I * input.grid.J;
  std::string current;
  auto fail = [&current](const std::string& why) {
    return ERROR_MSG("Parsing failed on line: '", current, "' with error: ", why);
  };
  auto check_index = [&](long int index) {
    if (static_cast<std::size_t>(static_cast<long int>(grid_size)) != grid_size)
      throw fail("Grid is too big.");
    if (index < 0 || index >= static_cast<long int>(grid_size))
      throw fail(utils::concat("Index ", index, " must be within grid boundaries."));
  };
  auto is_off = [&](const std::vector<std::size_t>& pos) {
    return find_grid_coord_in_list(input.grid.qubits_off, pos[0], pos[1]);
  };

  for (const auto& line : input.ordering.instructions) {
    if (line.empty() || line[0] == '#') continue;
    current = line;
    std::stringstream ss(line);
    std::string operation;
    ss >> operation;
    if (operation == "expand") {
      std::string patch;
      long int index = 0;
      ss >> patch >> index;
      check_index(index);
      const auto position = get_qubit_coords(index, input.grid.J);
      if (is_off(position)) throw fail("Index must specify an active qubit.");
      ordering->emplace_back(ExpandPatch(patch, position));
    } else if (operation == "cut") {
      std::string values_str;
      std::vector<std::size_t> values;
      ss >> values_str;
      if (!parse_cut_values(values_str, &values))
        throw fail("Cut values must be comma-separated ints, e.g. (0,1,3).");
      long int index_1 = 0, index_2 = 0;
      ss >> index_1;
      check_index(index_1);
      const auto position_1 = get_qubit_coords(index_1, input.grid.J);
      if (is_off(position_1)) throw fail("Index 1 must specify an active qubit.");
      if (ss.eof()) {
        ordering->emplace_back(CutIndex({position_1}, values));
      } else {
        ss >> index_2;
        check_index(index_2);
        const auto position_2 = get_qubit_coords(index_2, input.grid.J);
        if (is_off(position_2)) throw fail("Index 2 must specify an active qubit.");
        // the lower qubit id always comes first in a bond name
        if (index_1 < index_2)
          ordering->emplace_back(CutIndex({position_1, position_2}, values));
        else
          ordering->emplace_back(CutIndex({position_2, position_1}, values));
      }
    } else if (operation == "merge") {
      std::string patch_1, patch_2;
      ss >> patch_1 >> patch_2;
      ordering->emplace_back(MergePatches(patch_1, patch_2));
    } else {
      throw fail("Received an invalid operation in config.");
    }
  }
  ValidateOrdering(*ordering);
}

void ValidateOrdering(const std::list<ContractionOperation>& ordering) {
  struct PatchState {
    bool is_active = false;  // expanded, no cut since
    bool is_used = false;    // expanded and a cut happened afterwards
    bool is_merged = false;  // merged into another patch
  };
  std::unordered_map<std::string, PatchState> patches;
  std::unordered_set<std::string> cut_indices, used_tensors;
  std::string problems;
  for (const auto& op : ordering) {
    switch (op.op_type) {
      case ContractionOperation::EXPAND: {
        PatchState& p = patches[op.expand.id];
        if (p.is_used)
          problems = utils::concat(problems, "\nTensor at (", op.expand.tensor[0],
                                   ",", op.expand.tensor[1],
                                   ") is added to non-empty patch ",
                                   op.expand.id.c_str(), " after a cut.");
        if (p.is_merged)
          problems = utils::concat(problems, "\nTensor at (", op.expand.tensor[0],
                                   ",", op.expand.tensor[1],
                                   ") is added to non-empty patch ",
                                   op.expand.id.c_str(), " after a cut.");
        const std::string name =
            utils::concat("(", op.expand.tensor[0], ",", op.expand.tensor[1], ")");
        if (!used_tensors.insert(name).second)
          problems = utils::concat(problems, "\nTensor ", name,
                                   " is contracted multiple times.");
        p.is_active = true;
        break;
      }
      case ContractionOperation::CUT: {
        try {
          for (auto& kv : patches)
            if (kv.second.is_active) kv.second.is_used = true;
          const std::string index = index_name(op.cut.tensors);
          if (!cut_indices.insert(index).second)
            problems = utils::concat(problems, "\nIndex ", index.c_str(),
                                     " is cut multiple times.");
        } catch (const std::string& err_msg) {
          throw ERROR_MSG("Failed to during CUT. Error:\n\t[", err_msg, "]");
        }
        break;
      }
      case ContractionOperation::MERGE: {
        if (patches[op.merge.source_id].is_merged)
          problems = utils::concat(problems, "\nPatch ", op.merge.source_id.c_str(),
                                   " is merged multiple times.");
        if (patches[op.merge.target_id].is_used)
          problems = utils::concat(problems, "\nPatch ", op.merge.source_id.c_str(),
                                   " is merged into non-empty patch ",
                                   op.merge.target_id.c_str(), " after a cut.");
        patches[op.merge.source_id].is_merged = true;
        patches[op.merge.target_id].is_active = true;
        break;
      }
    }
  }
  if (!problems.empty()) throw problems;
}

}  // namespace qflex
