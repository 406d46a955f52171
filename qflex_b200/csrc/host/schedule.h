// Contraction schedule: the ordering file parsed into expand / cut / merge
// operations, with the reference's validation rules and index naming.
// Same public names as the reference (src/contraction_utils.h:28-112).
#ifndef QFLEX_HOST_SCHEDULE_H_
#define QFLEX_HOST_SCHEDULE_H_

#include <list>
#include <optional>
#include <string>
#include <vector>

#include "input.h"

namespace qflex {

using GridPos = std::vector<std::size_t>;  // {i, j}

// "(i,j),(k,l)" for a bond, "(i,j,c),(k,l,c)" for a virtual leg, "(i,j),(o)" for
// an open output leg (src/contraction_utils.cpp:477-519).
std::string index_name(const std::vector<std::size_t>& p1,
                       const std::vector<std::size_t>& p2);
std::string index_name(const std::vector<std::vector<std::size_t>>& tensors);

std::vector<std::size_t> get_qubit_coords(std::size_t q, std::size_t J);

bool find_grid_coord_in_list(
    const std::optional<std::vector<std::vector<std::size_t>>>& coord_list,
    const std::size_t i, const std::size_t j);

struct ExpandPatch {
  ExpandPatch() {}
  ExpandPatch(std::string id, std::vector<std::size_t> tensor)
      : id(std::move(id)), tensor(std::move(tensor)) {}
  std::string id;                   // patch being grown
  std::vector<std::size_t> tensor;  // grid tensor contracted into it
};

struct CutIndex {
  CutIndex() {}
  CutIndex(std::vector<std::vector<std::size_t>> tensors,
           std::vector<std::size_t> values = {})
      : tensors(std::move(tensors)), values(std::move(values)) {}
  // Two tensors: bond cut (values add into the same amplitude).  One tensor:
  // terminal cut (each value is a different output bitstring).
  std::vector<std::vector<std::size_t>> tensors;
  std::vector<std::size_t> values;  // empty = every value of the index
};

struct MergePatches {
  MergePatches() {}
  MergePatches(std::string source_id, std::string target_id)
      : source_id(std::move(source_id)), target_id(std::move(target_id)) {}
  std::string source_id;
  std::string target_id;  // the merged patch keeps this id
};

struct ContractionOperation {
  ContractionOperation(ExpandPatch e) : op_type(EXPAND), expand(std::move(e)) {}
  ContractionOperation(CutIndex c) : op_type(CUT), cut(std::move(c)) {}
  ContractionOperation(MergePatches m) : op_type(MERGE), merge(std::move(m)) {}
  enum OpType { EXPAND, CUT, MERGE };
  OpType op_type;
  ExpandPatch expand;
  CutIndex cut;
  MergePatches merge;
};

// Ordering text -> operations (+ ValidateOrdering).
// Reference: src/contraction_utils.cpp:358-475.
void ordering_data_to_contraction_ordering(
    const QflexInput& input, std::list<ContractionOperation>* ordering);

// Throws a std::string listing every violated rule
// (src/contraction_utils.h:189-204, src/contraction_utils.cpp:534-610).
void ValidateOrdering(const std::list<ContractionOperation>& ordering);

}  // namespace qflex

#endif  // QFLEX_HOST_SCHEDULE_H_
