// Forwarding header: the schedule IR (ContractionOperation, ExpandPatch, CutIndex,
// MergePatches, ordering_data_to_contraction_ordering, ValidateOrdering, index_name;
// reference src/contraction_utils.h:28-204) lives in schedule.h; what replaces
// ContractionData is the planner + Session of plan.h.
#ifndef QFLEX_HOST_FWD_CONTRACTION_UTILS_H_
#define QFLEX_HOST_FWD_CONTRACTION_UTILS_H_
#include "plan.h"
#include "schedule.h"
#endif
