// Contraction planner: symbolic replay of network build + flatten + schedule
// into a flat list of device steps (plan on host, execute on device).
//
// What it replaces in the reference, and how:
//   circuit_data_to_tensor_network (src/read_circuit.cpp:296-544)  -> gate
//       tensors with the same names/dims/data, uploaded once;
//   flatten_grid_of_tensors (src/read_circuit.cpp:548-689)         -> the same
//       world-line multiply chain per site (identical operand pairs, identical
//       left++right result order), then ONE permute per site that lands the
//       final [ (o), cut bonds, bonds in contraction order ] layout that the
//       reference reaches through bundle_between + reorder;
//   ContractionData::Initialize (src/contraction_utils.cpp:16-221)  -> arena
//       layout (one slot per patch + one spare per distinct patch size; no
//       reorder scratch and no cut copies are needed);
//   ContractionData::ContractGrid (src/contraction_utils.cpp:223-354) -> steps
//       tagged with their cut level; projections are operand base offsets,
//       patch seeding / merge-into-empty are aliases.
// Every tensor is described by named indices with power-of-two dimensions and
// explicit bit positions, so reorders never run as separate passes.
#ifndef QFLEX_HOST_PLAN_H_
#define QFLEX_HOST_PLAN_H_

#include <complex>
#include <cstdint>
#include <list>
#include <map>
#include <string>
#include <vector>

#include "gates.h"
#include "input.h"
#include "qflex_b200.h"
#include "schedule.h"

namespace qflex {

struct SymIndex {
  std::string name;
  int bits = 0;  // log2(dimension)
  int pos = 0;   // position of the index' lowest bit in the linear offset
};

// Slice-dependent part of an operand's base: value(cut) << shift.
// offset += ((value of cut `cut` >> from) & (2^width - 1)) << shift
struct VarTerm {
  int cut = 0;
  int shift = 0;
  int from = 0;    // cut-aware flatten: one virtual leg is a bit field of its bond's value
  int width = 64;
};

struct SymTensor {
  std::vector<SymIndex> idx;  // slowest-varying first, as in the reference
  uint64_t off = 0;           // arena offset (elements)
  std::vector<VarTerm> var;
  int span_bits = 0;          // log2 of the storage span addressed by `idx`+cuts

  int rank_bits() const {
    int b = 0;
    for (const auto& i : idx) b += i.bits;
    return b;
  }
  uint64_t size() const { return 1ull << rank_bits(); }
  int find(const std::string& name) const {
    for (std::size_t i = 0; i < idx.size(); ++i)
      if (idx[i].name == name) return (int)i;
    return -1;
  }
  // dense row-major positions from the index order
  void make_dense() {
    int p = 0;
    for (std::size_t i = idx.size(); i--;) {
      idx[i].pos = p;
      p += idx[i].bits;
    }
    span_bits = p;
  }
  std::vector<std::string> names() const {
    std::vector<std::string> n;
    for (const auto& i : idx) n.push_back(i.name);
    return n;
  }
};

struct PlanStep {
  enum Kind { CONTRACT, DOT_ACCUMULATE, SCALAR_ACCUMULATE, TENSOR_ACCUMULATE, PERMUTE, PROGRAM };
  Kind kind = CONTRACT;
  int level = 0;  // number of cuts that precede this step in the ordering
  int chain = -1; // flatten: world-line (grid site) this step belongs to
  qfx_contract_desc desc{};
  std::vector<VarTerm> var_a, var_b;
  // PERMUTE
  uint64_t src = 0, dst = 0;
  int n_bits = 0;
  int8_t dst_pos[QFX_MAX_BITS] = {0};
  // SCALAR_ACCUMULATE / TENSOR_ACCUMULATE (n_bits open output bits at scalar_off;
  // bit b of the element index moves the amplitude slot by slot_stride[b])
  uint64_t scalar_off = 0;
  std::vector<VarTerm> var_scalar;
  uint32_t slot_stride[QFX_MAX_BITS] = {0};
  // PROGRAM: index into Plan::programs (n_bits = its number of steps)
  int program = -1;
  // bookkeeping
  double flops = 0, bytes = 0;
  std::string label;
};

struct PlanCut {
  std::vector<std::size_t> values;
  bool terminal = false;
  std::string index;
  // terminal: value number v of this cut moves the amplitude slot by v * amp_stride
  // (mixed radix over the terminal cuts in ordering order, first = most significant:
  // output_index * |values| + i of src/contraction_utils.cpp:300-314)
  uint64_t amp_stride = 0;
  // bond cut: the low `fused_bits` bits of the bond are not sliced (BuildPlan's
  // fuse_bits); `values` then holds value >> fused_bits of every aligned block
  int fused_bits = 0;
};

// Chains of small contractions kept resident on the device (qfx_program_create) and run
// in one launch per slice with the slice's cut values (qfx_program_run): the small
// world-line steps of the grid tensors that are flattened inside the slice loop.
struct ChainProgram {
  std::vector<qfx_contract_desc> descs;
  std::vector<int32_t> chain_first;   // [n_chains + 1]
  std::vector<qfx_var_term> terms;
  std::vector<int32_t> term_first;    // [2 * n_steps + 1]: a-terms, b-terms per step
};

struct PlanUpload {
  uint64_t off = 0;
  std::vector<s_type> data;
};

struct Plan {
  std::vector<PlanUpload> uploads;       // gate tensors
  std::vector<PlanStep> flatten_steps;   // network -> one tensor per site
  std::vector<PlanStep> steps;           // schedule, ordered by level
  std::vector<PlanCut> cuts;
  std::vector<ChainProgram> programs;
  uint64_t arena_elems = 0;
  uint64_t n_amplitudes = 1;
  // final layout of every site tensor (for tests / parity with
  // read_circuit_test.cpp:281-345)
  std::vector<std::vector<SymTensor>> site;
  double flops_per_pass = 0, bytes_per_pass = 0;  // one execution of every step
  int precision = QFX_PREC_AUTO;
  // flatten: every site has its own ping-pong buffers, so the world-line chains are
  // independent and run as one qfx_contract_chains launch
  bool flatten_parallel = false;
  int n_chains = 0;
  // terminal cuts kept as open indices of the final tensor (see BuildPlan)
  int open_outputs = 0;
  // cut-aware flatten: the device computes amplitudes * 2^amp_scale_log2
  int amp_scale_log2 = 0;
  // Per active qubit (in the order of the state strings): index into `uploads` of the
  // delta tensor that closes its world line, or -1 when the qubit has a terminal cut
  // (open output).  Lets a resident circuit move to another final state by rewriting
  // two floats per qubit (Session::SetFinalState).
  std::vector<int> final_delta_upload;
  // sum of PlanCut::fused_bits: one leaf of this plan covers 2^fused_bits slices of
  // the ordering as written
  int fused_bits = 0;
};

struct ContractionOperation;
// Phase-1 results of BuildPlan kept between calls on the SAME (input, ordering,
// final_qubits); see BuildPlan.
struct PlanCache {
  bool valid = false;
  bool cut_aware = false;
  std::vector<const ContractionOperation*> key;
  Plan plan;  // uploads, flatten steps, site layouts
  std::map<std::string, int> pre_bond, pre_width;
  std::vector<PlanStep> dyn_steps;
  uint64_t work_base = 0, work_flatten = 0;
};

// Builds the full plan.  `final_qubits` = qubits with a terminal cut
// (EvaluateCircuit passes get_output_states' list).  Throws std::string.
//
// `open_outputs` = how many terminal cuts (counted from the end of the ordering) stay
// open: such a cut, when it runs over the full value range, is not projected value by
// value (the reference's 2^k leaf contractions per slice, src/contraction_utils.cpp:
// 300-314); its output index stays open, rides through the remaining expands and
// merges as one more free bit, and the schedule ends in a tensor of 2^k amplitudes
// that one TENSOR_ACCUMULATE step adds into the table.  Same amplitudes, same flops,
// 2^k times fewer launches after the cut, 2^k times larger patches after it -- the
// caller opens fewer cuts (down to none: the reference's plan) when that does not fit
// the memory limit.
//
// `cut_aware_flatten`: the bond cuts that open the ordering are applied
// to the GATE tensors -- every virtual leg of a cut bond is projected on its bit
// field of the cut value -- so a grid tensor next to a cut never carries that bond.
// Such sites are flattened inside the slice loop (their world-line products become
// ordinary steps after the leading cuts) instead of once; grid tensors that would
// not fit (Sycamore m >= 16: four bonds of 2^8..2^10) become representable.
//
// `fuse_bits`: how many bits of bond-cut slicing are undone (innermost cut first): the
// low j bits of a cut bond whose value list consists of aligned blocks of 2^j values
// stay an ordinary bond of its two grid tensors and are contracted when their patches
// meet, so 2^j slices of the ordering run as ONE pass of larger kernels (the batched
// form of the slice loop, src/contraction_utils.cpp:264-316) and steps that do not
// carry the bond run once instead of 2^j times.  Patches that carry it grow 2^j-fold;
// the caller hands out as many bits as fit the memory limit.
Plan BuildPlan(const QflexInput& input,
               const std::list<ContractionOperation>& ordering,
               const std::vector<std::vector<std::size_t>>& final_qubits,
               int open_outputs = 0, bool cut_aware_flatten = false, int fuse_bits = 0,
               PlanCache* cache = nullptr);

struct ExecStats {
  double seconds_upload = 0, seconds_flatten = 0, seconds_contract = 0;
  double device_ms_contract = 0;  // CUDA events around the schedule's launches
  uint64_t leaves = 0;            // scalar results accumulated (slices x outputs)
  qfx_stats device{};
};

// One plan resident on one GPU: the constructor reserves the arena, uploads the
// gate tensors and flattens the network into grid tensors; Contract() then runs
// the schedule (any number of times).  Only the cut-slices dealt to `rank` of
// `world` are executed (slice s -> rank s % world); partial amplitudes must be
// summed across ranks by the caller.  Everything throws std::string.
class Session {
 public:
  Session(Plan plan, int device);
  // The same plan resident on SEVERAL GPUs of this process (the reference's
  // EvaluateCircuit and CLI are single-process, src/main.cpp:129): every device gets
  // its own arena(s) and flattened network, one host thread per lane drives it, the
  // slices of this rank are dealt over (lane, device) and the partial amplitude tables
  // are summed on the host in a fixed order -- 16 bytes per amplitude, no collective.
  Session(Plan plan, std::vector<int> devices);
  ~Session();
  Session(const Session&) = delete;
  Session& operator=(const Session&) = delete;

  void Upload();   // host -> device copy of the gate tensors
  void Flatten();  // network -> grid tensors, on the device
  // Another output bitstring on the same resident circuit: rewrites the delta tensors
  // that close the world lines (positions of qubits with a terminal cut are ignored),
  // re-uploads the gate tensors and flattens again -- parse, plan, arena and the
  // contraction schedule are reused.  What a batch of final states needs
  // (src/pybind_main.cpp:49-73,111-113 parses such a list and rejects it).
  void SetFinalState(const std::string& final_state);
  // A projected cut with a single value moves to another value (resident final states:
  // the projection of an output leg; see DevicePlacement::resident_finals).
  void SetCutValue(int cut, std::size_t value);
  void Contract(int rank, int world, std::vector<std::complex<double>>* amplitudes);

  const Plan& plan() const { return plan_; }
  int lanes() const { return (int)lanes_.size(); }
  int devices() const { return (int)devices_.size(); }
  const ExecStats& stats() const { return stats_; }

  // Per-launch device timing of the next Contract() calls, by shape class
  // (reference analogue: the verbose>1 [M]/[R] lines, src/tensor.cpp:1110-1164).
  void set_profile(bool on) { profile_ = on; }
  const std::vector<qfx_profile_row>& profile_rows() const { return profile_rows_; }

 private:
  // Upload + flatten into one lane's arena; one lane's share of Contract().
  void PrepareLane(qfx_ctx* ctx);
  void ReleaseDevice();
  void ContractLane(int lane, int rank, int world, std::vector<double>* amps,
                    ExecStats* stats, std::vector<qfx_profile_row>* rows);

  Plan plan_;
  qfx_ctx* ctx_ = nullptr;          // lane 0
  // Lanes: extra contexts (own arena + stream) on the SAME device.  Small networks
  // (arena of MBs, contractions of microseconds) leave most SMs idle and are bound
  // by launch latency; their cut slices are independent, so the slices of this rank
  // are dealt once more over the lanes (lane l acts as rank l*world+rank of world*L:
  // lane-major, so that few slices still reach every rank),
  // each driven by its own host thread, and kernels of different slices overlap on
  // the GPU.  One lane when the arena is large (the kernels fill the GPU).
  std::vector<qfx_ctx*> lanes_;  // [lane * n_devices + device index]
  std::vector<std::vector<qfx_program*>> programs_;  // [lane][program of the plan]
  std::vector<int> devices_;
  ExecStats stats_;
  bool profile_ = false;
  std::vector<qfx_profile_row> profile_rows_;
};

}  // namespace qflex

#endif  // QFLEX_HOST_PLAN_H_
