// Planner: circuit + grid + schedule -> device steps.  See plan.h.
#include "plan.h"

#include <algorithm>
#include <cstdlib>
#include <functional>
#include <cstring>
#include <map>
#include <set>
#include <unordered_map>
#include <unordered_set>

#include "errors.h"

namespace qflex {
namespace {

constexpr uint64_t kAlignElems = 32;  // 256-byte alignment of every tensor

uint64_t align_up(uint64_t x) { return (x + kAlignElems - 1) & ~(kAlignElems - 1); }

int log2_exact(std::size_t dim) {
  int b = 0;
  while ((std::size_t(1) << b) < dim) ++b;
  if ((std::size_t(1) << b) != dim)
    throw ERROR_MSG("Dimension ", dim,
                    " is not a power of two; the B200 engine addresses tensors "
                    "by index bits.");
  return b;
}

std::string tensor_to_string(const SymTensor& t) {
  std::string s = "Tensor of rank " + std::to_string(t.idx.size()) + ", with indices {";
  for (std::size_t i = 0; i < t.idx.size(); ++i) {
    if (i) s += ", ";
    s += t.idx[i].name;
  }
  s += "} and dimensions {";
  for (std::size_t i = 0; i < t.idx.size(); ++i) {
    if (i) s += ", ";
    s += std::to_string(std::size_t(1) << t.idx[i].bits);
  }
  return s + "}.";
}

// C = A * B over every shared index; C indices = (A only, in A order) ++ (B only,
// in B order), dense at c_off (reference src/tensor.cpp:1065-1087,1167-1182).
SymTensor make_contract(const SymTensor& A, const SymTensor& B, uint64_t c_off,
                        PlanStep* step) {
  std::vector<const SymIndex*> left, right, common_a, common_b;
  for (const auto& ia : A.idx) {
    const int jb = B.find(ia.name);
    if (jb < 0) {
      left.push_back(&ia);
    } else {
      if (B.idx[jb].bits != ia.bits)
        throw ERROR_MSG("Common indices must have matching dimensions, but at index: ",
                        common_a.size(), ", the dimensions are A: ",
                        (std::size_t(1) << ia.bits), ", B: ",
                        (std::size_t(1) << B.idx[jb].bits), ".");
      common_a.push_back(&ia);
      common_b.push_back(&B.idx[jb]);
    }
  }
  for (const auto& ib : B.idx)
    if (A.find(ib.name) < 0) right.push_back(&ib);

  SymTensor C;
  for (const auto* l : left) C.idx.push_back(*l);
  for (const auto* r : right) C.idx.push_back(*r);
  C.off = c_off;
  C.make_dense();

  qfx_contract_desc& d = step->desc;
  std::memset(&d, 0, sizeof(d));
  d.off_a = A.off;
  d.off_b = B.off;
  d.off_c = c_off;
  int m = 0, n = 0, k = 0;
  for (std::size_t li = left.size(); li--;) {
    const SymIndex& ci = C.idx[li];
    for (int b = 0; b < left[li]->bits; ++b, ++m) {
      d.a_m[m] = (int8_t)(left[li]->pos + b);
      d.c_m[m] = (int8_t)(ci.pos + b);
    }
  }
  for (std::size_t ri = right.size(); ri--;) {
    const SymIndex& ci = C.idx[left.size() + ri];
    for (int b = 0; b < right[ri]->bits; ++b, ++n) {
      d.b_n[n] = (int8_t)(right[ri]->pos + b);
      d.c_n[n] = (int8_t)(ci.pos + b);
    }
  }
  for (std::size_t ki = common_a.size(); ki--;) {
    for (int b = 0; b < common_a[ki]->bits; ++b, ++k) {
      d.a_k[k] = (int8_t)(common_a[ki]->pos + b);
      d.b_k[k] = (int8_t)(common_b[ki]->pos + b);
    }
  }
  if (m + k > 32 || n + k > 32 || m + n > 32)
    throw ERROR_MSG("Tensor of more than 2^32 elements in contraction.");
  d.m_bits = m;
  d.n_bits = n;
  d.k_bits = k;
  step->kind = PlanStep::CONTRACT;
  step->var_a = A.var;
  step->var_b = B.var;
  const double M = (double)(1ull << m), N = (double)(1ull << n), K = (double)(1ull << k);
  step->flops = 8.0 * M * N * K;
  step->bytes = 8.0 * (M * K + K * N + M * N);
  return C;
}

struct GateTensor {
  SymTensor sym;             // dense
  std::vector<s_type> data;
};

struct SiteNet {
  std::vector<GateTensor> chain;                         // time order
  std::unordered_map<std::string, std::string> virt_to_bond;
};

std::string time_leg(std::size_t i, std::size_t j, std::size_t n) {
  return utils::concat("(", i, ",", j, "),(", n, ")");
}

GateTensor make_gate(const std::vector<std::string>& names,
                     const std::vector<std::size_t>& dims,
                     std::vector<s_type> data) {
  GateTensor g;
  std::size_t total = 1;
  for (std::size_t q = 0; q < names.size(); ++q) {
    SymIndex si;
    si.name = names[q];
    si.bits = log2_exact(dims[q]);
    g.sym.idx.push_back(si);
    total *= dims[q];
  }
  if (total != data.size())
    throw ERROR_MSG("The total allocated space: ", total,
                    ", is insufficient for the requested tensor dimensions: ",
                    data.size(), ".");
  g.sym.make_dense();
  g.data = std::move(data);
  return g;
}

// Gates -> per-site chains of small tensors.  Same names, dims, values and the
// same validation as src/read_circuit.cpp:296-544.
void build_network(const QflexInput& input,
                   const std::vector<std::vector<std::size_t>>& final_qubits,
                   std::vector<std::vector<SiteNet>>* net) {
  const QflexCircuit& circuit = input.circuit;
  const std::size_t I = input.grid.I, J = input.grid.J;
  const auto& off = input.grid.qubits_off;
  // (passing `off` to find_grid_coord_in_list would copy it into a std::optional on
  // every call -- thousands of calls per plan)
  std::vector<char> off_mask(I * J, 0);
  for (const auto& c : off)
    if (c.size() == 2 && c[0] < I && c[1] < J) off_mask[c[0] * J + c[1]] = 1;
  auto is_off = [&](std::size_t i, std::size_t j) {
    return i < I && j < J && off_mask[i * J + j] != 0;
  };
  const std::size_t active = I * J - off.size();
  if (circuit.num_active_qubits != active)
    throw ERROR_MSG("The number of active qubits read from the file: ",
                    circuit.num_active_qubits,
                    ", does not match the number of active qubits read from the grid: ",
                    active, ".");
  const std::string& initial_conf = input.initial_state;
  const std::string& final_conf = input.final_state;
  if (initial_conf.size() != active)
    throw ERROR_MSG("Size of initial_conf: ", initial_conf.size(),
                    ", must be equal to the number of qubits: ", active, ".");
  if (final_conf.size() != initial_conf.size())
    throw ERROR_MSG("Size of final_conf: ", final_conf.size(),
                    ", must be equal to size of initial_conf: ",
                    initial_conf.size(), ".");

  net->assign(I, std::vector<SiteNet>(J));
  std::vector<std::vector<std::size_t>> counter(I, std::vector<std::size_t>(J, 0));
  std::unordered_map<std::string, std::size_t> link_counter;

  std::size_t a = 0;
  for (std::size_t q = 0; q < I * J; ++q) {
    const std::size_t i = q / J, j = q % J;
    if (is_off(i, j)) continue;
    const char* delta = (initial_conf[a] == '0') ? "delta_0" : "delta_1";
    (*net)[i][j].chain.push_back(
        make_gate({time_leg(i, j, 0)}, {2}, gate_array(delta, {})));
    ++a;
  }
  for (const auto& w : off)
    if (w[0] >= I || w[1] >= J)
      throw ERROR_MSG("Off qubit '(", w[0], ", ", w[1], ")' is outside the grid.");

  std::unordered_set<std::size_t> busy;
  std::size_t last_cycle = 0;
  for (const auto& gate : circuit.gates) {
    if (last_cycle != gate.cycle) {
      last_cycle = gate.cycle;
      busy.clear();
    }
    for (const auto q : gate.qubits)
      if (!busy.insert(q).second)
        throw ERROR_MSG("Qubit ", q,
                        " has been used twice in the same cycle: ", gate.cycle);

    if (gate.qubits.size() == 1) {
      const std::size_t q1 = gate.qubits[0], i = q1 / J, j = q1 % J;
      if (is_off(i, j))
        throw ERROR_MSG("Qubit '", q1, "' in gate '", gate.raw,
                        "' must correspond to an active qubit.");
      const std::string in = time_leg(i, j, counter[i][j]);
      const std::string out = time_leg(i, j, ++counter[i][j]);
      try {
        (*net)[i][j].chain.push_back(
            make_gate({in, out}, {2, 2}, gate_array(gate.name, gate.params)));
      } catch (const std::string& err_msg) {
        throw ERROR_MSG("Failed to call Tensor(). Error:\n\t[", err_msg, "]");
      }
    } else if (gate.qubits.size() == 2) {
      const std::size_t q1 = gate.qubits[0], q2 = gate.qubits[1];
      const std::size_t i1 = q1 / J, j1 = q1 % J, i2 = q2 / J, j2 = q2 % J;
      if (is_off(i1, j1))
        throw ERROR_MSG("Qubit '", q1, "' in gate '", gate.raw,
                        "' must correspond to an active qubit.");
      if (is_off(i2, j2))
        throw ERROR_MSG("Qubit '", q2, "' in gate '", gate.raw,
                        "' must correspond to an active qubit.");
      const std::size_t di = i1 > i2 ? i1 - i2 : i2 - i1;
      const std::size_t dj = j1 > j2 ? j1 - j2 : j2 - j1;
      if (di + dj != 1)
        throw ERROR_MSG("Qubits ", q1, " and ", q2, " are not nearest neighbors.");

      auto [half_1, half_2, dims] = gate_arrays(gate.name, gate.params);
      const std::string link = index_name({i1, j1}, {i2, j2});
      const std::size_t c = ++link_counter[link];
      const std::size_t lo_i = std::min(i1, i2), lo_j = std::min(j1, j2);
      const std::size_t hi_i = std::max(i1, i2), hi_j = std::max(j1, j2);
      const std::string virt = index_name({lo_i, lo_j, c}, {hi_i, hi_j, c});
      const std::string bond = index_name({lo_i, lo_j}, {hi_i, hi_j});
      const std::string in_1 = time_leg(i1, j1, counter[i1][j1]);
      const std::string in_2 = time_leg(i2, j2, counter[i2][j2]);
      const std::string out_1 = time_leg(i1, j1, ++counter[i1][j1]);
      const std::string out_2 = time_leg(i2, j2, ++counter[i2][j2]);
      try {
        (*net)[i1][j1].chain.push_back(make_gate({in_1, virt, out_1}, dims, half_1));
        (*net)[i2][j2].chain.push_back(make_gate({in_2, virt, out_2}, dims, half_2));
      } catch (const std::string& err_msg) {
        throw ERROR_MSG("Failed to call Tensor(). Error:\n\t[", err_msg, "]");
      }
      (*net)[i1][j1].virt_to_bond[virt] = bond;
      (*net)[i2][j2].virt_to_bond[virt] = bond;
    } else {
      throw ERROR_MSG("k-qubit gates with k > 2 not yet implemented.");
    }
  }

  // close every world line: delta on fixed outputs, open "(i,j),(o)" leg on
  // qubits with a terminal cut
  a = 0;
  for (std::size_t q = 0; q < I * J; ++q) {
    const std::size_t i = q / J, j = q % J;
    if (is_off(i, j)) continue;
    const std::string last = time_leg(i, j, counter[i][j]);
    if (find_grid_coord_in_list(final_qubits, i, j)) {
      SymTensor& t = (*net)[i][j].chain.back().sym;
      const int p = t.find(last);
      if (p < 0) throw ERROR_MSG("old_name: ", last, ", has to be a valid index.");
      t.idx[p].name = index_name({i, j}, {});
    } else {
      const char* delta = (final_conf[a] == '0') ? "delta_0" : "delta_1";
      (*net)[i][j].chain.push_back(make_gate({last}, {2}, gate_array(delta, {})));
    }
    ++a;
  }
}

// Comparator that orders the neighbours of `local` by when their bond is
// consumed: cut bonds first (in cut order), then neighbours of the same patch
// by expand position, then other patches by merge order
// (semantics of order_func, src/read_circuit.cpp:191-291).
struct NeighbourOrder {
  const std::list<ContractionOperation>& ordering;
  GridPos local;

  std::vector<GridPos> bond_of(const GridPos& other) const {
    if (local[0] < other[0] || local[1] < other[1]) return {local, other};
    return {other, local};
  }
  bool find_expand(const GridPos& pos, std::string* patch, std::size_t* nth) const {
    std::size_t n = 0;
    for (const auto& op : ordering) {
      if (op.op_type != ContractionOperation::EXPAND) continue;
      if (op.expand.tensor == pos) {
        *patch = op.expand.id;
        *nth = n;
        return true;
      }
      ++n;
    }
    return false;
  }
  bool operator()(const GridPos& lhs, const GridPos& rhs) const {
    const auto lbond = bond_of(lhs), rbond = bond_of(rhs);
    for (const auto& op : ordering) {
      if (op.op_type != ContractionOperation::CUT) continue;
      if (op.cut.tensors == lbond) return true;
      if (op.cut.tensors == rbond) return false;
    }
    std::string lpatch, rpatch, home;
    std::size_t lnth = 0, rnth = 0, hnth = 0;
    if (!find_expand(lhs, &lpatch, &lnth))
      throw ERROR_MSG("Left hand side of pair not found: (", local[0], ',', local[1],
                      "),(", lhs[0], ',', lhs[1], ").");
    if (!find_expand(rhs, &rpatch, &rnth))
      throw ERROR_MSG("Right hand side of pair not found: (", local[0], ',', local[1],
                      "),(", rhs[0], ',', rhs[1], ").");
    if (lpatch == rpatch) return lnth < rnth;
    find_expand(local, &home, &hnth);
    if (lpatch == home) return true;
    if (rpatch == home) return false;
    for (const auto& op : ordering) {
      if (op.op_type != ContractionOperation::MERGE) continue;
      const auto& src = op.merge.source_id;
      const auto& dst = op.merge.target_id;
      if (home == src) {
        if (lpatch == dst) return true;
        if (rpatch == dst) return false;
        home = dst;
      } else if (home == dst) {
        if (lpatch == src) return true;
        if (rpatch == src) return false;
      } else if (lpatch == src) {
        lpatch = dst;
      } else if (rpatch == src) {
        rpatch = dst;
      }
    }
    throw ERROR_MSG("Failed to compare (", lhs[0], ',', lhs[1], ") and (", rhs[0],
                    ',', rhs[1], ") for local (", local[0], ',', local[1], ").");
  }
};

}  // namespace

namespace {
// Largest j such that the value list of a bond cut is a sequence of aligned blocks
// {u * 2^j, ..., u * 2^j + 2^j - 1}: the low j bits of the bond then run over their
// full range inside every block and need not be sliced.
int fusable_bits(std::vector<std::size_t> values) {
  // (the order of a cut's values only orders the fp64 sum over its slices)
  std::sort(values.begin(), values.end());
  int j = 0;
  while ((std::size_t(2) << j) <= values.size()) ++j;
  for (; j > 0; --j) {
    const std::size_t block = std::size_t(1) << j;
    if (values.size() % block) continue;
    bool ok = true;
    for (std::size_t t = 0; ok && t < values.size(); t += block) {
      if (values[t] % block) ok = false;
      for (std::size_t i = 1; ok && i < block; ++i)
        if (values[t + i] != values[t] + i) ok = false;
    }
    if (ok) return j;
  }
  return 0;
}

// a terminal cut whose values are the whole range in natural order can stay open
bool cut_can_open(const ContractionOperation& op) {
  if (op.op_type != ContractionOperation::CUT || op.cut.tensors.size() != 1) return false;
  const auto& v = op.cut.values;
  if (v.empty()) return true;
  if (v.size() != 2) return false;  // output indices have dimension 2
  return v[0] == 0 && v[1] == 1;
}
}  // namespace

Plan BuildPlan(const QflexInput& input,
               const std::list<ContractionOperation>& ordering,
               const std::vector<std::vector<std::size_t>>& final_qubits,
               int open_outputs, bool cut_aware_flatten, int fuse_bits, PlanCache* cache) {
  Plan plan;
  // the LAST `open_outputs` openable terminal cuts of the ordering stay open (a later
  // cut enlarges fewer of the remaining patches)
  std::vector<const ContractionOperation*> open_ops;
  {
    std::vector<const ContractionOperation*> can;
    for (const auto& op : ordering)
      if (cut_can_open(op)) can.push_back(&op);
    const std::size_t n = std::min<std::size_t>(can.size(), (std::size_t)std::max(open_outputs, 0));
    open_ops.assign(can.end() - n, can.end());
  }
  auto cut_stays_open = [&](const ContractionOperation& op, int) {
    return std::find(open_ops.begin(), open_ops.end(), &op) != open_ops.end();
  };
  const std::size_t I = input.grid.I, J = input.grid.J;
  const auto& off = input.grid.qubits_off;
  // (passing `off` to find_grid_coord_in_list would copy it into a std::optional on
  // every call -- thousands of calls per plan)
  std::vector<char> off_mask(I * J, 0);
  for (const auto& c : off)
    if (c.size() == 2 && c[0] < I && c[1] < J) off_mask[c[0] * J + c[1]] = 1;
  auto is_off = [&](std::size_t i, std::size_t j) {
    return i < I && j < J && off_mask[i * J + j] != 0;
  };
  // Phase 1 (network, gate uploads, flatten steps, grid-tensor layouts: ~95 % of the
  // planning time) does not depend on open_outputs / fuse_bits -- except, with the
  // cut-aware flatten, on which leading terminal cuts stay open -- so the searches of
  // PrepareCircuit over those two knobs reuse it through `cache`.
  std::vector<const ContractionOperation*> phase1_key;
  if (cut_aware_flatten) phase1_key = open_ops;
  std::map<std::string, int> pre_bond, pre_width;
  std::vector<PlanStep> dyn_steps;
  uint64_t work_base = 0, work_flatten = 0;
  if (cache != nullptr && cache->valid && cache->cut_aware == cut_aware_flatten &&
      cache->key == phase1_key) {
    plan = cache->plan;
    pre_bond = cache->pre_bond;
    pre_width = cache->pre_width;
    dyn_steps = cache->dyn_steps;
    work_base = cache->work_base;
    work_flatten = cache->work_flatten;
  } else {
    plan.flatten_steps.reserve(4096);  // PlanStep is ~300 bytes: avoid regrowth copies
    plan.steps.reserve(512);

    std::vector<std::vector<SiteNet>> net;
    build_network(input, final_qubits, &net);

    // ------------------------------------------ cut-aware flatten: project the gates
    // bond cuts that open the ordering -> id of the cut (position among the kept cuts)
    std::vector<std::vector<char>> dynamic(I, std::vector<char>(J, 0));
    if (cut_aware_flatten) {
      int id = 0;
      for (const auto& op : ordering) {
        if (op.op_type != ContractionOperation::CUT) break;
        if (cut_stays_open(op, 0)) continue;
        if (op.cut.tensors.size() == 2) pre_bond[index_name(op.cut.tensors)] = id;
        ++id;
      }
      for (std::size_t i = 0; i < I; ++i)
        for (std::size_t j = 0; j < J; ++j) {
          // virtual legs of every pre-cut bond, in chain (= chronological) order
          std::map<std::string, std::vector<std::pair<std::size_t, std::string>>> legs;
          auto& chain = net[i][j].chain;
          for (std::size_t k = 0; k < chain.size(); ++k)
            for (const auto& ix : chain[k].sym.idx) {
              auto it = net[i][j].virt_to_bond.find(ix.name);
              if (it != net[i][j].virt_to_bond.end() && pre_bond.count(it->second))
                legs[it->second].push_back({k, ix.name});
            }
          for (const auto& kv : legs) {
            int total = 0;
            for (const auto& leg : kv.second) {
              const SymTensor& t = chain[leg.first].sym;
              total += t.idx[t.find(leg.second)].bits;
            }
            pre_width[kv.first] = total;
            // A slice of a bond of w bits carries ~2^(-w/2) of the amplitude: with
            // hundreds of cut bits the partial amplitudes leave fp32's range (Sycamore
            // m=20: 2^-220).  One side of the bond (its first site) takes a factor 2 per
            // two projected bits -- exact in floating point -- undone in fp64 at the end.
            const bool scale_here = kv.first.rfind(utils::concat("(", i, ",", j, "),"), 0) == 0;
            int carry = 0;
            int seen = 0;  // the first leg is the most significant field of the bond value
            for (const auto& leg : kv.second) {
              SymTensor& t = chain[leg.first].sym;
              const int p = t.find(leg.second);
              seen += t.idx[p].bits;
              if (scale_here) {
                carry += t.idx[p].bits;
                for (; carry >= 2; carry -= 2) {
                  for (auto& x : chain[leg.first].data) x *= 2.0f;
                  ++plan.amp_scale_log2;
                }
              }
              VarTerm term;
              term.cut = pre_bond.at(kv.first);
              term.shift = t.idx[p].pos;
              term.from = total - seen;
              term.width = t.idx[p].bits;
              t.var.push_back(term);
              t.idx.erase(t.idx.begin() + p);
            }
            dynamic[i][j] = 1;
          }
        }
    }

    // ---------------------------------------------------------- arena: gates
    uint64_t cursor = 0;
    for (std::size_t i = 0; i < I; ++i)
      for (std::size_t j = 0; j < J; ++j) {
        for (auto& g : net[i][j].chain) {
          g.sym.off = cursor;
          PlanUpload up;
          up.off = cursor;
          up.data = g.data;
          plan.uploads.push_back(std::move(up));
          cursor = align_up(cursor + g.sym.size());
        }
        if (is_off(i, j)) continue;
        // build_network closes every world line without a terminal cut with a delta
        const bool open_leg = find_grid_coord_in_list(final_qubits, i, j);
        plan.final_delta_upload.push_back(open_leg ? -1 : (int)plan.uploads.size() - 1);
      }

    // ------------------------------------------------- flatten: sizes first
    plan.site.assign(I, std::vector<SymTensor>(J));
    uint64_t max_chain = 1, max_dyn_chain = 0;
    // The world-line products are derived ONCE here (bit tables, sizes); their arena
    // offsets are filled in below, when the ping-pong buffers are placed.
    std::vector<std::vector<std::vector<PlanStep>>> line_steps(
        I, std::vector<std::vector<PlanStep>>(J));
    std::vector<std::vector<SymTensor>> line_result(I, std::vector<SymTensor>(J));
    // tree-ordered world lines (dynamic sites): per step, whether operand a / b is an
    // intermediate (offset relative to the dyn region) or a gate tensor (absolute)
    // region of operand a / operand b / the result: 0 = a gate tensor (absolute
    // offset), 1 = the dyn region shared by all dynamic sites (large results, sites run
    // one after the other), 2 = the site's own small region (results <= kSmallLine
    // elements whose whole subtree is small: those steps of ALL dynamic sites run as
    // one resident program, qfx_program_run, one block per site)
    struct LineRel {
      uint8_t a, b, c;
    };
    constexpr uint64_t kSmallLine = uint64_t(1) << 13;
    std::vector<std::vector<std::vector<LineRel>>> line_rel(
        I, std::vector<std::vector<LineRel>>(J));
    std::vector<std::vector<uint64_t>> line_small_need(I, std::vector<uint64_t>(J, 0));
    uint64_t max_dyn_tree = 0;
    bool line_programs = true;
    if (const char* e = getenv("QFLEX_LINE_PROGRAMS")) line_programs = e[0] != '0';
    {
      std::size_t n_cut_ops = 0;
      for (const auto& op : ordering) n_cut_ops += op.op_type == ContractionOperation::CUT;
      if (n_cut_ops > QFX_MAX_CUT_VALUES) line_programs = false;  // values travel as kernel arguments
    }
    const bool tree_lines = [] {
      const char* e = getenv("QFLEX_WORLDLINE_TREE");
      return e == nullptr || e[0] != '0';
    }();
    for (std::size_t i = 0; i < I; ++i)
      for (std::size_t j = 0; j < J; ++j) {
        if (is_off(i, j)) continue;
        if (net[i][j].chain.empty()) throw ERROR_MSG("Time-direction cannot be empty");
        SymTensor a = net[i][j].chain[0].sym;
        line_steps[i][j].reserve(net[i][j].chain.size());
        if (dynamic[i][j] && tree_lines && net[i][j].chain.size() > 2) {
          // A site flattened inside the slice loop pays for every byte of every
          // intermediate once per slice.  Left to right, each of the last gates is a
          // pass over the full-size tensor (Sycamore m=16: ~80 passes over 2^24
          // elements per slice).  As a balanced tree -- split where the virtual-leg
          // bits left and right are closest -- only the root writes a full-size
          // tensor; the halves are about its square root.  Same factors, same
          // contracted legs, same index order of the result (concatenation is
          // associative); intermediates are bump-allocated in the dyn region.
          const auto& chain = net[i][j].chain;
          std::vector<int> bits_before(chain.size() + 1, 0);
          for (std::size_t k = 0; k < chain.size(); ++k) {
            int vb = 0;
            for (const auto& ix : chain[k].sym.idx)
              if (net[i][j].virt_to_bond.count(ix.name)) vb += ix.bits;
            bits_before[k + 1] = bits_before[k] + vb;
          }
          uint64_t bump = 0, bump_small = 0;
          auto& steps_out = line_steps[i][j];
          auto& rel_out = line_rel[i][j];
          const std::size_t n_gates = chain.size();
          std::function<SymTensor(std::size_t, std::size_t, int*)> build =
              [&](std::size_t lo, std::size_t hi, int* region) -> SymTensor {
            if (hi - lo == 1) {
              *region = 0;
              return chain[lo].sym;
            }
            std::size_t mid = lo + 1;
            int best = 1 << 30;
            for (std::size_t m = lo + 1; m < hi; ++m) {
              const int l = bits_before[m] - bits_before[lo], r = bits_before[hi] - bits_before[m];
              const int d = std::abs(l - r) * 64 + (int)std::min<std::size_t>(63, m > (lo + hi) / 2 ? m - (lo + hi) / 2 : (lo + hi) / 2 - m);
              if (d < best) best = d, mid = m;
            }
            int lreg = 0, rreg = 0;
            SymTensor L = build(lo, mid, &lreg);
            SymTensor R = build(mid, hi, &rreg);
            steps_out.emplace_back();
            // small = small result, small subtrees, a short contraction, not the root
            // (the root may be given the grid tensor's layout below)
            PlanStep probe;
            const SymTensor sized = make_contract(L, R, 0, &probe);
            const bool small = line_programs && sized.size() <= kSmallLine && lreg != 1 &&
                               rreg != 1 && probe.desc.k_bits <= 6 && !(lo == 0 && hi == n_gates);
            uint64_t& cursor_c = small ? bump_small : bump;
            SymTensor C = make_contract(L, R, cursor_c, &steps_out.back());
            rel_out.push_back({(uint8_t)lreg, (uint8_t)rreg, (uint8_t)(small ? 2 : 1)});
            cursor_c = align_up(cursor_c + C.size());
            *region = small ? 2 : 1;
            return C;
          };
          int top_region = 0;
          a = build(0, chain.size(), &top_region);
          max_dyn_tree = std::max(max_dyn_tree, bump);
          line_small_need[i][j] = bump_small;
        } else {
          for (std::size_t k = 1; k < net[i][j].chain.size(); ++k) {
            line_steps[i][j].emplace_back();
            a = make_contract(a, net[i][j].chain[k].sym, 0, &line_steps[i][j].back());
            if (dynamic[i][j]) max_dyn_chain = std::max(max_dyn_chain, a.size());
            else max_chain = std::max(max_chain, a.size());
          }
        }
        line_result[i][j] = a;
        // final layout: [(o)] ++ bonds ordered by consumption
        SymTensor& site = plan.site[i][j];
        if (find_grid_coord_in_list(final_qubits, i, j)) {
          SymIndex o;
          o.name = index_name({i, j}, {});
          o.bits = 1;
          site.idx.push_back(o);
        }
        std::vector<GridPos> pairs;
        if (i > 0 && !is_off(i - 1, j)) pairs.push_back({i - 1, j});
        if (j > 0 && !is_off(i, j - 1)) pairs.push_back({i, j - 1});
        if (i + 1 < I && !is_off(i + 1, j)) pairs.push_back({i + 1, j});
        if (j + 1 < J && !is_off(i, j + 1)) pairs.push_back({i, j + 1});
        NeighbourOrder cmp{ordering, {i, j}};
        std::sort(pairs.begin(), pairs.end(), cmp);
        for (const auto& p : pairs) {
          SymIndex b;
          b.name = (p[0] < i || p[1] < j) ? index_name(p, {i, j}) : index_name({i, j}, p);
          b.bits = 0;
          for (const auto& ix : a.idx) {
            auto it = net[i][j].virt_to_bond.find(ix.name);
            if (it != net[i][j].virt_to_bond.end() && it->second == b.name) b.bits += ix.bits;
          }
          if (b.bits == 0 && pre_bond.count(b.name)) continue;  // projected in the gates
          site.idx.push_back(b);
        }
        site.make_dense();
        if (site.rank_bits() != a.rank_bits())
          throw ERROR_MSG("Flattened tensor at (", i, ",", j,
                          ") has indices that are neither bonds nor outputs: ",
                          tensor_to_string(a));
        site.off = cursor;
        cursor = align_up(cursor + site.size());
      }
    // sites flattened inside the slice loop share one ping-pong pair of their own (the
    // work region below belongs to the patches by then)
    const uint64_t dyn_ping = cursor, dyn_pong = cursor + align_up(max_dyn_chain);
    if (max_dyn_chain > 0) cursor = dyn_pong + align_up(max_dyn_chain);
    cursor = std::max(cursor, dyn_ping + align_up(max_dyn_tree));  // (the two share the region)
    std::vector<std::vector<uint64_t>> line_small_base(I, std::vector<uint64_t>(J, 0));
    for (std::size_t i = 0; i < I; ++i)
      for (std::size_t j = 0; j < J; ++j)
        if (line_small_need[i][j] > 0) {
          line_small_base[i][j] = cursor;
          cursor = align_up(cursor + line_small_need[i][j]);
        }
    // the small world-line steps of all dynamic sites: one resident program
    ChainProgram line_program;
    line_program.chain_first.push_back(0);
    line_program.term_first.push_back(0);
    double program_flops = 0, program_bytes = 0;
    work_base = cursor;
    // Ping-pong buffers of the world-line products.  One pair PER SITE when that is
    // affordable (<= 2^26 elements = 512 MB in all): the chains are then independent
    // and run concurrently (qfx_contract_chains); otherwise one shared pair and the
    // sites are flattened one after the other.
    uint64_t n_sites = 0;
    for (std::size_t i = 0; i < I; ++i)
      for (std::size_t j = 0; j < J; ++j)
        if (!is_off(i, j) && !dynamic[i][j]) ++n_sites;
    const uint64_t pair_elems = 2 * align_up(max_chain);
    plan.flatten_parallel = n_sites * pair_elems <= (uint64_t(1) << 26);
    work_flatten = plan.flatten_parallel ? n_sites * pair_elems : pair_elems;

    // ------------------------------------------------- flatten: real steps
    for (std::size_t i = 0; i < I; ++i)
      for (std::size_t j = 0; j < J; ++j) {
        if (is_off(i, j)) continue;
        const auto& chain = net[i][j].chain;
        const bool dyn = dynamic[i][j] != 0;
        std::vector<PlanStep>& out_steps = dyn ? dyn_steps : plan.flatten_steps;
        const std::size_t first_step = out_steps.size();
        const int chain_id = dyn ? -1 : plan.n_chains++;
        const uint64_t ping =
            dyn ? dyn_ping
                : work_base + (plan.flatten_parallel ? (uint64_t)chain_id * pair_elems : 0);
        const uint64_t pong = dyn ? dyn_pong : ping + align_up(max_chain);
        SymTensor a = std::move(line_result[i][j]);
        if (chain.size() == 1) a = chain[0].sym;
        bool in_ping = false;
        uint64_t prev = chain[0].sym.off;
        const bool as_tree = !line_rel[i][j].empty();
        if (as_tree) {
          const uint64_t base_of[3] = {0, dyn_ping, line_small_base[i][j]};
          bool in_program = false;
          for (std::size_t k = 0; k < line_steps[i][j].size(); ++k) {
            PlanStep& st = line_steps[i][j][k];
            const LineRel rel = line_rel[i][j][k];
            st.level = 0;
            st.chain = chain_id;
            st.label = utils::concat("flatten (", i, ",", j, ") tree step ", k + 1);
            st.desc.off_a += base_of[rel.a];
            st.desc.off_b += base_of[rel.b];
            st.desc.off_c += base_of[rel.c];
            st.desc.precision = QFX_PREC_FFMA;
            if (rel.c == 2) {
              // into the program: descriptor + the slice-dependent base terms
              line_program.descs.push_back(st.desc);
              for (const std::vector<VarTerm>* var : {&st.var_a, &st.var_b}) {
                for (const VarTerm& v : *var)
                  line_program.terms.push_back(
                      qfx_var_term{v.cut, v.shift, v.from, std::min(v.width, 32)});
                line_program.term_first.push_back((int32_t)line_program.terms.size());
              }
              program_flops += st.flops;
              program_bytes += st.bytes;
              in_program = true;
              continue;
            }
            prev = st.desc.off_c;
            out_steps.push_back(std::move(st));
          }
          if (in_program) line_program.chain_first.push_back((int32_t)line_program.descs.size());
        }
        for (std::size_t k = 1; k < chain.size() && !as_tree; ++k) {
          PlanStep& st = line_steps[i][j][k - 1];
          st.level = 0;
          st.chain = chain_id;
          st.label = utils::concat("flatten (", i, ",", j, ") gate ", k);
          const uint64_t dst = in_ping ? pong : ping;
          st.desc.off_a = prev;  // the running product (first: the initial delta)
          st.desc.off_c = dst;
          st.desc.precision = QFX_PREC_FFMA;
          prev = dst;
          in_ping = !in_ping;
          out_steps.push_back(std::move(st));
        }
        if (chain.size() > 1) a.off = prev;
        // one permute into the final layout
        const SymTensor& site = plan.site[i][j];
        PlanStep pm;
        pm.kind = PlanStep::PERMUTE;
        pm.level = 0;
        pm.chain = chain_id;
        pm.label = utils::concat("flatten (", i, ",", j, ") layout");
        pm.src = a.off;
        pm.dst = site.off;
        pm.n_bits = a.rank_bits();
        // destination position of every source index: (o) keeps its name, virtual
        // legs land inside their bond in chronological order (first = slowest)
        std::map<std::string, int> bond_fill;  // bits already placed, from the top
        bool identity = true;
        for (const auto& ix : a.idx) {
          int dpos;
          auto it = net[i][j].virt_to_bond.find(ix.name);
          if (it == net[i][j].virt_to_bond.end()) {
            const int s = site.find(ix.name);
            if (s < 0)
              throw ERROR_MSG("Index ", ix.name, " of flattened tensor (", i, ",", j,
                              ") has no place in the grid tensor.");
            dpos = site.idx[s].pos;
          } else {
            const int s = site.find(it->second);
            int& used = bond_fill[it->second];
            used += ix.bits;
            dpos = site.idx[s].pos + site.idx[s].bits - used;
          }
          for (int b = 0; b < ix.bits; ++b) {
            pm.dst_pos[ix.pos + b] = (int8_t)(dpos + b);
            if (ix.pos + b != dpos + b) identity = false;
          }
        }
        static const bool fold_layout = [] {
          const char* e = getenv("QFLEX_FOLD_LAYOUT");
          return e == nullptr || e[0] != '0';
        }();
        if (identity && out_steps.size() > first_step && chain.size() > 1) {
          // last world-line product can write the grid tensor directly
          out_steps.back().desc.off_c = site.off;
        } else if (as_tree && fold_layout && out_steps.size() > first_step) {
          // the root of the tree writes the grid tensor in its final layout: the
          // permute (one more read + write of the full tensor per slice) is folded
          // into the root's C bit table
          qfx_contract_desc& d = out_steps.back().desc;
          for (int q = 0; q < d.m_bits; ++q) d.c_m[q] = pm.dst_pos[d.c_m[q]];
          for (int q = 0; q < d.n_bits; ++q) d.c_n[q] = pm.dst_pos[d.c_n[q]];
          d.off_c = site.off;
        } else {
          pm.bytes = 16.0 * (double)(1ull << pm.n_bits);
          out_steps.push_back(std::move(pm));
        }
      }

    if (!line_program.descs.empty()) {
      PlanStep run;
      run.kind = PlanStep::PROGRAM;
      run.program = (int)plan.programs.size();
      run.n_bits = (int)line_program.descs.size();
      run.flops = program_flops;
      run.bytes = program_bytes;
      run.label = utils::concat("world-line program: ", line_program.descs.size(), " steps of ",
                                line_program.chain_first.size() - 1, " sites");
      plan.programs.push_back(std::move(line_program));
      dyn_steps.insert(dyn_steps.begin(), std::move(run));
    }
    if (cache != nullptr) {
      cache->valid = true;
      cache->cut_aware = cut_aware_flatten;
      cache->key = phase1_key;
      cache->plan = plan;
      cache->pre_bond = pre_bond;
      cache->pre_width = pre_width;
      cache->dyn_steps = dyn_steps;
      cache->work_base = work_base;
      cache->work_flatten = work_flatten;
    }
  }

  // ------------------------------------------------- fused cut bits
  // A bond cut slices the contraction over the values of one bond so that the patches
  // fit the host memory the reference was written for; every slice re-runs the whole
  // schedule after the cut on tensors 2^bits smaller.  With 180 GB of HBM most of that
  // slicing is unnecessary: the LOW `j` bits of a cut bond may simply stay an ordinary
  // index of its two grid tensors (it is contracted when their patches meet, like any
  // other bond), and only the high bits are projected.  2^j slices then cost one pass of
  // 2^j-times larger kernels -- steps that do not carry the bond run once instead of
  // 2^j times -- and the sum over those slices happens inside the merge's K loop.  Same
  // pairs, same order; `fuse_bits` is the total handed out, innermost cut first.
  std::map<const ContractionOperation*, int> fuse_of;
  {
    int remaining = std::max(fuse_bits, 0);
    for (auto it = ordering.rbegin(); it != ordering.rend() && remaining > 0; ++it) {
      const ContractionOperation& op = *it;
      if (op.op_type != ContractionOperation::CUT || op.cut.tensors.size() != 2) continue;
      const std::string index = index_name(op.cut.tensors);
      if (pre_bond.count(index)) continue;  // projected in the gate tensors
      const auto& pos = op.cut.tensors[0];
      if (pos.size() != 2 || pos[0] >= I || pos[1] >= J) continue;
      const SymTensor& t = plan.site[pos[0]][pos[1]];
      const int p = t.find(index);
      if (p < 0) continue;
      std::vector<std::size_t> values = op.cut.values;
      if (values.empty())
        for (std::size_t v = 0; v < (std::size_t(1) << t.idx[p].bits); ++v) values.push_back(v);
      bool in_range = true;
      for (const auto v : values) in_range = in_range && v < (std::size_t(1) << t.idx[p].bits);
      if (!in_range) continue;  // reported by the cut itself below
      const int j = std::min(fusable_bits(values), remaining);
      if (j > 0) {
        fuse_of[&op] = j;
        remaining -= j;
      }
    }
  }
  auto fused_bits_of = [&](const ContractionOperation& op) {
    auto it = fuse_of.find(&op);
    return it == fuse_of.end() ? 0 : it->second;
  };

  // ------------------------------------------------- schedule: patch sizes
  // Symbolic replay (as ContractionData::Initialize does) to size patch slots.
  struct Patch {
    SymTensor cur;
    bool active = false;
    bool cur_in_own = false;
    uint64_t own = 0;       // arena offset of the patch's own slot
    uint64_t need = 0;      // largest contraction result stored for this patch
  };
  std::map<std::string, Patch> patches;
  {
    std::vector<std::vector<SymTensor>> site = plan.site;
    std::map<std::string, SymTensor> cur;
    std::map<std::string, bool> active;
    for (const auto& op : ordering) {
      if (op.op_type == ContractionOperation::EXPAND) {
        const SymTensor& t = site[op.expand.tensor[0]][op.expand.tensor[1]];
        patches[op.expand.id];
        if (!active[op.expand.id]) {
          cur[op.expand.id] = t;
          active[op.expand.id] = true;
        } else {
          PlanStep dummy;
          cur[op.expand.id] = make_contract(cur[op.expand.id], t, 0, &dummy);
          patches[op.expand.id].need =
              std::max(patches[op.expand.id].need, cur[op.expand.id].size());
        }
      } else if (op.op_type == ContractionOperation::MERGE) {
        patches[op.merge.source_id];
        patches[op.merge.target_id];
        if (!active[op.merge.target_id]) {
          cur[op.merge.target_id] = cur[op.merge.source_id];
          active[op.merge.target_id] = true;
        } else {
          PlanStep dummy;
          cur[op.merge.target_id] =
              make_contract(cur[op.merge.source_id], cur[op.merge.target_id], 0, &dummy);
          patches[op.merge.target_id].need =
              std::max(patches[op.merge.target_id].need, cur[op.merge.target_id].size());
        }
      } else if (!cut_stays_open(op, open_outputs)) {
        const std::string index = index_name(op.cut.tensors);
        const int keep = fused_bits_of(op);
        for (const auto& pos : op.cut.tensors) {
          SymTensor& t = site[pos[0]][pos[1]];
          const int p = t.find(index);
          if (p < 0) continue;
          if (keep > 0) t.idx[p].bits = keep;
          else t.idx.erase(t.idx.begin() + p);
        }
      }
    }
  }
  uint64_t cursor = work_base;
  std::map<uint64_t, uint64_t> spare;  // result size -> spare slot offset
  for (auto& kv : patches) {
    if (kv.second.need == 0) continue;
    kv.second.own = cursor;
    cursor = align_up(cursor + kv.second.need);
    spare[kv.second.need] = 0;
  }
  for (auto& kv : spare) {
    kv.second = cursor;
    cursor = align_up(cursor + kv.first);
  }
  plan.arena_elems = std::max(cursor, work_base + work_flatten);

  // ------------------------------------------------- schedule: real steps
  std::vector<std::vector<SymTensor>> site = plan.site;
  std::vector<std::vector<char>> site_used(I, std::vector<char>(J, 0));
  int level = 0;
  plan.n_amplitudes = 1;
  // terminal cuts in ordering order: projected (cut id >= 0) or open (index name)
  struct Terminal {
    int cut_id;
    std::string index;
    uint64_t count;
  };
  std::vector<Terminal> terminals;
  bool last_was_contract = false;
  SymTensor last_tensor;

  auto store_result = [&](Patch& p, const SymTensor& A, const SymTensor& B,
                          const std::string& label) {
    PlanStep st;
    st.level = level;
    st.label = label;
    uint64_t dst;
    bool via_spare = false;
    if (!p.cur_in_own) {
      dst = p.own;
    } else {
      // result goes to the spare slot of this size, which then becomes the
      // patch's slot (reference: scratch_map_ swap, contraction_utils.cpp:255-257)
      dst = spare.at(p.need);
      via_spare = true;
    }
    SymTensor c = make_contract(A, B, dst, &st);
    st.desc.precision = plan.precision;
    if (c.size() > p.need)
      throw ERROR_MSG("C: ", p.need,
                      " doesn't have enough space for the product of A*B: ", c.size(), ".");
    if (via_spare) std::swap(p.own, spare.at(p.need));
    p.cur = c;
    p.cur_in_own = true;
    plan.steps.push_back(std::move(st));
    last_was_contract = true;
    last_tensor = p.cur;
  };

  bool dyn_placed = dyn_steps.empty();
  for (const auto& op : ordering) {
    if (!dyn_placed && op.op_type != ContractionOperation::CUT) {
      // the leading cuts are fixed from here on: flatten the sites next to them
      for (auto& st : dyn_steps) {
        st.level = level;
        plan.steps.push_back(std::move(st));
      }
      dyn_placed = true;
    }
    switch (op.op_type) {
      case ContractionOperation::EXPAND: {
        const std::size_t i = op.expand.tensor[0], j = op.expand.tensor[1];
        Patch& p = patches[op.expand.id];
        const SymTensor& next = site[i][j];
        site_used[i][j] = 1;
        if (!p.active) {
          // first tensor of the patch: alias instead of the reference's copy
          p.cur = next;
          p.cur_in_own = false;
          p.active = true;
          last_was_contract = false;
          last_tensor = p.cur;
          break;
        }
        try {
          store_result(p, p.cur, next,
                       utils::concat("expand ", op.expand.id, " (", i, ",", j, ")"));
        } catch (const std::string& err_msg) {
          throw ERROR_MSG("Failed to call multiply(). Error:\n\t[", err_msg, "]");
        }
        break;
      }
      case ContractionOperation::MERGE: {
        Patch& src = patches[op.merge.source_id];
        Patch& dst = patches[op.merge.target_id];
        if (!dst.active) {
          dst.cur = src.cur;  // alias of the source patch' storage
          dst.cur_in_own = false;
          dst.active = true;
          last_was_contract = false;
          last_tensor = dst.cur;
          break;
        }
        try {
          store_result(dst, src.cur, dst.cur,
                       utils::concat("merge ", op.merge.source_id, " ", op.merge.target_id));
        } catch (const std::string& err_msg) {
          throw ERROR_MSG("Failed to call multiply(). Error:\n\t[", err_msg, "]");
        }
        break;
      }
      case ContractionOperation::CUT: {
        const std::string index = index_name(op.cut.tensors);
        PlanCut cut;
        cut.index = index;
        cut.terminal = (op.cut.tensors.size() == 1);
        cut.values = op.cut.values;
        const int cut_id = (int)plan.cuts.size();
        const bool stays_open = cut_stays_open(op, open_outputs);
        const int keep = fused_bits_of(op);
        int width = -1;
        const bool pre = pre_bond.count(index) != 0;  // already projected in the gates
        if (pre) width = pre_width.count(index) ? pre_width.at(index) : 0;
        for (const auto& pos : op.cut.tensors) {
          if (pre) break;
          SymTensor& t = site[pos[0]][pos[1]];
          const int p = t.find(index);
          if (p < 0 || site_used[pos[0]][pos[1]]) {
            // the reference projects the leading index of an untouched grid
            // tensor (src/tensor.cpp:280-285)
            throw ERROR_MSG("Failed to call project(). Error:\n\t[Index: '", index,
                            "' has to be equal to indices[0]: '",
                            t.idx.empty() ? std::string("") : t.idx[0].name, "'.]");
          }
          width = t.idx[p].bits;
          if (stays_open) continue;
          if (keep >= width) continue;  // the whole bond stays an index
          t.var.push_back(VarTerm{cut_id, t.idx[p].pos + keep});
          if (keep > 0) t.idx[p].bits = keep;  // low bits stay (same position)
          else t.idx.erase(t.idx.begin() + p);
        }
        if (stays_open) {
          terminals.push_back(Terminal{-1, index, uint64_t(1) << width});
          plan.n_amplitudes <<= width;
          ++plan.open_outputs;
          break;
        }
        if (cut.values.empty())
          for (std::size_t v = 0; v < (std::size_t(1) << width); ++v) cut.values.push_back(v);
        for (const auto v : cut.values)
          if (v >= (std::size_t(1) << width))
            throw ERROR_MSG("Failed to call project(). Error:\n\t[index_value: ", v,
                            " must be contained in [0, ", (std::size_t(1) << width), ").]");
        if (cut.terminal) {
          plan.n_amplitudes *= cut.values.size();
          terminals.push_back(Terminal{cut_id, index, (uint64_t)cut.values.size()});
        }
        if (keep > 0) {
          // one outer value per aligned block of 2^keep values
          std::vector<std::size_t> outer;
          std::sort(cut.values.begin(), cut.values.end());
          for (std::size_t v = 0; v < cut.values.size(); v += std::size_t(1) << keep)
            outer.push_back(cut.values[v] >> keep);
          cut.values = std::move(outer);
          cut.fused_bits = keep;
          plan.fused_bits += keep;
        }
        plan.cuts.push_back(std::move(cut));
        ++level;
        break;
      }
    }
  }

  // the schedule must end in a scalar, which is accumulated into the amplitude
  // of the current output bitstring (src/contraction_utils.cpp:347-352)
  if (ordering.empty()) throw ERROR_MSG("Empty contraction ordering.");
  // amplitude slot = mixed radix over the terminal cuts, first = most significant
  std::map<std::string, uint64_t> open_stride;
  {
    uint64_t stride = 1;
    for (std::size_t t = terminals.size(); t--;) {
      if (terminals[t].cut_id >= 0)
        plan.cuts[terminals[t].cut_id].amp_stride = stride;
      else
        open_stride[terminals[t].index] = stride;
      stride *= terminals[t].count;
    }
  }
  for (const auto& ix : last_tensor.idx)
    if (!open_stride.count(ix.name))
      throw ERROR_MSG("Contraction did not complete; final tensor is: ",
                      tensor_to_string(last_tensor));
  if (last_tensor.rank_bits() != 0) {
    // open outputs: the last contraction (if any) stays a CONTRACT that writes the
    // tensor of amplitudes, one more step adds it into the table
    if ((int)open_stride.size() * 1 != (int)last_tensor.idx.size() ||
        last_tensor.rank_bits() > QFX_MAX_BITS)
      throw ERROR_MSG("Contraction did not complete; final tensor is: ",
                      tensor_to_string(last_tensor));
    PlanStep st;
    st.kind = PlanStep::TENSOR_ACCUMULATE;
    st.level = level;
    st.scalar_off = last_tensor.off;
    st.var_scalar = last_tensor.var;
    st.n_bits = last_tensor.span_bits;
    st.label = "accumulate open outputs";
    for (const auto& ix : last_tensor.idx)
      for (int b = 0; b < ix.bits; ++b)
        st.slot_stride[ix.pos + b] = (uint32_t)(open_stride.at(ix.name) << b);
    if (last_tensor.span_bits != last_tensor.rank_bits())
      throw ERROR_MSG("Final tensor of open outputs is not dense: ",
                      tensor_to_string(last_tensor));
    plan.steps.push_back(std::move(st));
  } else if (last_was_contract) {
    plan.steps.back().kind = PlanStep::DOT_ACCUMULATE;
  } else {
    PlanStep st;
    st.kind = PlanStep::SCALAR_ACCUMULATE;
    st.level = level;
    st.scalar_off = last_tensor.off;
    st.var_scalar = last_tensor.var;
    st.label = "accumulate scalar";
    plan.steps.push_back(std::move(st));
  }
  if (plan.steps.back().level != level)
    throw ERROR_MSG("Contraction ordering ends with a cut; nothing to contract after it.");

  for (const auto& st : plan.steps) {
    plan.flops_per_pass += st.flops;
    plan.bytes_per_pass += st.bytes;
  }
  return plan;
}

}  // namespace qflex
