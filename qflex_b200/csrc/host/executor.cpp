// Plan executor: walks the cut tree and issues device steps through the C ABI.
// The recursion mirrors ContractionData::ContractGrid
// (reference src/contraction_utils.cpp:223-354): steps before a cut run once,
// everything after it runs once per cut value; a bond cut adds into the same
// amplitude, a terminal cut selects amplitude output_index*|values|+i (here: slot
// += i * amp_stride, so that projected and open terminal cuts can mix).
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <thread>

#include "errors.h"
#include "plan.h"

namespace qflex {
namespace {

void check(int rc, const char* call) {
  if (rc != QFX_OK)
    throw ERROR_MSG("Device call ", call, " failed: ", qfx_last_error());
}

double now_s() {
  return std::chrono::duration<double>(
             std::chrono::steady_clock::now().time_since_epoch())
      .count();
}

class Runner {
 public:
  Runner(const Plan& plan, qfx_ctx* ctx, int rank, int world,
         const std::vector<qfx_program*>* programs = nullptr)
      : plan_(plan), ctx_(ctx), rank_(rank), world_(world), programs_(programs),
        values_(plan.cuts.size(), 0) {
    // Slices are dealt to ranks at the shallowest cut levels whose combined
    // fan-out reaches the number of ranks.
    uint64_t fan = 1;
    share_levels_ = 0;
    while (share_levels_ < (int)plan.cuts.size() && fan < (uint64_t)world) {
      fan *= plan.cuts[share_levels_].values.size();
      ++share_levels_;
    }
  }

  uint64_t leaves() const { return leaves_; }

  void run(const std::vector<PlanStep>& steps) {
    // Nothing to deal (no projected cut at all, e.g. a cut-free ordering or one whose
    // only cuts are open outputs): the whole schedule, accumulate steps included, is
    // one share and belongs to rank 0 -- the other ranks contribute zeros to the sum.
    if (share_levels_ == 0 && world_ > 1 && rank_ != 0) return;
    walk(steps, 0, 0, 0, 0);
  }

 private:
  uint64_t resolve(uint64_t base, const std::vector<VarTerm>& var) const {
    for (const auto& v : var) {
      uint64_t x = (uint64_t)values_[v.cut] >> v.from;
      if (v.width < 64) x &= (uint64_t(1) << v.width) - 1;
      base += x << v.shift;
    }
    return base;
  }

  void issue(const PlanStep& st, uint64_t amp_index) {
    switch (st.kind) {
      case PlanStep::CONTRACT: {
        qfx_contract_desc d = st.desc;
        d.off_a = resolve(d.off_a, st.var_a);
        d.off_b = resolve(d.off_b, st.var_b);
        check(qfx_contract(ctx_, &d), "qfx_contract");
        break;
      }
      case PlanStep::DOT_ACCUMULATE: {
        qfx_contract_desc d = st.desc;
        d.off_a = resolve(d.off_a, st.var_a);
        d.off_b = resolve(d.off_b, st.var_b);
        check(qfx_dot_accumulate(ctx_, &d, amp_index), "qfx_dot_accumulate");
        ++leaves_;
        break;
      }
      case PlanStep::SCALAR_ACCUMULATE:
        check(qfx_scalar_accumulate(ctx_, resolve(st.scalar_off, st.var_scalar),
                                    amp_index),
              "qfx_scalar_accumulate");
        ++leaves_;
        break;
      case PlanStep::TENSOR_ACCUMULATE:
        check(qfx_tensor_accumulate(ctx_, resolve(st.scalar_off, st.var_scalar), st.n_bits,
                                    st.slot_stride, amp_index),
              "qfx_tensor_accumulate");
        leaves_ += uint64_t(1) << st.n_bits;
        break;
      case PlanStep::PERMUTE:
        check(qfx_permute(ctx_, st.src, st.dst, st.n_bits, st.dst_pos), "qfx_permute");
        break;
      case PlanStep::PROGRAM: {
        if (programs_ == nullptr || st.program < 0 || st.program >= (int)programs_->size())
          throw ERROR_MSG("Chain program ", st.program, " is not resident on this lane.");
        uint32_t values[QFX_MAX_CUT_VALUES];
        const int n = (int)std::min<std::size_t>(values_.size(), QFX_MAX_CUT_VALUES);
        for (int c = 0; c < n; ++c) values[c] = (uint32_t)values_[c];
        check(qfx_program_run(ctx_, (*programs_)[st.program], values, n), "qfx_program_run");
        break;
      }
    }
  }

  // `combo` numbers the path through the shared (rank-dealt) cut levels.
  void walk(const std::vector<PlanStep>& steps, std::size_t first, int level,
            uint64_t amp_index, uint64_t combo) {
    std::size_t i = first;
    // Work above the dealt levels is replicated on every rank; below them a
    // rank only descends into its own combos.
    for (; i < steps.size() && steps[i].level == level; ++i) issue(steps[i], amp_index);
    if (level == (int)plan_.cuts.size()) return;
    const PlanCut& cut = plan_.cuts[level];
    for (std::size_t v = 0; v < cut.values.size(); ++v) {
      const uint64_t next_combo = combo * cut.values.size() + v;
      if (level + 1 == share_levels_ && world_ > 1 &&
          (int)(next_combo % (uint64_t)world_) != rank_)
        continue;
      values_[level] = cut.values[v];
      walk(steps, i, level + 1, cut.terminal ? amp_index + v * cut.amp_stride : amp_index,
           level < share_levels_ ? next_combo : combo);
    }
  }

  const Plan& plan_;
  qfx_ctx* ctx_;
  int rank_, world_;
  const std::vector<qfx_program*>* programs_;
  int share_levels_ = 0;
  std::vector<std::size_t> values_;
  uint64_t leaves_ = 0;
};

}  // namespace

namespace {
// lanes per device
int choose_lanes(const Plan& plan, int n_devices) {
  if (const char* e = getenv("QFLEX_LANES")) return std::max(1, std::min(16, atoi(e)));
  uint64_t slices = 1;
  for (const auto& c : plan.cuts) slices *= std::max<uint64_t>(1, c.values.size());
  slices = std::max<uint64_t>(1, slices / (uint64_t)std::max(1, n_devices));
  const double arena_bytes = (double)plan.arena_elems * 8.0;
  // lanes only for networks whose arena is small next to the GPU (<= 2 GB each, 4 GB
  // in all): their kernels run for microseconds and do not fill the SMs
  int lanes = (int)std::min<double>(8.0, (4096.0 * 1048576.0) / std::max(arena_bytes, 1.0));
  if (arena_bytes > 2048.0 * 1048576.0) {
    lanes = 1;
    // larger arenas whose schedule is still a stream of small launches (cut-aware
    // flatten of Sycamore m=16: ~150 launches per slice of ~1e10 flop, tens of
    // microseconds each, with dependent launches back to back: 139 TFLOP/s on one
    // lane, 157-159 on 2, 4 or 8): a few lanes as long as they fit in 64 GB together.
    // The headline class (bristlecone_70 1-32-1, 7x7 1-32-1: ~8e10 flop per launch)
    // fills the GPU from one lane.
    const double per_launch = plan.flops_per_pass / std::max<double>(1.0, (double)plan.steps.size());
    if (per_launch < 3e10)
      lanes = (int)std::min<double>(4.0, (65536.0 * 1048576.0) / arena_bytes);
  }
  lanes = (int)std::min<uint64_t>((uint64_t)std::max(lanes, 1), std::max<uint64_t>(1, slices / 2));
  // the memory limit bounds ALL scratch of the run (reference --memory,
  // src/contraction_utils.cpp:205-214), not one lane's arena
  lanes = (int)std::min<double>(lanes, std::floor((double)global::memory_limit /
                                                  std::max(arena_bytes, 1.0)));
  return std::max(lanes, 1);
}
}  // namespace

Session::Session(Plan plan, int device) : Session(std::move(plan), std::vector<int>{device}) {}

Session::Session(Plan plan, std::vector<int> devices)
    : plan_(std::move(plan)), devices_(std::move(devices)) {
  if (devices_.empty()) devices_.push_back(0);
  try {
    const double t0 = now_s();
    // One lane per device first (each must get its arena), then further lanes device
    // by device while they fit: lanes_[l * n_devices + d] is lane l of device d, so the
    // dealing in Contract() reaches every device before any second lane.
    for (const int d : devices_) {
      qfx_ctx* c = nullptr;
      check(qfx_create(d, &c), "qfx_create");
      lanes_.push_back(c);
      check(qfx_arena_reserve(c, plan_.arena_elems), "qfx_arena_reserve");
      check(qfx_amps_reserve(c, plan_.n_amplitudes), "qfx_amps_reserve");
    }
    ctx_ = lanes_[0];
    const int n_lanes = choose_lanes(plan_, (int)devices_.size());
    for (int l = 1; l < n_lanes; ++l) {
      // an extra round of lanes that cannot get its arenas is dropped, not fatal
      std::vector<qfx_ctx*> round;
      bool ok = true;
      for (const int d : devices_) {
        qfx_ctx* c = nullptr;
        if (qfx_create(d, &c) != QFX_OK) {
          ok = false;
          break;
        }
        round.push_back(c);
        if (qfx_arena_reserve(c, plan_.arena_elems) != QFX_OK ||
            qfx_amps_reserve(c, plan_.n_amplitudes) != QFX_OK) {
          ok = false;
          break;
        }
      }
      if (!ok) {
        for (qfx_ctx* c : round) qfx_destroy(c);
        break;
      }
      lanes_.insert(lanes_.end(), round.begin(), round.end());
    }
    programs_.assign(lanes_.size(), {});
    for (std::size_t l = 0; l < lanes_.size(); ++l) {
      check(qfx_bind_thread(lanes_[l]), "qfx_bind_thread");
      for (const ChainProgram& p : plan_.programs) {
        qfx_program* h = nullptr;
        check(qfx_program_create(lanes_[l], p.descs.data(), p.chain_first.data(),
                                 (int32_t)p.chain_first.size() - 1, p.terms.data(),
                                 p.term_first.data(), &h),
              "qfx_program_create");
        programs_[l].push_back(h);
      }
    }
    Upload();
    const double t1 = now_s();
    stats_.seconds_upload = t1 - t0;
    Flatten();
    stats_.seconds_flatten = now_s() - t1;
  } catch (...) {
    ReleaseDevice();
    throw;
  }
}

void Session::ReleaseDevice() {
  for (std::size_t l = 0; l < programs_.size() && l < lanes_.size(); ++l)
    for (qfx_program* p : programs_[l]) qfx_program_destroy(lanes_[l], p);
  programs_.clear();
  for (qfx_ctx* c : lanes_) qfx_destroy(c);
  lanes_.clear();
  ctx_ = nullptr;
}

Session::~Session() { ReleaseDevice(); }

void Session::Upload() {
  // The gate tensors sit next to each other at the start of the arena (plan.cpp):
  // one staged host buffer and ONE host->device copy instead of ~2000 copies of 32
  // to 128 bytes (20 ms of the end-to-end call at ~10 us per cudaMemcpy).
  if (plan_.uploads.empty()) return;
  uint64_t lo = UINT64_MAX, hi = 0;
  for (const auto& up : plan_.uploads) {
    lo = std::min<uint64_t>(lo, up.off);
    hi = std::max<uint64_t>(hi, up.off + up.data.size());
  }
  if (hi - lo > (uint64_t(1) << 24)) {  // scattered uploads: keep them separate
    for (qfx_ctx* c : lanes_) {
      check(qfx_bind_thread(c), "qfx_bind_thread");
      for (const auto& up : plan_.uploads)
        check(qfx_upload(c, up.off, reinterpret_cast<const float*>(up.data.data()),
                         up.data.size()),
              "qfx_upload");
    }
    return;
  }
  std::vector<s_type> staged(hi - lo, s_type(0.f, 0.f));
  for (const auto& up : plan_.uploads)
    std::copy(up.data.begin(), up.data.end(), staged.begin() + (up.off - lo));
  for (qfx_ctx* c : lanes_) {
    check(qfx_bind_thread(c), "qfx_bind_thread");
    check(qfx_upload(c, lo, reinterpret_cast<const float*>(staged.data()), staged.size()),
          "qfx_upload");
  }
}

void Session::SetFinalState(const std::string& final_state) {
  if (final_state.size() != plan_.final_delta_upload.size())
    throw ERROR_MSG("Size of final_conf: ", final_state.size(),
                    ", must be equal to size of initial_conf: ",
                    plan_.final_delta_upload.size(), ".");
  for (std::size_t a = 0; a < final_state.size(); ++a) {
    const int u = plan_.final_delta_upload[a];
    if (u < 0) continue;  // terminal cut: an output, whatever the string says there
    const char c = final_state[a];
    if (c != '0' && c != '1')
      throw ERROR_MSG("Final state has non-binary character ", c, " at index ", a,
                      "despite having no terminal cut there.");
    auto& data = plan_.uploads[u].data;
    if (data.size() != 2) throw ERROR_MSG("Final delta tensor of qubit ", a, " is malformed.");
    data[0] = s_type(c == '0' ? 1.f : 0.f, 0.f);
    data[1] = s_type(c == '1' ? 1.f : 0.f, 0.f);
  }
  Upload();
  Flatten();
}

void Session::SetCutValue(int cut, std::size_t value) {
  if (cut < 0 || cut >= (int)plan_.cuts.size() || plan_.cuts[cut].values.size() != 1)
    throw ERROR_MSG("Cut ", cut, " is not a single-valued projection.");
  plan_.cuts[cut].values[0] = value;
}

void Session::Flatten() {
  // replicated on every rank and lane: site tensors are MB-scale.  Launches are
  // asynchronous, so the lanes' flattens overlap on the device.
  for (qfx_ctx* c : lanes_) {
    check(qfx_bind_thread(c), "qfx_bind_thread");
    PrepareLane(c);
  }
  for (qfx_ctx* c : lanes_) {
    check(qfx_bind_thread(c), "qfx_bind_thread");
    check(qfx_sync(c), "qfx_sync");
  }
}

void Session::PrepareLane(qfx_ctx* ctx) {
  Plan no_cuts;
  Runner r(no_cuts, ctx, 0, 1);
  if (!plan_.flatten_parallel) {
    r.run(plan_.flatten_steps);
    return;
  }
  // every world-line product of every site in one launch, then the layout permutes
  std::vector<qfx_contract_desc> descs;
  std::vector<int32_t> first(plan_.n_chains + 1, 0);
  std::vector<PlanStep> permutes;
  descs.reserve(plan_.flatten_steps.size());
  for (const auto& st : plan_.flatten_steps) {  // steps are ordered by chain
    if (st.kind == PlanStep::CONTRACT) {
      descs.push_back(st.desc);
      first[st.chain + 1] = (int32_t)descs.size();
    } else {
      permutes.push_back(st);
    }
  }
  for (int c = 0; c < plan_.n_chains; ++c) first[c + 1] = std::max(first[c + 1], first[c]);
  check(qfx_contract_chains(ctx, descs.data(), first.data(), plan_.n_chains),
        "qfx_contract_chains");
  r.run(permutes);
}

void Session::ContractLane(int lane, int rank, int world, std::vector<double>* amps,
                           ExecStats* stats, std::vector<qfx_profile_row>* rows) {
  qfx_ctx* ctx = lanes_[lane];
  check(qfx_bind_thread(ctx), "qfx_bind_thread");
  check(qfx_amps_reserve(ctx, plan_.n_amplitudes), "qfx_amps_reserve");
  check(qfx_stats_reset(ctx), "qfx_stats_reset");
  if (profile_) check(qfx_profile_begin(ctx), "qfx_profile_begin");
  check(qfx_timer_start(ctx), "qfx_timer_start");
  Runner runner(plan_, ctx, rank, world, &programs_[lane]);
  runner.run(plan_.steps);
  float ms = 0.f;
  check(qfx_timer_stop_ms(ctx, &ms), "qfx_timer_stop_ms");
  if (profile_) {
    int32_t n = 0;
    rows->assign(256, qfx_profile_row{});
    check(qfx_profile_end(ctx, rows->data(), 256, &n), "qfx_profile_end");
    rows->resize(std::min<int32_t>(n, 256));
  }
  stats->device_ms_contract = ms;
  // a leaf of a plan with fused cut bits covers 2^fused_bits slices of the ordering
  stats->leaves = runner.leaves() << plan_.fused_bits;
  check(qfx_stats_get(ctx, &stats->device), "qfx_stats_get");
  amps->assign(2 * plan_.n_amplitudes, 0.0);
  check(qfx_amps_download(ctx, amps->data(), plan_.n_amplitudes), "qfx_amps_download");
}

void Session::Contract(int rank, int world,
                       std::vector<std::complex<double>>* amplitudes) {
  if (amplitudes == nullptr) throw ERROR_MSG("Amplitude return vector must be non-null.");
  if (world < 1 || rank < 0 || rank >= world)
    throw ERROR_MSG("Invalid rank ", rank, " of ", world, ".");
  const double t0 = now_s();
  const int L = (int)lanes_.size();
  std::vector<std::vector<double>> amps(L);
  std::vector<ExecStats> st(L);
  std::vector<std::vector<qfx_profile_row>> rows(L);
  std::vector<std::string> errors(L);
  if (L == 1) {
    ContractLane(0, rank, world, &amps[0], &st[0], &rows[0]);
  } else {
    std::vector<std::thread> threads;
    for (int l = 0; l < L; ++l)
      threads.emplace_back([&, l]() {
        try {
          ContractLane(l, l * world + rank, world * L, &amps[l], &st[l], &rows[l]);
        } catch (const std::string& msg) {
          errors[l] = msg.empty() ? std::string("unknown error") : msg;
        } catch (const std::exception& ex) {
          errors[l] = ex.what();
        }
      });
    for (auto& t : threads) t.join();
    for (const auto& e : errors)
      if (!e.empty()) throw e;
  }
  // lanes summed in lane order (deterministic)
  amplitudes->assign(plan_.n_amplitudes, {0.0, 0.0});
  const double up = stats_.seconds_upload, fl = stats_.seconds_flatten;
  stats_ = ExecStats{};
  stats_.seconds_upload = up;
  stats_.seconds_flatten = fl;
  std::map<std::vector<int>, qfx_profile_row> merged;
  // cut-aware flatten scales projected gate halves by powers of two (plan.cpp) so that
  // partial amplitudes of heavily sliced circuits stay inside fp32's range; undone
  // here, exactly, in fp64
  const double unscale = std::ldexp(1.0, -plan_.amp_scale_log2);
  for (int l = 0; l < L; ++l) {
    for (uint64_t a = 0; a < plan_.n_amplitudes; ++a)
      (*amplitudes)[a] +=
          std::complex<double>(amps[l][2 * a] * unscale, amps[l][2 * a + 1] * unscale);
    stats_.device_ms_contract = std::max(stats_.device_ms_contract, st[l].device_ms_contract);
    stats_.leaves += st[l].leaves;
    const qfx_stats& d = st[l].device;
    stats_.device.launches_total += d.launches_total;
    stats_.device.launches_gemm_ffma += d.launches_gemm_ffma;
    stats_.device.launches_gemm_tc += d.launches_gemm_tc;
    stats_.device.launches_reduce += d.launches_reduce;
    stats_.device.launches_permute += d.launches_permute;
    stats_.device.launches_copy += d.launches_copy;
    stats_.device.launches_gemm_tc_fused += d.launches_gemm_tc_fused;
    stats_.device.launches_gemm_tc_pair += d.launches_gemm_tc_pair;
    stats_.device.flops += d.flops;
    stats_.device.bytes += d.bytes;
    for (const auto& r : rows[l]) {
      auto& m = merged[{r.kind, r.m_bits, r.n_bits, r.k_bits}];
      if (m.launches == 0) m = r;
      else {
        m.launches += r.launches;
        m.ms += r.ms;
        m.flops += r.flops;
        m.bytes += r.bytes;
      }
    }
  }
  profile_rows_.clear();
  for (const auto& kv : merged) profile_rows_.push_back(kv.second);
  stats_.seconds_contract = now_s() - t0;
}

}  // namespace qflex
