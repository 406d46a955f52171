// Plan executor: walks the cut tree and issues device steps through the C ABI.
// The recursion mirrors ContractionData::ContractGrid
// (reference src/contraction_utils.cpp:223-354): steps before a cut run once,
// everything after it runs once per cut value; a bond cut adds into the same
// amplitude, a terminal cut selects amplitude output_index*|values|+i.
#include <algorithm>
#include <chrono>
#include <cstring>

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
  Runner(const Plan& plan, qfx_ctx* ctx, int rank, int world)
      : plan_(plan), ctx_(ctx), rank_(rank), world_(world),
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
    walk(steps, 0, 0, 0, 0);
  }

 private:
  uint64_t resolve(uint64_t base, const std::vector<VarTerm>& var) const {
    for (const auto& v : var) base += (uint64_t)values_[v.cut] << v.shift;
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
      case PlanStep::PERMUTE:
        check(qfx_permute(ctx_, st.src, st.dst, st.n_bits, st.dst_pos), "qfx_permute");
        break;
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
    if (cut.terminal) amp_index *= cut.values.size();
    for (std::size_t v = 0; v < cut.values.size(); ++v) {
      const uint64_t next_combo = combo * cut.values.size() + v;
      if (level + 1 == share_levels_ && world_ > 1 &&
          (int)(next_combo % (uint64_t)world_) != rank_)
        continue;
      values_[level] = cut.values[v];
      walk(steps, i, level + 1, cut.terminal ? amp_index + v : amp_index,
           level < share_levels_ ? next_combo : combo);
    }
  }

  const Plan& plan_;
  qfx_ctx* ctx_;
  int rank_, world_;
  int share_levels_ = 0;
  std::vector<std::size_t> values_;
  uint64_t leaves_ = 0;
};

}  // namespace

Session::Session(Plan plan, int device) : plan_(std::move(plan)) {
  check(qfx_create(device, &ctx_), "qfx_create");
  try {
    const double t0 = now_s();
    check(qfx_arena_reserve(ctx_, plan_.arena_elems), "qfx_arena_reserve");
    check(qfx_amps_reserve(ctx_, plan_.n_amplitudes), "qfx_amps_reserve");
    Upload();
    const double t1 = now_s();
    stats_.seconds_upload = t1 - t0;
    Flatten();
    stats_.seconds_flatten = now_s() - t1;
  } catch (...) {
    qfx_destroy(ctx_);
    ctx_ = nullptr;
    throw;
  }
}

Session::~Session() {
  if (ctx_) qfx_destroy(ctx_);
}

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
    for (const auto& up : plan_.uploads)
      check(qfx_upload(ctx_, up.off, reinterpret_cast<const float*>(up.data.data()),
                       up.data.size()),
            "qfx_upload");
    return;
  }
  std::vector<s_type> staged(hi - lo, s_type(0.f, 0.f));
  for (const auto& up : plan_.uploads)
    std::copy(up.data.begin(), up.data.end(), staged.begin() + (up.off - lo));
  check(qfx_upload(ctx_, lo, reinterpret_cast<const float*>(staged.data()), staged.size()),
        "qfx_upload");
}

void Session::Flatten() {
  // replicated on every rank: site tensors are MB-scale
  Plan no_cuts;
  Runner r(no_cuts, ctx_, 0, 1);
  if (!plan_.flatten_parallel) {
    r.run(plan_.flatten_steps);
    check(qfx_sync(ctx_), "qfx_sync");
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
  check(qfx_contract_chains(ctx_, descs.data(), first.data(), plan_.n_chains),
        "qfx_contract_chains");
  r.run(permutes);
  check(qfx_sync(ctx_), "qfx_sync");
}

void Session::Contract(int rank, int world,
                       std::vector<std::complex<double>>* amplitudes) {
  if (amplitudes == nullptr) throw ERROR_MSG("Amplitude return vector must be non-null.");
  if (world < 1 || rank < 0 || rank >= world)
    throw ERROR_MSG("Invalid rank ", rank, " of ", world, ".");
  const double t0 = now_s();
  check(qfx_amps_reserve(ctx_, plan_.n_amplitudes), "qfx_amps_reserve");
  check(qfx_stats_reset(ctx_), "qfx_stats_reset");
  if (profile_) check(qfx_profile_begin(ctx_), "qfx_profile_begin");
  check(qfx_timer_start(ctx_), "qfx_timer_start");
  Runner runner(plan_, ctx_, rank, world);
  runner.run(plan_.steps);
  float ms = 0.f;
  check(qfx_timer_stop_ms(ctx_, &ms), "qfx_timer_stop_ms");
  if (profile_) {
    int32_t n = 0;
    profile_rows_.assign(256, qfx_profile_row{});
    check(qfx_profile_end(ctx_, profile_rows_.data(), 256, &n), "qfx_profile_end");
    profile_rows_.resize(std::min<int32_t>(n, 256));
  }
  stats_.device_ms_contract = ms;
  stats_.leaves = runner.leaves();
  check(qfx_stats_get(ctx_, &stats_.device), "qfx_stats_get");
  std::vector<double> host(2 * plan_.n_amplitudes);
  check(qfx_amps_download(ctx_, host.data(), plan_.n_amplitudes), "qfx_amps_download");
  amplitudes->assign(plan_.n_amplitudes, {0.0, 0.0});
  for (uint64_t a = 0; a < plan_.n_amplitudes; ++a)
    (*amplitudes)[a] = {host[2 * a], host[2 * a + 1]};
  stats_.seconds_contract = now_s() - t0;
}

}  // namespace qflex
