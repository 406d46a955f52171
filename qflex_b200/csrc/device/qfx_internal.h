// Internal declarations shared by the translation units of libqflex_b200.so.
// Nothing here is part of the ABI (see include/qflex_b200.h).
#ifndef QFX_INTERNAL_H_
#define QFX_INTERNAL_H_

#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <string>
#include <vector>

#include "qflex_b200.h"

struct qfx_ctx {
  int device = 0;
  int sm_count = 148;
  cudaStream_t stream = nullptr;
  float2* arena = nullptr;
  uint64_t arena_elems = 0;
  double2* amps = nullptr;
  uint64_t n_amps = 0;
  // split-K / reduction workspace (separate from the arena, grows on demand)
  void* ws = nullptr;
  size_t ws_bytes = 0;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  qfx_stats stats{};
  // per-launch profiling (qfx_profile_begin/end)
  int tc_chunk = 0;  // k blocks per TMEM accumulation chunk (0 = default)
  bool tc_debias = true;  // compensate the tensor core's accumulation bias (tc_common.cuh)
  bool profiling = false;
  struct ProfRec {
    int kind, m, n, k;
    double flops, bytes;
    cudaEvent_t e0, e1;
  };
  std::vector<ProfRec> prof;
  std::vector<cudaEvent_t> ev_pool;
};

// RAII scope: records an event pair around the launches issued inside it.
struct qfx_prof_scope {
  qfx_ctx* ctx;
  qfx_prof_scope(qfx_ctx* c, int kind, int m, int n, int k, double flops, double bytes);
  ~qfx_prof_scope();
};

// cudaFuncSetAttribute is per DEVICE: a once-flag per device (bit d = device d done).
// Two threads may both set the attribute once; that is harmless.
struct qfx_per_device_once {
  std::atomic<uint64_t> done{0};
  bool need(int device) const {
    return device < 0 || device >= 64 || !((done.load(std::memory_order_acquire) >> device) & 1u);
  }
  void mark(int device) {
    if (device >= 0 && device < 64) done.fetch_or(uint64_t(1) << device, std::memory_order_release);
  }
};

void qfx_set_error(const std::string& msg);
int qfx_fail_cuda(cudaError_t e, const char* what);
int qfx_ensure_ws(qfx_ctx* ctx, size_t bytes);

#define QFX_CUDA(expr)                                   \
  do {                                                   \
    cudaError_t _e = (expr);                             \
    if (_e != cudaSuccess) return qfx_fail_cuda(_e, #expr); \
  } while (0)

// contract.cu
int qfx_launch_contract(qfx_ctx* ctx, const qfx_contract_desc* d,
                        bool accumulate_amp, uint64_t slot);
// tc_gemm.cu (tcgen05 split-bf16 path; d has roles swapped so m_bits >= n_bits)
bool qfx_tc_eligible(const qfx_contract_desc& d);
int qfx_launch_contract_tc(qfx_ctx* ctx, const qfx_contract_desc& d, const float2* A,
                           const float2* B, float2* C);
// tc_fused.cu (same contract; *done = false -> not applicable, use the pre-split path)
int qfx_launch_contract_tc_fused(qfx_ctx* ctx, const qfx_contract_desc& d, const float2* A,
                                 const float2* B, float2* C, bool* done);
// tc_pair.cu (CTA-pair variant for k, n <= 8 bits and >= 74 256-row tiles; same contract)
int qfx_launch_contract_tc_pair(qfx_ctx* ctx, const qfx_contract_desc& d, const float2* A,
                                const float2* B, float2* C, bool* done);
int qfx_launch_contract_tc_pairlong(qfx_ctx* ctx, const qfx_contract_desc& d, const float2* A,
                                    const float2* B, float2* C, bool* done);
int qfx_launch_tc_splitk_finish(qfx_ctx* ctx, const float2* part, float2* C,
                                const qfx_contract_desc& d, int ksplit, int c_is_natural);
// permute.cu
int qfx_launch_permute(qfx_ctx* ctx, uint64_t src_off, uint64_t dst_off,
                       int n_bits, const int8_t* dst_pos);
int qfx_launch_permute_ptr(qfx_ctx* ctx, const float2* src, float2* dst, int n_bits,
                           const int8_t* dst_pos);

#endif  // QFX_INTERNAL_H_
