// C ABI of libqflex_b200.so: context, arena, transfers, amplitude table.
// See include/qflex_b200.h for the reference interfaces each entry replaces.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <string>
#include <mutex>
#include <vector>

#include "qfx_internal.h"

namespace {
thread_local std::string g_last_error = "";
}

void qfx_set_error(const std::string& msg) { g_last_error = msg; }

int qfx_fail_cuda(cudaError_t e, const char* what) {
  g_last_error = std::string("CUDA error: ") + cudaGetErrorString(e) + " in " + what;
  return QFX_ERR_CUDA;
}

// ---- device-buffer cache.  A destroyed context parks its arena and workspace
// here (one of each per device) and the next context of the process takes them
// over when they are big enough: repeated EvaluateCircuit calls (one per output
// bitstring in the reference's Python interface, src/pybind_main.cpp:75-138) then
// skip a multi-GB cudaMalloc + cudaFree per call.  qfx_trim_cache() frees them.
namespace {
struct ParkedBuffers {
  void* arena = nullptr;
  size_t arena_bytes = 0;
  void* ws = nullptr;
  size_t ws_bytes = 0;
};
std::mutex g_park_mu;
ParkedBuffers g_parked[64];

void* take_parked(int device, bool arena, size_t need, size_t* got) {
  if (device < 0 || device >= 64) return nullptr;
  std::lock_guard<std::mutex> lock(g_park_mu);
  ParkedBuffers& p = g_parked[device];
  void*& ptr = arena ? p.arena : p.ws;
  size_t& bytes = arena ? p.arena_bytes : p.ws_bytes;
  if (ptr == nullptr || bytes < need) return nullptr;
  void* out = ptr;
  *got = bytes;
  ptr = nullptr;
  bytes = 0;
  return out;
}

// keeps the larger of (parked, offered); frees the other.  Device must be current.
void park(int device, bool arena, void* ptr, size_t bytes) {
  if (ptr == nullptr) return;
  if (device < 0 || device >= 64) {
    cudaFree(ptr);
    return;
  }
  void* drop = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_park_mu);
    ParkedBuffers& p = g_parked[device];
    void*& slot = arena ? p.arena : p.ws;
    size_t& slot_bytes = arena ? p.arena_bytes : p.ws_bytes;
    if (slot == nullptr || slot_bytes < bytes) {
      drop = slot;
      slot = ptr;
      slot_bytes = bytes;
    } else {
      drop = ptr;
    }
  }
  if (drop) cudaFree(drop);
}
}  // namespace

extern "C" int qfx_trim_cache(int device) {
  if (device < 0 || device >= 64) return QFX_ERR_ARG;
  ParkedBuffers p;
  {
    std::lock_guard<std::mutex> lock(g_park_mu);
    p = g_parked[device];
    g_parked[device] = ParkedBuffers{};
  }
  if (p.arena || p.ws) {
    cudaSetDevice(device);
    if (p.arena) cudaFree(p.arena);
    if (p.ws) cudaFree(p.ws);
  }
  return QFX_OK;
}

int qfx_ensure_ws(qfx_ctx* ctx, size_t bytes) {
  if (bytes <= ctx->ws_bytes) return QFX_OK;
  // The workspace may still be in use by earlier launches on the stream.
  QFX_CUDA(cudaStreamSynchronize(ctx->stream));
  if (ctx->ws) QFX_CUDA(cudaFree(ctx->ws));
  ctx->ws = nullptr;
  ctx->ws_bytes = 0;
  size_t want = bytes < (1u << 20) ? (1u << 20) : bytes;
  size_t got = 0;
  if (void* p = take_parked(ctx->device, false, want, &got)) {
    ctx->ws = p;
    ctx->ws_bytes = got;
    return QFX_OK;
  }
  cudaError_t e = cudaMalloc(&ctx->ws, want);
  if (e != cudaSuccess) {
    // a parked buffer of an earlier context may be what is in the way
    cudaGetLastError();
    qfx_trim_cache(ctx->device);
    e = cudaMalloc(&ctx->ws, want);
  }
  if (e != cudaSuccess) {
    cudaGetLastError();
    ctx->ws = nullptr;
    qfx_set_error("out of device memory allocating the split-K workspace");
    return QFX_ERR_OOM;
  }
  ctx->ws_bytes = want;
  return QFX_OK;
}

namespace {
cudaEvent_t prof_event(qfx_ctx* ctx) {
  if (!ctx->ev_pool.empty()) {
    cudaEvent_t e = ctx->ev_pool.back();
    ctx->ev_pool.pop_back();
    return e;
  }
  cudaEvent_t e = nullptr;
  cudaEventCreate(&e);
  return e;
}
}  // namespace

qfx_prof_scope::qfx_prof_scope(qfx_ctx* c, int kind, int m, int n, int k, double flops,
                               double bytes)
    : ctx(c) {
  if (ctx == nullptr || !ctx->profiling) {
    ctx = nullptr;
    return;
  }
  qfx_ctx::ProfRec r{kind, m, n, k, flops, bytes, prof_event(c), prof_event(c)};
  cudaEventRecord(r.e0, c->stream);
  c->prof.push_back(r);
}

qfx_prof_scope::~qfx_prof_scope() {
  if (ctx != nullptr) cudaEventRecord(ctx->prof.back().e1, ctx->stream);
}

extern "C" {

int qfx_profile_begin(qfx_ctx* ctx) {
  if (ctx == nullptr) return QFX_ERR_ARG;
  ctx->profiling = true;
  return QFX_OK;
}

int qfx_profile_end(qfx_ctx* ctx, qfx_profile_row* rows, int32_t cap, int32_t* n_out) {
  if (ctx == nullptr || n_out == nullptr) return QFX_ERR_ARG;
  ctx->profiling = false;
  QFX_CUDA(cudaStreamSynchronize(ctx->stream));
  std::vector<qfx_profile_row> agg;
  for (auto& r : ctx->prof) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, r.e0, r.e1);
    ctx->ev_pool.push_back(r.e0);
    ctx->ev_pool.push_back(r.e1);
    qfx_profile_row* hit = nullptr;
    for (auto& a : agg)
      if (a.kind == r.kind && a.m_bits == r.m && a.n_bits == r.n && a.k_bits == r.k) hit = &a;
    if (hit == nullptr) {
      agg.push_back(qfx_profile_row{r.kind, r.m, r.n, r.k, 0, 0.0, 0.0, 0.0});
      hit = &agg.back();
    }
    hit->launches++;
    hit->ms += ms;
    hit->flops += r.flops;
    hit->bytes += r.bytes;
  }
  ctx->prof.clear();
  for (size_t i = 0; i < agg.size(); ++i)
    for (size_t j = i + 1; j < agg.size(); ++j)
      if (agg[j].ms > agg[i].ms) std::swap(agg[i], agg[j]);
  *n_out = (int32_t)agg.size();
  for (int32_t i = 0; i < cap && i < (int32_t)agg.size(); ++i)
    if (rows != nullptr) rows[i] = agg[i];
  return QFX_OK;
}

const char* qfx_last_error(void) { return g_last_error.c_str(); }

int qfx_device_count(int* n_out) {
  if (n_out == nullptr) {
    qfx_set_error("qfx_device_count: null argument");
    return QFX_ERR_ARG;
  }
  int n = 0;
  QFX_CUDA(cudaGetDeviceCount(&n));
  *n_out = n;
  return QFX_OK;
}

int qfx_device_memory(int device, uint64_t* free_bytes_out, uint64_t* total_bytes_out) {
  if (free_bytes_out == nullptr || total_bytes_out == nullptr) {
    qfx_set_error("qfx_device_memory: null argument");
    return QFX_ERR_ARG;
  }
  int prev = 0;
  QFX_CUDA(cudaGetDevice(&prev));
  if (device >= 0 && device != prev) QFX_CUDA(cudaSetDevice(device));
  size_t f = 0, t = 0;
  cudaError_t e = cudaMemGetInfo(&f, &t);
  if (device >= 0 && device != prev) cudaSetDevice(prev);
  if (e != cudaSuccess) return qfx_fail_cuda(e, "cudaMemGetInfo");
  *free_bytes_out = f;
  *total_bytes_out = t;
  return QFX_OK;
}

int qfx_create(int device, qfx_ctx** ctx_out) {
  if (ctx_out == nullptr) {
    qfx_set_error("qfx_create: null argument");
    return QFX_ERR_ARG;
  }
  *ctx_out = nullptr;
  int n = 0;
  QFX_CUDA(cudaGetDeviceCount(&n));
  if (device < 0 || device >= n) {
    qfx_set_error("qfx_create: no such CUDA device (this library has no CPU path)");
    return QFX_ERR_ARG;
  }
  QFX_CUDA(cudaSetDevice(device));
  // cudaGetDeviceProperties costs ~20 ms per call; two attributes are all we need,
  // cached per device (a context is created per EvaluateCircuit call)
  static std::mutex attr_mu;
  static int attr_major[64], attr_sms[64];
  static bool attr_known[64];
  int major = 0, sms = 0;
  {
    std::lock_guard<std::mutex> lock(attr_mu);
    if (device < 64 && attr_known[device]) {
      major = attr_major[device];
      sms = attr_sms[device];
    } else {
      QFX_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
      QFX_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
      if (device < 64) {
        attr_major[device] = major;
        attr_sms[device] = sms;
        attr_known[device] = true;
      }
    }
  }
  if (major < 10) {
    qfx_set_error("qfx_create: the device is not sm_100-class; this library is built for "
                  "sm_100a only");
    return QFX_ERR_ARG;
  }
  qfx_ctx* ctx = new (std::nothrow) qfx_ctx();
  if (ctx == nullptr) {
    qfx_set_error("qfx_create: host allocation failed");
    return QFX_ERR_OOM;
  }
  ctx->device = device;
  ctx->sm_count = sms;
  if (const char* e = getenv("QFLEX_TC_CHUNK")) ctx->tc_chunk = atoi(e);
  if (const char* e = getenv("QFLEX_TC_DEBIAS")) ctx->tc_debias = atoi(e) != 0;
  cudaError_t e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
  if (e == cudaSuccess) e = cudaEventCreate(&ctx->ev0);
  if (e == cudaSuccess) e = cudaEventCreate(&ctx->ev1);
  if (e != cudaSuccess) {
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
    return qfx_fail_cuda(e, "qfx_create (stream / events)");
  }
  *ctx_out = ctx;
  return QFX_OK;
}

int qfx_destroy(qfx_ctx* ctx) {
  if (ctx == nullptr) return QFX_OK;
  cudaSetDevice(ctx->device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  park(ctx->device, true, ctx->arena, ctx->arena_elems * sizeof(float2));
  if (ctx->amps) cudaFree(ctx->amps);
  park(ctx->device, false, ctx->ws, ctx->ws_bytes);
  if (ctx->ev0) cudaEventDestroy(ctx->ev0);
  if (ctx->ev1) cudaEventDestroy(ctx->ev1);
  for (auto& r : ctx->prof) {
    cudaEventDestroy(r.e0);
    cudaEventDestroy(r.e1);
  }
  for (auto e : ctx->ev_pool) cudaEventDestroy(e);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
  return QFX_OK;
}

int qfx_bind_thread(qfx_ctx* ctx) {
  if (ctx == nullptr) {
    qfx_set_error("qfx_bind_thread: null context");
    return QFX_ERR_ARG;
  }
  QFX_CUDA(cudaSetDevice(ctx->device));
  return QFX_OK;
}

int qfx_sync(qfx_ctx* ctx) {
  if (ctx == nullptr) {
    qfx_set_error("qfx_sync: null context");
    return QFX_ERR_ARG;
  }
  QFX_CUDA(cudaStreamSynchronize(ctx->stream));
  return QFX_OK;
}

int qfx_arena_reserve(qfx_ctx* ctx, uint64_t n_elems) {
  if (ctx == nullptr) {
    qfx_set_error("qfx_arena_reserve: null context");
    return QFX_ERR_ARG;
  }
  QFX_CUDA(cudaSetDevice(ctx->device));
  if (n_elems <= ctx->arena_elems) return QFX_OK;
  QFX_CUDA(cudaStreamSynchronize(ctx->stream));
  if (ctx->arena) QFX_CUDA(cudaFree(ctx->arena));
  ctx->arena = nullptr;
  ctx->arena_elems = 0;
  size_t got = 0;
  if (void* p = take_parked(ctx->device, true, n_elems * sizeof(float2), &got)) {
    ctx->arena = reinterpret_cast<float2*>(p);
    ctx->arena_elems = got / sizeof(float2);
    return QFX_OK;
  }
  cudaError_t e = cudaMalloc(&ctx->arena, n_elems * sizeof(float2));
  if (e != cudaSuccess) {
    // a parked buffer that was too small may be what is in the way
    cudaGetLastError();
    qfx_trim_cache(ctx->device);
    e = cudaMalloc(&ctx->arena, n_elems * sizeof(float2));
  }
  if (e != cudaSuccess) {
    cudaGetLastError();
    char buf[160];
    snprintf(buf, sizeof(buf),
             "out of device memory: arena of %.2f GB does not fit this GPU",
             (double)n_elems * 8.0 / (1024.0 * 1024.0 * 1024.0));
    qfx_set_error(buf);
    return QFX_ERR_OOM;
  }
  ctx->arena_elems = n_elems;
  return QFX_OK;
}

int qfx_arena_size(qfx_ctx* ctx, uint64_t* n_elems_out) {
  if (ctx == nullptr || n_elems_out == nullptr) {
    qfx_set_error("qfx_arena_size: null argument");
    return QFX_ERR_ARG;
  }
  *n_elems_out = ctx->arena_elems;
  return QFX_OK;
}

int qfx_upload(qfx_ctx* ctx, uint64_t off, const float* host, uint64_t n_elems) {
  if (ctx == nullptr || host == nullptr || off + n_elems > ctx->arena_elems) {
    qfx_set_error("qfx_upload: bad argument or span outside the arena");
    return QFX_ERR_ARG;
  }
  QFX_CUDA(cudaMemcpyAsync(ctx->arena + off, host, n_elems * sizeof(float2),
                           cudaMemcpyHostToDevice, ctx->stream));
  // pageable host memory: the copy is staged before return, but keep the
  // contract simple -- the host buffer may be reused once this returns.
  QFX_CUDA(cudaStreamSynchronize(ctx->stream));
  return QFX_OK;
}

int qfx_download(qfx_ctx* ctx, uint64_t off, float* host, uint64_t n_elems) {
  if (ctx == nullptr || host == nullptr || off + n_elems > ctx->arena_elems) {
    qfx_set_error("qfx_download: bad argument or span outside the arena");
    return QFX_ERR_ARG;
  }
  QFX_CUDA(cudaMemcpyAsync(host, ctx->arena + off, n_elems * sizeof(float2),
                           cudaMemcpyDeviceToHost, ctx->stream));
  QFX_CUDA(cudaStreamSynchronize(ctx->stream));
  return QFX_OK;
}

int qfx_copy(qfx_ctx* ctx, uint64_t src_off, uint64_t dst_off, uint64_t n_elems) {
  if (ctx == nullptr || src_off + n_elems > ctx->arena_elems ||
      dst_off + n_elems > ctx->arena_elems) {
    qfx_set_error("qfx_copy: bad argument or span outside the arena");
    return QFX_ERR_ARG;
  }
  if (n_elems == 0 || src_off == dst_off) return QFX_OK;
  QFX_CUDA(cudaMemcpyAsync(ctx->arena + dst_off, ctx->arena + src_off,
                           n_elems * sizeof(float2), cudaMemcpyDeviceToDevice,
                           ctx->stream));
  ctx->stats.launches_copy++;
  return QFX_OK;
}

int qfx_contract(qfx_ctx* ctx, const qfx_contract_desc* desc) {
  return qfx_launch_contract(ctx, desc, false, 0);
}

int qfx_dot_accumulate(qfx_ctx* ctx, const qfx_contract_desc* desc, uint64_t slot) {
  return qfx_launch_contract(ctx, desc, true, slot);
}

int qfx_permute(qfx_ctx* ctx, uint64_t src_off, uint64_t dst_off, int32_t n_bits,
                const int8_t* dst_pos) {
  qfx_prof_scope scope(ctx, 1, n_bits, 0, 0, 0.0,
                       16.0 * (double)(1ull << (n_bits < 0 ? 0 : n_bits > 40 ? 40 : n_bits)));
  return qfx_launch_permute(ctx, src_off, dst_off, n_bits, dst_pos);
}

int qfx_amps_reserve(qfx_ctx* ctx, uint64_t n_amps) {
  if (ctx == nullptr || n_amps == 0) {
    qfx_set_error("qfx_amps_reserve: bad argument");
    return QFX_ERR_ARG;
  }
  QFX_CUDA(cudaSetDevice(ctx->device));
  if (n_amps > ctx->n_amps) {
    QFX_CUDA(cudaStreamSynchronize(ctx->stream));
    if (ctx->amps) QFX_CUDA(cudaFree(ctx->amps));
    ctx->amps = nullptr;
    ctx->n_amps = 0;
    cudaError_t e = cudaMalloc(&ctx->amps, n_amps * sizeof(double2));
    if (e != cudaSuccess) {
      cudaGetLastError();
      qfx_trim_cache(ctx->device);
      e = cudaMalloc(&ctx->amps, n_amps * sizeof(double2));
    }
    if (e != cudaSuccess) {
      cudaGetLastError();
      ctx->amps = nullptr;
      qfx_set_error("out of device memory allocating the amplitude table");
      return QFX_ERR_OOM;
    }
    ctx->n_amps = n_amps;
  }
  QFX_CUDA(cudaMemsetAsync(ctx->amps, 0, ctx->n_amps * sizeof(double2), ctx->stream));
  return QFX_OK;
}

int qfx_amps_download(qfx_ctx* ctx, double* host, uint64_t n_amps) {
  if (ctx == nullptr || host == nullptr || n_amps > ctx->n_amps) {
    qfx_set_error("qfx_amps_download: bad argument");
    return QFX_ERR_ARG;
  }
  QFX_CUDA(cudaMemcpyAsync(host, ctx->amps, n_amps * sizeof(double2),
                           cudaMemcpyDeviceToHost, ctx->stream));
  QFX_CUDA(cudaStreamSynchronize(ctx->stream));
  return QFX_OK;
}

int qfx_timer_start(qfx_ctx* ctx) {
  if (ctx == nullptr) return QFX_ERR_ARG;
  QFX_CUDA(cudaEventRecord(ctx->ev0, ctx->stream));
  return QFX_OK;
}

int qfx_timer_stop_ms(qfx_ctx* ctx, float* ms_out) {
  if (ctx == nullptr || ms_out == nullptr) return QFX_ERR_ARG;
  QFX_CUDA(cudaEventRecord(ctx->ev1, ctx->stream));
  QFX_CUDA(cudaEventSynchronize(ctx->ev1));
  QFX_CUDA(cudaEventElapsedTime(ms_out, ctx->ev0, ctx->ev1));
  return QFX_OK;
}

int qfx_stats_get(qfx_ctx* ctx, qfx_stats* out) {
  if (ctx == nullptr || out == nullptr) return QFX_ERR_ARG;
  *out = ctx->stats;
  return QFX_OK;
}

int qfx_stats_reset(qfx_ctx* ctx) {
  if (ctx == nullptr) return QFX_ERR_ARG;
  ctx->stats = qfx_stats{};
  return QFX_OK;
}

}  // extern "C"
