/*
 * qflex_b200.h -- C ABI of the B200 contraction engine (libqflex_b200.so).
 *
 * The reference (ngnrsaa/qflex) has no FFI/plugin seam; its numeric hot path is
 * the C++ call boundary listed below.  This header is the thin C-ABI layer the
 * new host code (qflex_b200/csrc/host) calls instead; each entry point cites the
 * reference interface it replaces (paths relative to the reference repo).
 *
 * Conventions
 *   - plain pointers and sizes only; no C++/torch types cross this boundary;
 *   - every function returns 0 on success and a non-zero status on failure; the
 *     message is available from qfx_last_error() (thread-local).  The host layer
 *     turns it back into the reference's `throw std::string` convention
 *     (src/errors.h:7-9);
 *   - a tensor is a span of complex64 (2 floats, re then im) inside ONE device
 *     arena per context; positions are element offsets into that arena.  Device
 *     pointers never escape;
 *   - every tensor dimension is a power of two (true for every tensor qFlex
 *     builds: bond dims are 2^#cz or 4^#fsim, src/read_circuit.cpp:146-165), so
 *     a tensor is addressed by the BITS of its linear index.  Bit 0 is the
 *     fastest-varying (row-major, last index fastest, as src/tensor.h:34-60);
 *   - calls are asynchronous on the context's stream; qfx_sync() /
 *     qfx_download() / qfx_amps_download() synchronise.
 */
#ifndef QFLEX_B200_H_
#define QFLEX_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Capacity of the per-bit position tables.  An operand may have at most
 * QFX_MAX_OPERAND_BITS index bits (2^32 elements = 32 GB of complex64): every kernel
 * computes element offsets in 32 bits plus a 64-bit base. */
#define QFX_MAX_BITS 40
#define QFX_MAX_OPERAND_BITS 32
#define QFX_OK 0
#define QFX_ERR_CUDA 1
#define QFX_ERR_ARG 2
#define QFX_ERR_OOM 3

typedef struct qfx_ctx qfx_ctx;

/* Thread-local message of the last failing call (never NULL). */
const char* qfx_last_error(void);

/* Number of visible CUDA devices (fails loudly when there is no GPU/driver:
 * there is no CPU fallback anywhere in this library). */
int qfx_device_count(int* n_out);

/* Free and total memory of `device` in bytes (device < 0: the calling thread's current
 * device).  Replaces what --track-memory reports in the reference, the process'
 * resident memory (src/memory.h:27-83): here the scratch lives in HBM. */
int qfx_device_memory(int device, uint64_t* free_bytes_out, uint64_t* total_bytes_out);

/* One context = one GPU + one stream + one arena + one amplitude table.
 * Replaces the process-global state of the reference (global::*, _REORDER_MAPS,
 * src/tensor.cpp:83) with an explicit handle. */
int qfx_create(int device, qfx_ctx** ctx_out);
int qfx_destroy(qfx_ctx* ctx);
/* A destroyed context parks its arena and workspace (one of each per device)
 * for the next context of the process; this frees them.  (The reference
 * allocates its scratch once per EvaluateCircuit, src/contraction_utils.cpp:
 * 16-221; parking removes the per-call cudaMalloc/cudaFree of the arena.) */
int qfx_trim_cache(int device);
/* Makes the context's device current for the calling host thread.  A context may
 * be driven from any ONE thread at a time; a thread other than its creator calls
 * this first (the host layer runs one thread per lane, see Session in plan.h). */
int qfx_bind_thread(qfx_ctx* ctx);
int qfx_sync(qfx_ctx* ctx);

/* Arena of n_elems complex64.  Replaces the scratch arena that
 * ContractionData::Initialize sizes and allocates
 * (src/contraction_utils.cpp:117-218) and scratch_2D
 * (src/evaluate_circuit.cpp:207).  Re-reserving a larger size reallocates and
 * discards the contents. */
int qfx_arena_reserve(qfx_ctx* ctx, uint64_t n_elems);
int qfx_arena_size(qfx_ctx* ctx, uint64_t* n_elems_out);

/* Host <-> arena.  Replaces Tensor(indices, dims, data) construction and
 * Tensor::data() reads (src/tensor.h:62-118). */
int qfx_upload(qfx_ctx* ctx, uint64_t off, const float* host_re_im,
               uint64_t n_elems);
int qfx_download(qfx_ctx* ctx, uint64_t off, float* host_re_im,
                 uint64_t n_elems);

/* Flat device copy.  Replaces Tensor::operator= / _copy
 * (src/tensor.cpp:107-139,197-213) as used for patch seeding and
 * merge-into-empty (src/contraction_utils.cpp:238,323). */
int qfx_copy(qfx_ctx* ctx, uint64_t src_off, uint64_t dst_off,
             uint64_t n_elems);

/* Pairwise contraction C[l,r] = sum_c A[l,c] * B[c,r] with all three operands
 * addressed through per-bit position tables, so the two Tensor::reorder passes
 * of the reference are folded into the operand gathers.
 * Replaces qflex::multiply (src/tensor.cpp:1045-1194) = A.reorder + B.reorder
 * (src/tensor.cpp:866-911) + cblas_cgemm / cgemv / cdotu
 * (src/tensor.cpp:960-1042).
 *
 *   m (m_bits bits) enumerates the free indices of A, n the free indices of B,
 *   k the contracted indices.  For bit i of m: it is bit a_m[i] of A's linear
 *   offset and bit c_m[i] of C's.  Likewise b_n/c_n for n and a_k/b_k for k.
 *   All listed positions inside one operand must be distinct and together cover
 *   0..bits-1 of that operand.  Operands must not overlap C.
 */
typedef struct qfx_contract_desc {
  uint64_t off_a, off_b, off_c; /* element offsets into the arena */
  int32_t m_bits, n_bits, k_bits;
  int8_t a_m[QFX_MAX_BITS], c_m[QFX_MAX_BITS];
  int8_t b_n[QFX_MAX_BITS], c_n[QFX_MAX_BITS];
  int8_t a_k[QFX_MAX_BITS], b_k[QFX_MAX_BITS];
  int32_t precision; /* QFX_PREC_* */
} qfx_contract_desc;

#define QFX_PREC_AUTO 0 /* class by arithmetic intensity */
#define QFX_PREC_FFMA 1 /* fp32 FFMA everywhere */
#define QFX_PREC_BF16X3 2 /* tcgen05 split-bf16 emulation of fp32 where legal */

int qfx_contract(qfx_ctx* ctx, const qfx_contract_desc* desc);

/* The same contraction on the CTA-pair tensor-core kernel with NO device allocation:
 * the kernel's scratch (the smaller operand split into bf16 hi / mid) lives in a
 * static device array of 1 MiB instead of the context's workspace, so the call issues
 * no cudaMalloc / cudaFree (CUDA graph capture, serving loops, harnesses that forbid
 * allocation).  No counterpart in the reference.
 * Operands are arena offsets as for qfx_contract (the arena is the
 * caller's, reserved beforehand).  Limits: the smaller operand holds at most
 * 2^QFX_STATIC_MAX_B_BITS elements with n_bits <= 8 and k_bits <= 8 (after the larger
 * free side is made the rows), and the shape must be one the pair kernel takes
 * (m_bits >= 8 with 2^m_bits / 256 >= number of SM pairs, n_bits >= 4, k_bits >= 4);
 * anything else is QFX_ERR_ARG with the reason in qfx_last_error().  All calls on one
 * GPU share the array: calls that overlap on the device (different contexts or
 * streams) race; calls on one context serialise on its stream and are safe. */
#define QFX_STATIC_MAX_B_BITS 16
int qfx_contract_pair_static(qfx_ctx* ctx, const qfx_contract_desc* desc);

/* n_chains independent chains of small contractions in ONE launch: chain c is the
 * steps descs[chain_first[c] .. chain_first[c+1]) executed in order (a step may
 * read what earlier steps of ITS chain wrote; chains must not touch each
 * other's results).  k_bits <= 6, result <= 2^26 elements per step, fp32.
 * Replaces the per-site world-line products of flatten_grid_of_tensors
 * (src/read_circuit.cpp:548-689: one multiply() per gate, ~30 per qubit). */
int qfx_contract_chains(qfx_ctx* ctx, const qfx_contract_desc* descs,
                        const int32_t* chain_first, int32_t n_chains);

/* A RESIDENT program of chains whose operand bases depend on the cut values of the
 * slice being contracted.  qfx_program_create copies the steps to the device once;
 * qfx_program_run executes all chains in one launch with
 *   off_a(step s) += sum over terms[term_first[2s] .. term_first[2s+1]) of
 *                    ((cut_values[t.cut] >> t.from) & (2^t.width - 1)) << t.shift
 * and likewise off_b with terms[term_first[2s+1] .. term_first[2s+2]).
 * Replaces, for grid tensors next to cut bonds (flattened inside the slice loop by the
 * cut-aware flatten), the per-gate multiply() calls of flatten_grid_of_tensors
 * (src/read_circuit.cpp:560-580) applied to projected gate tensors
 * (Tensor::project, src/tensor.cpp:280-316): ~2000 launches per slice become one. */
typedef struct qfx_var_term {
  int32_t cut, shift, from, width;
} qfx_var_term;
typedef struct qfx_program qfx_program;
int qfx_program_create(qfx_ctx* ctx, const qfx_contract_desc* descs,
                       const int32_t* chain_first, int32_t n_chains,
                       const qfx_var_term* terms, const int32_t* term_first,
                       qfx_program** program_out);
#define QFX_MAX_CUT_VALUES 96
int qfx_program_run(qfx_ctx* ctx, qfx_program* program, const uint32_t* cut_values,
                    int32_t n_values);
int qfx_program_destroy(qfx_ctx* ctx, qfx_program* program);

/* Final scalar contraction (m_bits = n_bits = 0) accumulated in complex128 into
 * amplitude slot `slot`.  Replaces _multiply_vv + `(*amplitudes_)[output_index]
 * += *output->data()` (src/tensor.cpp:1027-1042,
 * src/contraction_utils.cpp:348-352). */
int qfx_dot_accumulate(qfx_ctx* ctx, const qfx_contract_desc* desc,
                       uint64_t slot);
/* amps[slot] += (complex128) arena[off]  (final tensor already a scalar). */
int qfx_scalar_accumulate(qfx_ctx* ctx, uint64_t off, uint64_t slot);
/* Open output indices: the final tensor holds 2^n_bits amplitudes at arena[off ..];
 * element i is added (complex128) to amplitude slot
 * `slot + sum_b ((i >> b) & 1) * slot_stride[b]`.  One call replaces the 2^k
 * projected leaf contractions of the terminal cuts, i.e. the `for i in values`
 * loop of src/contraction_utils.cpp:300-314 with output_index*|values|+i. */
int qfx_tensor_accumulate(qfx_ctx* ctx, uint64_t off, int32_t n_bits,
                          const uint32_t* slot_stride, uint64_t slot);

/* Amplitude table (complex128), replaces std::vector<std::complex<double>>
 * amplitudes (src/evaluate_circuit.cpp:226). */
int qfx_amps_reserve(qfx_ctx* ctx, uint64_t n_amps); /* also zeroes */
int qfx_amps_download(qfx_ctx* ctx, double* host_re_im, uint64_t n_amps);

/* Index permutation = bit permutation of the linear offset:
 * dst[sigma(i)] = src[i] where bit b of i becomes bit dst_pos[b] of sigma(i).
 * One pass (read once, write once).  Replaces Tensor::reorder ->
 * _fast_reorder/_left_reorder/_right_reorder (src/tensor.cpp:494-911).
 * src and dst spans must not overlap. */
int qfx_permute(qfx_ctx* ctx, uint64_t src_off, uint64_t dst_off,
                int32_t n_bits, const int8_t* dst_pos);

/* Index permutation for axes of ARBITRARY dimensions (not powers of two, unit axes):
 * the tensor at src_off is row-major over dims[0..rank) (last axis fastest); new axis j
 * of the result at dst_off is old axis new_order[j].  Replaces Tensor::reorder ->
 * _naive_reorder (src/tensor.cpp:377-492), the branch the reference takes when a
 * dimension is not 2^k.  The circuit pipeline never produces such tensors; provided
 * so that a Tensor-level port has the whole reorder family.  src and dst must not
 * overlap; rank <= QFX_MAX_RANK. */
#define QFX_MAX_RANK 24
int qfx_reorder_general(qfx_ctx* ctx, uint64_t src_off, uint64_t dst_off, int32_t rank,
                        const uint32_t* dims, const int32_t* new_order);

/* Timing helpers (CUDA events on the context stream). */
int qfx_timer_start(qfx_ctx* ctx);
int qfx_timer_stop_ms(qfx_ctx* ctx, float* ms_out);

/* Per-kernel-family accounting since the last reset: launches, device time is
 * NOT measured here (see qfx_timer_*); flops/bytes are the algorithmic ones
 * (8*M*N*K and 8*(MK+KN+MN)). */
typedef struct qfx_stats {
  uint64_t launches_total;
  uint64_t launches_gemm_ffma;
  uint64_t launches_gemm_tc;
  uint64_t launches_reduce;
  uint64_t launches_permute;
  uint64_t launches_copy;
  double flops;
  double bytes;
  uint64_t launches_gemm_tc_fused; /* subset of launches_gemm_tc: in-kernel fp32->bf16 split */
  uint64_t launches_gemm_tc_pair;  /* subset of launches_gemm_tc_fused: CTA-pair kernel (cta_group::2) */
} qfx_stats;
int qfx_stats_get(qfx_ctx* ctx, qfx_stats* out);
int qfx_stats_reset(qfx_ctx* ctx);

/* Per-launch device timing (CUDA events on the context stream around every
 * qfx_contract / qfx_dot_accumulate / qfx_permute), aggregated by shape class.
 * Plays the role of the reference's verbose>1 "[M] Time multiplying A*B" /
 * "[R] Time reordering" stopwatch lines (src/tensor.cpp:1059-1193).  Used by
 * bench.py for the live roofline of the dominant kernel. */
typedef struct qfx_profile_row {
  int32_t kind; /* 0 contract (gemm/dot), 1 permute */
  int32_t m_bits, n_bits, k_bits; /* as given by the caller (before role swap) */
  uint64_t launches;
  double ms;    /* summed device time */
  double flops; /* summed algorithmic flops */
  double bytes; /* summed algorithmic bytes */
} qfx_profile_row;
int qfx_profile_begin(qfx_ctx* ctx);
/* Synchronises, fills up to cap rows (sorted by ms, descending), stops
 * profiling.  *n_out = number of distinct classes seen. */
int qfx_profile_end(qfx_ctx* ctx, qfx_profile_row* rows, int32_t cap,
                    int32_t* n_out);

#ifdef __cplusplus
}
#endif
#endif /* QFLEX_B200_H_ */
