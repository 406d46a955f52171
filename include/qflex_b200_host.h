/*
 * qflex_b200_host.h -- C entry points of libqflex_host.so (the C++ host layer:
 * parsers, network builder, planner, executor).  These are what a language
 * binding (ctypes, pybind's qflex.simulate, cgo ...) calls; they wrap
 * qflex::EvaluateCircuit (reference src/evaluate_circuit.h:81-82) and its
 * resident variant.  Inputs are the TEXTS of the three qFlex files, exactly what
 * QflexCircuit/QflexOrdering/QflexGrid::load(std::istream&) consume
 * (reference src/circuit.h:26-35, src/ordering.h:13-19, src/grid.h:13-20).
 *
 * Errors: functions return -1 (or NULL) and copy the reference-style error
 * string ("ERROR (file: function:line) --> ...") into err.
 */
#ifndef QFLEX_B200_HOST_H_
#define QFLEX_B200_HOST_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define QFLEX_B200_N_STATS 20
/* stats_out layout (doubles):
 *  [0] seconds_upload   [1] seconds_flatten   [2] seconds_contract (host wall)
 *  [3] device_ms_contract (CUDA events)       [4] leaves executed on this rank
 *  [5] kernel launches in the last Contract() [6] algorithmic flops executed
 *  [7] algorithmic bytes executed             [8] arena bytes
 *  [9] flops of one pass over the schedule    [10] bytes of one pass
 *  [11] gemm (FFMA) launches [12] gemm (tensor core) launches
 *  [13] reduce launches      [14] permute launches   [15] number of amplitudes
 *  [16] fused cut bits (one pass covers 2^f slices of the ordering as written)
 *  [17] terminal cuts kept open          [18] lanes (contexts on this GPU)
 *  [19] algorithmic flops of the whole job executed once (all slices; what one rank
 *       alone would execute -- the useful work of a sharded run)
 */

/* One-shot EvaluateCircuit.  states_out receives the output bitstrings joined
 * by '\n'; amps_out receives 2*n doubles (re, im).  Returns n >= 0 or -1. */
long qflex_b200_evaluate(const char* circuit_text, const char* ordering_text,
                         const char* grid_text, const char* initial_state,
                         const char* final_state, uint64_t memory_limit_bytes,
                         int device, int rank, int world, int precision,
                         char* states_out, uint64_t states_cap, double* amps_out,
                         uint64_t amps_cap_complex, double* stats_out, char* err,
                         uint64_t err_cap);

/* Resident variant: parse + plan + upload + flatten once ... */
void* qflex_b200_prepare(const char* circuit_text, const char* ordering_text,
                         const char* grid_text, const char* initial_state,
                         const char* final_state, uint64_t memory_limit_bytes,
                         int device, int precision, char* err, uint64_t err_cap);
/* Same, planned for a run on `world` ranks: the planner undoes bond-cut slicing
 * (larger kernels, fewer passes) only as far as every rank keeps a share of the
 * remaining slices.  qflex_b200_prepare plans for world = 1. */
void* qflex_b200_prepare_sharded(const char* circuit_text, const char* ordering_text,
                                 const char* grid_text, const char* initial_state,
                                 const char* final_state, uint64_t memory_limit_bytes,
                                 int device, int precision, int world, char* err,
                                 uint64_t err_cap);
/* A resident circuit planned for a BATCH of final states (qflex_b200_set_final_state):
 * every qubit keeps its output leg on its grid tensor and the final state is a set of
 * projection offsets, so another bitstring costs the contraction alone -- no upload,
 * no flatten.  Falls back to the plain plan when the doubled grid tensors do not fit. */
void* qflex_b200_prepare_batch(const char* circuit_text, const char* ordering_text,
                               const char* grid_text, const char* initial_state,
                               const char* final_state, uint64_t memory_limit_bytes,
                               int device, int precision, int world, char* err,
                               uint64_t err_cap);
/* Same on SEVERAL GPUs driven by this one process: every listed device holds the
 * network, the slices of `rank` are dealt over them (one host thread per GPU), the
 * partial amplitudes are summed on the host.  The single-process form of multi-GPU
 * (the reference's EvaluateCircuit / CLI are single-process, src/main.cpp:129); the
 * one-process-per-GPU form is rank/world + one all-reduce (qflex_b200/sharded.py). */
void* qflex_b200_prepare_on(const char* circuit_text, const char* ordering_text,
                            const char* grid_text, const char* initial_state,
                            const char* final_state, uint64_t memory_limit_bytes,
                            const int* devices, int n_devices, int precision, int world,
                            char* err, uint64_t err_cap);
/* ... then run the contraction schedule any number of times.  Only the slices
 * dealt to `rank` of `world` are executed; sum the partial amplitudes across
 * ranks (NCCL all-reduce).  Returns the number of amplitudes or -1. */
long qflex_b200_contract(void* prepared, int rank, int world, double* amps_out,
                         uint64_t amps_cap_complex, double* stats_out, char* err,
                         uint64_t err_cap);
/* Re-run the host->device upload of the gate tensors and the device-side
 * flatten (used to time the full host-buffer path). */
long qflex_b200_reload(void* prepared, char* err, uint64_t err_cap);
/* The stats_out layout above for a resident circuit without running it (plan
 * figures: arena bytes, flops per pass, flops of the whole job, fused bits ...; the
 * execution figures are those of the last contract()). */
long qflex_b200_plan_stats(void* prepared, double* stats_out);
/* Another final state on the same resident circuit: only the delta tensors closing
 * the world lines are rewritten and the (ms-scale) flatten re-run; parse, plan, arena
 * and schedule are reused.  Positions of terminal-cut qubits are ignored.  This is the
 * batched form of the reference's `final_states` list, which src/pybind_main.cpp:49-73
 * parses and :111-113 rejects ("Not yet supported"), so that qflexcirq calls
 * simulate() once per bitstring (qflexcirq/interface/qflex_simulator.py:44-66). */
long qflex_b200_set_final_state(void* prepared, const char* final_state, char* err,
                                uint64_t err_cap);
long qflex_b200_states(void* prepared, char* states_out, uint64_t states_cap);
/* Text dump of the plan, one step per line:
 * "<phase> <level> <kind> <m_bits> <n_bits> <k_bits> <label>". Returns bytes
 * needed (including NUL). */
long qflex_b200_describe(void* prepared, char* out, uint64_t cap);
/* Per-launch device timing by shape class for the following contract() calls
 * (qfx_profile_begin/end).  Rows are 8 doubles: kind, m_bits, n_bits, k_bits,
 * launches, summed ms, summed algorithmic flops, summed algorithmic bytes;
 * sorted by ms descending.  Returns the number of rows available. */
void qflex_b200_profile(void* prepared, int enable);
long qflex_b200_profile_rows(void* prepared, double* out, uint64_t cap_rows);
void qflex_b200_release(void* prepared);

/* verbosity of the host layer (reference global::verbose, src/global.h:9). */
void qflex_b200_set_verbose(int level);

#ifdef __cplusplus
}
#endif
#endif /* QFLEX_B200_HOST_H_ */
