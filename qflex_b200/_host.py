"""ctypes binding of libqflex_host.so (include/qflex_b200_host.h).

Mirrors the reference's Python entry point ``qflex.simulate(options)``
(reference src/pybind_main.cpp:27-138): same option keys, same return type
(list of (bitstring, complex)), errors printed to stderr and an empty list
returned.  ``HostLibrary.evaluate`` is the raising variant used by tests.
"""
import ctypes
import os
import sys

import numpy as np

N_STATS = 20
STAT_NAMES = (
    "seconds_upload", "seconds_flatten", "seconds_contract", "device_ms_contract",
    "leaves", "launches", "flops", "bytes", "arena_bytes", "flops_per_pass",
    "bytes_per_pass", "launches_gemm_ffma", "launches_gemm_tc", "launches_reduce",
    "launches_permute", "n_amplitudes", "fused_bits", "open_outputs", "lanes", "flops_job",
)
PRECISIONS = {"auto": 0, "ffma": 1, "bf16x3": 2}

_HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_HOST_LIB = os.path.join(_HERE, "libqflex_host.so")


class QflexError(RuntimeError):
    pass


def _text(x):
    """Accepts a filename-less text: str, or an iterable of lines."""
    if x is None:
        return b""
    if isinstance(x, bytes):
        return x
    if isinstance(x, str):
        return x.encode()
    return "\n".join(str(l).rstrip("\n") for l in x).encode()


def _memory_bytes(m):
    if m is None:
        return 1 << 30
    if isinstance(m, int):
        if m <= 0:
            raise QflexError("'memory_limit' must be positive.")
        return m
    s = "".join(str(m).split()).upper()
    num = ""
    while s and s[0].isdigit():
        num, s = num + s[0], s[1:]
    scale = {"GB": 1 << 30, "MB": 1 << 20, "KB": 1 << 10, "B": 1, "": 1}
    if s not in scale:
        raise QflexError("Suffix %s not recognized." % s)
    return int(num) * scale[s]


def restrict_cuts(ordering_text, values_per_cut):
    """Rewrites the value list of the k-th `cut` line (k = 0, 1, ...) of an ordering
    with values_per_cut[k] (None keeps the line): a fixed subset of the slices /
    output bitstrings, selected the same way for this engine and for the reference
    (the grammar's `cut (v1,v2,...) a b`, reference docs/input_formats.md)."""
    import re
    out, k = [], 0
    for line in ordering_text.splitlines():
        if line.strip().startswith("cut"):
            if k < len(values_per_cut) and values_per_cut[k] is not None:
                vals = ",".join(str(v) for v in values_per_cut[k])
                line = re.sub(r"\([0-9,\s]*\)", "(" + vals + ")", line, count=1)
            k += 1
        out.append(line)
    return "\n".join(out) + "\n"


class Prepared:
    """A circuit resident on one GPU: contract() can be called repeatedly."""

    def __init__(self, lib, handle):
        self._lib = lib
        self._h = handle
        self._read_states()
        self.stats = {}

    def _read_states(self):
        n = self._lib._c.qflex_b200_states(self._h, None, 0)
        buf = ctypes.create_string_buffer(n)
        self._lib._c.qflex_b200_states(self._h, buf, n)
        self.states = buf.value.decode().split("\n")

    def contract(self, rank=0, world=1):
        n = len(self.states)
        amps = np.zeros(2 * n, dtype=np.float64)
        stats = np.zeros(N_STATS, dtype=np.float64)
        err = ctypes.create_string_buffer(1 << 16)
        rc = self._lib._c.qflex_b200_contract(
            self._h, rank, world, amps.ctypes.data_as(ctypes.c_void_p), n,
            stats.ctypes.data_as(ctypes.c_void_p), err, len(err))
        if rc < 0:
            raise QflexError(err.value.decode())
        self.stats = dict(zip(STAT_NAMES, stats.tolist()))
        return amps.view(np.complex128)

    def plan_stats(self):
        """Plan figures (arena_bytes, flops_per_pass, flops_job, fused_bits, ...) without
        contracting."""
        stats = np.zeros(N_STATS, dtype=np.float64)
        self._lib._c.qflex_b200_plan_stats(self._h, stats.ctypes.data_as(ctypes.c_void_p))
        return dict(zip(STAT_NAMES, stats.tolist()))

    def set_final_state(self, final_state):
        """Same circuit, another output bitstring: parse/plan/arena are reused."""
        err = ctypes.create_string_buffer(1 << 16)
        if self._lib._c.qflex_b200_set_final_state(self._h, _text(final_state), err, len(err)) < 0:
            raise QflexError(err.value.decode())
        self._read_states()

    def reload(self):
        err = ctypes.create_string_buffer(1 << 16)
        if self._lib._c.qflex_b200_reload(self._h, err, len(err)) < 0:
            raise QflexError(err.value.decode())

    def describe(self):
        n = self._lib._c.qflex_b200_describe(self._h, None, 0)
        buf = ctypes.create_string_buffer(n)
        self._lib._c.qflex_b200_describe(self._h, buf, n)
        rows = []
        for line in buf.value.decode().splitlines():
            phase, level, kind, m, n_, k, label = line.split(" ", 6)
            rows.append(dict(phase=phase, level=int(level), kind=kind, m_bits=int(m),
                             n_bits=int(n_), k_bits=int(k), label=label))
        return rows

    def profile(self, enable=True):
        self._lib._c.qflex_b200_profile(self._h, 1 if enable else 0)

    def profile_rows(self):
        out = np.zeros(8 * 256, dtype=np.float64)
        n = self._lib._c.qflex_b200_profile_rows(self._h, out.ctypes.data_as(ctypes.c_void_p), 256)
        keys = ("kind", "m_bits", "n_bits", "k_bits", "launches", "ms", "flops", "bytes")
        return [dict(zip(keys, out[8 * i: 8 * i + 8].tolist())) for i in range(max(0, min(n, 256)))]

    def close(self):
        if self._h:
            self._lib._c.qflex_b200_release(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class HostLibrary:
    def __init__(self, path=DEFAULT_HOST_LIB):
        if not os.path.exists(path):
            raise QflexError(
                "%s not found: build it with `python -c 'import __graft_entry__ as g; "
                "g.build()'` (there is no Python/CPU fallback)" % path)
        self.path = path
        c = ctypes.CDLL(path)
        c.qflex_b200_evaluate.restype = ctypes.c_long
        c.qflex_b200_evaluate.argtypes = [
            ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p,
            ctypes.c_char_p, ctypes.c_uint64, ctypes.c_int, ctypes.c_int, ctypes.c_int,
            ctypes.c_int, ctypes.c_char_p, ctypes.c_uint64, ctypes.c_void_p,
            ctypes.c_uint64, ctypes.c_void_p, ctypes.c_char_p, ctypes.c_uint64]
        c.qflex_b200_prepare.restype = ctypes.c_void_p
        c.qflex_b200_prepare.argtypes = [
            ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p,
            ctypes.c_char_p, ctypes.c_uint64, ctypes.c_int, ctypes.c_int,
            ctypes.c_char_p, ctypes.c_uint64]
        c.qflex_b200_prepare_sharded.restype = ctypes.c_void_p
        c.qflex_b200_prepare_sharded.argtypes = [
            ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p,
            ctypes.c_char_p, ctypes.c_uint64, ctypes.c_int, ctypes.c_int, ctypes.c_int,
            ctypes.c_char_p, ctypes.c_uint64]
        c.qflex_b200_prepare_batch.restype = ctypes.c_void_p
        c.qflex_b200_prepare_batch.argtypes = c.qflex_b200_prepare_sharded.argtypes
        c.qflex_b200_prepare_on.restype = ctypes.c_void_p
        c.qflex_b200_prepare_on.argtypes = [
            ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p, ctypes.c_char_p,
            ctypes.c_char_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
            ctypes.c_int, ctypes.c_char_p, ctypes.c_uint64]
        c.qflex_b200_contract.restype = ctypes.c_long
        c.qflex_b200_contract.argtypes = [
            ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
            ctypes.c_uint64, ctypes.c_void_p, ctypes.c_char_p, ctypes.c_uint64]
        c.qflex_b200_reload.restype = ctypes.c_long
        c.qflex_b200_reload.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_uint64]
        c.qflex_b200_plan_stats.restype = ctypes.c_long
        c.qflex_b200_plan_stats.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        c.qflex_b200_set_final_state.restype = ctypes.c_long
        c.qflex_b200_set_final_state.argtypes = [ctypes.c_void_p, ctypes.c_char_p,
                                                 ctypes.c_char_p, ctypes.c_uint64]
        c.qflex_b200_states.restype = ctypes.c_long
        c.qflex_b200_states.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_uint64]
        c.qflex_b200_describe.restype = ctypes.c_long
        c.qflex_b200_describe.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_uint64]
        c.qflex_b200_profile.restype = None
        c.qflex_b200_profile.argtypes = [ctypes.c_void_p, ctypes.c_int]
        c.qflex_b200_profile_rows.restype = ctypes.c_long
        c.qflex_b200_profile_rows.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64]
        c.qflex_b200_release.restype = None
        c.qflex_b200_release.argtypes = [ctypes.c_void_p]
        c.qflex_b200_set_verbose.restype = None
        c.qflex_b200_set_verbose.argtypes = [ctypes.c_int]
        self._c = c

    def set_verbose(self, level):
        self._c.qflex_b200_set_verbose(int(level))

    def prepare(self, circuit, ordering, grid, initial_state="", final_state="",
                memory_limit=None, device=0, precision="auto", world=1, devices=None,
                batch=False):
        """`world` = number of ranks the slices will be dealt over by contract(rank,
        world): the planner keeps at least that many slices apart.  `devices` = list of
        GPUs THIS process drives (one host thread each; overrides `device`).  `batch`:
        plan for many final states (set_final_state then costs no upload / flatten)."""
        err = ctypes.create_string_buffer(1 << 16)
        if devices:
            arr = (ctypes.c_int * len(devices))(*[int(d) for d in devices])
            h = self._c.qflex_b200_prepare_on(
                _text(circuit), _text(ordering), _text(grid), _text(initial_state),
                _text(final_state), _memory_bytes(memory_limit), arr, len(devices),
                PRECISIONS[precision], int(world), err, len(err))
            if not h:
                raise QflexError(err.value.decode())
            return Prepared(self, h)
        entry = self._c.qflex_b200_prepare_batch if batch else self._c.qflex_b200_prepare_sharded
        h = entry(
            _text(circuit), _text(ordering), _text(grid), _text(initial_state),
            _text(final_state), _memory_bytes(memory_limit), device,
            PRECISIONS[precision], int(world), err, len(err))
        if not h:
            raise QflexError(err.value.decode())
        return Prepared(self, h)

    def evaluate(self, circuit, ordering, grid, initial_state="", final_state="",
                 memory_limit=None, device=0, rank=0, world=1, precision="auto",
                 max_amplitudes=1 << 20):
        """EvaluateCircuit: returns ([(bitstring, complex)], stats dict)."""
        # output buffers sized for the common case first (a 2^20-entry table costs
        # ~80 MB of zeroed host memory per call); grown once when the engine says so
        for cap in sorted({min(4096, max_amplitudes), max_amplitudes}):
            amps = np.zeros(2 * cap, dtype=np.float64)
            stats = np.zeros(N_STATS, dtype=np.float64)
            states = ctypes.create_string_buffer(cap * 128 if cap <= 4096 else cap * 64)
            err = ctypes.create_string_buffer(1 << 16)
            n = self._c.qflex_b200_evaluate(
                _text(circuit), _text(ordering), _text(grid), _text(initial_state),
                _text(final_state), _memory_bytes(memory_limit), device, rank, world,
                PRECISIONS[precision], states, len(states),
                amps.ctypes.data_as(ctypes.c_void_p), cap,
                stats.ctypes.data_as(ctypes.c_void_p), err, len(err))
            if n >= 0 or not err.value.decode().endswith("_out too small"):
                break
        if n < 0:
            raise QflexError(err.value.decode())
        names = states.value.decode().split("\n")
        vals = amps[: 2 * n].view(np.complex128)
        return list(zip(names, [complex(v) for v in vals])), dict(zip(STAT_NAMES, stats.tolist()))


def _load_option(options, key):
    """options[key] = iterable of lines, or options[key + '_filename'] = path
    (reference LoadData, src/pybind_main.cpp:27-42)."""
    if key in options:
        v = options[key]
        if isinstance(v, (str, bytes)) or not hasattr(v, "__iter__"):
            raise QflexError("'%s' must be a list of strings." % key)
        return _text(v)
    fkey = key + "_filename"
    if fkey in options:
        if not isinstance(options[fkey], str):
            raise QflexError("'%s_filename' must be a string." % key)
        try:
            with open(options[fkey], "rb") as f:
                return f.read()
        except OSError:
            raise QflexError("Cannot open %s file %s." % (key, options[fkey]))
    return b""


def _load_state(options, key):
    if key + "_states" in options:
        v = options[key + "_states"]
        v = [v] if isinstance(v, str) else list(v)
        if not all(isinstance(x, str) for x in v):
            raise QflexError("states must be strings.")
        return v
    if key + "_state" in options:
        if not isinstance(options[key + "_state"], str):
            raise QflexError("states must be strings.")
        return [options[key + "_state"]]
    return [""]


def simulate(options, lib=None):
    """qflex.simulate(options) on the B200 engine.

    Keys (as in the reference): circuit / circuit_filename, ordering /
    ordering_filename, grid / grid_filename, initial_state(s), final_state(s),
    memory_limit.  Additive keys: device, rank, world, precision.
    Errors are printed to stderr and [] is returned (src/pybind_main.cpp:127-137).
    """
    try:
        lib = lib or default_library()
        initial = _load_state(options, "initial")
        final = _load_state(options, "final")
        if len(initial) != 1:
            raise QflexError("Not yet supported")
        # Several final states in one call: the reference answers "Not yet supported"
        # (src/pybind_main.cpp:111-113) and qflexcirq calls simulate once per bitstring.
        circuit = _load_option(options, "circuit")
        ordering = _load_option(options, "ordering")
        grid = _load_option(options, "grid")
        rank, world = options.get("rank", 0), options.get("world", 1)
        if len(final) == 1:
            return lib.evaluate(
                circuit, ordering, grid, initial[0], final[0],
                memory_limit=options.get("memory_limit"), device=options.get("device", 0),
                rank=rank, world=world, precision=options.get("precision", "auto"))[0]
        # batch: the circuit is parsed, planned, uploaded once and stays resident; each
        # further final state rewrites the closing delta tensors and re-runs the flatten
        prepared = lib.prepare(
            circuit, ordering, grid, initial[0], final[0],
            memory_limit=options.get("memory_limit"), device=options.get("device", 0),
            precision=options.get("precision", "auto"), world=world, batch=True)
        try:
            result = []
            for n, state in enumerate(final):
                if n:
                    prepared.set_final_state(state)
                amps = prepared.contract(rank, world)
                result.extend(zip(prepared.states, [complex(a) for a in amps]))
            return result
        finally:
            prepared.close()
    except QflexError as e:
        print(e, file=sys.stderr)
        return []


def trim_device_cache(device=0):
    """Returns the device buffers parked by closed circuits (one arena and one workspace
    per GPU are kept for the next evaluation: repeated EvaluateCircuit calls then skip a
    multi-GB cudaMalloc) to the driver.  Call it when other code needs the memory."""
    from . import _device
    return _device.load().qfx_trim_cache(int(device))


_default = None


def default_library():
    global _default
    if _default is None:
        _default = HostLibrary()
    return _default
