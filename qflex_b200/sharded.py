"""Slice sharding across the GPUs of one box: one process per GPU.

The ordering's bond cuts enumerate independent slices whose partial amplitudes
add up (reference: the `cut` recursion of ContractionData::ContractGrid,
src/contraction_utils.cpp:264-316, runs them one after the other on one host
and `+=`s into the same complex128 table).  Here rank r of `world` contracts
the slices the host layer deals to it (executor.cpp, round-robin at the
shallowest cut levels) and ONE all-reduce(sum) of the amplitude table -- 2*n
doubles -- combines the shares.  There is no other exchange on this path, so
the collective is torch.distributed's (NCCL over NVLink on the GPUs; gloo in
the CPU tests of this host logic).
"""
import numpy as np


def _dist():
    import torch.distributed as dist
    return dist


def rank_and_world(group=None):
    dist = _dist()
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


class AmplitudeReducer:
    """The one collective of the path: all-reduce(sum) of the per-rank partial amplitude
    tables (complex128 = 2 doubles per amplitude), through PERSISTENT buffers -- a
    pinned host staging tensor and a device tensor allocated once -- so that a step
    costs one H2D of 16*n bytes, one ncclAllReduce and one D2H, no allocation.
    `device` = the torch device the collective runs on (cuda:<local_rank> for NCCL,
    None for gloo / single process)."""

    def __init__(self, device=None, group=None):
        self.device, self.group = device, group
        self._host = self._dev = None

    def _buffers(self, n):
        import torch
        if self._host is None or self._host.numel() != n:
            pin = self.device is not None and torch.cuda.is_available()
            self._host = torch.empty(n, dtype=torch.float64, pin_memory=pin)
            self._dev = (torch.empty(n, dtype=torch.float64, device=self.device)
                         if self.device is not None else self._host)
        return self._host, self._dev

    def allreduce(self, part):
        dist = _dist()
        flat = np.ascontiguousarray(part, dtype=np.complex128).view(np.float64)
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(self.group) == 1:
            return flat.copy().view(np.complex128)
        import torch
        host, dev = self._buffers(flat.size)
        host.numpy()[:] = flat
        if dev is not host:
            dev.copy_(host, non_blocking=True)
        dist.all_reduce(dev, op=dist.ReduceOp.SUM, group=self.group)
        if dev is not host:
            host.copy_(dev, non_blocking=True)
            torch.cuda.current_stream(self.device).synchronize()
        return host.numpy().copy().view(np.complex128)


_reducers = {}


def allreduce_amplitudes(part, device=None, group=None):
    """Sum of the per-rank partial amplitude tables (complex128), on every rank."""
    key = (str(device), id(group))
    if key not in _reducers:
        _reducers[key] = AmplitudeReducer(device, group)
    return _reducers[key].allreduce(part)


def allgather_amplitudes(part, device=None, group=None):
    """Independent output bitstrings (separate final states) sharded over ranks: every
    rank contributes the amplitudes it computed, all ranks receive all of them in rank
    order ('for independent bitstrings one ncclAllGather').  `part` must
    have the same length on every rank."""
    import torch
    dist = _dist()
    flat = np.ascontiguousarray(part, dtype=np.complex128).view(np.float64)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return flat.copy().view(np.complex128)
    t = torch.from_numpy(flat.copy())
    if device is not None:
        t = t.to(device)
    out = torch.empty(dist.get_world_size(group) * t.numel(), dtype=t.dtype, device=t.device)
    dist.all_gather_into_tensor(out, t, group=group)
    return out.cpu().numpy().view(np.complex128)


def simulate_final_states(prepared, final_states, device=None, group=None):
    """A batch of output bitstrings on a resident circuit, the bitstrings dealt over the
    ranks (state i -> rank i % world), each rank contracting ALL slices of its states;
    one all-gather at the end.  Returns [(state, amplitude)] in the order of
    `final_states` x the circuit's open outputs, on every rank."""
    rank, world = rank_and_world(group)
    per_rank = -(-len(final_states) // world)
    padded = list(final_states) + [final_states[-1]] * (per_rank * world - len(final_states))
    mine, names = [], []
    for i in range(per_rank):
        prepared.set_final_state(padded[i * world + rank])
        mine.append(prepared.contract(0, 1).copy())
    n_out = len(prepared.states)
    allv = allgather_amplitudes(np.concatenate(mine), device, group).reshape(world, per_rank, n_out)
    out = []
    for i, state in enumerate(final_states):
        r, j = i % world, i // world
        prepared_states = _states_for(prepared, state)
        out.extend(zip(prepared_states, [complex(a) for a in allv[r, j]]))
    return out


def _states_for(prepared, final_state):
    """Output strings of `prepared` for another final state: the terminal-cut positions
    (where the current strings differ from each other or were open) keep their values."""
    cur = prepared.states
    if len(cur) == 1:
        return [final_state]
    varying = [p for p in range(len(cur[0])) if len({s[p] for s in cur}) > 1]
    res = []
    for s in cur:
        t = list(final_state)
        for p in varying:
            t[p] = s[p]
        res.append("".join(t))
    return res


def contract_sharded(prepared, device=None, group=None):
    """This rank's share of the slices of a resident circuit, then the all-reduce.
    Returns the full amplitude vector (same on every rank)."""
    rank, world = rank_and_world(group)
    return allreduce_amplitudes(prepared.contract(rank, world), device, group)


def evaluate_sharded(lib, circuit, ordering, grid, initial_state="", final_state="",
                     device_index=None, device=None, group=None, **kw):
    """EvaluateCircuit with the slices dealt over the ranks of `group`.
    Returns ([(state, amplitude)], stats of this rank)."""
    rank, world = rank_and_world(group)
    if device_index is not None:
        kw["device"] = device_index
    res, stats = lib.evaluate(circuit, ordering, grid, initial_state, final_state,
                              rank=rank, world=world, **kw)
    full = allreduce_amplitudes(np.array([a for _, a in res], dtype=np.complex128), device, group)
    return [(s, complex(a)) for (s, _), a in zip(res, full)], stats
