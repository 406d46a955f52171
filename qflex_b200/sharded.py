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


def allreduce_amplitudes(part, device=None, group=None):
    """Sum of the per-rank partial amplitude tables (complex128), on every rank.
    `device` = the torch device the collective runs on (cuda:<local_rank> for
    NCCL, None/cpu for gloo)."""
    import torch
    dist = _dist()
    flat = np.ascontiguousarray(part, dtype=np.complex128).view(np.float64)
    t = torch.from_numpy(flat.copy())
    if device is not None:
        t = t.to(device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t.cpu().numpy().view(np.complex128)


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
