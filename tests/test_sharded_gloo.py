"""The N>1 path on CPU: two processes (gloo), each contracting the slices dealt
to its rank on the CPU-emulated device ABI, then the one all-reduce of
qflex_b200/sharded.py.  Mirrors what bench.py / a torchrun launch does on the
GPUs with NCCL."""
import json
import os
import socket
import subprocess
import sys

import numpy as np

from tests import support

WORKER = r"""
import json, os, sys
sys.path.insert(0, %(root)r)
import numpy as np
import torch.distributed as dist
from tests import support
from qflex_b200 import sharded
dist.init_process_group("gloo")
lib = support.emulated_host_library()
case = support.CASES[%(case)r]
c, o, g = support.case_texts(case)
ini, fin = case.get("initial", ""), case.get("final", "")
p = lib.prepare(c, o, g, ini, fin, memory_limit=8 << 30, world=dist.get_world_size())
full = sharded.contract_sharded(p)
leaves = p.stats["leaves"]
p.close()
res, _ = sharded.evaluate_sharded(lib, c, o, g, ini, fin, memory_limit=8 << 30)
print("RESULT " + json.dumps({"rank": dist.get_rank(), "leaves": leaves,
      "re": full.real.tolist(), "im": full.imag.tolist(),
      "ev_states": [s for s, _ in res],
      "ev_re": [a.real for _, a in res], "ev_im": [a.imag for _, a in res]}))
dist.destroy_process_group()
"""


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run_world(case, world=2):
    support.emulated_host_library()  # build once, not concurrently in the workers
    port = _free_port()
    code = WORKER % dict(root=support.ROOT, case=case)
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE=str(world), LOCAL_RANK=str(r),
                   MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), OMP_NUM_THREADS="2")
        procs.append(subprocess.Popen([sys.executable, "-c", code], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = []
    for p in procs:
        out, err = p.communicate(timeout=300)
        assert p.returncode == 0, err[-3000:]
        line = [l for l in out.splitlines() if l.startswith("RESULT ")][-1]
        outs.append(json.loads(line[7:]))
    return outs


def test_two_ranks_allreduce_to_the_reference_amplitudes(golden):
    name = "bristlecone_48_1-16-1"   # 2 bond-cut slices x 4 terminal leaves
    outs = _run_world(name, 2)
    g = golden[name]
    want = np.array(g["re"]) + 1j * np.array(g["im"])
    assert sum(o["leaves"] for o in outs) == 8          # every leaf exactly once
    assert all(o["leaves"] > 0 for o in outs)           # and both ranks worked
    for o in outs:
        got = np.array(o["re"]) + 1j * np.array(o["im"])
        assert support.rel_err(got, want) < 1e-4
        ev = np.array(o["ev_re"]) + 1j * np.array(o["ev_im"])
        assert o["ev_states"] == g["states"]
        assert support.rel_err(ev, want) < 1e-4
    # the all-reduce leaves the same table on both ranks
    assert outs[0]["re"] == outs[1]["re"] and outs[0]["im"] == outs[1]["im"]


def test_orderings_without_a_dealt_cut_are_not_counted_twice(golden):
    """No cut at all: the whole schedule is one share; rank 0 runs it, rank 1 adds zeros."""
    outs = _run_world("rect2x2_cutfree_0110", 2)
    g = golden["rect2x2_cutfree_0110"]
    want = np.array(g["re"]) + 1j * np.array(g["im"])
    for o in outs:
        assert support.rel_err(np.array(o["re"]) + 1j * np.array(o["im"]), want) < 1e-4
        assert support.rel_err(np.array(o["ev_re"]) + 1j * np.array(o["ev_im"]), want) < 1e-4
    assert sorted(o["leaves"] for o in outs) == [0, len(want)]


def test_single_process_is_the_identity_collective(emul, golden):
    from qflex_b200 import sharded
    name = "test"
    p = emul.prepare(*support.case_texts(support.CASES[name]))
    got = sharded.contract_sharded(p)
    p.close()
    g = golden[name]
    assert support.rel_err(got, np.array(g["re"]) + 1j * np.array(g["im"])) < 1e-4


BATCH_WORKER = r"""
import json, os, sys
sys.path.insert(0, %(root)r)
import numpy as np
import torch.distributed as dist
from tests import support
from qflex_b200 import sharded
dist.init_process_group("gloo")
lib = support.emulated_host_library()
fix = json.load(open(os.path.join(%(root)r, "tests", "golden", "simulation_fixtures.json")))
finals = %(finals)r
p = lib.prepare(fix["circuit_test"], fix["ordering_with_cuts_test"], fix["grid_test"], "", finals[0])
res = sharded.simulate_final_states(p, finals)
p.close()
print("RESULT " + json.dumps({"rank": dist.get_rank(), "states": [s for s, _ in res],
      "re": [a.real for _, a in res], "im": [a.imag for _, a in res]}))
dist.destroy_process_group()
"""


def test_final_states_dealt_over_ranks_and_gathered():
    """A batch of output bitstrings on a resident circuit: bitstring i goes to rank
    i % world, one all-gather returns all amplitudes in request order (5 states on 2
    ranks: the padding of the short rank must not leak into the result)."""
    import json as _json
    from tests import statevector
    fix = _json.load(open(os.path.join(support.ROOT, "tests", "golden", "simulation_fixtures.json")))
    psi = statevector.final_state_vector(fix["circuit_test"])
    finals = [bin(x)[2:].zfill(16) for x in (3, 40000, 65535, 1, 777)]
    support.emulated_host_library()
    port = _free_port()
    code = BATCH_WORKER % dict(root=support.ROOT, finals=finals)
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", LOCAL_RANK=str(r),
                   MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), OMP_NUM_THREADS="2")
        procs.append(subprocess.Popen([sys.executable, "-c", code], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    for p in procs:
        out, err = p.communicate(timeout=300)
        assert p.returncode == 0, err[-3000:]
        got = json.loads([l for l in out.splitlines() if l.startswith("RESULT ")][-1][7:])
        assert got["states"] == finals
        for s, re_, im_ in zip(got["states"], got["re"], got["im"]):
            assert abs(complex(re_, im_) - psi[int(s, 2)]) < 1e-6
