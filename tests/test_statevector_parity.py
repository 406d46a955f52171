"""Engine vs an independent state-vector simulation, on the circuits the reference
checks against cirq (tests/python/simulation_test.py:405-487, 581-610; same 1e-6
bound).  Host logic on the CPU-emulated device here; the same cases run on the GPU
in test_gpu_e2e.py::test_statevector_cases_on_gpu."""
import json
import os

import numpy as np
import pytest

from tests import statevector, support

FIX = json.load(open(os.path.join(support.ROOT, "tests", "golden", "simulation_fixtures.json")))
RNG = np.random.default_rng(20)


def _amp(lib, circuit, ordering, grid, final):
    res, _ = lib.evaluate(circuit, ordering, grid, "", final)
    assert len(res) == 1 and res[0][0] == final
    return res[0][1]


@pytest.fixture(scope="module")
def psi16():
    return statevector.final_state_vector(FIX["circuit_test"])


@pytest.mark.parametrize("x", [int(v) for v in RNG.integers(0, 2 ** 16, 8)])
@pytest.mark.parametrize("ordering", ["ordering_test", "ordering_with_cuts_test"])
def test_16_qubit_circuit_matches_state_vector(emul, psi16, x, ordering):
    final = bin(x)[2:].zfill(16)
    got = _amp(emul, FIX["circuit_test"], FIX[ordering], FIX["grid_test"], final)
    assert abs(got - psi16[x]) < 1e-6


@pytest.mark.parametrize("x", range(16))
def test_fsim_circuit_matches_state_vector(emul, x):
    psi = statevector.final_state_vector(FIX["circuit_test_fsim"])
    final = bin(x)[2:].zfill(4)
    got = _amp(emul, FIX["circuit_test_fsim"], FIX["ordering_2x2_test"], FIX["grid_2x2_test"], final)
    assert abs(got - psi[x]) < 1e-6


def test_terminal_cuts_give_the_whole_distribution_of_the_cut_qubits(emul):
    # open terminal cuts on two qubits of the fSim circuit: 4 amplitudes per call
    ordering = FIX["ordering_2x2_test"].replace("cut () 1 3", "cut () 1 3\ncut () 2\ncut () 3")
    psi = statevector.final_state_vector(FIX["circuit_test_fsim"])
    res, _ = emul.evaluate(FIX["circuit_test_fsim"], ordering, FIX["grid_2x2_test"], "", "01xx")
    assert [s for s, _ in res] == ["0100", "0101", "0110", "0111"]
    for s, a in res:
        assert abs(a - psi[int(s, 2)]) < 1e-6


def test_norm_of_the_simulated_state_is_one(psi16):
    assert abs(np.vdot(psi16, psi16) - 1) < 1e-9


def test_simulate_accepts_a_batch_of_final_states(emul):
    import qflex_b200
    psi = statevector.final_state_vector(FIX["circuit_test_fsim"])
    states = ["0000", "0110", "1011", "1111"]
    out = qflex_b200.simulate({"circuit": FIX["circuit_test_fsim"].split("\n"),
                               "ordering": FIX["ordering_2x2_test"].split("\n"),
                               "grid": FIX["grid_2x2_test"].split("\n"),
                               "final_states": states}, lib=emul)
    assert [s for s, _ in out] == states
    for s, a in out:
        assert abs(a - psi[int(s, 2)]) < 1e-6
