"""A plain numpy state-vector simulator for the qFlex circuit file format: the
independent check the reference gets from cirq in tests/python/simulation_test.py
(cirq is not in this image).  Gates are written from their textbook definitions,
NOT from the engine's gate tables: H, T, X^1/2, Y^1/2, W^1/2 (hz_1_2), rz(theta*pi),
CZ, CNOT, fSim(theta*pi, phi*pi).  Qubit q of the file is bit q of the output string
(character 0 first), for grids without disabled qubits."""
import re

import numpy as np

_S = 1 / np.sqrt(2)
ONE_QUBIT = {
    "h": np.array([[_S, _S], [_S, -_S]], dtype=complex),
    "t": np.diag([1, np.exp(1j * np.pi / 4)]),
    "x_1_2": 0.5 * np.array([[1 + 1j, 1 - 1j], [1 - 1j, 1 + 1j]]),
    "y_1_2": 0.5 * np.array([[1 + 1j, -1 - 1j], [1 + 1j, 1 + 1j]]),
    # W = (X + Y)/sqrt(2); W^1/2 = Z^1/4 X^1/2 Z^-1/4
    "hz_1_2": np.diag([1, np.exp(1j * np.pi / 4)]) @ (0.5 * np.array([[1 + 1j, 1 - 1j], [1 - 1j, 1 + 1j]]))
    @ np.diag([1, np.exp(-1j * np.pi / 4)]),
}


def rz(theta):
    a = np.pi * theta
    return np.diag([np.exp(-0.5j * a), np.exp(0.5j * a)])


def fsim(theta, phi):
    t, p = np.pi * theta, np.pi * phi
    return np.array([[1, 0, 0, 0], [0, np.cos(t), -1j * np.sin(t), 0],
                     [0, -1j * np.sin(t), np.cos(t), 0], [0, 0, 0, np.exp(-1j * p)]])


TWO_QUBIT = {
    "cz": np.diag([1, 1, 1, -1]).astype(complex),
    "cx": np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 0, 1], [0, 0, 1, 0]], dtype=complex),
}


def parse(circuit_text):
    lines = [l.strip() for l in circuit_text.strip().splitlines() if l.strip()]
    n = int(lines[0])
    gates = []
    for l in lines[1:]:
        m = re.match(r"(\d+)\s+([a-z_0-9]+)(\(([^)]*)\))?\s+(.*)", l)
        cycle, name, params, qs = int(m.group(1)), m.group(2), m.group(4), m.group(5)
        params = [float(x) for x in params.split(",")] if params else []
        gates.append((cycle, name, params, [int(q) for q in qs.split()]))
    return n, gates


def final_state_vector(circuit_text, initial=None):
    """All 2^n amplitudes <s|U|initial>, index = int(s, 2) with qubit 0 first."""
    n, gates = parse(circuit_text)
    psi = np.zeros([2] * n, dtype=complex)
    psi[tuple(int(c) for c in (initial or "0" * n))] = 1
    for _, name, params, qs in gates:  # file order = cycle order
        if len(qs) == 1:
            u = rz(*params) if name == "rz" else ONE_QUBIT[name]
            psi = np.moveaxis(np.tensordot(u, psi, axes=([1], [qs[0]])), 0, qs[0])
        else:
            u = (fsim(*params) if name == "fsim" else TWO_QUBIT[name]).reshape(2, 2, 2, 2)
            psi = np.tensordot(u, psi, axes=([2, 3], qs))
            psi = np.moveaxis(psi, [0, 1], qs)
    return psi.reshape(-1)
