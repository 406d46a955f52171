"""qflex_b200.x: the reference binary's usage, output format and exit codes
(src/main.cpp:14-38, 140-158: `<initial> --> <state>: <re> <im>` per amplitude on
stdout; errors on stderr, exit code 2 for engine errors)."""
import os
import re
import subprocess

import numpy as np
import pytest

from tests import support

BIN = os.path.join(support.ROOT, "qflex_b200", "qflex_b200.x")


def _run(*args):
    return subprocess.run([BIN] + list(args), stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                          text=True, timeout=300)


def _cfg(kind, name):
    return os.path.join(support.CONFIG, kind, name)


def test_help_lists_the_reference_options():
    p = _run("--help")
    assert p.returncode == 0
    for needle in ("<circuit_filename> <ordering_filename> <grid_filename>", "--initial-conf",
                   "--final-conf", "--verbosity", "--memory", "--track-memory"):
        assert needle in p.stdout + p.stderr


def test_wrong_arguments_print_usage():
    p = _run("only_one_argument")
    assert "Usage:" in p.stdout + p.stderr


def test_engine_errors_exit_with_code_2_and_the_reference_wording():
    p = _run("-c", "/nonexistent/circuit.txt", "-o", _cfg("ordering", "test.txt"),
             "-g", _cfg("grid", "test.txt"))
    assert p.returncode == 2
    assert "Failed to call EvaluateCircuit()" in p.stderr


@pytest.mark.gpu
def test_output_lines_match_the_reference_format_and_golden(golden):
    p = _run(_cfg("circuits", "test.txt"), _cfg("ordering", "test.txt"), _cfg("grid", "test.txt"))
    assert p.returncode == 0, p.stderr
    lines = [l for l in p.stdout.splitlines() if "-->" in l]
    g = golden["test"]
    assert len(lines) == len(g["states"])
    for line, state, re_, im_ in zip(lines, g["states"], g["re"], g["im"]):
        m = re.match(r"^(\d+) --> (\d+): (\S+) (\S+)$", line)
        assert m, line
        assert m.group(1) == "00000" and m.group(2) == state
        assert abs(complex(float(m.group(3)), float(m.group(4))) - complex(re_, im_)) < 1e-5


@pytest.mark.gpu
def test_positional_and_flag_forms_agree():
    a = _run(_cfg("circuits", "test.txt"), _cfg("ordering", "test.txt"), _cfg("grid", "test.txt"))
    b = _run("-c", _cfg("circuits", "test.txt"), "-o", _cfg("ordering", "test.txt"),
             "-g", _cfg("grid", "test.txt"), "--memory=2GB")
    assert a.returncode == 0 and b.returncode == 0
    assert a.stdout == b.stdout


@pytest.mark.gpu
def test_devices_flag_deals_the_slices_over_contexts(golden):
    a = _run(_cfg("circuits", "test.txt"), _cfg("ordering", "test.txt"), _cfg("grid", "test.txt"))
    b = _run("-c", _cfg("circuits", "test.txt"), "-o", _cfg("ordering", "test.txt"),
             "-g", _cfg("grid", "test.txt"), "--devices=0,0")
    assert a.returncode == 0 and b.returncode == 0, b.stderr
    va = [complex(*map(float, l.split(": ")[1].split())) for l in a.stdout.splitlines() if "-->" in l]
    vb = [complex(*map(float, l.split(": ")[1].split())) for l in b.stdout.splitlines() if "-->" in l]
    assert len(va) == len(vb) == len(golden["test"]["states"])
    assert np.allclose(va, vb, rtol=0, atol=1e-6)
    bad = _run("-c", _cfg("circuits", "test.txt"), "-o", _cfg("ordering", "test.txt"),
               "-g", _cfg("grid", "test.txt"), "--devices=a,b")
    assert bad.returncode == 2 and "Device list" in bad.stderr
