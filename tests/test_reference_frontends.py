"""The drop-in boundary, tested the way a maintainer of the reference would use it: the
reference's UNMODIFIED front-ends (src/main.cpp, src/pybind_main.{h,cpp}) are compiled
against this repository's host headers and linked with its host library
(__graft_entry__.build_reference_frontends).  CPU tests do the compile here (needs
/root/reference) and run the results on the CPU-emulated device ABI; the -m gpu tests
run the copies linked against the real libraries, which travel to the GPU box prebuilt,
and the repo's own pybind module `qflex` (qflex_b200/pybind)."""
import json
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from tests import support

HAVE_REF = os.path.isdir("/root/reference/src")
REAL_DIR = os.path.join(support.ROOT, "tests", "_build", "refsrc")
OWN_MODULE_DIR = os.path.join(support.ROOT, "qflex_b200", "pybind")


def _cfg(kind, name):
    return os.path.join(support.CONFIG, kind, name)


def _check_cli(binary, golden, name="test"):
    p = subprocess.run([binary, _cfg("circuits", "test.txt"), _cfg("ordering", "test.txt"),
                        _cfg("grid", "test.txt")], stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                       text=True, timeout=300)
    assert p.returncode == 0, p.stderr
    g = golden[name]
    lines = [l for l in p.stdout.splitlines() if "-->" in l]
    assert len(lines) == len(g["states"])
    for line, state, re_, im_ in zip(lines, g["states"], g["re"], g["im"]):
        m = re.match(r"^(\d+) --> (\d+): (\S+) (\S+)$", line)
        assert m and m.group(2) == state, line
        assert abs(complex(float(m.group(3)), float(m.group(4))) - complex(re_, im_)) < 1e-5
    # the reference's own failure path: exit code 2, its wording
    bad = subprocess.run([binary, "/nonexistent", _cfg("ordering", "test.txt"), _cfg("grid", "test.txt")],
                         stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=60)
    assert bad.returncode == 2 and "Failed to call EvaluateCircuit()" in bad.stderr
    usage = subprocess.run([binary, "--help"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    assert "--track-memory" in usage.stdout + usage.stderr


_MODULE_CHECK = r"""
import json, sys
sys.path.insert(0, %(moddir)r)
import qflex
assert qflex.__file__.startswith(%(moddir)r), qflex.__file__
cfg = %(cfg)r
res = qflex.simulate({"circuit_filename": cfg + "/circuits/test.txt",
                      "ordering_filename": cfg + "/ordering/test.txt",
                      "grid_filename": cfg + "/grid/test.txt"})
lines = lambda name: open(cfg + name).read().splitlines()
res2 = qflex.simulate({"circuit": lines("/circuits/rectangular_2x2_1-2-1_0.txt"),
                       "ordering": lines("/ordering/rectangular_2x2.txt"),
                       "grid": lines("/grid/rectangular_2x2.txt"), "final_state": "0110",
                       "memory_limit": "2GB"})
bad = qflex.simulate({"circuit_filename": "/nonexistent", "ordering_filename": cfg + "/ordering/test.txt",
                      "grid_filename": cfg + "/grid/test.txt"})
extra = %(extra)s
print("RESULT " + json.dumps({"test": [[s, a.real, a.imag] for s, a in res],
                               "rect": [[s, a.real, a.imag] for s, a in res2],
                               "bad": len(bad), "extra": extra}))
"""


def _check_module(moddir, golden, extra="None"):
    code = _MODULE_CHECK % dict(moddir=moddir, cfg=support.CONFIG, extra=extra)
    p = subprocess.run([sys.executable, "-c", code], stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                       text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-3000:]
    out = json.loads([l for l in p.stdout.splitlines() if l.startswith("RESULT ")][-1][7:])
    for key, name in (("test", "test"), ("rect", "rect2x2_0110")):
        g = golden[name]
        assert [r[0] for r in out[key]] == g["states"]
        got = np.array([complex(r[1], r[2]) for r in out[key]])
        assert support.rel_err(got, np.array(g["re"]) + 1j * np.array(g["im"])) < 1e-4
    assert out["bad"] == 0  # error printed, empty list returned (src/pybind_main.cpp:127-137)
    return out["extra"]


@pytest.mark.skipif(not HAVE_REF, reason="needs the reference sources (build container only)")
def test_unmodified_reference_frontends_compile_and_run_on_the_emulated_abi(golden):
    import __graft_entry__ as ge
    cli, moddir = ge.build_reference_frontends(emulated=True)
    for f in ("main.cpp", "pybind_main.cpp", "pybind_main.h"):
        # symlinks to the reference's own files: nothing is copied or edited
        assert os.path.realpath(os.path.join(moddir, f)) == os.path.join("/root/reference/src", f)
    _check_cli(cli, golden)
    _check_module(moddir, golden)


@pytest.mark.skipif(not HAVE_REF, reason="needs the reference sources (build container only)")
def test_reference_frontends_link_against_the_product_libraries():
    import __graft_entry__ as ge
    cli, moddir = ge.build_reference_frontends(emulated=False)
    ldd = subprocess.run(["ldd", cli], stdout=subprocess.PIPE, text=True).stdout
    assert "libqflex_host.so" in ldd and "libqflex_b200.so" in ldd and "not found" not in ldd
    assert [f for f in os.listdir(moddir) if f.startswith("qflex.") and f.endswith(".so")]


def test_own_pybind_module_is_built_and_named_qflex():
    mods = [f for f in os.listdir(OWN_MODULE_DIR) if f.startswith("qflex.") and f.endswith(".so")]
    assert mods, "qflex_b200/pybind/qflex*.so missing: run __graft_entry__.build()"
    sym = subprocess.run(["nm", "-D", os.path.join(OWN_MODULE_DIR, mods[0])], stdout=subprocess.PIPE,
                         text=True).stdout
    assert "PyInit_qflex" in sym


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(os.path.join(REAL_DIR, "qflex_refmain.x")),
                    reason="tests/_build/refsrc did not travel (built where /root/reference exists)")
def test_reference_main_and_pybind_sources_on_the_gpu(golden):
    _check_cli(os.path.join(REAL_DIR, "qflex_refmain.x"), golden)
    _check_module(REAL_DIR, golden)


@pytest.mark.gpu
def test_own_qflex_module_on_the_gpu_with_a_batch_of_final_states(golden):
    """`import qflex; qflex.simulate({...})` as qflexcirq does, plus what the reference
    rejects: a list of final states served from one resident circuit."""
    fix = json.load(open(os.path.join(support.ROOT, "tests", "golden", "simulation_fixtures.json")))
    finals = [bin(x)[2:].zfill(16) for x in (0, 5, 1000, 65535, 12345, 4242)]
    extra = ("[[s, a.real, a.imag] for s, a in qflex.simulate({'circuit': %r.splitlines(), "
             "'ordering': %r.splitlines(), 'grid': %r.splitlines(), 'final_states': %r})]"
             % (fix["circuit_test"], fix["ordering_with_cuts_test"], fix["grid_test"], finals))
    out = _check_module(OWN_MODULE_DIR, golden, extra)
    from tests import statevector
    psi = statevector.final_state_vector(fix["circuit_test"])
    assert [r[0] for r in out] == finals
    for (s, re_, im_) in out:
        assert abs(complex(re_, im_) - psi[int(s, 2)]) < 1e-6
