"""The `--extract` training state machine, three ways (the product's host/stats_model.hpp and the oracle's
oracle.hpp are the same restatement typed twice, so a test that compares one with the other proves little): both are
compiled into tests/training_trace.cpp and driven, together with the independently written tests/pyref.py:TrainingTrace
(from include/classify_stats.hpp:34-114,341-368,463-578 alone), over sequences of unique proportions that reach every branch —
ties at the maximum, zeros (never negative training data), none / one / several categories above lo_hi_threshold,
data sets filling up in different orders, force_ready with half-trained models."""
import os
import subprocess
import numpy as np
import pytest
from tests import pyref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def harness(tmp_path_factory):
    d = tmp_path_factory.mktemp("trace")
    gxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    out = {}
    for impl in ("PRODUCT", "ORACLE"):
        exe = str(d / f"trace_{impl.lower()}")
        cmd = [gxx, "-O1", "-std=c++14", f"-DIMPL_{impl}", "-ffp-contract=off", os.path.join(ROOT, "tests", "training_trace.cpp"), "-o", exe]
        if impl == "PRODUCT":   # stats_model.hpp also holds the (unused here) model upload through the C-ABI
            cmd += ["-L", os.path.join(ROOT, "charon_b200"), "-lcharon_gpu", "-Wl,-rpath," + os.path.join(ROOT, "charon_b200")]
        else:                   # the built-in KDE data sets live in liboracle.so
            cmd += ["-L", os.path.join(ROOT, "oracle"), "-loracle", "-Wl,-rpath," + os.path.join(ROOT, "oracle")]
        subprocess.check_call(cmd)
        out[impl] = exe
    return out


def _cases():
    rng = np.random.default_rng(17)
    cases = []
    for ncat, cap, thr, n, force in [(2, 5, 0.1, 60, 0), (2, 3, 0.1, 12, 1), (3, 4, 0.1, 120, 0), (3, 6, 0.25, 40, 1), (2, 1, 0.1, 8, 0),
                                     (5, 3, 0.1, 200, 1), (2, 40, 0.1, 30, 1)]:
        rows = []
        for _ in range(n):
            kind = rng.integers(0, 7)
            p = np.zeros(ncat, np.float32)
            if kind == 0:                                   # one clear winner
                p[rng.integers(0, ncat)] = rng.uniform(0.2, 0.9)
                p += rng.uniform(0, 0.05, ncat).astype(np.float32)
            elif kind == 1:                                 # nothing above the threshold: negative data for everybody
                p = rng.uniform(0, thr * 0.9, ncat).astype(np.float32)
            elif kind == 2:                                 # two above the threshold: no training at all
                p = rng.uniform(thr * 1.1, 0.9, ncat).astype(np.float32)
            elif kind == 3:                                 # tie at the maximum
                v = np.float32(rng.uniform(0.2, 0.8)); p[:] = 0.01; p[0] = v; p[-1] = v
            elif kind == 4:                                 # exact zeros (add_neg wants val > 0)
                p[rng.integers(0, ncat)] = rng.uniform(0.3, 0.9)
            elif kind == 5:                                 # all zero: val == max_val (0.0) resets the winner
                pass
            else:                                           # a winner below the threshold
                p[rng.integers(0, ncat)] = np.float32(thr * 0.5)
            rows.append(p)
        cases.append((ncat, cap, thr, force, np.array(rows, np.float32)))
    return cases


def _run(exe, ncat, cap, thr, force, rows):
    text = f"{len(rows)} {ncat} {cap} {thr:.9g} {force}\n" + "\n".join(" ".join(f"{float(v):.9g}" for v in r) for r in rows) + "\n"
    return subprocess.check_output([exe], input=text.encode()).decode().splitlines()


def _python_trace(ncat, cap, thr, force, rows):
    t = pyref.TrainingTrace(ncat, float(np.float32(thr)), cap)
    lines = []
    for r in rows:
        ready = t.add([float(v) for v in r])
        st = t.state()
        lines.append(" ".join([str(int(ready))] + [f"{a} {b} {int(c)}" for a, b, c in st[1]]))
    if force:
        t.force_ready()
    return lines, t


def test_three_implementations_of_the_training_state_machine_agree(harness):
    from charon_b200 import kde_defaults
    dpos, dneg = kde_defaults.default_pos_sorted(), kde_defaults.default_neg_sorted()
    trained_something = 0
    for ncat, cap, thr, force, rows in _cases():
        py_lines, t = _python_trace(ncat, cap, thr, force, rows)
        for impl in ("PRODUCT", "ORACLE"):
            out = _run(harness[impl], ncat, cap, thr, force, rows)
            assert out[:len(rows)] == py_lines, (impl, ncat, cap, next(i for i, (a, b) in enumerate(zip(out, py_lines)) if a != b))
            fits = out[len(rows):]
            for c in range(ncat):
                for j, (key, default) in enumerate((("fit_pos", dpos), ("fit_neg", dneg))):
                    f = fits[2 * c + j].split()
                    got = np.array([float(x) for x in f[4:]], np.float32)
                    want = t.cat[c][key]
                    if want is None:                        # never fitted: the built-in data set, sorted (KDEParams ctor)
                        assert int(f[3]) == len(default) and np.array_equal(got, np.asarray(default, np.float32)), (impl, c, key)
                    else:
                        trained_something += 1
                        assert np.array_equal(got, np.array(want, np.float32)), (impl, c, key)
    assert trained_something >= 10
