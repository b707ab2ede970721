"""The KDE term the CUDA kernel evaluates (charon_b200/csrc/kde_fast_exp.inc) against the reference expression
float(exp(-z*z/2) / 2.5066282746310002) (include/classify_stats.hpp:242-252) on EVERY float z of [0, 15] and its negative
— the whole domain on which the term is not exactly zero.  Pure host C, a few seconds on all cores."""
import os
import shutil
import subprocess
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(shutil.which("gcc") is None, reason="needs gcc")
def test_kde_term_equals_reference_expression_for_every_float(tmp_path):
    exe = str(tmp_path / "kde_exp_exhaustive")
    flags = ["-O2", "-fopenmp"]
    try:
        if " fma " in open("/proc/cpuinfo").read():
            flags.append("-mfma")          # hardware fma; without it libm's exact software fma is used (same results, slower)
    except OSError:
        pass
    subprocess.check_call(["gcc", *flags, "-o", exe, os.path.join(ROOT, "tests", "kde_exp_exhaustive.c"), "-lm"])
    out = subprocess.run([exe], capture_output=True, text=True, timeout=1500)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "tested 1097859073 mismatches 0" in out.stdout, out.stdout


def test_table_is_reproducible():
    """kde_exp_table.inc is what tools/gen_kde_exp_table.py writes."""
    path = os.path.join(ROOT, "charon_b200", "csrc", "kde_exp_table.inc")
    before = open(path).read()
    subprocess.check_call(["python", os.path.join(ROOT, "tools", "gen_kde_exp_table.py")], stdout=subprocess.DEVNULL)
    assert open(path).read() == before
