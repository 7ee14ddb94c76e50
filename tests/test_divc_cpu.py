"""x / c by multiplication + two FMA residual corrections (csrc/irs_common.cuh div_const / DIVC / div_small) is the
hardware division bit for bit, for every divisor the kernels use it with.  tests/divc_check.c feeds adversarial
numerators (quotients next to representable values and next to rounding midpoints, random patterns, counts)."""
import os
import re
import shutil
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "..", "intent_radio_sched_multi_slice_b200", "csrc")
DIVISORS = ["1", "2", "3", "4", "5", "6", "7", "8", "9", "10", "40", "100", "120", "135", "1e6"]


@pytest.fixture(scope="module")
def checker(tmp_path_factory):
    if shutil.which("gcc") is None:
        pytest.skip("gcc not available")
    exe = str(tmp_path_factory.mktemp("divc") / "divc_check")
    subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-o", exe, os.path.join(HERE, "divc_check.c"), "-lm"])
    return exe


def test_sequence_matches_division(checker):
    out = subprocess.run([checker, "400000"] + DIVISORS, capture_output=True, text=True)
    assert out.returncode == 0, out.stdout
    assert out.stdout.count(" 0 mismatches") == len(DIVISORS)


def test_sequence_matches_division_for_runtime_divisors(checker):
    """Recip / div_by and the per-slice drift reciprocals: one IEEE division 1 / c, then the sequence per numerator"""
    out = subprocess.run([checker, "40", "rand:60000"], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout
    assert " 0 mismatches" in out.stdout


def test_small_integer_division_by_multiplication():
    """idiv_small (irs_common.cuh): (x * ceil(2^16 / m)) >> 16 == x // m for 0 <= x < 13,000 and m = 1..7"""
    import numpy as np
    c = np.array([0, 65536, 32768, 21846, 16384, 13108, 10923, 9363], np.uint64)
    x = np.arange(13000, dtype=np.uint64)
    for m in range(1, 8):
        assert c[m] == -(-65536 // m)
        assert np.array_equal((x * c[m]) >> np.uint64(16), x // np.uint64(m)), m


def test_numerator_set_is_adversarial(checker):
    """the bare product x * RN(1/c) must fail on the same numerators (otherwise the set proves nothing)"""
    out = subprocess.run([checker, "50000", "3", "5", "135", "1e6"], capture_output=True, text=True,
                         env=dict(os.environ, DIVC_WEAK="1"))
    assert out.returncode == 1
    for line in out.stdout.splitlines():
        assert int(line.split(":")[1].split()[0]) > 1000, line


def test_every_kernel_divisor_is_covered():
    """every constant the kernels pass to DIVC (and div_small's 1..10) is in DIVISORS"""
    consts = set()
    for f in os.listdir(CSRC):
        if f.endswith(".cuh"):
            for m in re.finditer(r"DIVC\((?:[^()]|\([^()]*(?:\([^()]*\))*[^()]*\))*,\s*([0-9.e]+|R|c)\)", open(os.path.join(CSRC, f)).read()):
                consts.add(m.group(1))
    consts.discard("c")  # the macro's own parameter
    vals = {135.0 if c == "R" else float(c) for c in consts}
    assert vals and vals <= {float(d) for d in DIVISORS}, vals
