"""Branch-free fp64 reciprocal / division / sqrt / atan2 of the innovation (csrc/ekf_fastmath.cuh): accuracy against
long double libm on the host. The header is the device code; only the ~2^-20 MUFU seeds are emulated."""
import math
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "slam-duckietown_b200", "csrc")


@pytest.fixture(scope="module")
def report(tmp_path_factory):
    if shutil.which("g++") is None:
        pytest.skip("g++ not available")
    exe = str(tmp_path_factory.mktemp("fm") / "fastmath_check")
    subprocess.run(["g++", "-O2", "-mfma", "-ffp-contract=off", "-I", CSRC,
                    os.path.join(ROOT, "tests", "fastmath_check.cpp"), "-o", exe], check=True)
    return subprocess.run([exe], check=True, capture_output=True, text=True).stdout


def test_fastmath_ulp_bounds(report):
    m = re.search(r"atan2 ([\d.]+) .* div ([\d.]+) rcp ([\d.]+) sqrt ([\d.]+) rsqrt ([\d.]+) ulp", report)
    assert m, report
    atan2_ulp, div_ulp, rcp_ulp, sqrt_ulp, rsqrt_ulp = map(float, m.groups())
    assert atan2_ulp <= 2.5
    assert div_ulp <= 1.0 and rcp_ulp <= 1.0 and sqrt_ulp <= 1.0
    assert rsqrt_ulp <= 2.0


def test_fastmath_atan2_special_values(report):
    rows = re.findall(r"atan2\((\S+),(\S+)\) = (\S+) libm (\S+)", report)
    assert len(rows) >= 10
    for y, x, got, want in rows:
        got, want = float(got), float(want)
        assert got == want or abs(got - want) <= 2 * math.ulp(want), (y, x, got, want)
        assert math.copysign(1.0, got) == math.copysign(1.0, want), (y, x, got, want)
