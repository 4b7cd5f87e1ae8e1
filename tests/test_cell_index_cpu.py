"""The divide-free cell index of the kernels (csrc/bvx_kernels.cuh: cell_of) restated in C and checked on the
CPU against the reference expression floor((v - lo) / size) in IEEE fp32: 35 M axis values over 117 (size,
origin) pairs -- uniform values, values within a few ulps of cell boundaries, random bit patterns (NaN, Inf,
denormals).  The GPU parity tests check the same thing end to end; this one runs without a GPU."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_fast_cell_index_equals_ieee_divide(tmp_path):
    exe = str(tmp_path / "cell_axis_check")
    subprocess.run(["gcc", "-O2", "-ffp-contract=off", "-fno-fast-math", "-o", exe,
                    os.path.join(ROOT, "tests", "csrc", "cell_axis_check.c"), "-lm"], check=True)
    r = subprocess.run([exe, "300000"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert " bad 0 " in r.stdout, r.stdout
