"""fit-slam_b200/csrc/atan2_cr.h (the angles of the median sort, chunk_kernel) against this machine's libm, on the host: the
header compiles for both sides.  The contract the search relies on: wherever the `risky` flag is clear, glibc's atan2 returns
the same bits (so the device may decide the sort); where it is set, glibc may return the neighbour, and such a value only
matters next to another angle or a comparator threshold, in which case the chunk's sort goes to the host's libm."""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_atan2_cr_equals_libm_where_not_flagged(tmp_path):
    exe = str(tmp_path / "atan2_check")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-I", os.path.join(ROOT, "fit-slam_b200", "csrc"),
                           os.path.join(ROOT, "tests", "cpp", "atan2_check.cpp"), "-o", exe])
    r = subprocess.run([exe, "1500000"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    m = re.search(r"calls (\d+) disagreements (\d+) unflagged (\d+) risky (\d+)", r.stdout)
    calls, bad, unflagged, risky = map(int, m.groups())
    assert calls > 5_000_000 and unflagged == 0
    assert bad < calls * 2e-3          # glibc is off the correctly rounded value about once in 3 000 calls (0.523 ulp bound)
    assert risky < calls * 0.09        # the flag covers 0.03 ulp on either side of a boundary: ~6 % of all values
