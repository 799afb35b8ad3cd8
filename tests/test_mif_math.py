"""The fp64 exp10 / log2 replacements of the MIF0 kernel (ipk_b200/csrc/mif_math.cuh), compiled for the HOST and
checked against long-double libm over the argument ranges the filter produces (filter.cpp:20-23,55-58)."""
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_mif_math_accuracy(tmp_path):
    exe = str(tmp_path / "mif_math_check")
    subprocess.run(["g++", "-O2", "-I", os.path.join(ROOT, "ipk_b200", "csrc"), "-o", exe,
                    os.path.join(ROOT, "tests", "mif_math_check.cpp")], check=True)
    out = subprocess.run([exe], check=True, stdout=subprocess.PIPE, text=True).stdout
    m = re.search(r"log2: max abs err ([0-9.e+-]+), max rel err \(\|log2\|>1e-3\) ([0-9.e+-]+)", out)
    assert m, out
    assert float(m.group(1)) < 1.5e-14 and float(m.group(2)) < 5e-14          # 1 ulp at |log2| ~ 40; no cancellation near w = 1
    m = re.search(r"exp10: max rel err ([0-9.e+-]+), float mismatches (\d+) of (\d+)", out)
    assert m, out
    assert float(m.group(1)) < 2e-14
    assert int(m.group(2)) <= 2                                              # rounding to float agrees with pow(10, s)
