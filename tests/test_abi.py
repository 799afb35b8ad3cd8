"""CPU tests of the drop-in boundary: libipkgpu.so loads without a GPU, exports every symbol include/ipkgpu.h declares,
refuses to create a context without a device (no CPU fallback), and the ctypes mirror of the structs has the C layout."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "ipkgpu.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ipkgpu_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported():
    from ipk_b200 import capi
    lib = capi.load()
    names = _declared()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/ipkgpu.h but not exported by libipkgpu.so"


def test_header_compiles_as_plain_c_and_struct_sizes_match(tmp_path):
    from ipk_b200 import capi
    src = tmp_path / "t.c"
    src.write_text('#include <stdio.h>\n#include "ipkgpu.h"\nint main(void){printf("%zu %zu %zu\\n", sizeof(ipkgpu_input), sizeof(ipkgpu_stats), sizeof(ipkgpu_result));return 0;}\n')
    exe = tmp_path / "t"
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(src)])
    a, b, c = map(int, subprocess.check_output([str(exe)]).split())
    assert (a, b, c) == (C.sizeof(capi.Input), C.sizeof(capi.Stats), C.sizeof(capi.Result))


def test_score_threshold_helper():
    from ipk_b200 import capi
    import numpy as np
    for omega, sigma, k in [(1.5, 4, 12), (2.0, 4, 7), (1.5, 20, 6), (10.0, 20, 4)]:
        t, lt = capi.score_threshold(omega, sigma, k)
        assert abs(t - (omega / sigma) ** k) <= 1e-5 * (omega / sigma) ** k
        assert abs(lt - k * np.log10(omega / sigma)) < 1e-5


def test_no_cpu_fallback():
    import torch
    from ipk_b200 import capi
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(capi.IpkGpuError) as e:
        capi.Context(0)
    assert e.value.code == -2          # IPKGPU_ENODEV


def test_product_does_not_touch_the_oracle():
    """nothing under ipk_b200/ may import, link or execute oracle/ (test infrastructure)"""
    for d, _, files in os.walk(os.path.join(ROOT, "ipk_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".c", ".cpp", ".h")):
                text = open(os.path.join(d, f), errors="ignore").read()
                assert "oraclelib" not in text and "libipk_oracle" not in text and "libipkref" not in text and "oracle/" not in text, os.path.join(d, f)
