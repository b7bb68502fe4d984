"""The C ABI of the kernel library: every exported `cxn_*` function is declared in
include/cxxnet_b200.h and vice versa, and the header is plain C."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "cxxnet_b200", "lib", "libcxxnet_b200.so")
HDR = os.path.join(ROOT, "include", "cxxnet_b200.h")


def _declared():
    text = open(HDR).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return set(re.findall(r"\b(cxn_[a-z0-9_]+)\s*\(", text))


@pytest.mark.skipif(not os.path.exists(LIB) or shutil.which("nm") is None, reason="kernel library not built")
def test_exported_symbols_match_the_header():
    out = subprocess.run(["nm", "-D", "--defined-only", LIB], stdout=subprocess.PIPE, text=True, check=True).stdout
    exported = set(l.split()[-1] for l in out.splitlines() if l.split() and l.split()[-2] == "T" and l.split()[-1].startswith("cxn_"))
    declared = _declared()
    assert exported - declared == set(), "exported but not declared: %s" % sorted(exported - declared)
    assert declared - exported == set(), "declared but not exported: %s" % sorted(declared - exported)


@pytest.mark.skipif(shutil.which("gcc") is None or not os.path.isdir("/usr/local/cuda/include"), reason="no C compiler / CUDA headers")
def test_header_compiles_as_c():
    r = subprocess.run(["gcc", "-fsyntax-only", "-x", "c", "-I/usr/local/cuda/include", HDR], stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0, r.stdout
