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


IO_LIB = os.path.join(ROOT, "cxxnet_b200", "lib", "libcxxnet_io.so")
IO_HDR = os.path.join(ROOT, "include", "cxxnet_io.h")


@pytest.mark.skipif(not os.path.exists(IO_LIB) or shutil.which("nm") is None or shutil.which("gcc") is None,
                    reason="data pipeline library not built")
def test_io_library_matches_its_header(tmp_path):
    """include/cxxnet_io.h: every exported cxnio_* function is declared and vice versa, the header is plain C, and the
    batch item struct has the size the library was built with"""
    out = subprocess.run(["nm", "-D", "--defined-only", IO_LIB], stdout=subprocess.PIPE, text=True, check=True).stdout
    exported = set(l.split()[-1] for l in out.splitlines() if l.split() and l.split()[-2] == "T" and l.split()[-1].startswith("cxnio_"))
    text = re.sub(r"/\*.*?\*/", "", open(IO_HDR).read(), flags=re.S)
    declared = set(re.findall(r"\b(cxnio_[a-z0-9_]+)\s*\(", text))
    assert exported == declared, (sorted(exported - declared), sorted(declared - exported))
    src = tmp_path / "t.c"
    src.write_text('#include "cxxnet_io.h"\n#include <stdio.h>\nint main(void) { printf("%d %d\\n", (int)sizeof(CxnioAugItem), '
                   'cxnio_aug_item_bytes()); return 0; }\n')
    exe = tmp_path / "t"
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe),
                        IO_LIB, "-Wl,-rpath," + os.path.dirname(IO_LIB)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert r.returncode == 0, r.stdout
    a, b = subprocess.run([str(exe)], stdout=subprocess.PIPE, text=True, check=True).stdout.split()
    assert a == b
