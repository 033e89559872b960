"""CPU: the reference-side binding of INTEGRATION.md is real code — integration/shims/b200_shims.cxx defines the guarded
members of Advec_2, Diff_smag2, Pres_2 and Timeloop (TF = double and float) on top of include/mhh_b200.h /
mhh_b200_f32.h and must type-check against the reference's own headers (g++ -fsyntax-only, every shim instantiated)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/include"


@pytest.mark.skipif(not os.path.isdir(REF), reason="needs the reference tree (not present on the GPU box)")
def test_shims_compile_against_the_reference_headers():
    cmd = ["g++", "-std=c++17", "-fsyntax-only", "-w", "-DRESTRICTKEYWORD=__restrict__", "-I" + os.path.join(ROOT, "include"),
           "-I" + os.path.join(ROOT, "integration", "shims"), "-I" + os.path.join(ROOT, "oracle", "shim"),
           "-I/usr/local/cuda/include", "-I" + REF, os.path.join(ROOT, "integration", "shims", "b200_shims.cxx")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]


def test_c_headers_parse_as_c_and_cxx():
    for lang, comp in (("c", "gcc"), ("c++", "g++")):
        for h in ("mhh_b200.h", "mhh_b200_f32.h"):
            r = subprocess.run([comp, "-fsyntax-only", "-x", lang, os.path.join(ROOT, "include", h)], capture_output=True, text=True)
            assert r.returncode == 0, (h, lang, r.stderr[-2000:])
