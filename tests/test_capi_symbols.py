"""The C-ABI library builds for sm_100a, loads on a machine without a GPU and exports exactly what
include/signals_b200.h declares.  No compute call is made here."""
import os
import re
import subprocess

from signals_b200 import _lib, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    h = open(os.path.join(ROOT, "include", "signals_b200.h")).read()
    return sorted(set(re.findall(r"SGB_API[^;(]*?\b(sgb_[a-z0-9_]+)\s*\(", h)))


def test_library_builds_and_exports_every_declared_symbol():
    so = build.build()
    assert os.path.exists(so)
    declared = _declared()
    assert len(declared) >= 15
    nm = subprocess.check_output(["nm", "-D", "--defined-only", so], text=True)
    exported = set(re.findall(r"\bT (sgb_[a-z0-9_]+)", nm))
    assert set(declared) <= exported, set(declared) - exported
    # nothing else leaks out of the library under the sgb_ prefix
    assert exported == set(declared), exported - set(declared)
    # the ctypes binding covers the same set
    assert set(_lib.SIGNATURES) == set(declared)


def test_library_loads_and_reports_errors_as_text():
    lib = _lib.load()
    n = lib.sgb_device_count()
    assert n >= 0
    if n == 0:
        assert b"CUDA" in lib.sgb_last_error()      # no silent CPU fallback: the absence of a device is an error text


def test_sass_is_sm100a_only():
    so = build.build()
    out = subprocess.check_output(["cuobjdump", "-lelf", so], text=True)
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_product_never_touches_the_oracle():
    """no file of the product imports, includes, links or dlopens anything under oracle/"""
    pat = re.compile(r"(import\s+oracle|from\s+oracle|liboracle|#include\s*[\"<][^\">]*oracle|oracle/|cpu_oracle)")
    for dp, _, files in os.walk(os.path.join(ROOT, "signals_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".c")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                hits = [m.group(0) for m in pat.finditer(txt)]
                # comments may NAME oracle/oracle.c as the other side of the arithmetic contract
                hits = [h for h in hits if h != "oracle/"]
                assert not hits, (f, hits)
