"""CPU suite: the C-ABI library builds, loads, exports exactly what include/twtml.h declares, and
refuses to work without a GPU (no CPU fallback).  No compute calls here."""
import ctypes as C
import os
import re
import subprocess

import pytest

import twtml_b200
from twtml_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    src = open(os.path.join(ROOT, "include", "twtml.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\bint32_t\s+(twtml_\w+)\s*\(", src)))


def test_header_and_binding_agree():
    names = declared_functions()
    assert len(names) >= 30
    assert sorted(_lib.SIGNATURES) == names          # the ctypes table mirrors the header one to one


def test_library_exports_every_declared_symbol():
    so = _lib.build()
    out = subprocess.check_output(["nm", "-D", "--defined-only", so], text=True)
    exported = set(re.findall(r" T (twtml_\w+)", out))
    assert exported == set(declared_functions())     # nothing missing, nothing extra leaks out
    L = _lib.lib()
    assert L.twtml_abi_version() == 1


def test_header_compiles_as_plain_c(tmp_path):
    c = tmp_path / "t.c"
    c.write_text('#include "twtml.h"\nint main(void){ twtml_stats s; (void)s; return TWTML_ABI_VERSION - 1; }\n')
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-c", str(c),
                           "-o", str(tmp_path / "t.o")])


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    rc = _lib.lib().twtml_create(0, C.byref(h))
    assert rc == _lib.E_CUDA and not h.value
    assert "no CPU fallback" in _lib.last_error()
    with pytest.raises(twtml_b200.TwtmlError):
        twtml_b200.Context(0)


def test_product_never_touches_the_oracle():
    """No import, link, dlopen or call of anything under oracle/ from the product tree."""
    pkg = os.path.join(ROOT, "twitter-stream-ml_b200")
    bad = re.compile(r"import\s+oracle|from\s+oracle|oracle_c\b|oracle_np\b|libtwtml_oracle|twtml_oracle_\w+\s*\(")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                assert not bad.search(open(os.path.join(dp, f)).read()), f


def test_stats_wire_format_matches_the_web_client():
    """web/src/test/.../WebTestSuite.scala:10-11 uses Stats(1000, 10, 2000, 15, 25); web/src/main/assets/js/api.js:55-63
    reads count / batch / mse / realStddev / predStddev from the same JSON."""
    import json
    body = twtml_b200.stats_payload(1000, dict(batch=10, mse=2000.0, real_stdev=15.0, pred_stdev=25.0))
    assert json.loads(body) == {"jsonClass": "Stats", "count": 1000, "batch": 10, "mse": 2000, "realStddev": 15, "predStddev": 25}
