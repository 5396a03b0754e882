"""CPU tests (no GPU): the C-ABI library builds, loads and exports every symbol include/feab.h declares; the
ctypes structs match the header; without a CUDA device the product fails loudly instead of falling back."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from fea_visualizer_b200 import build as fbuild
from fea_visualizer_b200 import feab

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    fbuild.build()
    return feab.load_library()


def header_symbols():
    text = open(os.path.join(ROOT, "include", "feab.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(feab_[a-z_0-9]+)\s*\(", text)))


def test_header_and_binding_agree():
    assert header_symbols() == sorted(feab.ABI_SYMBOLS)


def test_library_exports_every_declared_symbol(lib):
    for name in header_symbols():
        assert hasattr(lib, name), name


def header_struct_fields(name):
    """Field names of `typedef struct { ... } <name>;` in include/feab.h, in declaration order."""
    text = open(os.path.join(ROOT, "include", "feab.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    body = re.search(r"typedef struct\s*\{([^}]*)\}\s*" + name + r"\s*;", text).group(1)
    return [re.sub(r"\[.*?\]", "", d.split()[-1]) for d in body.split(";") if d.strip()]


def test_struct_fields_follow_the_header_order():
    for ct in (feab.FeabParams, feab.FeabSolverOptions, feab.FeabStats):
        assert header_struct_fields(ct.__name__) == [f[0] for f in ct._fields_], ct.__name__


def test_struct_layouts_match_header():
    assert C.sizeof(feab.FeabParams) == 20
    assert C.sizeof(feab.FeabSolverOptions) == 32
    assert C.sizeof(feab.FeabStats) == 2 * 8 + 2 * 8 + 8 + 8 + 8 + 8 + 8 + 2 * 8


def test_no_gpu_means_error_not_fallback(lib):
    h = C.c_void_p()
    rc = lib.feab_create(C.c_int(0), C.c_int(8), C.byref(h))
    if rc == 0:   # a GPU is present: fine, just clean up
        lib.feab_destroy(h)
        return
    assert rc == -3, rc                                  # FEAB_ENOGPU
    assert b"no CPU fallback" in lib.feab_last_error(None)
    with pytest.raises(feab.FeabError):
        feab.Feab(real=64)


def test_bad_real_size_rejected(lib):
    h = C.c_void_p()
    assert lib.feab_create(C.c_int(0), C.c_int(2), C.byref(h)) == -1


def test_null_handle_rejected(lib):
    assert lib.feab_step(None, None, C.c_int(1), C.c_int(0)) == -1
    assert lib.feab_clear(None) == -1


def test_product_never_touches_the_oracle():
    """The product path must not import, link or execute anything under oracle/."""
    pkg = os.path.join(ROOT, "fea_visualizer_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "liboracle" not in text and "from oracle" not in text and "import oracle" not in text, f
    for f in os.listdir(os.path.join(ROOT, "include")):
        pass
    out = os.popen(f"ldd {feab.LIB_PATH}").read()
    assert "oracle" not in out and "fea_ref" not in out
