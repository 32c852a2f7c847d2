"""CPU tests of the boundary: the C-ABI library loads, exports every symbol include/gl_commit.h declares,
and fails loudly (no fallback) when there is no CUDA device."""
import ctypes
import os
import re

import numpy as np
import pytest

import __graft_entry__ as ge
from conftest import ROOT


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "gl_commit.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(gl_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    mod = ge.load()
    if not os.path.exists(mod.LIB_PATH):
        mod.build()
    L = ctypes.CDLL(mod.LIB_PATH)
    syms = header_symbols()
    assert len(syms) >= 30
    for s in syms:
        assert hasattr(L, s), "libglb200.so does not export %s" % s
    # and the Python binding knows a prototype for each of them
    assert sorted(mod.SIGNATURES) == syms


def test_layout_helpers_need_no_gpu():
    mod = ge.load()
    assert mod.tree_num_elements(1) == 4 and mod.tree_num_elements(8) == 60
    assert mod.proof_size(1 << 10, 7) == 7 + 40
    assert mod.proof_size(1, 3) == 3


def test_no_cpu_fallback():
    """Without a device every compute call must raise, never silently compute on the host."""
    mod = ge.load()
    if mod.device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(mod.GLError):
        mod.init(0)
    with pytest.raises(mod.GLError):
        mod.ntt_host(np.ones((4, 1), dtype=np.uint64))
    with pytest.raises(mod.GLError):
        mod.merkletree_host(np.ones((4, 8), dtype=np.uint64))


def test_product_does_not_reference_oracle():
    """The shipped package must not import, link or execute anything under oracle/."""
    pkg = os.path.join(ROOT, "merlin-zkevm-prover_b200")
    for dp, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".hpp", ".cpp", ".h", "Makefile")):
                txt = open(os.path.join(dp, fn), errors="ignore").read()
                assert "gl_oracle" not in txt and "pyoracle" not in txt and "libglref" not in txt, os.path.join(dp, fn)
