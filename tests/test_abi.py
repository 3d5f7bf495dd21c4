"""CPU tests: the C-ABI library builds for sm_100a, loads, exports every symbol include/gasma_b200.h
declares, and refuses to work without a GPU (no CPU fallback).  No compute calls here."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built(pkg):
    import importlib.util
    spec = importlib.util.spec_from_file_location("gasma_build", os.path.join(ROOT, "approximate-string-matching_b200", "build.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m.build_library()


def test_header_symbols_are_exported(pkg, built):
    hdr = open(os.path.join(ROOT, "include", "gasma_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(gasma_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(pkg.EXPORTS)
    L = C.CDLL(built)
    for name in declared:
        assert hasattr(L, name), name


def test_library_is_sm100a_only(built):
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", built], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_defaults_match_reference_constructor(pkg, built):
    L = pkg.lib()
    p = pkg.Params()
    L.gasma_default_params(C.byref(p))
    # hurdle_matrix.h:552-562 and benchmark.cpp:22
    assert (p.alignment_type, p.x, p.o, p.e, p.k, p.max_length) == (0, 1, 1, 1, 3, 128)
    assert (p.match_prob, p.mismatch_prob, p.indel_prob) == (0.80, 0.20 / 3, 0.40 / 3)


def test_no_cpu_fallback(pkg, built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pkg.GasmaError) as ei:
        pkg.Aligner()
    assert ei.value.code == -5  # GASMA_E_CUDA


def test_bad_params_rejected_before_touching_cuda(pkg, built):
    for kw, code in [(dict(max_length=100), -6), (dict(k=51), -3), (dict(alignment_type=2), -6), (dict(match_prob=0.0), -1)]:
        with pytest.raises(pkg.GasmaError) as ei:
            pkg.Aligner(**kw)
        assert ei.value.code == code, kw


def test_render_cigar(pkg, built):
    import numpy as np
    ops = np.array([(22 << 3) | 0, (1 << 3) | 2, (50 << 3) | 0, (1 << 3) | 1, (3 << 3) | 3, (2 << 3) | 4], np.uint16)
    buf = C.create_string_buffer(64)
    n = pkg.lib().gasma_render_cigar(ops.ctypes.data_as(C.c_void_p), len(ops), buf, 64)
    assert buf.value == b"22M1D50M1I3=2X" and n == 14
    assert pkg.render_ops(ops) == "22M1D50M1I3=2X"
