"""The C++ host side (approximate-string-matching_b200/host/): the reference-shaped adapters hurdle_matrix_gpu /
benchmark_gpu and the benchmark.cpp counterpart, built with g++ against the C ABI."""
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "approximate-string-matching_b200", "host")
BIN = os.path.join(HOST, "_build")


@pytest.fixture(scope="module")
def drivers(pkg):
    import importlib.util
    spec = importlib.util.spec_from_file_location("gasma_b200_build", os.path.join(os.path.dirname(HOST), "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod.build_library()
    subprocess.run(["make", "-C", HOST], check=True, stdout=subprocess.DEVNULL)
    return BIN


def read_seq_file(path):
    reads, refs = [], []
    with open(path) as f:
        lines = f.read().split("\n")
    for i in range(0, len(lines) - 1, 2):
        assert lines[i][0] == ">" and lines[i + 1][0] == "<"
        reads.append(lines[i][1:])
        refs.append(lines[i + 1][1:])
    return reads, refs


def test_dataset_file_format(drivers, ol, tmp_path):
    """benchmark_dataset.h writes the reference's '>read / <ref' format with the requested edit budget."""
    out = subprocess.run([os.path.join(drivers, "gasma-benchmark"), "--dataset-only", "--pairs", "300", "--length", "100"],
                         cwd=tmp_path, check=True, capture_output=True, text=True).stdout.split()
    assert out == ["simulated_300_100_%s_eq.seq" % r for r in ("0.050000", "0.100000", "0.150000", "0.200000")]
    for name, e in zip(out, (0.05, 0.10, 0.15, 0.20)):
        reads, refs = read_seq_file(os.path.join(tmp_path, name))
        assert len(reads) == 300 and all(len(r) == 100 and set(r) <= set("ACGT") for r in reads)
        ne = int(np.ceil(np.float32(100) * np.float32(e)))
        assert all(abs(len(f) - 100) <= ne for f in refs)
        b = ol.batch_from_strings(reads, refs)
        pen = ol.oracle_nw(b, 1, 1, 1, threads=2, want_ops=False)["penalty"]
        assert pen.max() <= ne and pen.mean() > 0.5 * ne      # exactly ne edits applied, some cancel out


def test_drivers_fail_loudly_without_gpu(drivers):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    r = subprocess.run([os.path.join(drivers, "gasma-demo")], capture_output=True, text=True)
    assert r.returncode == 1 and "no CPU path" in r.stderr


@pytest.mark.gpu
def test_demo_pair_is_golden_g1(drivers):
    """GASMA/main.cpp:7-8 through hurdle_matrix_gpu: cost 6, 22M1D50M1D28M."""
    r = subprocess.run([os.path.join(drivers, "gasma-demo")], capture_output=True, text=True, check=True)
    assert "CIGAR: 22M1D50M1D28M" in r.stdout and "cost: 6" in r.stdout


@pytest.mark.gpu
def test_benchmark_driver_report_matches_oracle(drivers, ol, tmp_path):
    """gasma-benchmark on a simulated file: the report's percentages equal the oracle's counters on the same file."""
    subprocess.run([os.path.join(drivers, "gasma-benchmark"), "--dataset-only", "--pairs", "20000", "--length", "100"],
                   cwd=tmp_path, check=True, capture_output=True)
    name = "simulated_20000_100_0.200000_eq.seq"
    r = subprocess.run([os.path.join(drivers, "gasma-benchmark"), name, "--pairs", "20000"], cwd=tmp_path,
                       capture_output=True, text=True, check=True)
    reads, refs = read_seq_file(os.path.join(tmp_path, name))
    b = ol.batch_from_strings(reads, refs)
    og = ol.oracle_greedy(b, ol.make_params(), threads=8)
    on = ol.oracle_nw(b, 1, 1, 1, threads=8)
    total, nwc, gc, cov = [int(v) for v in ol.oracle_metrics(b, og, on)]
    txt = r.stdout
    assert "Total number of alignments: %d\n" % total in txt
    acc = re.search(r"\[Accuracy\].*?=> Needleman-Wunsch \| ([\d.]+) %.*?=> Greedy           \| ([\d.]+) %", txt, re.S)
    covm = re.search(r"\[Coverage\].*?=> Greedy           \| ([\d.]+) %", txt, re.S)
    assert acc.group(1) == "%.3f" % (nwc / total * 100)
    assert acc.group(2) == "%.3f" % (gc / total * 100)
    assert covm.group(1) == "%.3f" % (cov / total * 100)
    # the LEAP rows: penalties of the LV restatement (pinned to the unmodified class) against the NW optimum
    leap = ol.oracle_leap(b, threads=8)
    lacc = re.search(r"\[Accuracy\].*?=> LEAP             \| ([\d.]+) %", txt, re.S)
    assert lacc.group(1) == "%.3f" % (int((leap == on["penalty"]).sum()) / total * 100)
    assert re.search(r"\[Time\].*?=> LEAP             \| \d+\.\d{3} s", txt, re.S)
