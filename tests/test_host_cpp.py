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


@pytest.mark.gpu
def test_benchmark_driver_sharded_over_devices(drivers, tmp_path):
    """Batches dealt to several contexts (one host thread per device entry) give the single-context report: the only
    cross-device quantity is the counter sum.  Uses two real GPUs when the box has them, else device 0 twice."""
    import torch
    exe = os.path.join(drivers, "gasma-benchmark")
    subprocess.run([exe, "--dataset-only", "--pairs", "20000", "--length", "100"], cwd=tmp_path, check=True, capture_output=True)
    name = "simulated_20000_100_0.150000_eq.seq"

    def report(*extra):
        txt = subprocess.run([exe, name, "--pairs", "20000", *extra], cwd=tmp_path, capture_output=True, text=True, check=True).stdout
        report.tail = txt[txt.index("[Device time, ms]"):]
        return txt[txt.index("Total number of alignments"):txt.index("[Time]")] + txt[txt.index("[Accuracy]"):txt.index("[Device time, ms]")]

    one = report()
    assert report("--batch", "3000", "--devices", "0,0,0") == one
    assert "counters summed by host (contexts share a device)" in report.tail
    if torch.cuda.device_count() >= 2:
        assert report("--batch", "3000", "--gpus", "2") == one
        assert "2 GPUs, counters summed by NCCL 2" in report.tail      # the path's only collective, from C++


def _mapper_inputs(tmp_path, n_reads=400, seed=11):
    """A two-contig reference, reads of 100 bp sampled from it with <= 3 edits, as FASTA / FASTQ files."""
    rng = np.random.default_rng(seed)
    contigs = ["".join(rng.choice(list("ACGT"), n)) for n in (60000, 25000)]
    with open(tmp_path / "ref.fa", "w") as f:
        for i, c in enumerate(contigs):
            f.write(">chr%d some description\n" % (i + 1))
            for j in range(0, len(c), 70):
                f.write(c[j:j + 70] + "\n")
    truth = []
    with open(tmp_path / "reads.fq", "w") as f:
        for r in range(n_reads):
            cid = int(rng.integers(0, 2))
            pos = int(rng.integers(1, len(contigs[cid]) - 110))
            seq = list(contigs[cid][pos:pos + 100])
            for _ in range(int(rng.integers(0, 4))):
                kind = rng.random()
                j = int(rng.integers(0, len(seq)))
                if kind < 0.7:
                    seq[j] = "ACGT"[int(rng.integers(0, 4))]
                elif kind < 0.85:
                    del seq[j]
                else:
                    seq.insert(j, "ACGT"[int(rng.integers(0, 4))])
            seq = "".join(seq)
            f.write("@read%d extra\n%s\n+\n%s\n" % (r, seq, "I" * len(seq)))
            truth.append((cid, pos, seq))
    return contigs, truth


def test_mapper_search_stage(drivers, tmp_path):
    """gasma-mapper --hits-only (FASTA/FASTQ parsing + seed index; no GPU involved) finds the sampled placements."""
    contigs, truth = _mapper_inputs(tmp_path)
    out = subprocess.run([os.path.join(drivers, "gasma-mapper"), "-r", "ref.fa", "-q", "reads.fq", "-e", "3", "--hits-only"],
                         cwd=tmp_path, check=True, capture_output=True, text=True).stdout.strip().split("\n")
    assert len(out) == len(truth)
    good = 0
    for line, (cid, pos, _) in zip(out, truth):
        name, rid, begin = line.split()
        if rid != "-" and int(rid) == cid and abs(int(begin) - pos) <= 3:
            good += 1
    assert good >= 0.98 * len(truth)
    r = subprocess.run([os.path.join(drivers, "gasma-mapper"), "-r", "ref.fa"], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 255 and "PARSER ERROR" in r.stderr


@pytest.mark.gpu
def test_mapper_sam_matches_oracle(drivers, ol, tmp_path):
    """gasma-mapper end to end: every SAM record's MAPQ is 60 + the greedy cost of (read, window) as the reference's
    mapper computes it (mapper/main.cpp:75-91), checked against the oracle; the CIGAR is the oracle's, SAM-merged;
    two chunks give the same file as one."""
    contigs, truth = _mapper_inputs(tmp_path, n_reads=600, seed=12)
    exe = os.path.join(drivers, "gasma-mapper")
    subprocess.run([exe, "-r", "ref.fa", "-q", "reads.fq", "-e", "3", "-o", "a.sam"], cwd=tmp_path, check=True)
    subprocess.run([exe, "-r", "ref.fa", "-q", "reads.fq", "-e", "3", "-o", "b.sam", "--chunk", "250"], cwd=tmp_path, check=True)
    a = open(tmp_path / "a.sam").read()
    assert a == open(tmp_path / "b.sam").read()
    lines = a.split("\n")
    assert lines[0].startswith("@HD") and lines[1] == "@SQ\tSN:chr1\tLN:60000" and lines[2] == "@SQ\tSN:chr2\tLN:25000"
    recs = [l.split("\t") for l in lines if l and not l.startswith("@")]
    assert len(recs) >= 0.98 * len(truth)
    pairs = []
    for rec in recs:
        cid = int(rec[2][3:]) - 1
        start = int(rec[3]) - 3                      # POS = start + 2, written 1-based
        pairs.append((rec[9], contigs[cid][start:start + len(rec[9]) + 1]))
    b = ol.batch_from_strings([p[0] for p in pairs], [p[1] for p in pairs])
    r = ol.oracle_greedy(b, ol.make_params(3, 1, 1, 1), threads=4)
    for i, rec in enumerate(recs):
        assert int(rec[4]) == 60 + int(r["cost"][i])
        merged, run = [], None
        for op in r["ops"][r["op_off"][i]:r["op_off"][i + 1]]:
            code, ln = int(op) & 7, int(op) >> 3
            if run and run[0] == code:
                run[1] += ln
            else:
                run = [code, ln]
                merged.append(run)
        assert rec[5] == ("".join("%d%s" % (ln, "MID=X"[c]) for c, ln in merged) or "*")
        assert rec[0].startswith("read") and rec[10] == "I" * len(rec[9])


# ---- the reference's OWN drivers, unmodified, against include/compat (drop-in boundary) ---------------------------
REF = "/root/reference/GASMA"
REFDRV = os.path.join(ROOT, "tests", "_build", "refdrv")


def build_reference_drivers():
    """Compile scratch copies of GASMA/main.cpp and GASMA/benchmark/benchmark.cpp -- byte for byte the reference's
    files, placed away from the reference's headers because a quoted #include looks beside the source first -- with
    -I include/compat.  Outputs stay under tests/_build (git-ignored; they travel to the GPU box)."""
    os.makedirs(REFDRV, exist_ok=True)
    built = {}
    for name, src in (("ref_main", "main.cpp"), ("ref_benchmark", "benchmark/benchmark.cpp")):
        text = open(os.path.join(REF, src), "rb").read()
        scratch = os.path.join(REFDRV, name + ".cpp")
        with open(scratch, "wb") as f:
            f.write(text)
        out = os.path.join(REFDRV, name)
        subprocess.run(["g++", "-std=c++17", "-O2", "-pthread", "-I", os.path.join(ROOT, "include", "compat"), "-o", out, scratch,
                        "-L" + os.path.dirname(HOST), "-lgasma_b200", "-Wl,-rpath," + os.path.dirname(HOST)], check=True)
        os.remove(scratch)          # no reference source is kept in the tree, not even in the scratch directory
        built[name] = out
    return built


def test_reference_drivers_compile_unmodified(drivers):
    if not os.path.isdir(REF):
        pytest.skip("/root/reference absent (GPU box): the binaries were built in the container")
    built = build_reference_drivers()
    assert all(os.path.exists(p) for p in built.values())
    import torch
    if not torch.cuda.is_available():
        r = subprocess.run([built["ref_main"]], capture_output=True, text=True)
        assert r.returncode != 0 and "no CPU path" in r.stderr      # fails loudly, no fallback


@pytest.mark.gpu
def test_reference_demo_runs_unmodified(drivers, ol):
    """GASMA/main.cpp:5-19 compiled unmodified: prints the 7 lanes, then the CIGAR, the cost and the LCM line."""
    exe = os.path.join(REFDRV, "ref_main")
    if not os.path.exists(exe):
        if not os.path.isdir(REF):
            pytest.skip("tests/_build/refdrv/ref_main was not built in the container")
        exe = build_reference_drivers()["ref_main"]
    out = subprocess.run([exe], check=True, capture_output=True, text=True).stdout.split("\n")
    read = "TGCGTCGCAGTGTCGAAAAAAATCTAAAGGGCCTACCGGGCCCCTTCATGAAGAGCAGGGAAAGAATCTCTATACTATTACTAATGGAAAGCAGTCTTGG"
    ref = "TGCGTCGCAGTGTCGAAATAAAATCTAAAGGGCCTACCGGGCCCCTTCATGAAGAGTAGGTAAAGAATCTCTAGTACTACTACTAATGGAAAGCAGTCTTGG"
    # lanes: _construct_hurdles + flip_short_hurdles(1) restated on python ints from the oracle's planes
    a0, a1 = [int(v[0]) | (int(v[1]) << 64) for v in ol.oracle_pack(read.encode(), 2)]
    b0, b1 = [int(v[0]) | (int(v[1]) << 64) for v in ol.oracle_pack(ref.encode(), 2)]
    M = (1 << 128) - 1
    for i, l in enumerate(range(-3, 4)):
        v = (((a0 >> -l) ^ b0) | ((a1 >> -l) ^ b1)) if l < 0 else (((b0 >> l) ^ a0) | ((b1 >> l) ^ a1))
        f = v & ((v >> 1) | ((v << 1) & M))
        assert out[i] == "lane %d:" % l + "".join("1" if (f >> j) & 1 else "0" for j in range(128)), l
    assert out[7] == "" and out[8] == "22M1D50M1D28M" and out[9] == "cost: 6"      # golden G1
    s1 = b"GCCCCGTCCCAGCACAGGCAGCGGGGATGCTCTAGAGCGATGTCGACCTGGGAAAGGCGCTGGGGCGCGCCGAATTCCAAAGGAGTTCCGTAAGGTTCAG"
    ops = np.array([(22 << 3), (1 << 3) | 2, (50 << 3), (1 << 3) | 2, (28 << 3)], np.uint16)
    buf = np.zeros(256, np.uint8)
    n = ol.oracle().go_lcm(ol._ptr(np.frombuffer(s1, np.uint8)), len(s1), ol._ptr(ops), len(ops), 3, ol._ptr(buf), 256)
    assert out[10] == buf[:n].tobytes().decode()


@pytest.mark.gpu
def test_reference_benchmark_driver_links_and_reports_missing_file(drivers):
    """GASMA/benchmark/benchmark.cpp:27-31 (USE_SIMULATED_DATA false) names a file that only exists on its author's
    machine: like the reference, the harness reports that it cannot open it and prints an empty report."""
    exe = os.path.join(REFDRV, "ref_benchmark")
    if not os.path.exists(exe):
        pytest.skip("tests/_build/refdrv/ref_benchmark was not built in the container")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert "SRR611076.data" in r.stdout + r.stderr
