"""Build libgasma_b200.so in-tree with nvcc for sm_100a only (cross-compiles without a GPU)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB = os.path.join(HERE, "libgasma_b200.so")
SOURCES = [os.path.join(HERE, "csrc", "gasma_b200.cu")]
HEADERS = [os.path.join(HERE, "csrc", f) for f in os.listdir(os.path.join(HERE, "csrc")) if f.endswith((".cuh", ".inl"))] + [
    os.path.join(ROOT, "include", "gasma_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared",
    # libstdc++ is only available as a static archive in this image: keep its symbols private so they do
    # not interpose with the libstdc++.so.6 that python extensions load
    "-Xlinker", "-Bsymbolic", "-Xlinker", "--exclude-libs,ALL",
    "-ldl",         # NCCL is bound at run time with dlopen (csrc/nccl_entry.inl)
    "-lpthread",    # host threads of the asynchronous call pair (csrc/async_entry.inl)
    "--fmad=true",  # FP64 on the parity path uses explicit __dmul_rn/__dadd_rn, never contracted
]


def nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(p) > t for p in SOURCES + HEADERS + [os.path.abspath(__file__)])


def source_sha():
    """SHA-1 (16 hex digits) over the CUDA sources, the public header and the compiler flags: identifies WHAT was built.
    nvcc does not produce the same bytes twice, so the library file's own hash changes with every rebuild; profiles
    taken from one build (profiles/traffic.json) are matched to a later build of the same sources by this hash."""
    import hashlib
    h = hashlib.sha1()
    for path in sorted(SOURCES + HEADERS):
        h.update(os.path.basename(path).encode() + b"\0")
        with open(path, "rb") as f:
            h.update(f.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()[:16]


def build_library(force=False, verbose=False):
    """Compile the CUDA extension if sources are newer than the .so.  Returns the library path."""
    if not force and not stale():
        return LIB
    cmd = [nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB] + SOURCES
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("nvcc failed building libgasma_b200.so")
    if verbose:
        sys.stderr.write(res.stderr)
    return LIB


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
