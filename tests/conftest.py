"""pytest plumbing: markers, the importlib loader for the (hyphenated) package, checker-library builds."""
import importlib.util
import json
import os
import subprocess
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
PKG_DIR = os.path.join(ROOT, "approximate-string-matching_b200")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


def load_pkg():
    """Import approximate-string-matching_b200/ under the module name gasma_b200."""
    if "gasma_b200" in sys.modules:
        return sys.modules["gasma_b200"]
    spec = importlib.util.spec_from_file_location("gasma_b200", os.path.join(PKG_DIR, "__init__.py"),
                                                  submodule_search_locations=[PKG_DIR])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["gasma_b200"] = mod
    spec.loader.exec_module(mod)
    return mod


@pytest.fixture(scope="session")
def pkg():
    return load_pkg()


@pytest.fixture(scope="session")
def ol():
    import oracle_lib
    oracle_lib.build_oracle()
    return oracle_lib


@pytest.fixture(scope="session")
def host_sim():
    """g++ build of the kernels' __host__ __device__ per-pair code (tests/host_sim.cpp)."""
    import ctypes as C
    out_dir = os.path.join(HERE, "_build")
    os.makedirs(out_dir, exist_ok=True)
    lib = os.path.join(out_dir, "libhost_sim.so")
    src = os.path.join(HERE, "host_sim.cpp")
    deps = [src] + [os.path.join(PKG_DIR, "csrc", f) for f in os.listdir(os.path.join(PKG_DIR, "csrc"))]
    if not os.path.exists(lib) or any(os.path.getmtime(d) > os.path.getmtime(lib) for d in deps):
        subprocess.run(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-x", "c++",
                        "-Wl,-Bsymbolic", "-Wl,--exclude-libs,ALL", "-o", lib, src], check=True)
    return C.CDLL(lib)


@pytest.fixture(scope="session")
def golden():
    def load(name):
        with open(os.path.join(HERE, "golden", name)) as f:
            return json.load(f)
    return load
