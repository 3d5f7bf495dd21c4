"""One pass over every kernel and entry point at several (max_length, k) shapes, checked against the oracle
(tools/sanitize_smoke.py, also the script run under a memory checker when one is available)."""
import os
import runpy

import pytest

pytestmark = pytest.mark.gpu


def test_every_kernel_small_shapes():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    runpy.run_path(os.path.join(root, "tools", "sanitize_smoke.py"), run_name="__main__")
