"""One pass over every kernel and entry point at several (max_length, k) shapes, checked against the oracle
(tools/sanitize_smoke.py, also the script run under a memory checker when one is available)."""
import os
import runpy

import pytest

pytestmark = pytest.mark.gpu


def test_every_kernel_small_shapes():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    runpy.run_path(os.path.join(root, "tools", "sanitize_smoke.py"), run_name="__main__")


@pytest.mark.gpu
def test_allreduce_counters_over_nccl(pkg):
    """gasma_allreduce_counters: the path's only collective through the C ABI (NCCL bound at run time).  One rank per
    visible GPU (a clique of one on a single-GPU box): every device array ends up holding the sum."""
    import torch
    assert pkg.lib().gasma_nccl_version() >= 20000
    g = min(torch.cuda.device_count(), 8)
    als = [pkg.Aligner(device=d) for d in range(g)]
    for rep in range(2):      # the second call reuses the cached communicators
        ts = [torch.tensor([d + 1, 10 * (d + 1), 2**40 + d, rep], dtype=torch.int64, device="cuda:%d" % d) for d in range(g)]
        for d in range(g):
            torch.cuda.synchronize(d)
        pkg.allreduce_counters(als, ts)
        want = [g * (g + 1) // 2, 10 * g * (g + 1) // 2, g * 2**40 + g * (g - 1) // 2, rep * g]
        for t in ts:
            assert t.cpu().tolist() == want
    with pytest.raises(pkg.GasmaError):      # two ranks on one device are refused, not deadlocked
        pkg.allreduce_counters([als[0], pkg.Aligner(device=0)], [ts[0], ts[0].clone()])
    for a in als:
        a.close()
