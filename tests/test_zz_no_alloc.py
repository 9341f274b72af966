"""b200_lg_assignment_static: the allocation-free variant of b200_lg_assignment (scratch in a file-scope device
array of the library).  Collected last on purpose.

The entry has no counterpart in the reference; it must reproduce the existing entry (same kernels on a different
scratch address => identical results; the comparison uses the tolerance of test_log_assignment_and_filter), on
two calls in a row, the second with different inputs and a different shape so stale scratch contents would show.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


def _inputs(B, M, N, seed, dev):
    g = torch.Generator().manual_seed(seed)
    sim = (torch.randn(B, M, N, generator=g) * 4).to(dev)
    z0 = (torch.randn(B, M, generator=g) * 3).to(dev)
    z1 = (torch.randn(B, N, generator=g) * 3).to(dev)
    return sim, z0, z1


def _call_static(lib, sim, z0, z1, threshold):
    """Straight through the C ABI: the caller owns inputs and outputs, the library owns the scratch."""
    B, M, N = sim.shape
    dev = sim.device
    out = (torch.empty(B, M + 1, N + 1, dtype=torch.float32, device=dev),
           torch.empty(B, M, dtype=torch.int64, device=dev), torch.empty(B, N, dtype=torch.int64, device=dev),
           torch.empty(B, M, dtype=torch.float32, device=dev), torch.empty(B, N, dtype=torch.float32, device=dev))
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.b200_lg_assignment_static(sim.data_ptr(), z0.data_ptr(), z1.data_ptr(), B, M, N, float(threshold),
                                       *[t.data_ptr() for t in out], stream)
    return rc, out


def test_static_scratch_matches_workspace_entry_twice(dev):
    from glue_factory_b200 import _lib, ops

    lib = _lib.lib()
    assert lib.b200_lg_assignment_static_capacity() == 4 << 20
    torch.cuda.set_device(dev)
    for call, (B, M, N) in enumerate([(2, 700, 1100), (3, 515, 300)]):
        sim, z0, z1 = _inputs(B, M, N, 100 + call, dev)
        assert lib.b200_lg_assignment_workspace(B, M, N) <= lib.b200_lg_assignment_static_capacity()
        want = ops.lg_assignment(sim, z0, z1, 0.1)  # existing entry, torch-allocated workspace
        rc, got = _call_static(lib, sim, z0, z1, 0.1)
        assert rc == 0, lib.b200_last_error()
        torch.cuda.synchronize()
        np.testing.assert_allclose(got[0].cpu().numpy(), want[0].cpu().numpy(), rtol=1e-5, atol=2e-5)
        assert torch.equal(got[1], want[1]) and torch.equal(got[2], want[2])
        np.testing.assert_allclose(got[3].cpu().numpy(), want[3].cpu().numpy(), rtol=1e-5)
        np.testing.assert_allclose(got[4].cpu().numpy(), want[4].cpu().numpy(), rtol=1e-5)
        assert (got[1] >= 0).any()


def test_static_scratch_rejects_oversized_problems(dev):
    from glue_factory_b200 import _lib

    lib = _lib.lib()
    B, M, N = 8, 4096, 4096
    assert lib.b200_lg_assignment_workspace(B, M, N) > lib.b200_lg_assignment_static_capacity()
    one = torch.zeros(1, device=dev)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = one.data_ptr()  # never dereferenced: the size check comes first
    rc = lib.b200_lg_assignment_static(p, p, p, B, M, N, 0.1, p, p, p, p, p, stream)
    assert rc != 0
    msg = lib.b200_last_error().decode()
    assert "assignment_static" in msg and "bytes" in msg
