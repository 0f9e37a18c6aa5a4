"""sevenn_colsum_f64_noalloc (static device scratch, no allocation) against sevenn_colsum_f64, through the C ABI.

Collected last on purpose (file name).  The entry point has no counterpart in the
reference; see include/sevenn_b200.h.
"""
import ctypes

import pytest

pytestmark = pytest.mark.gpu


def test_colsum_noalloc_matches_colsum_twice_in_a_row():
    import torch
    if not torch.cuda.is_available():
        pytest.skip('needs a CUDA device')
    from sevennet_exp_b200 import _lib
    lib = _lib.load()
    dev = torch.device('cuda:0')
    torch.cuda.set_device(dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    g = torch.Generator(device='cpu').manual_seed(11)
    # two calls in a row with different sizes and contents: the second must not see the first one's partials
    for n, D, scale in ((100_003, 6, 1.0), (4_099, 3, 250.0), (0, 1, 1.0)):
        a = (torch.randn(max(n, 1), D, generator=g, dtype=torch.float32) * scale)[:n].contiguous().to(dev)
        part = torch.full((296 * D,), float('nan'), dtype=torch.float64, device=dev)
        want = torch.full((D,), float('nan'), dtype=torch.float64, device=dev)
        got = torch.full((D,), float('nan'), dtype=torch.float64, device=dev)
        ap = ctypes.c_void_p(a.data_ptr() if n else 0)
        lib.call('sevenn_colsum_f64', ap, n, D, ctypes.c_void_p(part.data_ptr()), ctypes.c_void_p(want.data_ptr()),
                 stream)
        lib.call('sevenn_colsum_f64_noalloc', ap, n, D, ctypes.c_void_p(got.data_ptr()), stream)
        torch.cuda.synchronize()
        # same kernels, same fixed summation order: bit-exact
        assert torch.equal(got, want), (n, D, got, want)
        ref = a.double().sum(0) if n else torch.zeros(D, dtype=torch.float64, device=dev)
        assert torch.allclose(got, ref, rtol=1e-12, atol=1e-9 * scale)
    # sizes beyond the stated maximum are rejected, not truncated
    rc = lib.cdll.sevenn_colsum_f64_noalloc(ap, 0, 9, ctypes.c_void_p(got.data_ptr()), stream)
    assert rc == 10001
