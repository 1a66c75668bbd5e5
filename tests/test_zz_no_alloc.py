"""wt_select_abs_noalloc_f32 (no counterpart in the reference): the select
with its scratch in a fixed device array inside the library, called through the C ABI, against the allocating entry point
wt_select_abs_f32 -- bit for bit (the tolerance of every integer / order-statistic test here), on two calls in a row that
share the scratch.  Collected last on purpose."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _both(L, x, zeros, k0, k1):
    s = int(torch.cuda.current_stream().cuda_stream)
    out = []
    for fn in (L.wt_select_abs_f32, L.wt_select_abs_noalloc_f32):
        ab = torch.full((2,), -1.0, dtype=torch.float32, device=x.device)
        rc = fn(x.data_ptr(), x.numel(), int(zeros), int(k0), int(k1), ab.data_ptr(), ctypes.c_void_p(s))
        assert rc == 0, L.wt_last_error()
        torch.cuda.synchronize()
        out.append(ab.cpu().numpy().view(np.uint32).copy())
    return out


def test_noalloc_select_matches_allocating_select_twice_in_a_row():
    from wavelettransforms_b200 import _lib
    L = _lib.lib()
    dev = torch.device('cuda', 0)
    torch.cuda.set_device(0)
    g = torch.Generator(device=dev).manual_seed(7)
    # two calls in a row over different populations: the second finds the first one's histograms and candidates in the scratch
    cases = [(torch.randn(2_359_296, device=dev, generator=g) * 0.02, 0, 0.5),
             (torch.randn(700_001, device=dev, generator=g).abs() * 3.0, 123_456, 0.9),
             (torch.randn(4097, device=dev, generator=g), 0, 0.1)]
    for x, zeros, q in cases:
        pop = x.numel() + zeros
        k0 = int((pop - 1) * q)
        k1 = min(k0 + 1, pop - 1)
        want, got = _both(L, x, zeros, k0, k1)
        assert np.array_equal(want, got), (x.numel(), zeros, q)
        srt = torch.sort(torch.cat([x.abs(), torch.zeros(zeros, device=dev)])).values
        assert got.view(np.float32)[0] == srt[k0].item() and got.view(np.float32)[1] == srt[k1].item()


def test_noalloc_select_rejects_sizes_beyond_its_scratch():
    from wavelettransforms_b200 import _lib
    L = _lib.lib()
    dev = torch.device('cuda', 0)
    x = torch.zeros(8, device=dev)
    ab = torch.zeros(2, device=dev)
    n = (1 << 22) + 1
    rc = L.wt_select_abs_noalloc_f32(x.data_ptr(), n, 0, 0, 1, ab.data_ptr(), ctypes.c_void_p(0))
    assert rc == _lib.WT_E_INVALID
    assert b'exceeds the fixed scratch' in L.wt_last_error()
