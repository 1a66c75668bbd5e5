"""The kernels' peer mirror (wt_plan_set_peers: every pruned weight is also stored at out + delta) on ONE GPU: the "peers"
are two more buffers of the same device, so the mirrored stores can be checked without a second rank.  Covers the
level-0 apply of the select with a MISALIGNED input view (the apply loop then runs scalar while the mirror
vectorises on the output's alignment), the small-plane synthesis and the fused kernels."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(monkeypatch, mode, shapes, levels, wavelet, kind, value, misalign):
    from wavelettransforms_b200 import _lib, engine
    monkeypatch.setenv('WT_FUSED', mode)
    dev = torch.device('cuda', 0)
    g = torch.Generator(device=dev).manual_seed(11)
    xs = []
    for sh in shapes:
        n = int(np.prod(sh))
        buf = torch.randn(n + 8, device=dev, generator=g) * 0.05
        xs.append(buf[1:1 + n].view(sh) if misalign else buf[:n].view(sh))     # 4-byte offset: not 16-byte aligned
    total = sum(((x.numel() + 63) // 64) * 64 for x in xs)
    bufs = [torch.full((total,), float('nan'), device=dev) for _ in range(3)]       # out + two "peers"
    outs, off = [], 0
    for x in xs:
        outs.append(bufs[0][off:off + x.numel()].view(x.shape))
        off += ((x.numel() + 63) // 64) * 64
    items = [(x.data_ptr(), o.data_ptr()) + engine.plane_geometry(x.shape) + (lv,) for x, o, lv in zip(xs, outs, levels)]
    plan = _lib.Plan(items, wavelet, 'periodization', kind, value, 'legacy_1_22')
    plan.set_peers([bufs[1].data_ptr() - bufs[0].data_ptr(), bufs[2].data_ptr() - bufs[0].data_ptr()])
    s = int(torch.cuda.current_stream().cuda_stream)
    for _ in range(2):
        plan.run(s)
    torch.cuda.synchronize()
    plan.close()
    a, b, c = (t.cpu().numpy().view(np.uint32) for t in bufs)
    assert not np.isnan(bufs[0][:xs[0].numel()].cpu().numpy()).any()
    assert np.array_equal(a, b) and np.array_equal(a, c)


@pytest.mark.parametrize('misalign', [False, True])
@pytest.mark.parametrize('mode', ['off', 'all'])
def test_level0_apply_mirror(monkeypatch, mode, misalign):
    from wavelettransforms_b200 import _lib
    shapes = [(64, 32, 3, 3), (300, 301), (4099,)]
    _run(monkeypatch, mode, shapes, [0, 0, 0], 'bior4.4', _lib.THRESHOLD_PERCENTILE, 50.0, misalign)


@pytest.mark.parametrize('misalign', [False, True])
@pytest.mark.parametrize('mode', ['off', 'all'])
def test_small_plane_mirror(monkeypatch, mode, misalign):
    from wavelettransforms_b200 import _lib
    shapes = [(64, 3, 7, 7), (70, 33, 3, 3), (128, 64, 1, 1)]
    _run(monkeypatch, mode, shapes, [2, 1, 1], 'haar', _lib.THRESHOLD_PERCENTILE, 60.0, misalign)


def test_absolute_mirror_on_the_fused_path(monkeypatch):
    """absolute thresholds mirror on the fused kernels only: level-0 layers by default, small planes with WT_FUSED=all"""
    from wavelettransforms_b200 import _lib
    _run(monkeypatch, 'abs', [(50, 50), (4099,), (16, 8, 3, 3)], [0, 0, 0], 'db2', _lib.THRESHOLD_ABSOLUTE, 0.02, False)
    _run(monkeypatch, 'all', [(70, 33, 3, 3), (128, 64, 1, 1), (50, 50)], [1, 1, 0], 'db2', _lib.THRESHOLD_ABSOLUTE, 0.02, False)
