"""Level-0 percentile prune of N tensors of 512x512x3x3 (bank slice) and of ResNet-152 bior2.2 L3 clamped: fused vs legacy (tuning aid)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from wavelettransforms_b200 import _lib, engine


def run(shapes, label, pct=50.0):
    dev = torch.device('cuda', 0)
    g = torch.Generator(device=dev).manual_seed(1000)
    xs = [torch.randn(sh, device=dev, generator=g) * 0.02 for sh in shapes]
    outs = [torch.empty_like(x) for x in xs]
    items = [(x.data_ptr(), o.data_ptr()) + engine.plane_geometry(x.shape) + (0,) for x, o in zip(xs, outs)]
    plan = _lib.Plan(items, 'db4', 'periodization', _lib.THRESHOLD_PERCENTILE, pct, 'legacy_1_22')
    stream = int(torch.cuda.current_stream().cuda_stream)
    for _ in range(3):
        plan.run(stream)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    n = 20
    for _ in range(n):
        plan.run(stream)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    elems = sum(x.numel() for x in xs)
    plan.results(stream)
    print(f'{label} p={pct}: {ms * 1e3:.1f} us, {8 * elems / (ms * 1e-3) / 1e9:.0f} GB/s ({8 * elems / (ms * 1e-3) / 1e9 / 6546.6:.2f} of measured HBM)', flush=True)
    plan.close()


if __name__ == '__main__':
    for n in (4, 32, 128):
        run([(512, 512, 3, 3)] * n, f'bank {n} x 512x512x3x3')
    run([(512, 512, 3, 3)] * 32, 'bank 32 x 512x512x3x3', 90.0)
