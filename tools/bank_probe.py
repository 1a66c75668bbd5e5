"""Tuning aid: BASELINE config 5 shapes -- 4096x4096 planes (generic tiled path, db4 level 5) and a slice of the
512x512x3x3 bank (db4 clamps to level 0 on 3x3) -- device-resident, per-phase times."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from wavelettransforms_b200 import _lib, engine

dev = torch.device('cuda', 0)
torch.cuda.set_device(0)
stream = int(torch.cuda.current_stream().cuda_stream)


def run(tensors, wavelet, level, label, reps=5):
    outs = [torch.empty_like(t) for t in tensors]
    items = [(t.data_ptr(), o.data_ptr()) + engine.plane_geometry(t.shape) + (level,) for t, o in zip(tensors, outs)]
    plan = _lib.Plan(items, wavelet, 'periodization', _lib.THRESHOLD_PERCENTILE, 50.0, 'legacy_1_22')
    for _ in range(2):
        plan.run(stream)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        plan.run(stream)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    plan.set_timing(True)   # per-phase events serialise the stream lanes: after the step time
    for _ in range(reps):
        plan.run(stream)
    torch.cuda.synchronize()
    n = sum(t.numel() for t in tensors)
    print(f'{label}: {ms:.3f} ms/step, {n / ms / 1e6:.2f} G elems/s, {8 * n / ms / 1e6:.1f} GB/s algorithmic, launches {plan.num_launches}')
    ph = plan.phase_times()
    agg = {}
    for name, t, ln in ph:
        key = name.split('_layer')[0]
        agg[key] = agg.get(key, 0.0) + t
    print('   ', {k: round(v, 4) for k, v in agg.items()})
    plan.close()


g = torch.Generator(device='cuda').manual_seed(1)
big = [torch.randn(4096, 4096, device=dev, generator=g) * 0.02 for _ in range(4)]
run(big, 'db4', 5, '4 x 4096x4096 db4 L5 (generic path)')
run(big, 'haar', 5, '4 x 4096x4096 haar L5 (generic path)')
bank = [torch.randn(512, 512, 3, 3, device=dev, generator=g) * 0.02 for _ in range(64)]
run(bank, 'db4', 0, '64 x 512x512x3x3 level 0 (db4 clamped)')
run(bank, 'haar', 1, '64 x 512x512x3x3 haar L1')
