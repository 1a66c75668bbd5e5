"""Times one plan per workload (tuning aid): python tools/fused_probe.py [workload ...]
ABS=1: absolute threshold (the apply pass alone); STEPS=n; WT_LEGACY_PATH=1 for the round-1 kernels."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import bench
from wavelettransforms_b200 import _lib, engine


def main():
    names = sys.argv[1:] or ['resnet18_bior4.4_L5_p50_clamped', 'resnet18_haar_L5_p50_clamped', 'resnet18_bior4.4_L5_p50_unclamped']
    dev = torch.device('cuda', 0)
    torch.cuda.set_device(0)
    steps = int(os.environ.get('STEPS', '50'))
    for name in names:
        wl = bench.DeviceWorkload(name, 8, dev)
        if os.environ.get('ABS'):
            for i, p in enumerate(wl.plans):
                p.close()
                items = [(x.data_ptr(), o.data_ptr()) + engine.plane_geometry(x.shape) + (lv,)
                         for x, o, lv in zip(wl.ins[i], wl.outs[i], wl.levels)]
                wl.plans[i] = _lib.Plan(items, wl.wavelet, 'periodization', _lib.THRESHOLD_ABSOLUTE, 0.01, 'legacy_1_22')
        ms, _ = bench.time_device_steps(wl, steps, 5, 1, dev)
        stream = int(torch.cuda.current_stream().cuda_stream)
        wl.plans[0].run(stream)
        wl.plans[0].results(stream)
        print(f'{name}: {ms / steps * 1e3:.1f} us/step, {8 * wl.elems / (ms / steps * 1e-3) / 1e9:.0f} GB/s', flush=True)
        wl.close()


if __name__ == '__main__':
    main()
