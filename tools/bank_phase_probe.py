"""Per-launch-group device times of the 4096 x 4096 db4 L5 plan (generic tiled path): python tools/bank_phase_probe.py [n_planes] [wavelet] [level]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from wavelettransforms_b200 import _lib, engine

if __name__ == '__main__':
    n_planes = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    wavelet = sys.argv[2] if len(sys.argv) > 2 else 'db4'
    level = int(sys.argv[3]) if len(sys.argv) > 3 else 5
    dev = torch.device('cuda', 0)
    torch.cuda.set_device(0)
    g = torch.Generator(device=dev).manual_seed(1000)
    xs = [torch.randn((4096, 4096), device=dev, generator=g) * 0.02 for _ in range(n_planes)]
    outs = [torch.empty_like(x) for x in xs]
    items = [(x.data_ptr(), o.data_ptr()) + engine.plane_geometry(x.shape) + (level,) for x, o in zip(xs, outs)]
    plan = _lib.Plan(items, wavelet, 'periodization', _lib.THRESHOLD_PERCENTILE, 50.0, 'legacy_1_22')
    if os.environ.get('WT_PROBE_ARITH'):
        plan.set_arithmetic(os.environ['WT_PROBE_ARITH'])
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        st = int(s.cuda_stream)
        for graph in (False, True):
            plan.set_graph(graph)
            for _ in range(4):
                plan.run(st)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20):
                plan.run(st)
            e1.record()
            torch.cuda.synchronize()
            print(f"{'graph' if graph else 'plain'}: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us/step, {plan.num_launches} launches", flush=True)
        plan.set_graph(False)
        plan.set_timing(True)
        acc = {}
        for _ in range(10):
            plan.run(st)
            for name, ms, ln in plan.phase_times():
                a = acc.setdefault(name, [0.0, ln])
                a[0] += ms
        print(' '.join(f'{k}={v[0] / 10 * 1e3:.1f}({v[1]})' for k, v in acc.items()), flush=True)
    plan.close()
