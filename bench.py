#!/usr/bin/env python
"""bench.py -- BASELINE.json's metric: pruned weight elems/s of the DWT-prune hot path
(wavedec2 -> percentile threshold -> waverec2 on every conv weight), on synthetic random-init
ResNet conv weights.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

One "step" = one pass of the path over every conv layer of the workload's model (one grouped
plan: a fixed sequence of kernel launches, no host synchronisation).  Legs of the `ours` arm:

  value        device-resident: weights already in HBM, K steps back to back over ROTATING
               replicas of the model (their total size exceeds the 126 MB L2, so no step finds its
               input in L2), CUDA events on the launching stream, max over ranks.
  roofline     a second pass over the same K steps with libwtprune's per-phase cudaEvents on
               (wt_plan_set_timing): live average duration of every launch group; the dominant
               kernel's algorithmic bytes (8 B per weight element it processes) / its duration.
  e2e          the reference-facing call with HOST buffers: pinned host weights -> C-ABI host-I/O
               plan (H2D inside) -> pruned weights back in pinned host memory, every step.
  cpu_baseline the oracle's restatement of the reference flow (oracle/reference_flow.py) on ONE
               host core over a bounded sample of the same layers (N=1, rank 0 only).

`--impl reference` times that CPU flow on all host cores (a process pool over layers); there is
no compiled reference (it is pure Python over PyWavelets, which is absent: see DESIGN.md).

Multi-GPU (`--gpus N` under torchrun): models are independent objects, every rank prunes its own
replicas with no data-path collective ("scaling": "weak"); the layer-sharded + all-gather mode of
one model lives in wavelettransforms_b200/sharded.py and is reported under "layer_sharded".
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np
import torch

METRIC = 'pruned weight elems/s (DWT fwd+threshold+inv)'
UNIT = 'elems/s'
BYTES_PER_ELEM = 8  # read the fp32 weight once + write the pruned fp32 weight once

# name -> (arch, wavelet, level, percentile, clamp_level)
WORKLOADS = {
    # north_star target: all ResNet-18 conv layers through a level-5 bior decomposition.  Unclamped =
    # the decomposition dwt_pruning_NoEntropy.py:41 runs (3x3 planes: 3->2->1->1->1->1), with the
    # percentile threshold of dwt_pruning.py:25-32.  The real transform on every layer.
    'resnet18_bior4.4_L5_p50_unclamped': ('resnet18', 'bior4.4', 5, 50.0, False),
    # the same call through dwt_pruning.py's clamp (:64-65): dec_len 10 => level 0 on every ResNet
    # plane => select + zero, no transform.  What the reference's stored runs executed.
    'resnet18_bior4.4_L5_p50_clamped': ('resnet18', 'bior4.4', 5, 50.0, True),
    # deepest real transform the clamped reference ever runs: haar, 3x3 -> L1, 7x7 -> L2
    'resnet18_haar_L5_p50_clamped': ('resnet18', 'haar', 5, 50.0, True),
    'resnet18_bior2.2_L3_p50_unclamped': ('resnet18', 'bior2.2', 3, 50.0, False),
    # BASELINE.json configs[0]: the reference's own CPU-runnable case (dec_len 6 => clamps to level 0)
    'resnet18_bior1.3_L1_p50_clamped': ('resnet18', 'bior1.3', 1, 50.0, True),
    'resnet50_bior2.2_L3_p50_unclamped': ('resnet50', 'bior2.2', 3, 50.0, False),
    'resnet152_bior2.2_L3_p50_unclamped': ('resnet152', 'bior2.2', 3, 50.0, False),
}
DEFAULT_WORKLOAD = 'resnet18_bior4.4_L5_p50_unclamped'
VARIANTS = ['resnet18_bior4.4_L5_p50_clamped', 'resnet18_haar_L5_p50_clamped', 'resnet18_bior2.2_L3_p50_unclamped',
            'resnet18_bior1.3_L1_p50_clamped']
NUMPY_SEMANTICS = 'legacy_1_22'  # the reference's pin (requirements.txt:2)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=200)
    ap.add_argument('--warmup', type=int, default=10)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--workload', default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument('--replicas', type=int, default=8, help='rotating model replicas per GPU (defeats L2 reuse)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-variants', action='store_true')
    ap.add_argument('--sharded-workload', default='resnet152_bior2.2_L3_p50_unclamped', choices=sorted(WORKLOADS))
    ap.add_argument('--cpu-sample-seconds', type=float, default=15.0)
    return ap.parse_args()


def dist_env():
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    return rank, world, local


def bind_to_gpu_numa_node(local):
    """Pins this rank to the host cores NVML reports as local to its GPU, BEFORE any pinned allocation, so that the
    staging buffers of the end-to-end leg are first-touched on the GPU's own NUMA node (eight ranks sharing one node's
    memory controllers is what caps e2e at N = 8).  Best effort: any failure leaves the affinity alone."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = [64 * i + b for i, w in enumerate(mask) for b in range(64) if (w >> b) & 1]
        allowed = os.sched_getaffinity(0)
        cpus = [c for c in cpus if c in allowed]
        if cpus:
            os.sched_setaffinity(0, cpus)
        return len(cpus)
    except Exception:
        return 0


def effective_levels(census, wavelet, level, clamp):
    """per-layer level as the reference computes it: each layer is its own
    multi_resolution_analysis([weight], ...) call (dwt_pruning.py:116-117), clamp at :64-65"""
    from wavelettransforms_b200 import _lib
    _, F = _lib.wavelet_lookup(wavelet)
    out = []
    for _, shp in census:
        out.append(min(level, _lib.dwt_max_level(min(shp[-2:]), F)) if clamp else level)
    return out


# ---- clocks -------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi sampled every 100 ms while the timed region runs (B200_PROFILING.md clocks line)"""
    Q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
         'clocks_event_reasons.sw_power_cap')

    def __init__(self, index=0):
        self.index = index
        self.proc = None
        self.path = None

    def start(self):
        try:
            f = tempfile.NamedTemporaryFile('w', suffix='.csv', delete=False)
            self.path = f.name
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), f'--query-gpu={self.Q}',
                                          '--format=csv,noheader,nounits', '-lms', '100'], stdout=f,
                                         stderr=subprocess.DEVNULL)
            f.close()
        except Exception:
            self.proc = None

    def stop(self):
        out = {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': [], 'samples': 0}
        if self.proc is None:
            return out
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        try:
            for line in open(self.path):
                p = [x.strip() for x in line.split(',')]
                if len(p) < 7:
                    continue
                try:
                    sm.append(float(p[0]))
                    mx.append(float(p[1]))
                except ValueError:
                    continue
                for nm, v in zip(names, p[3:7]):
                    if v.lower().startswith('active'):
                        reasons.add(nm)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons), samples=len(sm))
        return out


# ---- the `ours` arm ----------------------------------------------------------------------------------
class DeviceWorkload:
    """R replicas of the model's conv weights on the GPU, one device-I/O plan per replica."""

    def __init__(self, name, replicas, dev):
        from wavelettransforms_b200 import _lib, engine, workloads
        self.name = name
        arch, self.wavelet, self.level, self.pct, self.clamp = WORKLOADS[name]
        self.census = workloads.conv_census(arch)
        self.levels = effective_levels(self.census, self.wavelet, self.level, self.clamp)
        self.elems = sum(int(np.prod(s)) for _, s in self.census)
        self.host = workloads.random_init_weights(self.census, seed=0, pin_memory=True)
        self.dev = dev
        self.ins, self.outs, self.plans = [], [], []
        for _ in range(replicas):
            xs = [w.to(dev, non_blocking=True) for w in self.host]
            os_ = [torch.empty_like(x) for x in xs]
            items = [(x.data_ptr(), o.data_ptr()) + engine.plane_geometry(x.shape) + (lv,)
                     for x, o, lv in zip(xs, os_, self.levels)]
            self.plans.append(_lib.Plan(items, self.wavelet, 'periodization', _lib.THRESHOLD_PERCENTILE, self.pct,
                                        NUMPY_SEMANTICS, host_io=False))
            self.ins.append(xs)
            self.outs.append(os_)
        torch.cuda.synchronize()

    def step(self, i, stream):
        self.plans[i % len(self.plans)].run(stream)

    def close(self):
        for p in self.plans:
            p.close()
        self.plans = []


def barrier(world):
    if world > 1:
        torch.distributed.barrier()


def max_over_ranks(x, world, dev):
    if world == 1:
        return x
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    return float(t.item())


def time_device_steps(wl, steps, warmup, world, dev):
    """ms for exactly `steps` steps (max over ranks), launches enqueued in the timed region"""
    stream = int(torch.cuda.current_stream().cuda_stream)
    for i in range(warmup):
        wl.step(i, stream)
    torch.cuda.synchronize()
    barrier(world)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    launches = 0
    for i in range(steps):
        wl.step(warmup + i, stream)
        launches += wl.plans[0].num_launches
    e1.record()
    torch.cuda.synchronize()
    barrier(world)
    ms = e0.elapsed_time(e1)
    return max_over_ranks(ms, world, dev), launches


def phase_profile(wl, steps):
    """second pass over the same steps with per-phase cudaEvents: {phase: (avg ms, launches)}"""
    stream = int(torch.cuda.current_stream().cuda_stream)
    for p in wl.plans:
        p.set_timing(True)
    acc = {}
    order = []
    for i in range(steps):
        p = wl.plans[i % len(wl.plans)]
        p.run(stream)
        for name, ms, ln in p.phase_times():
            if name not in acc:
                acc[name] = [0.0, ln, 0]
                order.append(name)
            acc[name][0] += ms
            acc[name][2] += 1
    for p in wl.plans:
        p.set_timing(False)
    return [(n, acc[n][0] / acc[n][2], acc[n][1]) for n in order]


def phase_elems(wl, phase):
    """weight elements the launch group `phase` processes"""
    def numel(sel):
        return sum(int(np.prod(s)) for (_, s), lv in zip(wl.census, wl.levels) if sel(s, lv))
    if phase == 'fused_flat':
        return numel(lambda s, lv: lv == 0)
    if phase.startswith('fused_'):
        h, w = (int(v) for v in phase.split('_', 1)[1].split('x'))
        return numel(lambda s, lv: lv > 0 and s[-2] == h and s[-1] == w)
    if phase.startswith('small_'):
        hw = phase.rsplit('_', 1)[1]
        h, w = (int(v) for v in hw.split('x'))
        return numel(lambda s, lv: lv > 0 and s[-2] == h and s[-1] == w)
    if phase.startswith('select'):
        return wl.elems
    return 0


def time_e2e(name, steps, warmup, world, dev):
    """host buffers in, host buffers out, through the C-ABI host-I/O plan; returns (ms, h2d, d2h bytes/step)"""
    from wavelettransforms_b200 import _lib, engine, workloads
    arch, wavelet, level, pct, clamp = WORKLOADS[name]
    census = workloads.conv_census(arch)
    levels = effective_levels(census, wavelet, level, clamp)
    host_in = workloads.random_init_weights(census, seed=0, pin_memory=True)
    host_out = [torch.empty_like(w).pin_memory() for w in host_in]
    items = [(x.data_ptr(), o.data_ptr()) + engine.plane_geometry(x.shape) + (lv,)
             for x, o, lv in zip(host_in, host_out, levels)]
    plan = _lib.Plan(items, wavelet, 'periodization', _lib.THRESHOLD_PERCENTILE, pct, NUMPY_SEMANTICS, host_io=True)
    stream = int(torch.cuda.current_stream().cuda_stream)
    nbytes = sum(w.numel() * 4 for w in host_in)
    for _ in range(warmup):
        plan.run(stream)
        torch.cuda.current_stream().synchronize()
    barrier(world)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        plan.run(stream)
        torch.cuda.current_stream().synchronize()   # the pruned weights are now readable on the host
    ms = (time.perf_counter() - t0) * 1e3
    barrier(world)
    counts, _, _ = plan.results(stream)
    plan.close()
    return max_over_ranks(ms, world, dev), nbytes, nbytes, host_in, host_out, counts


# ---- CPU legs (the oracle's restatement of the reference flow; the only place bench.py runs oracle/) ----
def _cpu_prune_layer(args):
    """worker: one layer through oracle.reference_flow.multi_resolution_analysis (dwt_pruning.py:35-95)"""
    idx, shape, seed_w, wavelet, level, pct, clamp = args
    torch.set_num_threads(1)
    from oracle import reference_flow as rf
    w = torch.from_numpy(seed_w).view(shape)
    t0 = time.perf_counter()
    out, zeros = rf.multi_resolution_analysis([w], wavelet, level, pct, numpy_semantics='live', clamp_level=clamp)
    return idx, time.perf_counter() - t0, zeros


def cpu_sample(name, budget_s):
    """layers (in census order, largest-first trimmed) whose single-core time fits `budget_s`"""
    from wavelettransforms_b200 import workloads
    arch, wavelet, level, pct, clamp = WORKLOADS[name]
    census = workloads.conv_census(arch)
    ws = workloads.random_init_weights(census, seed=0)
    return census, ws, (wavelet, level, pct, clamp)


def run_cpu_baseline(name, budget_s):
    """ONE core, bounded sample: walk the layers in model order until ~budget_s of CPU time is spent"""
    import oracle
    oracle.build()
    census, ws, (wavelet, level, pct, clamp) = cpu_sample(name, budget_s)
    torch.set_num_threads(1)
    # warm the library / allocator on a small layer
    _cpu_prune_layer((0, tuple(ws[0].shape), ws[0].numpy(), wavelet, level, pct, clamp))
    spent, elems, used = 0.0, 0, []
    order = sorted(range(len(ws)), key=lambda i: ws[i].numel())   # small layers first, so several plane sizes are covered
    for i in order:
        if spent >= budget_s:
            break
        _, dt, _ = _cpu_prune_layer((i, tuple(ws[i].shape), ws[i].numpy(), wavelet, level, pct, clamp))
        spent += dt
        elems += ws[i].numel()
        used.append(i)
    return {
        'value': elems / spent, 'unit': UNIT, 'cores': 1, 'kind': 'port',
        'sample': f'{len(used)} of {len(ws)} conv layers of {name} ({elems} elems, smallest first), '
                  f'{spent:.1f} s on one core; oracle C DWT + live numpy percentile/where '
                  f'(oracle/reference_flow.py)',
    }


def run_reference_arm(args):
    """--impl reference: the CPU flow on all host cores, one step = every conv layer of the workload
    (or a bounded subset when a single-core probe says the run would not end within minutes)"""
    import multiprocessing as mp
    import oracle
    oracle.build()
    name = args.workload
    census, ws, (wavelet, level, pct, clamp) = cpu_sample(name, 0)
    cores = os.cpu_count() or 1
    # probe: one mid-size layer on one core -> elems/s/core -> pick how many layers a step may hold
    probe_i = min(range(len(ws)), key=lambda i: abs(ws[i].numel() - 150000))
    _, dt, _ = _cpu_prune_layer((probe_i, tuple(ws[probe_i].shape), ws[probe_i].numpy(), wavelet, level, pct, clamp))
    _, dt, _ = _cpu_prune_layer((probe_i, tuple(ws[probe_i].shape), ws[probe_i].numpy(), wavelet, level, pct, clamp))
    rate = ws[probe_i].numel() / dt
    total_steps = args.steps + args.warmup
    budget_elems = rate * min(cores, len(ws)) * 150.0 / max(total_steps, 1)    # whole run ~<= 150 s
    order = list(range(len(ws)))
    chosen, acc = [], 0
    for i in sorted(order, key=lambda i: -ws[i].numel()):
        if acc + ws[i].numel() <= budget_elems or not chosen and ws[i].numel() <= budget_elems:
            chosen.append(i)
            acc += ws[i].numel()
    if not chosen:
        chosen = [min(order, key=lambda i: ws[i].numel())]
        acc = ws[chosen[0]].numel()
    chosen.sort()
    jobs = [(i, tuple(ws[i].shape), ws[i].numpy(), wavelet, level, pct, clamp) for i in chosen]
    jobs.sort(key=lambda j: -int(np.prod(j[1])))    # largest first: LPT over the pool
    nproc = min(cores, len(jobs))
    ctx = mp.get_context('fork')
    with ctx.Pool(nproc) as pool:
        for _ in range(args.warmup):
            pool.map(_cpu_prune_layer, jobs, chunksize=1)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            pool.map(_cpu_prune_layer, jobs, chunksize=1)
        dt = time.perf_counter() - t0
    value = acc * args.steps / dt
    sample = (f'{len(chosen)} of {len(ws)} conv layers per step ({acc} elems), process pool of {nproc} on '
              f'{cores} host cores; oracle C DWT + live numpy percentile/where (oracle/reference_flow.py); '
              f'no compiled reference exists (pure Python over PyWavelets, absent here)')
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': dt / args.steps * 1e3, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': name, 'numpy_semantics': 'live'},
        'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': nproc, 'kind': 'port', 'sample': sample},
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line), flush=True)


def verify_e2e_against_device(wl, host_out, counts):
    """the host-I/O plan and the device-I/O plan must agree bit for bit (same kernels, same inputs)"""
    stream = int(torch.cuda.current_stream().cuda_stream)
    wl.plans[0].run(stream)
    c_dev, _, _ = wl.plans[0].results(stream)
    ok = bool(np.array_equal(c_dev, counts))
    for o_dev, o_host in zip(wl.outs[0], host_out):
        ok = ok and bool(torch.equal(o_dev.cpu(), o_host))
    return ok


def run_ours(args):
    rank, world, local = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit('bench.py (impl ours) needs a CUDA device: libwtprune has no CPU fallback')
    numa_cpus = bind_to_gpu_numa_node(local) if world > 1 else 0
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        torch.distributed.init_process_group('nccl', device_id=dev)
    from wavelettransforms_b200 import _lib
    _lib.lib()   # fail loudly if the CUDA library is missing
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except Exception:
        pass
    peak_gbs = float(peaks.get('hbm_gbs', 6650.0))
    peak_src = 'measured (MEASURED_PEAKS.json hbm_gbs)' if 'hbm_gbs' in peaks else 'fallback 6650 GB/s (B200_PROFILING.md)'

    wl = DeviceWorkload(args.workload, args.replicas, dev)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms, launches = time_device_steps(wl, args.steps, args.warmup, world, dev)
    value = wl.elems * args.steps * world / (ms * 1e-3)
    phases = phase_profile(wl, args.steps)
    e2e_ms, h2d, d2h, host_in, host_out, e2e_counts = time_e2e(args.workload, args.steps, args.warmup, world, dev)
    clocks = sampler.stop() if rank == 0 else {}
    e2e_value = wl.elems * args.steps * world / (e2e_ms * 1e-3)
    consistent = verify_e2e_against_device(wl, host_out, e2e_counts)

    # dominant kernel = the launch group with the largest live average duration
    kern = [(n, t, ln) for n, t, ln in phases if n not in ('memset', 'h2d', 'd2h')]
    dom = max(kern, key=lambda x: x[1])
    dom_elems = phase_elems(wl, dom[0])
    achieved = BYTES_PER_ELEM * dom_elems / (dom[1] * 1e-3) / 1e9
    step_gbs = BYTES_PER_ELEM * wl.elems / (ms / args.steps * 1e-3) / 1e9
    roofline = {
        'bound': 'hbm', 'kernel': dom[0], 'achieved': achieved, 'peak': peak_gbs, 'unit': 'GB/s',
        'frac': achieved / peak_gbs, 'traffic': load_traffic(dom[0]),
        'peak_source': peak_src, 'algorithmic_bytes_per_launch': BYTES_PER_ELEM * dom_elems,
        'kernel_ms': dom[1], 'binding_resource': binding_resource(args.workload),
        'whole_step': {'achieved': step_gbs, 'frac': step_gbs / peak_gbs,
                       'note': '8 B x all weight elems / device ms_per_step (all launch groups)'},
        'phases_ms': {n: round(t, 5) for n, t, _ in phases},
        'timing': 'second pass over the same K steps with cudaEvents between launch groups',
    }
    variants = {}
    if not args.no_variants and world == 1:
        for vn in VARIANTS:
            if vn == args.workload:
                continue
            v = DeviceWorkload(vn, args.replicas, dev)
            vms, _ = time_device_steps(v, args.steps, args.warmup, world, dev)
            vph = phase_profile(v, min(args.steps, 50))
            vgb = BYTES_PER_ELEM * v.elems / (vms / args.steps * 1e-3) / 1e9
            variants[vn] = {'value': v.elems * args.steps / (vms * 1e-3), 'ms_per_step': vms / args.steps,
                            'hbm_frac_whole_step': vgb / peak_gbs, 'launches_per_step': v.plans[0].num_launches,
                            'phases_ms': {n: round(t, 5) for n, t, _ in vph}}
            v.close()
    line = {
        'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup,
        'ms_per_step': ms / args.steps, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
        'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': args.workload, 'model': None, 'layers': len(wl.census), 'elems_per_step_per_gpu': wl.elems,
                   'levels': sorted(set(wl.levels)), 'mode': 'periodization', 'numpy_semantics': NUMPY_SEMANTICS,
                   'l2_policy': f'{args.replicas} rotating replicas per GPU, inputs+outputs '
                                f'{args.replicas * 2 * wl.elems * 4 / 1e6:.0f} MB > 126 MB L2',
                   'parallelism': f'{world} independent replicas (no data-path collective)' if world > 1 else 'single GPU'},
        'clocks': clocks,
        'e2e': {'value': e2e_value, 'unit': UNIT, 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h,
                'ms_per_step': e2e_ms / args.steps, 'matches_device_path_bitwise': consistent,
                'host_cores_bound_to_gpu_numa_node': numa_cpus},
        'gpu_launches': launches,
        'launches_per_step': wl.plans[0].num_launches,
        'roofline': roofline,
        'variants': variants,
    }
    del line['config']['model']
    if not args.no_variants and world == 1:
        line['siblings'] = time_siblings(args.steps, dev)
    if not args.no_variants:
        line['bank'] = time_bank(args.steps, dev, peak_gbs)
        line['layer_sharded'] = time_layer_sharded(args.sharded_workload, min(args.steps, 100), min(args.warmup, 5), world, dev)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        line['cpu_baseline'] = run_cpu_baseline(args.workload, args.cpu_sample_seconds)
    wl.close()
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


def time_layer_sharded(name, steps, warmup, world, dev):
    """one model, conv layers dealt to the ranks (LPT), each rank prunes its own, ONE all-gather reassembles
    the pruned weights on every rank (north_star / SURVEY 8(e)); strong scaling, device time, max over ranks"""
    from wavelettransforms_b200 import _lib, engine, sharded, workloads
    arch, wavelet, level, pct, clamp = WORKLOADS[name]
    census = workloads.conv_census(arch)
    levels = effective_levels(census, wavelet, level, clamp)
    rank = torch.distributed.get_rank() if world > 1 else 0
    layout = sharded.ShardLayout([s for _, s in census], world)
    mine = layout.owned(rank)
    host = workloads.random_init_weights(census, seed=0)
    elems = sum(layout.numels)
    R = 4   # rotating replicas of the OWNED layers (L2 policy as in the main leg)
    flats, plans = [], []
    for _ in range(R):
        flat = torch.zeros(world * layout.shard_elems, dtype=torch.float32, device=dev)
        xs = [host[i].to(dev) for i in mine]
        items = [(x.data_ptr(), layout.view(flat, i).data_ptr()) + engine.plane_geometry(x.shape) + (levels[i],)
                 for x, i in zip(xs, mine)]
        plans.append((_lib.Plan(items, wavelet, 'periodization', _lib.THRESHOLD_PERCENTILE, pct, NUMPY_SEMANTICS), xs))
        flats.append(flat)
    stream = int(torch.cuda.current_stream().cuda_stream)

    def step(i, gather):
        r = i % R
        if mine:
            plans[r][0].run(stream)
        if gather and world > 1:
            torch.distributed.all_gather_into_tensor(flats[r], flats[r][rank * layout.shard_elems:(rank + 1) * layout.shard_elems])

    # third mode: the kernels store every pruned weight into all ranks' buffers themselves (symmetric memory over NVLink,
    # Plan.set_peers); the collective shrinks to one device-side barrier
    peer = None
    if world > 1 and not os.environ.get('WT_BENCH_NO_PEER_STORE'):
        try:
            peer = []
            for r in range(R):
                pb = sharded.PeerBuffer(layout, dev)
                items = [(x.data_ptr(), layout.view(pb.flat, i).data_ptr()) + engine.plane_geometry(x.shape) + (levels[i],)
                         for x, i in zip(plans[r][1], mine)]
                pl = _lib.Plan(items, wavelet, 'periodization', _lib.THRESHOLD_PERCENTILE, pct, NUMPY_SEMANTICS)
                pl.set_peers(pb.deltas)
                peer.append((pl, pb))
        except Exception as e:      # symmetric memory unavailable on this box, or a layer off the mirrored paths
            print(f'[bench] peer-store mode unavailable: {e}', file=sys.stderr)
            peer = None
        ok = torch.tensor([1 if peer else 0], device=dev)
        torch.distributed.all_reduce(ok, op=torch.distributed.ReduceOp.MIN)      # all ranks or none
        if not int(ok.item()):
            peer = None

    def peer_step(i, _):
        pl, pb = peer[i % R]
        if mine:
            pl.run(stream)
        pb.barrier()

    out = {}
    modes = [('prune_only', False, step), ('prune_allgather', True, step)]
    if peer:
        modes.append(('prune_peer_store', True, peer_step))
    for label, gather, step in modes:
        for i in range(warmup):
            step(i, gather)
        torch.cuda.synchronize()
        barrier(world)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            step(warmup + i, gather)
        e1.record()
        torch.cuda.synchronize()
        barrier(world)
        ms = max_over_ranks(e0.elapsed_time(e1), world, dev)
        out[label] = {'value': elems * steps / (ms * 1e-3), 'ms_per_step': ms / steps}
    if peer:
        # the peer-stored buffer of every rank must equal the all-gathered one, bit for bit
        step0 = modes[0][2]
        step0(0, True)
        peer_step(0, True)
        torch.cuda.synchronize()
        barrier(world)
        out['prune_peer_store']['matches_allgather_bitwise'] = bool(torch.equal(flats[0], peer[0][1].flat))
        for pl, _ in peer:
            pl.close()
    for p, _ in plans:
        p.close()
    out.update(workload=name, scaling='strong', layers=len(census), elems=elems, lpt_balance=layout.balance,
               allgather_bytes_per_rank=layout.shard_elems * 4 * max(world - 1, 0))
    return out


def time_siblings(steps, dev):
    """BASELINE config 3: the two sibling maskers on ResNet-18 at 50 % (device-resident, in place, every step on a
    fresh copy of the weights; per-call allocation and the host synchronisation of the entry points included)"""
    from wavelettransforms_b200 import engine, workloads
    census = workloads.conv_census('resnet18')
    host = workloads.random_init_weights(census, seed=0)
    elems = sum(w.numel() for w in host)
    n = max(4, min(steps, 32))
    out = {}
    for which in ('min_weight', 'random'):
        copies = [[w.to(dev) for w in host] for _ in range(n + 2)]
        torch.cuda.synchronize()

        def one(ws):
            if which == 'min_weight':    # min_weight_pruning.py:66-74 for every layer, one grouped call
                engine.min_weight_prune_(ws, [int(w.numel() * 0.5) for w in ws])
            else:                        # random_pruning.py:51-58 per layer (torch.randperm is the RNG)
                for w in ws:
                    flat = w.view(-1)
                    engine.scatter_zero_(flat, torch.randperm(flat.numel(), device=dev)[: flat.numel() // 2])
        one(copies[0])
        one(copies[1])
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(n):
            one(copies[2 + i])
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        out[which] = {'value': elems * n / dt, 'unit': UNIT, 'ms_per_step': dt / n * 1e3, 'steps': n,
                      'sparsity': 0.5, 'workload': 'resnet18 conv layers, 50 %'}
        del copies
    return out


def time_bank(steps, dev, peak_gbs):
    """BASELINE config 5 (a slice of the synthetic weight bank per GPU, seeds 1000+i): 32 tensors of 512x512x3x3 and
    2 matrices of 4096x4096, level-5 db4 at p = 50 through dwt_pruning's clamp (3x3 planes: level 0 = select + zero;
    4096x4096: the full five-level transform on the generic tiled path).  Device-resident, cudaEvents."""
    from wavelettransforms_b200 import _lib, engine
    g = torch.Generator(device=dev).manual_seed(1000)
    stream = int(torch.cuda.current_stream().cuda_stream)
    out = {}
    for label, shapes in (('512x512x3x3 x32 (db4 L5 clamps to level 0)', [(512, 512, 3, 3)] * 32),
                          ('4096x4096 x2 (db4 L5)', [(4096, 4096)] * 2)):
        xs = [torch.randn(sh, device=dev, generator=g) * 0.02 for sh in shapes]
        outs = [torch.empty_like(x) for x in xs]
        levels = [min(5, _lib.dwt_max_level(min(sh[-2:]), 8)) for sh in shapes]
        items = [(x.data_ptr(), o.data_ptr()) + engine.plane_geometry(x.shape) + (lv,) for x, o, lv in zip(xs, outs, levels)]
        plan = _lib.Plan(items, 'db4', 'periodization', _lib.THRESHOLD_PERCENTILE, 50.0, NUMPY_SEMANTICS)
        n = max(3, min(steps, 20))
        for _ in range(3):
            plan.run(stream)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            plan.run(stream)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        elems = sum(x.numel() for x in xs)
        gbs = BYTES_PER_ELEM * elems / (ms * 1e-3) / 1e9
        out[label] = {'value': elems / (ms * 1e-3), 'unit': UNIT, 'ms_per_step': ms, 'elems': elems, 'steps': n,
                      'launches_per_step': plan.num_launches, 'hbm_frac_whole_step': gbs / peak_gbs,
                      'l2_policy': f'inputs + outputs {2 * elems * 4 / 1e6:.0f} MB > 126 MB L2'}
        plan.close()
        del xs, outs
        torch.cuda.empty_cache()
    return out


def load_traffic(kernel):
    """dram bytes per launch of `kernel` from the committed ncu --set full summary, or None"""
    try:
        t = json.load(open(os.path.join(ROOT, 'profiles', 'traffic.json')))
        return t.get(kernel)
    except Exception:
        return None


def binding_resource(workload):
    """which roofline actually bounds the configuration"""
    _, wavelet, level, _, clamp = WORKLOADS[workload]
    if clamp and wavelet not in ('haar', 'db1', 'db2', 'sym2'):
        return 'hbm (level-0 select+zero)'
    if wavelet in ('haar', 'db1'):
        return 'hbm'
    return 'fp32_issue (faithful no-FMA arithmetic, every tap wraps on 3x3 planes)'


def main():
    args = parse_args()
    rank, world, _ = dist_env()
    if args.impl == 'reference':
        if rank == 0:
            run_reference_arm(args)
        return 0
    run_ours(args)
    return 0


if __name__ == '__main__':
    sys.exit(main())
