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

# Workload names: <arch>_<wavelet>_L<level>_p<percentile>_<clamped|unclamped>
#   unclamped = the decomposition dwt_pruning_NoEntropy.py:41 runs (3x3 planes: 3->2->1->1->1->1) with the percentile
#               threshold of dwt_pruning.py:25-32: the real transform on every layer (the north-star target);
#   clamped   = the same call through dwt_pruning.py's clamp (:64-65): every wavelet with dec_len >= 6 gets level 0 on
#               ResNet planes => select + zero, no transform -- what the reference's stored runs executed.
import re


class Workloads(dict):
    """name -> (arch, wavelet, level, percentile, clamp_level); any well-formed name is accepted"""
    PAT = re.compile(r'^(resnet18|resnet50|resnet152)_([a-z0-9.]+)_L(\d+)_p([0-9.]+)_(clamped|unclamped)$')

    def __missing__(self, name):
        m = self.PAT.match(name)
        if not m:
            raise KeyError(name)
        v = (m.group(1), m.group(2), int(m.group(3)), float(m.group(4)), m.group(5) == 'clamped')
        self[name] = v
        return v

    def __contains__(self, name):
        return bool(self.PAT.match(str(name)))


WORKLOADS = Workloads()
# Headline: all ResNet-18 conv layers through a level-5 bior4.4 decomposition at the 90th percentile.  (Round 1 quoted
# p = 50, where the threshold of this decomposition is 3.5e-19 -- 24 of the 49 packed cells of a 3x3 plane are padding
# zeros and the level-3..5 details are rounding noise -- and NOTHING is pruned; at p = 90 the threshold is ~0.02, 49 % of
# the weights go, and the select works on long candidate lists.  p = 50 stays as a variant.)
DEFAULT_WORKLOAD = 'resnet18_bior4.4_L5_p90_unclamped'
VARIANTS = ['resnet18_bior4.4_L5_p90_clamped',      # what dwt_pruning.py itself executes for this call: level 0
            'resnet18_haar_L5_p90_clamped',         # deepest real transform the clamped reference runs: 3x3 -> L1, 7x7 -> L2
            'resnet18_bior2.2_L3_p90_unclamped',
            'resnet18_bior4.4_L5_p50_unclamped',    # round 1's headline (prunes nothing)
            'resnet18_bior1.3_L1_p50_clamped']      # BASELINE.json configs[0]: the reference's own CPU-runnable case
NUMPY_SEMANTICS = 'legacy_1_22'  # the reference's pin (requirements.txt:2)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=200)
    ap.add_argument('--warmup', type=int, default=10)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--workload', default=DEFAULT_WORKLOAD)
    ap.add_argument('--replicas', type=int, default=8, help='rotating model replicas per GPU (defeats L2 reuse)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-variants', action='store_true')
    ap.add_argument('--sharded-workload', default='resnet152_bior2.2_L3_p90_unclamped')
    ap.add_argument('--no-sweep', action='store_true')
    ap.add_argument('--cpu-sample-seconds', type=float, default=15.0)
    return ap.parse_args()


def base_config(name, replicas=8):
    """the `config` object of BOTH arms (ours and --impl reference): identical keys and values, so the driver can see that
    the two lines describe the same workload"""
    from wavelettransforms_b200 import workloads
    arch, wavelet, level, pct, clamp = WORKLOADS[name]
    census = workloads.conv_census(arch)
    elems = sum(int(np.prod(s)) for _, s in census)
    levels = host_effective_levels(census, wavelet, level, clamp)
    return {'workload': name, 'layers': len(census), 'elems_per_step_per_gpu': elems, 'levels': sorted(set(levels)),
            'mode': 'periodization', 'numpy_semantics': NUMPY_SEMANTICS, 'arithmetic': 'faithful (fp32 multiply then add, no FMA)',
            'l2_policy': f'GPU arm: {replicas} rotating replicas of the model per GPU, inputs+outputs '
                         f'{replicas * 2 * elems * 4 / 1e6:.0f} MB > 126 MB L2, so no step finds its input in L2; the CPU arm has none'}


# dec_len of the 17 flag wavelets: the clamp needs no GPU library
DEC_LEN = {'haar': 2, 'db1': 2, 'db2': 4, 'sym2': 4, 'db4': 8, 'sym4': 8, 'db6': 12, 'sym6': 12, 'coif1': 6, 'coif2': 12, 'coif3': 18,
           'bior1.3': 6, 'rbio1.3': 6, 'bior2.2': 6, 'rbio2.2': 6, 'bior4.4': 10, 'rbio4.4': 10}


def host_effective_levels(census, wavelet, level, clamp):
    """effective_levels without the CUDA library (pywt.dwt_max_level)"""
    F = DEC_LEN[wavelet]
    out = []
    for _, shp in census:
        n = min(shp[-2:])
        mx = 0 if (F <= 1 or n < F - 1) else int(np.floor(np.log2(n // (F - 1))))
        out.append(min(level, mx) if clamp else level)
    return out


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

    def __init__(self, name, replicas, dev, arithmetic='faithful'):
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
            if arithmetic != 'faithful':
                self.plans[-1].set_arithmetic(arithmetic)
            self.ins.append(xs)
            self.outs.append(os_)
        torch.cuda.synchronize()

    def step(self, i, stream):
        self.plans[i % len(self.plans)].run(stream)

    def pruned_fraction(self):
        """zeros in the output of one step / elements (a headline that prunes nothing is not a headline)"""
        stream = int(torch.cuda.current_stream().cuda_stream)
        self.plans[0].run(stream)
        counts, _, _ = self.plans[0].results(stream)
        return float(counts[:, 2].sum()) / float(self.elems)

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


def phase_bytes(wl, phase):
    """bytes the launch group `phase` has to move for its own job in this pipeline (what it reads once + what it writes
    once): the denominator of its own roofline fraction.  (The whole path's algorithmic bytes are 8 B per weight.)"""
    from wavelettransforms_b200 import _lib
    def per_layer(sel):
        w = c = 0
        for (_, s), lv in zip(wl.census, wl.levels):
            if not sel(s, lv):
                continue
            n = int(np.prod(s))
            w += n
            if lv > 0:
                b = int(np.prod(s[:-2]))
                c += b * _lib.packed_shape(s[-2], s[-1], wl.wavelet, lv)[2]
            else:
                c += n
        return w, c
    if phase.startswith('fused_'):
        w, _ = per_layer(lambda s, lv: (lv == 0) if phase == 'fused_flat' else (lv > 0 and f'{s[-2]}x{s[-1]}' == phase.split('_')[1]))
        return 8 * w
    if phase.startswith('small_'):
        hw = phase.rsplit('_', 1)[1]
        w, c = per_layer(lambda s, lv: lv > 0 and f'{s[-2]}x{s[-1]}' == hw)
        return 4 * w + 4 * c                  # analysis: weights in, coefficients out; synthesis: the reverse
    if phase.startswith('select'):
        w0, _ = per_layer(lambda s, lv: lv == 0)
        _, c1 = per_layer(lambda s, lv: lv > 0)
        return 4 * c1 + 8 * w0                # one read of the coefficients; level-0 layers: read + pruned write
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
    out, zeros = rf.multi_resolution_analysis([w], wavelet, level, pct, numpy_semantics=NUMPY_SEMANTICS, clamp_level=clamp)
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
                  f'{spent:.1f} s on one core; oracle C DWT + numpy partition / where, numpy-1.22 percentile arithmetic '
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
              f'{cores} host cores; oracle C DWT + numpy partition / where with the numpy-1.22 percentile arithmetic of the reference pin (oracle/reference_flow.py); '
              f'no compiled reference exists (pure Python over PyWavelets, absent here)')
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': dt / args.steps * 1e3, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': base_config(name, args.replicas),
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

    # roofline: `frac` is the WHOLE STEP on the path's algorithmic bytes (8 B per weight: read it once, write the pruned
    # weight once); the launch groups are listed with their own bytes; the dominant one = largest live average duration
    kern = [(n, t, ln) for n, t, ln in phases if n not in ('memset', 'h2d', 'd2h')]
    dom = max(kern, key=lambda x: x[1])
    step_ms = ms / args.steps
    step_gbs = BYTES_PER_ELEM * wl.elems / (step_ms * 1e-3) / 1e9
    kernels = {}
    for n, t, ln in kern:
        b = phase_bytes(wl, n)
        kernels[n] = {'ms': round(t, 5), 'launches': ln, 'own_bytes': b, 'own_gbs': round(b / (t * 1e-3) / 1e9, 1) if t > 0 else None,
                      'own_frac': round(b / (t * 1e-3) / 1e9 / peak_gbs, 4) if t > 0 else None,
                      'share_of_step': round(t / sum(x[1] for x in phases), 4)}
    roofline = {
        'bound': 'hbm', 'achieved': step_gbs, 'peak': peak_gbs, 'unit': 'GB/s', 'frac': step_gbs / peak_gbs,
        'what': 'whole step: 8 B x all weight elements / device ms_per_step (every launch group of the plan)',
        'algorithmic_bytes_per_step': BYTES_PER_ELEM * wl.elems,
        'traffic': load_traffic('step:' + args.workload), 'peak_source': peak_src,
        'binding_resource': binding_resource(args.workload),
        'dominant_kernel': dict(kernels[dom[0]], name=dom[0], traffic=load_traffic(dom[0])),
        'kernels': kernels,
        'timing': 'second pass over the same K steps with cudaEvents between launch groups (groups run serially there; the '
                  'timed steps fork independent groups onto side streams)',
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
                            'select_path': v.plans[0].select_path, 'pruned_fraction': v.pruned_fraction(),
                            'binding_resource': binding_resource(vn),
                            'phases_ms': {n: round(t, 5) for n, t, _ in vph}}
            v.close()
    # the named fast mode beside the faithful arithmetic (never mixed into `value`): same workload, every DWT multiply-add fused
    fast = None
    if not args.no_variants and world == 1 and max(wl.levels) > 0:
        fw = DeviceWorkload(args.workload, args.replicas, dev, arithmetic='fma_taporder')
        fms, _ = time_device_steps(fw, args.steps, args.warmup, world, dev)
        fph = phase_profile(fw, min(args.steps, 50))
        fast = {'arithmetic': 'fma_taporder (same tap order, FFMA; oracle twin orc_set_arithmetic(1); masks may differ from the reference)',
                'value': fw.elems * args.steps / (fms * 1e-3), 'ms_per_step': fms / args.steps,
                'speedup_vs_faithful_step': (ms / args.steps) / (fms / args.steps),
                'hbm_frac_whole_step': BYTES_PER_ELEM * fw.elems / (fms / args.steps * 1e-3) / 1e9 / peak_gbs,
                'pruned_fraction': fw.pruned_fraction(), 'phases_ms': {n: round(t, 5) for n, t, _ in fph}}
        fw.close()
    line = {
        'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup,
        'ms_per_step': ms / args.steps, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
        'dtype': 'f32', 'data': 'synthetic',
        'config': base_config(args.workload, args.replicas),
        'parallelism': f'{world} independent replicas (no data-path collective)' if world > 1 else 'single GPU',
        'select_path': wl.plans[0].select_path, 'pruned_fraction': wl.pruned_fraction(),
        'clocks': clocks,
        'e2e': {'value': e2e_value, 'unit': UNIT, 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h,
                'ms_per_step': e2e_ms / args.steps, 'matches_device_path_bitwise': consistent,
                'host_cores_bound_to_gpu_numa_node': numa_cpus},
        'gpu_launches': launches,
        'launches_per_step': wl.plans[0].num_launches,
        'roofline': roofline,
        'fast_mode': fast,
        'variants': variants,
    }
    if not args.no_variants and world == 1:
        line['siblings'] = time_siblings(args.steps, dev)
        line['noentropy_absolute'] = time_noentropy(dev, peak_gbs, args.steps)
        line['dropin_call'] = time_dropin_call()
        if not args.no_sweep:
            line['sweep'] = time_sweep(dev, peak_gbs, args.steps)
    if not args.no_variants:
        line['bank'] = time_bank(args.steps, dev, peak_gbs, world)
        line['layer_sharded'] = {}
        for swl in dict.fromkeys([args.sharded_workload, 'resnet50_bior2.2_L3_p90_unclamped', 'resnet18_bior2.2_L3_p90_unclamped']):
            line['layer_sharded'][swl] = time_layer_sharded(swl, min(args.steps, 50), min(args.warmup, 5), world, dev)
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
                pm = None
                if pb.multicast_delta is not None:     # NVLS: the same plan storing once through the multicast mapping
                    pm = _lib.Plan(items, wavelet, 'periodization', _lib.THRESHOLD_PERCENTILE, pct, NUMPY_SEMANTICS)
                    pm.set_multicast(pb.multicast_delta)
                peer.append((pl, pb, pm))
        except Exception as e:      # symmetric memory unavailable on this box, or a layer off the mirrored paths
            print(f'[bench] peer-store mode unavailable: {e}', file=sys.stderr)
            peer = None
        ok = torch.tensor([1 if peer else 0], device=dev)
        torch.distributed.all_reduce(ok, op=torch.distributed.ReduceOp.MIN)      # all ranks or none
        if not int(ok.item()):
            peer = None

    def peer_step(i, _):
        pl, pb, _pm = peer[i % R]
        if mine:
            pl.run(stream)
        pb.barrier()

    def mc_step(i, _):
        _pl, pb, pm = peer[i % R]
        if mine:
            pm.run(stream)
        pb.barrier()

    # fourth mode: the gather overlapped with the pruning -- layers in G size-ordered groups, each with its own LPT layout,
    # flat buffer and plan; group g's all-gather runs on a side stream while group g + 1 is pruned (sharded.prune_sharded_pipelined)
    pipe = None
    if world > 1:
        G = 4
        groups = sharded.layer_groups(layout.numels, G)
        side = torch.cuda.Stream(device=dev)
        pipe = []
        for _ in range(R):
            rep = []
            for idx in groups:
                lay = sharded.ShardLayout([census[i][1] for i in idx], world)
                fl = torch.zeros(world * lay.shard_elems, dtype=torch.float32, device=dev)
                own = lay.owned(rank)
                xs = [host[idx[k]].to(dev) for k in own]
                items = [(x.data_ptr(), lay.view(fl, k).data_ptr()) + engine.plane_geometry(x.shape) + (levels[idx[k]],)
                         for x, k in zip(xs, own)]
                pl = _lib.Plan(items, wavelet, 'periodization', _lib.THRESHOLD_PERCENTILE, pct, NUMPY_SEMANTICS) if own else None
                rep.append((pl, xs, fl, fl[rank * lay.shard_elems:(rank + 1) * lay.shard_elems]))
            pipe.append(rep)

    def pipe_step(i, _):
        main = torch.cuda.current_stream()
        for pl, _, fl, shard in pipe[i % R]:
            if pl is not None:
                pl.run(stream)
            side.wait_stream(main)
            with torch.cuda.stream(side):
                torch.distributed.all_gather_into_tensor(fl, shard)
        main.wait_stream(side)

    out = {}
    # the unsharded model on ONE GPU, timed on every rank in the same run: the numerator of the strong-scaling efficiency
    single_ms = None
    if world > 1:
        full = []
        for _ in range(2):
            xs = [w.to(dev) for w in host]
            os_ = [torch.empty_like(x) for x in xs]
            items = [(x.data_ptr(), o.data_ptr()) + engine.plane_geometry(x.shape) + (lv,) for x, o, lv in zip(xs, os_, levels)]
            full.append((_lib.Plan(items, wavelet, 'periodization', _lib.THRESHOLD_PERCENTILE, pct, NUMPY_SEMANTICS), xs, os_))
        single_ms = max_over_ranks(_time_plans([f[0] for f in full], steps, warmup), world, dev)
        for f in full:
            f[0].close()
        del full
    modes = [('prune_only', False, step), ('prune_allgather', True, step)]
    if pipe:
        modes.append(('prune_allgather_pipelined', True, pipe_step))
    if peer:
        modes.append(('prune_peer_store', True, peer_step))
        has_mc = torch.tensor([1 if all(p[2] is not None for p in peer) else 0], device=dev)
        torch.distributed.all_reduce(has_mc, op=torch.distributed.ReduceOp.MIN)
        if int(has_mc.item()):
            modes.append(('prune_multicast_store', True, mc_step))
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
        if single_ms:
            out[label]['efficiency_vs_1gpu'] = round(single_ms / (world * ms / steps), 4)
    if single_ms:
        out['single_gpu_ms_per_step'] = single_ms
    if peer:
        # the peer-stored buffer of every rank must equal the all-gathered one, bit for bit
        step0 = modes[0][2]
        step0(0, True)
        peer_step(0, True)
        torch.cuda.synchronize()
        barrier(world)
        out['prune_peer_store']['matches_allgather_bitwise'] = bool(torch.equal(flats[0], peer[0][1].flat))
        if 'prune_multicast_store' in out:
            peer[0][1].flat.zero_()
            torch.cuda.synchronize()
            barrier(world)
            mc_step(0, True)
            torch.cuda.synchronize()
            barrier(world)
            out['prune_multicast_store']['matches_allgather_bitwise'] = bool(torch.equal(flats[0], peer[0][1].flat))
            out['prune_multicast_store']['how'] = 'one multimem.st per pruned weight to the NVSwitch multicast mapping (egress = own shard)'
        for pl, _, pm in peer:
            pl.close()
            if pm is not None:
                pm.close()
    if pipe:
        # the pipelined buffers against the single all-gather, layer by layer
        modes[0][2](0, True)
        pipe_step(0, True)
        torch.cuda.synchronize()
        barrier(world)
        ok, gi = True, 0
        for idx, (pl, _, fl, _) in zip(sharded.layer_groups(layout.numels, 4), pipe[0]):
            lay = sharded.ShardLayout([census[i][1] for i in idx], world)
            for k, i in enumerate(idx):
                ok = ok and bool(torch.equal(lay.view(fl, k), layout.view(flats[0], i)))
        out['prune_allgather_pipelined']['matches_allgather_bitwise'] = ok
        out['prune_allgather_pipelined']['groups'] = 4
        for rep in pipe:
            for pl, _, _, _ in rep:
                if pl is not None:
                    pl.close()
    for p, _ in plans:
        p.close()
    out.update(workload=name, scaling='strong', n_gpus=world, layers=len(census), elems=elems, lpt_balance=layout.balance,
               allgather_bytes_per_rank=layout.shard_elems * 4 * max(world - 1, 0))
    return out


def _time_plans(plans, steps, warmup):
    stream = int(torch.cuda.current_stream().cuda_stream)
    for i in range(warmup):
        plans[i % len(plans)].run(stream)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        plans[(warmup + i) % len(plans)].run(stream)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def time_sweep(dev, peak_gbs, steps):
    """BASELINE.json configs[1]: ResNet-18, {bior1.3, 2.2, 4.4, rbio1.3, 2.2, 4.4} x levels 1..5 x p in {10, 30, 50, 70, 90},
    unclamped (the real transform) and clamped (dwt_pruning.py:64-65: dec_len >= 6 => level 0 for EVERY point of this grid,
    so the clamped leg is wavelet- and level-independent and is run once per p).  Device-resident, 3 rotating replicas
    (inputs + outputs 268 MB > L2), CUDA events."""
    from wavelettransforms_b200 import _lib, engine, workloads
    census = workloads.conv_census('resnet18')
    host = workloads.random_init_weights(census, seed=0)
    elems = sum(w.numel() for w in host)
    R = 3
    ins = [[w.to(dev) for w in host] for _ in range(R)]
    outs = [[torch.empty_like(x) for x in xs] for xs in ins]
    n = max(4, min(steps, 10))
    stream = int(torch.cuda.current_stream().cuda_stream)

    def point(wavelet, level, pct, clamp):
        levels = host_effective_levels(census, wavelet, level, clamp)
        plans = [_lib.Plan([(x.data_ptr(), o.data_ptr()) + engine.plane_geometry(x.shape) + (lv,) for x, o, lv in zip(xs, os_, levels)],
                           wavelet, 'periodization', _lib.THRESHOLD_PERCENTILE, float(pct), NUMPY_SEMANTICS) for xs, os_ in zip(ins, outs)]
        ms = _time_plans(plans, n, 2)
        plans[0].run(stream)
        counts, _, _ = plans[0].results(stream)
        path = plans[0].select_path
        for p in plans:
            p.close()
        return {'wavelet': wavelet, 'level': level, 'p': pct, 'ms': round(ms, 5), 'gelems_s': round(elems / (ms * 1e-3) / 1e9, 2),
                'hbm_frac': round(BYTES_PER_ELEM * elems / (ms * 1e-3) / 1e9 / peak_gbs, 4),
                'pruned_fraction': round(float(counts[:, 2].sum()) / elems, 5), 'select_path': path}

    def summary(points):
        v = sorted(points, key=lambda q: q['gelems_s'])
        med = v[len(v) // 2]
        tag = lambda q: f"{q['wavelet']} L{q['level']} p{q['p']}"
        return {'n': len(v), 'min': {'config': tag(v[0]), 'gelems_s': v[0]['gelems_s'], 'hbm_frac': v[0]['hbm_frac']},
                'median': {'config': tag(med), 'gelems_s': med['gelems_s'], 'hbm_frac': med['hbm_frac']},
                'max': {'config': tag(v[-1]), 'gelems_s': v[-1]['gelems_s'], 'hbm_frac': v[-1]['hbm_frac']},
                'configs_that_prune_nothing': sum(1 for q in v if q['pruned_fraction'] == 0.0),
                'select_paths': {k: sum(1 for q in v if q['select_path'] == k) for k in sorted({str(q['select_path']) for q in v})}}

    ps = [10, 30, 50, 70, 90]
    unclamped = [point(w, L, p, False) for w in ('bior1.3', 'bior2.2', 'bior4.4', 'rbio1.3', 'rbio2.2', 'rbio4.4')
                 for L in (1, 2, 3, 4, 5) for p in ps]
    clamped = [point('bior4.4', 5, p, True) for p in ps]
    return {'workload': 'resnet18 conv layers', 'elems': elems, 'unit': 'G elems/s', 'steps_per_point': n,
            'unclamped': dict(summary(unclamped), points=unclamped),
            'clamped_level0': dict(summary(clamped), points=clamped,
                                   note='every (wavelet, level) of the grid clamps to level 0 on ResNet planes: one run per p')}


def time_noentropy(dev, peak_gbs, steps):
    """dwt_pruning_NoEntropy.py's flow (absolute threshold, no clamp) on ResNet-18: the fused recompute kernels (WT_FUSED=all:
    one pass per plane class) against the round-1 kernels (WT_FUSED=off) and the library's default routing (fused for
    level-0 layers only)"""
    from wavelettransforms_b200 import _lib, engine, workloads
    census = workloads.conv_census('resnet18')
    host = workloads.random_init_weights(census, seed=0)
    elems = sum(w.numel() for w in host)
    R = 3
    ins = [[w.to(dev) for w in host] for _ in range(R)]
    outs = [[torch.empty_like(x) for x in xs] for xs in ins]
    out = {}
    prev = os.environ.get('WT_FUSED')
    for label, wavelet, level in (('level0 (any wavelet, L0)', 'bior4.4', 0), ('haar L1', 'haar', 1), ('db2 L2', 'db2', 2), ('bior4.4 L5', 'bior4.4', 5)):
        row = {}
        for mode in ('default', 'all', 'off'):
            if mode == 'default':
                os.environ.pop('WT_FUSED', None)
            else:
                os.environ['WT_FUSED'] = mode
            plans = [_lib.Plan([(x.data_ptr(), o.data_ptr()) + engine.plane_geometry(x.shape) + (level,) for x, o in zip(xs, os_)],
                               wavelet, 'periodization', _lib.THRESHOLD_ABSOLUTE, 0.02, NUMPY_SEMANTICS) for xs, os_ in zip(ins, outs)]
            ms = _time_plans(plans, max(4, min(steps, 20)), 3)
            row[{'default': 'default_routing', 'all': 'fused_kernels', 'off': 'round1_kernels'}[mode]] = {
                'ms_per_step': round(ms, 5), 'launches_per_step': plans[0].num_launches,
                'hbm_frac_whole_step': round(BYTES_PER_ELEM * elems / (ms * 1e-3) / 1e9 / peak_gbs, 4)}
            for p in plans:
                p.close()
        out[label] = row
    if prev is None:
        os.environ.pop('WT_FUSED', None)
    else:
        os.environ['WT_FUSED'] = prev
    return out


def time_dropin_call():
    """the call the reference's driver makes (main_pruning.py:185-186): dwt_pruning.wavelet_pruning(model_on_cpu, ...) --
    pageable host weights, per-call plan, results back on the model's device, per-layer CSV, save_pretrained.  Wall clock."""
    import shutil
    from transformers import ResNetConfig, ResNetForImageClassification
    from wavelettransforms_b200 import dwt_pruning, utils
    arch, wavelet, level, pct, clamp = WORKLOADS[DEFAULT_WORKLOAD]
    tmp = tempfile.mkdtemp(prefix='wt_dropin_')
    old_root, old_save = utils.OUTPUT_ROOT, dwt_pruning.save_model
    out = {}
    try:
        utils.OUTPUT_ROOT = os.path.join(tmp, 'SavedModels')
        torch.manual_seed(0)
        base = ResNetForImageClassification(ResNetConfig(layer_type='basic', hidden_sizes=[64, 128, 256, 512], depths=[2, 2, 2, 2]))
        import copy
        for label, save in (('save_excluded', False), ('save_included', True)):
            dwt_pruning.save_model = old_save if save else (lambda model, d: None)
            ts = []
            for rep in range(3):
                m = copy.deepcopy(base)
                t0 = time.perf_counter()
                dwt_pruning.wavelet_pruning(m, wavelet, level, pct, os.path.join(tmp, 'experiment_log.csv'), f'{rep:08x}')
                ts.append(time.perf_counter() - t0)
            out[label] = {'ms': round(min(ts[1:]) * 1e3, 2), 'first_call_ms': round(ts[0] * 1e3, 2)}
        out['call'] = f"dwt_pruning.wavelet_pruning(cpu_model, '{wavelet}', {level}, {pct}, csv, guid) -- the clamped flow of the driver"
        out['elems'] = 11166912
    finally:
        utils.OUTPUT_ROOT, dwt_pruning.save_model = old_root, old_save
        shutil.rmtree(tmp, ignore_errors=True)
    return out


def time_siblings(steps, dev):
    """BASELINE config 3: the two sibling maskers on ResNet-18 at 10..90 % (device-resident, in place, every step on a
    fresh copy of the weights; per-call allocation and the host synchronisation of the entry points included)"""
    from wavelettransforms_b200 import engine, workloads
    census = workloads.conv_census('resnet18')
    host = workloads.random_init_weights(census, seed=0)
    elems = sum(w.numel() for w in host)
    n = max(4, min(steps, 16))
    out = {}
    for which in ('min_weight', 'random'):
        rows = {}
        for sparsity in (0.1, 0.3, 0.5, 0.7, 0.9):
            copies = [[w.to(dev) for w in host] for _ in range(n + 2)]
            torch.cuda.synchronize()

            def one(ws):
                if which == 'min_weight':    # min_weight_pruning.py:66-74 for every layer, one grouped call
                    engine.min_weight_prune_(ws, [int(w.numel() * sparsity) for w in ws])
                else:                        # random_pruning.py:51-58: torch.randperm per layer (the RNG of record), grouped scatter + count
                    engine.scatter_zero_group_([w.view(-1) for w in ws],
                                               [torch.randperm(w.numel(), device=dev)[: int(w.numel() * sparsity)] for w in ws])
            one(copies[0])
            one(copies[1])
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for i in range(n):
                one(copies[2 + i])
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            rows[f'{int(sparsity * 100)}%'] = {'value': elems * n / dt, 'ms_per_step': dt / n * 1e3}
            del copies
        out[which] = dict(rows['50%'], unit=UNIT, steps=n, sparsity=0.5, workload='resnet18 conv layers, 50 %', sweep=rows)
    # what torch.randperm alone costs (it stays the RNG so masks match the reference's bit for bit)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(4):
        for w in host:
            torch.randperm(w.numel(), device=dev)
    torch.cuda.synchronize()
    out['random']['randperm_only_ms'] = (time.perf_counter() - t0) / 4 * 1e3
    return out


def time_bank(steps, dev, peak_gbs, world=1):
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
        ms_local = e0.elapsed_time(e1) / n
        barrier(world)
        ms = max_over_ranks(ms_local, world, dev)      # every rank prunes its own slice of the bank: no collective, weak scaling
        elems = sum(x.numel() for x in xs)
        gbs = BYTES_PER_ELEM * elems / (ms * 1e-3) / 1e9
        out[label] = {'value': elems * world / (ms * 1e-3), 'unit': UNIT, 'ms_per_step': ms, 'elems': elems * world,
                      'elems_per_gpu': elems, 'n_gpus': world, 'scaling': 'weak', 'steps': n, 'binding_resource':
                      'hbm (level-0 select+zero)' if len(shapes[0]) == 4 else 'fp32_issue + shared-memory latency (db4 round trip, no FMA)',
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
