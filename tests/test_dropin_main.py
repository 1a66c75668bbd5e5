"""The reference's own driver against the drop-in modules (ResNet/main_pruning.py:52-55 flat imports, :166-213 one
main-thread call + two concurrent threading.Threads + a log-queue thread).

CPU: the UNMODIFIED /root/reference/ResNet/main_pruning.py is imported in a subprocess with
wavelettransforms_b200/dropin first on sys.path: its flat imports must resolve to this package (skipped where the
reference tree is absent, e.g. on the GPU box).  GPU: the body of its main() is replayed statement by statement with
real threads on a small random-init HF ResNet, and the three per-layer logs are held against the oracle flow."""
import csv
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DROPIN = os.path.join(ROOT, 'wavelettransforms_b200', 'dropin')
REF_MAIN = '/root/reference/ResNet/main_pruning.py'


@pytest.mark.skipif(not os.path.exists(REF_MAIN), reason='reference tree not on this machine')
def test_unmodified_reference_driver_imports_the_dropin_modules():
    pytest.importorskip('absl')
    code = f'''
import importlib.util, sys
sys.path.insert(0, {DROPIN!r}); sys.path.insert(1, {ROOT!r})
spec = importlib.util.spec_from_file_location("ref_main_pruning", {REF_MAIN!r})
m = importlib.util.module_from_spec(spec); spec.loader.exec_module(m)
for fn in (m.wavelet_pruning, m.random_pruning, m.min_weight_pruning, m.load_model, m.append_to_experiment_log):
    assert fn.__module__.startswith("wavelettransforms_b200."), (fn, fn.__module__)
assert callable(m.main) and callable(m.threaded_pruning) and callable(m.log_worker)
names = sorted(f for f in m.FLAGS if f in ("model_path", "config_path", "csv_path", "wavelet", "level", "threshold", "output_dir"))
print("OK", names)
'''
    r = subprocess.run([sys.executable, '-c', code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert 'OK' in r.stdout and 'wavelet' in r.stdout


def _small_resnet(tmp_path):
    from transformers import ResNetConfig, ResNetForImageClassification
    torch.manual_seed(0)
    cfg = ResNetConfig(layer_type='basic', embedding_size=32, hidden_sizes=[32, 64, 128], depths=[1, 2, 1], num_labels=10)
    d = str(tmp_path / 'model')
    ResNetForImageClassification(cfg).save_pretrained(d)
    return d


@pytest.mark.gpu
@pytest.mark.parametrize('wavelet,level,threshold', [('bior4.4', 5, 0.5), ('haar', 5, 0.3)])
def test_main_sequence_with_threads(tmp_path, monkeypatch, oracle_lib, wavelet, level, threshold):
    """main_pruning.py:166-213 with the drop-in modules imported the way the driver imports them"""
    import copy
    import queue
    import threading
    monkeypatch.syspath_prepend(DROPIN)
    for name in ('utils', 'dwt_pruning', 'min_weight_pruning', 'random_pruning'):
        sys.modules.pop(name, None)
    from utils import load_model, append_to_experiment_log                # noqa: E402  (main_pruning.py:53)
    from random_pruning import random_pruning                             # :54
    from dwt_pruning import wavelet_pruning                               # :55
    from min_weight_pruning import min_weight_pruning                     # :56
    import wavelettransforms_b200.utils as wt_utils
    from oracle import reference_flow as rf
    monkeypatch.setattr(wt_utils, 'OUTPUT_ROOT', str(tmp_path / 'SavedModels'))
    model_path = _small_resnet(tmp_path)
    csv_path = str(tmp_path / 'experiment_log.csv')
    log_queue = queue.Queue()
    errors = []

    def log_worker(path):                                                 # :108-115
        while True:
            rec = log_queue.get()
            if rec is None:
                break
            append_to_experiment_log(path, *rec)

    def threaded_pruning(fn, model, log_path, guid, wv, lv, thr, path, method, q):   # :118-127
        try:
            fn(model, log_path, guid, wv, lv, thr, path, q)
        except Exception as e:  # the reference prints and swallows; the test wants to know
            errors.append((method, repr(e)))

    model = load_model(model_path, os.path.join(model_path, 'config.json'))
    ref = {n: m.weight.detach().clone() for n, m in model.named_modules() if isinstance(m, torch.nn.Conv2d)}
    guid = os.urandom(4).hex()
    dwt_model, random_model, min_weight_model = (copy.deepcopy(model) for _ in range(3))
    log_thread = threading.Thread(target=log_worker, args=(csv_path,))
    log_thread.start()
    selective_log_path = wavelet_pruning(dwt_model, wavelet, level, threshold * 100, csv_path, guid)
    t1 = threading.Thread(target=threaded_pruning, args=(random_pruning, random_model, selective_log_path, guid, wavelet, level,
                                                         threshold, csv_path, 'Random', log_queue))
    t2 = threading.Thread(target=threaded_pruning, args=(min_weight_pruning, min_weight_model, selective_log_path, guid, wavelet,
                                                         level, threshold, csv_path, 'Minimum Weight', log_queue))
    t1.start(); t2.start(); t1.join(); t2.join()
    log_queue.put(None)
    log_thread.join()
    assert not errors, errors

    # the selective log against the reference flow (oracle DWT + numpy percentile), layer by layer
    rows = list(csv.DictReader(open(selective_log_path)))
    assert [r['Layer Name'] for r in rows] == list(ref)
    total_pruned = total = 0
    for r, (name, w) in zip(rows, ref.items()):
        want, wz = rf.multi_resolution_analysis([w], wavelet, level, threshold * 100, numpy_semantics='legacy_1_22')
        got = dict(dwt_model.named_modules())[name].weight.data.cpu()
        assert np.array_equal(got.numpy().view(np.uint32), want[0].numpy().view(np.uint32)), name
        assert int(r['Total Pruned Count']) == wz and int(r['Original Parameter Count']) == w.numel()
        total_pruned += wz
        total += w.numel()
    base = os.path.dirname(os.path.dirname(selective_log_path))
    rnd = list(csv.DictReader(open(os.path.join(base, 'random_pruned', 'log.csv'))))
    mnw = list(csv.DictReader(open(os.path.join(base, 'min_pruned', 'log.csv'))))
    pct = total_pruned / total                                            # min_weight_pruning.py:54-63
    for r, a, b, (name, w) in zip(rows, rnd, mnw, ref.items()):
        k = int(r['Total Pruned Count'])
        assert a['Layer Name'] == b['Layer Name'] == name
        zeros_before = int((w == 0).sum())
        assert k <= int(a['Total Pruned Count']) <= k + zeros_before      # random_pruning.py:45-58: k positions zeroed
        assert int(b['Total Pruned Count']) == int(w.numel() * pct)       # min_weight_pruning.py:70 (tie-free random init)
    phases = [r['Phase'] for r in csv.DictReader(open(csv_path))]
    assert phases[0] == 'selective' and len(phases) == 3 and len(set(phases)) == 3     # one experiment row per method
    for sub in ('selective_pruned', 'random_pruned', 'min_pruned'):
        assert os.path.exists(os.path.join(base, sub, 'config.json'))


@pytest.mark.gpu
def test_percentile_based_thresholding_and_prune_layer_weights(oracle_lib):
    """rows a3 / a4: ResNet/dwt_pruning.py:25-32 and :98-127 called directly"""
    from oracle import reference_flow as rf
    from wavelettransforms_b200 import dwt_pruning
    rng = np.random.default_rng(5)
    arr = (rng.standard_normal((37, 19)) * 0.1).astype(np.float32)
    arr[3, :5] = 0.0
    for p in (0, 10, 50, 90, 100):
        got = dwt_pruning.percentile_based_thresholding(arr, p)
        t = np.percentile(np.abs(arr), p)
        want = np.where(np.abs(arr) < t, 0, arr)
        assert isinstance(got, np.ndarray) and got.shape == arr.shape and got.dtype == arr.dtype
        assert (got != want).sum() <= 1          # numpy 2.x here vs the pinned 1.22 semantics: at most the rank-k element
    torch.manual_seed(1)
    layer = torch.nn.Conv2d(16, 24, 3, bias=False)
    w0 = layer.weight.detach().clone()
    storage_before = layer.weight.data.data_ptr()
    numel, nonzero, zeros = dwt_pruning.prune_layer_weights(layer, 'haar', 5, 50.0)
    want, wz = rf.multi_resolution_analysis([w0], 'haar', 5, 50.0, numpy_semantics='legacy_1_22')
    assert (numel, nonzero, zeros) == (w0.numel(), w0.numel() - wz, wz)
    assert np.array_equal(layer.weight.data.cpu().numpy().view(np.uint32), want[0].numpy().view(np.uint32))
    assert layer.weight.data.data_ptr() != storage_before            # `.data` is REBOUND (:125), not copied into
    assert layer.weight.device == w0.device and layer.weight.dtype == torch.float32
