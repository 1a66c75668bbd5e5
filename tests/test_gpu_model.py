"""GPU tests at model level: whole ResNet-18 (random init, the BASELINE.json workload) through the
drop-in walkers, the two sibling maskers against live torch.topk / torch.randperm, and the
size-independent properties at full size."""
import csv
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def resnet18():
    from transformers import ResNetConfig, ResNetForImageClassification
    torch.manual_seed(0)
    cfg = ResNetConfig(layer_type='basic', hidden_sizes=[64, 128, 256, 512], depths=[2, 2, 2, 2])
    return ResNetForImageClassification(cfg)


def conv_layers(model):
    return [(n, m) for n, m in model.named_modules() if isinstance(m, torch.nn.Conv2d)]


@pytest.fixture(scope='module')
def mods(wtlib, oracle_lib):
    from oracle import reference_flow as rf
    from wavelettransforms_b200 import dwt_pruning, dwt_pruning_NoEntropy, engine, min_weight_pruning, random_pruning, utils
    return dict(rf=rf, dp=dwt_pruning, ne=dwt_pruning_NoEntropy, eng=engine, mw=min_weight_pruning, rp=random_pruning,
                utils=utils)


def test_resnet18_census():
    layers = conv_layers(resnet18())
    assert len(layers) == 20
    assert sum(m.weight.numel() for _, m in layers) == 11166912
    assert layers[0][0] == 'resnet.embedder.embedder.convolution'


@pytest.mark.parametrize('wavelet,level,p', [('bior4.4', 5, 50.0), ('haar', 5, 50.0), ('db2', 5, 10.0), ('bior1.3', 1, 90.0)])
def test_resnet18_wavelet_pruning_matches_reference_flow(mods, tmp_path, monkeypatch, wavelet, level, p):
    dp, rf, utils = mods['dp'], mods['rf'], mods['utils']
    monkeypatch.setattr(utils, 'OUTPUT_ROOT', str(tmp_path / 'SavedModels'))
    monkeypatch.setattr(dp, 'save_model', lambda model, d: None)
    model = resnet18()
    ref = {n: m.weight.detach().clone() for n, m in conv_layers(model)}
    log_path = dp.wavelet_pruning(model, wavelet, level, p, str(tmp_path / 'experiment_log.csv'), 'deadbeef')
    rows = list(csv.DictReader(open(log_path)))
    assert [r['Layer Name'] for r in rows] == [n for n, _ in conv_layers(model)]
    total = 0
    for (name, m), row in zip(conv_layers(model), rows):
        want, wz = rf.multi_resolution_analysis([ref[name]], wavelet, level, p, numpy_semantics='legacy_1_22')
        assert torch.equal(m.weight.data.cpu(), want[0]), name
        assert np.array_equal(m.weight.data.cpu().numpy().view(np.uint32), want[0].numpy().view(np.uint32)), name
        assert int(row['Total Pruned Count']) == wz and int(row['Original Parameter Count']) == ref[name].numel()
        assert int(row['Non-zero Params']) == ref[name].numel() - wz
        total += wz
    exp = list(csv.DictReader(open(tmp_path / 'experiment_log.csv')))
    assert int(exp[0]['Total Pruned Count']) == total and exp[0]['Phase'] == 'selective'
    if wavelet == 'bior4.4' and p == 50.0:
        # count law of the stored run bior4.4 L5 p=50 (experiment_log.csv:772): N/2 per layer, -1 where the
        # float32 threshold lands on sorted[k]
        for (name, _), row in zip(conv_layers(model), rows):
            n = ref[name].numel()
            assert int(row['Total Pruned Count']) in (n // 2, n // 2 - 1)


def test_resnet18_noentropy_walker(mods, tmp_path, monkeypatch):
    ne, rf, utils = mods['ne'], mods['rf'], mods['utils']
    monkeypatch.setattr(utils, 'OUTPUT_ROOT', str(tmp_path / 'SavedModels'))
    monkeypatch.setattr(ne, 'save_model', lambda model, d: None)
    model = resnet18()
    ref = {n: m.weight.detach().clone() for n, m in conv_layers(model)}
    log_path = ne.wavelet_pruning(model, 'bior2.2', 3, 0.02, str(tmp_path / 'e.csv'), 'cafef00d')
    rows = list(csv.DictReader(open(log_path)))
    for (name, m), row in zip(conv_layers(model), rows):
        want, wp = rf.multi_resolution_analysis_absolute([ref[name]], 'bior2.2', 3, 0.02)
        assert np.array_equal(m.weight.data.cpu().numpy().view(np.uint32), want[0].numpy().view(np.uint32)), name
        assert int(row['Total Pruned Count']) == wp
        assert int(row['Non-zero Params']) == int(torch.count_nonzero(ref[name]))


def test_min_weight_vs_torch_topk(mods):
    mw, rf = mods['mw'], mods['rf']
    g = torch.Generator().manual_seed(5)
    for shape in ((64, 3, 7, 7), (128, 128, 3, 3), (512, 256, 1, 1), (7,), (1, 1, 1, 1)):
        for pct in (0.0, 0.1, 0.4999999, 0.5, 0.9, 1.0):
            w = torch.randn(shape, generator=g) * 0.02
            a, b = w.clone(), w.clone()
            ra = mw.percentage_min_pruning(a, pct)
            rb = rf.percentage_min_pruning(b, pct)
            assert ra.data_ptr() == a.data_ptr() and ra.shape == a.shape   # in place, view of the input
            assert torch.equal(a, b), (shape, pct)
            assert int((a == 0).sum()) == int(a.numel() * pct)
    # ties: exactly k zeroed, lowest flat index first among equals
    w = torch.tensor([0.5, -0.25, 0.25, 0.25, -0.25, 1.0, 0.25, 0.0])
    out = mw.percentage_min_pruning(w.clone(), 0.5)   # k = 4: the 0.0 and the first three of |0.25|
    assert out.tolist() == [0.5, 0.0, 0.0, 0.0, -0.25, 1.0, 0.25, 0.0]
    big = (torch.randint(0, 64, (1 << 20,), generator=g).float() - 32) / 64
    for pct in (0.13, 0.5, 0.77):
        o = mw.percentage_min_pruning(big.clone().cuda(), pct)
        k = int(big.numel() * pct)
        s = torch.sort(big.abs()).values
        assert int((o == 0).sum().item()) == k      # exactly k even with heavy ties (zeros are among the k smallest)
        kept = o.cpu()[o.cpu() != 0].abs()
        assert kept.min().item() >= s[k - 1].item()


def test_random_pruning_matches_randperm(mods):
    rp, rf = mods['rp'], mods['rf']
    g = torch.Generator().manual_seed(9)
    for shape, k in (((64, 3, 7, 7), 4704), ((128, 128, 3, 3), 1000), ((16, 16, 1, 1), 0), ((16, 16, 1, 1), 256)):
        w = torch.randn(shape, generator=g)
        a, b = w.clone(), w.clone()
        torch.manual_seed(123)
        nz_a = rp.random_prune_(a, k)
        torch.manual_seed(123)
        _, nz_b = rf.random_prune_(b, k)
        assert torch.equal(a, b) and nz_a == nz_b == w.numel() - k
    # device-resident weights draw from the CUDA generator, as the reference would
    w = torch.randn(64, 64, 3, 3, generator=g).cuda()
    a, b = w.clone(), w.clone()
    torch.manual_seed(77)
    nz_a = rp.random_prune_(a, 5000)
    torch.manual_seed(77)
    _, nz_b = rf.random_prune_(b, 5000)
    assert torch.equal(a, b) and nz_a == nz_b


def test_three_walkers_end_to_end(mods, tmp_path, monkeypatch):
    """the call sequence of ResNet/main_pruning.py:185-209 against our modules"""
    import copy
    dp, mw, rp, utils = mods['dp'], mods['mw'], mods['rp'], mods['utils']
    monkeypatch.setattr(utils, 'OUTPUT_ROOT', str(tmp_path / 'SavedModels'))
    for mod in (dp, mw, rp):
        monkeypatch.setattr(mod, 'save_model', lambda model, d: None)
    model = resnet18()
    dwt_model, random_model, min_model = copy.deepcopy(model), copy.deepcopy(model), copy.deepcopy(model)
    csv_path = str(tmp_path / 'experiment_log.csv')
    log = dp.wavelet_pruning(dwt_model, 'bior4.4', 5, 0.5 * 100, csv_path, 'abcd1234')
    rp.random_pruning(random_model, log, 'abcd1234', 'bior4.4', 5, 0.5, csv_path)
    mw.min_weight_pruning(min_model, log, 'abcd1234', 'bior4.4', 5, 0.5, csv_path)
    exp = list(csv.DictReader(open(csv_path)))
    assert [r['Phase'] for r in exp] == ['selective', 'random', 'min']
    sel = list(csv.DictReader(open(log)))
    pct = sum(int(r['Total Pruned Count']) for r in sel) / sum(int(r['Original Parameter Count']) for r in sel)
    for (name, m), (_, r), row in zip(conv_layers(min_model), conv_layers(random_model), sel):
        n = m.weight.numel()
        assert int((m.weight == 0).sum()) == int(n * pct)
        assert int((r.weight == 0).sum()) == int(row['Total Pruned Count'])
    assert int(exp[1]['Total Pruned Count']) == int(exp[0]['Total Pruned Count'])


def test_full_size_properties(mods):
    """largest ResNet-18 layer (512,512,3,3): idempotence of level-0 pruning, zero-threshold round trip is
    within 1e-5, count laws"""
    dp, ne = mods['dp'], mods['ne']
    g = torch.Generator().manual_seed(2)
    w = torch.randn(512, 512, 3, 3, generator=g) * 0.02
    out, z = dp.multi_resolution_analysis([w], 'bior4.4', 5, 50.0)
    assert z in (w.numel() // 2, w.numel() // 2 - 1)
    out2, z2 = dp.multi_resolution_analysis(out, 'bior4.4', 5, float(100.0 * z / w.numel()) * 0.5)
    assert z2 == z and torch.equal(out2[0], out[0])      # thresholding below the zero mass changes nothing
    rec, pruned = ne.multi_resolution_analysis([w], 'bior4.4', 5, 0.0)
    assert pruned <= 0 or pruned < 10
    err = (rec[0] - w).abs().max().item()
    assert err <= 1e-6 + 1e-5 * w.abs().max().item(), err


@pytest.mark.parametrize('wavelet,level,p,clamp', [('bior4.4', 5, 50.0, False), ('haar', 5, 30.0, True)])
def test_synthetic_logits_and_sparsity_parity(mods, wavelet, level, p, clamp):
    """north star, third criterion: the model's outputs match the reference's to 1e-5 relative / 1e-6 absolute.
    ImageNet is not in the reference tree, so: one ResNet-18 pruned by the oracle's reference flow on
    the CPU, one by the CUDA walker, the same fixed-seed batch through both on the same device (eval mode, fp32),
    logits compared at that tolerance; calculate_sparsity (testing_suite/eval_model.py:7-20) of both agrees."""
    import copy
    from wavelettransforms_b200 import eval_model
    dp, rf = mods['dp'], mods['rf']
    base = resnet18().eval()
    ours, theirs = copy.deepcopy(base), copy.deepcopy(base)
    dp.prune_model_layers(ours, wavelet, level, p, clamp_level=clamp)
    for name, m in conv_layers(theirs):
        want, _ = rf.multi_resolution_analysis([m.weight.detach()], wavelet, level, p, numpy_semantics='legacy_1_22',
                                               clamp_level=clamp)
        m.weight.data = want[0]
    torch.manual_seed(123)
    batch = torch.randn(4, 3, 224, 224)
    prev = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        with torch.no_grad():
            lo = ours.cuda()(batch.cuda()).logits.float().cpu()
            lt = theirs.cuda()(batch.cuda()).logits.float().cpu()
            lb = base.cuda()(batch.cuda()).logits.float().cpu()
    finally:
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = prev
    err = (lo - lt).abs().max().item()
    assert torch.allclose(lo, lt, rtol=1e-5, atol=1e-6), f'max |dlogit| = {err:.3e}'
    assert not torch.allclose(lo, lb, rtol=1e-3, atol=1e-4)          # the pruning did change the function
    # reference calculate_sparsity, verbatim arithmetic, on the oracle-pruned model
    tot = nz = 0
    for prm in theirs.parameters():
        if prm.dim() > 1:
            tot += prm.numel()
            nz += torch.sum(torch.abs(prm) < 1e-6).item()
    assert eval_model.calculate_sparsity(ours) == nz / tot
    assert eval_model.calculate_sparsity(ours.cpu()) == nz / tot


def test_background_save_of_a_cuda_model_overlaps_the_next_method(tmp_path, monkeypatch):
    """save_model(async) on a model that lives on the GPU snapshots it through pinned host memory, returns,
    and the serialisation runs while the next pruning method already works on (and changes) the same weights"""
    import os

    from wavelettransforms_b200 import config, engine, utils
    dev = torch.device('cuda', 0)
    g = torch.Generator().manual_seed(5)
    m = torch.nn.Sequential(torch.nn.Conv2d(8, 16, 3, bias=False), torch.nn.Conv2d(16, 16, 1, bias=False))
    with torch.no_grad():
        for p in m.parameters():
            p.copy_(torch.randn(p.shape, generator=g) * 0.05)
    m.to(dev)
    want = {k: v.detach().cpu().clone() for k, v in m.state_dict().items()}
    monkeypatch.setattr(config, 'async_save', True)
    fut = utils.save_model(m, str(tmp_path / 'bg'))
    # the caller moves on: the next masker prunes the live weights in place while the save is (or may be) still writing
    ws = [p.data for p in m.parameters()]
    nz = engine.min_weight_prune_(ws, [w.numel() // 2 for w in ws])
    torch.cuda.synchronize()
    utils.wait_for_saves()
    assert fut.done() and fut.exception() is None
    sd = torch.load(os.path.join(tmp_path, 'bg', 'pytorch_model.bin'))
    for k in want:
        assert not sd[k].is_cuda and torch.equal(sd[k], want[k]), k      # the snapshot, not the pruned weights
    for w, n in zip(ws, nz):
        assert int(torch.count_nonzero(w)) == n == w.numel() - w.numel() // 2


def test_three_methods_on_three_streams_equal_the_three_walkers_arithmetic():
    """three_streams.prune_three_ways (one upload, DWT on the caller's stream, the two comparison maskers
    forked onto two more streams) against the same three steps done one after the other on one stream"""
    from wavelettransforms_b200 import dwt_pruning, engine, three_streams
    dev = torch.device('cuda', 0)
    g = torch.Generator().manual_seed(12)
    ws = [torch.randn(64, 3, 7, 7, generator=g) * 0.05, torch.randn(128, 64, 3, 3, generator=g) * 0.02,
          torch.randn(256, 128, 1, 1, generator=g) * 0.02, torch.randn(96, 96, 3, 3, generator=g) * 0.03]
    keep = [w.clone() for w in ws]
    torch.manual_seed(77)
    got = three_streams.prune_three_ways(ws, 'haar', 5, 60.0)
    torch.cuda.synchronize()
    for w, k in zip(ws, keep):
        assert torch.equal(w, k)                                                     # inputs untouched
    # selective: the drop-in's own function
    want, _ = dwt_pruning.multi_resolution_analysis(ws, 'haar', 5, 60.0)
    sel, sel_counts = got['selective']
    for a, b in zip(sel, want):
        assert np.array_equal(a.cpu().numpy().view(np.uint32), b.cpu().numpy().view(np.uint32))
    numel = np.array([w.numel() for w in ws])
    pruned = numel - sel_counts[:, 1]
    assert np.array_equal(sel_counts[:, 0], numel) and pruned.sum() > 0
    # min-weight: the DWT's overall fraction, grouped call, one stream
    pct = float(pruned.sum()) / float(numel.sum())
    mn = [w.to(dev).clone() for w in ws]
    nz = engine.min_weight_prune_(mn, [int(m.numel() * pct) for m in mn])
    for a, b in zip(got['min'][0], mn):
        assert torch.equal(a, b)
    assert list(got['min'][1][:, 1]) == nz
    # random: the DWT's per-layer counts, torch.randperm drawn in layer order from the same seed
    torch.manual_seed(77)
    for w, k, a, row in zip(ws, pruned, got['random'][0], got['random'][1]):
        r = w.to(dev).clone().view(-1)
        idx = torch.randperm(r.numel(), device=dev)[:int(k)]
        r[idx] = 0
        assert torch.equal(a.view(-1), r)
        assert row[1] == int(torch.count_nonzero(r)) and row[2] == row[0] - row[1]
