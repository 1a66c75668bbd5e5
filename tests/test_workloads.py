"""The conv-layer census bench.py and the sharding use must equal transformers' own modules."""
import pytest
import torch


@pytest.mark.parametrize('arch,n_layers,n_elems', [('resnet18', 20, 11166912), ('resnet50', 53, 23454912),
                                                   ('resnet152', 155, 57992384)])
def test_census_matches_transformers(arch, n_layers, n_elems):
    from wavelettransforms_b200 import workloads
    census = workloads.conv_census(arch)
    assert len(census) == n_layers
    assert sum(torch.Size(s).numel() for _, s in census) == n_elems
    tr = pytest.importorskip('transformers')
    c = workloads.ARCHS[arch]
    m = tr.ResNetForImageClassification(tr.ResNetConfig(layer_type=c['layer_type'], hidden_sizes=c['hidden_sizes'],
                                                        depths=c['depths']))
    ref = [(n, tuple(x.weight.shape)) for n, x in m.named_modules() if isinstance(x, torch.nn.Conv2d)]
    assert ref == census


def test_random_init_is_deterministic_and_kaiming_scaled():
    from wavelettransforms_b200 import workloads
    census = workloads.conv_census('resnet18')[:3]
    a = workloads.random_init_weights(census, seed=0)
    b = workloads.random_init_weights(census, seed=0)
    assert all(torch.equal(x, y) for x, y in zip(a, b))
    w = a[2]
    fan_out = w.shape[0] * w.shape[2] * w.shape[3]
    assert abs(float(w.std()) / (2.0 / fan_out) ** 0.5 - 1.0) < 0.05
