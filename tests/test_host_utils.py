"""Host-side helpers of the drop-in modules (no GPU): the two CSV wire formats of /root/reference/ResNet/utils.py
(:55-58 per-layer, :127-128 experiment) that min_weight_pruning / random_pruning read back by column name, and the
opt-in background save."""
import csv
import os

import torch

from wavelettransforms_b200 import config, utils


def test_csv_wire_formats(tmp_path):
    w, f = utils.setup_csv_writer(str(tmp_path / 'a' / 'log.csv'), mode='w')
    utils.log_pruning_details(w, 'g', 'haar', 1, 0.5, 'selective', 100, 60, 40, 'resnet.embedder.embedder.convolution')
    utils.log_pruning_details(w, 'g', 'haar', 1, 0.5, 'selective', 1, 1, 0, '')      # no layer name: dropped (utils.py:85)
    f.close()
    rows = list(csv.DictReader(open(tmp_path / 'a' / 'log.csv')))
    assert list(rows[0].keys()) == utils.PER_LAYER_FIELDS and len(rows) == 1
    assert rows[0]['Total Pruned Count'] == '40' and rows[0]['Original Parameter Count'] == '100'
    p = str(tmp_path / 'experiment_log.csv')
    utils.append_to_experiment_log(p, 'g', 'haar', 1, 0.5, 'selective', 40, 60, 'x/y')
    utils.append_to_experiment_log(p, 'g', 'haar', 1, 0.5, 'random', 40, 60, 'x/z')
    rows = list(csv.DictReader(open(p)))
    assert list(rows[0].keys()) == utils.EXPERIMENT_FIELDS and [r['Phase'] for r in rows] == ['selective', 'random']


def test_get_layer_dotted_lookup():
    m = torch.nn.Sequential(torch.nn.Conv2d(1, 2, 3), torch.nn.Sequential(torch.nn.Conv2d(2, 2, 1)))
    assert utils.get_layer(m, '1.0') is m[1][0]
    assert utils.get_layer(m, '1.7') is None and utils.get_layer(m, 'nope') is None


def test_background_save_snapshots_then_writes(tmp_path, monkeypatch):
    m = torch.nn.Conv2d(2, 3, 3, bias=False)
    want = m.weight.detach().clone()
    monkeypatch.setattr(config, 'async_save', True)
    fut = utils.save_model(m, str(tmp_path / 'bg'))
    with torch.no_grad():
        m.weight.zero_()                      # the caller moves on and mutates the model: the snapshot must not see it
    utils.wait_for_saves()
    assert fut.done() and fut.exception() is None
    sd = torch.load(os.path.join(tmp_path, 'bg', 'pytorch_model.bin'))
    assert torch.equal(sd['weight'], want)
    monkeypatch.setattr(config, 'async_save', False)
    assert utils.save_model(m, str(tmp_path / 'fg')) is None
    assert torch.equal(torch.load(os.path.join(tmp_path, 'fg', 'pytorch_model.bin'))['weight'], torch.zeros_like(want))


def test_get_layer_prefix_and_print_helpers(capsys):
    """ResNet/utils.py:220-223 (class-name prefix), :165-206 (print helpers)"""
    m = torch.nn.Sequential(torch.nn.Conv2d(1, 2, 3), torch.nn.Sequential(torch.nn.Conv2d(2, 2, 1)))
    assert utils.get_layer(m, 'ResNetForImageClassification.1.0') is m[1][0]
    assert utils.get_layer(m, '1..0') is m[1][0]
    utils.print_model_summary(m)
    utils.print_model_structure(m)
    out = capsys.readouterr().out
    assert 'Total Params: 20' in out and '  0 - Conv2d' in out


def test_load_model_reads_config_path_and_rejects_files(tmp_path):
    """ResNet/utils.py:17-22"""
    import pytest
    from transformers import ResNetConfig, ResNetForImageClassification
    cfg = ResNetConfig(layer_type='basic', embedding_size=8, hidden_sizes=[8, 16], depths=[1, 1], num_labels=3)
    ResNetForImageClassification(cfg).save_pretrained(str(tmp_path / 'm'))
    model = utils.load_model(str(tmp_path / 'm'), str(tmp_path / 'm' / 'config.json'))
    assert model.config.hidden_sizes == [8, 16]
    assert utils.load_model(str(tmp_path / 'm')).config.num_labels == 3
    with pytest.raises(ValueError, match='not a valid directory'):
        utils.load_model(str(tmp_path / 'm' / 'config.json'), None)
