"""Synthetic workloads of BASELINE.json: the conv-layer census of the HF ResNets the reference prunes
(/root/reference/ResNet/utils.py:6-25 loads microsoft/resnet-18 through
AutoModelForImageClassification) and random-init weights of those shapes, without importing
transformers (tests/test_workloads.py checks the census against transformers' own modules).

Layer names and order are those of model.named_modules() filtered to nn.Conv2d, i.e. the order
/root/reference/ResNet/dwt_pruning.py:155-158 walks and the names in the stored per-layer logs
(ResNet/StoredModels/*/selective_pruned/log.csv).
"""
import math
from typing import List, Sequence, Tuple

import torch

ARCHS = {
    'resnet18': dict(layer_type='basic', hidden_sizes=[64, 128, 256, 512], depths=[2, 2, 2, 2]),
    'resnet34': dict(layer_type='basic', hidden_sizes=[64, 128, 256, 512], depths=[3, 4, 6, 3]),
    'resnet50': dict(layer_type='bottleneck', hidden_sizes=[256, 512, 1024, 2048], depths=[3, 4, 6, 3]),
    'resnet152': dict(layer_type='bottleneck', hidden_sizes=[256, 512, 1024, 2048], depths=[3, 8, 36, 3]),
}


def conv_census(arch: str = 'resnet18', embedding_size: int = 64, num_channels: int = 3) -> List[Tuple[str, Tuple[int, ...]]]:
    """[(module name, weight shape [O, I, kH, kW])] of every nn.Conv2d, named_modules() order."""
    cfg = ARCHS[arch]
    out = [('resnet.embedder.embedder.convolution', (embedding_size, num_channels, 7, 7))]
    cin = embedding_size
    for s, (cout, depth) in enumerate(zip(cfg['hidden_sizes'], cfg['depths'])):
        for l in range(depth):
            # stage 0 keeps stride 1 (downsample_in_first_stage=False); later stages stride 2 in layer 0
            stride = 2 if (l == 0 and s > 0) else 1
            pre = f'resnet.encoder.stages.{s}.layers.{l}'
            if cin != cout or stride != 1:
                out.append((f'{pre}.shortcut.convolution', (cout, cin, 1, 1)))
            if cfg['layer_type'] == 'basic':
                out.append((f'{pre}.layer.0.convolution', (cout, cin, 3, 3)))
                out.append((f'{pre}.layer.1.convolution', (cout, cout, 3, 3)))
            else:
                mid = cout // 4
                out.append((f'{pre}.layer.0.convolution', (mid, cin, 1, 1)))
                out.append((f'{pre}.layer.1.convolution', (mid, mid, 3, 3)))
                out.append((f'{pre}.layer.2.convolution', (cout, mid, 1, 1)))
            cin = cout
    return out


def random_init_weights(census: Sequence[Tuple[str, Tuple[int, ...]]], seed: int = 0, device='cpu',
                        pin_memory: bool = False) -> List[torch.Tensor]:
    """kaiming_normal_(mode='fan_out', nonlinearity='relu') -- the initialiser HF's
    ResNetPreTrainedModel._init_weights applies to nn.Conv2d -- from one seeded CPU generator, so every
    rank and both bench arms see the same values."""
    g = torch.Generator(device='cpu').manual_seed(seed)
    ws = []
    for _, shp in census:
        fan_out = shp[0] * shp[2] * shp[3]
        w = torch.randn(shp, generator=g, dtype=torch.float32) * math.sqrt(2.0 / fan_out)
        if pin_memory:
            w = w.pin_memory()
        ws.append(w.to(device) if str(device) != 'cpu' else w)
    return ws


def lpt_assignment(sizes: Sequence[int], n_ranks: int) -> List[int]:
    """Longest-processing-time greedy: item -> rank, balancing the summed sizes.
    Deterministic (ties: lowest index first, lowest rank first) so every rank computes the same map."""
    order = sorted(range(len(sizes)), key=lambda i: (-sizes[i], i))
    load = [0] * n_ranks
    owner = [0] * len(sizes)
    for i in order:
        r = min(range(n_ranks), key=lambda q: (load[q], q))
        owner[i] = r
        load[r] += sizes[i]
    return owner
