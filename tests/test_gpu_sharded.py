"""GPU twin of tests/test_sharded.py: the layer-sharded path with libwtprune's kernels storing straight into
the all-gather send buffer.  One GPU => world_size 1 over NCCL (the collective code path still runs); with
N GPUs the same test can be launched under torchrun."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def pg():
    created = False
    if not dist.is_initialized():
        s = socket.socket()
        s.bind(('127.0.0.1', 0))
        port = s.getsockname()[1]
        s.close()
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        os.environ.setdefault('MASTER_PORT', str(port))
        torch.cuda.set_device(int(os.environ.get('LOCAL_RANK', '0')))
        dist.init_process_group('nccl', rank=int(os.environ.get('RANK', '0')), world_size=int(os.environ.get('WORLD_SIZE', '1')))
        created = True
    yield dist.group.WORLD
    if created:
        dist.destroy_process_group()


@pytest.mark.parametrize('wavelet,level,clamp', [('bior4.4', 5, True), ('haar', 5, True), ('bior2.2', 3, False)])
def test_sharded_gpu_matches_reference_flow(wtlib, oracle_lib, pg, wavelet, level, clamp):
    from oracle import reference_flow as rf
    from wavelettransforms_b200 import sharded, workloads
    census = workloads.conv_census('resnet18')[:9]
    host = workloads.random_init_weights(census, seed=3)
    dev = torch.device('cuda', torch.cuda.current_device())
    ws = [w.to(dev) for w in host]
    pruned, counts, flat, layout = sharded.prune_model_sharded(ws, wavelet, level, 50.0, clamp_level=clamp,
                                                               numpy_semantics='legacy_1_22', group=pg)
    torch.cuda.synchronize()
    for i, w in enumerate(host):
        want, zeros = rf.multi_resolution_analysis([w], wavelet, level, 50.0, numpy_semantics='legacy_1_22', clamp_level=clamp)
        assert np.array_equal(pruned[i].cpu().numpy().view(np.uint32), want[0].numpy().view(np.uint32)), census[i][0]
        assert counts[i, 0] == w.numel() and counts[i, 2] == zeros and counts[i, 1] == w.numel() - zeros
        assert counts[i, 3] == int(torch.count_nonzero(w))
    # inputs untouched (multi_resolution_analysis never modifies its input, dwt_pruning.py:84-85)
    for w, x in zip(host, ws):
        assert torch.equal(w, x.cpu())


def test_sharded_gpu_peer_store_equals_allgather(wtlib, oracle_lib, pg):
    """gather='peer_store': the synthesis kernels and the select's apply phase store every pruned weight into all
    ranks' symmetric-memory buffers themselves; the result must be the all-gathered buffer bit for bit, on every rank.
    Needs world_size > 1: `torchrun --nproc-per-node 2 -m pytest tests/test_gpu_sharded.py -m gpu`."""
    if dist.get_world_size(pg) < 2:
        pytest.skip('one rank: there is no peer to store into')
    from oracle import reference_flow as rf
    from wavelettransforms_b200 import sharded, workloads
    census = workloads.conv_census('resnet18')[:12]
    host = workloads.random_init_weights(census, seed=5)
    dev = torch.device('cuda', torch.cuda.current_device())
    ws = [w.to(dev) for w in host]
    for wavelet, level, clamp in (('bior4.4', 5, False), ('bior4.4', 5, True), ('haar', 5, True)):
        a, ca, flat_a, _ = sharded.prune_model_sharded(ws, wavelet, level, 50.0, clamp_level=clamp,
                                                       numpy_semantics='legacy_1_22', group=pg)
        b, cb, flat_b, _ = sharded.prune_model_sharded(ws, wavelet, level, 50.0, clamp_level=clamp,
                                                       numpy_semantics='legacy_1_22', group=pg, gather='peer_store')
        c, cc, flat_c, _ = sharded.prune_model_sharded(ws, wavelet, level, 50.0, clamp_level=clamp,
                                                       numpy_semantics='legacy_1_22', group=pg, gather='multicast')
        torch.cuda.synchronize()
        assert torch.equal(flat_a, flat_b) and np.array_equal(ca, cb)
        assert torch.equal(flat_a, flat_c) and np.array_equal(ca, cc)      # NVLS multicast stores (peer stores without NVLS)
        for i, w in enumerate(host):
            want, _ = rf.multi_resolution_analysis([w], wavelet, level, 50.0, numpy_semantics='legacy_1_22', clamp_level=clamp)
            assert np.array_equal(b[i].cpu().numpy().view(np.uint32), want[0].numpy().view(np.uint32)), census[i][0]


@pytest.mark.parametrize('wavelet,level,shape,pct', [('bior2.2', 2, (37, 8, 3, 3), 70.0), ('haar', 1, (64, 16, 3, 3), 50.0),
                                                     ('db4', 3, (5, 200, 264), 83.0), ('bior4.4', 0, (300, 7, 1, 1), 35.0),
                                                     ('db2', 2, (1, 130, 130), 61.8), ('haar', 0, (513,), 90.0)])
def test_intra_layer_split_equals_the_single_gpu_plan(pg, wavelet, level, shape, pct):
    """one tensor's planes over the ranks, threshold agreed by the distributed radix select (wt_key_hist_f32 +
    all-reduce): the gathered result, the threshold and the zero count equal the single-GPU percentile plan, bit for bit.
    World size 1 under plain pytest; `torchrun --nproc-per-node N -m pytest tests/test_gpu_sharded.py` runs it on N GPUs."""
    from wavelettransforms_b200 import _lib, dwt_pruning, engine, sharded
    g = torch.Generator().manual_seed(21)
    w = torch.randn(shape, generator=g) * 0.05
    w.view(-1)[::11] = 0.0
    got, t, zeros = sharded.prune_tensor_split(w, wavelet, level, pct)
    want, wz = dwt_pruning.multi_resolution_analysis([w], wavelet, level, pct, clamp_level=False)
    assert zeros == wz
    assert np.array_equal(got.cpu().numpy().view(np.uint32), want[0].cpu().numpy().view(np.uint32))
    dev = torch.device('cuda', torch.cuda.current_device())
    ref = engine.prune_device_tensors([w.to(dev)], [level if w.ndim >= 2 else 0], wavelet, 'periodization', _lib.THRESHOLD_PERCENTILE, pct)
    assert np.float32(t).view(np.uint32) == np.float32(ref.thresholds[0]).view(np.uint32)
    # without the gather every rank holds its own planes
    part, _, _ = sharded.prune_tensor_split(w, wavelet, level, pct, gather=False)
    b, H, W = engine.plane_geometry(w.shape)
    lo, hi = sharded.plane_slice(b, dist.get_rank(), dist.get_world_size())
    assert np.array_equal(part.cpu().numpy().view(np.uint32), want[0].reshape(b, H, W)[lo:hi].cpu().numpy().view(np.uint32))
