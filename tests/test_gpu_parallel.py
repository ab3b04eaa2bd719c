"""GPU tests of the data-parallel training step: GradSync behind the fused backward.
  * world size 1 (always runs on the 1-GPU box): the chunked backward + side-stream path leaves the same gradients;
  * world size 2 over NCCL (needs 2 GPUs, skipped otherwise): sharded batch + reduced gradients == the
    single-process gradients of the whole batch, and ActNorm data init is rank-consistent."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import ipoke_inn_b200 as P
from ipoke_inn_b200.parallel import GradSync, shard_bounds
from test_host_logic import build

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _data(B, C, E, seed=11):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(B, C, generator=g), torch.randn(B, E, 1, 1, generator=g)


def _grads(net, x, cond, B_global):
    net.zero_grad(set_to_none=True)
    z, ld = net(x, cond)
    # FlowLoss of the GLOBAL batch restricted to this shard: mean over B_global
    loss = (0.5 * (z.reshape(z.shape[0], -1) ** 2).sum() - ld.sum()) / B_global
    loss.backward()
    return [p.grad.detach().clone() for p in net.parameters()]


def test_gradsync_world1_matches_plain_backward():
    dev = torch.device("cuda:0")
    net = build("cfgS_cond").to(dev)
    x, cond = _data(64, 8, 16)
    x, cond = x.to(dev), cond.to(dev)
    with torch.no_grad():
        net(x, cond)                                   # ActNorm data init
    g0 = _grads(net, x, cond, 64)
    sync = GradSync(net.flow, chunk_blocks=3)
    g1 = _grads(net, x, cond, 64)
    sync.remove()
    assert len(sync.messages) == 3
    for a, b in zip(g0, g1):
        assert torch.allclose(a, b, rtol=0, atol=1e-6 * float(a.abs().max()) + 1e-12)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        B = 96
        net = build("cfgS_cond").to(dev)               # same seed on every rank, like the reference
        x, cond = _data(B, 8, 16)
        lo, hi = shard_bounds(B, rank, world)
        xs, cs = x[lo:hi].to(dev), cond[lo:hi].to(dev)
        with torch.no_grad():
            net(xs, cs)                                # ActNorm init from all-reduced moments
        sync = GradSync(net.flow, average=False, chunk_blocks=2)
        g = _grads(net, xs, cs, B)
        torch.cuda.synchronize()
        ret[rank] = ([t.cpu().numpy() for t in g],
                     net.flow.sub_layers[5].norm_layer.scale.detach().cpu().numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_n_ranks_equal_single_process():
    world = min(torch.cuda.device_count(), 8)           # every GPU of the box (2 on a 2-GPU lease, 8 on the scaling box)
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    # single process, whole batch
    dev = torch.device("cuda:0")
    net = build("cfgS_cond").to(dev)
    x, cond = _data(96, 8, 16)
    x, cond = x.to(dev), cond.to(dev)
    with torch.no_grad():
        net(x, cond)
    ref = [t.cpu().numpy() for t in _grads(net, x, cond, 96)]
    scale_ref = net.flow.sub_layers[5].norm_layer.scale.detach().cpu().numpy()
    for r in range(world):
        grads, scale = ret[r]
        assert np.allclose(scale, scale_ref, rtol=1e-5), "ActNorm init differs from the single-process init"
        for a, b in zip(grads, ref):
            assert np.max(np.abs(a - b)) <= 2e-4 * np.max(np.abs(b)) + 1e-9
