"""World-size-2 `gloo` tests (CPU) of the data-parallel host logic: batch sharding, parameter broadcast, and the
chunked gradient reduction GradSync performs behind the fused backward (the reference relies on Lightning's
implicit DDP, experiments/experiment.py:86).  No kernel is launched here; the GPU path of the same class is
exercised by tests/test_gpu_parallel.py."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import inn_oracle as O
from conftest import load_golden


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, fn, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ret[rank] = fn(rank, world)
    finally:
        dist.destroy_process_group()


def _run(fn, world=2):
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, fn, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0, f"worker exited with {p.exitcode}"
    return [ret[r] for r in range(world)]


def test_shard_bounds_cover_the_batch():
    from ipoke_inn_b200.parallel import shard_bounds
    for n in (0, 1, 7, 8, 100, 257):
        for world in (1, 2, 3, 8):
            spans = [shard_bounds(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


class _FakeEngine:
    """Just what GradSync reads from engine.FlowEngine."""

    def __init__(self, flow):
        self.flow = flow
        self.n_flows = len(flow.sub_layers)
        self.grad_ready_hook = None
        self.grad_chunk_blocks = 2
        offs, cur, ranges = {}, 0, []
        for blk in flow.sub_layers:
            start = cur
            for p in blk.parameters():
                offs[id(p)] = (cur, p.numel(), tuple(p.shape))
                cur += p.numel()
            ranges.append((start, cur))
        self._layout = (offs, cur, ranges, [p for blk in flow.sub_layers for p in blk.parameters()])

    def grad_layout(self):
        return self._layout


def _gradsync_job(rank, world):
    import ipoke_inn_b200 as P
    from ipoke_inn_b200.parallel import GradSync
    from test_host_logic import build
    net = build("cfgS_cond")
    flow = net.flow
    eng = _FakeEngine(flow)
    object.__setattr__(flow, "_engine", eng)
    sync = GradSync(flow, average=True, chunk_blocks=3)
    total = eng.grad_layout()[1]
    flat = torch.arange(total, dtype=torch.float32) * (rank + 1)          # rank-dependent "gradients"
    ke = eng.n_flows
    while ke > 0:                                                            # what backward.FlowFunction does
        kb = max(0, ke - eng.grad_chunk_blocks)
        eng.grad_ready_hook(eng, flat, kb, ke)
        ke = kb
    expect = torch.arange(total, dtype=torch.float32) * (sum(range(1, world + 1)) / world)
    covered = sorted(sync.messages)
    contiguous = covered[0][0] == 0 and all(a[0] + a[1] == b[0] for a, b in zip(covered, covered[1:])) \
        and covered[-1][0] + covered[-1][1] == total
    return bool(torch.allclose(flat, expect, rtol=1e-6)), contiguous, len(sync.messages)


def test_gradsync_reduces_every_block_chunk_to_the_mean():
    for ok, contiguous, n_msgs in _run(_gradsync_job):
        assert ok and contiguous
        assert n_msgs == 3          # 8 blocks in chunks of 3: [5,8) [2,5) [0,2)


def _broadcast_job(rank, world):
    from ipoke_inn_b200.parallel import broadcast_parameters
    from test_host_logic import build
    net = build("cfgS_cond", seed=100 + rank)        # deliberately different weights and permutations per rank
    broadcast_parameters(net, src=0)
    sd = net.state_dict()
    digest = float(sum(v.double().sum() for v in sd.values()))
    perm = sd["flow.sub_layers.3.shuffle.forward_shuffle_idx"].tolist()
    return digest, perm


def test_broadcast_parameters_makes_ranks_identical():
    (d0, p0), (d1, p1) = _run(_broadcast_job)
    assert d0 == d1 and p0 == p1


def _sharded_oracle_job(rank, world):
    """Sharded forward / summed gradients equal the single-process result (samples are independent)."""
    from ipoke_inn_b200.parallel import shard_bounds
    sd, d = load_golden("cfgS_cond")
    B = d["x"].shape[0]
    lo, hi = shard_bounds(B, rank, world)
    x, cond = d["x"][lo:hi], d["cond"].reshape(B, -1)[lo:hi]
    z, ld = O.flow_forward(sd, x, cond)
    dz, dld = z / B, -np.ones(hi - lo) / B                  # FlowLoss gradients with the GLOBAL batch size
    grads, _, _ = O.flow_backward(sd, x, cond, dz, dld)
    key = "flow.sub_layers.5.coupling.t.1.main.2.weight"
    g = torch.from_numpy(np.ascontiguousarray(grads[key]))
    dist.all_reduce(g)
    return z, ld, g.numpy(), (lo, hi)


def test_sharded_oracle_equals_single_process():
    sd, d = load_golden("cfgS_cond")
    outs = _run(_sharded_oracle_job)
    z = np.concatenate([o[0] for o in outs])
    ld = np.concatenate([o[1] for o in outs])
    B = d["x"].shape[0]
    assert np.allclose(z, d["z"].reshape(B, -1), atol=2e-6) and np.allclose(ld, d["logdet"], atol=2e-5)
    key = "flow.sub_layers.5.coupling.t.1.main.2.weight"
    assert np.allclose(outs[0][2], d["grad::" + key], rtol=1e-4, atol=1e-7)
    assert np.allclose(outs[0][2], outs[1][2])


def _macow_sharded(rank, world):
    """MaCow flow, batch-sharded over 2 ranks through the CPU emulation of its C ABI: rank 0 runs the data-dependent
    first call, its parameters are broadcast, every rank transforms its shard -- no collective on the data path."""
    import ctypes
    import ipoke_inn_b200 as P
    from ipoke_inn_b200 import macow
    from ipoke_inn_b200.parallel import broadcast_parameters, shard_batch
    import test_macow_flow_emulation as E
    L = macow._bind(ctypes.CDLL(E.OUT))
    macow._lib, macow._stream = (lambda: L), (lambda dev: None)
    macow._check_input = lambda x, channels: x.contiguous()
    g, sd, cfg = E.load("macow_stack_a")
    torch.manual_seed(100 + rank)                                  # ranks start with different parameters
    net = P.UnsupervisedMaCowTransformer3(cfg).eval()
    x = torch.from_numpy(np.concatenate([g["x"], g["z_in"]]))      # the same 6-sample batch on every rank
    with torch.no_grad():
        net(x if rank == 0 else x + 1.0)                           # first call: every rank initialises on ITS data and
        if rank == 0:                                              # packs its own parameters ...
            for k, p in net.state_dict().items():                  # (rank 0: non-trivial couplings)
                if k.endswith("weight_g"):
                    p.fill_(0.2)
        broadcast_parameters(net)                                  # ... which rank 0's state replaces, incl. the
                                                                   # `initialized` flags and permutations; stale packs dropped
        mine = shard_batch(x, rank=rank, world=world)
        z, ld = net(mine)
        whole = net(x) if rank == 0 else None
    return (z.numpy(), ld.numpy(), None if whole is None else (whole[0].numpy(), whole[1].numpy()),
            {k: v.numpy() for k, v in net.state_dict().items() if k.endswith(("log_scale", "shuffle_idx"))})


def test_macow_flow_shards_without_a_collective():
    import test_macow_flow_emulation as E
    import subprocess
    os.makedirs(os.path.dirname(E.OUT), exist_ok=True)
    if not os.path.exists(E.OUT) or os.path.getmtime(E.OUT) < max(os.path.getmtime(f) for f in E.DEPS):
        subprocess.run(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-Wno-unknown-pragmas", "-Wno-attributes",
                        "-I", os.path.join(ROOT, "include"), "-o", E.OUT, E.SRC], check=True)
    (z0, l0, whole, s0), (z1, l1, _, s1) = _run(_macow_sharded)
    for k in s0:
        assert np.array_equal(s0[k], s1[k]), k                     # rank 1 took rank 0's initialisation and permutations
    np.testing.assert_allclose(np.concatenate([z0, z1]), whole[0], rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(np.concatenate([l0, l1]), whole[1], rtol=1e-6, atol=1e-5)
