"""GPU tests of the sharded optimizer step (optim_sharded.ShardedAdam, packed-master training).
  * one GPU: fed the same gradients, ShardedAdam equals torch.optim.Adam step for step, and the bf16 hi / lo weights it
    writes into the packed buffer are bit-identical to a plain repack of the fp32 parameters (same forward outputs);
  * N ranks over NCCL (all the GPUs of the box, skipped on one GPU): reduce-scatter -> shard update -> all-gather leaves
    every rank with the same packed weights, equal to the single-process whole-batch training, in both wire formats."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import ipoke_inn_b200 as P
from ipoke_inn_b200.parallel import shard_bounds
from test_host_logic import CFG, build

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HYPER = dict(lr=1e-3, betas=(0.9, 0.999), weight_decay=1e-5, amsgrad=True)


def _data(name, B, seed=11):
    kind, kw = CFG[name]
    C, E = kw["flow_in_channels"], kw.get("flow_embedding_channels", 0)
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, C, generator=g)
    cond = torch.randn(B, E, 1, 1, generator=g) if kind == "cond" else None
    return x, cond


def _call(net, x, cond):
    return net(x, cond) if cond is not None else net(x)


@pytest.mark.parametrize("name", ["cfgA_mini", "cfgB_mini"])
@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_sharded_adam_equals_torch_adam_on_the_same_gradients(name, precision):
    dev = torch.device("cuda:0")
    net = build(name, precision=precision).to(dev)
    ref = build(name, precision=precision).to(dev)
    x, cond = _data(name, 96)
    x, cond = x.to(dev), (cond.to(dev) if cond is not None else None)
    with torch.no_grad():
        _call(net, x, cond)                                    # ActNorm data init
    ref.load_state_dict(net.state_dict())
    keys = list(net.state_dict().keys())
    opt = P.ShardedAdam(net, **HYPER)
    assert list(net.state_dict().keys()) == keys               # adoption keeps names / shapes
    topt = torch.optim.Adam(ref.parameters(), **HYPER)
    lf = P.FlowLoss()
    for step in range(4):
        opt.zero_grad(set_to_none=True)
        z, ld = _call(net, x, cond)
        loss, _ = lf(z, ld)
        loss.backward()
        for p, q in zip(net.parameters(), ref.parameters()):
            q.grad = p.grad.detach().clone()
        opt.param_groups[0]["lr"] = topt.param_groups[0]["lr"] = 1e-3 * (step + 1) / 4     # warm-up ramp
        opt.step()
        topt.step()
        for (k, p), q in zip(net.named_parameters(), ref.parameters()):
            d = float((p.detach() - q.detach()).abs().max())
            assert d <= 1e-6 + 1e-5 * float(q.detach().abs().max()), (k, step, d)
    # the packed bf16 images written by the optimizer == a plain repack of the same fp32 values
    fresh = build(name, precision=precision).to(dev)
    fresh.load_state_dict(net.state_dict())
    with torch.no_grad():
        za, la = _call(net, x, cond)
        zb, lb = _call(fresh, x, cond)
    # (logdet sums partial row sums with fp32 atomics: equal up to the summation order)
    assert torch.equal(za, zb) and torch.allclose(la, lb, rtol=1e-6, atol=1e-5)
    # out-of-band edit of the master (load_state_dict) re-splits it: outputs follow
    net.load_state_dict(ref.state_dict())
    with torch.no_grad():
        zc, _ = _call(net, x, cond)
        zd, _ = _call(ref, x, cond)
    assert torch.equal(zc, zd)


def test_sharded_adam_refuses_padded_shapes_and_cpu_modules():
    with pytest.raises(RuntimeError):
        P.ShardedAdam(build("cfgA_mini"))                      # still on the CPU
    with pytest.raises(NotImplementedError):
        P.ShardedAdam(build("cfgS_cond").to("cuda:0"))         # E = 16, C/2 = 4 are padded inside the packed arena


def test_module_survives_deepcopy_and_pickle_after_cuda_calls():
    import copy
    import io
    dev = torch.device("cuda:0")
    net = build("cfgA_mini").to(dev)
    x, cond = _data("cfgA_mini", 32)
    x, cond = x.to(dev), cond.to(dev)
    with torch.no_grad():
        z0, _ = net(x, cond)
        twin = copy.deepcopy(net)
        z1, _ = twin(x, cond)
        buf = io.BytesIO()
        torch.save(net, buf)
        buf.seek(0)
        back = torch.load(buf, weights_only=False)
        z2, _ = back(x, cond)
    assert torch.equal(z0, z1) and torch.equal(z0, z2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _train(net, opt, x, cond, B_global, steps):
    for _ in range(steps):
        opt.zero_grad(set_to_none=True)
        z, ld = _call(net, x, cond)
        # FlowLoss of the GLOBAL batch restricted to this shard, times the world size: the mean over ranks (DDP
        # semantics of the exchange) is then the gradient of the whole-batch loss
        loss = (0.5 * (z.reshape(z.shape[0], -1) ** 2).sum() - ld.sum()) / x.shape[0]
        loss.backward()
        opt.step()


def _worker(rank, world, port, wire, name, B, steps, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        net = build(name, precision="fp32").to(dev)    # same seed on every rank, like the reference
        x, cond = _data(name, B)
        lo, hi = shard_bounds(B, rank, world)
        xs, cs = x[lo:hi].to(dev), (cond[lo:hi].to(dev) if cond is not None else None)
        with torch.no_grad():
            _call(net, xs, cs)                         # ActNorm init from all-reduced moments
        opt = P.ShardedAdam(net, wire=wire, chunk_blocks=1, eps=1e-3, **HYPER)
        assert all(c["sharded"] for c in opt.pieces)
        _train(net, opt, xs, cs, B, steps)
        packed = net.flow.engine._packed.clone()
        sd = net.state_dict()                          # gathers the fp32 master
        torch.cuda.synchronize()
        ret[rank] = ({k: v.detach().cpu().numpy() for k, v in sd.items() if v.dtype == torch.float32},
                     packed.cpu().numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("wire", ["fp32", "bf16"])
def test_n_ranks_equal_single_process(wire):
    world = min(torch.cuda.device_count(), 8)
    name, B, steps = "cfgA_mini", 64 * world, 3
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, wire, name, B, steps, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(600)
        assert p.exitcode == 0
    # single process, whole batch, replicated fused Adam
    dev = torch.device("cuda:0")
    net = build(name, precision="fp32").to(dev)
    x, cond = _data(name, B)
    x, cond = x.to(dev), cond.to(dev)
    with torch.no_grad():
        net(x, cond)
    before = {k: v.detach().cpu().numpy().copy() for k, v in net.state_dict().items() if v.dtype == torch.float32}
    opt = P.FusedAdam(net.parameters(), eps=1e-3, **HYPER)
    _train(net, opt, x, cond, B, steps)
    ref = {k: v.detach().cpu().numpy() for k, v in net.state_dict().items() if v.dtype == torch.float32}
    sd0, packed0 = ret[0]
    for r in range(1, world):
        sdr, packedr = ret[r]
        # the small arena holds this process's own pointers (job tables): compare the weight arenas
        n16 = net.flow.engine.adopt_packed_master().n * 2 * 2
        assert np.array_equal(packedr[:n16], packed0[:n16]), "ranks ended with different packed weights"
        for k in sd0:
            assert np.array_equal(sdr[k], sd0[k]), (k, r)
    # bf16 on the wire rounds every RANK's gradient to 3 significant digits before the sum; where the ranks' gradients
    # cancel, the error relative to the (small) mean is larger than 2^-9: measured 2.1 % of the largest update
    tol = 2e-4 if wire == "fp32" else 5e-2
    for k in ref:
        upd = np.max(np.abs(ref[k] - before[k]))
        err = np.max(np.abs(sd0[k] - ref[k]))
        assert err <= tol * max(upd, 1e-12) + 1e-7, (k, err, upd)
