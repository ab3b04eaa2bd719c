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
        assert all(p.grad is None for p in net.parameters())       # the optimizer owns the one flat gradient buffer
        for g, q in zip(opt.gradients(), ref.parameters()):
            q.grad = g.detach().clone()
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


def _train(net, opt, x, cond, steps):
    for _ in range(steps):
        opt.zero_grad(set_to_none=True)
        z, ld = _call(net, x, cond)
        # FlowLoss restricted to this process's rows: with equal shards the mean over ranks (DDP semantics of the
        # exchange) is the gradient of the whole-batch loss
        loss = (0.5 * (z.reshape(z.shape[0], -1) ** 2).sum() - ld.sum()) / x.shape[0]
        loss.backward()
        opt.step()


def _kink_clear_batch(net, name, B, dev, shard=None):
    """ActNorm-initialise `net` on a first batch, then pick B samples whose hidden pre-activations all stay clear of
    LeakyReLU's kink under the initialised weights (LeakyReLU' is discontinuous: different tile shapes at different
    batch sizes would otherwise flip single units between the sharded and the whole-batch run)."""
    from oracle import inn_oracle as O
    x0, c0 = _data(name, B, seed=3)
    lo, hi = shard if shard is not None else (0, B)
    with torch.no_grad():
        _call(net, x0[lo:hi].to(dev), c0[lo:hi].to(dev) if c0 is not None else None)
    n = 2 * B
    x, cond = _data(name, n, seed=11)
    sd = O.state_dict_to_numpy(net.state_dict())
    d = O.kink_distance(sd, x.numpy(), None if cond is None else cond.numpy().reshape(n, -1))
    keep = torch.from_numpy(np.nonzero(d > 1e-4)[0][:B])       # 4096 hidden units: 72 % of the samples are clear
    assert len(keep) == B
    return x[keep].contiguous(), (cond[keep].contiguous() if cond is not None else None)


def _worker(rank, world, port, wire, name, B, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        net = build(name, precision="fp32").to(dev)    # same seed on every rank, like the reference
        lo, hi = shard_bounds(B, rank, world)
        x, cond = _kink_clear_batch(net, name, B, dev, (lo, hi))      # ActNorm init from all-reduced moments
        xs, cs = x[lo:hi].to(dev), (cond[lo:hi].to(dev) if cond is not None else None)
        opt = P.ShardedAdam(net, wire=wire, chunk_blocks=1, eps=1e-3, **HYPER)
        assert all(c["sharded"] for c in opt.pieces)
        _train(net, opt, xs, cs, 1)
        sd1 = {k: v.detach().cpu().numpy().copy() for k, v in net.state_dict().items() if v.dtype == torch.float32}
        _train(net, opt, xs, cs, 2)
        packed = net.flow.engine._packed.clone()
        sd3 = net.state_dict()                         # gathers the fp32 master
        torch.cuda.synchronize()
        ret[rank] = (sd1, {k: v.detach().cpu().numpy() for k, v in sd3.items() if v.dtype == torch.float32},
                     packed.cpu().numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
@pytest.mark.parametrize("wire", ["fp32", "bf16"])
def test_n_ranks_equal_single_process(wire):
    world = min(torch.cuda.device_count(), 8)
    name, B = "cfgA_mini", 64 * world
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, wire, name, B, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(600)
        assert p.exitcode == 0
    # single process, whole batch, replicated fused Adam
    dev = torch.device("cuda:0")
    net = build(name, precision="fp32").to(dev)
    x, cond = _kink_clear_batch(net, name, B, dev)
    x, cond = x.to(dev), cond.to(dev)

    def snap():
        return {k: v.detach().cpu().numpy().copy() for k, v in net.state_dict().items() if v.dtype == torch.float32}
    before = snap()
    opt = P.FusedAdam(net.parameters(), eps=1e-3, **HYPER)
    _train(net, opt, x, cond, 1)
    ref1 = snap()
    _train(net, opt, x, cond, 2)
    ref3 = snap()
    sd1_0, sd3_0, packed0 = ret[0]
    n16 = net.flow.engine.adopt_packed_master().n * 2 * 2       # bytes of the hi + lo weight arenas
    for r in range(1, world):
        _, sd3_r, packed_r = ret[r]
        # (the small arena holds each process's own pointers in its job tables: compare the weight arenas)
        assert np.array_equal(packed_r[:n16], packed0[:n16]), "ranks ended with different packed weights"
        for k in sd3_0:
            assert np.array_equal(sd3_r[k], sd3_0[k]), (k, r)
    # step 1 on kink-clear samples.  fp32 wire: strict, per tensor.  bf16 wire: every RANK's gradient is rounded to 3
    # significant digits before the sum, so where the ranks' gradients cancel the error relative to the (small) mean
    # exceeds 2^-9 (measured up to 8 % of a tensor's largest update): the update as a whole must agree to 1 %
    if wire == "fp32":
        for k in ref1:
            upd = np.max(np.abs(ref1[k] - before[k]))
            err = np.max(np.abs(sd1_0[k] - ref1[k]))
            assert err <= 2e-4 * max(upd, 1e-12) + 1e-7, (k, err, upd)
    num = sum(float(((sd1_0[k].astype(np.float64) - ref1[k]) ** 2).sum()) for k in ref1)
    den = sum(float(((ref1[k].astype(np.float64) - before[k]) ** 2).sum()) for k in ref1)
    assert np.sqrt(num / den) < (1e-4 if wire == "fp32" else 1e-2), np.sqrt(num / den)
    # steps 2-3 (samples no longer kink-filtered for the updated weights): the whole update, globally
    num = sum(float(((sd3_0[k].astype(np.float64) - ref3[k]) ** 2).sum()) for k in ref3)
    den = sum(float(((ref3[k].astype(np.float64) - before[k]) ** 2).sum()) for k in ref3)
    assert np.sqrt(num / den) < (2e-2 if wire == "fp32" else 8e-2), np.sqrt(num / den)


def test_sharded_adam_with_direct_param_grads_and_bf16_wire_on_one_gpu():
    """engine.direct_param_grads (gradients handed to p.grad inside backward) and wire="bf16" (no exchange at N = 1) leave
    the step unchanged."""
    dev = torch.device("cuda:0")
    x, cond = _data("cfgA_mini", 64)
    x, cond = x.to(dev), cond.to(dev)
    res = []
    for direct, wire in ((False, "fp32"), (True, "bf16")):
        net = build("cfgA_mini", precision="bf16").to(dev)
        with torch.no_grad():
            net(x, cond)
        net.flow.engine.direct_param_grads = direct
        opt = P.ShardedAdam(net, wire=wire, **HYPER)
        _train(net, opt, x, cond, 3)
        res.append({k: v.detach().cpu().numpy() for k, v in net.state_dict().items() if v.dtype == torch.float32})
    for k in res[0]:
        d = np.max(np.abs(res[0][k] - res[1][k]))
        assert d <= 1e-6 + 1e-5 * np.max(np.abs(res[0][k])), (k, d)
