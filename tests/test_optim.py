"""Optimizer step of the training path: oracle vs torch.optim.Adam golden vectors on CPU;
the fused CUDA step (through the C ABI) vs the same vectors on the GPU."""
import os

import numpy as np
import pytest
import torch

from oracle import adam_oracle as AO

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "adam_golden.npz")
HYPER = {"ams": dict(lr=1e-3, betas=(0.9, 0.999), weight_decay=1e-5, amsgrad=True),
         "plain": dict(lr=3e-4, betas=(0.8, 0.99), weight_decay=0.0, amsgrad=False, eps=1e-6)}
# fp32 rounding of one fused expression vs torch's op-by-op kernels: a few ulp of the update, relative to |p| ~ 1
TOL = 2e-6


def _close(a, b, tol=TOL):
    return np.max(np.abs(a - b)) <= tol * max(1.0, float(np.max(np.abs(b))))


@pytest.mark.parametrize("tag", ["ams", "plain"])
def test_oracle_matches_torch_adam_golden(tag):
    G = np.load(GOLD)
    kw = HYPER[tag]
    for i in range(int(G["n_tensors"])):
        p = G[f"{tag}_p0_{i}"].copy()
        m, v, vm = np.zeros_like(p), np.zeros_like(p), np.zeros_like(p)
        for step in range(4):
            AO.adam_step(p, G[f"{tag}_g{step}_{i}"], m, v, vm, step + 1, lr=kw["lr"], beta1=kw["betas"][0],
                         beta2=kw["betas"][1], eps=kw.get("eps", 1e-8), weight_decay=kw["weight_decay"],
                         amsgrad=kw["amsgrad"])
            assert _close(p, G[f"{tag}_p{step + 1}_{i}"]), (tag, i, step)
        assert _close(m, G[f"{tag}_m_{i}"]) and _close(v, G[f"{tag}_v_{i}"], 1e-6)
        if kw["amsgrad"]:
            assert _close(vm, G[f"{tag}_vmax_{i}"], 1e-6)


def test_linear_var_matches_reference():
    import ipoke_inn_b200 as P
    G = np.load(GOLD)
    for it, w, d in zip(G["lv_its"], G["lv_warm"], G["lv_decay"]):
        for f in (AO.linear_var, P.linear_var):
            assert f(int(it), 0, 500, 0.0, 1e-3, 0.0, 1e-3) == pytest.approx(float(w), abs=1e-15)
            assert f(int(it), 500, 10000, 1e-3, 0.0, 0.0, 1e-3) == pytest.approx(float(d), abs=1e-15)


def test_fused_adam_refuses_cpu_parameters():
    import ipoke_inn_b200 as P
    p = torch.nn.Parameter(torch.zeros(4))
    opt = P.FusedAdam([p], lr=1e-3)
    p.grad = torch.ones(4)
    with pytest.raises(RuntimeError):
        opt.step()


def _run_gpu(tag, interrupt_at=None, grad_scale=1.0):
    import ipoke_inn_b200 as P
    G = np.load(GOLD)
    kw = HYPER[tag]
    n = int(G["n_tensors"])
    dev = torch.device("cuda:0")
    params = [torch.nn.Parameter(torch.from_numpy(G[f"{tag}_p0_{i}"]).to(dev)) for i in range(n)]
    opt = P.FusedAdam(params, grad_scale=grad_scale, **kw)
    for step in range(4):
        if interrupt_at == step:                       # checkpoint / resume through the torch state_dict layout
            sd = opt.state_dict()
            sd = {"state": {k: {a: (b.clone() if torch.is_tensor(b) else b) for a, b in s.items()}
                            for k, s in sd["state"].items()}, "param_groups": sd["param_groups"]}
            opt = P.FusedAdam(params, grad_scale=grad_scale, **kw)
            opt.load_state_dict(sd)
        for i, p in enumerate(params):
            p.grad = torch.from_numpy(G[f"{tag}_g{step}_{i}"]).to(dev) / grad_scale
        opt.step()
        for i, p in enumerate(params):
            assert _close(p.detach().cpu().numpy(), G[f"{tag}_p{step + 1}_{i}"]), (tag, i, step)
    for i, p in enumerate(params):
        st = opt.state[p]
        assert _close(st["exp_avg"].cpu().numpy(), G[f"{tag}_m_{i}"])
        assert _close(st["exp_avg_sq"].cpu().numpy(), G[f"{tag}_v_{i}"], 1e-6)
        if kw["amsgrad"]:
            assert _close(st["max_exp_avg_sq"].cpu().numpy(), G[f"{tag}_vmax_{i}"], 1e-6)
        assert int(st["step"]) == 4
    assert set(opt.state_dict()["state"][0].keys()) >= {"step", "exp_avg", "exp_avg_sq"}


@pytest.mark.gpu
@pytest.mark.parametrize("tag", ["ams", "plain"])
def test_fused_adam_matches_torch_adam_golden(tag):
    _run_gpu(tag)


@pytest.mark.gpu
def test_fused_adam_checkpoint_resume_and_grad_scale():
    _run_gpu("ams", interrupt_at=2)
    _run_gpu("ams", grad_scale=0.125)


@pytest.mark.gpu
def test_fused_adam_trains_the_flow():
    """End to end: forward(train) -> FlowLoss -> fused backward -> FusedAdam lowers the loss like torch.optim.Adam."""
    import ipoke_inn_b200 as P
    dev = torch.device("cuda:0")
    cfg = dict(flow_in_channels=8, flow_mid_channels_factor=8, flow_hidden_depth=2, n_flows=4,
               flow_conditioning_option="none", flow_embedding_channels=16)
    g = torch.Generator().manual_seed(3)
    x, c = torch.randn(64, 8, generator=g).to(dev), torch.randn(64, 16, 1, 1, generator=g).to(dev)
    losses = {}
    for name in ("fused", "torch"):
        torch.manual_seed(42)
        net = P.SupervisedTransformer(cfg).to(dev)
        kw = dict(lr=1e-3, betas=(0.9, 0.999), weight_decay=1e-5, amsgrad=True)
        opt = P.FusedAdam(net.parameters(), **kw) if name == "fused" else torch.optim.Adam(net.parameters(), **kw)
        lf = P.FlowLoss()
        hist = []
        for _ in range(6):
            opt.zero_grad(set_to_none=True)
            z, ld = net(x, c)
            loss, _ = lf(z, ld)
            loss.backward()
            opt.step()
            hist.append(float(loss.detach()))
        losses[name] = hist
    assert losses["fused"][-1] < losses["fused"][0]
    assert np.allclose(losses["fused"], losses["torch"], rtol=2e-4, atol=2e-4), losses
