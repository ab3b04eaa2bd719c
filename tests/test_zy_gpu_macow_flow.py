"""GPU parity of the full MaCow flow (row N3): ``UnsupervisedMaCowTransformer3`` and its parts through the CUDA kernels
behind include/ipoke_macow.h, against the fixtures of the live reference (tests/golden/macow_stack_*.npz) and the float64
oracle (oracle/macow_stack_oracle.py) on seeded inputs.

Tolerance: the reference computes this path in float32; the kernels are float32 with a different summation order:
max |a - b| / max |b| <= 2e-5 forward and log-determinant, 2e-4 through the inverse (the units' sequential line loops).
Collected late (file name) so that a failure here cannot hide the tests of the flat-INN path under ``pytest -x``."""
import os

import numpy as np
import pytest
import torch

import ipoke_inn_b200 as P
from oracle import macow_stack_oracle as so
from conftest import max_rel
from test_macow_flow_emulation import load

pytestmark = pytest.mark.gpu
TOL_FWD, TOL_REV = 2e-5, 2e-4
DEV = torch.device("cuda:0")       # tests/test_macow_flow_emulation.py re-runs some of these functions on its CPU emulation


def make(cfg, sd, dev):
    net = P.UnsupervisedMaCowTransformer3(cfg)
    net.load_state_dict(sd)
    return net.to(dev).eval()


@pytest.mark.parametrize("name", ["macow_stack_a", "macow_stack_b"])
def test_reference_fixtures(name):
    dev = DEV
    g, sd, cfg = load(name)
    net = make(cfg, sd, dev)
    x = torch.from_numpy(g["x"]).to(dev)
    with torch.no_grad():
        out, ld = net(x)
        smp = net(torch.from_numpy(g["z_in"]).to(dev), reverse=True)
        back = net(out, reverse=True)
        out2, ld2 = net(x)
    assert torch.equal(x.cpu(), torch.from_numpy(g["x"]))
    assert max_rel(out.cpu().numpy(), g["out64"]) <= TOL_FWD and max_rel(out.cpu().numpy(), g["out"]) <= TOL_FWD
    assert max_rel(ld.cpu().numpy(), g["logdet64"]) <= TOL_FWD
    assert max_rel(smp.cpu().numpy(), g["x_smp64"]) <= TOL_REV and max_rel(smp.cpu().numpy(), g["x_smp"]) <= TOL_REV
    assert max_rel(back.cpu().numpy(), g["x"]) <= TOL_REV
    assert torch.equal(out, out2) and torch.equal(ld, ld2)             # deterministic: fixed summation orders


def test_components_called_like_the_reference_calls_them():
    dev = DEV
    g, sd, cfg = load("macow_stack_a")
    net = make(cfg, sd, dev)
    x, z = torch.from_numpy(g["x"]).to(dev), torch.from_numpy(g["z_in"]).to(dev)
    step = net.flow.layers[0][0]
    with torch.no_grad():
        for cn in ("coupling1_up", "coupling1_dn", "coupling2_up", "coupling2_dn"):
            o, l = getattr(step, cn)(x)
            assert max_rel(o.cpu().numpy(), g[cn + "_out"]) <= TOL_FWD, cn
            assert max_rel(l.cpu().numpy(), g[cn + "_logdet"]) <= TOL_FWD, cn
            assert max_rel(getattr(step, cn)(z, reverse=True).cpu().numpy(), g[cn + "_rev"]) <= TOL_FWD, cn
        o, l = step(x)
        assert max_rel(o.cpu().numpy(), g["step_out"]) <= TOL_FWD and max_rel(l.cpu().numpy(), g["step_logdet"]) <= TOL_FWD
        assert max_rel(step(z, reverse=True).cpu().numpy(), g["step_rev"]) <= TOL_REV
        o, l = net.flow.priors[0](x)
        assert max_rel(o.cpu().numpy(), g["prior_out"]) <= TOL_FWD and max_rel(l.cpu().numpy(), g["prior_logdet"]) <= TOL_FWD
        assert max_rel(net.flow.priors[0](z, reverse=True).cpu().numpy(), g["prior_rev"]) <= TOL_FWD


def test_first_call_initialisation_matches_reference():
    dev = DEV
    g, sd, cfg = load("macow_stack_b")
    torch.manual_seed(42)
    net = P.UnsupervisedMaCowTransformer3(cfg).to(dev).eval()
    with torch.no_grad():
        out, ld = net(torch.from_numpy(g["x"]).to(dev))
    assert max_rel(out.cpu().numpy(), g["init_out"]) <= TOL_FWD and max_rel(ld.cpu().numpy(), g["init_logdet"]) <= TOL_FWD
    new = {k: v.cpu().numpy() for k, v in net.state_dict().items()}
    for k in g.files:
        if k.startswith("init::"):
            np.testing.assert_allclose(new[k[6:]], g[k], rtol=1e-4, atol=2e-5, err_msg=k)
    assert all(int(v) == 1 for k, v in new.items() if k.endswith("initialized"))


def test_conditional_flow_reference_fixture():
    """SupervisedMacowTransformer (INN.py:447-482, condition_nice false): forward / inverse with the conditioning, a
    conditional unit and a step on their own, the first-call initialisation."""
    dev = DEV
    g, sd, cfg = load("macow_stack_c")
    x, z, h = (torch.from_numpy(g[k]).to(dev) for k in ("x", "z_in", "h"))
    torch.manual_seed(42)
    net = P.SupervisedMacowTransformer(cfg).to(dev).eval()
    with torch.no_grad():
        out, ld = net(x, h)
    assert max_rel(out.cpu().numpy(), g["init_out"]) <= TOL_FWD and max_rel(ld.cpu().numpy(), g["init_logdet"]) <= TOL_FWD
    for k in g.files:
        if k.startswith("init::"):
            np.testing.assert_allclose(net.state_dict()[k[6:]].cpu().numpy(), g[k], rtol=1e-4, atol=2e-5, err_msg=k)
    net.load_state_dict(sd)
    with torch.no_grad():
        out, ld = net(x, h)
        smp = net(z, h, reverse=True)
        back = net(out, h, reverse=True)
        assert max_rel(out.cpu().numpy(), g["out64"]) <= TOL_FWD and max_rel(ld.cpu().numpy(), g["logdet64"]) <= TOL_FWD
        assert max_rel(smp.cpu().numpy(), g["x_smp64"]) <= TOL_REV and max_rel(back.cpu().numpy(), g["x"]) <= TOL_REV
        step = net.flow.layers[0][0]
        o, l = step.units1[0](x, h=h)
        assert max_rel(o.cpu().numpy(), g["unit_out"]) <= TOL_FWD and max_rel(l.cpu().numpy(), g["unit_logdet"]) <= TOL_FWD
        assert max_rel(step.units1[0](z, h=h, reverse=True).cpu().numpy(), g["unit_rev"]) <= TOL_REV
        o, l = step(x, h=h)
        assert max_rel(o.cpu().numpy(), g["step_out"]) <= TOL_FWD and max_rel(l.cpu().numpy(), g["step_logdet"]) <= TOL_FWD
        assert max_rel(step(z, h=h, reverse=True).cpu().numpy(), g["step_rev"]) <= TOL_REV
        assert not torch.equal(net(x, h + 0.5)[0], out)
        smp2 = net.sample(tuple(x.shape), h, device=dev)
        assert smp2.shape == x.shape
        with pytest.raises(ValueError):
            net.flow(x)                                                # a conditional flow needs its conditioning


def test_conditional_flow_against_float64_oracle():
    """z_dim-32 latents at 8 x 8 with 128 conditioning channels (conditioner + poke embedding), batch 20 (the reference's
    test batch): two conditional unit chains per step, 1024-column conditioning product."""
    dev = DEV
    C, hc, B, H, W = 32, 128, 20, 8, 8
    cfg = dict(num_steps=[1], flow_in_channels=C, flow_mid_channels=64, factor=16, transform="affine",
               prior_transform="affine", kernel_size=[2, 3], coupling_type="conv", activation="elu", h_channels=hc,
               flow_attn_heads=4, cond_conv_hidden_channels=16)
    torch.manual_seed(5)
    net = randomise(P.SupervisedMacowTransformer(cfg), 55).to(dev).eval()
    sd = so.sub({k: v.cpu().numpy() for k, v in net.state_dict().items()}, "flow.")
    rng = np.random.default_rng(9)
    x = rng.standard_normal((B, C, H, W)).astype(np.float32)
    z = rng.standard_normal((B, C, H, W)).astype(np.float32)
    h = rng.standard_normal((B, hc, H, W)).astype(np.float32)
    want, want_ld = so.flow_forward(sd, x, [1], 16, "elu", h=h)
    want_rev = so.flow_reverse(sd, z, [1], 16, "elu", h=h)
    with torch.no_grad():
        out, ld = net(torch.from_numpy(x).to(dev), torch.from_numpy(h).to(dev))
        got = net(torch.from_numpy(z).to(dev), torch.from_numpy(h).to(dev), reverse=True)
    assert max_rel(out.cpu().numpy(), want) <= TOL_FWD and max_rel(ld.cpu().numpy(), want_ld) <= TOL_FWD
    assert max_rel(got.cpu().numpy(), want_rev) <= TOL_REV


def randomise(net, seed):
    """Non-trivial gains / biases for every zero-initialised convolution and ActNorm (a fresh flow is the identity).
    The gains are kept small: with alpha = 1 the affine scale tanh(log_scale / 2) + 1 reaches 0 when a random net
    saturates the tanh, and the inverse of such a flow amplifies float32 rounding without bound -- in the reference's own
    float32 arithmetic just the same (checked with oracle/macow_stack_oracle_torch.py: 6e-3 at gains up to 0.4)."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for k, p in net.state_dict().items():
            if k.endswith("weight_g"):
                p.copy_(0.02 + 0.06 * torch.rand(p.shape, generator=g))
            elif k.endswith("conv.bias") or k.endswith("actnorm.bias") or k.endswith("actnorm1.bias") or k.endswith("actnorm2.bias"):
                p.copy_(0.1 * torch.randn(p.shape, generator=g))
            elif k.endswith("log_scale"):
                p.copy_(0.1 * torch.randn(p.shape, generator=g))
            elif k.endswith("initialized"):
                p.fill_(1)
    return net


SHAPES = [  # channels, hidden of the couplings, factor, num_steps, H, W, batch, activation, coupling-pass budget (MB)
    (32, 96, 16, [1, 1], 8, 8, 37, "elu", None),         # z_dim-32 latents at 8 x 8 (config/second_stage.yaml), 19 row tiles
    (16, 136, 4, [1, 1, 1], 16, 16, 5, "elu", "1"),      # 16 x 16 latents; 1 MB budget: 3 samples per coupling pass (2 passes)
    (12, 20, 4, [2], 5, 7, 3, "leaky_relu", None),       # non-square, odd level widths
]


@pytest.mark.parametrize("C,mid,factor,steps,H,W,B,act,mb", SHAPES)
def test_against_float64_oracle(C, mid, factor, steps, H, W, B, act, mb, monkeypatch):
    dev = DEV
    if mb is not None:
        monkeypatch.setenv("IPOKE_MACOW_NICE_MB", mb)
    cfg = dict(num_steps=steps, flow_in_channels=C, flow_mid_channels=mid, factor=factor, transform="affine",
               prior_transform="affine", kernel_size=[2, 3], coupling_type="conv", activation=act)
    torch.manual_seed(C)
    net = randomise(P.UnsupervisedMaCowTransformer3(cfg), 100 + C).to(dev).eval()
    sd = so.sub({k: v.cpu().numpy() for k, v in net.state_dict().items()}, "flow.")
    rng = np.random.default_rng(C)
    x = rng.standard_normal((B, C, H, W)).astype(np.float32)
    z = rng.standard_normal((B, C, H, W)).astype(np.float32)
    want, want_ld = so.flow_forward(sd, x, steps, factor, act)
    want_rev = so.flow_reverse(sd, z, steps, factor, act)
    with torch.no_grad():
        out, ld = net(torch.from_numpy(x).to(dev))
        got = net(torch.from_numpy(z).to(dev), reverse=True)
        back = net(out, reverse=True)
    assert max_rel(out.cpu().numpy(), want) <= TOL_FWD and max_rel(ld.cpu().numpy(), want_ld) <= TOL_FWD
    assert max_rel(got.cpu().numpy(), want_rev) <= TOL_REV
    assert max_rel(back.cpu().numpy(), x) <= TOL_REV


def test_inn_test_stack_against_float64_restatement():
    """The benchmark shape itself (tools/macow_flow_bench.py --workload inn_test: config/INN_test.yaml, 45 steps over 7
    levels, couplings 1536 wide, 467 M parameters): 4 samples against the float64 run of the torch restatement of the
    reference's op sequence (oracle/macow_stack_oracle_torch.py, pinned to the live-reference fixtures) on the same GPU,
    the round trip, and rows of a larger batch against the same rows alone.
    Bounds: 1e-4 forward / log-determinant over the 277 stacked layers (the reference's own float32 arithmetic is 2e-6
    from float64 here, measured on the CPU), 2e-4 for the round trip (float64 itself: 1e-5, every ActNorm2dFlow inverse
    divides by exp(log_scale) + 1e-8)."""
    from oracle.macow_stack_oracle_torch import TorchFlow
    dev = DEV
    cfg = dict(num_steps=[20, 8, 5, 5, 3, 3, 1], flow_in_channels=8, flow_mid_channels=1536, factor=8, transform="affine",
               prior_transform="affine", kernel_size=[2, 3], coupling_type="conv", activation="elu")
    torch.manual_seed(1)
    net = P.UnsupervisedMaCowTransformer3(cfg).eval()
    g = torch.Generator().manual_seed(17)
    with torch.no_grad():                                    # the bench tool's "trained-looking" parameters
        for k, v in net.state_dict().items():
            if k.endswith("initialized"):
                v.fill_(1)
            elif k.endswith("weight_g"):
                v.copy_(0.02 + 0.06 * torch.rand(v.shape, generator=g))
            elif k.endswith("bias") or k.endswith("log_scale"):
                v.copy_(0.05 * torch.randn(v.shape, generator=g))
    sd = so.sub({k: v.numpy() for k, v in net.state_dict().items()}, "flow.")
    ref = TorchFlow(sd, 8, cfg["num_steps"], 8, "elu", 1.0, device=dev, dtype=torch.float64)
    net = net.to(dev)
    x = torch.randn(24, 8, 8, 8, generator=torch.Generator().manual_seed(3)).to(dev)
    with torch.no_grad():
        want, want_ld = ref.forward(x[:4].double())
        out, ld = net(x)
        back = net(out, reverse=True)
        out4, ld4 = net(x[:4])
    assert max_rel(out[:4].cpu().numpy(), want.cpu().numpy()) <= 1e-4
    assert max_rel(ld[:4].cpu().numpy(), want_ld.cpu().numpy()) <= 1e-4
    assert max_rel(back.cpu().numpy(), x.cpu().numpy()) <= TOL_REV
    assert max_rel(out4.cpu().numpy(), out[:4].cpu().numpy()) <= 1e-6 and max_rel(ld4.cpu().numpy(), ld[:4].cpu().numpy()) <= 1e-6


def test_large_batch_round_trip_and_sample():
    dev = DEV
    cfg = dict(num_steps=[1, 1], flow_in_channels=8, flow_mid_channels=64, factor=4, transform="affine",
               prior_transform="affine", kernel_size=[2, 3], coupling_type="conv", activation="elu")
    torch.manual_seed(3)
    net = randomise(P.UnsupervisedMaCowTransformer3(cfg), 7).to(dev).eval()
    x = torch.randn(2048, 8, 8, 8, device=dev)
    with torch.no_grad():
        z, ld = net(x)
        back = net(z, reverse=True)
        smp = net.sample((4, 8, 8, 8), device=dev)
    assert torch.isfinite(ld).all() and torch.isfinite(smp).all() and smp.shape == (4, 8, 8, 8)
    assert max_rel(back.cpu().numpy(), x.cpu().numpy()) <= TOL_REV
    # rows of a batch are independent: the first 16 samples alone give the same result
    with torch.no_grad():
        z16, ld16 = net(x[:16])
    assert max_rel(z16.cpu().numpy(), z[:16].cpu().numpy()) <= 1e-6 and max_rel(ld16.cpu().numpy(), ld[:16].cpu().numpy()) <= 1e-6
    with pytest.raises(NotImplementedError):
        net(x.requires_grad_())
