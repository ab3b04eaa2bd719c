"""GPU tests of the multi-draw sampling call (ipoke_inn_reverse_draws / ipoke_inn_forward_draws): n_draws latents per
conditioning row in one call must equal the reference's per-draw loop (second_stage_video_fc.py:356-378: one
``flow(randn, cond, reverse=True)`` per draw with the same cond) -- checked against
(1) the per-draw calls of the same module (same operands and MMAs; fp32 summation order of split-K tails may differ),
(2) the numpy oracle on the row-repeated conditioning (<= 1e-4 in fp32 mode, the stated bf16 bound otherwise),
(3) the forward -> reverse round trip across draws.
"""
import numpy as np
import pytest
import torch

import ipoke_inn_b200 as P
from oracle import inn_oracle as O
from conftest import max_rel

pytestmark = pytest.mark.gpu

CFG_A_SMALL = dict(flow_in_channels=128, flow_mid_channels_factor=2, flow_hidden_depth=3, n_flows=4,
                   flow_conditioning_option="none", flow_embedding_channels=128)
CFG_S = dict(flow_in_channels=8, flow_mid_channels_factor=8, flow_hidden_depth=2, n_flows=8,
             flow_conditioning_option="none", flow_embedding_channels=16)


def _net(cfg, precision, B_init=64):
    dev = torch.device("cuda:0")
    torch.manual_seed(42)
    net = P.SupervisedTransformer(dict(cfg), precision=precision).to(dev)
    g = torch.Generator().manual_seed(7)
    x0 = torch.randn(B_init, cfg["flow_in_channels"], generator=g).to(dev)
    c0 = torch.randn(B_init, cfg["flow_embedding_channels"], 1, 1, generator=g).to(dev)
    with torch.no_grad():
        net(x0, c0)                       # ActNorm data init
    return net, dev


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("cfg,B,n", [(CFG_A_SMALL, 128, 5), (CFG_A_SMALL, 20, 5), (CFG_A_SMALL, 384, 3),
                                     (CFG_S, 10, 4), (CFG_A_SMALL, 100, 1)])
def test_reverse_draws_equals_per_draw_loop(cfg, B, n, precision):
    net, dev = _net(cfg, precision)
    C, E = cfg["flow_in_channels"], cfg["flow_embedding_channels"]
    g = torch.Generator().manual_seed(1234)
    z = torch.randn(n, B, C, generator=g).to(dev)
    cond = torch.randn(B, E, 1, 1, generator=g).to(dev)
    with torch.no_grad():
        loop = torch.stack([net(z[j], cond, reverse=True) for j in range(n)])       # the reference's loop
        one = net.flow.reverse_draws(z, cond)
    assert one.shape == (n, B, C, 1, 1)
    assert torch.isfinite(one).all()
    # same operands through the same MMAs; what may differ is the launch shape picked for n*B rows instead of B (the
    # K-split of the small-batch coupling tails changes the fp32 summation order, and in bf16 mode a last-bit change
    # can flip the bf16 rounding of the next subnetwork input)
    assert max_rel(one.cpu().numpy(), loop.cpu().numpy()) < (2e-5 if precision == "fp32" else 1e-2)
    # oracle on the repeated conditioning
    sd = O.state_dict_to_numpy(net.state_dict())
    zr = z.reshape(n * B, C).cpu().numpy()
    cr = cond.reshape(B, E).repeat(n, 1).cpu().numpy()
    ref = O.flow_reverse(sd, zr, cr)
    tol = 1e-4 if precision == "fp32" else 1e-2
    assert max_rel(one.reshape(n * B, C).cpu().numpy(), ref) < tol


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_forward_draws_round_trip_and_oracle(precision):
    cfg, B, n = CFG_A_SMALL, 256, 4
    net, dev = _net(cfg, precision)
    C, E = cfg["flow_in_channels"], cfg["flow_embedding_channels"]
    g = torch.Generator().manual_seed(99)
    x = torch.randn(n, B, C, generator=g).to(dev)
    cond = torch.randn(B, E, 1, 1, generator=g).to(dev)
    z, ld = net.flow.forward_draws(x, cond)
    assert z.shape == (n, B, C, 1, 1) and ld.shape == (n, B)
    sd = O.state_dict_to_numpy(net.state_dict())
    z_ref, ld_ref = O.flow_forward(sd, x.reshape(n * B, C).cpu().numpy(), cond.reshape(B, E).repeat(n, 1).cpu().numpy())
    tol = 1e-4 if precision == "fp32" else 1e-2
    assert max_rel(z.reshape(n * B, C).cpu().numpy(), z_ref) < tol
    ld_tol = 1e-4 * np.max(np.abs(ld_ref)) if precision == "fp32" else 2e-3 * np.sqrt(cfg["n_flows"] * C)
    assert np.max(np.abs(ld.reshape(-1).cpu().numpy() - ld_ref)) < ld_tol
    back = net.flow.reverse_draws(z.reshape(n, B, C), cond)
    assert float((back.reshape(n, B, C) - x).abs().max()) < (1e-4 if precision == "fp32" else 2e-2)


def test_sample_many_shapes_and_draw_independence():
    net, dev = _net(CFG_A_SMALL, "bf16")
    cond = torch.randn(20, 128, 1, 1, device=dev)
    g = torch.Generator(device=dev).manual_seed(5)
    s = net.sample_many(5, cond, generator=g)
    assert s.shape == (5, 20, 128, 1, 1) and torch.isfinite(s).all()
    assert float((s[0] - s[1]).abs().max()) > 1e-3            # different latents -> different samples


def test_draws_error_paths():
    net, dev = _net(CFG_A_SMALL, "bf16")
    with pytest.raises(RuntimeError):
        net.flow.reverse_draws(torch.randn(3, 16, 128, device=dev), torch.randn(8, 128, 1, 1, device=dev))
    torch.manual_seed(0)
    unc = P.UnsupervisedTransformer3(flow_in_channels=16, flow_mid_channels=64, flow_hidden_depth=2, n_flows=2).to(dev)
    with pytest.raises(RuntimeError):
        unc.flow.engine.reverse_draws(torch.randn(2, 4, 16, device=dev), torch.randn(4, 8, device=dev), 2)
