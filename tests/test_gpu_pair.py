"""GPU parity of the CTA-pair hidden-layer GEMM (hidden_pair_kernel.cuh: tcgen05 cta_group::2, used for multi-wave
launches) against the numpy oracle at batch sizes that select it: 256- and 128-column pair tiles, an odd number of
128-row tiles (the last pair's second CTA works on rows past the batch), conditional and unconditional stacks, both
precisions; plus equality with the single-CTA kernels (IPOKE_INN_PAIR=0 plan) to fp32 rounding."""
import os

import numpy as np
import pytest
import torch

import ipoke_inn_b200 as P
from oracle import inn_oracle as O
from conftest import max_rel

pytestmark = pytest.mark.gpu

CASES = {
    "cond_bn256": ("cond", dict(flow_in_channels=64, flow_mid_channels_factor=4, flow_hidden_depth=2, n_flows=2,
                                flow_conditioning_option="none", flow_embedding_channels=32), 16500),
    "uncond_bn128": ("uncond", dict(flow_in_channels=64, flow_mid_channels=128, flow_hidden_depth=1, n_flows=2), 32768),
    "uncond_bn256_k1024": ("uncond", dict(flow_in_channels=128, flow_mid_channels=1024, flow_hidden_depth=2, n_flows=1), 12000),
}


def _build(kind, kw, precision):
    torch.manual_seed(42)
    if kind == "cond":
        return P.SupervisedTransformer(dict(kw), precision=precision)
    return P.UnsupervisedTransformer3(**kw, precision=precision)


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("name", list(CASES))
def test_pair_kernel_matches_oracle_and_single_cta(name, precision):
    kind, kw, B = CASES[name]
    dev = torch.device("cuda:0")
    net = _build(kind, kw, precision).to(dev)
    C, E = kw["flow_in_channels"], kw.get("flow_embedding_channels", 0)
    g = torch.Generator().manual_seed(3)
    x = torch.randn(B, C, generator=g)
    cond = torch.randn(B, E, 1, 1, generator=g) if kind == "cond" else None
    xd, cd = x.to(dev), (cond.to(dev) if cond is not None else None)
    call = (lambda inp, **k: net(inp, cd, **k)) if kind == "cond" else (lambda inp, **k: net(inp, **k))
    os.environ["IPOKE_INN_PAIR"] = "2"                 # read at plan creation (first call): pairs for every multi-wave launch
    try:
        with torch.no_grad():
            z, ld = call(xd)
            xr = call(z, reverse=True)
    finally:
        del os.environ["IPOKE_INN_PAIR"]
    sd = O.state_dict_to_numpy(net.state_dict())
    z_ref, ld_ref = O.flow_forward(sd, x.numpy(), cond.numpy().reshape(B, E) if cond is not None else None)
    tol = 1e-4 if precision == "fp32" else 1e-2
    assert max_rel(z.reshape(B, C).cpu().numpy(), z_ref) < tol
    n_flows = kw["n_flows"]
    ld_tol = 1e-4 * np.max(np.abs(ld_ref)) if precision == "fp32" else 2e-3 * np.sqrt(n_flows * C)
    assert np.max(np.abs(ld.cpu().numpy() - ld_ref)) < ld_tol
    assert float((xr.reshape(B, C).cpu() - x).abs().max()) < (1e-4 if precision == "fp32" else 3e-2)
    # the same module through the single-CTA kernels
    os.environ["IPOKE_INN_PAIR"] = "0"
    try:
        net1 = _build(kind, kw, precision).to(dev)
        net1.load_state_dict(net.state_dict())
        call1 = (lambda inp, **k: net1(inp, cd, **k)) if kind == "cond" else (lambda inp, **k: net1(inp, **k))
        with torch.no_grad():
            z1, ld1 = call1(xd)
    finally:
        del os.environ["IPOKE_INN_PAIR"]
    assert max_rel(z.cpu().numpy(), z1.cpu().numpy()) < (2e-5 if precision == "fp32" else 1e-2)
