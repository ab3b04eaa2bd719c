"""The numpy oracle of the full MaCow flow (oracle/macow_stack_oracle.py) against fixtures generated from the live
reference (tests/golden/macow_stack_*.npz, tests/golden/make_macow_stack_golden.py).  CPU only.

Bounds: float64 oracle vs the reference's float64 run <= 1e-12; float32 reference outputs vs the float64 oracle
<= 2e-5 forward, 5e-4 through the sequential inverses (max |a - b| / max |b|)."""
import os

import numpy as np
import pytest

from oracle import macow_stack_oracle as so
from conftest import max_rel

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
CASES = ["macow_stack_a", "macow_stack_b", "macow_stack_c"]           # c: conditional (SupervisedMacowTransformer)


def load(name):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    sd = {k[4:]: g[k] for k in g.files if k.startswith("sd::")}
    C, mid, factor, k0, k1 = (int(v) for v in g["meta"])
    return g, sd, dict(C=C, mid=mid, factor=factor, ks=[k0, k1], act=str(g["activation"]),
                       num_steps=[int(v) for v in g["num_steps"]], h=g["h"] if "h" in g.files else None)


@pytest.mark.parametrize("name", CASES)
def test_full_flow_forward_and_inverse(name):
    g, sd, m = load(name)
    fsd = so.sub(sd, "flow.")
    out, ld = so.flow_forward(fsd, g["x"], m["num_steps"], m["factor"], m["act"], h=m["h"])
    assert max_rel(out, g["out64"]) <= 1e-12 and max_rel(ld, g["logdet64"]) <= 1e-12
    assert max_rel(g["out"], out) <= 2e-5 and max_rel(g["logdet"], ld) <= 2e-5
    smp = so.flow_reverse(fsd, g["z_in"], m["num_steps"], m["factor"], m["act"], h=m["h"])
    assert max_rel(smp, g["x_smp64"]) <= 1e-10
    assert max_rel(g["x_smp"], smp) <= 5e-4
    back = so.flow_reverse(fsd, out, m["num_steps"], m["factor"], m["act"], h=m["h"])
    # not exact even in float64: every ActNorm2dFlow inverse divides by exp(log_scale) + 1e-8 (macow2.py:520)
    assert max_rel(back, g["x"].astype(np.float64)) <= 5e-6


@pytest.mark.parametrize("name", CASES)
def test_components(name):
    g, sd, m = load(name)
    step = so.sub(sd, "flow.layers.0.0.")
    x, z = g["x"], g["z_in"]
    for cn, up, st in so.STEP_COUPLINGS:
        csd = so.sub(step, cn + ".")
        o, l = so.nice_apply(csd, x, up, st, 2, m["act"])
        assert max_rel(g[cn + "_out"], o) <= 2e-5 and max_rel(g[cn + "_logdet"], l) <= 2e-5, cn
        assert max_rel(g[cn + "_rev"], so.nice_apply(csd, z, up, st, 2, m["act"], reverse=True)) <= 2e-5, cn
    o, l = so.step_forward(step, x, m["act"], h=m["h"])
    assert max_rel(g["step_out"], o) <= 2e-5 and max_rel(g["step_logdet"], l) <= 2e-5
    assert max_rel(g["step_rev"], so.step_reverse(step, z, m["act"], h=m["h"])) <= 5e-4
    if m["h"] is not None:                                            # one conditional MaCow unit on its own
        from oracle import macow_oracle as mo
        usd = so.sub(step, "units1.0.")
        o, l = mo.unit_forward(usd, x, m["act"], h=m["h"])
        assert max_rel(g["unit_out"], o) <= 2e-5 and max_rel(g["unit_logdet"], l) <= 2e-5
        assert max_rel(g["unit_rev"], mo.unit_reverse(usd, z, m["act"], h=m["h"])) <= 5e-4
    prior = so.sub(sd, "flow.priors.0.")
    o, l = so.prior_forward(prior, x, m["factor"], m["act"])
    assert max_rel(g["prior_out"], o) <= 2e-5 and max_rel(g["prior_logdet"], l) <= 2e-5
    assert max_rel(g["prior_rev"], so.prior_reverse(prior, z, m["factor"], m["act"])) <= 2e-5


def test_level_bookkeeping_matches_the_state_dict():
    g, sd, m = load("macow_stack_b")
    for i, (C, f, n) in enumerate(so.levels(m["C"], m["num_steps"], m["factor"])):
        assert sd[f"flow.shuffle_layers.{i}.forward_shuffle_idx"].shape == (C,)
        assert sd[f"flow.priors.{i}.actnorm.log_scale"].shape[0] == C // f
        c_t, c_in, _, _ = so.nice_geometry(C, f)
        assert sd[f"flow.priors.{i}.coupling.net.conv1.weight"].shape[:2] == (m["mid"], c_in)
        assert sd[f"flow.priors.{i}.coupling.net.conv3.conv.weight_v"].shape[0] == 2 * c_t
        assert f"flow.layers.{i}.{n - 1}.actnorm1.bias" in sd and f"flow.layers.{i}.{n}.actnorm1.bias" not in sd


@pytest.mark.parametrize("name", CASES[:2])
def test_torch_restatement_matches_the_reference(name):
    """oracle/macow_stack_oracle_torch.py (the op sequence the bench times as the reference's stock path)."""
    import torch
    from oracle.macow_stack_oracle_torch import TorchFlow
    g, sd, m = load(name)
    flow = TorchFlow(so.sub(sd, "flow."), m["C"], m["num_steps"], m["factor"], m["act"])
    with torch.no_grad():
        out, ld = flow.forward(torch.from_numpy(g["x"]))
        smp = flow.reverse(torch.from_numpy(g["z_in"]))
    assert max_rel(out.numpy(), g["out"]) <= 2e-6 and max_rel(ld.numpy(), g["logdet"]) <= 2e-6
    assert max_rel(smp.numpy(), g["x_smp"]) <= 1e-5
