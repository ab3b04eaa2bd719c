"""Pin the numpy oracle against the live-reference fixtures (tests/golden/make_golden.py).

CPU only.  Tolerances: the oracle runs the same fp32 arithmetic as the reference with a
different BLAS / summation order, so agreement is to fp32 rounding (<= 2e-6 relative);
permutation gathers are bit-exact.
"""
import numpy as np
import pytest

from oracle import inn_oracle as O
from conftest import load_golden, max_rel, rel_l2

TOL = 2e-6


def _cond(data):
    c = data.get("cond")
    return None if c is None else c.reshape(c.shape[0], -1)


def test_forward_matches_reference(golden):
    name, sd, d = golden
    z, ld, outs, lds = O.flow_forward(sd, d["x"], _cond(d), return_blocks=True)
    B, C = d["x"].shape
    assert max_rel(z, d["z"].reshape(B, C)) < TOL
    assert max_rel(ld, d["logdet"]) < TOL
    assert max_rel(np.stack(outs), d["last_outs"]) < TOL
    assert max_rel(np.stack(lds), d["last_logdets"]) < TOL


def test_forward_fp64_matches_reference_fp64(golden):
    name, sd, d = golden
    z, ld = O.flow_forward(sd, d["x"], _cond(d), dtype=np.float64)
    B, C = d["x"].shape
    assert max_rel(z, d["z64"].reshape(B, C)) < 1e-12
    assert max_rel(ld, d["logdet64"]) < 1e-12


def test_reverse_matches_reference(golden):
    name, sd, d = golden
    B, C = d["x"].shape
    x_rec = O.flow_reverse(sd, d["z"].reshape(B, C), _cond(d))
    assert max_rel(x_rec, d["x_rec"].reshape(B, C)) < 5e-6
    x_smp = O.flow_reverse(sd, d["z_in"].reshape(B, C), _cond(d))
    assert max_rel(x_smp, d["x_smp"].reshape(B, C)) < 5e-6
    # round trip (reference achieves ~1e-6)
    assert np.max(np.abs(x_rec - d["x"])) < 2e-5


def test_permutation_bit_exact(golden):
    name, sd, d = golden
    for i in range(O.n_flows_of(sd)):
        p = f"flow.sub_layers.{i}.shuffle"
        f, b = sd[p + ".forward_shuffle_idx"], sd[p + ".backward_shuffle_idx"]
        assert f.dtype == np.int64 and b.dtype == np.int64
        assert np.array_equal(np.argsort(f), b)               # flow_blocks.py:320
        x = np.arange(3 * f.size, dtype=np.float32).reshape(3, f.size)
        assert np.array_equal(O.shuffle_reverse(sd, p, O.shuffle_forward(sd, p, x)), x)


def test_actnorm_init_reproduces_reference_state(golden):
    """Reset loc/scale/initialized, replay the first call, recover the stored state."""
    name, sd, d = golden
    sd0 = dict(sd)
    for i in range(O.n_flows_of(sd)):
        p = f"flow.sub_layers.{i}.norm_layer"
        sd0[p + ".loc"] = np.zeros_like(sd[p + ".loc"])
        sd0[p + ".scale"] = np.ones_like(sd[p + ".scale"])
        sd0[p + ".initialized"] = np.array(0, dtype=np.uint8)
    z, ld = O.flow_forward_with_actnorm_init(sd0, d["x"], _cond(d))
    B, C = d["x"].shape
    for i in range(O.n_flows_of(sd)):
        p = f"flow.sub_layers.{i}.norm_layer"
        assert max_rel(sd0[p + ".loc"], sd[p + ".loc"]) < 2e-5
        assert max_rel(sd0[p + ".scale"], sd[p + ".scale"]) < 2e-5
        assert int(sd0[p + ".initialized"]) == 1
    assert max_rel(z, d["z"].reshape(B, C)) < 2e-5
    assert max_rel(ld, d["logdet"]) < 2e-5


@pytest.mark.parametrize("name", ["cfgS_cond", "uncond_small"])
def test_loss_and_backward_match_autograd(name):
    sd, d = load_golden(name)
    B, C = d["x"].shape
    cond = _cond(d)
    z, ld = O.flow_forward(sd, d["x"], cond, dtype=np.float64)
    loss, nll_loss, nlogdet = O.flow_loss(z, ld)
    assert abs(loss - float(d["loss"])) < 1e-4 * abs(float(d["loss"]))
    assert abs(nll_loss - float(d["nll_loss"])) < 1e-4 * abs(float(d["nll_loss"]))
    # d loss / dz = z / B ; d loss / dlogdet = -1/B   (loss.py:18-25, 90-97)
    grads, dx, dcond = O.flow_backward(sd, d["x"], cond, z / B, -np.ones(B) / B)
    assert rel_l2(dx, d["dx"]) < 1e-4
    if cond is not None:
        assert rel_l2(dcond, d["dcond"].reshape(B, -1)) < 1e-4
    checked = 0
    for k, v in d.items():
        if k.startswith("grad::"):
            assert rel_l2(grads[k[6:]], v) < 2e-4, k
            checked += 1
    assert checked == sum(1 for k in sd if k.endswith((".weight", ".bias", ".loc", ".scale")))


def test_macs_per_sample():
    sd, _ = load_golden("cfgS_cond")
    assert O.macs_per_sample(sd) == 307200            # BASELINE.md section 3, cfg S


def test_torch_port_matches_reference(golden):
    """The torch-operator restatement timed by bench.py's CPU arm (oracle/inn_oracle_torch.py) against the same
    live-reference fixtures."""
    import torch
    from oracle import inn_oracle_torch as OT
    name, sd, d = golden
    sdt = {k: torch.from_numpy(np.ascontiguousarray(v)) for k, v in sd.items()}
    B, C = d["x"].shape
    c = _cond(d)
    ct = None if c is None else torch.from_numpy(np.ascontiguousarray(c))
    z, ld = OT.flow_forward(sdt, torch.from_numpy(d["x"]), ct)
    assert max_rel(z.numpy(), d["z"].reshape(B, C)) < TOL
    assert max_rel(ld.numpy(), d["logdet"]) < TOL
    x_smp = OT.flow_reverse(sdt, torch.from_numpy(np.ascontiguousarray(d["z_in"].reshape(B, C))), ct)
    assert max_rel(x_smp.numpy(), d["x_smp"].reshape(B, C)) < 5e-6
