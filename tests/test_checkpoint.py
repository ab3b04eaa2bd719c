"""Checkpoint compatibility on the CPU: the reference's state_dict -- taken from the committed
live-reference fixtures -- wrapped the way its LightningModules save it ('flow.' / 'INN.' prefixes next to other
sub-models' keys) loads strictly; export round-trips; checkpoint selection follows experiments/experiment.py:107-142."""
import os

import pytest
import torch
import yaml

from ipoke_inn_b200 import checkpoint as C
from conftest import load_golden
from test_host_logic import build


@pytest.mark.parametrize("name,prefix", [("cfgA_mini", "flow."), ("cfgB_mini", "INN.")])
def test_lightning_checkpoint_round_trip(tmp_path, name, prefix):
    sd_ref, _ = load_golden(name)
    ckpt = {"epoch": 3, "global_step": 1234,
            "state_dict": {**{prefix + k: torch.from_numpy(v) for k, v in sd_ref.items()},
                           "first_stage_model.enc.weight": torch.zeros(2, 2), "loss_func.dummy": torch.ones(1)}}
    path = tmp_path / "epoch=3.ckpt"
    torch.save(ckpt, path)
    net = build(name, seed=7)                                   # different init
    res = C.load_flow_from_lightning(net, str(path), prefix=prefix)
    assert not res.missing_keys and not res.unexpected_keys
    for k, v in net.state_dict().items():
        assert torch.equal(v, torch.from_numpy(sd_ref[k])), k
    out = C.export_flow_to_lightning(net, ckpt, prefix=prefix, path=str(tmp_path / "out.ckpt"))
    assert out["epoch"] == 3 and "first_stage_model.enc.weight" in out["state_dict"]
    back = torch.load(tmp_path / "out.ckpt", map_location="cpu", weights_only=False)["state_dict"]
    assert set(back) == set(ckpt["state_dict"])
    for k in ckpt["state_dict"]:
        assert torch.equal(back[k], ckpt["state_dict"][k]), k
    with pytest.raises(KeyError):
        C.load_flow_from_lightning(net, ckpt, prefix="nothing.")


def test_select_checkpoint_like_the_reference(tmp_path):
    d = tmp_path
    with pytest.raises(ValueError, match="Not enough"):
        C.select_checkpoint(str(d))
    a, b, c = [str(d / n) for n in ("a.ckpt", "b.ckpt", "gone.ckpt")]
    for p in (a, b):
        torch.save({"state_dict": {}}, p)
    with open(d / "best_k_models.yaml", "w") as f:
        yaml.safe_dump({a: 0.7, b: 0.4, c: 0.1}, f)             # best score belongs to a file that no longer exists
    assert C.select_checkpoint(str(d)) == b
    with pytest.raises(ValueError, match="no last ckpt"):
        C.select_checkpoint(str(d), last_ckpt=True)
    torch.save({"state_dict": {}}, d / "last.ckpt")
    assert C.select_checkpoint(str(d), last_ckpt=True) == str(d / "last.ckpt")
    with open(d / "second.yaml", "w") as f:
        yaml.safe_dump({a: 0.1}, f)
    with pytest.raises(ValueError, match="Too many"):
        C.select_checkpoint(str(d))
    os.remove(d / "second.yaml")
    os.remove(a); os.remove(b)
    with pytest.raises(ValueError, match="No valid files"):
        C.select_checkpoint(str(d))


@pytest.mark.parametrize("name", ["macow_stack_a", "macow_stack_c"])
def test_macow_lightning_checkpoint_round_trip(tmp_path, name):
    """The conv second stage / FlowMotion save their MaCow flow under 'flow.' (models/second_stage_video.py:142,
    models/flow_motion.py): the live-reference state_dict of the fixtures loads strictly into the drop-in wrappers."""
    import ipoke_inn_b200 as P
    from test_macow_flow_emulation import load
    g, sd_ref, cfg = load(name)
    ckpt = {"state_dict": {**{"flow." + k: v for k, v in sd_ref.items()}, "first_stage_model.enc.weight": torch.zeros(2, 2)}}
    path = tmp_path / "macow.ckpt"
    torch.save(ckpt, path)
    torch.manual_seed(3)
    net = (P.SupervisedMacowTransformer if "h_channels" in cfg else P.UnsupervisedMaCowTransformer3)(cfg)
    res = C.load_flow_from_lightning(net, str(path), prefix="flow.")
    assert not res.missing_keys and not res.unexpected_keys
    for k, v in net.state_dict().items():
        assert torch.equal(v, sd_ref[k]), k
    out = C.export_flow_to_lightning(net, ckpt, prefix="flow.")
    assert set(out["state_dict"]) == set(ckpt["state_dict"])
