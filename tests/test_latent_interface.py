"""Host logic of the latent interface on the CPU, with a stand-in flow that records
its inputs: the interface must hand the flow what the reference's make_flow_input / forward_sample / forward_density
hand it (models/second_stage_video_fc.py:273-387; experiments/fully_connected_ae_INN.py:104-127), draw latents in
the reference's RNG order, and encode the conditioning once per batch instead of once per draw.  The reference's
LightningModules cannot be imported here (pytorch_lightning / lpips / kornia are absent), so the expected values
are restated in the test from the cited lines.  The GPU counterpart (real flow) is test_gpu_latent_interface.py."""
import types

import pytest
import torch

from ipoke_inn_b200.latent_interface import FlowAELatentInterface, SecondStageLatentInterface

B, C, EI, EP, T = 6, 8, 5, 3, 4


class _Flow:
    """Stand-in for SupervisedTransformer: affine map so results are checkable, call log for the assertions."""

    def __init__(self):
        self.flow = types.SimpleNamespace(reshape="none", reverse_draws=self._reverse_draws)
        self.calls = []

    def __call__(self, x, cond, reverse=False):
        self.calls.append(("call", tuple(x.shape), tuple(cond.shape), reverse))
        bias = cond.reshape(cond.shape[0], -1).sum(1).reshape(-1, 1, 1, 1)
        if reverse:
            return x.reshape(x.shape[0], -1, 1, 1) * 2 + bias
        return x.reshape(x.shape[0], -1, 1, 1) * 0.5, torch.zeros(x.shape[0])

    def _reverse_draws(self, z, cond, n):
        self.calls.append(("draws", tuple(z.shape), tuple(cond.shape), n))
        return torch.stack([self(z[j], cond, reverse=True) for j in range(n)])


def _parts(count):
    torch.manual_seed(3)
    wp, wi, wm = torch.randn(2, EP), torch.randn(3, EI), torch.randn(3 * T, C)

    def poke_enc(p):
        count["poke"] += 1
        return p.mean((2, 3)) @ wp

    def cond_enc(x0):
        count["cond"] += 1
        return x0.mean((2, 3)) @ wi

    def enc_fs(X):
        return X.mean((3, 4)).reshape(X.shape[0], -1) @ wm, None

    def dec_fs(motion, X):
        return motion[:, None, :3, None, None].expand(-1, X.shape[1] - 1, -1, 4, 4) + motion.sum(1).reshape(-1, 1, 1, 1, 1)
    return poke_enc, cond_enc, enc_fs, dec_fs


def _batch():
    g = torch.Generator().manual_seed(11)
    return {"images": torch.randn(B, T, 3, 4, 4, generator=g), "poke": [torch.randn(B, 2, 4, 4, generator=g), 0]}


def _iface(flow, count, **kw):
    pe, ce, ef, df = _parts(count)
    return SecondStageLatentInterface(flow, pe, ce, ef, df, flow_in_channels=C, **kw)


def test_make_flow_input_reverse_matches_reference_recipe():
    count = {"poke": 0, "cond": 0}
    it = _iface(_Flow(), count)
    batch = _batch()
    torch.manual_seed(42)
    flow_input, cond = it.make_flow_input(batch, reverse=True)
    # restated from second_stage_video_fc.py:320-322, 338-343
    torch.manual_seed(42)
    expect = torch.randn((B, C, 1, 1)).type_as(batch["images"])
    assert torch.equal(flow_input, expect) and not flow_input.requires_grad
    pe, ce, _, _ = _parts({"poke": 0, "cond": 0})
    c_ref = torch.cat([ce(batch["images"][:, 0]), pe(batch["poke"][0])], dim=1)[:, :, None, None]
    assert cond.shape == (B, EI + EP, 1, 1) and torch.equal(cond, c_ref)           # [image emb | poke emb], 4-D


def test_make_flow_input_forward_and_augment():
    count = {"poke": 0, "cond": 0}
    scale, shift = torch.tensor([2.0, 3.0]), torch.tensor([0.5, -0.5])
    it = _iface(_Flow(), count, augment_channels=2, scale_augment=scale, shift_augment=shift)
    batch = _batch()
    it.encode_first_stage = lambda X: (torch.ones(B, C - 2, 1, 1), None)
    torch.manual_seed(5)
    flow_input, cond = it.make_flow_input(batch)
    torch.manual_seed(5)
    aug = scale[None, :, None, None] * torch.randn((B, 2, 1, 1)) + shift[None, :, None, None]   # :333-335
    assert flow_input.shape == (B, C, 1, 1)
    assert torch.equal(flow_input[:, :C - 2], torch.ones(B, C - 2, 1, 1)) and torch.equal(flow_input[:, C - 2:], aug)


def test_radial_base_distribution():
    it = _iface(_Flow(), {"poke": 0, "cond": 0}, base_distribution="radial")
    X = _batch()["images"]
    torch.manual_seed(9)
    z = it.draw_latent(X)
    torch.manual_seed(9)
    raw = torch.randn((B, C, 1, 1))
    r = torch.abs(torch.randn(B))                                                   # :325-328
    expect = (torch.nn.functional.normalize(raw.view(B, -1)).T * r).T.view(B, C, 1, 1)
    assert torch.allclose(z, expect) and torch.allclose(z.view(B, -1).norm(dim=1), r, atol=1e-6)


def test_forward_sample_equals_reference_loop_and_encodes_once():
    count = {"poke": 0, "cond": 0}
    flow = _Flow()
    it = _iface(flow, count)
    batch = _batch()
    n = 5
    torch.manual_seed(123)
    vids = it.forward_sample(batch, n_samples=n, n_logged_vids=2, add_first_frame=True)
    assert count == {"poke": 1, "cond": 1}                        # the reference encodes n times (:362)
    assert flow.calls[0] == ("draws", (n, B, C, 1, 1), (B, EI + EP, 1, 1), n)
    # the reference's loop, restated (:360-376)
    ref_flow, ref_it = _Flow(), _iface(_Flow(), {"poke": 0, "cond": 0})
    torch.manual_seed(123)
    X = batch["images"]
    for k in range(n):
        flow_input, cond = ref_it.make_flow_input(batch, reverse=True)
        out = ref_flow(flow_input, cond, reverse=True).squeeze(-1).squeeze(-1)
        vid = ref_it.decode_first_stage(out, X)
        vid = torch.cat([X[:, 0].unsqueeze(1), vid], dim=1)[:2]
        assert vids[k].shape == vid.shape == (2, T, 3, 4, 4)
        assert torch.equal(vids[k], vid), k


def test_forward_density_and_spatial_latents_rejected():
    flow = _Flow()
    it = _iface(flow, {"poke": 0, "cond": 0})
    out, logdet, cond = it.forward_density(_batch(), return_cond=True)
    assert out.shape == (B, C, 1, 1) and logdet.shape == (B,) and cond.shape == (B, EI + EP, 1, 1)
    assert flow.calls[-1] == ("call", (B, C), (B, EI + EP, 1, 1), False)
    it2 = _iface(_Flow(), {"poke": 0, "cond": 0}, min_spatial_size=8)
    with pytest.raises(NotImplementedError):
        it2.sample_latents(_batch(), 2)


def test_flow_ae_interface_draw_order_and_shapes():
    class Dist:
        def __init__(self, m): self.m = m
        def sample(self): return self.m + 1
        def mode(self): return self.m
    enc_calls = []
    ae = types.SimpleNamespace(encode=lambda x: (enc_calls.append(1), Dist(x.mean((2, 3))))[1], decode=lambda z: z * 3)

    class INN:
        def __call__(self, x, reverse=False):
            return x.reshape(x.shape[0], -1, 1, 1) * 2 if reverse else (x * 0.5, torch.zeros(x.shape[0]))
    it = FlowAELatentInterface(INN(), ae)
    batch = {"flow": torch.randn(4, 7, 3, 3)}
    torch.manual_seed(77)
    imgs = it.forward_sample(batch, n_samples=3, n_logged_imgs=2)
    assert len(enc_calls) == 1 and len(imgs) == 3
    torch.manual_seed(77)
    for k in range(3):                                             # fully_connected_ae_INN.py:108-113
        z = torch.randn(4, 7)
        assert torch.equal(imgs[k], ((z * 2).reshape(4, 7, 1, 1) * 3)[:2])
    out, ld = it.forward_density(batch)
    assert torch.equal(out, (batch["flow"].mean((2, 3)) + 1) * 0.5)


# ----------------------------------------------------------------------------- third stage
from ipoke_inn_b200.latent_interface import ThirdStageLatentInterface  # noqa: E402

ZD, CF, BS = 5, 8, 6          # flow-AE z_dim, flow_in_channels, config batch size


class _Dist:
    def __init__(self, m): self.m = m
    def sample(self): return self.m + 1
    def mode(self): return self.m


def _flow_ae():
    calls = []
    ae = types.SimpleNamespace(encode=lambda x: (calls.append("enc"), _Dist(x.mean((2, 3))[:, :ZD]))[1],
                               decode=lambda z: (calls.append("dec"), z * 3)[1])
    return ae, calls


class _UncondINN:
    def __init__(self): self.calls = []

    def __call__(self, x, reverse=False):
        self.calls.append((tuple(x.shape), reverse))
        if reverse:
            return x.reshape(x.shape[0], -1, 1, 1) * 2
        return x.reshape(x.shape[0], -1, 1, 1) * 0.5, torch.zeros(x.shape[0])


def test_third_stage_make_flow_input_pads_with_noise_and_keeps_rng_order():
    ae, _ = _flow_ae()
    it = ThirdStageLatentInterface(_UncondINN(), ae, flow_in_channels=CF, flow_z_dim=ZD, batch_size=BS)
    batch = {"flow": torch.randn(4, 7, 3, 3)}
    torch.manual_seed(21)
    fi = it.make_flow_input(batch)
    torch.manual_seed(21)
    enc = batch["flow"].mean((2, 3))[:, :ZD] + 1
    rest = torch.randn([4, CF - ZD])                                         # :192-196
    assert torch.equal(fi, torch.cat((enc, rest), 1))
    torch.manual_seed(22)
    z = it.make_flow_input(batch, reverse=True)
    torch.manual_seed(22)
    assert torch.equal(z, torch.randn([BS, CF, 1, 1])) and z.shape[0] == BS   # config batch size, not the batch's (:183)
    same = ThirdStageLatentInterface(_UncondINN(), ae, flow_in_channels=ZD, flow_z_dim=ZD, batch_size=BS)
    assert torch.equal(same.make_flow_input(batch), enc)                      # no padding channels (:198-199)
    with pytest.raises(AssertionError):
        ThirdStageLatentInterface(_UncondINN(), ae, flow_in_channels=ZD - 1, flow_z_dim=ZD, batch_size=BS)


def test_third_stage_density_returns_untouched_copy_and_poke_half_of_cond():
    ae, _ = _flow_ae()
    flow = _Flow()
    second = types.SimpleNamespace(forward_density=lambda b, return_cond=False: (
        torch.full((4, CF, 1, 1), 7.0), torch.zeros(4), torch.arange(4 * (EI + EP), dtype=torch.float32).reshape(4, EI + EP, 1, 1)))
    it = ThirdStageLatentInterface(flow, ae, second, flow_in_channels=CF, flow_z_dim=ZD, batch_size=BS, poke_z_dim=EP)
    batch = {"flow": torch.randn(4, 7, 3, 3)}
    torch.manual_seed(5)
    out, logdet, vec, out_hat = it.density_with_second_stage(batch)
    torch.manual_seed(5)
    expect_in = it.make_flow_input(batch)
    assert torch.equal(vec, expect_in) and not vec.requires_grad
    assert flow.calls[-1] == ("call", (4, CF), (4, EP, 1, 1), False)          # cond[:, -poke_z_dim:] (:246)
    assert torch.equal(out, expect_in.reshape(4, CF, 1, 1) * 0.5) and torch.equal(out_hat, torch.full((4, CF, 1, 1), 7.0))


def test_third_stage_forward_sample_equals_reference_loop():
    ae, _ = _flow_ae()
    batch = {"flow": torch.randn(BS, 7, 3, 3)}
    nl, n = 3, 4
    # conditional INN: one reverse_draws call on the shared conditioning
    flow = _Flow()
    it = ThirdStageLatentInterface(flow, ae, flow_in_channels=CF, flow_z_dim=ZD, batch_size=BS, poke_z_dim=EP)
    cond = torch.randn(BS, EP, 1, 1)
    torch.manual_seed(31)
    imgs, residual = it.forward_sample(batch, cond, n_samples=n, n_logged_imgs=nl)
    assert [c[0] for c in flow.calls].count("draws") == 1 and flow.calls[0] == ("draws", (n, nl, CF, 1, 1), (nl, EP, 1, 1), n)
    ref = _Flow()
    torch.manual_seed(31)
    for k in range(n):                                                        # :219-230 restated
        fi = torch.randn([BS, CF, 1, 1])
        out = ref(fi[:nl], cond[:nl], reverse=True)
        o, res = out[:, :ZD], out[:, ZD:]
        assert torch.equal(imgs[k], o.squeeze(-1).squeeze(-1) * 3), k
    assert torch.equal(residual, res) and residual.shape == (nl, CF - ZD, 1, 1)
    # unconditional INN: one stacked reverse call; a given flow_input replaces the single draw
    inn = _UncondINN()
    it2 = ThirdStageLatentInterface(inn, ae, flow_in_channels=CF, flow_z_dim=ZD, batch_size=BS)
    torch.manual_seed(32)
    imgs2, _ = it2.forward_sample(batch, None, n_samples=n, n_logged_imgs=nl)
    assert inn.calls == [((n * nl, CF, 1, 1), True)]
    torch.manual_seed(32)
    for k in range(n):
        fi = torch.randn([BS, CF, 1, 1])
        assert torch.equal(imgs2[k], (fi[:nl] * 2)[:, :ZD].squeeze(-1).squeeze(-1) * 3)
    given = torch.randn(BS, CF, 1, 1)
    imgs3, _ = it2.forward_sample(None, None, 1, 2, given)                    # batch may be None then (:269)
    assert torch.equal(imgs3[0], (given[:2] * 2)[:, :ZD].squeeze(-1).squeeze(-1) * 3)
    with pytest.raises(AssertionError):
        it2.forward_sample(batch, None, 2, 2, given)


def test_second_stage_keypoint_pokes():
    count = {"poke": 0, "cond": 0}
    it = _iface(_Flow(), count)
    batch = _batch()
    kp = torch.randint(0, 5, (B, 2, 4, 4))
    batch["keypoint_poke"] = (kp, "ids")
    _, cond = it.make_flow_input(batch, reverse=True, use_kp_poke=True)       # :276-278
    pe, ce, _, _ = _parts({"poke": 0, "cond": 0})
    c_ref = torch.cat([ce(batch["images"][:, 0]), pe(kp.to(torch.float))], dim=1)[:, :, None, None]
    assert torch.equal(cond, c_ref)
    vids = it.forward_sample(batch, n_samples=2, n_logged_vids=1, use_keypoint_pokes=True)
    assert len(vids) == 2
