"""GPU test of the latent interface with the real fused flow: forward_sample (conditioning encoded once, all draws in one
ipoke_inn_reverse_draws call) must reproduce the reference's loop -- one make_flow_input + flow(.., reverse=True) per
draw, models/second_stage_video_fc.py:360-376 -- under the same seed, and forward_density must equal a direct call."""
import pytest
import torch

import ipoke_inn_b200 as P
from ipoke_inn_b200.latent_interface import SecondStageLatentInterface
from conftest import max_rel

pytestmark = pytest.mark.gpu

CFG = dict(flow_in_channels=128, flow_mid_channels_factor=2, flow_hidden_depth=3, n_flows=4,
           flow_conditioning_option="none", flow_embedding_channels=128)


@pytest.mark.parametrize("precision,B", [("fp32", 20), ("bf16", 20), ("fp32", 128)])
def test_forward_sample_equals_reference_loop(precision, B):
    dev = torch.device("cuda:0")
    torch.manual_seed(42)
    flow = P.SupervisedTransformer(dict(CFG), precision=precision).to(dev)
    wp, wi = torch.randn(2, 64, device=dev), torch.randn(3, 64, device=dev)
    wm = torch.randn(3 * 4, 128, device=dev)
    poke_enc = lambda p: p.mean((2, 3)) @ wp
    cond_enc = lambda x0: x0.mean((2, 3)) @ wi
    enc_fs = lambda X: ((X.mean((3, 4)).reshape(X.shape[0], -1) @ wm)[:, :, None, None], None)
    dec_fs = lambda m, X: m[:, None, :3, None, None].expand(-1, X.shape[1] - 1, -1, 4, 4)
    it = SecondStageLatentInterface(flow, poke_enc, cond_enc, enc_fs, dec_fs, flow_in_channels=128)
    batch = {"images": torch.randn(B, 4, 3, 4, 4, device=dev), "poke": torch.randn(B, 2, 4, 4, device=dev)}
    out, logdet, cond = it.forward_density(batch, return_cond=True)          # also the ActNorm data init
    with torch.no_grad():
        z_direct, ld_direct = flow(enc_fs(batch["images"])[0], cond)
    # forward_density runs under autograd (training forward, per-block outputs kept), the direct call under no_grad
    # (block boundaries folded into the coupling tails): same arithmetic, different fp32 log-det summation order
    assert max_rel(out.detach().cpu().numpy(), z_direct.cpu().numpy()) < 1e-5
    assert max_rel(logdet.detach().cpu().numpy(), ld_direct.cpu().numpy()) < 1e-5
    n = 5
    torch.manual_seed(7)
    motion = it.sample_latents(batch, n)
    assert motion.shape == (n, B, 128)
    torch.manual_seed(7)
    with torch.no_grad():
        for k in range(n):                                                   # the reference's loop
            flow_input, c = it.make_flow_input(batch, reverse=True)
            ref = flow(flow_input, c, reverse=True).squeeze(-1).squeeze(-1)
            assert max_rel(motion[k].cpu().numpy(), ref.cpu().numpy()) < (2e-5 if precision == "fp32" else 1e-2)
    torch.manual_seed(7)
    vids = it.forward_sample(batch, n_samples=n, n_logged_vids=3, add_first_frame=True)
    assert len(vids) == n and vids[0].shape == (3, 4, 3, 4, 4) and vids[0].device.type == "cpu"


def test_third_stage_interface_with_the_real_flows():
    """ThirdStageLatentInterface (third_stage_video_fc_conditional.py:180-255; unconditional :166-231) around the fused
    INNs: forward_sample in one reverse call == the reference's per-draw loop under the same seed; forward_density leaves
    its input untouched and feeds the INN the poke half of the second stage's conditioning."""
    import types
    from ipoke_inn_b200.latent_interface import ThirdStageLatentInterface
    dev = torch.device("cuda:0")
    torch.manual_seed(42)
    cfg = dict(CFG, flow_embedding_channels=64)
    inn_c = P.SupervisedTransformer(cfg).to(dev)
    inn_u = P.UnsupervisedTransformer3(flow_in_channels=128, flow_mid_channels=256, flow_hidden_depth=2, n_flows=4).to(dev)
    wa = torch.randn(6, 96, device=dev)

    class Dist:
        def __init__(self, m): self.m = m
        def sample(self): return self.m
    ae = types.SimpleNamespace(encode=lambda X: Dist(X.mean((2, 3)) @ wa), decode=lambda z: z[:, :, None, None] * 2)
    BS, nl, n = 24, 10, 3
    batch = {"flow": torch.randn(BS, 6, 8, 8, device=dev)}
    cond2 = torch.randn(BS, 128, 1, 1, device=dev)                           # second stage: [image emb 64 | poke emb 64]
    second = types.SimpleNamespace(forward_density=lambda b, return_cond=False: (
        (torch.zeros(BS, 128, 1, 1, device=dev), torch.zeros(BS, device=dev), cond2) if return_cond else
        (torch.zeros(BS, 128, 1, 1, device=dev), torch.zeros(BS, device=dev))))
    for inn, pz in ((inn_c, 64), (inn_u, 0)):
        it = ThirdStageLatentInterface(inn, ae, second, flow_in_channels=128, flow_z_dim=96, batch_size=BS, poke_z_dim=pz)
        torch.manual_seed(3)
        out, logdet, vec, out_hat = it.density_with_second_stage(batch)       # ActNorm data init happens here
        torch.manual_seed(3)
        assert torch.equal(vec, it.make_flow_input(batch)) and out.shape == (BS, 128, 1, 1) and logdet.shape == (BS,)
        cond = it.poke_encoding(cond2) if pz else None
        torch.manual_seed(9)
        imgs, residual = it.forward_sample(batch, cond, n_samples=n, n_logged_imgs=nl)
        assert len(imgs) == n and imgs[0].shape == (nl, 96, 1, 1) and residual.shape == (nl, 32, 1, 1)
        torch.manual_seed(9)
        with torch.no_grad():
            for k in range(n):                                               # the reference's loop (:219-230)
                fi = it.make_flow_input(batch, reverse=True)
                o = inn(fi[:nl], cond[:nl], reverse=True) if pz else inn(fi[:nl], reverse=True)
                ref = ae.decode(o[:, :96].squeeze(-1).squeeze(-1))
                assert max_rel(imgs[k].cpu().numpy(), ref.cpu().numpy()) < 2e-5, (pz, k)
