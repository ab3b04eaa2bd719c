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
