"""GPU test of the host-buffer entry points (ipoke_inn_forward_host / _reverse_host: H2D copy, kernels, D2H copy enqueued
on one stream; with pinned buffers the whole request replays as one CUDA graph): results equal the device-pointer
calls, new buffer contents are picked up by the replayed graph, pageable buffers work (copied outside the graph)."""
import pytest
import torch

import ipoke_inn_b200 as P

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("pinned", [True, False])
@pytest.mark.parametrize("kind", ["cond", "uncond"])
def test_host_entry_points_match_device_calls(kind, pinned):
    dev = torch.device("cuda:0")
    torch.manual_seed(42)
    if kind == "cond":
        net = P.SupervisedTransformer(dict(flow_in_channels=64, flow_mid_channels_factor=2, flow_hidden_depth=2, n_flows=4,
                                           flow_conditioning_option="none", flow_embedding_channels=32), precision="bf16").to(dev)
        C, E = 64, 32
    else:
        net = P.UnsupervisedTransformer3(flow_in_channels=64, flow_mid_channels=128, flow_hidden_depth=2, n_flows=4,
                                         precision="bf16").to(dev)
        C, E = 64, 0
    B = 200
    eng = net.flow.engine
    mk = (lambda *s: torch.empty(*s).pin_memory()) if pinned else (lambda *s: torch.empty(*s))
    xh, zh, ldh, xoh = mk(B, C), mk(B, C), mk(B), mk(B, C)
    ch = mk(B, E) if E else None
    g = torch.Generator().manual_seed(0)
    with torch.no_grad():
        x0 = torch.randn(B, C, generator=g)
        c0 = torch.randn(B, E, 1, 1, generator=g) if E else None
        net(x0.to(dev), c0.to(dev)) if E else net(x0.to(dev))                  # ActNorm init, plan, pack
        for rep in range(4):                                                    # repeated calls: graph replay
            xh.copy_(torch.randn(B, C, generator=g))
            if E:
                ch.copy_(torch.randn(B, E, generator=g))
            eng.forward_host(xh, ch, zh, ldh)
            torch.cuda.synchronize()
            cd = ch.to(dev).reshape(B, E, 1, 1) if E else None
            z_ref, ld_ref = net(xh.to(dev), cd) if E else net(xh.to(dev))
            assert torch.equal(zh, z_ref.reshape(B, C).cpu()), rep
            assert torch.allclose(ldh, ld_ref.cpu(), rtol=1e-6, atol=1e-5), rep
            eng.reverse_host(zh, ch, xoh)
            torch.cuda.synchronize()
            x_ref = net(zh.to(dev), cd, reverse=True) if E else net(zh.to(dev), reverse=True)
            assert torch.equal(xoh, x_ref.reshape(B, C).cpu()), rep
            assert float((xoh - xh).abs().max()) < 3e-2
