#!/usr/bin/env python
"""Small workload for compute-sanitizer (tools/sanitize.sh): every kernel family once at tiny sizes -- forward /
reverse (folded and unfolded programs, tail clusters), training forward + fused backward (dgrad, wgrad, EPI_DU, ActNorm
gradients), FlowLoss, fused Adam, multi-draw sampling, host-buffer entry points, the sharded-optimizer kernels, and the
MaCow conv flow (unit chains both directions, coupling GEMMs in both variants with split-K taps, glue kernels, the
conditional product)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ipoke_inn_b200 as P  # noqa: E402
from test_host_logic import build  # noqa: E402

dev = torch.device("cuda:0")
os.environ.setdefault("IPOKE_INN_GRAPH", "0")           # plain launches: the sanitizer attributes errors per kernel
which = sys.argv[1:] or ["infer", "train", "draws", "host", "sharded", "macow"]
g = torch.Generator().manual_seed(1)

for prec in ("fp32", "bf16"):
    if "infer" in which:
        for name, B in (("cfgS_cond", 10), ("cfgS_cond", 130), ("uncond_small", 33), ("cfgB_mini", 600)):
            net = build(name, precision=prec).to(dev)
            kind = "cond" if hasattr(net, "emb_channels") else "uncond"
            C = net.flow.engine.C
            x = torch.randn(B, C, generator=g).to(dev)
            c = torch.randn(B, net.flow.engine.E, 1, 1, generator=g).to(dev) if net.flow.engine.E else None
            with torch.no_grad():
                z, ld = net(x, c) if c is not None else net(x)
                xr = net(z, c, reverse=True) if c is not None else net(z, reverse=True)
            assert float((xr.reshape(B, C) - x).abs().max()) < 5e-2
    if "train" in which:
        for name, B in (("cfgS_cond", 48), ("cfgA_mini", 130)):
            net = build(name, precision=prec).to(dev)
            C, E = net.flow.engine.C, net.flow.engine.E
            x = torch.randn(B, C, generator=g).to(dev).requires_grad_(True)
            c = torch.randn(B, E, 1, 1, generator=g).to(dev).requires_grad_(True)
            opt = P.FusedAdam(net.parameters(), lr=1e-3, amsgrad=True, weight_decay=1e-5)
            for _ in range(2):
                opt.zero_grad(set_to_none=True)
                z, ld = net(x, c)
                loss, _ = P.FlowLoss()(z, ld)
                loss.backward()
                opt.step()
    if "draws" in which:
        net = build("cfgA_mini", precision=prec).to(dev)
        c = torch.randn(128, 128, 1, 1, generator=g).to(dev)
        zs = torch.randn(3, 128, 128, generator=g).to(dev)
        with torch.no_grad():
            net(zs[0], c)
            xs = net.flow.reverse_draws(zs, c)
            zz, ll = net.flow.forward_draws(xs, c)
    if "host" in which:
        net = build("cfgS_cond", precision=prec).to(dev)
        B = 20
        x = torch.randn(B, 8, generator=g)
        c = torch.randn(B, 16, generator=g)
        with torch.no_grad():
            net(x.to(dev), c.reshape(B, 16, 1, 1).to(dev))
        eng = net.flow.engine
        xh, ch = x.pin_memory(), c.pin_memory()
        zh, lh, xo = torch.empty(B, 8).pin_memory(), torch.empty(B).pin_memory(), torch.empty(B, 8).pin_memory()
        eng.forward_host(xh, ch, zh, lh)
        torch.cuda.synchronize()
        eng.reverse_host(zh, ch, xo)
        torch.cuda.synchronize()
    if "sharded" in which:
        net = build("cfgA_mini", precision=prec).to(dev)
        x = torch.randn(64, 128, generator=g).to(dev)
        c = torch.randn(64, 128, 1, 1, generator=g).to(dev)
        with torch.no_grad():
            net(x, c)
        opt = P.ShardedAdam(net, lr=1e-3, amsgrad=True, weight_decay=1e-5)
        for _ in range(2):
            opt.zero_grad(set_to_none=True)
            z, ld = net(x, c)
            loss, _ = P.FlowLoss()(z, ld)
            loss.backward()
            opt.step()
torch.cuda.synchronize()
print("sanitize_run: done", which)

if "macow" in which:
    base = dict(num_steps=[1, 1], flow_in_channels=8, flow_mid_channels=160, factor=4, transform="affine",
                prior_transform="affine", kernel_size=[2, 3], coupling_type="conv", activation="elu")
    for async_gemm in ("1", "0"):
        os.environ["IPOKE_MACOW_GEMM_ASYNC"] = async_gemm
        for cfg, cls in ((base, P.UnsupervisedMaCowTransformer3),
                         (dict(base, h_channels=12, flow_attn_heads=4, cond_conv_hidden_channels=16),
                          P.SupervisedMacowTransformer)):
            torch.manual_seed(2)
            net = cls(cfg).to(dev).eval()
            x = torch.randn(5, 8, 6, 7, generator=g).to(dev)                    # ragged tiles, non-square image
            h = (torch.randn(5, 12, 6, 7, generator=g).to(dev),) if "h_channels" in cfg else ()
            with torch.no_grad():
                z, ld = net(x, *h)                                                  # first call: data-dependent inits
                for k, v in net.state_dict().items():
                    if k.endswith("weight_g"):
                        v.fill_(0.05)
                z, ld = net(x, *h)
                back = net(z, *h, reverse=True)
            assert float((back - x).abs().max()) < 1e-3
