#!/usr/bin/env python
"""Generate golden input/output vectors from the LIVE reference (build container only).

Imports the reference INN modules from /root/reference (read-only; with the
``opt_einsum`` stub of SURVEY.md's appendix), runs them on seeded inputs and stores
weights (state_dict after ActNorm data init), inputs, forward / reverse outputs, the
per-block lists and autograd gradients as ``tests/golden/<name>.npz``.

The reference cannot travel to the GPU box, so these fixtures (and this script) are
committed.  Re-run:  python tests/golden/make_golden.py
"""
import os
import sys
import types
import warnings

import numpy as np
import torch

REF = os.environ.get("IPOKE_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))

CASES = {
    # name: (kind, ctor kwargs, batch, with_grads)
    # cfg S of SURVEY.md section 8(d): config/INN_test.yaml:17 flow_in_channels=8
    "cfgS_cond": ("cond", dict(flow_in_channels=8, flow_mid_channels_factor=8, flow_hidden_depth=2,
                               n_flows=8, flow_conditioning_option="none",
                               flow_embedding_channels=16), 10, True),
    "uncond_small": ("uncond", dict(flow_in_channels=16, flow_mid_channels=64, flow_hidden_depth=2,
                                    n_flows=4), 12, True),
    # cfg A / cfg B widths (C, E, depth) with a narrow hidden layer so the fixture stays small
    "cfgA_mini": ("cond", dict(flow_in_channels=128, flow_mid_channels_factor=1, flow_hidden_depth=3,
                               n_flows=2, flow_conditioning_option="none",
                               flow_embedding_channels=128), 32, False),
    "cfgB_mini": ("uncond", dict(flow_in_channels=512, flow_mid_channels=64, flow_hidden_depth=2,
                                 n_flows=2), 16, False),
}


def import_reference():
    sys.path.insert(0, REF)
    sys.dont_write_bytecode = True
    stub = types.ModuleType("opt_einsum")
    stub.contract = torch.einsum                      # modules.py:9 imports it; unused on this path
    sys.modules["opt_einsum"] = stub
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        from models.modules.INN.INN import SupervisedTransformer, UnsupervisedTransformer3
        from models.modules.INN.loss import FlowLoss
    return SupervisedTransformer, UnsupervisedTransformer3, FlowLoss


def build(kind, kw, ST, UT3):
    torch.manual_seed(42)                             # INN_FCAE.py:27
    if kind == "cond":
        return ST(kw)
    return UT3(**kw)


def main():
    ST, UT3, FlowLoss = import_reference()
    for name, (kind, kw, B, with_grads) in CASES.items():
        net = build(kind, kw, ST, UT3)
        C = kw["flow_in_channels"]
        E = kw.get("flow_embedding_channels", 0)
        g = torch.Generator().manual_seed(1234)
        x = torch.randn(B, C, generator=g)
        cond = torch.randn(B, E, 1, 1, generator=g) if kind == "cond" else None
        z_in = torch.randn(B, C, 1, 1, generator=g)
        args = (x, cond) if kind == "cond" else (x,)
        out = {}
        # perms are buffers drawn at construction under seed 42 (flow_blocks.py:318)
        with torch.no_grad():
            z0, ld0 = net(*args)                      # first call: ActNorm data init
        sd = {k: v.detach().clone() for k, v in net.state_dict().items()}
        with torch.no_grad():
            z, ld = net(*args)
            outs, lds = net(*args, train=True)
            rargs = (z, cond) if kind == "cond" else (z,)
            x_rec = net(*rargs, reverse=True)
            sargs = (z_in, cond) if kind == "cond" else (z_in,)
            x_smp = net(*sargs, reverse=True)
        assert torch.equal(z0, z) and torch.equal(ld0, ld)
        out.update({"sd::" + k: v.numpy() for k, v in sd.items()})
        out.update(x=x.numpy(), z=z.numpy(), logdet=ld.numpy(), x_rec=x_rec.numpy(),
                   z_in=z_in.numpy(), x_smp=x_smp.numpy(),
                   last_outs=np.stack([o.reshape(B, C).numpy() for o in outs]),
                   last_logdets=np.stack([l.numpy() for l in lds]))
        if cond is not None:
            out["cond"] = cond.numpy()
        # fp64 truth (ranks fp32 / split-bf16 / bf16 variants)
        net64 = build(kind, kw, ST, UT3).double()
        net64.load_state_dict({k: (v.double() if v.is_floating_point() else v) for k, v in sd.items()})
        with torch.no_grad():
            a64 = tuple(a.double() for a in args)
            z64, ld64 = net64(*a64)
        out.update(z64=z64.numpy(), logdet64=ld64.numpy())
        if with_grads:
            xg = x.clone().requires_grad_(True)
            cg = cond.clone().requires_grad_(True) if cond is not None else None
            net.zero_grad()
            zz, ll = net(*((xg, cg) if cg is not None else (xg,)))
            torch.manual_seed(7)
            loss, log = FlowLoss()(zz, ll)            # loss.py:17-42
            loss.backward()
            out["loss"] = loss.detach().numpy()
            out["nll_loss"] = log["nll_loss"].numpy()
            out["nlogdet_loss"] = log["nlogdet_loss"].numpy()
            out["dx"] = xg.grad.numpy()
            if cg is not None:
                out["dcond"] = cg.grad.numpy()
            for k, p in net.named_parameters():
                out["grad::" + k] = p.grad.numpy()
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **out)
        print(f"{name}: wrote {path} ({os.path.getsize(path)/1e6:.2f} MB), "
              f"params={sum(p.numel() for p in net.parameters())}, "
              f"round-trip max|x_rec-x|={float((x_rec.reshape(B, C) - x).abs().max()):.2e}")


if __name__ == "__main__":
    main()
