#!/usr/bin/env python
"""Golden vectors of the full MaCow flow (row N3 of SURVEY.md section 8f) from the LIVE reference (build container only).

Imports ``UnsupervisedMaCowTransformer3`` (models/modules/INN/INN.py:365-390 -> ``MultiScaleInternal(MaCowStep, ...)``,
macow2.py:821-918, 999-1117) from /root/reference, runs the first forward (data-dependent ActNorm2dFlow and
Conv2dWeightNorm initialisations), then gives every zero-initialised weight-normed convolution (the 1x1 of the MCF
blocks, the 3x3 of the NICE coupling nets) non-trivial gains / biases -- a freshly initialised flow is the identity up
to ActNorm and permutations -- and records forward (out, logdet), the inverse of ``out`` and of a fresh N(0, 1) tensor,
the float64 run, and the outputs of single components (one NICE2d per split type / order, one MaCowStep, one
MultiScalePrior) as ``tests/golden/<name>.npz``.  A second file holds a freshly constructed model's state after its first
forward (the data-dependent initialisation) for the same seed.
Re-run:  python tests/golden/make_macow_stack_golden.py
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
    # name: (config, batch, H, W)
    "macow_stack_a": (dict(num_steps=[2, 1], flow_in_channels=8, flow_mid_channels=32, factor=4, transform="affine",
                           prior_transform="affine", kernel_size=[2, 3], coupling_type="conv", activation="elu"),
                      3, 6, 6),
    # odd channel counts at the later levels (skip split falls back to continuous), leaky_relu, non-square image
    "macow_stack_b": (dict(num_steps=[1, 1, 1], flow_in_channels=12, flow_mid_channels=24, factor=4, transform="affine",
                           prior_transform="affine", kernel_size=[2, 3], coupling_type="conv",
                           activation="leaky_relu"), 2, 4, 5),
    # conditional: SupervisedMacowTransformer (INN.py:447-482) as models/second_stage_video.py:122 builds it, condition_nice
    # false -- the conditioning h [B, h_channels, H, W] enters the 1x1 convolution of every masked flow
    "macow_stack_c": (dict(num_steps=[1, 1], flow_in_channels=8, flow_mid_channels=24, factor=4, transform="affine",
                           prior_transform="affine", kernel_size=[2, 3], coupling_type="conv", activation="elu",
                           h_channels=6, flow_attn_heads=4, cond_conv_hidden_channels=16), 3, 5, 6),
}


def main():
    sys.dont_write_bytecode = True
    sys.path.insert(0, REF)
    stub = types.ModuleType("opt_einsum")
    stub.contract = torch.einsum
    sys.modules["opt_einsum"] = stub
    warnings.simplefilter("ignore")
    from models.modules.INN.INN import SupervisedMacowTransformer, UnsupervisedMaCowTransformer3
    for name, (cfg, B, H, W) in CASES.items():
        cond = "h_channels" in cfg
        cls = SupervisedMacowTransformer if cond else UnsupervisedMaCowTransformer3
        torch.manual_seed(42)
        net = cls(cfg).eval()
        C = cfg["flow_in_channels"]
        g = torch.Generator().manual_seed(11)
        x = torch.randn(B, C, H, W, generator=g)
        if cond:                                                          # the wrapper's forward takes the conditioning second
            hcond = torch.randn(B, cfg["h_channels"], H, W, generator=g)
            plain = net
            class _Bound:                                                 # same call shapes as the unconditional wrapper below
                flow = plain.flow
                def __call__(self, t, reverse=False):
                    return plain(t, hcond, reverse=reverse)
            call = _Bound()
        else:
            call = net
        with torch.no_grad():
            init_out, init_ld = call(x)                                   # data-dependent inits
            init_sd = {k: v.clone() for k, v in net.state_dict().items()}
            for k, p in net.state_dict().items():                         # state_dict tensors alias the parameters
                if k.endswith("weight_g"):
                    p.copy_(0.1 + 0.3 * torch.rand(p.shape, generator=g))
                elif k.endswith("conv.bias"):
                    p.copy_(0.1 * torch.randn(p.shape, generator=g))
            out, logdet = call(x)
            x_rec = call(out, reverse=True)
            z_in = torch.randn(B, C, H, W, generator=g)
            x_smp = call(z_in, reverse=True)
            step = net.flow.layers[0][0]
            comp = {}
            hk = dict(h=hcond) if cond else {}
            for cn in ("coupling1_up", "coupling1_dn", "coupling2_up", "coupling2_dn"):
                o, l = getattr(step, cn)(x)
                comp[cn + "_out"], comp[cn + "_logdet"] = o.numpy(), l.numpy()
                comp[cn + "_rev"] = getattr(step, cn)(z_in, reverse=True).numpy()
            o, l = step(x, **hk)
            comp["step_out"], comp["step_logdet"] = o.numpy(), l.numpy()
            comp["step_rev"] = step(z_in, reverse=True, **hk).numpy()
            o, l = net.flow.priors[0](x)
            comp["prior_out"], comp["prior_logdet"] = o.numpy(), l.numpy()
            comp["prior_rev"] = net.flow.priors[0](z_in, reverse=True).numpy()
            if cond:
                unit = step.units1[0]
                o, l = unit(x, h=hcond)
                comp["unit_out"], comp["unit_logdet"] = o.numpy(), l.numpy()
                comp["unit_rev"] = unit(z_in, h=hcond, reverse=True).numpy()
                comp["h"] = hcond.numpy()
            fp64 = cls(cfg).double().eval()
            fp64.load_state_dict({k: (v.double() if v.is_floating_point() else v) for k, v in net.state_dict().items()})
            if cond:
                out64, ld64 = fp64(x.double(), hcond.double())
                smp64 = fp64(z_in.double(), hcond.double(), reverse=True)
            else:
                out64, ld64 = fp64(x.double())
                smp64 = fp64(z_in.double(), reverse=True)
        arrays = {"sd::" + k: v.detach().cpu().numpy() for k, v in net.state_dict().items()}
        arrays.update({"init::" + k: v.cpu().numpy() for k, v in init_sd.items()
                       if k.endswith(("log_scale", "bias", "weight_g")) and "units" not in k})
        arrays.update(comp)
        arrays.update(x=x.numpy(), out=out.numpy(), logdet=logdet.numpy(), x_rec=x_rec.numpy(), z_in=z_in.numpy(),
                      x_smp=x_smp.numpy(), out64=out64.numpy(), logdet64=ld64.numpy(), x_smp64=smp64.numpy(),
                      init_out=init_out.numpy(), init_logdet=init_ld.numpy(),
                      num_steps=np.array(cfg["num_steps"]), meta=np.array([C, cfg["flow_mid_channels"], cfg["factor"],
                                                                           cfg["kernel_size"][0], cfg["kernel_size"][1]]),
                      activation=np.array(cfg["activation"]), h_channels=np.array(cfg.get("h_channels", 0)))
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **arrays)
        print(name, "->", path, os.path.getsize(path) // 1024, "KiB;", len(net.state_dict()), "state_dict entries")


if __name__ == "__main__":
    main()
