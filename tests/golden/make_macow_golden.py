#!/usr/bin/env python
"""Golden vectors of the MaCow unit (row N3 of SURVEY.md section 8f) from the LIVE reference (build container only).

Imports ``MaCowUnit`` (models/modules/INN/macow2.py:925-996: MaskedConvFlow A / B, ActNorm2dFlow, MaskedConvFlow C / D,
ActNorm2dFlow) from /root/reference, runs the first forward (data-dependent ActNorm2dFlow init, macow2.py:526-541; the
zero-initialised weight-normed 1x1 convolutions of the MCF blocks, macow_utils.py:230-251), then gives the 1x1
convolutions non-trivial gains / biases (a freshly initialised unit is the identity up to ActNorm) and records forward
(out, logdet), the sequential inverse of ``out`` and of a fresh N(0, 1) tensor as ``tests/golden/<name>.npz``.
Re-run:  python tests/golden/make_macow_golden.py
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
    # name: (in_channels, kernel_size, activation, alpha, batch, H, W)
    "macow_elu": (4, [2, 3], "elu", 1.0, 3, 6, 6),            # config/second_stage.yaml:58, 66 (kernel [2,3], elu)
    "macow_relu": (8, [2, 3], "relu", 1.0, 2, 5, 7),
    "macow_lrelu": (16, [3, 3], "leaky_relu", 0.9, 4, 8, 8),
}


def main():
    sys.dont_write_bytecode = True
    sys.path.insert(0, REF)
    stub = types.ModuleType("opt_einsum")
    stub.contract = torch.einsum
    sys.modules["opt_einsum"] = stub
    warnings.simplefilter("ignore")
    from models.modules.INN.macow2 import MaCowUnit
    for name, (C, ks, act, alpha, B, H, W) in CASES.items():
        torch.manual_seed(42)
        unit = MaCowUnit(C, ks, activation=act, alpha=alpha).eval()
        g = torch.Generator().manual_seed(7)
        x = torch.randn(B, C, H, W, generator=g)
        with torch.no_grad():
            unit(x)                                                   # data-dependent inits
            for k in (1, 2, 3, 4):
                conv = getattr(unit, f"conv{k}").net.conv1x1.conv
                conv.weight_g.copy_(0.1 + 0.3 * torch.rand(conv.weight_g.shape, generator=g))
                conv.bias.copy_(0.1 * torch.randn(conv.bias.shape, generator=g))
            out, logdet = unit(x)
            x_rec = unit(out, reverse=True)
            z_in = torch.randn(B, C, H, W, generator=g)
            x_smp = unit(z_in, reverse=True)
            # the first masked-conv flow alone (order A) and the column flow (order C), for unit tests of one flow
            o1, l1 = unit.conv1(x)
            o3, l3 = unit.conv3(x)
            fp64 = MaCowUnit(C, ks, activation=act, alpha=alpha).double().eval()
            fp64.load_state_dict({k: (v.double() if v.is_floating_point() else v) for k, v in unit.state_dict().items()})
            out64, ld64 = fp64(x.double())
        arrays = {"sd::" + k: v.detach().cpu().numpy() for k, v in unit.state_dict().items()}
        arrays.update(x=x.numpy(), out=out.numpy(), logdet=logdet.numpy(), x_rec=x_rec.numpy(), z_in=z_in.numpy(),
                      x_smp=x_smp.numpy(), mcf1_out=o1.numpy(), mcf1_logdet=l1.numpy(), mcf3_out=o3.numpy(),
                      mcf3_logdet=l3.numpy(), out64=out64.numpy(), logdet64=ld64.numpy(),
                      meta=np.array([C, ks[0], ks[1], B, H, W]), alpha=np.float64(alpha), activation=np.array(act))
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **arrays)
        print(name, "round trip", float((x_rec - x).abs().max()), "logdet", logdet.numpy()[:2])


if __name__ == "__main__":
    main()
