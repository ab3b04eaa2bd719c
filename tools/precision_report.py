#!/usr/bin/env python
"""Measured numerical error of both precisions against the fp64 oracle: z, logdet, reverse reconstruction, as
max-abs error / max-abs of the tensor ("max_rel", the tests' yardstick) and as rel-L2, at the fixture sizes, at the
benchmark shapes (row subsets of large batches: rows are independent) and with every Linear weight scaled by 3
("trained-scale": larger pre-activations, saturated tanh, stronger error amplification than at random init).
Output is committed under profiles/ and quoted in DESIGN.md / BASELINE.md."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ipoke_inn_b200 as P  # noqa: E402
from oracle import inn_oracle as O  # noqa: E402
from conftest import max_rel, rel_l2  # noqa: E402
from test_host_logic import CFG, FULL_CFG  # noqa: E402

dev = torch.device("cuda:0")


def case(name, B, precision, rows=256, wscale=1.0, seed=1234):
    kind, kw = (CFG.get(name) or FULL_CFG[name])
    torch.manual_seed(42)
    net = P.SupervisedTransformer(dict(kw), precision=precision) if kind == "cond" else \
        P.UnsupervisedTransformer3(**kw, precision=precision)
    if wscale != 1.0:
        with torch.no_grad():
            for k, p in net.named_parameters():
                if ".main." in k and k.endswith(".weight"):
                    p.mul_(wscale)
    C, E = kw["flow_in_channels"], kw.get("flow_embedding_channels", 0)
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, C, generator=g)
    cond = torch.randn(B, E, 1, 1, generator=g) if kind == "cond" else None
    sd = O.state_dict_to_numpy(net.state_dict())
    n0 = min(B, 256)
    c2 = None if cond is None else cond.numpy().reshape(B, -1)
    O.flow_forward_with_actnorm_init(sd, x[:n0].numpy(), None if c2 is None else c2[:n0])
    net.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in sd.items()})
    net = net.to(dev)
    with torch.no_grad():
        xd, cd = x.to(dev), (cond.to(dev) if cond is not None else None)
        z, ld = net(xd, cd) if cd is not None else net(xd)
        sel = np.linspace(0, B - 1, min(rows, B)).astype(np.int64)
        xs, cs = x[sel].numpy().astype(np.float64), (None if c2 is None else c2[sel].astype(np.float64))
        z_ref, ld_ref = O.flow_forward(sd, xs, cs, dtype=np.float64)
        # the reference's own arithmetic (plain fp32, numpy) against the same fp64 truth: its rounding noise
        z32, ld32 = O.flow_forward(sd, xs.astype(np.float32), None if cs is None else cs.astype(np.float32), dtype=np.float32)
        zr = torch.from_numpy(z_ref.astype(np.float32))
        zin = torch.zeros(B, C)
        zin[sel] = zr
        xr = net(zin.to(dev), cd, reverse=True) if cd is not None else net(zin.to(dev), reverse=True)
        x_ref = O.flow_reverse(sd, zr.numpy().astype(np.float64), cs, dtype=np.float64)
    zg = z.cpu().numpy().reshape(B, C)[sel]
    lg = ld.cpu().numpy()[sel]
    xg = xr.cpu().numpy().reshape(B, C)[sel]
    out = dict(z=(max_rel(zg, z_ref), rel_l2(zg, z_ref)), logdet=(max_rel(lg, ld_ref), rel_l2(lg, ld_ref)),
               logdet_abs=float(np.max(np.abs(lg - ld_ref))), rev=(max_rel(xg, x_ref), rel_l2(xg, x_ref)),
               ld_scale=float(np.max(np.abs(ld_ref))),
               ref32=(max_rel(z32, z_ref), max_rel(ld32, ld_ref)) if z32 is not None else (float("nan"), float("nan")))
    del net
    torch.cuda.empty_cache()
    return out


if __name__ == "__main__":
    cases = [("cfgS_cond", 256, 1.0), ("uncond_small", 256, 1.0), ("cfgA_mini", 256, 1.0), ("cfgB_mini", 256, 1.0),
             ("cfgA", 64, 1.0), ("cfgA", 4096, 1.0), ("cfgA", 32768, 1.0), ("cfgB", 256, 1.0), ("cfgB", 4096, 1.0),
             ("cfgB", 32768, 1.0), ("cfgA_mini", 256, 2.0), ("cfgA", 64, 2.0), ("cfgB", 256, 2.0),
             ("cfgA_mini", 256, 3.0), ("cfgA", 64, 3.0), ("cfgA", 4096, 3.0), ("cfgB", 256, 3.0)]
    if len(sys.argv) > 1 and sys.argv[1] == "scaled":
        cases = [c for c in cases if c[2] != 1.0]
    print("%-13s %6s %5s %5s | %-21s | %-21s %9s | %-21s | %s" % ("config", "batch", "wx", "prec", "z max_rel / rel_l2",
                                                                   "logdet max_rel/rel_l2", "abs", "reverse max_rel/rel_l2",
                                                                   "plain-fp32 numpy vs fp64: z, logdet max_rel"))
    for name, B, ws in cases:
        for prec in ("fp32", "bf16"):
            r = case(name, B, prec, wscale=ws)
            print("%-13s %6d %5.1f %5s | %9.2e / %9.2e | %9.2e / %9.2e %9.2e | %9.2e / %9.2e | %9.2e %9.2e" % (
                name, B, ws, prec, *r["z"], *r["logdet"], r["logdet_abs"], *r["rev"], *r["ref32"]), flush=True)
