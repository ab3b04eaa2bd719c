#!/usr/bin/env python
"""Short driver for ncu captures: a few forward(+reverse) passes of one config."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ipoke_inn_b200 as P
from bench import CONFIGS

ap = argparse.ArgumentParser()
ap.add_argument("--config", default="cfgA")
ap.add_argument("--batch", type=int, default=32768)
ap.add_argument("--precision", default="bf16")
ap.add_argument("--iters", type=int, default=2)
ap.add_argument("--reverse", action="store_true")
a = ap.parse_args()
kind, kw, _, _ = CONFIGS[a.config]
dev = torch.device("cuda:0")
torch.manual_seed(42)
net = (P.SupervisedTransformer(dict(kw), precision=a.precision) if kind == "cond" else
       P.UnsupervisedTransformer3(**kw, precision=a.precision)).to(dev).eval()
C, E = kw["flow_in_channels"], kw.get("flow_embedding_channels", 0)
x = torch.randn(a.batch, C, device=dev)
cond = torch.randn(a.batch, E, 1, 1, device=dev) if kind == "cond" else None
with torch.no_grad():
    f = (lambda inp, r: net(inp, cond, reverse=r)) if kind == "cond" else (lambda inp, r: net(inp, reverse=r))
    z, _ = f(x, False)
    torch.cuda.synchronize()
    for _ in range(a.iters):
        z, _ = f(x, False)
        if a.reverse:
            f(z, True)
    torch.cuda.synchronize()
print("ok", float(z.abs().mean()))
