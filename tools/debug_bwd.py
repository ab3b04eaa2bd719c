#!/usr/bin/env python
"""Per-tensor gradient error table of the fused backward vs the fp64 oracle (GPU box)."""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ipoke_inn_b200 as P
from oracle import inn_oracle as O
from conftest import load_golden, max_rel
from test_host_logic import CFG, build

ap = argparse.ArgumentParser()
ap.add_argument("--name", default="cfgS_cond")
ap.add_argument("--precision", default="bf16")
ap.add_argument("--batch", type=int, default=10)
ap.add_argument("--pdl", type=int, default=1)
ap.add_argument("--thresh", type=float, default=0.02)
a = ap.parse_args()
os.environ["IPOKE_INN_PDL"] = str(a.pdl)
dev = torch.device("cuda:0")
sd_ref, d = load_golden(a.name)
net = build(a.name, precision=a.precision)
net.load_state_dict({k: torch.from_numpy(v) for k, v in sd_ref.items()}, strict=True)
net = net.to(dev)
kind, kw = CFG[a.name]
B = a.batch
C, E = kw["flow_in_channels"], kw.get("flow_embedding_channels", 0)
g = torch.Generator().manual_seed(5)
x = torch.randn(B, C, generator=g)
cond = torch.randn(B, E, 1, 1, generator=g) if kind == "cond" else None
dz = torch.randn(B, C, generator=g)
dld = torch.randn(B, generator=g)
xg = x.to(dev).requires_grad_(True)
cg = cond.to(dev).requires_grad_(True) if cond is not None else None
z, ld = net(xg, cg) if cg is not None else net(xg)
((z.reshape(B, -1) * dz.to(dev)).sum() + (ld * dld.to(dev)).sum()).backward()
torch.cuda.synchronize()
with O.gemm_rounding(a.precision if a.precision == "bf16" else None):
    grads, dx_ref, dcond_ref = O.flow_backward(sd_ref, x.numpy(), None if cond is None else cond.numpy().reshape(B, -1),
                                               dz.numpy(), dld.numpy())
print(f"{a.name} {a.precision} B={B} pdl={a.pdl}: dx err {max_rel(xg.grad.cpu().numpy(), dx_ref):.2e}",
      f"dcond err {max_rel(cg.grad.cpu().numpy().reshape(B, -1), dcond_ref):.2e}" if cg is not None else "")
n_bad = 0
for k, p in net.named_parameters():
    e = max_rel(p.grad.cpu().numpy(), np.asarray(grads[k]).reshape(p.shape))
    if e > a.thresh:
        n_bad += 1
        print(f"  {k:58s} {e:.3e}  |ref|max {np.abs(grads[k]).max():.3e}")
print("bad tensors:", n_bad)
