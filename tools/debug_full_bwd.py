#!/usr/bin/env python
"""Per-tensor gradient error of the fused backward against the fp64 oracle at a full benchmark shape."""
import os, sys
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from oracle import inn_oracle as O
import test_gpu_backward as T
from conftest import max_rel, rel_l2

name, B = sys.argv[1], int(sys.argv[2])
margin = float(sys.argv[3]) if len(sys.argv) > 3 else T.FULL_KINK_MARGIN
import time; t0=time.time()
net, sd_ref, x, cond, dz, dld = T._full_case(name, B, "fp32", margin=margin)
print("case built in %.1f s" % (time.time()-t0))
dev = torch.device("cuda:0")
xg = x.to(dev).requires_grad_(True)
cg = cond.to(dev).requires_grad_(True) if cond is not None else None
z, ld = net(xg, cg) if cg is not None else net(xg)
((z.reshape(B, -1) * dz.to(dev)).sum() + (ld * dld.to(dev)).sum()).backward()
grads, dx_ref, dcond_ref = O.flow_backward(sd_ref, x.numpy(), None if cond is None else cond.numpy().reshape(B, -1),
                                           dz.numpy(), dld.numpy())
zr, ldr = O.flow_forward(sd_ref, x.numpy().astype(np.float64), None if cond is None else cond.numpy().reshape(B, -1).astype(np.float64))
print("forward z max_rel", max_rel(z.detach().cpu().numpy().reshape(B, -1), zr), "logdet", max_rel(ld.detach().cpu().numpy(), ldr))
print("dx max_rel", max_rel(xg.grad.cpu().numpy(), dx_ref), "rel_l2", rel_l2(xg.grad.cpu().numpy(), dx_ref))
rows = np.max(np.abs(xg.grad.cpu().numpy() - dx_ref), axis=1) / np.max(np.abs(dx_ref))
print("dx per-row max err (sorted, top 8):", np.sort(rows)[-8:], "rows", np.argsort(rows)[-8:])
if cg is not None:
    print("dcond max_rel", max_rel(cg.grad.cpu().numpy().reshape(B, -1), dcond_ref))
errs = []
for k, p in net.named_parameters():
    g = p.grad.cpu().numpy()
    r = np.asarray(grads[k]).reshape(p.shape)
    errs.append((max_rel(g, r), rel_l2(g, r), k))
errs.sort(reverse=True)
for e in errs[:25]:
    print("%.3e %.3e %s" % e)
print("... median max_rel %.3e" % np.median([e[0] for e in errs]), "n>2e-4:", sum(1 for e in errs if e[0] > 2e-4), "of", len(errs))
