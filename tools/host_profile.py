#!/usr/bin/env python
"""cProfile of the Python side of the training step (forward + FlowLoss + backward, no optimizer) at small batch:
where the host time goes once the kernels are graph-replayed.  python tools/host_profile.py --config cfgA --batch 64"""
import argparse, cProfile, os, pstats, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ipoke_inn_b200 as P
from bench import CONFIGS

ap = argparse.ArgumentParser()
ap.add_argument("--config", default="cfgA"); ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--steps", type=int, default=50); ap.add_argument("--direct-grads", action="store_true")
a = ap.parse_args()
kind, kw, _, _ = CONFIGS[a.config]
dev = torch.device("cuda:0")
torch.manual_seed(42)
net = (P.SupervisedTransformer(dict(kw)) if kind == "cond" else P.UnsupervisedTransformer3(**kw)).to(dev)
net.flow.engine.direct_param_grads = a.direct_grads
x = torch.randn(a.batch, kw["flow_in_channels"], device=dev)
cond = torch.randn(a.batch, kw.get("flow_embedding_channels", 0), 1, 1, device=dev) if kind == "cond" else None
loss_fn = P.FlowLoss()
params = list(net.parameters())


def step():
    for p in params:
        p.grad = None
    z, ld = net(x, cond) if cond is not None else net(x)
    loss, _ = loss_fn(z, ld)
    loss.backward()


for _ in range(5):
    step()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(a.steps):
    step()
torch.cuda.synchronize()
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(22)
