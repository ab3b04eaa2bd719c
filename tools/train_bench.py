#!/usr/bin/env python
"""Training-step timing: forward(train) + FlowLoss + fused backward [+ optimizer], per-kind kernel times."""
import argparse
import ctypes
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ipoke_inn_b200 as P
from ipoke_inn_b200 import _capi
from bench import CONFIGS, macs_per_sample

ap = argparse.ArgumentParser()
ap.add_argument("--config", default="cfgA")
ap.add_argument("--batch", type=int, default=4096)
ap.add_argument("--precision", default="bf16")
ap.add_argument("--reps", type=int, default=10)
ap.add_argument("--opt", default="none", choices=["none", "adam", "fused", "sharded"])
ap.add_argument("--tag", default="")
ap.add_argument("--direct-grads", action="store_true", help="engine.direct_param_grads = True")
ap.add_argument("--host-times", action="store_true", help="host time of each part of the step (no syncs)")
a = ap.parse_args()
kind, kw, _, _ = CONFIGS[a.config]
dev = torch.device("cuda:0")
torch.manual_seed(42)
net = (P.SupervisedTransformer(dict(kw), precision=a.precision) if kind == "cond" else
       P.UnsupervisedTransformer3(**kw, precision=a.precision)).to(dev)
C, E = kw["flow_in_channels"], kw.get("flow_embedding_channels", 0)
B = a.batch
x = torch.randn(B, C, device=dev)
cond = torch.randn(B, E, 1, 1, device=dev) if kind == "cond" else None
loss_fn = P.FlowLoss()
opt = (torch.optim.Adam(net.parameters(), lr=1e-4, amsgrad=True) if a.opt == "adam" else
       P.FusedAdam(net.parameters(), lr=1e-4, amsgrad=True) if a.opt == "fused" else
       P.ShardedAdam(net, lr=1e-4, amsgrad=True) if a.opt == "sharded" else None)
eng = net.flow.engine
eng.direct_param_grads = bool(a.direct_grads)


def step():
    (opt if opt is not None else net).zero_grad(set_to_none=True)
    z, ld = net(x, cond) if cond is not None else net(x)
    loss, _ = loss_fn(z, ld)
    loss.backward()
    if opt is not None:
        opt.step()
    return loss


for _ in range(3):
    loss = step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0 = time.perf_counter()
e0.record()
for _ in range(a.reps):
    loss = step()
e1.record()
torch.cuda.synchronize()
t1 = time.perf_counter()
ms = e0.elapsed_time(e1) / a.reps
pw = macs_per_sample(kind, kw)
tf = 6.0 * pw * B / (ms * 1e-3) / 1e12
print(f"{a.tag or 'train'}: {a.config} B={B} {a.precision} opt={a.opt}: {ms:.3f} ms/step (wall {(t1 - t0) / a.reps * 1e3:.3f}), "
      f"{B / ms * 1e3:.0f} samples/s, {tf:.1f} TFLOP/s algorithmic (6*P_w), loss {float(loss):.4f}, "
      f"launches fwd {eng.launches} bwd {eng.bwd_launches}")

if a.host_times:
    # host time of each part of one step, GPU queue drained first and no sync inside: what the CPU spends issuing
    acc = [0.0] * 5
    for _ in range(a.reps):
        torch.cuda.synchronize()
        t = [time.perf_counter()]
        (opt if opt is not None else net).zero_grad(set_to_none=True); t.append(time.perf_counter())
        z, ld = net(x, cond) if cond is not None else net(x); t.append(time.perf_counter())
        loss, _ = loss_fn(z, ld); t.append(time.perf_counter())
        loss.backward(); t.append(time.perf_counter())
        if opt is not None:
            opt.step()
        t.append(time.perf_counter())
        for i in range(5):
            acc[i] += (t[i + 1] - t[i]) / a.reps * 1e3
    print("  host ms: zero_grad %.3f forward %.3f loss %.3f backward %.3f optimizer %.3f" % tuple(acc))

# per-kind kernel time of one forward(train)+backward
L = _capi.lib()
L.ipoke_inn_profile_enable(eng._plan, 1)
ms5 = (ctypes.c_double * 5)()
n5 = (ctypes.c_int * 5)()
net.zero_grad(set_to_none=True)
z, ld = net(x, cond) if cond is not None else net(x)
L.ipoke_inn_profile_read_n(eng._plan, ms5, n5, 5)
fwd = list(ms5), list(n5)
loss, _ = loss_fn(z, ld)
loss.backward()
L.ipoke_inn_profile_read_n(eng._plan, ms5, n5, 5)
bwd = list(ms5), list(n5)
L.ipoke_inn_profile_enable(eng._plan, 0)
names = ["elementwise", "hidden", "tail", "dgrad", "wgrad"]
print("  forward :", {n: (round(m, 3), c) for n, m, c in zip(names, *fwd) if c})
print("  backward:", {n: (round(m, 3), c) for n, m, c in zip(names, *bwd) if c})
