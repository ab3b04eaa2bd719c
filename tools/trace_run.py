#!/usr/bin/env python
"""Device timeline of the GEMM launches of one forward pass (ipoke_inn_debug_trace): where a launch's time goes."""
import argparse, ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ipoke_inn_b200 as P
from ipoke_inn_b200 import _capi
from bench import CONFIGS
ap = argparse.ArgumentParser()
ap.add_argument("--config", default="cfgB"); ap.add_argument("--batch", type=int, default=256)
ap.add_argument("--precision", default="bf16"); ap.add_argument("--ops", type=int, default=12)
a = ap.parse_args()
kind, kw, _, _ = CONFIGS[a.config]
dev = torch.device("cuda:0"); torch.manual_seed(42)
net = (P.SupervisedTransformer(dict(kw), precision=a.precision) if kind == "cond" else
       P.UnsupervisedTransformer3(**kw, precision=a.precision)).to(dev).eval()
C, E = kw["flow_in_channels"], kw.get("flow_embedding_channels", 0)
x = torch.randn(a.batch, C, device=dev); cond = torch.randn(a.batch, E, 1, 1, device=dev) if kind == "cond" else None
f = (lambda i: net(i, cond)) if kind == "cond" else (lambda i: net(i))
L = _capi.lib()
with torch.no_grad():
    f(x); eng = net.flow.engine; eng.assume_frozen = True
    n = 256
    buf = torch.zeros(n * 32, dtype=torch.int64, device=dev)
    _capi.check(L.ipoke_inn_debug_trace(eng._plan, ctypes.c_void_p(buf.data_ptr()), n))
    for _ in range(4): f(x)
    torch.cuda.synchronize()
    t = buf.cpu().view(n, 2, 16)
    _capi.check(L.ipoke_inn_debug_trace(eng._plan, None, 0))
names = ["entry", "prolog", "dep_ok", "1st_full", "lastMMA", "acc_rdy", "epi_done", "exit", "staged", "bar1", "math", "bar2", "wb", "cc", "r1", "r8"]
t0 = int(t[0, 0, 0])
print("ns relative to the first GEMM's entry; per launch: first CTA | last CTA")
prev_exit = None
for i in range(n):
    if int(t[i, 0, 0]) == 0: break
    if i < a.ops or i % 16 == 0:
        for c in range(2):
            r = [int(v) - t0 if int(v) else -1 for v in t[i, c]]
            base = r[0]
            print(f"op{i:3d} cta{'F' if c == 0 else 'L'} entry@{base:8d} " + " ".join(f"{names[k]}+{r[k]-base:5d}" for k in range(1, 16) if r[k] >= 0)
                  + (f" | gap_from_prev_exit {base - prev_exit:5d}" if (prev_exit is not None and c == 0) else ""))
    prev_exit = max(int(t[i, 0, 7]), int(t[i, 1, 7])) - t0
tot = max(int(t[i - 1, 0, 7]), int(t[i - 1, 1, 7])) - t0
print(f"{i} GEMM launches, {tot/1e3:.1f} us from first entry to last exit = {tot/i:.0f} ns per GEMM launch")
