#!/usr/bin/env python
"""Per-kind kernel time (CUDA events around every launch) of one forward and one reverse pass."""
import argparse, ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ipoke_inn_b200 as P
from ipoke_inn_b200 import _capi
from bench import CONFIGS, macs_per_sample
ap = argparse.ArgumentParser()
ap.add_argument("--config", default="cfgB"); ap.add_argument("--batch", type=int, default=32768)
ap.add_argument("--precision", default="bf16")
a = ap.parse_args()
kind, kw, _, _ = CONFIGS[a.config]
dev = torch.device("cuda:0"); torch.manual_seed(42)
net = (P.SupervisedTransformer(dict(kw), precision=a.precision) if kind == "cond" else
       P.UnsupervisedTransformer3(**kw, precision=a.precision)).to(dev).eval()
C, E = kw["flow_in_channels"], kw.get("flow_embedding_channels", 0)
B = a.batch
x = torch.randn(B, C, device=dev); cond = torch.randn(B, E, 1, 1, device=dev) if kind == "cond" else None
f = (lambda i, r: net(i, cond, reverse=r)) if kind == "cond" else (lambda i, r: net(i, reverse=r))
L = _capi.lib()
names = ["elementwise", "hidden", "tail", "dgrad", "wgrad"]
with torch.no_grad():
    z, _ = f(x, False); f(z, True)
    eng = net.flow.engine
    for rev in (False, True):
        f(z if rev else x, rev); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); f(z if rev else x, rev); e1.record(); torch.cuda.synchronize()
        L.ipoke_inn_profile_enable(eng._plan, 1)
        f(z if rev else x, rev)
        ms5 = (ctypes.c_double * 5)(); n5 = (ctypes.c_int * 5)()
        L.ipoke_inn_profile_read_n(eng._plan, ms5, n5, 5)
        L.ipoke_inn_profile_enable(eng._plan, 0)
        pw = macs_per_sample(kind, kw)
        print(f"{a.config} B={B} {a.precision} {'rev' if rev else 'fwd'}: total {e0.elapsed_time(e1):.3f} ms "
              f"({2*pw*B/e0.elapsed_time(e1)/1e9:.0f} TF/s) | " +
              ", ".join(f"{names[i]} {ms5[i]:.3f} ms/{n5[i]}" for i in range(5) if n5[i]))
