#!/usr/bin/env python
"""Quick A/B timing of forward+reverse (env: IPOKE_INN_GRAPH / IPOKE_INN_PDL / IPOKE_INN_* knobs)."""
import argparse, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ipoke_inn_b200 as P
from bench import CONFIGS
ap = argparse.ArgumentParser()
ap.add_argument("--config", default="cfgB"); ap.add_argument("--batch", type=int, default=256)
ap.add_argument("--precision", default="bf16"); ap.add_argument("--reps", type=int, default=50)
ap.add_argument("--tag", default=""); ap.add_argument("--dual", action="store_true")
a = ap.parse_args()
kind, kw, _, _ = CONFIGS[a.config]
dev = torch.device("cuda:0"); torch.manual_seed(42)
net = (P.SupervisedTransformer(dict(kw), precision=a.precision) if kind == "cond" else
       P.UnsupervisedTransformer3(**kw, precision=a.precision)).to(dev).eval()
C, E = kw["flow_in_channels"], kw.get("flow_embedding_channels", 0)
x = torch.randn(a.batch, C, device=dev); cond = torch.randn(a.batch, E, 1, 1, device=dev) if kind == "cond" else None
f = (lambda i, r: net(i, cond, reverse=r)) if kind == "cond" else (lambda i, r: net(i, reverse=r))
with torch.no_grad():
    z, _ = f(x, False); net.flow.engine.assume_frozen = True
    eng = net.flow.engine
    B = a.batch
    zb = torch.empty(B, C, device=dev); ld = torch.empty(B, device=dev); xb = torch.empty(B, C, device=dev)
    x2 = x.contiguous(); c2 = cond.reshape(B, E).contiguous() if cond is not None else None
    def step():
        eng._forward_raw(x2, c2, zb, ld, None, None, 0, eng.n_flows)
        import ctypes
        from ipoke_inn_b200 import _capi
        _capi.lib().ipoke_inn_reverse(eng._plan, eng._packed.data_ptr(), zb.data_ptr(), c2.data_ptr() if c2 is not None else None,
                                      xb.data_ptr(), B, 0, eng.n_flows, eng._workspace(B).data_ptr(), eng._stream())
    if a.dual:      # forward and reverse (independent inputs) on two streams
        sf, sr = torch.cuda.Stream(), torch.cuda.Stream()
        zin = torch.randn(B, C, device=dev)
        import ctypes
        from ipoke_inn_b200 import _capi
        def step():
            cur = torch.cuda.current_stream()
            sf.wait_stream(cur); sr.wait_stream(cur)
            with torch.cuda.stream(sf):
                eng._forward_raw(x2, c2, zb, ld, None, None, 0, eng.n_flows)
            with torch.cuda.stream(sr):
                _capi.lib().ipoke_inn_reverse(eng._plan, eng._packed.data_ptr(), zin.data_ptr(), c2.data_ptr() if c2 is not None else None,
                                              xb.data_ptr(), B, 0, eng.n_flows, eng._workspace(B).data_ptr(), eng._stream())
            cur.wait_stream(sf); cur.wait_stream(sr)
    for _ in range(5): step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    for _ in range(a.reps): step()
    e1.record(); torch.cuda.synchronize(); t1 = time.perf_counter()
    err = float((xb - x).abs().max())
print(f"{a.tag or 'run'}: {a.config} B={a.batch} {a.precision} GRAPH={os.environ.get('IPOKE_INN_GRAPH','1')} "
      f"PDL={os.environ.get('IPOKE_INN_PDL','1')} gpu {e0.elapsed_time(e1)/a.reps*1e3:.1f} us/step, wall {(t1-t0)/a.reps*1e6:.1f} us/step, "
      f"launches {eng.launches}x2, roundtrip err {err:.2e}")
