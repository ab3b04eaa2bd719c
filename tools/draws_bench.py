#!/usr/bin/env python
"""Sampling loop timing: n_draws latents per conditioning row (reference forward_sample, second_stage_video_fc.py:356-378)
as (a) the reference's loop of per-draw reverse calls, (b) one reverse call on the row-repeated conditioning,
(c) ipoke_inn_reverse_draws (cond-only subnetworks evaluated once per conditioning row)."""
import argparse, json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ipoke_inn_b200 as P
from bench import CONFIGS

ap = argparse.ArgumentParser()
ap.add_argument("--config", default="cfgA"); ap.add_argument("--cond-rows", type=int, default=4096)
ap.add_argument("--draws", type=int, default=5); ap.add_argument("--precision", default="bf16")
ap.add_argument("--reps", type=int, default=20)
a = ap.parse_args()
kind, kw, _, _ = CONFIGS[a.config]
assert kind == "cond"
dev = torch.device("cuda:0"); torch.manual_seed(42)
net = P.SupervisedTransformer(dict(kw), precision=a.precision).to(dev).eval()
C, E, B, n = kw["flow_in_channels"], kw["flow_embedding_channels"], a.cond_rows, a.draws
cond = torch.randn(B, E, 1, 1, device=dev); z = torch.randn(n, B, C, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timed(fn):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(a.reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / a.reps


with torch.no_grad():
    net(torch.randn(min(B, 4096), C, device=dev), cond[:min(B, 4096)])      # ActNorm init
    net.flow.engine.assume_frozen = True
    cond_rep = cond.repeat(n, 1, 1, 1); z_flat = z.reshape(n * B, C)
    t_loop = timed(lambda: [net(z[j], cond, reverse=True) for j in range(n)])
    t_rep = timed(lambda: net(z_flat, cond_rep, reverse=True))
    t_draws = timed(lambda: net.flow.reverse_draws(z, cond))
    ref = net(z_flat, cond_rep, reverse=True).reshape(n, B, C)
    got = net.flow.reverse_draws(z, cond).reshape(n, B, C)
    err = float((ref - got).abs().max() / ref.abs().max())
Pw = sum(p.numel() for nme, p in net.named_parameters() if nme.endswith("weight"))
out = dict(config=a.config, cond_rows=B, draws=n, precision=a.precision,
           ms_loop=t_loop, ms_repeated=t_rep, ms_draws=t_draws,
           samples_per_s_loop=n * B / t_loop * 1e3, samples_per_s_repeated=n * B / t_rep * 1e3,
           samples_per_s_draws=n * B / t_draws * 1e3, speedup_vs_repeated=t_rep / t_draws,
           tflops_algorithmic_draws=2 * Pw * n * B / t_draws * 1e-9, max_rel_diff=err)
print(json.dumps(out))
