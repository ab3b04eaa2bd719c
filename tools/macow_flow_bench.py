#!/usr/bin/env python
"""Measurement of the full MaCow flow (row N3): ``UnsupervisedMaCowTransformer3`` through the CUDA path vs the
reference's op sequence.

Workload ``inn_test`` is config/INN_test.yaml:1-18 as INN_test.py / opticalFlowINN.py run it: 8-channel VAE latents at
8 x 8, kernel [2, 3], elu, num_steps [20, 8, 5, 5, 3, 3, 1], factor 8, coupling nets 1536 wide; ``small`` is a reduced
stack for quick checks.  One JSON line per workload with, for forward (out, logdet) and for the inverse (sampling):
  * ms per call (CUDA events on the launching stream, >= 3 warm-ups; the flow's packed weights -- 1.7 GB for inn_test --
    exceed L2, and 256 MB are written between timed iterations), samples/s, kernel launches per call;
  * ``roofline``: the coupling nets' GEMM FLOPs (2 * MACs of the three convolutions, > 99 % of the flow's arithmetic)
    per call / time against the fp32-FMA peak SMs x 128 lanes x 2 x the SM clock sampled under load -- a whole-call
    figure (host launch gaps included), stated for the float32 parity path this is;
  * ``eager_gpu``: the reference's own op sequence (oracle/macow_stack_oracle_torch.py: F.conv2d, index_select, cat, and
    for the inverse every unit line by line) with torch's default settings on the same GPU, and the error against it;
  * ``cpu``: the same op sequence on the host cores (all threads) on a bounded sample.
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import ipoke_inn_b200 as P  # noqa: E402
from bench import ClockSampler  # noqa: E402


def levels(C, num_steps, factor):
    """Channels, prior factor and steps per level (MultiScaleInternal.__init__, macow2.py:841-868)."""
    step, out = C // factor, []
    for n in num_steps:
        out.append((C, factor, n))
        C, factor = C - step, factor - 1
    return out


WORKLOADS = {
    "inn_test": dict(num_steps=[20, 8, 5, 5, 3, 3, 1], flow_in_channels=8, flow_mid_channels=1536, factor=8, H=8, W=8),
    "small": dict(num_steps=[2, 1], flow_in_channels=8, flow_mid_channels=256, factor=4, H=8, W=8),
}


def nice_macs_per_sample(cfg):
    """MACs of the coupling nets per sample: per NICE2d H W (9 c_in Hd + Hd^2 + 9 Hd 2 c_t)."""
    Hd, HW, total = cfg["flow_mid_channels"], cfg["H"] * cfg["W"], 0
    for C, f, n in levels(cfg["flow_in_channels"], cfg["num_steps"], cfg["factor"]):
        per = lambda ct: HW * (9 * (C - ct) * Hd + Hd * Hd + 9 * Hd * 2 * ct)
        total += n * 4 * per(C // 2) + per(C // f)
    return total


def timed(fn, steps, warmup, flush):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(steps):
        flush.add_(1.0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    return float(np.mean(ms)), float(np.min(ms))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="inn_test")
    ap.add_argument("--batch", type=int, default=40)          # config/INN_test.yaml:32
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--no-baselines", action="store_true")
    ap.add_argument("--cpu-samples", type=int, default=4)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    w = WORKLOADS[args.workload]
    cfg = dict(num_steps=w["num_steps"], flow_in_channels=w["flow_in_channels"], flow_mid_channels=w["flow_mid_channels"],
               factor=w["factor"], transform="affine", prior_transform="affine", kernel_size=[2, 3], coupling_type="conv",
               activation="elu")
    torch.manual_seed(1)
    net = P.UnsupervisedMaCowTransformer3(cfg)
    with torch.no_grad():                                     # a trained-looking flow: non-zero gains, ActNorms set
        for k, v in net.state_dict().items():
            if k.endswith("initialized"):
                v.fill_(1)
            elif k.endswith("weight_g"):
                v.copy_(0.02 + 0.06 * torch.rand_like(v))          # small gains: a well-conditioned inverse (tests/test_zy_gpu_macow_flow.py::randomise)
            elif k.endswith("bias") or k.endswith("log_scale"):
                v.copy_(0.05 * torch.randn_like(v))
    if not args.no_baselines:          # the baseline legs are the only users of oracle/ here
        from oracle.macow_stack_oracle import sub
        from oracle.macow_stack_oracle_torch import TorchFlow
    sd = None if args.no_baselines else sub({k: v.clone().numpy() for k, v in net.state_dict().items()}, "flow.")
    net = net.to(dev).eval()
    B, C, H, W = args.batch, w["flow_in_channels"], w["H"], w["W"]
    x = torch.randn(B, C, H, W, device=dev)
    flush = torch.zeros(64 * 1024 * 1024, device=dev)
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    flops = 2.0 * nice_macs_per_sample(w) * B
    line = {"workload": args.workload, "config": {k: w[k] for k in ("num_steps", "flow_in_channels", "flow_mid_channels", "factor")},
            "H": H, "W": W, "batch": B, "dtype": "f32", "coupling_flops_per_call": flops, "steps": args.steps,
            "warmup": args.warmup, "l2": "packed weights exceed L2; 256 MB written between timed iterations"}
    with torch.no_grad():
        z = net(x)[0]                                         # pack + first call; the inverse is timed on the flow's own
        net(z, reverse=True)                                  # outputs (a deep random flow inverts N(0, 1) noise to inf)
        clk = ClockSampler(0)
        t0 = time.time()
        f_ms, f_min = timed(lambda: net(x), args.steps, args.warmup, flush)
        r_ms, r_min = timed(lambda: net(z, reverse=True), args.steps, args.warmup, flush)
        clocks = clk.stop(t0, time.time())
    mhz = clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965.0
    peak = sms * 128 * 2 * mhz * 1e6 / 1e12
    launches = net.flow._program.launches
    for tag, ms, mn in (("forward", f_ms, f_min), ("reverse", r_ms, r_min)):
        line[tag] = {"ms": round(ms, 3), "ms_min": round(mn, 3), "samples_per_s": round(B / ms * 1e3, 1),
                     "gpu_launches": launches,
                     "roofline": {"bound": "fp32 fma", "achieved": round(flops / ms / 1e9, 2), "peak": round(peak, 1),
                                  "unit": "TFLOP/s", "frac": round(flops / ms / 1e9 / peak, 4)}}
    line["clocks"] = clocks
    if not args.no_baselines:
        ref = TorchFlow(sd, C, w["num_steps"], w["factor"], "elu", 1.0, device=dev)
        with torch.no_grad():
            ef, _ = timed(lambda: ref.forward(x), 2, 1, flush)
            er, _ = timed(lambda: ref.reverse(z), 2, 1, flush)
            a, b = ref.forward(x)[0], net(x)[0]
            err_f = float((a - b).abs().max() / a.abs().max())
            a, b = ref.reverse(z), net(z, reverse=True)
            err_r = float((a - b).abs().max() / a.abs().max())
            round_trip = float((b - x).abs().max() / x.abs().max())
        line["eager_gpu"] = {"forward_ms": round(ef, 2), "reverse_ms": round(er, 2), "speedup_forward": round(ef / f_ms, 2),
                             "speedup_reverse": round(er / r_ms, 2), "max_rel_forward": err_f, "max_rel_reverse": err_r, "round_trip_max_rel": round_trip,
                             "tf32": {"cudnn": torch.backends.cudnn.allow_tf32, "matmul": torch.backends.cuda.matmul.allow_tf32},
                             "what": "reference op sequence (oracle/macow_stack_oracle_torch.py), torch eager defaults, same GPU"}
        nb = max(1, min(B, args.cpu_samples))
        cpu = TorchFlow(sd, C, w["num_steps"], w["factor"], "elu", 1.0, device="cpu")
        xc, zc = x[:nb].cpu(), z[:nb].cpu()
        with torch.no_grad():
            t = time.perf_counter(); cpu.forward(xc); cf = time.perf_counter() - t
            t = time.perf_counter(); cpu.reverse(zc); cr = time.perf_counter() - t
        line["cpu"] = {"kind": "port", "cores": torch.get_num_threads(), "sample": f"{nb} samples, one call each",
                       "forward_samples_per_s": round(nb / cf, 2), "reverse_samples_per_s": round(nb / cr, 2)}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
