#!/usr/bin/env python
"""Measurement of the MaCow unit (row N3, first slice): fused kernel vs the reference's op sequence.

For each workload prints one JSON line with, for forward (out, logdet) and for the sequential inverse:
  * kernel ms per call (CUDA events on the launching stream, >= 3 warm-ups, L2 flushed between iterations),
    samples/s, and the fp32-FMA roofline of the launch: achieved TFLOP/s = 2 * MACs / time against
    SMs x 128 FMA lanes x 2 x the SM clock sampled under load;
  * "eager_gpu": the reference's own op sequence (oracle/macow_oracle_torch.py: F.pad, F.conv2d, activation,
    1x1 conv, tanh / log, and for the inverse all of that once per line, macow2.py:222-289) on the same GPU --
    the stock path the reference would run on this box;
  * "cpu": the same op sequence on the host cores (torch CPU, all threads) on a bounded sample.
Workloads: z_dim-32 latents at 8 x 8 with kernel [2, 3] / elu (config/second_stage.yaml:53-66,
config/first_stage.yaml:53,61), reference test batch 20 (second_stage.yaml:79) and throughput batches.
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
from ipoke_inn_b200 import macow  # noqa: E402
from bench import ClockSampler  # noqa: E402
from oracle.macow_oracle_torch import TorchUnit  # noqa: E402  (baseline legs only)

WORKLOADS = [  # name, C, kernel, H, W, batch
    ("z32_8x8_b20", 32, [2, 3], 8, 8, 20),
    ("z32_8x8_b256", 32, [2, 3], 8, 8, 256),
    ("z32_8x8_b4096", 32, [2, 3], 8, 8, 4096),
    ("z16_16x16_b1024", 16, [2, 3], 16, 16, 1024),
]


def macs_per_sample(C, Hd, k0, k1, H, W):
    return 4 * H * W * (Hd * C * k0 * k1 + 2 * C * Hd)


def timed(fn, steps, warmup, flush):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(steps):
        flush.add_(1.0)                                    # > L2 (126 MB): the next call starts cold
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    return float(np.mean(ms)), float(np.min(ms))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--only", default="")
    ap.add_argument("--no-baselines", action="store_true")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    flush = torch.zeros(64 * 1024 * 1024, device=dev)      # 256 MB
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    for name, C, ks, H, W, B in WORKLOADS:
        if args.only and args.only not in name:
            continue
        torch.manual_seed(1)
        unit = macow.MaCowUnit(C, ks, activation="elu")
        Hd = unit.conv1.hidden_channels
        with torch.no_grad():                              # a trained-looking unit: non-zero gains, ActNorms set
            for k, v in unit.state_dict().items():
                if k.endswith("initialized"):
                    v.fill_(1)
                elif k.endswith("weight_g"):
                    v.copy_(0.15 + 0.3 * torch.rand_like(v))
                elif k.endswith("bias") or k.endswith("log_scale"):
                    v.copy_(0.1 * torch.randn_like(v))
        sd = {k: v.clone().numpy() for k, v in unit.state_dict().items()}
        unit = unit.to(dev).eval()
        x = torch.randn(B, C, H, W, device=dev)
        z = torch.randn(B, C, H, W, device=dev)
        flops = 2.0 * macs_per_sample(C, Hd, ks[0], ks[1], H, W) * B
        line = {"workload": name, "C": C, "hidden": Hd, "kernel": ks, "H": H, "W": W, "batch": B,
                "flops_per_call": flops, "steps": args.steps, "warmup": args.warmup,
                "l2": "256 MB written between timed iterations"}
        with torch.no_grad():
            unit(x)                                        # pack + first call
            clk = ClockSampler(0)
            t0 = time.time()
            f_ms, f_min = timed(lambda: unit(x), args.steps, args.warmup, flush)
            r_ms, r_min = timed(lambda: unit(z, reverse=True), args.steps, args.warmup, flush)
            clocks = clk.stop(t0, time.time())
        mhz = clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965.0
        peak = sms * 128 * 2 * mhz * 1e6 / 1e12
        for tag, ms, mn, rev in (("forward", f_ms, f_min, False), ("reverse", r_ms, r_min, True)):
            line[tag] = {"ms": round(ms, 4), "ms_min": round(mn, 4), "samples_per_s": round(B / ms * 1e3, 1),
                         "roofline": {"bound": "fp32 fma", "achieved": round(flops / ms / 1e9, 3), "peak": round(peak, 1),
                                      "unit": "TFLOP/s", "frac": round(flops / ms / 1e9 / peak, 4)},
                         "launch": unit.launch_info(H, W, rev), "gpu_launches": 1}
        line["clocks"] = clocks
        if not args.no_baselines:
            ref = TorchUnit(sd, "elu", 1.0, device=dev)
            with torch.no_grad():
                ef, _ = timed(lambda: ref.forward(x), max(3, args.steps // 4), 3, flush)
                er, _ = timed(lambda: ref.reverse(z), max(3, args.steps // 4), 3, flush)
                err_f = float((ref.forward(x)[0] - unit(x)[0]).abs().max() / ref.forward(x)[0].abs().max())
                err_r = float((ref.reverse(z) - unit(z, reverse=True)).abs().max() / ref.reverse(z).abs().max())
            line["eager_gpu"] = {"forward_ms": round(ef, 3), "reverse_ms": round(er, 3), "speedup_forward": round(ef / f_ms, 1),
                                 "speedup_reverse": round(er / r_ms, 1), "max_rel_forward": err_f, "max_rel_reverse": err_r,
                                 "what": "reference op sequence (oracle/macow_oracle_torch.py), torch eager, same GPU"}
            nb = min(B, 64)                                # bounded CPU sample
            cpu = TorchUnit(sd, "elu", 1.0, device="cpu")
            xc, zc = x[:nb].cpu(), z[:nb].cpu()
            with torch.no_grad():
                cpu.forward(xc)
                t = time.perf_counter(); cpu.forward(xc); cf = time.perf_counter() - t
                t = time.perf_counter(); cpu.reverse(zc); cr = time.perf_counter() - t
            line["cpu"] = {"kind": "port", "cores": torch.get_num_threads(), "sample": f"{nb} samples, one call each",
                           "forward_samples_per_s": round(nb / cf, 1), "reverse_samples_per_s": round(nb / cr, 1)}
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
