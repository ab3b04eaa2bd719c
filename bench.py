#!/usr/bin/env python
"""bench.py -- INN samples/s (forward+logdet then reverse sampling) on N B200s, one process per GPU.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                  [--config cfgB|cfgA|cfgS] [--batch B] [--precision bf16|fp32] [--no-sweep]

A "step" = one forward (z, logdet) + one reverse pass over a batch of B synthetic latents per GPU
(weights: reference default init under seed 42, ActNorm data-initialised on the first batch).
Default workload = BASELINE.json configs[1]: cfg B (INN_FCAE.yaml: unconditional C=512 H=1024 depth=2
n_flows=8), batch 256 per GPU.  Prints ONE JSON line on rank 0 (contract in the task statement).
`--impl reference` times the CPU port of the reference path (oracle/, numpy + OpenBLAS on all host cores).
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIGS = {
    # name: (kind, ctor kwargs, default batch per GPU, source)
    "cfgB": ("uncond", dict(flow_in_channels=512, flow_mid_channels=1024, flow_hidden_depth=2, n_flows=8), 256,
             "INN_FCAE.yaml (config/INN_FCAE.yaml:40-42, z_dim config/BigGanAE.yaml:49): UnsupervisedTransformer3"),
    "cfgA": ("cond", dict(flow_in_channels=128, flow_mid_channels_factor=12, flow_hidden_depth=3, n_flows=8,
                          flow_conditioning_option="none", flow_embedding_channels=128), 4096,
             "FullyConnected/second_stage.yaml:55-59: SupervisedTransformer"),
    "cfgS": ("cond", dict(flow_in_channels=8, flow_mid_channels_factor=8, flow_hidden_depth=2, n_flows=8,
                          flow_conditioning_option="none", flow_embedding_channels=16), 256,
             "INN_test.yaml:17 (smoke)"),
}
METRIC = "INN samples/sec (fwd+logdet, reverse sampling)"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], bf16=d["bf16_tflops"], bf16_sus=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    src="MEASURED_PEAKS.json")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sus=1400.0, src="fallback (B200_PROFILING.md)")


def workload_config(cfg_name, B, world, precision):
    """The `config` object of the JSON line -- identical for our arm and the reference arm."""
    kind, kw, _, src = CONFIGS[cfg_name]
    C, E = kw["flow_in_channels"], kw.get("flow_embedding_channels", 0)
    H = kw["flow_mid_channels"] if kind == "uncond" else kw["flow_mid_channels_factor"] * C
    wbytes = 2 if precision == "bf16" else 4
    return {"workload": f"{cfg_name}: {src}; {B} latents per GPU per step, forward+logdet then reverse",
            "batch_per_gpu": B, "global_batch": B * world, "C": C, "E": E, "H": H, "depth": kw["flow_hidden_depth"],
            "n_flows": kw["n_flows"], "precision": precision, "parallelism": f"dp{world}",
            "l2": "256 MiB buffer written between timed iterations (L2 flush); weights "
                  f"{macs_per_sample(kind, kw) * wbytes / 1e6:.0f} MB",
            "streams": "the forward pass and the (independent) reverse pass of a step run concurrently on 2 CUDA streams"}


def traffic_from_profiles(cfg_name, B, precision):
    """dram read+write bytes per hidden-GEMM launch from the committed `ncu --set full` capture of this workload
    (profiles/r01_traffic.json, written by tools/ncu_summary.py); None when that workload was not captured."""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "r01_traffic.json")))
        return d.get(f"{cfg_name}:{B}:{precision}:streams2", d.get(f"{cfg_name}:{B}:{precision}"))["hidden_gemm_dram_bytes_per_launch"]
    except Exception:  # noqa: BLE001
        return None


def macs_per_sample(kind, kw):
    C = kw["flow_in_channels"]
    d, n = kw["flow_hidden_depth"], kw["n_flows"]
    if kind == "cond":
        H, E = kw["flow_mid_channels_factor"] * C, kw["flow_embedding_channels"]
        tot = 0
        for k in range(n):
            din = E if k % 2 == 0 else C // 2 + E
            tot += 4 * (din * H + d * H * H + H * (C // 2))
        return tot
    H = kw["flow_mid_channels"]
    return n * 4 * ((C // 2) * H + d * H * H + H * (C // 2))


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region: an NVML polling thread (5 ms period; the
    timed region of the default workload lasts tens of milliseconds, too short for `nvidia-smi -lms`)."""
    BAD = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}

    def __init__(self, gpu_index):
        self.rows, self.ok, self._stop = [], False, False
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[gpu_index]) if vis and vis.split(",")[gpu_index].strip().isdigit() else gpu_index
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.ok = True
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
        except Exception as e:  # noqa: BLE001
            self.err = str(e)[:100]

    def _poll(self):
        nv = self.nv
        while not self._stop:
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                try:
                    reasons = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:  # noqa: BLE001
                    reasons = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.rows.append((time.time(), float(mhz), int(reasons)))
            except Exception:  # noqa: BLE001
                pass
            time.sleep(0.005)

    def stop(self, t0, t1):
        if not self.ok:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable: " + getattr(self, "err", "")]}
        self._stop = True
        self.t.join(1.0)
        rows = [r for r in self.rows if t0 <= r[0] <= t1] or [r for r in self.rows if t0 - 0.05 <= r[0] <= t1 + 0.05]
        sm = sorted(r[1] for r in rows)
        reasons = sorted(n for n, bit in self.BAD.items() if any(r[2] & bit for r in rows))
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_mhz, "reasons": reasons,
                "samples": len(sm), "source": "NVML polled every 5 ms inside the timed region"}


# ----------------------------------------------------------------------------- CPU arm (oracle port)
def cpu_port_run(cfg_name, B, steps, warmup, budget_s=20.0):
    """Times the CPU restatement of the reference path on the host cores.  One step = forward + reverse on a sample
    of the workload; the sample batch is bounded so a run stays short.  The timed code is oracle/inn_oracle_torch.py:
    the reference's own torch CPU operators (F.linear / leaky_relu / tanh / exp / cat / index) on all host threads --
    the reference module itself is Python under /root/reference and cannot travel to the GPU box."""
    import torch
    from oracle import inn_oracle as O
    from oracle import inn_oracle_torch as OT
    import ipoke_inn_b200 as P
    kind, kw, _, _ = CONFIGS[cfg_name]
    torch.manual_seed(42)
    torch.set_num_threads(os.cpu_count())
    net = P.SupervisedTransformer(dict(kw)) if kind == "cond" else P.UnsupervisedTransformer3(**kw)
    sd = O.state_dict_to_numpy(net.state_dict())
    C, E = kw["flow_in_channels"], kw.get("flow_embedding_channels", 0)
    g = torch.Generator().manual_seed(1234)
    x = torch.randn(B, C, generator=g)
    cond = torch.randn(B, E, generator=g) if kind == "cond" else None
    zin = torch.randn(B, C, generator=g)
    O.flow_forward_with_actnorm_init(sd, x.numpy(), None if cond is None else cond.numpy())   # ActNorm data init
    sdt = {k: torch.from_numpy(__import__("numpy").ascontiguousarray(v)) for k, v in sd.items()}
    t_first = time.perf_counter()
    OT.flow_forward(sdt, x, cond)
    OT.flow_reverse(sdt, zin, cond)
    t_first = time.perf_counter() - t_first
    steps = max(1, min(steps, int(budget_s / max(t_first, 1e-6))))
    for _ in range(max(warmup, 1)):
        OT.flow_forward(sdt, x, cond)
        OT.flow_reverse(sdt, zin, cond)
    t0 = time.perf_counter()
    for _ in range(steps):
        OT.flow_forward(sdt, x, cond)
        OT.flow_reverse(sdt, zin, cond)
    dt = time.perf_counter() - t0
    return {"value": B * steps / dt, "unit": "samples/s", "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{steps} step(s) of forward+reverse on batch {B} ({cfg_name}, torch {torch.__version__} CPU "
                      f"operators as in the reference, fp32, {torch.get_num_threads()} threads of {os.cpu_count()} cores)",
            "ms_per_step": 1e3 * dt / steps, "steps": steps}


# ----------------------------------------------------------------------------- GPU arm
def gpu_run(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import ipoke_inn_b200 as P
    from ipoke_inn_b200 import _capi

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"          # keep stdout to the one JSON line
        dist.init_process_group("nccl", device_id=dev)
    kind, kw, defB, src = CONFIGS[args.config]
    B = args.batch or defB
    C, E = kw["flow_in_channels"], kw.get("flow_embedding_channels", 0)
    pk = peaks()

    def make(precision):
        torch.manual_seed(42)
        net = P.SupervisedTransformer(dict(kw), precision=precision) if kind == "cond" else \
            P.UnsupervisedTransformer3(**kw, precision=precision)
        return net.to(dev).eval()

    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    x = torch.randn(B, C, device=dev, generator=g)
    cond = torch.randn(B, E, 1, 1, device=dev, generator=g) if kind == "cond" else None
    zin = torch.randn(B, C, 1, 1, device=dev, generator=g)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)       # > 126 MB L2

    def call(net, inp, reverse=False):
        with torch.no_grad():
            if kind == "cond":
                return net(inp, cond, reverse=reverse)
            return net(inp, reverse=reverse)

    def timed(net, steps, warmup, step_fn):
        for _ in range(warmup):
            step_fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        t0 = time.time()
        for s in range(steps):
            flush.fill_(s & 0xFF)                                       # L2 flush between timed iterations
            evs[s][0].record()
            step_fn()
            evs[s][1].record()
        torch.cuda.synchronize()
        t1 = time.time()
        ms = sum(a.elapsed_time(b) for a, b in evs)
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
            dist.barrier()
        return ms, t0, t1

    net = make(args.precision)
    call(net, x)                                                        # ActNorm data init + pack
    eng = net.flow.engine
    eng.assume_frozen = True

    # A step = one forward (z, logdet) pass over B latents + one reverse (sampling) pass over B independent N(0,1)
    # draws.  The two passes share nothing but the weights, so they are issued as two concurrent requests on two CUDA
    # streams (one scratch workspace per stream); the serial single-stream figure is measured first and reported too.
    s_f, s_r = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def step_serial():
        call(net, x)
        call(net, zin, reverse=True)

    def step_dev():
        cur = torch.cuda.current_stream(dev)
        s_f.wait_stream(cur)
        s_r.wait_stream(cur)
        with torch.cuda.stream(s_f):
            call(net, x)
        with torch.cuda.stream(s_r):
            call(net, zin, reverse=True)
        cur.wait_stream(s_f)
        cur.wait_stream(s_r)

    ms_serial, _, _ = timed(net, args.steps, args.warmup, step_serial)
    eng.set_concurrency(2)
    sampler = ClockSampler(local) if rank == 0 else None
    ms, t0, t1 = timed(net, args.steps, args.warmup, step_dev)
    clocks = sampler.stop(t0, t1) if sampler else None
    call(net, x)
    launches = eng.launches
    call(net, zin, reverse=True)
    launches += eng.launches
    value = world * B * args.steps / (ms * 1e-3)

    # ---- e2e: host (pinned) buffers through the C-ABI host entry points, copies inside the timed region
    xh = x.cpu().pin_memory()
    zinh = zin.reshape(B, C).cpu().pin_memory()
    ch = cond.reshape(B, E).cpu().pin_memory() if cond is not None else None
    zh = torch.empty(B, C).pin_memory()
    ldh = torch.empty(B).pin_memory()
    xoh = torch.empty(B, C).pin_memory()

    def step_host():
        cur = torch.cuda.current_stream(dev)
        s_f.wait_stream(cur)
        s_r.wait_stream(cur)
        with torch.cuda.stream(s_f):
            eng.forward_host(xh, ch, zh, ldh)
        with torch.cuda.stream(s_r):
            eng.reverse_host(zinh, ch, xoh)
        s_f.synchronize()                                               # results are on the host
        s_r.synchronize()
        cur.wait_stream(s_f)
        cur.wait_stream(s_r)

    ms_e2e, _, _ = timed(net, args.steps, args.warmup, step_host)
    e2e_value = world * B * args.steps / (ms_e2e * 1e-3)
    h2d = 2 * B * C * 4 + (2 * B * E * 4 if cond is not None else 0)
    d2h = 2 * B * C * 4 + B * 4

    out = None
    if rank == 0:
        # ---- roofline of the dominant kernel (hidden-layer GEMM), CUDA events around each launch
        L = _capi.lib()
        L.ipoke_inn_profile_enable(eng._plan, 1)
        tot = np.zeros(3)
        cnt = np.zeros(3, dtype=np.int64)
        ms3 = (ctypes.c_double * 3)()
        n3 = (ctypes.c_int * 3)()
        for s in range(args.steps):
            flush.fill_(s & 0xFF)
            for rev in (False, True):
                call(net, zin if rev else x, reverse=rev)
                L.ipoke_inn_profile_read(eng._plan, ms3, n3)
                tot += np.array(list(ms3))
                cnt += np.array(list(n3))
        L.ipoke_inn_profile_enable(eng._plan, 0)
        H = kw["flow_mid_channels"] if kind == "uncond" else kw["flow_mid_channels_factor"] * C
        d = kw["flow_hidden_depth"]
        P_w = macs_per_sample(kind, kw)
        tail_macs = kw["n_flows"] * 4 * H * (C // 2)
        hidden_flops_per_step = 2 * 2.0 * (P_w - tail_macs) * B             # fwd + rev
        wbytes = 2 if args.precision == "bf16" else 4
        hidden_launches = int(cnt[1] // args.steps)
        # algorithmic bytes of the hidden GEMMs per step: weights once + activations in/out (bf16) per launch
        hidden_bytes_per_step = 2.0 * ((P_w - tail_macs) * wbytes) + hidden_launches * (2.0 * B * 2 * H * wbytes)
        # duration of the hidden-layer launches inside the step: the event-bracketed pass serialises the launches
        # (no graph replay, no programmatic dependent launch, ~5 us of launch latency each), so the kernel's share
        # of that pass is applied to the serial single-stream step time of the timed region
        t_hidden_isolated = tot[1] / args.steps * 1e-3
        share = float(tot[1] / max(tot.sum(), 1e-9))
        t_hidden = share * (ms_serial / args.steps) * 1e-3
        tf = hidden_flops_per_step / t_hidden / 1e12 * (3 if args.precision == "fp32" else 1)
        gbs = hidden_bytes_per_step / t_hidden / 1e9
        tensor_bound = (hidden_flops_per_step / (pk["bf16"] * 1e12)) > (hidden_bytes_per_step / (pk["hbm"] * 1e9))
        roofline = {
            "kernel": "inn_gemm_kernel<EPI_HIDDEN> (hidden Linear + bias + LeakyReLU)",
            "bound": "tensor" if tensor_bound else "hbm",
            "achieved": tf if tensor_bound else gbs,
            "peak": pk["bf16"] if tensor_bound else pk["hbm"],
            "unit": "TFLOP/s" if tensor_bound else "GB/s",
            "frac": (tf / pk["bf16"]) if tensor_bound else (gbs / pk["hbm"]),
            "traffic": traffic_from_profiles(args.config, B, args.precision),
            "peak_source": pk["src"] + " (burst)",
            "launches_per_step": hidden_launches,
            "avg_launch_us": 1e6 * t_hidden / max(hidden_launches, 1),
            "avg_launch_us_isolated": 1e3 * tot[1] / max(cnt[1], 1),
            "share_of_step": share,
            "tflops_bf16_equiv": tf, "gbs": gbs,
            "algorithmic_bytes_per_launch": hidden_bytes_per_step / max(hidden_launches, 1),
            "per_kind_ms_per_step": {"boundary": tot[0] / args.steps, "hidden": tot[1] / args.steps,
                                     "tail": tot[2] / args.steps},
            "note": "share_of_step from per-launch CUDA events in a separate pass right after the timed region (same "
                    "inputs, L2 flushed each step; that pass serialises the launches: avg_launch_us_isolated); "
                    "avg_launch_us = share x serial single-stream step time / launches, i.e. the duration inside the "
                    "graph-replayed, PDL-overlapped step; fp32-split mode counts 3 MMAs per algorithmic MAC in "
                    "tflops_bf16_equiv",
        }
        cpu = cpu_port_run(args.config, min(B, 256), 20, 2, budget_s=15.0)
        out = {
            "metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16" if args.precision == "bf16" else "bf16x2-split(fp32 parity)",
            "data": "synthetic",
            "config": workload_config(args.config, B, world, args.precision),
            "e2e": {"value": e2e_value, "unit": "samples/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / args.steps,
                    "path": "ipoke_inn_forward_host / ipoke_inn_reverse_host (pinned host buffers)"},
            "serial": {"value": world * B * args.steps / (ms_serial * 1e-3), "unit": "samples/s",
                       "ms_per_step": ms_serial / args.steps,
                       "what": "same step with the forward and the reverse pass back to back on ONE stream"},
            "gpu_launches": launches * args.steps,
            "launches_per_step": launches,
            "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu,
            "flops_per_sample_fwd_rev": 4 * P_w,
            "tflops_algorithmic": value * 4 * P_w / 1e12,
            "frac_of_bf16_peak": value * 4 * P_w / 1e12 / (pk["bf16"] * world),
        }
    # ---- optional sweep (rank 0, N=1): the other precision at this size + large-batch regime
    if rank == 0 and world == 1 and not args.no_sweep:
        sweep = []
        del net
        torch.cuda.empty_cache()
        cases = [(args.config, B, "fp32" if args.precision == "bf16" else "bf16")]
        cases += [("cfgB", 4096, "bf16"), ("cfgB", 32768, "bf16"), ("cfgA", 4096, "bf16"), ("cfgA", 32768, "bf16"),
                  ("cfgA", 32768, "fp32")]
        for cname, cb, prec in cases:
            try:
                k2, kw2, _, _ = CONFIGS[cname]
                torch.manual_seed(42)
                n2 = (P.SupervisedTransformer(dict(kw2), precision=prec) if k2 == "cond" else
                      P.UnsupervisedTransformer3(**kw2, precision=prec)).to(dev).eval()
                C2, E2 = kw2["flow_in_channels"], kw2.get("flow_embedding_channels", 0)
                xs = torch.randn(cb, C2, device=dev, generator=g)
                cs = torch.randn(cb, E2, 1, 1, device=dev, generator=g) if k2 == "cond" else None
                with torch.no_grad():
                    f = (lambda inp, r: n2(inp, cs, reverse=r)) if k2 == "cond" else (lambda inp, r: n2(inp, reverse=r))
                    z0, _ = f(xs, False)
                    n2.flow.engine.assume_frozen = True
                    for _ in range(3):
                        f(xs, False); f(z0, True)
                    torch.cuda.synchronize()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    reps = 5
                    e0.record()
                    for _ in range(reps):
                        f(xs, False); f(z0, True)
                    e1.record()
                    torch.cuda.synchronize()
                    xr = f(z0, True)
                    rt = float((xr.reshape(cb, C2) - xs).abs().max())
                t = e0.elapsed_time(e1) / reps * 1e-3
                pw2 = macs_per_sample(k2, kw2)
                sweep.append({"config": cname, "batch": cb, "precision": prec, "samples_per_s": cb / t,
                              "ms_per_step": t * 1e3, "tflops_algorithmic": cb * 4 * pw2 / t / 1e12,
                              "frac_of_bf16_peak": cb * 4 * pw2 / t / 1e12 / pk["bf16"], "round_trip_max_abs": rt})
                del n2, xs, cs, z0
                torch.cuda.empty_cache()
            except Exception as e:      # the sweep is informational; never lose the headline line
                sweep.append({"config": cname, "batch": cb, "precision": prec, "error": str(e)[:200]})
        out["sweep"] = sweep
        # ---- second-stage sampling loop (second_stage_video_fc.py:356-378): n draws per conditioning row, informational
        loops = []
        for cb, nd in [(4096, 5), (20, 5)]:         # 20 x 5 is the reference's test setting (second_stage.yaml:78-79)
            try:
                k2, kw2, _, _ = CONFIGS["cfgA"]
                torch.manual_seed(42)
                n2 = P.SupervisedTransformer(dict(kw2), precision="bf16").to(dev).eval()
                C2, E2 = kw2["flow_in_channels"], kw2["flow_embedding_channels"]
                cs = torch.randn(cb, E2, 1, 1, device=dev, generator=g)
                zs = torch.randn(nd, cb, C2, device=dev, generator=g)
                with torch.no_grad():
                    n2(zs[0], cs)                                  # ActNorm data init
                    n2.flow.engine.assume_frozen = True
                    cs_rep, zs_flat = cs.repeat(nd, 1, 1, 1), zs.reshape(nd * cb, C2)
                    variants = {"per_draw_loop": lambda: [n2(zs[j], cs, reverse=True) for j in range(nd)],
                                "one_call_repeated_cond": lambda: n2(zs_flat, cs_rep, reverse=True),
                                "reverse_draws": lambda: n2.flow.reverse_draws(zs, cs)}
                    res = {}
                    for vname, fn in variants.items():
                        for _ in range(3):
                            fn()
                        torch.cuda.synchronize()
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        reps = 10
                        e0.record()
                        for _ in range(reps):
                            fn()
                        e1.record()
                        torch.cuda.synchronize()
                        res[vname] = e0.elapsed_time(e1) / reps
                pw2 = macs_per_sample(k2, kw2)
                t = res["reverse_draws"] * 1e-3
                loops.append({"config": "cfgA", "cond_rows": cb, "draws": nd, "precision": "bf16", "ms": res,
                              "samples_per_s": nd * cb / t, "tflops_algorithmic": nd * cb * 2 * pw2 / t / 1e12,
                              "speedup_vs_per_draw_loop": res["per_draw_loop"] / res["reverse_draws"],
                              "what": "reverse sampling of `draws` latents per conditioning row; reverse_draws runs the "
                                      "cond-only subnetworks of the 'cond' couplings once per conditioning row "
                                      "(algorithmic FLOPs counted as the reference's 2*P_w per draw)"})
                del n2, cs, zs
                torch.cuda.empty_cache()
            except Exception as e:  # noqa: BLE001
                loops.append({"config": "cfgA", "cond_rows": cb, "draws": nd, "error": str(e)[:200]})
        out["sampling_loop"] = loops
        # ---- training step (forward with saved state + FlowLoss + fused backward), informational
        train = []
        for cname, cb, prec in [("cfgB", 256, "bf16"), ("cfgA", 64, "bf16"), ("cfgA", 4096, "bf16"), ("cfgA", 4096, "fp32")]:
            try:
                k2, kw2, _, _ = CONFIGS[cname]
                torch.manual_seed(42)
                n2 = (P.SupervisedTransformer(dict(kw2), precision=prec) if k2 == "cond" else
                      P.UnsupervisedTransformer3(**kw2, precision=prec)).to(dev)
                C2, E2 = kw2["flow_in_channels"], kw2.get("flow_embedding_channels", 0)
                xs = torch.randn(cb, C2, device=dev, generator=g)
                cs = torch.randn(cb, E2, 1, 1, device=dev, generator=g) if k2 == "cond" else None
                lf = P.FlowLoss()
                # the reference's optimizer (second_stage_video_fc.py:685-687; FullyConnected/second_stage.yaml:92-93)
                opt = P.FusedAdam(n2.parameters(), lr=1e-3, betas=(0.9, 0.999), weight_decay=1e-5, amsgrad=True)

                def tstep():
                    opt.zero_grad(set_to_none=True)
                    zz, ll = n2(xs, cs) if cs is not None else n2(xs)
                    lo, _ = lf(zz, ll)
                    lo.backward()
                    opt.step()
                for _ in range(3):
                    tstep()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                reps = 5
                e0.record()
                for _ in range(reps):
                    tstep()
                e1.record()
                torch.cuda.synchronize()
                t = e0.elapsed_time(e1) / reps * 1e-3
                pw2 = macs_per_sample(k2, kw2)
                eng2 = n2.flow.engine
                # the optimizer kernel alone: 36 bytes per parameter (p, g, m, v, vmax read; p, m, v, vmax written)
                n_par = sum(p.numel() for p in n2.parameters())
                e0.record()
                for _ in range(reps):
                    opt.step()
                e1.record()
                torch.cuda.synchronize()
                t_adam = e0.elapsed_time(e1) / reps * 1e-3
                train.append({"config": cname, "batch": cb, "precision": prec, "ms_per_step": t * 1e3,
                              "samples_per_s": cb / t, "tflops_algorithmic_6Pw": cb * 6 * pw2 / t / 1e12,
                              "frac_of_bf16_peak": cb * 6 * pw2 / t / 1e12 / pk["bf16"],
                              "launches": eng2.launches + eng2.bwd_launches + 2,
                              "adam": {"ms": t_adam * 1e3, "params": n_par, "gbs": 36.0 * n_par / t_adam / 1e9,
                                       "frac_of_hbm_peak": 36.0 * n_par / t_adam / 1e9 / pk["hbm"]},
                              "what": "forward(train) + FlowLoss + fused backward (dgrad, wgrad, ActNorm grads) + "
                                      "fused Adam(amsgrad, weight decay) step"})
                del n2, xs, cs
                torch.cuda.empty_cache()
            except Exception as e:  # noqa: BLE001
                train.append({"config": cname, "batch": cb, "precision": prec, "error": str(e)[:200]})
        out["train_step"] = train
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="cfgB", choices=list(CONFIGS))
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--no-sweep", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    if args.impl == "reference":
        if int(os.environ.get("RANK", 0)) != 0:
            return
        kind, kw, defB, src = CONFIGS[args.config]
        B = args.batch or defB
        cpu = cpu_port_run(args.config, min(B, 256), args.steps, args.warmup, budget_s=60.0)
        line = {"impl": "reference", "metric": METRIC, "value": cpu["value"], "unit": "samples/s",
                "n_gpus": args.gpus, "steps": cpu["steps"], "warmup": max(args.warmup, 1), "ms_per_step": cpu["ms_per_step"],
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
                "data": "synthetic",
                "config": dict(workload_config(args.config, B, 1, "fp32"),
                               arm=f"CPU port of the reference path (oracle/inn_oracle_torch.py: the reference's torch "
                                   f"CPU operators, all host threads); each step = forward+reverse on a bounded sample "
                                   f"of {min(B, 256)} latents"),
                "cpu_baseline": {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")},
                "e2e": {"value": cpu["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return

    # stdout carries exactly one JSON line: whatever libraries print (NCCL's version banner, warnings) goes to stderr
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        out = gpu_run(args)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
    if out is not None:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
