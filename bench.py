#!/usr/bin/env python
"""bench.py -- INN samples/s (forward+logdet then reverse sampling) on N B200s, one process per GPU.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                  [--config cfgB|cfgA|cfgS] [--batch B | --global-batch G] [--precision fp32|bf16]
                  [--no-regimes] [--no-train]

A "step" = one forward (z, logdet) + one reverse pass over a batch of B synthetic latents per GPU
(weights: reference default init under seed 42, ActNorm data-initialised on the first batch).
Default workload = BASELINE.json configs[1]: cfg B (INN_FCAE.yaml: unconditional C=512 H=1024 depth=2
n_flows=8), batch 256 per GPU, in the PARITY precision ("fp32": bf16 hi+lo split operands, <= 1e-4 of the fp32
reference -- the reference computes in fp32); the single-pass bf16 variant is reported inside `roofline.variants`.
Prints ONE JSON line on rank 0 (contract in the task statement).  Everything the line needs to be re-derived is
inside `roofline` (the driver keeps that object): the dominant kernel's figures for the timed step, `step_frac`,
`variants`, the large-batch `regimes` and the `train_step` results (with the gradient sync at N > 1).
`--global-batch G` shards a fixed global batch over the ranks (strong scaling; BASELINE configs 3 and 5).
`--impl reference` times the CPU port of the reference path (oracle/, torch CPU operators on all host cores).
"""
import argparse
import ctypes
import json
import os
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
DTYPE_NAME = {"fp32": "bf16x2-split(fp32 parity)", "bf16": "bf16"}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], bf16=d["bf16_tflops"], bf16_sus=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    src="MEASURED_PEAKS.json")
    return dict(hbm=6650.0, bf16=1590.0, bf16_sus=1400.0, src="fallback (B200_PROFILING.md)")


def dims(cfg_name):
    kind, kw, _, _ = CONFIGS[cfg_name]
    C, E = kw["flow_in_channels"], kw.get("flow_embedding_channels", 0)
    H = kw["flow_mid_channels"] if kind == "uncond" else kw["flow_mid_channels_factor"] * C
    return kind, kw, C, E, H


def macs_per_sample(kind, kw):
    """P_w of SURVEY.md 8d: Linear weight elements = MACs per sample per direction."""
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


def tail_macs(cfg_name):
    kind, kw, C, E, H = dims(cfg_name)
    return kw["n_flows"] * 4 * H * (C // 2)


def pass_bytes(cfg_name, B, precision):
    """SURVEY.md 8d "minimum HBM bytes per pass (one direction, per GPU)": weights once + I/O B*(2C+E+1)*4."""
    kind, kw, C, E, H = dims(cfg_name)
    wbytes = 2 if precision == "bf16" else 4
    return macs_per_sample(kind, kw) * wbytes + B * (2 * C + E + 1) * 4


def workload_config(cfg_name, B, world, precision, scaling="weak"):
    """The `config` object of the JSON line -- identical for our arm and the reference arm (except `arm`)."""
    kind, kw, C, E, H = dims(cfg_name)
    src = CONFIGS[cfg_name][3]
    wbytes = 2 if precision == "bf16" else 4
    return {"workload": f"{cfg_name}: {src}; {B} latents per GPU per step, forward+logdet then reverse",
            "batch_per_gpu": B, "global_batch": B * world, "C": C, "E": E, "H": H, "depth": kw["flow_hidden_depth"],
            "n_flows": kw["n_flows"], "precision": precision, "parallelism": f"dp{world}", "scaling": scaling,
            "l2": "256 MiB buffer written between timed iterations (L2 flush); weights "
                  f"{macs_per_sample(kind, kw) * wbytes / 1e6:.0f} MB",
            "streams": "the forward pass and the (independent) reverse pass of a step run concurrently on 2 CUDA streams"}


def traffic_from_profiles(cfg_name, B, precision):
    """dram read+write bytes per hidden-GEMM launch from the committed `ncu --set full` capture of this workload
    (profiles/r0*_traffic.json, written by tools/ncu_summary.py); None when that workload was not captured."""
    for name in ("r02_traffic.json", "r01_traffic.json"):
        try:
            d = json.load(open(os.path.join(ROOT, "profiles", name)))
            key = f"{cfg_name}:{B}:{precision}"
            ent = d.get(key + ":streams2", d.get(key))
            if ent is not None:
                return ent["hidden_gemm_dram_bytes_per_launch"]
        except Exception:  # noqa: BLE001
            pass
    return None


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
def macow_flow_line(timeout_s=240):
    """The MaCow conv flow at the config/INN_test.yaml stack, batch 40: the line of
    tools/macow_flow_bench.py (its own process, after this bench's timed regions; CUDA events, 3 warm-ups, the flow's
    1.7 GB of packed weights exceed L2) reduced to the figures DESIGN.md section 4b quotes.  Informational: any failure
    becomes an "error" entry, the headline line is never lost."""
    import subprocess
    try:
        cmd = [sys.executable, os.path.join(ROOT, "tools", "macow_flow_bench.py"), "--workload", "inn_test", "--batch", "40",
               "--steps", "3"]
        res = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout_s)
        lines = [ln for ln in res.stdout.splitlines() if ln.startswith("{")]
        if res.returncode != 0 or not lines:
            return {"error": ("rc %d: " % res.returncode) + (res.stderr or res.stdout)[-200:]}
        m = json.loads(lines[-1])
        out = {"workload": "UnsupervisedMaCowTransformer3, config/INN_test.yaml stack, batch %d, %dx%d latents" % (m["batch"], m["H"], m["W"]),
               "dtype": m["dtype"], "coupling_flops_per_call": m["coupling_flops_per_call"]}
        for tag in ("forward", "reverse"):
            out[tag] = {k: m[tag][k] for k in ("ms", "samples_per_s", "gpu_launches")}
            out[tag]["fp32_fma_frac"] = m[tag]["roofline"]["frac"]
            out[tag]["tflops"] = m[tag]["roofline"]["achieved"]
        out["fp32_fma_peak_tflops"] = m["forward"]["roofline"]["peak"]
        if "eager_gpu" in m:
            out["eager_reference_ops_same_gpu"] = {k: m["eager_gpu"][k] for k in ("forward_ms", "reverse_ms", "max_rel_forward", "round_trip_max_rel")}
        if "cpu" in m:
            out["cpu"] = m["cpu"]
        return out
    except Exception as e:      # noqa: BLE001
        return {"error": str(e)[:200]}


def cpu_port_run(cfg_name, B, steps, warmup, budget_s=20.0):
    """Times the CPU restatement of the reference path on the host cores.  One step = forward + reverse on a sample
    of the workload; the sample batch is bounded so a run stays short.  The timed code is oracle/inn_oracle_torch.py:
    the reference's own torch CPU operators (F.linear / leaky_relu / tanh / exp / cat / index) on all host threads --
    the reference module itself is Python under /root/reference and cannot travel to the GPU box."""
    import numpy as np
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
    sdt = {k: torch.from_numpy(np.ascontiguousarray(v)) for k, v in sd.items()}
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
class Ctx:
    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank = int(os.environ.get("RANK", 0))
        self.world = int(os.environ.get("WORLD_SIZE", 1))
        self.local = int(os.environ.get("LOCAL_RANK", 0))
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
                os.environ["NCCL_DEBUG"] = "WARN"          # keep stdout to the one JSON line
            dist.init_process_group("nccl", device_id=self.dev)
        self.pk = peaks()
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)       # > 126 MB L2
        self.gen = torch.Generator(device=self.dev).manual_seed(1234 + self.rank)

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, v):
        if self.world == 1:
            return float(v)
        t = self.torch.tensor([v], device=self.dev, dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def timed(self, step_fn, steps, warmup, flush=True):
        """W untimed steps, barrier + synchronize, K steps each bracketed by CUDA events on the current stream (L2
        flushed before each), synchronize.  -> (sum of the K step times in ms as the MAX over ranks, this rank's
        per-step list, host wall-clock window)."""
        torch = self.torch
        for _ in range(warmup):
            step_fn()
        self.barrier()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        t0 = time.time()
        for s in range(steps):
            if flush:
                self.flush.fill_(s & 0xFF)                              # L2 flush between timed iterations
            evs[s][0].record()
            step_fn()
            evs[s][1].record()
        torch.cuda.synchronize()
        t1 = time.time()
        per = [a.elapsed_time(b) for a, b in evs]
        ms = self.max_over_ranks(sum(per))
        if self.world > 1:
            self.dist.barrier()
        return ms, per, (t0, t1)


def make_net(ctx, cfg_name, precision, train=False):
    import ipoke_inn_b200 as P
    kind, kw, _, _ = CONFIGS[cfg_name]
    ctx.torch.manual_seed(42)
    net = P.SupervisedTransformer(dict(kw), precision=precision) if kind == "cond" else \
        P.UnsupervisedTransformer3(**kw, precision=precision)
    net = net.to(ctx.dev)
    return net if train else net.eval()


def stats(per):
    s = sorted(per)
    return {"min": s[0], "median": s[len(s) // 2], "max": s[-1]}


def fwd_rev_workload(ctx, cfg_name, B, precision, steps, warmup, want_clocks=False, want_e2e=True, want_kinds=True):
    """The headline measurement for one (config, batch, precision): serial step, 2-stream step, e2e step with pinned
    host buffers, launches, and (rank 0) the per-kind kernel times of an event-bracketed pass."""
    import numpy as np
    torch = ctx.torch
    from ipoke_inn_b200 import _capi
    kind, kw, C, E, H = dims(cfg_name)
    dev = ctx.dev
    net = make_net(ctx, cfg_name, precision)
    x = torch.randn(B, C, device=dev, generator=ctx.gen)
    cond = torch.randn(B, E, 1, 1, device=dev, generator=ctx.gen) if kind == "cond" else None
    zin = torch.randn(B, C, 1, 1, device=dev, generator=ctx.gen)

    def call(inp, reverse=False):
        with torch.no_grad():
            if kind == "cond":
                return net(inp, cond, reverse=reverse)
            return net(inp, reverse=reverse)

    call(x)                                                             # ActNorm data init + pack
    eng = net.flow.engine
    eng.assume_frozen = True
    # A step = one forward (z, logdet) pass over B latents + one reverse (sampling) pass over B independent N(0,1)
    # draws.  The two passes share nothing but the weights, so they are issued as two concurrent requests on two CUDA
    # streams (one scratch workspace per stream); the serial single-stream figure is measured first and reported too.
    s_f, s_r = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def step_serial():
        call(x)
        call(zin, reverse=True)

    def step_dual():
        cur = torch.cuda.current_stream(dev)
        s_f.wait_stream(cur)
        s_r.wait_stream(cur)
        with torch.cuda.stream(s_f):
            call(x)
        with torch.cuda.stream(s_r):
            call(zin, reverse=True)
        cur.wait_stream(s_f)
        cur.wait_stream(s_r)

    res = {"B": B, "precision": precision}
    ms_serial, per_serial, _ = ctx.timed(step_serial, steps, warmup)
    res["ms_serial"], res["serial_stats"] = ms_serial / steps, stats(per_serial)
    eng.set_concurrency(2)
    sampler = ClockSampler(ctx.local) if (want_clocks and ctx.rank == 0) else None
    ms, per, (t0, t1) = ctx.timed(step_dual, steps, warmup)
    res["clocks"] = sampler.stop(t0, t1) if sampler else None
    res["ms"], res["stats"] = ms / steps, stats(per)
    call(x)
    launches = eng.launches
    call(zin, reverse=True)
    res["launches"] = launches + eng.launches

    if want_e2e:
        # e2e: host (pinned) buffers through the C-ABI host entry points, copies inside the timed region
        xh = x.cpu().pin_memory()
        zinh = zin.reshape(B, C).cpu().pin_memory()
        ch = cond.reshape(B, E).cpu().pin_memory() if cond is not None else None
        zh, ldh, xoh = torch.empty(B, C).pin_memory(), torch.empty(B).pin_memory(), torch.empty(B, C).pin_memory()

        def step_host():
            cur = torch.cuda.current_stream(dev)
            s_f.wait_stream(cur)
            s_r.wait_stream(cur)
            with torch.cuda.stream(s_f):
                eng.forward_host(xh, ch, zh, ldh)
            with torch.cuda.stream(s_r):
                eng.reverse_host(zinh, ch, xoh)
            s_f.synchronize()                                           # results are on the host
            s_r.synchronize()
            cur.wait_stream(s_f)
            cur.wait_stream(s_r)

        ms_e2e, per_e2e, _ = ctx.timed(step_host, steps, warmup)
        res["ms_e2e"], res["e2e_stats"] = ms_e2e / steps, stats(per_e2e)
        res["h2d"] = 2 * B * C * 4 + (2 * B * E * 4 if cond is not None else 0)
        res["d2h"] = 2 * B * C * 4 + B * 4

    if want_kinds and ctx.rank == 0:
        # per-kind kernel time: library CUDA events around every launch of a forward + a reverse pass (this pass
        # serialises the launches: no graph replay, no programmatic dependent launch); used for the kernel's SHARE.
        # Once with the launch sizing of the 2-stream step (64-CTA launches) and once with the one-request sizing.
        L = _capi.lib()
        res["kind_ms"], res["kind_launches"] = {}, {}
        for conc in (2, 1):
            eng.set_concurrency(conc)
            L.ipoke_inn_profile_enable(eng._plan, 1)
            tot, cnt = np.zeros(3), np.zeros(3, dtype=np.int64)
            ms3, n3 = (ctypes.c_double * 3)(), (ctypes.c_int * 3)()
            reps = max(3, min(steps, 10))
            for s in range(reps):
                ctx.flush.fill_(s & 0xFF)
                for rev in (False, True):
                    call(zin if rev else x, reverse=rev)
                    L.ipoke_inn_profile_read(eng._plan, ms3, n3)
                    tot += np.array(list(ms3))
                    cnt += np.array(list(n3))
            L.ipoke_inn_profile_enable(eng._plan, 0)
            res["kind_ms"][conc] = (tot / reps).tolist()              # per step: boundary, hidden, tail
            res["kind_launches"][conc] = (cnt // reps).tolist()
    del net
    torch.cuda.empty_cache()
    return res


def hidden_roofline(ctx, cfg_name, r, streams):
    """Roofline of the dominant kernel (hidden-layer GEMM) for the step `r` times.
    Algorithmic bytes: the hidden layers' weights once per direction (2 B per weight in bf16, 4 B as
    hi+lo in the parity mode); activations live in L2 between launches and are not counted.  Duration: the kernel's
    share of the event-bracketed pass x the timed step; with `streams` concurrent passes every stream is busy for the
    whole step, so the hidden kernels run for share x ms_per_step of wall time on `streams` streams at once."""
    pk = ctx.pk
    kind, kw, C, E, H = dims(cfg_name)
    B, prec = r["B"], r["precision"]
    P_w = macs_per_sample(kind, kw)
    wbytes = 2 if prec == "bf16" else 4
    hid_w = P_w - tail_macs(cfg_name)
    tot = r["kind_ms"][streams]                                # measured with this step's launch sizing
    share = tot[1] / max(sum(tot), 1e-12)
    n_launch = int(r["kind_launches"][streams][1])             # per step (forward + reverse)
    ms_step = r["ms"] if streams > 1 else r["ms_serial"]
    t_hidden = share * ms_step * 1e-3                           # wall time with hidden kernels running
    bytes_step = 2.0 * hid_w * wbytes                           # both directions
    flops_step = 2 * 2.0 * hid_w * B * (3 if prec == "fp32" else 1)   # MMA FLOPs issued (3 passes in the split mode)
    gbs = bytes_step / t_hidden / 1e9
    tf = flops_step / t_hidden / 1e12
    tensor_bound = (flops_step / (pk["bf16"] * 1e12)) > (bytes_step / (pk["hbm"] * 1e9))
    step_bytes = pass_bytes(cfg_name, B, prec) * 2.0
    out = {
        "kernel": "inn_gemm_kernel<EPI_HIDDEN> (hidden Linear + bias + LeakyReLU)",
        "bound": "tensor" if tensor_bound else "hbm",
        "achieved": tf if tensor_bound else gbs,
        "peak": pk["bf16"] if tensor_bound else pk["hbm"],
        "unit": "TFLOP/s" if tensor_bound else "GB/s",
        "frac": (tf / pk["bf16"]) if tensor_bound else (gbs / pk["hbm"]),
        "traffic": traffic_from_profiles(cfg_name, B, prec),
        "peak_source": pk["src"] + " (burst)",
        "step": f"{cfg_name} batch {B} {prec}, " + ("forward and reverse concurrently on 2 streams (the step `value` times)"
                                                     if streams > 1 else "forward then reverse on one stream"),
        "launches_per_step": n_launch,
        "concurrent_streams": streams,
        "share_of_step": share,
        "ms_per_step": ms_step,
        "algorithmic_bytes_per_launch": bytes_step / max(n_launch, 1),
        "issued_flops_per_launch": flops_step / max(n_launch, 1),
        "gbs": gbs, "tflops_issued": tf,
        "avg_launch_us": 1e6 * t_hidden * streams / max(n_launch, 1),
        "avg_launch_us_isolated": 1e3 * tot[1] / max(n_launch, 1),
        "per_kind_ms_isolated": {"boundary": tot[0], "hidden": tot[1], "tail": tot[2]},
        "step_algorithmic_bytes": step_bytes,
        "step_frac": step_bytes / (ms_step * 1e-3) / 1e9 / pk["hbm"],
        "how": "bound = whichever of issued FLOPs / bf16 peak and algorithmic bytes / HBM peak is larger (the parity mode "
               "issues 3 bf16 MMAs per MAC and reads 4 B per weight: past the ridge from ~170 rows); achieved = "
               "launches_per_step x (issued_flops_per_launch | algorithmic_bytes_per_launch) / (share_of_step x "
               "ms_per_step); algorithmic bytes = hidden-layer weights once per direction (SURVEY.md 8d; activations stay "
               "in L2 and are not counted); share_of_step from per-launch CUDA events of a separate serialised pass; "
               "avg_launch_us = share x ms_per_step x concurrent_streams / launches; step_frac = whole-step bytes "
               "(weights + I/O, both directions) / ms_per_step / peak",
    }
    return out


def regime_workload(ctx, cfg_name, B, precision, reps=5):
    """Large-batch (tensor-bound) regime: forward+reverse on one stream, per-kind kernel times, round trip."""
    import numpy as np
    torch = ctx.torch
    from ipoke_inn_b200 import _capi
    pk = ctx.pk
    kind, kw, C, E, H = dims(cfg_name)
    net = make_net(ctx, cfg_name, precision)
    xs = torch.randn(B, C, device=ctx.dev, generator=ctx.gen)
    cs = torch.randn(B, E, 1, 1, device=ctx.dev, generator=ctx.gen) if kind == "cond" else None
    f = (lambda inp, r: net(inp, cs, reverse=r)) if kind == "cond" else (lambda inp, r: net(inp, reverse=r))
    with torch.no_grad():
        z0, _ = f(xs, False)
        eng = net.flow.engine
        eng.assume_frozen = True
        for _ in range(3):
            f(xs, False); f(z0, True)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            f(xs, False); f(z0, True)
        e1.record()
        torch.cuda.synchronize()
        xr = f(z0, True)
        rt = float((xr.reshape(B, C) - xs).abs().max())
        t = e0.elapsed_time(e1) / reps * 1e-3
        L = _capi.lib()
        L.ipoke_inn_profile_enable(eng._plan, 1)
        tot, cnt = np.zeros(3), np.zeros(3, dtype=np.int64)
        ms3, n3 = (ctypes.c_double * 3)(), (ctypes.c_int * 3)()
        for rev in (False, True):
            f(z0 if rev else xs, rev)
            L.ipoke_inn_profile_read(eng._plan, ms3, n3)
            tot += np.array(list(ms3))
            cnt += np.array(list(n3))
        L.ipoke_inn_profile_enable(eng._plan, 0)
    P_w = macs_per_sample(kind, kw)
    hid_flops = 2 * 2.0 * (P_w - tail_macs(cfg_name)) * B
    tail_flops = 2 * 2.0 * tail_macs(cfg_name) * B
    tf_step = B * 4 * P_w / t / 1e12
    out = {"config": cfg_name, "batch": B, "precision": precision, "samples_per_s": B / t, "ms_per_step": t * 1e3,
           "tflops": tf_step, "frac_burst": tf_step / pk["bf16"], "frac_sustained": tf_step / pk["bf16_sus"],
           "kernel": "hidden GEMM (inn_hidden_pair_kernel / inn_gemm_kernel<EPI_HIDDEN>)",
           "avg_us": 1e3 * tot[1] / max(cnt[1], 1), "launches": int(cnt[1]),
           "kernel_tflops": hid_flops / (tot[1] * 1e-3) / 1e12,
           "kernel_frac_burst": hid_flops / (tot[1] * 1e-3) / 1e12 / pk["bf16"],
           "tail_avg_us": 1e3 * tot[2] / max(cnt[2], 1), "tail_tflops": tail_flops / (tot[2] * 1e-3) / 1e12,
           "boundary_ms": tot[0], "round_trip_max_abs": rt,
           "note": "algorithmic FLOPs (2 per MAC); the fp32 parity mode issues 3 bf16 MMAs per MAC"
                   if precision == "fp32" else "algorithmic FLOPs (2 per MAC)"}
    del net, xs, cs, z0
    torch.cuda.empty_cache()
    return out


def train_workload(ctx, cfg_name, B, precision, reps=8):
    """cfg A training step (BASELINE config 4): forward(train) + FlowLoss + fused backward + the optimizer of the
    reference (Adam amsgrad, lr 1e-3, weight decay 1e-5; FullyConnected/second_stage.yaml:92-93), B rows per GPU.
    At N > 1 with the gradient exchange: (a) chunked NCCL all-reduce overlapped with the backward + replicated fused
    Adam, (b) reduce-scatter -> Adam on the rank's shard writing the packed bf16 weights -> all-gather (ShardedAdam)."""
    import ipoke_inn_b200 as P
    torch, dist, pk = ctx.torch, ctx.dist, ctx.pk
    kind, kw, C, E, H = dims(cfg_name)
    world = ctx.world
    P_w = macs_per_sample(kind, kw)
    out = {"config": cfg_name, "batch_per_gpu": B, "precision": precision, "n_gpus": world}

    def run(variant):
        net = make_net(ctx, cfg_name, precision, train=True)
        xs = torch.randn(B, C, device=ctx.dev, generator=ctx.gen)
        cs = torch.randn(B, E, 1, 1, device=ctx.dev, generator=ctx.gen) if kind == "cond" else None
        lf = P.FlowLoss()
        hyper = dict(lr=1e-3, betas=(0.9, 0.999), weight_decay=1e-5, amsgrad=True)
        sync = None
        if variant in ("sharded", "sharded_bf16_wire"):
            opt = P.ShardedAdam(net, wire="bf16" if variant == "sharded_bf16_wire" else "fp32", **hyper)
        else:
            opt = P.FusedAdam(net.parameters(), **hyper)
            if variant == "allreduce":
                sync = P.GradSync(net.flow, average=True, chunk_blocks=2)

        def tstep():
            opt.zero_grad(set_to_none=True)
            zz, ll = net(xs, cs) if cs is not None else net(xs)
            lo, _ = lf(zz, ll)
            lo.backward()
            opt.step()
            return lo
        ms, per, _ = ctx.timed(tstep, reps, 3, flush=False)
        n_par = sum(p.numel() for p in net.parameters())
        res = {"ms_per_step": ms / reps, "ms_min": min(per)}
        if variant == "local":
            # the optimizer kernel alone: 36 bytes per parameter (p, g, m, v, vmax read; p, m, v, vmax written)
            ms_a, _, _ = ctx.timed(opt.step, reps, 1, flush=False)
            # ... and the fp32 -> packed bf16 repack the next forward pays after every replicated step
            eng = net.flow.engine
            ms_p, _, _ = ctx.timed(lambda: eng.pack(force=True), reps, 1, flush=False)
            res["adam_ms"], res["adam_frac_of_hbm_peak"] = ms_a / reps, 36.0 * n_par / (ms_a / reps * 1e-3) / 1e9 / pk["hbm"]
            res["pack_ms"] = ms_p / reps
            res["launches"] = eng.launches + eng.bwd_launches + 2
        if sync is not None:
            sync.remove()
        del net, opt, xs, cs
        torch.cuda.empty_cache()
        return res, n_par

    try:
        out["local"], n_par = run("local")
        out["params"] = n_par
        t = out["local"]["ms_per_step"] * 1e-3
        out["local"]["tflops_6Pw_per_gpu"] = B * 6 * P_w / t / 1e12
        out["what_local"] = "every rank trains on its own rows, no gradient exchange (the 1-GPU step)"
        if world > 1:
            # the collective alone: fp32 all-reduce of the flat gradient buffer (what DDP moves per step)
            flat = torch.zeros(n_par, dtype=torch.float32, device=ctx.dev)
            ms_c, _, _ = ctx.timed(lambda: dist.all_reduce(flat), 5, 2, flush=False)
            t_ar = ms_c / 5 * 1e-3
            out["allreduce_alone"] = {"bytes": 4 * n_par, "ms": t_ar * 1e3,
                                      "busbw_gbs": 2.0 * (world - 1) / world * 4 * n_par / t_ar / 1e9}
            del flat
            torch.cuda.empty_cache()
            out["allreduce"], _ = run("allreduce")
            extra = out["allreduce"]["ms_per_step"] - out["local"]["ms_per_step"]
            out["allreduce"]["exposed_ms"] = extra
            out["allreduce"]["hidden_fraction"] = max(0.0, 1.0 - extra / (t_ar * 1e3))
        if hasattr(P, "ShardedAdam"):
            out["sharded"], _ = run("sharded")
            out["sharded"]["vs_local"] = out["sharded"]["ms_per_step"] / out["local"]["ms_per_step"]
            if world > 1:
                out["sharded_bf16_wire"], _ = run("sharded_bf16_wire")      # gradients rounded to bf16 per rank on the wire
            out["what_sharded"] = ("reduce-scatter of the flat gradient -> Adam on the rank's 1/N shard, writing the fp32 "
                                   "master and the packed bf16 (hi/lo) weights in one pass -> all-gather of the bf16 "
                                   "weights straight into the packed layout (no repack pass)")
        # (the bf16-wire variant changes the numerics of the exchange and is reported beside, not as the best step)
        best = min(out[k]["ms_per_step"] for k in ("allreduce", "sharded", "local") if k in out and (k != "local" or world == 1))
        out["best_ms_per_step"] = best
        out["samples_per_s"] = world * B / (best * 1e-3)
        out["tflops_6Pw"] = world * B * 6 * P_w / (best * 1e-3) / 1e12
    except Exception as e:  # noqa: BLE001 -- informational section: never lose the headline line
        out["error"] = f"{type(e).__name__}: {str(e)[:300]}"
    return out


def gpu_run(args):
    ctx = Ctx()
    torch, dist, pk = ctx.torch, ctx.dist, ctx.pk
    world, rank = ctx.world, ctx.rank
    kind, kw, C, E, H = dims(args.config)
    scaling = "weak"
    if args.global_batch:
        if args.global_batch % world:
            raise SystemExit("--global-batch must be a multiple of the number of GPUs")
        B, scaling = args.global_batch // world, "strong"
    else:
        B = args.batch or CONFIGS[args.config][2]
    P_w = macs_per_sample(kind, kw)

    head = fwd_rev_workload(ctx, args.config, B, args.precision, args.steps, args.warmup, want_clocks=True)
    value = world * B / (head["ms"] * 1e-3)
    out = None
    if rank == 0:
        roofline = hidden_roofline(ctx, args.config, head, streams=2)
        roofline["serial"] = hidden_roofline(ctx, args.config, head, streams=1)
        for k in ("kernel", "peak", "unit", "peak_source", "how", "traffic", "per_kind_ms_isolated"):
            roofline["serial"].pop(k, None)
        out = {
            "metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": head["ms"], "higher_is_better": True, "scaling": scaling,
            "vs_baseline": None, "dtype": DTYPE_NAME[args.precision], "data": "synthetic",
            "config": workload_config(args.config, B, world, args.precision, scaling),
            "e2e": {"value": world * B / (head["ms_e2e"] * 1e-3), "unit": "samples/s",
                    "h2d_bytes_per_step": head["h2d"], "d2h_bytes_per_step": head["d2h"], "ms_per_step": head["ms_e2e"],
                    "ms_per_step_stats": head["e2e_stats"],
                    "path": "ipoke_inn_forward_host / ipoke_inn_reverse_host (pinned host buffers)"},
            "ms_per_step_stats": head["stats"],
            "serial": {"value": world * B / (head["ms_serial"] * 1e-3), "unit": "samples/s",
                       "ms_per_step": head["ms_serial"], "ms_per_step_stats": head["serial_stats"],
                       "what": "same step with the forward and the reverse pass back to back on ONE stream"},
            "gpu_launches": head["launches"] * args.steps,
            "launches_per_step": head["launches"],
            "clocks": head["clocks"], "roofline": roofline,
            "flops_per_sample_fwd_rev": 4 * P_w,
            "tflops_algorithmic": value * 4 * P_w / 1e12,
            "frac_of_bf16_peak": value * 4 * P_w / 1e12 / (pk["bf16"] * world),
        }
    # ---- the other precision at the headline workload (all ranks take part: it has barriers)
    other = "bf16" if args.precision == "fp32" else "fp32"
    var = fwd_rev_workload(ctx, args.config, B, other, max(5, args.steps // 2), args.warmup, want_kinds=True)
    if rank == 0:
        v = hidden_roofline(ctx, args.config, var, streams=2)
        out["roofline"]["variants"] = {other: {
            "dtype": DTYPE_NAME[other], "value": world * B / (var["ms"] * 1e-3), "ms_per_step": var["ms"],
            "serial_ms_per_step": var["ms_serial"], "e2e_value": world * B / (var["ms_e2e"] * 1e-3),
            "hidden_frac": v["frac"], "step_frac": v["step_frac"], "avg_launch_us": v["avg_launch_us"],
            "what": ("single-pass bf16 variant (looser stated bound, DESIGN.md section 2)" if other == "bf16"
                     else "fp32 parity mode (bf16 hi+lo split operands)")}}
    # ---- large-batch regimes (rank 0 at N = 1; at N > 1 the sharded sweep is `--global-batch`)
    if world == 1 and not args.no_regimes:
        regimes = []
        for cname, cb, prec in [("cfgA", 32768, "bf16"), ("cfgB", 32768, "bf16"), ("cfgA", 32768, "fp32"),
                                ("cfgA", 4096, "bf16"), ("cfgB", 4096, "bf16")]:
            try:
                regimes.append(regime_workload(ctx, cname, cb, prec))
            except Exception as e:      # informational; never lose the headline line
                regimes.append({"config": cname, "batch": cb, "precision": prec, "error": str(e)[:200]})
        out["roofline"]["regimes"] = regimes
    # ---- training step (every rank takes part; the gradient exchange runs at N > 1)
    if not args.no_train:
        train = [train_workload(ctx, "cfgA", 64, "bf16"), train_workload(ctx, "cfgA", 4096, "bf16")]
        if rank == 0:
            out["roofline"]["train_step"] = train
    # ---- MaCow conv flow (row N3): its own tool in its own process, rank 0 at N = 1
    if rank == 0 and world == 1 and not args.no_macow:
        out["roofline"]["macow_flow"] = macow_flow_line()
    # ---- CPU baseline: the oracle port on the host cores, rank 0 at N = 1 only (contract); at N > 1 torchrun pins
    # OMP_NUM_THREADS=1 and the other ranks would spin in a barrier, so the figure is not taken there
    if rank == 0:
        if world == 1:
            out["cpu_baseline"] = cpu_port_run(args.config, min(B, 256), 20, 2, budget_s=15.0)
        else:
            out["cpu_baseline"] = None
            out["cpu_baseline_note"] = "measured at N = 1 only (bench.py --gpus 1, and the --impl reference arm)"
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
    ap.add_argument("--batch", type=int, default=0, help="latents per GPU per step (weak scaling)")
    ap.add_argument("--global-batch", type=int, default=0, help="latents per step over all GPUs (strong scaling)")
    ap.add_argument("--precision", default="fp32", choices=["bf16", "fp32"],
                    help="fp32 = parity mode (bf16 hi+lo split operands, <= 1e-4 of the fp32 reference); bf16 = one pass")
    ap.add_argument("--no-regimes", action="store_true")
    ap.add_argument("--no-train", action="store_true")
    ap.add_argument("--no-macow", action="store_true", help="skip the MaCow conv-flow section (roofline.macow_flow)")
    ap.add_argument("--no-sweep", action="store_true", help="same as --no-regimes --no-train")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.no_sweep:
        args.no_regimes = args.no_train = True

    if args.impl == "reference":
        if int(os.environ.get("RANK", 0)) != 0:
            return
        kind, kw, defB, src = CONFIGS[args.config]
        world = int(os.environ.get("WORLD_SIZE", 1))
        B = (args.global_batch // max(world, 1)) if args.global_batch else (args.batch or defB)
        cpu = cpu_port_run(args.config, min(B, 256), args.steps, args.warmup, budget_s=60.0)
        line = {"impl": "reference", "metric": METRIC, "value": cpu["value"], "unit": "samples/s",
                "n_gpus": args.gpus, "steps": cpu["steps"], "warmup": max(args.warmup, 1), "ms_per_step": cpu["ms_per_step"],
                "higher_is_better": True, "scaling": "strong" if args.global_batch else "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic",
                "config": dict(workload_config(args.config, B, 1, "fp32", "strong" if args.global_batch else "weak"),
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
