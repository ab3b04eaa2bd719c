"""Optimizer side of the training step.

``FusedAdam`` has the constructor and ``param_groups`` / ``state_dict`` layout of ``torch.optim.Adam`` -- the reference
builds ``Adam(trainable_params, lr, betas=(0.9, 0.999), weight_decay, amsgrad=True)`` in ``configure_optimizers``
(models/second_stage_video_fc.py:670-687) and rewrites ``pg["lr"]`` every step during warm-up (:256-271) -- but
``step()`` is ONE kernel launch per param group (``ipoke_inn_adam_step``) instead of ~8 eager kernels per tensor.
Moments live in three flat fp32 buffers (views are exposed per parameter under the usual state keys
``exp_avg`` / ``exp_avg_sq`` / ``max_exp_avg_sq`` / ``step``).  CUDA only: there is no eager fallback.

``linear_var`` is the reference's warm-up / decay ramp (utils/general.py:218-225).
"""
import ctypes

import torch

from ._nvtx import annotate

from . import _capi


def linear_var(act_it, start_it, end_it, start_val, end_val, clip_min, clip_max):
    """utils/general.py:218-225: linear ramp from start_val (at start_it) to end_val (at end_it), clipped."""
    act_val = float(end_val - start_val) / (end_it - start_it) * (act_it - start_it) + start_val
    return min(max(act_val, clip_min), clip_max)


class _AdamJob(ctypes.Structure):
    _fields_ = [("p", ctypes.c_void_p), ("g", ctypes.c_void_p), ("m", ctypes.c_void_p), ("v", ctypes.c_void_p),
                ("vmax", ctypes.c_void_p), ("n", ctypes.c_longlong)]


class _GroupPlan:
    """Flat moment buffers and the device-side job / chunk tables of one param group."""

    def __init__(self, params, amsgrad):
        self.params = params
        dev = params[0].device
        self.device = dev
        self.numels = [p.numel() for p in params]
        # every parameter starts on a 16-byte boundary of the flat buffers so the kernel's float4 path applies
        self.offsets, cur = [], 0
        for n in self.numels:
            self.offsets.append(cur)
            cur += (n + 3) // 4 * 4
        self.total = cur
        self.m = torch.zeros(cur, dtype=torch.float32, device=dev)
        self.v = torch.zeros(cur, dtype=torch.float32, device=dev)
        self.vmax = torch.zeros(cur, dtype=torch.float32, device=dev) if amsgrad else None
        chunk = _capi.lib().ipoke_inn_adam_chunk_elems()
        cj, co = [], []
        for j, n in enumerate(self.numels):
            for off in range(0, n, chunk):
                cj.append(j)
                co.append(off)
        self.n_chunks = len(cj)
        self.chunk_job = torch.tensor(cj, dtype=torch.int32, device=dev)
        self.chunk_off = torch.tensor(co, dtype=torch.int64, device=dev)
        self.jobs_host = (_AdamJob * len(params))()
        self.jobs_pinned = torch.empty(ctypes.sizeof(self.jobs_host), dtype=torch.uint8).pin_memory()
        self.jobs_dev = torch.empty(ctypes.sizeof(self.jobs_host), dtype=torch.uint8, device=dev)
        self.last_ptrs = None
        self.copied = None                      # event after the last upload of the job table

    def views(self, i):
        o, n, shape = self.offsets[i], self.numels[i], self.params[i].shape
        return (self.m[o:o + n].view(shape), self.v[o:o + n].view(shape),
                None if self.vmax is None else self.vmax[o:o + n].view(shape))

    def upload_jobs(self, grads):
        ptrs = tuple((p.data_ptr(), g.data_ptr()) for p, g in zip(self.params, grads))
        if ptrs == self.last_ptrs:
            return
        if self.copied is not None:
            self.copied.synchronize()           # the pinned staging buffer is free again
        for i, (pp, gp) in enumerate(ptrs):
            o = 4 * self.offsets[i]
            j = self.jobs_host[i]
            j.p, j.g, j.n = pp, gp, self.numels[i]
            j.m, j.v = self.m.data_ptr() + o, self.v.data_ptr() + o
            j.vmax = (self.vmax.data_ptr() + o) if self.vmax is not None else None
        ctypes.memmove(self.jobs_pinned.data_ptr(), ctypes.addressof(self.jobs_host), ctypes.sizeof(self.jobs_host))
        self.jobs_dev.copy_(self.jobs_pinned, non_blocking=True)
        self.copied = torch.cuda.Event()
        self.copied.record()
        self.last_ptrs = ptrs


class FusedAdam(torch.optim.Optimizer):
    """Drop-in for ``torch.optim.Adam`` on CUDA fp32 parameters (dense gradients)."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False, grad_scale=1.0):
        if lr < 0.0 or eps < 0.0 or not 0.0 <= betas[0] < 1.0 or not 0.0 <= betas[1] < 1.0 or weight_decay < 0.0:
            raise ValueError("FusedAdam: invalid hyper-parameter")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=amsgrad))
        self.grad_scale = float(grad_scale)     # e.g. 1 / world_size when gradients were sum-all-reduced
        self._plans = {}

    def _plan(self, gi, group):
        params = [p for p in group["params"] if p.requires_grad]
        key = (gi, tuple(id(p) for p in params), bool(group["amsgrad"]))
        plan = self._plans.get(gi)
        if plan is None or plan[0] != key:
            for p in params:
                if p.device.type != "cuda" or p.dtype != torch.float32 or not p.is_contiguous():
                    raise RuntimeError("FusedAdam needs contiguous fp32 CUDA parameters (no CPU / eager fallback)")
            gp = _GroupPlan(params, bool(group["amsgrad"]))
            # one step counter shared by the group's parameters (torch keeps one tensor per parameter)
            gp.step_t = torch.tensor(float(self.state[params[0]].get("step", 0.0)))
            for i, p in enumerate(params):
                st = self.state[p]
                m, v, vm = gp.views(i)
                if "exp_avg" in st:                       # state loaded from a checkpoint: adopt its values
                    m.copy_(st["exp_avg"])
                    v.copy_(st["exp_avg_sq"])
                    if vm is not None and "max_exp_avg_sq" in st:
                        vm.copy_(st["max_exp_avg_sq"])
                st["step"] = gp.step_t
                st["exp_avg"], st["exp_avg_sq"] = m, v
                if vm is not None:
                    st["max_exp_avg_sq"] = vm
            plan = (key, gp)
            self._plans[gi] = plan
        return plan[1]

    @torch.no_grad()
    @annotate("optim.step")
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        L = _capi.lib()
        for gi, group in enumerate(self.param_groups):
            if not any(p.requires_grad for p in group["params"]):
                continue
            gp = self._plan(gi, group)
            grads = []
            for p in gp.params:
                g = p.grad
                if g is None:
                    raise RuntimeError("FusedAdam: every trainable parameter of a group needs a gradient "
                                       "(the fused backward always produces all of them)")
                if g.is_sparse:
                    raise RuntimeError("FusedAdam does not support sparse gradients")
                if g.dtype != torch.float32 or not g.is_contiguous():
                    g = g.float().contiguous()
                grads.append(g)
            gp.upload_jobs(grads)
            gp.step_t += 1
            step = int(gp.step_t)
            b1, b2 = group["betas"]
            stream = torch.cuda.current_stream(gp.device).cuda_stream
            with torch.cuda.device(gp.device):
                _capi.check(L.ipoke_inn_adam_step(gp.jobs_dev.data_ptr(), gp.chunk_job.data_ptr(),
                                                  gp.chunk_off.data_ptr(), gp.n_chunks, float(group["lr"]), float(b1),
                                                  float(b2), float(group["eps"]), float(group["weight_decay"]), step,
                                                  int(bool(group["amsgrad"])), self.grad_scale, stream), "adam_step")
            # the kernel wrote through raw pointers: bump the version counters like an in-place torch op would
            # (the flow engine repacks its bf16 weight copy when a parameter's version changes)
            torch.autograd.graph.increment_version(gp.params)
        return loss

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        self._plans = {}                 # the loaded moments are copied into fresh flat buffers at the next step
