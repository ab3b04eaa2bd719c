#!/usr/bin/env python
"""Golden vectors for the optimizer step: torch.optim.Adam (the implementation the reference calls,
models/second_stage_video_fc.py:685-687) run on CPU for 4 steps on tensors of odd sizes, with the reference's
hyper-parameters (config/FullyConnected/second_stage.yaml:92-93: lr 1e-3, weight_decay 1e-5, amsgrad) and a
no-amsgrad / no-decay variant.  Also the reference's linear_var ramp (utils/general.py:218-225), imported from
/root/reference.   Run in the build container:  python tests/golden/make_adam_golden.py"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
out = {}
torch.manual_seed(7)
shapes = [(64, 24), (24,), (1, 8, 1, 1), (13, 7), (5,), (1536, 12)]
for tag, kw in (("ams", dict(lr=1e-3, betas=(0.9, 0.999), weight_decay=1e-5, amsgrad=True)),
                ("plain", dict(lr=3e-4, betas=(0.8, 0.99), weight_decay=0.0, amsgrad=False, eps=1e-6))):
    params = [torch.nn.Parameter(torch.randn(*s)) for s in shapes]
    opt = torch.optim.Adam(params, **kw)
    for i, p in enumerate(params):
        out[f"{tag}_p0_{i}"] = p.detach().numpy().copy()
    for step in range(4):
        for i, p in enumerate(params):
            # gradients shrink over the steps so that max_exp_avg_sq != exp_avg_sq matters (amsgrad)
            p.grad = torch.randn_like(p) * (1.0 / (1 + 2 * step))
            out[f"{tag}_g{step}_{i}"] = p.grad.numpy().copy()
        opt.step()
        for i, p in enumerate(params):
            out[f"{tag}_p{step + 1}_{i}"] = p.detach().numpy().copy()
    for i, p in enumerate(params):
        st = opt.state[p]
        out[f"{tag}_m_{i}"] = st["exp_avg"].numpy().copy()
        out[f"{tag}_v_{i}"] = st["exp_avg_sq"].numpy().copy()
        if kw["amsgrad"]:
            out[f"{tag}_vmax_{i}"] = st["max_exp_avg_sq"].numpy().copy()
out["n_tensors"] = np.array(len(shapes))

# linear_var from the reference itself (utils/general.py imports heavy deps at module level: exec the function only)
src = open("/root/reference/utils/general.py").read()
start = src.index("def linear_var(")
end = src.index("\ndef ", start + 10)
ns = {"np": np}
exec(src[start:end], ns)
its = np.array([0, 1, 100, 250, 499, 500, 501, 10000], dtype=np.int64)
out["lv_its"] = its
out["lv_warm"] = np.array([ns["linear_var"](int(i), 0, 500, 0.0, 1e-3, 0.0, 1e-3) for i in its])
out["lv_decay"] = np.array([ns["linear_var"](int(i), 500, 10000, 1e-3, 0.0, 0.0, 1e-3) for i in its])
np.savez_compressed(os.path.join(HERE, "adam_golden.npz"), **out)
print("wrote adam_golden.npz", len(out), "arrays")
