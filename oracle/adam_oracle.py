"""TEST INFRASTRUCTURE -- CPU restatement (numpy) of the optimizer step the reference's training loop runs on the
flow parameters.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this.

The reference calls torch.optim.Adam(params, lr, betas=(0.9, 0.999), weight_decay=wd, amsgrad=True)
(models/second_stage_video_fc.py:685-687; AdaBelief is off in every config).  The arithmetic lives in PyTorch
(third-party: pinned pytorch=1.7.1 in ipoke.yml:21; 2.11 in this image), `torch/optim/adam.py::_single_tensor_adam`;
this restates that published algorithm.  Pinned by tests/golden/adam_golden.npz, generated from torch.optim.Adam
itself on CPU by tests/golden/make_adam_golden.py.
"""
import numpy as np


def adam_step(p, g, m, v, vmax, step, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, amsgrad=True,
              grad_scale=1.0):
    """One step on fp32 arrays, in place; `step` counts from 1.  Mirrors _single_tensor_adam line by line."""
    f = np.float32
    g = g.astype(f) * f(grad_scale)
    g = g + f(weight_decay) * p                                    # grad.add(param, alpha=weight_decay)
    m += (g - m) * f(1.0 - beta1)                                  # exp_avg.lerp_(grad, 1 - beta1)
    v *= f(beta2)
    v += f(1.0 - beta2) * g * g                                    # addcmul_(grad, grad, value=1 - beta2)
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    step_size = lr / bc1
    bc2_sqrt = bc2 ** 0.5
    if amsgrad:
        np.maximum(vmax, v, out=vmax)
        denom = np.sqrt(vmax) / f(bc2_sqrt) + f(eps)
    else:
        denom = np.sqrt(v) / f(bc2_sqrt) + f(eps)
    p -= f(step_size) * (m / denom)                                # addcdiv_(exp_avg, denom, value=-step_size)
    return p


def linear_var(act_it, start_it, end_it, start_val, end_val, clip_min, clip_max):
    """utils/general.py:218-225."""
    act_val = float(end_val - start_val) / (end_it - start_it) * (act_it - start_it) + start_val
    return float(np.clip(act_val, a_min=clip_min, a_max=clip_max))
