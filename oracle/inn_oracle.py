"""CPU oracle for the iPOKE flat-INN hot path  --  TEST INFRASTRUCTURE, NOT PRODUCT.

A numpy restatement of the reference algorithm (altmann144/ipoke-main-opticalflow).
Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module, and only as the checker or the
timed CPU baseline.  The shipped path (``ipoke-main-opticalflow_b200``) never does.

Pinning: the reference ships no tests or golden vectors for this path (SURVEY.md
section 4), so the oracle is pinned against outputs of the reference itself, run in
the build container by ``tests/golden/make_golden.py`` (imports ``/root/reference``)
and committed as ``tests/golden/*.npz``; ``tests/test_oracle_golden.py`` checks
every function below against those fixtures.

All tensors are numpy arrays.  ``sd`` is a flat dict with the reference's
``state_dict`` key names relative to the wrapper module, e.g.
``flow.sub_layers.3.coupling.s.0.main.2.weight``.  Every function cites the
reference file:line it follows (paths relative to /root/reference).
"""
from __future__ import annotations

import numpy as np

LEAKY_MLP_SLOPE = 0.01   # nn.LeakyReLU() default, models/modules/INN/modules.py:174,179
INV_LEAKY_ALPHA = 0.9    # InvLeakyRelu(alpha=0.9), models/modules/INN/flow_blocks.py:373


# ------------------------------------------------- optional emulation of the GPU's bf16 GEMM operands
# The CUDA path's precision="bf16" variant rounds every GEMM operand (weights, MLP inputs, hidden
# activations, and in the backward pass the output-gradient operands) to bfloat16 and accumulates in fp32.
# With ``gemm_rounding("bf16")`` active the oracle rounds at exactly those points, so the bf16 variant can
# be held to a tight bound against "the same arithmetic on the CPU" (LeakyReLU' is discontinuous: against
# the exact fp32 reference a bf16 forward flips the mask of units whose pre-activation is ~1e-3 from zero,
# which moves single-sample gradient contributions by O(1) -- see tests/test_gpu_backward.py).
def round_bf16(a):
    """Round-to-nearest-even to bfloat16, returned in the input's dtype."""
    a = np.asarray(a)
    u = np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)
    r = ((u >> np.uint32(16)) & np.uint32(1)) + np.uint32(0x7FFF)
    return ((u + r) & np.uint32(0xFFFF0000)).view(np.float32).astype(a.dtype)


_ROUND = None


class gemm_rounding:
    """Context manager: ``with gemm_rounding("bf16"): ...`` (None / "none" = exact)."""

    def __init__(self, mode):
        self.fn = round_bf16 if mode == "bf16" else None

    def __enter__(self):
        global _ROUND
        self.prev, _ROUND = _ROUND, self.fn

    def __exit__(self, *exc):
        global _ROUND
        _ROUND = self.prev


def _R(a):
    return a if _ROUND is None else _ROUND(a)


# --------------------------------------------------------------------------- MLP
def _linear_indices(sd, prefix):
    """Indices of the nn.Linear entries inside ``<prefix>.main`` (0, 2, 4, ...)."""
    idx = set()
    plen = len(prefix) + len(".main.")
    for k in sd:
        if k.startswith(prefix + ".main.") and k.endswith(".weight"):
            idx.add(int(k[plen:].split(".")[0]))
    return sorted(idx)


def basic_fc_net(sd, prefix, x, use_tanh):
    """BasicFullyConnectedNet.forward -- models/modules/INN/modules.py:165-191.

    Linear -> LeakyReLU(0.01) -> [Linear -> LeakyReLU] x depth -> Linear -> [Tanh].
    """
    idx = _linear_indices(sd, prefix)
    h = x
    for n, i in enumerate(idx):
        w = sd[f"{prefix}.main.{i}.weight"].astype(x.dtype, copy=False)
        b = sd[f"{prefix}.main.{i}.bias"].astype(x.dtype, copy=False)
        h = _R(h) @ _R(w).T + b                          # nn.Linear: y = x W^T + b
        if n + 1 < len(idx):
            h = np.where(h >= 0, h, h * x.dtype.type(LEAKY_MLP_SLOPE))
    if use_tanh:
        h = np.tanh(h)
    return h


# ----------------------------------------------------------------------- ActNorm
def actnorm_initialize(x):
    """ActNorm.initialize -- models/modules/INN/modules.py:255-275.

    loc = -mean_c, scale = 1/(std_c + 1e-6), std unbiased over the batch.
    x: [B, C].  Returns (loc[C], scale[C]).
    """
    mean = x.mean(axis=0)
    std = x.std(axis=0, ddof=1)
    return -mean, 1.0 / (std + x.dtype.type(1e-6))


def actnorm_forward(loc, scale, x):
    """ActNorm.forward -- modules.py:292-301.  h = scale*(x+loc); logdet = sum log|scale|."""
    h = scale[None, :] * (x + loc[None, :])
    logdet = np.sum(np.log(np.abs(scale))) * np.ones(x.shape[0], dtype=x.dtype)
    return h, logdet


def actnorm_reverse(loc, scale, y):
    """ActNorm.reverse -- modules.py:305-316.  h = y/scale - loc."""
    return y / scale[None, :] - loc[None, :]


# ------------------------------------------------------------------ InvLeakyRelu
def inv_leaky_forward(x, alpha=INV_LEAKY_ALPHA):
    """InvLeakyRelu.forward -- flow_blocks.py:377-382.  logdet contribution is 0.0 (sic)."""
    scaling = (x >= 0).astype(x.dtype) + (x < 0).astype(x.dtype) * x.dtype.type(alpha)
    return x * scaling


def inv_leaky_reverse(y, alpha=INV_LEAKY_ALPHA):
    """InvLeakyRelu.reverse -- flow_blocks.py:384-387."""
    scaling = (y >= 0).astype(y.dtype) + (y < 0).astype(y.dtype) * y.dtype.type(alpha)
    return y / scaling


# ----------------------------------------------------------------------- Shuffle
def shuffle_forward(sd, prefix, x):
    """Shuffle.forward -- flow_blocks.py:322-324.  x[:, forward_shuffle_idx]."""
    return x[:, sd[f"{prefix}.forward_shuffle_idx"]]


def shuffle_reverse(sd, prefix, x):
    """Shuffle.forward(reverse=True) -- flow_blocks.py:325-326."""
    return x[:, sd[f"{prefix}.backward_shuffle_idx"]]


# --------------------------------------------------------------------- couplings
def _coupling_input(xa, cond, mode):
    # 'normal': cat(x_apply, cond)  flow_blocks.py:137 ; 'cond': cond only  :224 ;
    # 'uncond': x_apply only  flow_blocks.py:600
    if mode == "normal":
        return np.concatenate((xa, cond), axis=1)
    if mode == "cond":
        return cond
    return xa


def double_coupling_forward(sd, prefix, x, cond, mode):
    """ConditionalDoubleVectorCouplingBlock(_cond) / DoubleVectorCouplingBlock2 forward.

    flow_blocks.py:125-143 ('normal'), :212-230 ('cond'), :591-606 ('uncond').
    Returns (x_out [B,C] in the reference's output order, logdet [B]).
    """
    logdet = np.zeros(x.shape[0], dtype=x.dtype)
    half = x.shape[1] // 2
    for i in range(2):
        if i % 2 != 0:                                   # swap halves, :134-135
            x = np.concatenate((x[:, half:], x[:, :half]), axis=1)
        xa, xk = x[:, :half], x[:, half:]
        u = _coupling_input(xa, cond, mode)
        scale = basic_fc_net(sd, f"{prefix}.s.{i}", u, use_tanh=True)
        t = basic_fc_net(sd, f"{prefix}.t.{i}", u, use_tanh=False)
        xk = xk * np.exp(scale) + t                      # :139
        x = np.concatenate((xa, xk), axis=1)
        logdet = logdet + scale.sum(axis=1)              # :141-142
    return x, logdet


def double_coupling_reverse(sd, prefix, x, cond, mode):
    """Inverse coupling -- flow_blocks.py:144-153 / :231-240 / :607-615."""
    half = x.shape[1] // 2
    for i in reversed(range(2)):
        if i % 2 == 0:                                   # :147-148
            x = np.concatenate((x[:, half:], x[:, :half]), axis=1)
        xa, xk = x[:, :half], x[:, half:]
        u = _coupling_input(xa, cond, mode)
        scale = basic_fc_net(sd, f"{prefix}.s.{i}", u, use_tanh=True)
        t = basic_fc_net(sd, f"{prefix}.t.{i}", u, use_tanh=False)
        xk = (xk - t) * np.exp(-scale)                   # :151
        x = np.concatenate((xa, xk), axis=1)
    return x


# ------------------------------------------------------------------------ blocks
def _block_mode(i, conditional):
    # ConditionalFlow: mode = 'cond' if flow % 2 == 0 else 'normal'  flow_blocks.py:27
    if not conditional:
        return "uncond"
    return "cond" if i % 2 == 0 else "normal"


def block_forward(sd, i, x, cond, conditional):
    """ConditionalFlatDoubleCouplingFlowBlock.forward (flow_blocks.py:255-267) or
    UnconditionalFlatDoubleCouplingFlowBlock2.forward (:513-523)."""
    p = f"flow.sub_layers.{i}"
    loc = sd[f"{p}.norm_layer.loc"].reshape(-1).astype(x.dtype, copy=False)
    scale = sd[f"{p}.norm_layer.scale"].reshape(-1).astype(x.dtype, copy=False)
    h, logdet = actnorm_forward(loc, scale, x)
    if conditional:                                      # activation='lrelu', :13,33,254
        h = inv_leaky_forward(h)
    h, ld = double_coupling_forward(sd, f"{p}.coupling", h, cond, _block_mode(i, conditional))
    logdet = logdet + ld
    h = shuffle_forward(sd, f"{p}.shuffle", h)
    return h, logdet


def block_reverse(sd, i, y, cond, conditional):
    """Block inverse -- flow_blocks.py:268-274 / :524-529."""
    p = f"flow.sub_layers.{i}"
    loc = sd[f"{p}.norm_layer.loc"].reshape(-1).astype(y.dtype, copy=False)
    scale = sd[f"{p}.norm_layer.scale"].reshape(-1).astype(y.dtype, copy=False)
    h = shuffle_reverse(sd, f"{p}.shuffle", y)
    h = double_coupling_reverse(sd, f"{p}.coupling", h, cond, _block_mode(i, conditional))
    if conditional:
        h = inv_leaky_reverse(h)
    return actnorm_reverse(loc, scale, h)


def n_flows_of(sd):
    n = 0
    while f"flow.sub_layers.{n}.norm_layer.loc" in sd:
        n += 1
    return n


# ------------------------------------------------------------------------ stacks
def flow_forward(sd, x, cond=None, dtype=np.float32, return_blocks=False):
    """ConditionalFlow.forward (flow_blocks.py:37-57) / UnconditionalFlow2.forward (:448-459).

    x: [B,C]; cond: [B,E] or None (unconditional).  Returns z [B,C], logdet [B]
    (and the per-block lists ``last_outs`` / ``last_logdets`` when asked).
    The wrappers' 4-D reshapes (INN.py:65-82, 278-286) are pure views and are
    left to the caller.
    """
    conditional = cond is not None
    x = np.asarray(x, dtype=dtype)
    cond = None if cond is None else np.asarray(cond, dtype=dtype)
    logdet = np.zeros(x.shape[0], dtype=dtype)
    outs, lds = [], []
    for i in range(n_flows_of(sd)):
        x, ld = block_forward(sd, i, x, cond, conditional)
        logdet = logdet + ld
        outs.append(x)
        lds.append(logdet)
    if return_blocks:
        return x, logdet, outs, lds
    return x, logdet


def flow_reverse(sd, z, cond=None, dtype=np.float32):
    """ConditionalFlow.forward(reverse=True) (flow_blocks.py:58-63) / UnconditionalFlow2 (:460-462)."""
    conditional = cond is not None
    x = np.asarray(z, dtype=dtype)
    cond = None if cond is None else np.asarray(cond, dtype=dtype)
    for i in reversed(range(n_flows_of(sd))):
        x = block_reverse(sd, i, x, cond, conditional)
    return x


def flow_forward_with_actnorm_init(sd, x, cond=None, dtype=np.float32):
    """First-call behaviour: every ActNorm with ``initialized == 0`` takes its
    statistics from its own input before being applied (modules.py:288-290), block
    after block.  Mutates ``sd`` (loc/scale/initialized) like the reference does."""
    conditional = cond is not None
    x = np.asarray(x, dtype=dtype)
    cond = None if cond is None else np.asarray(cond, dtype=dtype)
    logdet = np.zeros(x.shape[0], dtype=dtype)
    for i in range(n_flows_of(sd)):
        p = f"flow.sub_layers.{i}.norm_layer"
        if int(sd[f"{p}.initialized"]) == 0:
            loc, scale = actnorm_initialize(x)
            sd[f"{p}.loc"] = loc.reshape(1, -1, 1, 1).astype(np.float32)
            sd[f"{p}.scale"] = scale.reshape(1, -1, 1, 1).astype(np.float32)
            sd[f"{p}.initialized"] = np.array(1, dtype=np.uint8)
        x, ld = block_forward(sd, i, x, cond, conditional)
        logdet = logdet + ld
    return x, logdet


# -------------------------------------------------------------------------- loss
def nll(z):
    """nll -- models/modules/INN/loss.py:86-97 (radial=False, spatial_mean=False)."""
    return 0.5 * np.sum(z.reshape(z.shape[0], -1) ** 2, axis=1)


def flow_loss(z, logdet, nll_weight=1.0, logdet_weight=1.0):
    """FlowLoss.forward -- loss.py:17-42 (deterministic part; the reference also draws a
    randn_like sample for a logged reference NLL, which does not enter the loss)."""
    nll_loss = np.mean(nll(z))
    nlogdet_loss = -np.mean(logdet)
    return nll_weight * nll_loss + logdet_weight * nlogdet_loss, nll_loss, nlogdet_loss


# ---------------------------------------------------------------------- backward
def _fc_forward_cache(sd, prefix, x, use_tanh):
    idx = _linear_indices(sd, prefix)
    h = _R(x)
    acts = [h]
    for n, i in enumerate(idx):
        w = sd[f"{prefix}.main.{i}.weight"].astype(x.dtype, copy=False)
        b = sd[f"{prefix}.main.{i}.bias"].astype(x.dtype, copy=False)
        h = h @ _R(w).T + b
        if n + 1 < len(idx):
            if _KINK_LOG is not None:      # per-sample distance of the closest pre-activation to LeakyReLU's kink
                rms = np.sqrt(np.mean(h * h)) + 1e-30
                _KINK_LOG.append(np.min(np.abs(h), axis=1) / rms)
            h = _R(np.where(h >= 0, h, h * x.dtype.type(LEAKY_MLP_SLOPE)))
            acts.append(h)
    out = np.tanh(h) if use_tanh else h
    return out, acts


def _fc_backward(sd, prefix, acts, out, dout, use_tanh, grads):
    """Reverse-mode through BasicFullyConnectedNet (autograd over modules.py:190-191)."""
    idx = _linear_indices(sd, prefix)
    dt = dout.dtype
    g = _R(dout * (1.0 - out * out) if use_tanh else dout)
    for n in reversed(range(len(idx))):
        i = idx[n]
        w = sd[f"{prefix}.main.{i}.weight"].astype(dt, copy=False)
        a = acts[n]
        grads[f"{prefix}.main.{i}.weight"] = grads.get(f"{prefix}.main.{i}.weight", 0) + g.T @ a
        grads[f"{prefix}.main.{i}.bias"] = grads.get(f"{prefix}.main.{i}.bias", 0) + g.sum(axis=0)
        g = g @ _R(w)
        if n > 0:
            # torch's leaky_relu backward: input > 0 ? grad : grad * slope (a has the sign of the pre-activation)
            g = _R(g * np.where(a > 0, dt.type(1.0), dt.type(LEAKY_MLP_SLOPE)))
    return g


def flow_backward(sd, x, cond, dz, dlogdet, dtype=np.float64):
    """Gradients of  sum(dz*z) + sum(dlogdet*logdet)  w.r.t. every parameter, x and cond.

    The reference has no backward code of its own (autograd over the forward ops,
    SURVEY.md section 8a); this is the hand-derived reverse pass of flow_forward, checked
    against torch.autograd on the reference modules in tests/golden/make_golden.py.
    Returns (grads: dict keyed like sd, dx [B,C], dcond [B,E] or None).
    """
    conditional = cond is not None
    dt = np.dtype(dtype)
    x = np.asarray(x, dtype=dt)
    cond = None if cond is None else np.asarray(cond, dtype=dt)
    n = n_flows_of(sd)
    half = x.shape[1] // 2
    # ---- forward with caches
    tape = []
    h = x
    for i in range(n):
        p = f"flow.sub_layers.{i}"
        loc = sd[f"{p}.norm_layer.loc"].reshape(-1).astype(dt)
        scale = sd[f"{p}.norm_layer.scale"].reshape(-1).astype(dt)
        x_in = h
        a = scale[None] * (x_in + loc[None])
        if conditional:
            lmask = np.where(a >= 0, 1.0, INV_LEAKY_ALPHA).astype(dt)
            a = a * lmask
        else:
            lmask = None
        mode = _block_mode(i, conditional)
        steps = []
        cur = a
        for j in range(2):
            if j % 2 != 0:
                cur = np.concatenate((cur[:, half:], cur[:, :half]), axis=1)
            xa, xk = cur[:, :half], cur[:, half:]
            u = _coupling_input(xa, cond, mode)
            s_out, s_acts = _fc_forward_cache(sd, f"{p}.coupling.s.{j}", u, True)
            t_out, t_acts = _fc_forward_cache(sd, f"{p}.coupling.t.{j}", u, False)
            es = np.exp(s_out)
            steps.append((xk, es, s_out, s_acts, t_out, t_acts))
            cur = np.concatenate((xa, xk * es + t_out), axis=1)
        fidx = sd[f"{p}.shuffle.forward_shuffle_idx"]
        h = cur[:, fidx]
        tape.append((x_in, loc, scale, lmask, mode, steps, fidx))
    # ---- backward
    grads = {}
    g = np.asarray(dz, dtype=dt).reshape(x.shape)
    dld = np.asarray(dlogdet, dtype=dt)
    dcond = None if cond is None else np.zeros_like(cond)
    for i in reversed(range(n)):
        p = f"flow.sub_layers.{i}"
        x_in, loc, scale, lmask, mode, steps, fidx = tape[i]
        gc = np.zeros_like(g)
        gc[:, fidx] = g                                   # un-shuffle the gradient
        for j in reversed(range(2)):
            xk, es, s_out, s_acts, t_out, t_acts = steps[j]
            ga, gk = gc[:, :half], gc[:, half:]
            d_t = gk
            d_s = gk * xk * es + dld[:, None]             # y = xk*exp(s)+t ; logdet += sum s
            d_xk = gk * es
            du = _fc_backward(sd, f"{p}.coupling.s.{j}", s_acts, s_out, d_s, True, grads)
            du = du + _fc_backward(sd, f"{p}.coupling.t.{j}", t_acts, t_out, d_t, False, grads)
            if mode == "normal":
                ga = ga + du[:, :half]
                dcond += du[:, half:]
            elif mode == "cond":
                dcond += du
            else:
                ga = ga + du
            gc = np.concatenate((ga, d_xk), axis=1)
            if j % 2 != 0:
                gc = np.concatenate((gc[:, half:], gc[:, :half]), axis=1)
        if lmask is not None:
            gc = gc * lmask
        # a = scale*(x+loc) ; logdet += sum_c log|scale_c|
        grads[f"{p}.norm_layer.scale"] = (np.sum(gc * (x_in + loc[None]), axis=0)
                                          + dld.sum() / scale).reshape(1, -1, 1, 1)
        grads[f"{p}.norm_layer.loc"] = np.sum(gc * scale[None], axis=0).reshape(1, -1, 1, 1)
        g = gc * scale[None]
    return grads, g, dcond


_KINK_LOG = None


def kink_distance(sd, x, cond=None, dtype=np.float64, chunk=0):
    """Per sample: min over every hidden unit of every subnetwork of |pre-activation| / rms(layer).

    LeakyReLU' is discontinuous at 0, so a gradient comparison between two arithmetics is only meaningful for
    samples whose pre-activations all stay clear of 0 by more than the forward error of the less exact one.
    Samples of the flow are independent (ActNorm is already initialised), so tests simply drop the others.
    `chunk` > 0 walks the rows in slices (bounded memory at the full shapes); `dtype` float32 is enough for picking
    samples (its error is two orders below any useful margin)."""
    global _KINK_LOG
    x = np.asarray(x, dtype=dtype)
    n = x.shape[0]
    step = chunk if chunk > 0 else n
    out = []
    try:
        for lo in range(0, n, step):
            _KINK_LOG = []
            xs = x[lo:lo + step]
            cs = None if cond is None else np.asarray(cond[lo:lo + step], dtype=dtype)
            flow_backward(sd, xs, cs, np.zeros_like(xs), np.zeros(xs.shape[0], dtype=dtype), dtype=dtype)
            out.append(np.min(np.stack(_KINK_LOG, axis=0), axis=0))
        return np.concatenate(out)
    finally:
        _KINK_LOG = None


# --------------------------------------------------------------------- utilities
def state_dict_to_numpy(module_state_dict):
    """torch state_dict -> {name: np.ndarray} (detached CPU copies)."""
    return {k: v.detach().cpu().numpy() for k, v in module_state_dict.items()}


def macs_per_sample(sd):
    """P_w of SURVEY.md section 8(d): number of nn.Linear weight elements."""
    return int(sum(v.size for k, v in sd.items() if ".main." in k and k.endswith(".weight")))
