"""CPU oracle for the full MaCow flow  --  TEST INFRASTRUCTURE, NOT PRODUCT.

A numpy restatement of the reference's multi-scale masked-convolution flow
(altmann144/ipoke-main-opticalflow, models/modules/INN): ``NICE2d`` with the convolutional coupling net
(``NICEConvBlock``), ``MaCowStep``, ``MultiScalePrior``, ``MultiScaleInternal`` and the wrapper
``UnsupervisedMaCowTransformer3`` (INN.py:365-390), unconditional (``h_channels = 0``), affine transforms,
``Shuffle`` permutations (``use1x1`` false), no normalisation layers, no attention, dropout 0 -- what
``config/second_stage.yaml:53-80`` selects.  The MaCow units come from ``oracle/macow_oracle.py``.
Only ``tests/`` and the CPU-baseline leg of ``tools/macow_bench.py`` may import this module.

Pinning: the reference ships no tests for this path; the functions below are pinned against outputs of the
reference itself (``tests/golden/make_macow_stack_golden.py`` -> ``tests/golden/macow_stack_*.npz``) by
``tests/test_macow_stack_oracle.py``.

``sd`` is a flat dict of numpy arrays keyed like the reference's ``state_dict()`` below the prefix given.
Paths in citations are relative to /root/reference/models/modules/INN.
"""
from __future__ import annotations

import numpy as np

from . import macow_oracle as mo


def sub(sd, prefix):
    """The entries of ``sd`` below ``prefix`` (which ends with a dot), prefix removed."""
    n = len(prefix)
    return {k[n:]: v for k, v in sd.items() if k.startswith(prefix)}


def conv2d_same(x, w):
    """nn.Conv2d with an odd square kernel and padding = k // 2, no bias: x [B, C, H, W], w [O, C, k, k]."""
    B, C, H, W = x.shape
    k = w.shape[2]
    p = k // 2
    xp = np.zeros((B, C, H + 2 * p, W + 2 * p), dtype=x.dtype)
    xp[:, :, p:p + H, p:p + W] = x
    out = np.zeros((B, w.shape[0], H, W), dtype=x.dtype)
    for dy in range(k):
        for dx in range(k):
            out += np.einsum("oc,bchw->bohw", w[:, :, dy, dx], xp[:, :, dy:dy + H, dx:dx + W])
    return out


def weight_norm(v, g):
    """nn.utils.weight_norm, dim = 0 (macow_utils.py:230-233): w = g * v / ||v||, the norm over all axes but the first."""
    nrm = np.sqrt((v.reshape(v.shape[0], -1) ** 2).sum(1)).reshape((-1,) + (1,) * (v.ndim - 1))
    return v * (g / nrm)


def nice_geometry(C, factor=2, split_type="continuous", up=True):
    """NICE2d.__init__ (macow2.py:303-320): channels transformed, channels the net reads, z1_channels, split type."""
    if split_type == "skip" and C % factor == 1:
        split_type = "continuous"
    c_t = C // factor
    c_in = C - c_t
    return c_t, c_in, (c_in if up else c_t), split_type


def nice_split(x, z1_channels, split_type):
    """NICE2d.split (macow2.py:363-376)."""
    if split_type == "continuous":
        return x[:, :z1_channels], x[:, z1_channels:]
    return x[:, 0::2], x[:, 1::2]


def nice_unsplit(z1, z2, split_type):
    """NICE2d.unsplit (macow2.py:378-389)."""
    z = np.concatenate([z1, z2], axis=1)
    if split_type == "continuous":
        return z
    dim = z1.shape[1]
    idx = [i // 2 if i % 2 == 0 else i // 2 + dim for i in range(dim * 2)]
    return z[:, idx]


def nice_params(sd, z, act, alpha, dtype):
    """NICEConvBlock.forward without norms / attention / conditioning (macow_utils.py:312-338): 3x3 conv (no bias)
    -> activation -> 1x1 conv (no bias) -> activation -> weight-normed 3x3 conv with bias; then
    Affine.calc_params (macow_utils.py:49-52)."""
    h = mo.activation(conv2d_same(z, sd["net.conv1.weight"].astype(dtype)), act)
    h = mo.activation(conv2d_same(h, sd["net.conv2.weight"].astype(dtype)), act)
    w3 = weight_norm(sd["net.conv3.conv.weight_v"].astype(dtype), sd["net.conv3.conv.weight_g"].astype(dtype))
    p = conv2d_same(h, w3) + sd["net.conv3.conv.bias"].astype(dtype).reshape(1, -1, 1, 1)
    c_t = p.shape[1] // 2
    mu, log_scale = p[:, :c_t], p[:, c_t:]
    return mu, np.tanh(log_scale * dtype(0.5)) * dtype(alpha) + dtype(1.0)


def nice_apply(sd, x, up, split_type, factor=2, act="elu", alpha=1.0, dtype=np.float64, reverse=False):
    """NICE2d.forward / backward_analytic (macow2.py:397-437) with Affine.fwd / bwd (macow_utils.py:54-66)."""
    x = np.asarray(x, dtype=dtype)
    _, _, z1c, split_type = nice_geometry(x.shape[1], factor, split_type, up)
    z1, z2 = nice_split(x, z1c, split_type)
    z, zp = (z1, z2) if up else (z2, z1)
    mu, scale = nice_params(sd, z, act, alpha, dtype)
    if reverse:
        zp = (zp - mu) / (scale + dtype(mo.AFFINE_BWD_EPS))
    else:
        zp = scale * zp + mu
    z1, z2 = (z, zp) if up else (zp, z)
    out = nice_unsplit(z1, z2, split_type)
    if reverse:
        return out
    return out, np.log(scale).reshape(x.shape[0], -1).sum(1)


def shuffle(sd, x, reverse=False):
    """Shuffle on [B, C, H, W] (flow_blocks.py:322-326): x[:, idx]; log-determinant 0."""
    return x[:, sd["backward_shuffle_idx" if reverse else "forward_shuffle_idx"]]


STEP_COUPLINGS = (("coupling1_up", True, "continuous"), ("coupling1_dn", False, "continuous"),
                  ("coupling2_up", True, "skip"), ("coupling2_dn", False, "skip"))        # macow2.py:1025-1057


def step_forward(sd, x, act="elu", alpha=1.0, dtype=np.float64, h=None):
    """MaCowStep.forward (macow2.py:1063-1091): actnorm1, Shuffle, 2 units, NICE up / down (continuous split),
    actnorm2, 2 units, NICE up / down (skip split)."""
    out, logdet = mo.actnorm2d_forward(sd, "actnorm1", x, dtype)
    out = shuffle(sub(sd, "conv1x1."), out)
    for part, an in ((1, None), (2, "actnorm2")):
        if an:
            out, ld = mo.actnorm2d_forward(sd, an, out, dtype)
            logdet = logdet + ld
        for u in range(2):
            out, ld = mo.unit_forward(sub(sd, f"units{part}.{u}."), out, act, alpha, dtype, h=h)
            logdet = logdet + ld
        for name, up, st in STEP_COUPLINGS[2 * (part - 1):2 * part]:
            out, ld = nice_apply(sub(sd, name + "."), out, up, st, 2, act, alpha, dtype)
            logdet = logdet + ld
    return out, logdet


def step_reverse(sd, z, act="elu", alpha=1.0, dtype=np.float64, h=None):
    """MaCowStep.forward(reverse=True) (macow2.py:1092-1117)."""
    out = np.asarray(z, dtype=dtype)
    for part, an in ((2, "actnorm2"), (1, None)):
        for name, up, st in reversed(STEP_COUPLINGS[2 * (part - 1):2 * part]):
            out = nice_apply(sub(sd, name + "."), out, up, st, 2, act, alpha, dtype, reverse=True)
        for u in (1, 0):
            out = mo.unit_reverse(sub(sd, f"units{part}.{u}."), out, act, alpha, dtype, h=h)
        if an:
            out = mo.actnorm2d_reverse(sd, an, out, dtype)
    out = shuffle(sub(sd, "conv1x1."), out, reverse=True)
    return mo.actnorm2d_reverse(sd, "actnorm1", out, dtype)


def prior_forward(sd, x, factor, act="elu", alpha=1.0, dtype=np.float64):
    """MultiScalePrior.forward (macow2.py:571-582): Shuffle, NICE up (continuous, C // factor channels transformed),
    ActNorm2dFlow on the transformed channels."""
    out = shuffle(sub(sd, "conv1x1."), np.asarray(x, dtype=dtype))
    out, logdet = nice_apply(sub(sd, "coupling."), out, True, "continuous", factor, act, alpha, dtype)
    z1c = x.shape[1] - x.shape[1] // factor
    out2, ld = mo.actnorm2d_forward(sd, "actnorm", out[:, z1c:], dtype)
    return np.concatenate([out[:, :z1c], out2], axis=1), logdet + ld


def prior_reverse(sd, z, factor, act="elu", alpha=1.0, dtype=np.float64):
    """MultiScalePrior.forward(reverse=True) (macow2.py:583-594)."""
    z = np.asarray(z, dtype=dtype)
    z1c = z.shape[1] - z.shape[1] // factor
    out = np.concatenate([z[:, :z1c], mo.actnorm2d_reverse(sd, "actnorm", z[:, z1c:], dtype)], axis=1)
    out = nice_apply(sub(sd, "coupling."), out, True, "continuous", factor, act, alpha, dtype, reverse=True)
    return shuffle(sub(sd, "conv1x1."), out, reverse=True)


def levels(C, num_steps, factor):
    """Channel bookkeeping of MultiScaleInternal.__init__ (macow2.py:841-868): per level (channels in, prior factor)."""
    step = C // factor
    out = []
    for n in num_steps:
        out.append((C, factor, n))
        C, factor = C - step, factor - 1
    return out


def flow_forward(sd, x, num_steps, factor, act="elu", alpha=1.0, dtype=np.float64, h=None):
    """MultiScaleInternal.forward (macow2.py:872-897) below the prefix of its ``layers / priors / shuffle_layers``.
    ``h``: the conditioning of SupervisedMacowTransformer (INN.py:447-482) with ``condition_nice`` false: it reaches the
    masked flows of the units only (the couplings ignore it, macow_utils.py:330)."""
    out = np.asarray(x, dtype=dtype)
    logdet = np.zeros(out.shape[0], dtype=dtype)
    outputs = []
    for i, (C, f, n) in enumerate(levels(out.shape[1], num_steps, factor)):
        for s in range(n):
            out, ld = step_forward(sub(sd, f"layers.{i}.{s}."), out, act, alpha, dtype, h=h)
            logdet = logdet + ld
        out, ld = prior_forward(sub(sd, f"priors.{i}."), out, f, act, alpha, dtype)
        logdet = logdet + ld
        out = shuffle(sub(sd, f"shuffle_layers.{i}."), out)
        z1c = C - C // f
        outputs.append(out[:, z1c:])
        out = out[:, :z1c]
    outputs.append(out)
    outputs.reverse()
    return np.concatenate(outputs, axis=1), logdet


def flow_reverse(sd, z, num_steps, factor, act="elu", alpha=1.0, dtype=np.float64, h=None):
    """MultiScaleInternal.forward(reverse=True) (macow2.py:898-918)."""
    out = np.asarray(z, dtype=dtype)
    lv = levels(out.shape[1], num_steps, factor)
    outputs = []
    for C, f, _ in lv:
        z1c = C - C // f
        outputs.append(out[:, z1c:])
        out = out[:, :z1c]
    for i in range(len(lv) - 1, -1, -1):
        C, f, n = lv[i]
        out = np.concatenate([out, outputs.pop()], axis=1)
        out = shuffle(sub(sd, f"shuffle_layers.{i}."), out, reverse=True)
        out = prior_reverse(sub(sd, f"priors.{i}."), out, f, act, alpha, dtype)
        for s in range(n - 1, -1, -1):
            out = step_reverse(sub(sd, f"layers.{i}.{s}."), out, act, alpha, dtype, h=h)
    return out
