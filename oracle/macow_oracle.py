"""CPU oracle for the MaCow unit  --  TEST INFRASTRUCTURE, NOT PRODUCT.

A numpy restatement of the reference's masked-convolution flow unit
(altmann144/ipoke-main-opticalflow, models/modules/INN/macow2.py and macow_utils.py):
``MaskedConvFlow`` orders A / B / C / D with the affine transform, ``ActNorm2dFlow`` and their
composition ``MaCowUnit``, forward (out, logdet) and the sequential line-by-line inverse.
Only ``tests/`` and the CPU-baseline leg of ``tools/macow_bench.py`` may import this module; the shipped
path (``ipoke-main-opticalflow_b200/macow.py`` -> CUDA) never does.

Pinning: the reference ships no tests for this path, so the oracle is pinned against outputs of the
reference itself, generated in the build container by ``tests/golden/make_macow_golden.py`` and committed
as ``tests/golden/macow_*.npz``; ``tests/test_macow_oracle.py`` checks every function below against them.

``sd`` is a flat dict of numpy arrays keyed like the reference's ``MaCowUnit.state_dict()``
(``conv1.net.shift_conv.weight`` ...).  Paths in citations are relative to /root/reference/models/modules/INN.
"""
from __future__ import annotations

import numpy as np

ORDERS = {"conv1": "A", "conv2": "B", "conv3": "C", "conv4": "D"}       # macow2.py:938-956
LEAKY_SLOPE = 0.1                                                        # macow_utils.py:424
AFFINE_BWD_EPS = 1e-12                                                   # macow_utils.py:64
ACTNORM_BWD_EPS = 1e-8                                                   # macow2.py:520


def activation(h, kind):
    """MCFBlock activation (macow_utils.py:417-424): ReLU, ELU(alpha=1) or LeakyReLU(0.1)."""
    if kind == "relu":
        return np.maximum(h, 0)
    if kind == "elu":
        return np.where(h > 0, h, np.expm1(np.minimum(h, 0)))
    if kind == "leaky_relu":
        return np.where(h > 0, h, LEAKY_SLOPE * h)
    raise ValueError(kind)


def weight_norm_1x1(sd, prefix, dtype):
    """Effective 1x1 weights [2C, Hd] of Conv2dWeightNorm (macow_utils.py:230-233: nn.utils.weight_norm with
    dim=0, i.e. w = g * v / ||v|| with the norm over every axis but the first) and its bias [2C]."""
    v = sd[prefix + ".conv1x1.conv.weight_v"].astype(dtype)
    g = sd[prefix + ".conv1x1.conv.weight_g"].astype(dtype)
    nrm = np.sqrt((v.reshape(v.shape[0], -1) ** 2).sum(1)).reshape(-1, 1, 1, 1)
    w = v * (g / nrm)
    return w.reshape(w.shape[0], -1), sd[prefix + ".conv1x1.conv.bias"].astype(dtype)


def tap_origin(order, kh, kw):
    """(oy, ox): the input pixel under kernel tap (0, 0) for the output pixel (0, 0).

    ShiftedConv2d (macow_utils.py:460-497) pads and cuts the image so that a valid kH x kW convolution sees,
    for output (y, x), rows y-kH .. y-1 (A), y+1 .. y+kH (B), or columns x-kW .. x-1 (C), x+1 .. x+kW (D),
    centred in the other axis."""
    if order == "A":
        return -kh, -((kw - 1) // 2)
    if order == "B":
        return 1, -((kw - 1) // 2)
    if order == "C":
        return -((kh - 1) // 2), -kw
    if order == "D":
        return -((kh - 1) // 2), 1
    raise ValueError(order)


def shifted_conv(x, w, order, rows=None, cols=None):
    """ShiftedConv2d.forward(shifted=True) (macow_utils.py:489-497), no bias (macow_utils.py:415).
    x [B, C, H, W], w [Hd, C, kH, kW]  ->  [B, Hd, H, W]; pixels outside the image read as zero.
    ``rows`` / ``cols`` = (first, count) restrict the OUTPUT pixels (the inverse needs one line at a time)."""
    B, C, H, W = x.shape
    hd, _, kh, kw = w.shape
    oy, ox = tap_origin(order, kh, kw)
    y0, nh = rows if rows is not None else (0, H)
    x0, nw = cols if cols is not None else (0, W)
    pad = max(kh, kw) + 1
    xp = np.zeros((B, C, H + 2 * pad, W + 2 * pad), dtype=x.dtype)
    xp[:, :, pad:pad + H, pad:pad + W] = x
    out = np.zeros((B, hd, nh, nw), dtype=x.dtype)
    for dy in range(kh):
        for dx in range(kw):
            ys, xs = pad + oy + dy + y0, pad + ox + dx + x0
            out += np.einsum("jc,bchw->bjhw", w[:, :, dy, dx], xp[:, :, ys:ys + nh, xs:xs + nw])
    return out


def mcf_params(sd, name, x, act, alpha, dtype, rows=None, cols=None, h=None):
    """MCFBlock.forward (macow_utils.py:429-436: shifted conv -> [concatenate the conditioning h on the channel axis]
    -> activation -> weight-normed 1x1 conv; dropout p = 0) followed by Affine.calc_params (macow_utils.py:49-52):
    mu, scale = tanh(log_scale / 2) * alpha + 1.  ``h`` [B, h_channels, H, W] is cut to the same output pixels."""
    prefix = name + ".net"
    w = sd[prefix + ".shift_conv.weight"].astype(dtype)
    hid = shifted_conv(x, w, ORDERS[name], rows, cols)
    if h is not None:
        hh = np.asarray(h, dtype=dtype)
        if rows is not None:
            hh = hh[:, :, rows[0]:rows[0] + rows[1], :]
        if cols is not None:
            hh = hh[:, :, :, cols[0]:cols[0] + cols[1]]
        hid = np.concatenate([hid, hh], axis=1)
    hid = activation(hid, act)
    w1, b1 = weight_norm_1x1(sd, prefix, dtype)
    p = np.einsum("oj,bjhw->bohw", w1, hid) + b1.reshape(1, -1, 1, 1)
    C = x.shape[1]
    mu, log_scale = p[:, :C], p[:, C:]
    return mu, np.tanh(log_scale * dtype(0.5)) * dtype(alpha) + dtype(1.0)


def mcf_forward(sd, name, x, act="elu", alpha=1.0, dtype=np.float64, h=None):
    """MaskedConvFlow.forward(reverse=False) (macow2.py:116-119) with Affine.fwd (macow_utils.py:54-59)."""
    x = np.asarray(x, dtype=dtype)
    mu, scale = mcf_params(sd, name, x, act, alpha, dtype, h=h)
    return scale * x + mu, np.log(scale).reshape(x.shape[0], -1).sum(1)


def mcf_reverse(sd, name, z, act="elu", alpha=1.0, dtype=np.float64, h=None):
    """MaskedConvFlow.backward_analytic (macow2.py:139-151): one line at a time, rows top-down (A), bottom-up
    (B), columns left-right (C), right-left (D) (backward_height macow2.py:222-246, backward_width :271-289);
    each line is Affine.bwd (macow_utils.py:61-66) of the input line with parameters computed from the lines
    already reconstructed."""
    z = np.asarray(z, dtype=dtype)
    order = ORDERS[name]
    B, C, H, W = z.shape
    out = np.zeros_like(z)
    n = H if order in "AB" else W
    lines = range(n) if order in "AC" else range(n - 1, -1, -1)
    for l in lines:                                                   # causal: line l sees finished lines only
        if order in "AB":
            mu, scale = mcf_params(sd, name, out, act, alpha, dtype, rows=(l, 1), h=h)
            out[:, :, l:l + 1, :] = (z[:, :, l:l + 1, :] - mu) / (scale + dtype(AFFINE_BWD_EPS))
        else:
            mu, scale = mcf_params(sd, name, out, act, alpha, dtype, cols=(l, 1), h=h)
            out[:, :, :, l:l + 1] = (z[:, :, :, l:l + 1] - mu) / (scale + dtype(AFFINE_BWD_EPS))
    return out


def actnorm2d_forward(sd, name, x, dtype=np.float64):
    """ActNorm2dFlow.forward (macow2.py:502-514): x * exp(log_scale) + bias, logdet = H * W * sum(log_scale)."""
    x = np.asarray(x, dtype=dtype)
    ls = sd[name + ".log_scale"].astype(dtype)
    out = x * np.exp(ls)[None] + sd[name + ".bias"].astype(dtype)[None]
    return out, np.full(x.shape[0], ls.sum() * (x.shape[2] * x.shape[3]), dtype=dtype)


def actnorm2d_reverse(sd, name, z, dtype=np.float64):
    """ActNorm2dFlow.forward(reverse=True) (macow2.py:515-521): (z - bias) / (exp(log_scale) + 1e-8)."""
    z = np.asarray(z, dtype=dtype)
    ls = sd[name + ".log_scale"].astype(dtype)
    return (z - sd[name + ".bias"].astype(dtype)[None]) / (np.exp(ls)[None] + dtype(ACTNORM_BWD_EPS))


def actnorm2d_init(x, dtype=np.float64):
    """ActNorm2dFlow.init (macow2.py:526-541) for freshly constructed parameters is applied to
    data * exp(log_scale) + bias; this helper returns (log_scale, bias) for data already in that form:
    per-channel mean / unbiased std over batch and pixels, log_scale = log(1 / (std + 1e-6)), bias = -mean / (std + 1e-6)."""
    x = np.asarray(x, dtype=dtype)
    flat = x.transpose(1, 0, 2, 3).reshape(x.shape[1], -1)
    inv = 1.0 / (flat.std(axis=1, ddof=1) + 1e-6)
    return np.log(inv).reshape(-1, 1, 1), (-flat.mean(axis=1) * inv).reshape(-1, 1, 1)


STAGES = ("conv1", "conv2", "actnorm1", "conv3", "conv4", "actnorm2")     # macow2.py:965-982


def unit_forward(sd, x, act="elu", alpha=1.0, dtype=np.float64, stages=STAGES, h=None):
    """MaCowUnit.forward(reverse=False) (macow2.py:963-982); ``h``: the conditioning every masked flow receives."""
    out = np.asarray(x, dtype=dtype)
    logdet = np.zeros(out.shape[0], dtype=dtype)
    for s in stages:
        if s.startswith("conv"):
            out, ld = mcf_forward(sd, s, out, act, alpha, dtype, h=h)
        else:
            out, ld = actnorm2d_forward(sd, s, out, dtype)
        logdet = logdet + ld
    return out, logdet


def unit_reverse(sd, z, act="elu", alpha=1.0, dtype=np.float64, stages=STAGES, h=None):
    """MaCowUnit.forward(reverse=True) (macow2.py:983-996): the stages in the opposite order, each inverted."""
    out = np.asarray(z, dtype=dtype)
    for s in reversed(stages):
        if s.startswith("conv"):
            out = mcf_reverse(sd, s, out, act, alpha, dtype, h=h)
        else:
            out = actnorm2d_reverse(sd, s, out, dtype)
    return out
