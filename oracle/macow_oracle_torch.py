"""torch restatement of the reference's MaCow unit, op for op  --  TEST / BENCH INFRASTRUCTURE, NOT PRODUCT.

The sequence of tensor operations the reference launches for ``MaCowUnit`` (models/modules/INN/macow2.py:925-996):
F.pad + slice + F.conv2d, activation, weight-normed 1x1 F.conv2d, tanh / log / affine map per flow, and for the
inverse the same chain once PER LINE (macow2.py:222-289).  It exists so that ``tools/macow_bench.py`` can time "what
the reference's stock path launches" on the box's GPU and host cores next to the fused kernel (the reference itself
cannot travel to the GPU box); ``tests/test_macow_oracle.py`` pins it to the reference fixtures.  Takes the same flat
numpy / tensor state dict as oracle/macow_oracle.py.
"""
import torch
import torch.nn.functional as F

ORDERS = {"conv1": "A", "conv2": "B", "conv3": "C", "conv4": "D"}
STAGES = ("conv1", "conv2", "actnorm1", "conv3", "conv4", "actnorm2")


def _act(h, kind):
    return {"relu": F.relu, "elu": F.elu, "leaky_relu": lambda t: F.leaky_relu(t, 0.1)}[kind](h)


def _shift(x, order, kh, kw):
    """ShiftedConv2d's pad-and-cut (macow_utils.py:460-494)."""
    if order == "A":
        return F.pad(x, ((kw - 1) // 2, (kw - 1) // 2, kh, 0))[:, :, :-1]
    if order == "B":
        return F.pad(x, ((kw - 1) // 2, (kw - 1) // 2, 0, kh))[:, :, 1:]
    if order == "C":
        return F.pad(x, (kw, 0, (kh - 1) // 2, (kh - 1) // 2))[:, :, :, :-1]
    return F.pad(x, (0, kw, (kh - 1) // 2, (kh - 1) // 2))[:, :, :, 1:]


class TorchUnit:
    def __init__(self, sd, act, alpha, device="cpu", dtype=torch.float32):
        t = lambda a: torch.as_tensor(a).to(device=device, dtype=dtype)
        self.act, self.alpha = act, alpha
        self.flows = {}
        for n in ORDERS:
            v, g = t(sd[n + ".net.conv1x1.conv.weight_v"]), t(sd[n + ".net.conv1x1.conv.weight_g"])
            self.flows[n] = (t(sd[n + ".net.shift_conv.weight"]), v, g, t(sd[n + ".net.conv1x1.conv.bias"]))
        self.norms = {n: (t(sd[n + ".log_scale"]), t(sd[n + ".bias"])) for n in ("actnorm1", "actnorm2")}

    def _params(self, n, x, shifted=True):
        w, v, g, b = self.flows[n]
        if shifted:
            x = _shift(x, ORDERS[n], w.shape[2], w.shape[3])
        c = _act(F.conv2d(x, w), self.act)
        w1 = torch._weight_norm(v, g, 0)                      # recomputed per call, like the weight_norm hook
        mu, ls = F.conv2d(c, w1, b).chunk(2, dim=1)
        return mu, torch.tanh(ls * 0.5) * self.alpha + 1.0

    def forward(self, x):
        logdet = x.new_zeros(x.shape[0])
        for s in STAGES:
            if s.startswith("conv"):
                mu, scale = self._params(s, x)
                x = scale * x + mu
                logdet = logdet + scale.log().reshape(x.shape[0], -1).sum(1)
            else:
                ls, b = self.norms[s]
                x = x * ls.exp() + b
                logdet = logdet + ls.sum() * (x.shape[2] * x.shape[3])
        return x, logdet

    def _reverse_flow(self, n, z):
        w = self.flows[n][0]
        kh, kw = w.shape[2], w.shape[3]
        order = ORDERS[n]
        B, C, H, W = z.shape
        if order in "AB":                                     # backward_height, macow2.py:222-246
            cw = kw // 2
            out = z.new_zeros(B, C, H + kh, W + 2 * cw)
            for h in (range(H) if order == "A" else reversed(range(H))):
                s = h if order == "A" else h + 1
                mu, scale = self._params(n, out[:, :, s:s + kh], shifted=False)
                new = (z[:, :, h] - mu.squeeze(2)) / (scale.squeeze(2) + 1e-12)
                out[:, :, h + kh if order == "A" else h, cw:W + cw] = new
            return out[:, :, kh:, cw:cw + W] if order == "A" else out[:, :, :H, cw:cw + W]
        ch = kh // 2                                          # backward_width, macow2.py:271-289
        out = z.new_zeros(B, C, H + 2 * ch, W + kw)
        for w_ in (range(W) if order == "C" else reversed(range(W))):
            s = w_ if order == "C" else w_ + 1
            mu, scale = self._params(n, out[:, :, :, s:s + kw], shifted=False)
            new = (z[:, :, :, w_] - mu.squeeze(3)) / (scale.squeeze(3) + 1e-12)
            out[:, :, ch:H + ch, w_ + kw if order == "C" else w_] = new
        return out[:, :, ch:ch + H, kw:] if order == "C" else out[:, :, ch:ch + H, :W]

    def reverse(self, z):
        for s in reversed(STAGES):
            if s.startswith("conv"):
                z = self._reverse_flow(s, z)
            else:
                ls, b = self.norms[s]
                z = (z - b) / (ls.exp() + 1e-8)
        return z
