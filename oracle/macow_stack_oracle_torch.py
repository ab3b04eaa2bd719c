"""torch restatement of the reference's full MaCow flow, op for op  --  TEST / BENCH INFRASTRUCTURE, NOT PRODUCT.

The tensor operations the reference launches for ``UnsupervisedMaCowTransformer3`` (models/modules/INN/INN.py:365-390;
``MultiScaleInternal`` macow2.py:872-918, ``MaCowStep`` :1063-1117, ``MultiScalePrior`` :571-594, ``NICE2d`` :397-437
with ``NICEConvBlock`` macow_utils.py:312-338): F.conv2d 3x3 / 1x1 / weight-normed 3x3, activations, channel
index_select / cat for the splits and Shuffles, and the units of oracle/macow_oracle_torch.py.  It exists so that
``tools/macow_flow_bench.py`` can time "what the reference's stock path launches" on the box's GPU and host cores next to
the CUDA path (the reference itself cannot travel to the GPU box); ``tests/test_macow_stack_oracle.py`` pins it to the
reference fixtures.  Takes the flat state dict below the ``flow.`` prefix, like oracle/macow_stack_oracle.py.
"""
import torch
import torch.nn.functional as F

from .macow_oracle_torch import TorchUnit, _act
from .macow_stack_oracle import STEP_COUPLINGS, levels, nice_geometry, sub


class TorchNice:
    def __init__(self, sd, C, up, split_type, factor, act, alpha, device, dtype):
        t = lambda a: torch.as_tensor(a).to(device=device, dtype=dtype)
        self.w1, self.w2 = t(sd["net.conv1.weight"]), t(sd["net.conv2.weight"])
        self.v, self.g, self.b = t(sd["net.conv3.conv.weight_v"]), t(sd["net.conv3.conv.weight_g"]), t(sd["net.conv3.conv.bias"])
        self.up, self.act, self.alpha = up, act, alpha
        _, _, self.z1c, self.split_type = nice_geometry(C, factor, split_type, up)
        if self.split_type == "skip":                                        # macow2.py:368-373, 383-386
            self.idx1 = torch.arange(0, C, 2, device=device)
            self.idx2 = torch.arange(1, C, 2, device=device)
            self.unidx = torch.tensor([i // 2 if i % 2 == 0 else i // 2 + C // 2 for i in range(C)], device=device)

    def _split(self, x):
        if self.split_type == "continuous":
            return x.split([self.z1c, x.shape[1] - self.z1c], dim=1)
        return x.index_select(1, self.idx1), x.index_select(1, self.idx2)

    def _unsplit(self, z1, z2):
        z = torch.cat([z1, z2], dim=1)
        return z if self.split_type == "continuous" else z.index_select(1, self.unidx)

    def _params(self, z):
        h = _act(F.conv2d(z, self.w1, padding=1), self.act)
        h = _act(F.conv2d(h, self.w2), self.act)
        p = F.conv2d(h, torch._weight_norm(self.v, self.g, 0), self.b, padding=1)
        mu, ls = p.chunk(2, dim=1)
        return mu, torch.tanh(ls * 0.5) * self.alpha + 1.0

    def forward(self, x):
        z1, z2 = self._split(x)
        z, zp = (z1, z2) if self.up else (z2, z1)
        mu, scale = self._params(z)
        zp = scale * zp + mu
        z1, z2 = (z, zp) if self.up else (zp, z)
        return self._unsplit(z1, z2), scale.log().reshape(x.shape[0], -1).sum(1)

    def reverse(self, x):
        z1, z2 = self._split(x)
        z, zp = (z1, z2) if self.up else (z2, z1)
        mu, scale = self._params(z)
        zp = (zp - mu) / (scale + 1e-12)
        z1, z2 = (z, zp) if self.up else (zp, z)
        return self._unsplit(z1, z2)


class TorchFlow:
    """``flow`` of UnsupervisedMaCowTransformer3 from a state dict: forward(x) -> (out, logdet), reverse(z) -> x."""

    def __init__(self, sd, C, num_steps, factor, act="elu", alpha=1.0, device="cpu", dtype=torch.float32):
        t = lambda a: torch.as_tensor(a).to(device=device, dtype=dtype)
        ti = lambda a: torch.as_tensor(a).to(device=device)
        self.levels = []
        for i, (Cl, f, n) in enumerate(levels(C, num_steps, factor)):
            steps = []
            for s in range(n):
                ssd = sub(sd, f"layers.{i}.{s}.")
                steps.append(dict(
                    an1=(t(ssd["actnorm1.log_scale"]), t(ssd["actnorm1.bias"])),
                    an2=(t(ssd["actnorm2.log_scale"]), t(ssd["actnorm2.bias"])),
                    perm=(ti(ssd["conv1x1.forward_shuffle_idx"]), ti(ssd["conv1x1.backward_shuffle_idx"])),
                    units=[[TorchUnit(sub(ssd, f"units{p}.{u}."), act, alpha, device, dtype) for u in range(2)] for p in (1, 2)],
                    nice=[TorchNice(sub(ssd, cn + "."), Cl, up, st, 2, act, alpha, device, dtype) for cn, up, st in STEP_COUPLINGS]))
            psd = sub(sd, f"priors.{i}.")
            prior = dict(perm=(ti(psd["conv1x1.forward_shuffle_idx"]), ti(psd["conv1x1.backward_shuffle_idx"])),
                         nice=TorchNice(sub(psd, "coupling."), Cl, True, "continuous", f, act, alpha, device, dtype),
                         an=(t(psd["actnorm.log_scale"]), t(psd["actnorm.bias"])), z1c=Cl - Cl // f)
            lsd = sub(sd, f"shuffle_layers.{i}.")
            self.levels.append((steps, prior, (ti(lsd["forward_shuffle_idx"]), ti(lsd["backward_shuffle_idx"]))))

    @staticmethod
    def _an(x, p):
        return x * p[0].exp() + p[1], p[0].sum() * (x.shape[2] * x.shape[3])

    @staticmethod
    def _an_inv(x, p):
        return (x - p[1]) / (p[0].exp() + 1e-8)

    def forward(self, x):
        logdet = x.new_zeros(x.shape[0])
        outputs = []
        for steps, prior, lperm in self.levels:
            for st in steps:
                x, ld = self._an(x, st["an1"]); logdet = logdet + ld
                x = x[:, st["perm"][0]]
                for part in (0, 1):
                    if part:
                        x, ld = self._an(x, st["an2"]); logdet = logdet + ld
                    for u in st["units"][part]:
                        x, ld = u.forward(x); logdet = logdet + ld
                    for nc in st["nice"][2 * part:2 * part + 2]:
                        x, ld = nc.forward(x); logdet = logdet + ld
            x = x[:, prior["perm"][0]]
            x, ld = prior["nice"].forward(x); logdet = logdet + ld
            o1, o2 = x[:, :prior["z1c"]], x[:, prior["z1c"]:]
            o2, ld = self._an(o2, prior["an"]); logdet = logdet + ld
            x = torch.cat([o1, o2], dim=1)[:, lperm[0]]
            outputs.append(x[:, prior["z1c"]:])
            x = x[:, :prior["z1c"]]
        outputs.append(x)
        outputs.reverse()
        return torch.cat(outputs, dim=1), logdet

    def reverse(self, z):
        outputs = []
        for _, prior, _ in self.levels:
            outputs.append(z[:, prior["z1c"]:])
            z = z[:, :prior["z1c"]]
        for steps, prior, lperm in reversed(self.levels):
            z = torch.cat([z, outputs.pop()], dim=1)[:, lperm[1]]
            z = torch.cat([z[:, :prior["z1c"]], self._an_inv(z[:, prior["z1c"]:], prior["an"])], dim=1)
            z = prior["nice"].reverse(z)[:, prior["perm"][1]]
            for st in reversed(steps):
                for part in (1, 0):
                    for nc in reversed(st["nice"][2 * part:2 * part + 2]):
                        z = nc.reverse(z)
                    for u in reversed(st["units"][part]):
                        z = u.reverse(z)
                    if part:
                        z = self._an_inv(z, st["an2"])
                z = self._an_inv(z[:, st["perm"][1]], st["an1"])
        return z
