"""The multi-scale MaCow flow on B200: drop-ins for the reference's ``NICE2d`` / ``NICEConvBlock``
(models/modules/INN/macow2.py:291-437, macow_utils.py:253-338), ``MaCowStep`` (macow2.py:999-1117),
``MultiScalePrior`` (macow2.py:543-594), ``MultiScaleInternal`` (macow2.py:821-918) and the wrapper
wrappers ``UnsupervisedMaCowTransformer3`` (INN.py:365-390) and ``SupervisedMacowTransformer`` (INN.py:447-482) with the
same constructors, attribute names, state_dict keys and RNG draw order, computing through the CUDA kernels behind
``include/ipoke_macow.h``.  Row N3 of SURVEY.md section 8(f).

Built: what ``config/second_stage.yaml:53-80`` selects -- 'affine' transforms, 'conv' couplings, Shuffle permutations
(``use1x1`` false), no normalisation layers / attention, ``cond_conv`` and ``condition_nice`` false, dropout 0;
unconditional, or conditional on ``h`` [B, h_channels, H, W] (which then reaches the masked flows of the units, as in
the reference with ``condition_nice`` false).  Anything else raises NotImplementedError.  Inference only (forward with logdet and the inverse
used for sampling); there is no CPU or eager fallback.

How a call runs: the image lives in ONE [B, C, H, W] buffer for the whole flow.  Every level works in place on the first
channels of each sample, so ``split2d`` / ``unsplit2d`` (macow2.py:15-23, 886-896) move no data: the channels a level
factors out already sit where the final concatenation wants them.  The module tree is flattened once into a program of
three kinds of kernel calls -- per-channel glue (ActNorm2dFlow and / or Shuffle, one launch), a chain of MaCow units
(one launch per pair), a NICE coupling (three GEMM launches + the tap / affine kernel) -- and the inverse walks
the same program backwards.
"""
import ctypes

import torch
import torch.nn as nn

from . import _capi
from . import macow as _m
from ._nvtx import annotate
from .macow import (ACTIVATIONS, ActNorm2dFlow, Conv2dWeightNorm, MaCowUnit, NiceDesc, ProgramOwner,
                    _invalidate_after_load)
from .modules import Shuffle


def _unsupported(**options):
    for name, (value, allowed) in options.items():
        if value not in allowed:
            raise NotImplementedError(f"MaCow flow: {name}={value!r} is not built (supported: {allowed})")


class NICEConvBlock(nn.Module):
    """Parameter holder of macow_utils.py:253-338 (no norms, no attention, no conditioning): ``conv1`` 3x3 without bias,
    ``conv2`` 1x1 without bias, ``conv3`` the weight-normed zero-initialised 3x3 convolution with bias."""

    def __init__(self, in_channels, out_channels, hidden_channels, activation, normalize=None, num_groups=None,
                 h_channels=0, attention=False, spatial_size=8, heads=4, cond_conv=False,
                 cond_conv_hidden_channels=None, p_dropout=0.):
        super().__init__()
        assert activation in ACTIVATIONS
        _unsupported(normalize=(normalize, (None,)), h_channels=(h_channels, (0, None)),
                     attention=(bool(attention), (False,)), p_dropout=(float(p_dropout), (0.,)))
        self.cond = False
        self.attention = False
        self.cond_conv = cond_conv
        self.conv1 = nn.Conv2d(in_channels, hidden_channels, kernel_size=3, padding=1, bias=False)
        self.conv2 = nn.Conv2d(hidden_channels, hidden_channels, kernel_size=1, bias=False)
        self.conv3 = Conv2dWeightNorm(hidden_channels, out_channels, zero_init=True, kernel_size=3)

    def forward(self, x, h=None):
        raise RuntimeError("NICEConvBlock is a parameter container here: the coupling runs fused through NICE2d")


class NICE2d(ProgramOwner, nn.Module):
    """macow2.py:291-437 with the 'affine' transform.  ``forward(input, h=None, reverse=False)`` returns
    ``(out, logdet)`` or the inverse image, like the reference; inside a flow it is called through the flow's program."""

    def __init__(self, in_channels, hidden_channels=None, h_channels=0, split_type="continuous", order="up", factor=2,
                 transform="affine", alpha=1.0, type="conv", h_type=None, activation="relu", normalize=None,
                 num_groups=None, attention=False, heads=4, spatial_size=8, cond_conv=False,
                 cond_conv_hidden_channels=None, p_dropout=0.):
        super().__init__()
        _unsupported(transform=(transform, ("affine",)), type=(type, ("conv",)), h_type=(h_type, (None,)))
        self.in_channels = in_channels
        self.factor = factor
        assert split_type in ["continuous", "skip"]
        if split_type == "skip":
            assert factor == 2
            if in_channels % factor == 1:
                split_type = "continuous"
        assert order in ["up", "down"]
        self.split_type = split_type
        self.up = order == "up"
        if hidden_channels is None:
            hidden_channels = min(8 * in_channels, 512)
        c_t = in_channels // factor
        c_in = in_channels - c_t
        self.z1_channels = c_in if self.up else c_t
        self.analytic_bwd = True
        self.alpha = alpha
        self.net = NICEConvBlock(c_in, 2 * c_t, hidden_channels, activation, normalize=normalize, num_groups=num_groups,
                                 h_channels=h_channels, attention=attention, heads=heads, spatial_size=spatial_size,
                                 cond_conv=cond_conv, cond_conv_hidden_channels=cond_conv_hidden_channels,
                                 p_dropout=p_dropout)
        if split_type == "continuous":                      # macow2.py:366-367, 410-414
            geo = (c_in, 0, 1, c_t, c_in, 1) if self.up else (c_in, c_t, 1, c_t, 0, 1)
        else:                                               # even / odd channels, macow2.py:368-373
            geo = (c_in, 0, 2, c_t, 1, 2) if self.up else (c_in, 1, 2, c_t, 0, 2)
        self._desc = NiceDesc(*geo, hidden_channels, ACTIVATIONS[activation], alpha)
        self._program = None
        self.register_load_state_dict_post_hook(_invalidate_after_load)

    def params(self):
        n = self.net
        return [n.conv1.weight, n.conv2.weight, n.conv3.conv.bias, n.conv3.conv.weight_g, n.conv3.conv.weight_v]

    def forward(self, input, h=None, reverse=False):
        # built with h_channels = 0: like the reference's NICEConvBlock (macow_utils.py:330) it ignores a conditioning
        if self._program is None:
            self._program = Program([("nice", self)], self.in_channels, [self])
        return self._program.run(input, reverse)


class MaCowStep(ProgramOwner, nn.Module):
    """macow2.py:999-1117."""

    def __init__(self, in_channels, kernel_size, hidden_channels, h_channels, transform="affine", alpha=1.0,
                 coupling_type="conv", h_type=None, activation="relu", normalize=None, num_groups=None, **kwargs):
        super().__init__()
        _unsupported(condition_nice=(bool(kwargs.get("condition_nice", False)), (False,)),
                     cond_conv=(bool(kwargs.get("cond_conv", False)) and bool(h_channels), (False,)))
        self.h_channels = int(h_channels or 0)
        attention = bool(kwargs.get("attention", False))
        heads = kwargs.get("heads", 4)
        ssize = kwargs.get("spatial_size", 8)
        cond_conv = kwargs.get("cond_conv", False)
        cchc = kwargs.get("cond_conv_hidden_channels", None)
        p_dropout = kwargs.get("p_dropout", 0.)
        unit_kw = dict(h_channels=self.h_channels, transform=transform, alpha=alpha, h_type=h_type, activation=activation,
                       cond_conv=cond_conv, cond_conv_hidden_channels=cchc, p_dropout=p_dropout)
        nice_kw = dict(hidden_channels=hidden_channels, h_channels=0, transform=transform, alpha=alpha,
                       type=coupling_type, h_type=h_type, activation=activation, normalize=normalize,
                       num_groups=num_groups, attention=attention, heads=heads, spatial_size=ssize, cond_conv=cond_conv,
                       cond_conv_hidden_channels=cchc, p_dropout=p_dropout)
        self.in_channels = in_channels
        self.actnorm1 = ActNorm2dFlow(in_channels)
        self.conv1x1 = Shuffle(in_channels)
        self.units1 = nn.ModuleList([MaCowUnit(in_channels, kernel_size, **unit_kw) for _ in range(2)])
        self.coupling1_up = NICE2d(in_channels, split_type="continuous", order="up", **nice_kw)
        self.coupling1_dn = NICE2d(in_channels, split_type="continuous", order="down", **nice_kw)
        self.actnorm2 = ActNorm2dFlow(in_channels)
        self.units2 = nn.ModuleList([MaCowUnit(in_channels, kernel_size, **unit_kw) for _ in range(2)])
        self.coupling2_up = NICE2d(in_channels, split_type="skip", order="up", **nice_kw)
        self.coupling2_dn = NICE2d(in_channels, split_type="skip", order="down", **nice_kw)
        self._program = None
        self.register_load_state_dict_post_hook(_invalidate_after_load)

    def ops(self):
        """The step as forward kernel calls (macow2.py:1063-1091)."""
        C = self.in_channels
        return [("glue", self.conv1x1, C, self.actnorm1, 0, C),
                ("units", list(self.units1)), ("nice", self.coupling1_up), ("nice", self.coupling1_dn),
                ("glue", None, 0, self.actnorm2, 0, C),
                ("units", list(self.units2)), ("nice", self.coupling2_up), ("nice", self.coupling2_dn)]

    def forward(self, input, h=None, reverse=False):
        if self._program is None:
            self._program = Program(self.ops(), self.in_channels, [self])
        return self._program.run(input, reverse, h)


class MultiScalePrior(ProgramOwner, nn.Module):
    """macow2.py:543-594."""

    def __init__(self, in_channels, hidden_channels, h_channels, factor, transform, alpha, coupling_type, h_type,
                 activation, normalize, num_groups, use_1x1=False, condition_nice=False, attention=False, heads=4,
                 spatial_size=8, cond_conv=False, cond_conv_hidden_channels=None, p_dropout=0.):
        super().__init__()
        _unsupported(use_1x1=(bool(use_1x1), (False,)), condition_nice=(bool(condition_nice), (False,)))
        self.in_channels = in_channels
        self.conv1x1 = Shuffle(in_channels)
        self.coupling = NICE2d(in_channels, hidden_channels=hidden_channels, h_channels=0, transform=transform,
                               alpha=alpha, factor=factor, type=coupling_type, h_type=h_type, split_type="continuous",
                               order="up", activation=activation, normalize=normalize, num_groups=num_groups,
                               attention=attention, heads=heads, spatial_size=spatial_size, cond_conv=cond_conv,
                               cond_conv_hidden_channels=cond_conv_hidden_channels, p_dropout=p_dropout)
        out_channels = in_channels // factor
        self.z1_channels = self.coupling.z1_channels
        assert out_channels + self.z1_channels == in_channels
        self.actnorm = ActNorm2dFlow(out_channels)
        self._program = None
        self.register_load_state_dict_post_hook(_invalidate_after_load)

    def ops(self, level_shuffle=None):
        """Forward calls (macow2.py:571-582); ``level_shuffle`` folds the permutation that follows the prior in
        MultiScaleInternal (macow2.py:889) into the launch that applies the prior's ActNorm."""
        C = self.in_channels
        return [("glue", self.conv1x1, C, None, 0, 0), ("nice", self.coupling),
                ("glue", level_shuffle, C if level_shuffle is not None else 0, self.actnorm, self.z1_channels,
                 C - self.z1_channels)]

    def forward(self, input, h=None, reverse=False):
        # with condition_nice false the prior's coupling is built without conditioning channels and ignores h
        # (macow2.py:556, macow_utils.py:330)
        if self._program is None:
            self._program = Program(self.ops(), self.in_channels, [self])
        return self._program.run(input, reverse)


class MultiScaleInternal(ProgramOwner, nn.Module):
    """macow2.py:821-918."""

    def __init__(self, flow_step, num_steps, in_channels, hidden_channels, h_channels, factor=2, transform="affine",
                 prior_transform="affine", alpha=1.0, kernel_size=(2, 3), coupling_type="conv", h_type=None,
                 activation="relu", normalize=None, num_groups=None, use_1x1=False, condition_nice=False,
                 attention=False, heads=4, spatial_size=8, cond_conv=False, cond_conv_hidden_channels=None,
                 p_dropout=0.):
        super().__init__()
        _unsupported(use_1x1=(bool(use_1x1), (False,)), condition_nice=(bool(condition_nice), (False,)),
                     cond_conv=(bool(cond_conv) and bool(h_channels), (False,)))
        self.h_channels = int(h_channels or 0)
        self.reshape = "none"
        num_layers = len(num_steps)
        assert num_layers < factor
        self.in_channels = in_channels
        self.layers = nn.ModuleList()
        self.priors = nn.ModuleList()
        self.shuffle_layers = nn.ModuleList()
        channel_step = in_channels // factor
        for num_step in num_steps:
            layer = [flow_step(in_channels, hidden_channels=hidden_channels, h_channels=h_channels, transform=transform,
                               alpha=alpha, coupling_type=coupling_type, h_type=h_type, activation=activation,
                               normalize=normalize, num_groups=num_groups, kernel_size=kernel_size,
                               condition_nice=condition_nice, attention=attention, heads=heads,
                               spatial_size=spatial_size, cond_conv=cond_conv,
                               cond_conv_hidden_channels=cond_conv_hidden_channels, p_dropout=p_dropout)
                     for _ in range(num_step)]
            self.layers.append(nn.ModuleList(layer))
            prior = MultiScalePrior(in_channels, hidden_channels=hidden_channels, h_channels=h_channels,
                                    transform=prior_transform, alpha=alpha, factor=factor, coupling_type=coupling_type,
                                    h_type=h_type, activation=activation, normalize=normalize, num_groups=num_groups,
                                    condition_nice=condition_nice, attention=attention, heads=heads,
                                    spatial_size=spatial_size, cond_conv=cond_conv,
                                    cond_conv_hidden_channels=cond_conv_hidden_channels, p_dropout=p_dropout)
            self.priors.append(prior)
            self.shuffle_layers.append(Shuffle(in_channels))
            in_channels = in_channels - channel_step
            assert in_channels == prior.z1_channels
            factor = factor - 1
        self.z_channels = in_channels
        self._program = None
        self.register_load_state_dict_post_hook(_invalidate_after_load)

    def ops(self):
        ops = []
        for layer, prior, shuffle in zip(self.layers, self.priors, self.shuffle_layers):
            for step in layer:
                ops += step.ops()
            ops += prior.ops(level_shuffle=shuffle)        # the split that follows keeps the first z1_channels: no data moves
        return ops

    def forward(self, input, h=None, reverse=False):
        if self._program is None:
            self._program = Program(self.ops(), self.in_channels, [self])
        return self._program.run(input, reverse, h)


class Program:
    """A flattened list of forward kernel calls over one in-place [B, C, H, W] buffer, its packed parameters (one segment
    per call, repacked when a tensor of the segment changed) and the reference's first-call initialisations.  A program
    owns its workspaces: calls on one module are for one host thread and one stream at a time (like the reference's
    modules, whose first call mutates parameters)."""

    def __init__(self, ops, channels, roots):
        self.ops = ops
        self.channels = channels
        self.roots = roots
        self.packed = [None] * len(ops)           # per op: (tensor, key)
        self.params = [None] * len(ops)           # per op: its parameter tensors (walked once)
        self.init_done = False
        self.ws = None
        self.ws_need = {}                         # (B, H, W) -> workspace floats of the widest coupling
        self.hb = None
        self.h_channels = max([op[1][0].h_channels for op in ops if op[0] == "units"] + [0])

    def invalidate(self):
        self.packed = [None] * len(self.ops)
        self.params = [None] * len(self.ops)
        self.init_done = False

    # ---- packed parameters
    def _segment(self, i, dev):
        op = self.ops[i]
        units = op[1] if op[0] == "units" else None
        ps = self.params[i]
        if ps is None:                            # load_state_dict keeps the Parameter objects; invalidate() drops this
            ps = [p for u in units for p in u._param_list()] if units is not None else op[1].params()
            self.params[i] = ps
        key = tuple((p.data_ptr(), p._version) for p in ps)
        cur = self.packed[i]
        if cur is not None and cur[1] == key and cur[0].device == dev:
            return cur[0] if op[0] != "units" else (cur[0], cur[2])
        L = _m._lib()
        if any(p.dtype != torch.float32 for p in ps):
            raise TypeError("MaCow parameters must be float32")
        tensors = [p.detach().contiguous() for p in ps]
        if op[0] == "units":
            d = units[0]._desc
            n = L.ipoke_macow_unit_floats(ctypes.byref(d))
            if n == 0:
                raise RuntimeError("ipoke_macow: unsupported unit shape (hidden % 4, kernel_size[1] in {1, 3, 5})")
            buf = torch.empty(n * len(units), dtype=torch.float32, device=dev)
            wh = None
            if d.h_channels:                       # conditional chain: the conditioning matrix of all its flows
                wh = torch.empty(L.ipoke_macow_cond_floats(ctypes.byref(d), len(units)), dtype=torch.float32, device=dev)
            for u in range(len(units)):
                ptrs = (ctypes.c_void_p * 20)(*[t.data_ptr() for t in tensors[20 * u:20 * u + 20]])
                if wh is None:
                    _capi.check(L.ipoke_macow_pack_unit(ctypes.byref(d), ptrs, buf.data_ptr() + 4 * n * u,
                                                        _m._stream(dev)), "ipoke_macow_pack_unit")
                else:
                    _capi.check(L.ipoke_macow_pack_unit_cond(ctypes.byref(d), ptrs, buf.data_ptr() + 4 * n * u,
                                                             wh.data_ptr(), len(units), u, _m._stream(dev)),
                                "ipoke_macow_pack_unit_cond")
            self.packed[i] = (buf, key, wh)
            return buf, wh
        else:
            d = op[1]._desc
            n = L.ipoke_macow_nice_floats(ctypes.byref(d))
            buf = torch.empty(n, dtype=torch.float32, device=dev)
            ptrs = (ctypes.c_void_p * 5)(*[t.data_ptr() for t in tensors])
            _capi.check(L.ipoke_macow_pack_nice(ctypes.byref(d), ptrs, buf.data_ptr(), _m._stream(dev)),
                        "ipoke_macow_pack_nice")
        self.packed[i] = (buf, key)
        return buf

    def _workspace(self, B, H, W, dev):
        need = self.ws_need.get((B, H, W))
        if need is None:
            L = _m._lib()
            need = max([L.ipoke_macow_nice_workspace_floats(ctypes.byref(op[1]._desc), B, H, W)
                        for op in self.ops if op[0] == "nice"] + [0])
            self.ws_need[(B, H, W)] = need
        if self.ws is None or self.ws.numel() < need or self.ws.device != dev:
            self.ws = torch.empty(max(need, 1), dtype=torch.float32, device=dev)
        return self.ws

    # ---- kernel calls
    @annotate("macow.glue")
    def _glue(self, op, buf, stride, B, H, W, logdet, reverse):
        _, shuffle, n_perm, actnorm, c0, n = op
        perm = None
        if shuffle is not None:
            perm = shuffle.backward_shuffle_idx if reverse else shuffle.forward_shuffle_idx
            if perm.dtype != torch.int64:
                raise TypeError("Shuffle indices must be int64")
        ls = actnorm.log_scale.detach() if actnorm is not None else None
        bias = actnorm.bias.detach() if actnorm is not None else None
        _capi.check(_m._lib().ipoke_macow_chanmap(
            buf.data_ptr(), stride, B, H, W, perm.data_ptr() if perm is not None else None, n_perm,
            ls.data_ptr() if ls is not None else None, bias.data_ptr() if bias is not None else None, c0, n,
            int(reverse), int(not reverse), logdet.data_ptr() if (logdet is not None and not reverse) else None,
            _m._stream(buf.device)), "ipoke_macow_chanmap")

    @annotate("macow.units")
    def _units(self, i, buf, stride, B, H, W, logdet, reverse, h=None, s0=None, s1=None):
        units = self.ops[i][1]
        packed, wh = self._segment(i, buf.device)
        d = units[0]._desc
        s0 = 0 if s0 is None else s0
        s1 = 6 * len(units) if s1 is None else s1
        L = _m._lib()
        if wh is not None:
            # the conditioning's contribution to every (shift, log-scale) of the chain: hb = act(h) x Wh, one launch
            need = B * H * W * (wh.numel() // d.h_channels)
            if self.hb is None or self.hb.numel() < need or self.hb.device != buf.device:
                self.hb = torch.empty(need, dtype=torch.float32, device=buf.device)
            st = _m._stream(buf.device)
            _capi.check(L.ipoke_macow_cond_bias(ctypes.byref(d), wh.data_ptr(), len(units), h.data_ptr(),
                                                d.h_channels * H * W, self.hb.data_ptr(), B, H, W, st),
                        "ipoke_macow_cond_bias")
            if reverse:
                _capi.check(L.ipoke_macow_units_reverse_cond(ctypes.byref(d), packed.data_ptr(), len(units),
                                                             self.hb.data_ptr(), buf.data_ptr(), stride, buf.data_ptr(),
                                                             stride, B, H, W, s0, s1, st), "ipoke_macow_units_reverse_cond")
            else:
                _capi.check(L.ipoke_macow_units_forward_cond(ctypes.byref(d), packed.data_ptr(), len(units),
                                                             self.hb.data_ptr(), buf.data_ptr(), stride, buf.data_ptr(),
                                                             stride, logdet.data_ptr(), 1, B, H, W, s0, s1, st),
                            "ipoke_macow_units_forward_cond")
            return
        if reverse:
            _capi.check(L.ipoke_macow_units_reverse(ctypes.byref(d), packed.data_ptr(), len(units), buf.data_ptr(), stride,
                                                    buf.data_ptr(), stride, B, H, W, s0, s1, _m._stream(buf.device)),
                        "ipoke_macow_units_reverse")
        else:
            _capi.check(L.ipoke_macow_units_forward(ctypes.byref(d), packed.data_ptr(), len(units), buf.data_ptr(), stride,
                                                    buf.data_ptr(), stride, logdet.data_ptr(), 1, B, H, W, s0, s1,
                                                    _m._stream(buf.device)), "ipoke_macow_units_forward")

    @annotate("macow.nice")
    def _nice(self, i, buf, stride, B, H, W, logdet, reverse):
        nice = self.ops[i][1]
        packed = self._segment(i, buf.device)
        ws = self._workspace(B, H, W, buf.device)
        _capi.check(_m._lib().ipoke_macow_nice(ctypes.byref(nice._desc), packed.data_ptr(), buf.data_ptr(), stride,
                                               logdet.data_ptr() if not reverse else None, int(reverse), B, H, W,
                                               ws.data_ptr(), ws.numel(), _m._stream(buf.device)), "ipoke_macow_nice")

    # ---- first call: the reference's data-dependent initialisations, in its order
    def _needs_init(self):
        if self.init_done:
            return False
        flags = []
        for op in self.ops:
            if op[0] == "glue" and op[3] is not None:
                flags.append(op[3].initialized)
            elif op[0] == "nice":
                flags.append(op[1].net.conv3.initialized)
            elif op[0] == "units":
                for u in op[1]:
                    flags += [getattr(u, n).net.conv1x1.initialized for n in ("conv1", "conv2", "conv3", "conv4")]
                    flags += [u.actnorm1.initialized, u.actnorm2.initialized]
        if flags and int(torch.stack([f.reshape(()) for f in flags]).min().item()) == 0:
            return True
        self.init_done = True
        return False

    def _forward_with_init(self, buf, stride, B, H, W, logdet, h=None):
        """ActNorm2dFlow.init on the data that reaches it (macow2.py:503-505, 526-541); the zero_init weight-normed
        convolutions get gain 0 and bias 0 (macow_utils.py:235-251 with init_scale = 0)."""
        for i, op in enumerate(self.ops):
            if op[0] == "glue":
                an = op[3]
                if an is not None and int(an.initialized.item()) == 0:
                    an.init(buf[:, op[4]:op[4] + op[5]])
                    an.initialized.fill_(1)
                self._glue(op, buf, stride, B, H, W, logdet, False)
            elif op[0] == "nice":
                conv3 = op[1].net.conv3
                if int(conv3.initialized.item()) == 0:
                    conv3.data_init()
                self._nice(i, buf, stride, B, H, W, logdet, False)
            else:
                C = op[1][0].in_channels
                for u, unit in enumerate(op[1]):
                    for s, name in enumerate(_m.STAGES):
                        mod = getattr(unit, name)
                        if name.startswith("conv"):
                            if int(mod.net.conv1x1.initialized.item()) == 0:
                                mod.net.conv1x1.data_init()
                        elif int(mod.initialized.item()) == 0:
                            mod.init(buf[:, :C])
                            mod.initialized.fill_(1)
                        self._units(i, buf, stride, B, H, W, logdet, False, h, 6 * u + s, 6 * u + s + 1)

    def _check_h(self, h, x):
        if not self.h_channels:
            return None                                          # no conditional unit in this program: h is ignored
        if h is None:
            raise ValueError(f"this flow is conditional: pass the conditioning h [batch, {self.h_channels}, H, W]")
        want = (x.shape[0], self.h_channels, x.shape[2], x.shape[3])
        if tuple(h.shape) != want or h.dtype != torch.float32 or h.device != x.device:
            raise ValueError(f"conditioning: expected float32 {want} on {x.device}, got {h.dtype} {tuple(h.shape)} on {h.device}")
        return h.contiguous()

    @annotate("macow.flow")
    def run(self, input, reverse=False, h=None):
        if torch.is_grad_enabled() and (input.requires_grad or any(p.requires_grad for r in self.roots
                                                                   for p in r.parameters())):
            raise NotImplementedError("the MaCow path is inference only: call it under torch.no_grad()")
        x = _m._check_input(input, self.channels)
        B, C, H, W = x.shape
        where = next(self.roots[0].parameters()).device
        if where != x.device:                                  # the kernels would dereference the other device's pointers
            raise RuntimeError(f"module parameters are on {where}, the input is on {x.device}")
        h = self._check_h(h, x)
        buf = x.clone() if x.data_ptr() == input.data_ptr() else x
        stride = C * H * W
        if B == 0:
            return buf if reverse else (buf, torch.zeros(0, dtype=torch.float32, device=buf.device))
        if reverse:
            for i in range(len(self.ops) - 1, -1, -1):
                op = self.ops[i]
                if op[0] == "glue":
                    self._glue(op, buf, stride, B, H, W, None, True)
                elif op[0] == "units":
                    self._units(i, buf, stride, B, H, W, None, True, h)
                else:
                    self._nice(i, buf, stride, B, H, W, None, True)
            return buf
        logdet = torch.zeros(B, dtype=torch.float32, device=buf.device)
        if self._needs_init():
            self._forward_with_init(buf, stride, B, H, W, logdet, h)
            return buf, logdet
        for i, op in enumerate(self.ops):
            if op[0] == "glue":
                self._glue(op, buf, stride, B, H, W, logdet, False)
            elif op[0] == "units":
                self._units(i, buf, stride, B, H, W, logdet, False, h)
            else:
                self._nice(i, buf, stride, B, H, W, logdet, False)
        return buf, logdet

    @property
    def launches(self):
        """Kernel launches of one call (either direction) at a batch that needs one pass per coupling: a coupling is
        three GEMMs and the tap / affine kernel."""
        return sum({"glue": 1, "units": 2 if self.h_channels else 1, "nice": 4}[op[0]] for op in self.ops)


class UnsupervisedMaCowTransformer3(nn.Module):
    """INN.py:365-390."""

    def __init__(self, config):
        super().__init__()
        self.config = config
        self.flow = MultiScaleInternal(MaCowStep, num_steps=self.config["num_steps"],
                                       in_channels=self.config["flow_in_channels"],
                                       hidden_channels=self.config["flow_mid_channels"], h_channels=0,
                                       factor=self.config["factor"], transform=self.config["transform"],
                                       prior_transform=self.config["prior_transform"],
                                       kernel_size=self.config["kernel_size"],
                                       coupling_type=self.config["coupling_type"], activation=self.config["activation"],
                                       use_1x1=self.config["use1x1"] if "use1x1" in self.config else False)

    def forward(self, input, reverse=False):
        if reverse:
            return self.reverse(input)
        out, logdet = self.flow(input)
        return out, logdet

    def reverse(self, out):
        return self.flow(out, reverse=True)

    def sample(self, shape, device="cpu"):
        z_tilde = torch.randn(shape).to(device)
        sample = self.reverse(z_tilde)
        return sample


class SupervisedMacowTransformer(nn.Module):
    """INN.py:447-482: the conditional flow of the conv second stage (models/second_stage_video.py:122).  Reads the same
    config keys as the reference (``flow_attn_heads`` and ``cond_conv_hidden_channels`` are required there too)."""

    def __init__(self, config):
        super().__init__()
        self.config = config
        condition_nice = "condition_nice" in self.config and self.config["condition_nice"]
        attention = "attention" in self.config and self.config["attention"]
        heads = self.config["flow_attn_heads"]
        ssize = "ssize" in self.config and self.config["ssize"]
        cond_conv = "cond_conv" in config and config["cond_conv"]
        cond_conv_hidden_channels = config["cond_conv_hidden_channels"]
        p_drop = config["p_dropout"] if "p_dropout" in config else 0.
        self.flow = MultiScaleInternal(MaCowStep, num_steps=self.config["num_steps"],
                                       in_channels=self.config["flow_in_channels"],
                                       hidden_channels=self.config["flow_mid_channels"],
                                       h_channels=self.config["h_channels"], factor=self.config["factor"],
                                       transform=self.config["transform"], prior_transform=self.config["prior_transform"],
                                       kernel_size=self.config["kernel_size"], coupling_type=self.config["coupling_type"],
                                       activation=self.config["activation"],
                                       use_1x1=self.config["use1x1"] if "use1x1" in self.config else False,
                                       condition_nice=condition_nice, attention=attention, heads=heads, spatial_size=ssize,
                                       cond_conv=cond_conv, cond_conv_hidden_channels=cond_conv_hidden_channels,
                                       p_dropout=p_drop)

    def forward(self, input, cond, reverse=False):
        if reverse:
            return self.reverse(input, cond)
        out, logdet = self.flow(input, cond)
        return out, logdet

    def reverse(self, out, cond):
        return self.flow(out, cond, reverse=True)

    def sample(self, shape, cond, device="cpu"):
        z_tilde = torch.randn(shape).to(device)
        sample = self.reverse(z_tilde, cond)
        return sample
