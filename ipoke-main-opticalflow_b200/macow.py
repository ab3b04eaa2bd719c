"""MaCow unit on B200: the reference's ``MaCowUnit`` / ``MaskedConvFlow`` / ``ActNorm2dFlow`` modules
(models/modules/INN/macow2.py:25-289, 476-541, 925-996; macow_utils.py:42-66, 211-252, 407-497) with the same
constructors, state_dict keys and RNG draw order, computing through the CUDA kernels behind
``include/ipoke_macow.h``.  Row N3 of SURVEY.md section 8(f); the couplings, steps and the multi-scale flow built from
these units are in ``macow_flow.py``.  Affine transform, dropout 0, unconditional or conditional (``h_channels > 0``)
-- what ``config/second_stage.yaml:53-80`` selects.  Inference only (forward with logdet, and the sequential inverse
used for sampling); anything else raises NotImplementedError.  There is no CPU or eager fallback: CPU tensors raise.
"""
import ctypes
import math

import torch
import torch.nn as nn

from . import _capi

ACTIVATIONS = {"relu": 0, "elu": 1, "leaky_relu": 2}          # macow_utils.py:417-424
STAGES = ("conv1", "conv2", "actnorm1", "conv3", "conv4", "actnorm2")


class MacowDesc(ctypes.Structure):
    _fields_ = [("C", ctypes.c_int32), ("hidden", ctypes.c_int32), ("k0", ctypes.c_int32), ("k1", ctypes.c_int32),
                ("activation", ctypes.c_int32), ("alpha", ctypes.c_float), ("h_channels", ctypes.c_int32)]


class NiceDesc(ctypes.Structure):
    _fields_ = [("c_in", ctypes.c_int32), ("z_off", ctypes.c_int32), ("z_stride", ctypes.c_int32),
                ("c_t", ctypes.c_int32), ("p_off", ctypes.c_int32), ("p_stride", ctypes.c_int32),
                ("hidden", ctypes.c_int32), ("activation", ctypes.c_int32), ("alpha", ctypes.c_float)]


def _bind(L):
    """Argument types of the entry points of include/ipoke_macow.h on a loaded library."""
    vp, i32, i64, sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_size_t
    dp, ndp = ctypes.POINTER(MacowDesc), ctypes.POINTER(NiceDesc)
    L.ipoke_macow_unit_floats.restype = sz
    L.ipoke_macow_unit_floats.argtypes = [dp]
    L.ipoke_macow_pack_unit.restype = i32
    L.ipoke_macow_pack_unit.argtypes = [dp, ctypes.POINTER(vp), vp, vp]
    L.ipoke_macow_forward.restype = i32
    L.ipoke_macow_forward.argtypes = [dp, vp, i32, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp]
    L.ipoke_macow_reverse.restype = i32
    L.ipoke_macow_reverse.argtypes = [dp, vp, i32, vp, vp, i32, i32, i32, i32, i32, vp]
    L.ipoke_macow_units_forward.restype = i32
    L.ipoke_macow_units_forward.argtypes = [dp, vp, i32, vp, i64, vp, i64, vp, i32, i32, i32, i32, i32, i32, vp]
    L.ipoke_macow_units_reverse.restype = i32
    L.ipoke_macow_units_reverse.argtypes = [dp, vp, i32, vp, i64, vp, i64, i32, i32, i32, i32, i32, vp]
    L.ipoke_macow_cond_floats.restype = sz
    L.ipoke_macow_cond_floats.argtypes = [dp, i32]
    L.ipoke_macow_pack_unit_cond.restype = i32
    L.ipoke_macow_pack_unit_cond.argtypes = [dp, ctypes.POINTER(vp), vp, vp, i32, i32, vp]
    L.ipoke_macow_cond_bias.restype = i32
    L.ipoke_macow_cond_bias.argtypes = [dp, vp, i32, vp, i64, vp, i32, i32, i32, vp]
    L.ipoke_macow_units_forward_cond.restype = i32
    L.ipoke_macow_units_forward_cond.argtypes = [dp, vp, i32, vp, vp, i64, vp, i64, vp, i32, i32, i32, i32, i32, i32, vp]
    L.ipoke_macow_units_reverse_cond.restype = i32
    L.ipoke_macow_units_reverse_cond.argtypes = [dp, vp, i32, vp, vp, i64, vp, i64, i32, i32, i32, i32, i32, vp]
    L.ipoke_macow_launch_info.restype = i32
    L.ipoke_macow_launch_info.argtypes = [dp, i32, i32, i32, ctypes.POINTER(i32), ctypes.POINTER(i32),
                                          ctypes.POINTER(i32)]
    L.ipoke_macow_nice_floats.restype = sz
    L.ipoke_macow_nice_floats.argtypes = [ndp]
    L.ipoke_macow_pack_nice.restype = i32
    L.ipoke_macow_pack_nice.argtypes = [ndp, ctypes.POINTER(vp), vp, vp]
    L.ipoke_macow_nice_workspace_floats.restype = sz
    L.ipoke_macow_nice_workspace_floats.argtypes = [ndp, i32, i32, i32]
    L.ipoke_macow_nice.restype = i32
    L.ipoke_macow_nice.argtypes = [ndp, vp, vp, i64, vp, i32, i32, i32, i32, vp, sz, vp]
    L.ipoke_macow_chanmap.restype = i32
    L.ipoke_macow_chanmap.argtypes = [vp, i64, i32, i32, i32, vp, i32, vp, vp, i32, i32, i32, i32, vp, vp]
    return L


_bound = False


def _lib():
    global _bound
    L = _capi.lib()
    if not _bound:
        _bind(L)
        _bound = True
    return L


def _stream(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _check_input(x, channels):
    if not x.is_cuda:
        raise RuntimeError("the MaCow path runs on CUDA (sm_100a) only: there is no CPU fallback")
    if x.dim() != 4 or x.shape[1] != channels:
        raise ValueError(f"expected [batch, {channels}, H, W], got {tuple(x.shape)}")
    if x.dtype != torch.float32:
        raise TypeError("the MaCow path computes in float32 like the reference")
    return x.contiguous()


def _invalidate_after_load(module, incompatible):
    """load_state_dict post hook (a module-level function, so that modules stay picklable)."""
    module.invalidate()


class ProgramOwner:
    """Mixin of the modules that run through a lazily built ``macow_flow.Program``: the program (packed parameters,
    workspaces) is not part of a copy or a pickle -- ``copy.deepcopy`` / ``torch.save(module)`` carry the parameters and
    buffers, the copy rebuilds its program at its first call."""

    def __getstate__(self):
        state = self.__dict__.copy()
        state["_program"] = None
        return state

    def invalidate(self):
        """Force a repack and a re-read of the ``initialized`` flags (needed after writing parameters through ``.data``,
        which the version counters used to detect changes do not see)."""
        if getattr(self, "_program", None) is not None:
            self._program.invalidate()


class ShiftedConv2d(nn.Module):
    """Parameter holder of macow_utils.py:439-497 (``weight`` [hidden, C, kH, kW], no bias in MCFBlock)."""

    def __init__(self, in_channels, out_channels, kernel_size, order="A"):
        super().__init__()
        assert order in {"A", "B", "C", "D"}, "unknown order: {}".format(order)
        if order in {"A", "B"}:
            assert kernel_size[1] % 2 == 1, "kernel width cannot be even number: {}".format(kernel_size)
        else:
            assert kernel_size[0] % 2 == 1, "kernel height cannot be even number: {}".format(kernel_size)
        self.order = order
        self.weight = nn.Parameter(torch.empty(out_channels, in_channels, kernel_size[0], kernel_size[1]))
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))          # nn.Conv2d.reset_parameters


class _WeightNormedConv(nn.Module):
    """``nn.utils.weight_norm(nn.Conv2d(hidden, out, 1))`` as plain parameters bias / weight_g / weight_v in the
    reference's registration order, drawing the same random numbers (macow_utils.py:216-233)."""

    def __init__(self, in_channels, out_channels, kernel_size=1):
        super().__init__()
        w = torch.empty(out_channels, in_channels, kernel_size, kernel_size)
        nn.init.kaiming_uniform_(w, a=math.sqrt(5))                    # nn.Conv2d.reset_parameters: weight ...
        bound = 1 / math.sqrt(in_channels * kernel_size * kernel_size)
        self.bias = nn.Parameter(torch.empty(out_channels))
        nn.init.uniform_(self.bias, -bound, bound)                     # ... and bias
        nn.init.normal_(w, mean=0.0, std=0.05)                         # macow_utils.py:227
        nn.init.constant_(self.bias, 0)                                # macow_utils.py:229
        self.weight_g = nn.Parameter(w.reshape(out_channels, -1).norm(dim=1).reshape(out_channels, 1, 1, 1))
        self.weight_v = nn.Parameter(w)


class Conv2dWeightNorm(nn.Module):
    """macow_utils.py:211-251 for the zero_init cases: the 1x1 convolution of MCFBlock and the 3x3 one (padding 1) that
    ends NICEConvBlock (macow_utils.py:283)."""

    def __init__(self, in_channels, out_channels, zero_init=True, kernel_size=1):
        super().__init__()
        self.conv = _WeightNormedConv(in_channels, out_channels, kernel_size)
        self.register_buffer("initialized", torch.tensor(0, dtype=torch.uint8))
        self.zero_init = zero_init

    def data_init(self):
        """Conv2dWeightNorm.init with init_scale = 0 (macow_utils.py:235-251, zero_init): the gain becomes
        0 / (std + 1e-6) = 0 and the bias -mean * 0 = 0 whatever the (finite) activations are."""
        if not self.zero_init:
            raise NotImplementedError("only the zero_init convolutions of MCFBlock / NICEConvBlock are built")
        with torch.no_grad():
            self.conv.weight_g.zero_()
            self.conv.bias.zero_()
            self.initialized.fill_(1)


class MCFBlock(nn.Module):
    """macow_utils.py:407-436 (parameter holder; the arithmetic is in the kernel)."""

    def __init__(self, in_channels, out_channels, kernel_size, hidden_channels, order, activation, h_channels=None,
                 p_dropout=0.):
        super().__init__()
        if p_dropout != 0.:
            raise NotImplementedError("dropout inside the MCF block is not built")
        assert activation in ACTIVATIONS
        self.shift_conv = ShiftedConv2d(in_channels, hidden_channels, kernel_size, order=order)
        # the conditioning is concatenated to the hidden activations in front of the 1x1 convolution (macow_utils.py:409-412)
        self.conv1x1 = Conv2dWeightNorm(hidden_channels + (h_channels or 0), out_channels, zero_init=True)


class ActNorm2dFlow(nn.Module):
    """macow2.py:476-541."""

    def __init__(self, in_channels):
        super().__init__()
        self.in_channels = in_channels
        self.log_scale = nn.Parameter(torch.Tensor(in_channels, 1, 1))
        self.bias = nn.Parameter(torch.Tensor(in_channels, 1, 1))
        nn.init.normal_(self.log_scale, mean=0, std=0.05)
        nn.init.constant_(self.bias, 0.)
        self.register_buffer("initialized", torch.tensor(0, dtype=torch.uint8))

    def init(self, data, init_scale=1.0):
        """Data-dependent initialisation (macow2.py:526-541), host-side tensor statistics, first call only."""
        with torch.no_grad():
            out = data * self.log_scale.exp() + self.bias
            out = out.transpose(0, 1).contiguous().view(self.in_channels, -1)
            mean = out.mean(dim=1).view(self.in_channels, 1, 1)
            std = out.std(dim=1).view(self.in_channels, 1, 1)
            inv_stdv = init_scale / (std + 1e-6)
            self.log_scale.copy_(inv_stdv.log())          # in place on the Parameter: bumps its version -> repack
            self.bias.copy_(-mean * inv_stdv)


class MaskedConvFlow(nn.Module):
    """macow2.py:25-289 (parameter holder with the reference's attributes; run through its MaCowUnit)."""

    def __init__(self, in_channels, kernel_size, hidden_channels=None, h_channels=None, h_type=None,
                 activation="relu", order="A", transform="affine", alpha=1.0, p_dropout=0.):
        super().__init__()
        self.in_channels = in_channels
        self.cond = h_channels is not None and h_channels > 0
        if transform != "affine":
            raise NotImplementedError(f"transform {transform!r}: only 'affine' is built")
        if hidden_channels is None:
            hidden_channels = 4 * in_channels if in_channels <= 96 else min(2 * in_channels, 512)
        self.hidden_channels = hidden_channels
        self.alpha = alpha
        self.kernel_size = kernel_size
        self.order = order
        self.analytic_bwd = True
        self.net = MCFBlock(in_channels, in_channels * 2, kernel_size, hidden_channels, order, activation,
                            h_channels=h_channels if self.cond else None, p_dropout=p_dropout)


class MaCowUnit(ProgramOwner, nn.Module):
    """Drop-in for macow2.py:925-996: ``forward(input, h=None, reverse=False)`` returns ``(out, logdet)`` or,
    with ``reverse=True``, the inverse image.  One kernel launch per call (one CTA per sample); a conditional unit
    (``h_channels > 0``, conditioning ``h`` [B, h_channels, H, W]) adds the launch of its conditioning product."""

    def __init__(self, in_channels, kernel_size, h_channels=0, transform="affine", alpha=1.0, h_type=None,
                 activation="relu", cond_conv=False, cond_conv_hidden_channels=None, p_dropout=0.):
        super().__init__()
        if cond_conv and h_channels:
            raise NotImplementedError("the gated convolution on the conditioning (cond_conv) is not built")
        self.cond_conv = False
        self.h_channels = int(h_channels or 0)
        k0, k1 = kernel_size[0], kernel_size[1]
        kw = dict(h_channels=h_channels, transform=transform, alpha=alpha, h_type=h_type, activation=activation,
                  p_dropout=p_dropout)
        self.conv1 = MaskedConvFlow(in_channels, (k0, k1), order="A", **kw)
        self.conv2 = MaskedConvFlow(in_channels, (k0, k1), order="B", **kw)
        self.actnorm1 = ActNorm2dFlow(in_channels)
        self.conv3 = MaskedConvFlow(in_channels, (k1, k0), order="C", **kw)
        self.conv4 = MaskedConvFlow(in_channels, (k1, k0), order="D", **kw)
        self.actnorm2 = ActNorm2dFlow(in_channels)
        self.in_channels = in_channels
        self._desc = MacowDesc(in_channels, self.conv1.hidden_channels, k0, k1, ACTIVATIONS[activation], alpha,
                               self.h_channels)
        self._program = None
        self._packed = None
        self._packed_key = None
        self._init_checked = False
        self.register_load_state_dict_post_hook(_invalidate_after_load)

    def invalidate(self):
        """Force a repack and a re-read of the ``initialized`` flags on the next call.  Needed only after writing
        parameters through ``.data`` (which the version counters used to detect changes do not see)."""
        self._packed_key = None
        self._init_checked = False
        ProgramOwner.invalidate(self)

    # ---- parameters -> packed kernel layout
    def _param_list(self):
        ps = []
        for name in ("conv1", "conv2", "conv3", "conv4"):
            net = getattr(self, name).net
            ps += [net.shift_conv.weight, net.conv1x1.conv.bias, net.conv1x1.conv.weight_g, net.conv1x1.conv.weight_v]
        return ps + [self.actnorm1.log_scale, self.actnorm1.bias, self.actnorm2.log_scale, self.actnorm2.bias]

    def packed(self):
        """The unit's parameters in the kernel layout; repacked (one small kernel) when any tensor changed."""
        ps = self._param_list()
        dev = ps[0].device
        if dev.type != "cuda":
            raise RuntimeError("the MaCow path runs on CUDA (sm_100a) only: move the module to the GPU")
        key = tuple((p.data_ptr(), p._version) for p in ps)
        if self._packed is None or self._packed_key != key or self._packed.device != dev:
            L = _lib()
            n = L.ipoke_macow_unit_floats(ctypes.byref(self._desc))
            if n == 0:
                raise RuntimeError("ipoke_macow: unsupported unit shape (hidden % 4, kernel_size[1] in {1, 3, 5})")
            if self._packed is None or self._packed.device != dev:
                self._packed = torch.empty(n, dtype=torch.float32, device=dev)
            tensors = [p.detach().contiguous() for p in ps]
            if any(t.dtype != torch.float32 for t in tensors):
                raise TypeError("MaCow parameters must be float32")
            ptrs = (ctypes.c_void_p * 20)(*[t.data_ptr() for t in tensors])
            _capi.check(L.ipoke_macow_pack_unit(ctypes.byref(self._desc), ptrs, self._packed.data_ptr(), _stream(dev)),
                        "ipoke_macow_pack_unit")
            self._packed_key = key
        return self._packed

    # ---- kernel calls over a stage range
    def _forward_stages(self, x, s0, s1, logdet=None):
        B, _, H, W = x.shape
        out = torch.empty_like(x)
        acc = logdet is not None
        if logdet is None:
            logdet = torch.empty(B, dtype=torch.float32, device=x.device)
        _capi.check(_lib().ipoke_macow_forward(ctypes.byref(self._desc), self.packed().data_ptr(), 1, x.data_ptr(),
                                               out.data_ptr(), logdet.data_ptr(), int(acc), B, H, W, s0, s1,
                                               _stream(x.device)), "ipoke_macow_forward")
        return out, logdet

    def _reverse_stages(self, z, s0, s1):
        B, _, H, W = z.shape
        out = torch.empty_like(z)
        _capi.check(_lib().ipoke_macow_reverse(ctypes.byref(self._desc), self.packed().data_ptr(), 1, z.data_ptr(),
                                               out.data_ptr(), B, H, W, s0, s1, _stream(z.device)),
                    "ipoke_macow_reverse")
        return out

    def _needs_init(self):
        if self._init_checked:                        # flags only go 0 -> 1; one device read-back, then cached
            return False
        flags = [getattr(self, n).net.conv1x1.initialized for n in ("conv1", "conv2", "conv3", "conv4")]
        flags += [self.actnorm1.initialized, self.actnorm2.initialized]
        if any(int(f.item()) == 0 for f in flags):
            return True
        self._init_checked = True
        return False

    def _forward_with_init(self, x):
        """First forward call: the data-dependent initialisations of the reference, in its order
        (macow_utils.py:247-250 per masked flow, macow2.py:503-505 per ActNorm2dFlow), running the kernel stage
        by stage in between."""
        out, logdet = x, None
        for s, name in enumerate(STAGES):
            mod = getattr(self, name)
            if name.startswith("conv"):
                if int(mod.net.conv1x1.initialized.item()) == 0:
                    mod.net.conv1x1.data_init()
            elif int(mod.initialized.item()) == 0:
                mod.init(out)
                mod.initialized.fill_(1)
            out, logdet = self._forward_stages(out, s, s + 1, logdet)
        return out, logdet

    def forward(self, input, h=None, reverse=False):
        if self.h_channels:                       # conditional: through a one-call program (macow_flow.Program)
            if self._program is None:
                from .macow_flow import Program
                self._program = Program([("units", [self])], self.in_channels, [self])
            return self._program.run(input, reverse, h)
        if h is not None:
            raise ValueError("this unit was built without conditioning channels (h_channels = 0)")
        if torch.is_grad_enabled() and (input.requires_grad or any(p.requires_grad for p in self.parameters())):
            raise NotImplementedError("the MaCow path is inference only: call it under torch.no_grad()")
        x = _check_input(input, self.in_channels)
        if not reverse:
            if self._needs_init():
                return self._forward_with_init(x)
            return self._forward_stages(x, 0, len(STAGES))
        return self._reverse_stages(x, 0, len(STAGES))

    def launch_info(self, H, W, reverse=False):
        smem, threads, tile = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        _capi.check(_lib().ipoke_macow_launch_info(ctypes.byref(self._desc), H, W, int(reverse), ctypes.byref(smem),
                                                   ctypes.byref(threads), ctypes.byref(tile)))
        return {"smem_bytes": smem.value, "threads": threads.value, "tile_lines": tile.value}
