"""Flat flow stacks with the reference's module API, executed by the fused sm_100a kernels.

Mirrors models/modules/INN/flow_blocks.py of the reference: ConditionalFlow (:10-66), UnconditionalFlow2
(:432-465), their blocks (:243-277, :503-532) and couplings (:113-153, :200-240, :572-615) -- same constructor
arguments, attribute names, parameter/buffer names and construction (RNG) order.  ``forward`` of the two
stacks calls ``libipoke_inn.so`` through ``engine.FlowEngine``; sub-modules are parameter containers.
"""
import torch
import torch.nn as nn

from .engine import FlowEngine
from .modules import ActNorm, BasicFullyConnectedNet, IgnoreLeakyRelu, InvLeakyRelu, Shuffle, _fused_only

# default numeric mode of newly built flows: "fp32" = split-bf16 tensor-core passes (<= 1e-4 vs the fp32
# reference), "bf16" = single pass (looser stated bound).  Per-instance: flow.set_precision(...).
DEFAULT_PRECISION = "fp32"


class ConditionalDoubleVectorCouplingBlock(nn.Module):
    """'normal' coupling: s,t read cat(x_apply, cond).  Reference flow_blocks.py:113-153."""

    def __init__(self, in_channels, cond_channels, hidden_dim, depth=2):
        super().__init__()
        self.s = nn.ModuleList([
            BasicFullyConnectedNet(dim=in_channels // 2 + cond_channels, depth=depth, hidden_dim=hidden_dim,
                                   use_tanh=True, out_dim=in_channels // 2) for _ in range(2)])
        self.t = nn.ModuleList([
            BasicFullyConnectedNet(dim=in_channels // 2 + cond_channels, depth=depth, hidden_dim=hidden_dim,
                                   use_tanh=False, out_dim=in_channels // 2) for _ in range(2)])

    def forward(self, x, xc, reverse=False):
        _fused_only("ConditionalDoubleVectorCouplingBlock")


class ConditionalDoubleVectorCouplingBlock_cond(nn.Module):
    """'cond' coupling: s,t read the conditioning only.  Reference flow_blocks.py:200-240."""

    def __init__(self, in_channels, cond_channels, hidden_dim, depth=2):
        super().__init__()
        self.s = nn.ModuleList([
            BasicFullyConnectedNet(cond_channels, depth=depth, hidden_dim=hidden_dim, use_tanh=True,
                                   out_dim=in_channels // 2) for _ in range(2)])
        self.t = nn.ModuleList([
            BasicFullyConnectedNet(cond_channels, depth=depth, hidden_dim=hidden_dim, use_tanh=False,
                                   out_dim=in_channels // 2) for _ in range(2)])

    def forward(self, x, xc, reverse=False):
        _fused_only("ConditionalDoubleVectorCouplingBlock_cond")


class DoubleVectorCouplingBlock2(nn.Module):
    """Unconditional coupling.  Reference flow_blocks.py:572-615 (note the s,s,t,t construction order)."""

    def __init__(self, in_channels, hidden_dim, hidden_depth=2):
        super().__init__()
        dim1 = (in_channels // 2) + (in_channels % 2)
        dim2 = in_channels // 2
        self.s = nn.ModuleList([
            BasicFullyConnectedNet(dim=dim1, out_dim=dim2, depth=hidden_depth, hidden_dim=hidden_dim, use_tanh=True),
            BasicFullyConnectedNet(dim=dim1, out_dim=dim2, depth=hidden_depth, hidden_dim=hidden_dim, use_tanh=True),
        ])
        self.t = nn.ModuleList([
            BasicFullyConnectedNet(dim=dim1, out_dim=dim2, depth=hidden_depth, hidden_dim=hidden_dim, use_tanh=False),
            BasicFullyConnectedNet(dim=dim1, out_dim=dim2, depth=hidden_depth, hidden_dim=hidden_dim, use_tanh=False),
        ])

    def forward(self, x, reverse=False):
        _fused_only("DoubleVectorCouplingBlock2")


class ConditionalFlatDoubleCouplingFlowBlock(nn.Module):
    """ActNorm -> InvLeakyRelu -> coupling -> Shuffle.  Reference flow_blocks.py:243-277."""

    def __init__(self, in_channels, cond_channels, hidden_dim, hidden_depth, activation="lrelu", mode="normal"):
        super().__init__()
        acts = {"lrelu": InvLeakyRelu, "none": IgnoreLeakyRelu}
        self.norm_layer = ActNorm(in_channels, logdet=True)
        if mode == "cond":
            self.coupling = ConditionalDoubleVectorCouplingBlock_cond(in_channels, cond_channels, hidden_dim, hidden_depth)
        else:
            self.coupling = ConditionalDoubleVectorCouplingBlock(in_channels, cond_channels, hidden_dim, hidden_depth)
        self.activation = acts[activation]()
        self.shuffle = Shuffle(in_channels)

    def forward(self, x, xcond, reverse=False):
        _fused_only("ConditionalFlatDoubleCouplingFlowBlock")


class UnconditionalFlatDoubleCouplingFlowBlock2(nn.Module):
    """ActNorm -> coupling -> Shuffle.  Reference flow_blocks.py:503-532."""

    def __init__(self, in_channels, hidden_dim, hidden_depth, data_init=False):
        super().__init__()
        if data_init:
            raise NotImplementedError("data_init=True (DoubleVectorCouplingBlock3) is unused: "
                                      "UnsupervisedTransformer3 passes data_init=False (reference INN.py:268)")
        self.norm_layer = ActNorm(in_channels, logdet=True)
        self.coupling = DoubleVectorCouplingBlock2(in_channels, hidden_dim, hidden_depth)
        self.shuffle = Shuffle(in_channels)

    def forward(self, x, reverse=False):
        _fused_only("UnconditionalFlatDoubleCouplingFlowBlock2")


class _FusedFlowMixin:
    """Shared host logic of the two stacks (engine ownership, precision knob, state hooks)."""

    def _make_engine(self, conditional, E, alpha, precision):
        object.__setattr__(self, "_engine", FlowEngine(self, conditional, self.in_channels, E, self.mid_channels,
                                                        self.num_blocks, self.n_flows, alpha, precision))

    @property
    def engine(self):
        return self._engine

    @property
    def precision(self):
        return self._engine.precision

    def set_precision(self, precision):
        """'fp32' (split-bf16, parity mode) or 'bf16' (single pass)."""
        old = self._engine
        object.__setattr__(self, "_engine", FlowEngine(self, bool(old.desc.conditional), old.C, old.E, old.desc.H,
                                                        old.desc.depth, old.n_flows, old.desc.alpha, precision))
        old.release()
        return self

    # ``last_outs`` / ``last_logdets`` (reference flow_blocks.py:48-57) are only read by the wrappers' ``train=True``
    # return (INN.py:80-81, 284-285).  Under no_grad the fused forward therefore skips the per-block outputs (so
    # the block boundaries stay folded into the coupling tails) and the lists are materialised on first access
    # by one extra pass over the same input tensors (which are kept alive, not copied, until the next call).
    def _reset_last(self):
        self._last_outs, self._last_logdets, self._last_lazy = [], [], None

    def _fill_last(self, outs, lds):
        four_d = not bool(self._engine.desc.conditional)            # unconditional blocks return 4-D (reference :606)
        self._last_outs = [o.unsqueeze(-1).unsqueeze(-1) for o in outs.unbind(0)] if four_d else list(outs.unbind(0))
        self._last_logdets = list(lds.unbind(0))
        self._last_lazy = None

    def _materialise_last(self):
        if getattr(self, "_last_lazy", None) is not None:
            x, cond = self._last_lazy
            with torch.no_grad():
                _, _, outs, lds = self._engine.forward(x, cond, want_blocks=True)
            self._fill_last(outs, lds)

    @property
    def last_outs(self):
        self._materialise_last()
        return self._last_outs

    @last_outs.setter
    def last_outs(self, v):
        self._last_outs, self._last_lazy = v, None

    @property
    def last_logdets(self):
        self._materialise_last()
        return self._last_logdets

    @last_logdets.setter
    def last_logdets(self, v):
        self._last_logdets, self._last_lazy = v, None

    def _forward_fused(self, x, cond):
        from .autograd import flow_forward
        self._reset_last()
        z, logdet, outs, lds = flow_forward(self._engine, x, cond)
        if outs is not None:
            self._fill_last(outs, lds)
            if z.requires_grad:
                # the last block's output IS the flow output: hand out the differentiable tensors there, so that a loss
                # built from ``last_outs[-1]`` / ``last_logdets[-1]`` (the wrappers' train=True return, reference
                # INN.py:80-81) reaches the parameters; the earlier blocks' entries are detached (fused backward)
                self._last_outs[-1] = z.unsqueeze(-1).unsqueeze(-1) if not bool(self._engine.desc.conditional) else z
                self._last_logdets[-1] = logdet
        else:
            self._last_lazy = (x, cond)
        return z.unsqueeze(-1).unsqueeze(-1), logdet                # reference :57

    def _apply(self, fn, *args, **kwargs):      # .to() / .cuda() / .float(): parameters move, repack
        out = super()._apply(fn, *args, **kwargs)
        self._engine.invalidate()
        return out


class ConditionalFlow(_FusedFlowMixin, nn.Module):
    """Flat conditional flow.  Reference flow_blocks.py:10-66."""

    def __init__(self, in_channels, embedding_dim, hidden_dim, hidden_depth, n_flows, conditioning_option="none",
                 activation="lrelu", precision=None):
        super().__init__()
        self.in_channels = in_channels
        self.cond_channels = embedding_dim
        self.mid_channels = hidden_dim
        self.num_blocks = hidden_depth
        self.n_flows = n_flows
        self.conditioning_option = conditioning_option
        self.reshape = "none"
        if self.conditioning_option.lower() != "none":
            raise NotImplementedError("conditioning_option 'parallel'/'sequential' (per-block 1x1 conv on the "
                                      "condition, reference flow_blocks.py:34-35) is 'none' in every reference config "
                                      "and is outside the fused path")
        self.sub_layers = nn.ModuleList()
        for flow in range(self.n_flows):
            mode = "cond" if flow % 2 == 0 else "normal"          # reference flow_blocks.py:27
            self.sub_layers.append(ConditionalFlatDoubleCouplingFlowBlock(
                self.in_channels, self.cond_channels, self.mid_channels, self.num_blocks,
                activation=activation, mode=mode))
        self._reset_last()
        alpha = 0.9 if activation == "lrelu" else 1.0              # 'none' -> identity (IgnoreLeakyRelu)
        self._make_engine(True, embedding_dim, alpha, precision or DEFAULT_PRECISION)

    def forward(self, x, embedding, reverse=False):
        assert len(embedding.shape) == 4                            # reference flow_blocks.py:127
        if not reverse:
            return self._forward_fused(x, embedding)                # fills last_outs ([B,C] each, reference :55)
        self._reset_last()
        from .autograd import flow_reverse
        return flow_reverse(self._engine, x, embedding).unsqueeze(-1).unsqueeze(-1)   # reference :153

    def reverse(self, out, xcond):
        return self(out, xcond, reverse=True)

    def reverse_draws(self, out, xcond, n_draws=None):
        """Sampling loop in one call: ``out`` [n_draws, B, C(,1,1)] latents for ``xcond`` [B, E, 1, 1] ->
        [n_draws, B, C, 1, 1], equal to ``stack([self(out[j], xcond, reverse=True) for j in range(n_draws)])``
        (the loop of the reference's forward_sample, second_stage_video_fc.py:356-378).  The cond-only
        subnetworks of the 'cond' blocks (reference :200-240) are evaluated once per conditioning row.
        Inference only (no autograd graph)."""
        assert len(xcond.shape) == 4                                # reference flow_blocks.py:127
        n_draws = out.shape[0] if n_draws is None else n_draws
        self._reset_last()
        with torch.no_grad():
            x = self._engine.reverse_draws(out, xcond, n_draws)
        return x.unsqueeze(-1).unsqueeze(-1)

    def forward_draws(self, x, xcond, n_draws=None):
        """-> (z [n_draws, B, C, 1, 1], logdet [n_draws, B]) for n_draws inputs per conditioning row (no autograd)."""
        assert len(xcond.shape) == 4
        n_draws = x.shape[0] if n_draws is None else n_draws
        self._reset_last()
        with torch.no_grad():
            z, ld = self._engine.forward_draws(x, xcond, n_draws)
        return z.unsqueeze(-1).unsqueeze(-1), ld


class UnconditionalFlow2(_FusedFlowMixin, nn.Module):
    """Flat unconditional flow.  Reference flow_blocks.py:432-465."""

    def __init__(self, in_channels, hidden_dim, hidden_depth, n_flows, data_init=False, precision=None):
        super().__init__()
        self.in_channels = in_channels
        self.mid_channels = hidden_dim
        self.num_blocks = hidden_depth
        self.n_flows = n_flows
        self.sub_layers = nn.ModuleList()
        for flow in range(self.n_flows):
            self.sub_layers.append(UnconditionalFlatDoubleCouplingFlowBlock2(
                self.in_channels, self.mid_channels, self.num_blocks, data_init))
        self._reset_last()
        self._make_engine(False, 0, 1.0, precision or DEFAULT_PRECISION)

    def forward(self, x, reverse=False):
        assert len(x.shape) == 4                                    # reference flow_blocks.py:592
        if not reverse:
            return self._forward_fused(x, None)                     # last_outs are 4-D here, reference :606
        self._reset_last()
        from .autograd import flow_reverse
        return flow_reverse(self._engine, x, None).unsqueeze(-1).unsqueeze(-1)

    def reverse(self, out):
        return self(out, reverse=True)
