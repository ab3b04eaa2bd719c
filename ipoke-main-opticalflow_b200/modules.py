"""Parameter containers mirroring the reference's sub-modules of the flat INN.

Same class names, constructor arguments, parameter / buffer names, shapes, dtypes and RNG consumption
order as the reference, so a reference ``state_dict`` loads with ``strict=True`` and the same seed gives
bit-identical weights and permutations.  The arithmetic does NOT live here: every block of the stack is
executed by the fused sm_100a kernels behind ``libipoke_inn.so`` through the parent flow
(``flow_blocks.ConditionalFlow`` / ``UnconditionalFlow2``); calling one of these sub-modules directly
raises -- there is no eager or CPU fallback.
"""
import torch
import torch.nn as nn


def _fused_only(name):
    raise RuntimeError(
        f"{name}.forward is not a standalone op in ipoke_inn_b200: it runs fused inside the parent flow's "
        "CUDA kernels (call the ConditionalFlow / UnconditionalFlow2 / *Transformer module instead).")


class BasicFullyConnectedNet(nn.Module):
    """Reference: models/modules/INN/modules.py:165-191.
    Linear(dim,H) LeakyReLU [Linear(H,H) LeakyReLU]*depth Linear(H,out) [Tanh] -- kept as ``main`` (nn.Sequential)
    so that keys are ``main.{0,2,...}.{weight,bias}`` and ``main[-1]`` is the last Linear of a t-net."""

    def __init__(self, dim, depth, hidden_dim=256, use_tanh=False, use_bn=False, out_dim=None,
                 data_init=False, last_zero=False):
        super().__init__()
        if use_bn or data_init:
            raise NotImplementedError("use_bn / data_init variants are not used on the flat INN path "
                                      "(reference modules.py:170,173; UnconditionalFlow2 passes data_init=False)")
        layers = [nn.Linear(dim, hidden_dim), nn.LeakyReLU()]
        for _ in range(depth):
            layers += [nn.Linear(hidden_dim, hidden_dim), nn.LeakyReLU()]
        layers.append(nn.Linear(hidden_dim, dim if out_dim is None else out_dim))
        if use_tanh:
            layers.append(nn.Tanh())
        self.main = nn.Sequential(*layers)
        self.use_tanh = use_tanh

    def linears(self):
        return [m for m in self.main if isinstance(m, nn.Linear)]

    def forward(self, x):
        _fused_only("BasicFullyConnectedNet")


class ActNorm(nn.Module):
    """Reference: models/modules/INN/modules.py:245-316.  loc/scale [1,C,1,1], buffer ``initialized`` (uint8).
    Data-dependent initialisation is performed by the parent flow (engine.FlowEngine._actnorm_init)."""

    def __init__(self, num_features, logdet=False, affine=True):
        assert affine
        super().__init__()
        self.logdet = logdet
        self.loc = nn.Parameter(torch.zeros(1, num_features, 1, 1))
        self.scale = nn.Parameter(torch.ones(1, num_features, 1, 1))
        self.register_buffer("initialized", torch.tensor(0, dtype=torch.uint8))
        self._state_epoch = 0      # bumped whenever a state_dict is loaded (the host mirror of `initialized` is stale)

    def _load_from_state_dict(self, *args, **kwargs):
        self._state_epoch += 1
        return super()._load_from_state_dict(*args, **kwargs)

    def forward(self, input, reverse=False):
        _fused_only("ActNorm")


class Shuffle(nn.Module):
    """Reference: models/modules/INN/flow_blocks.py:314-326.  idx = torch.randperm(C) at construction."""

    def __init__(self, in_channels, **kwargs):
        super().__init__()
        self.in_channels = in_channels
        idx = torch.randperm(in_channels)
        self.register_buffer("forward_shuffle_idx", idx)
        self.register_buffer("backward_shuffle_idx", torch.argsort(idx))

    def forward(self, x, reverse=False, conditioning=None):
        _fused_only("Shuffle")


class InvLeakyRelu(nn.Module):
    """Reference: flow_blocks.py:372-387 (alpha = 0.9; contributes 0 to logdet)."""

    def __init__(self, alpha=0.9):
        super().__init__()
        self.alpha = alpha

    def forward(self, input, reverse=False):
        _fused_only("InvLeakyRelu")


class IgnoreLeakyRelu(nn.Module):
    """Reference: flow_blocks.py:356-369 (identity)."""

    def __init__(self, alpha=0.9):
        super().__init__()

    def forward(self, input, reverse=False):
        _fused_only("IgnoreLeakyRelu")
