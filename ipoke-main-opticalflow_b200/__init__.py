"""ipoke_inn_b200 -- B200-native (sm_100a) implementation of iPOKE's flat conditional-INN hot path.

Drop-in for the reference's ``models/modules/INN/INN.py`` wrappers and ``flow_blocks.py`` stacks; all
arithmetic runs in hand-written CUDA kernels behind the C ABI of ``include/ipoke_inn.h``.
"""
import sys
import types

from . import _capi
from .INN import SupervisedTransformer, UnsupervisedTransformer3
from .flow_blocks import (ConditionalFlow, UnconditionalFlow2, ConditionalFlatDoubleCouplingFlowBlock,
                          UnconditionalFlatDoubleCouplingFlowBlock2, ConditionalDoubleVectorCouplingBlock,
                          ConditionalDoubleVectorCouplingBlock_cond, DoubleVectorCouplingBlock2)
from .loss import FlowLoss
from .latent_interface import SecondStageLatentInterface, FlowAELatentInterface, ThirdStageLatentInterface
from .optim import FusedAdam, linear_var
from .optim_sharded import ShardedAdam, plan_pieces
from .parallel import GradSync, broadcast_parameters, shard_batch, shard_bounds
from .macow import MaCowUnit, MaskedConvFlow, ActNorm2dFlow
from .macow_flow import (UnsupervisedMaCowTransformer3, SupervisedMacowTransformer, MultiScaleInternal, MultiScalePrior, MaCowStep, NICE2d,
                         NICEConvBlock)
from .modules import ActNorm, BasicFullyConnectedNet, IgnoreLeakyRelu, InvLeakyRelu, Shuffle

__all__ = ["SupervisedTransformer", "UnsupervisedTransformer3", "ConditionalFlow", "UnconditionalFlow2",
           "FlowLoss", "ActNorm", "BasicFullyConnectedNet", "Shuffle", "InvLeakyRelu", "IgnoreLeakyRelu",
           "FusedAdam", "linear_var", "install_as_reference", "SecondStageLatentInterface",
           "FlowAELatentInterface", "ThirdStageLatentInterface", "MaCowUnit", "MaskedConvFlow", "ActNorm2dFlow", "UnsupervisedMaCowTransformer3", "SupervisedMacowTransformer", "MultiScaleInternal", "MultiScalePrior", "MaCowStep", "NICE2d", "NICEConvBlock", "GradSync", "ShardedAdam", "broadcast_parameters", "shard_batch", "shard_bounds"]


def install_as_reference():
    """Shadow the reference's import paths so its LightningModules pick up these classes unchanged:
    ``from models.modules.INN.INN import SupervisedTransformer`` (models/second_stage_video_fc.py:16) etc.
    The flat-INN wrappers, the MaCow wrappers (``UnsupervisedMaCowTransformer3`` for INN_test.py / opticalFlowINN.py /
    models/flow_motion.py, ``SupervisedMacowTransformer`` for models/second_stage_video.py:122) and FlowLoss are replaced;
    other names of those modules keep resolving to the reference (if it is importable)."""
    from . import INN as _inn, loss as _loss, macow_flow as _mf
    for name, ours, attrs in (("models.modules.INN.INN", _inn, ("SupervisedTransformer", "UnsupervisedTransformer3")),
                              ("models.modules.INN.INN", _mf, ("UnsupervisedMaCowTransformer3", "SupervisedMacowTransformer")),
                              ("models.modules.INN.loss", _loss, ("FlowLoss",))):
        mod = sys.modules.get(name)
        if mod is None:
            try:
                mod = __import__(name, fromlist=["_"])
            except Exception:
                mod = types.ModuleType(name)
                sys.modules[name] = mod
        for a in attrs:
            setattr(mod, a, getattr(ours, a))
