"""Checkpoint / wire-format compatibility: the fused flows keep the reference's parameter and
buffer names, so a Lightning ``.ckpt`` written by the reference's runs loads into them unchanged.

  * :func:`select_checkpoint` -- the reference's checkpoint choice (experiments/experiment.py:107-142): ``last.ckpt``
    when asked for, else the single ``*.yaml`` best-k index (ModelCheckpoint.to_yaml, utils/callbacks.py:10-23) whose
    best-scored existing file wins.
  * :func:`load_flow_from_lightning` -- pull the INN's entries out of a LightningModule ``state_dict``
    (``flow.flow.sub_layers...`` under ``PokeMotionModelFC.flow`` = SupervisedTransformer, second_stage_video_fc.py:121;
    ``INN.flow.sub_layers...`` under ``FCAEINNModel.INN``, fully_connected_ae_INN.py:81) and load them strictly.
  * :func:`export_flow_to_lightning` -- write the INN's entries back under the same prefix (frozen sub-models are
    omitted by the reference's ``on_save_checkpoint``, second_stage_video_fc.py:1233-1263, so only the flow travels).
"""
import glob
import os

import torch
import yaml


def select_checkpoint(ckpt_dir, last_ckpt=False):
    """-> path of the checkpoint the reference would load from ``ckpt_dir`` (experiments/experiment.py:107-142)."""
    names = glob.glob(os.path.join(ckpt_dir, "*.yaml"))
    last = os.path.join(ckpt_dir, "last.ckpt")
    if last_ckpt and os.path.isfile(last):
        names = [last]
    elif last_ckpt:
        raise ValueError("intending to load last ckpt, but no last ckpt found. Aborting....")
    if len(names) != 1:
        raise ValueError(("Not enough" if len(names) < 1 else "Too many") + " checkpoint files found! Aborting...")
    name = names[0]
    if name.endswith(".yaml"):
        with open(name, "r") as f:
            ckpts = yaml.load(f, Loader=yaml.FullLoader) or {}
        ckpts = dict(ckpts)
        chosen = None
        while ckpts:                                       # best (lowest) monitored value whose file still exists
            best = min(ckpts, key=lambda k: ckpts[k])
            if os.path.isfile(best):
                chosen = best
                break
            del ckpts[best]
        if chosen is None:
            raise ValueError(f'No valid files contained in ckpt-name-holding file "{name}"')
        name = chosen
    return name


def _state_dict_of(ckpt):
    if isinstance(ckpt, (str, os.PathLike)):
        ckpt = torch.load(ckpt, map_location="cpu", weights_only=False)
    return ckpt["state_dict"] if isinstance(ckpt, dict) and "state_dict" in ckpt else ckpt


def flow_entries(state_dict, prefix):
    """Entries of ``state_dict`` under ``prefix`` (e.g. 'flow.' or 'INN.'), prefix stripped."""
    return {k[len(prefix):]: v for k, v in state_dict.items() if k.startswith(prefix)}


def load_flow_from_lightning(module, ckpt, prefix="flow.", strict=True):
    """Load the INN part of a LightningModule checkpoint (path, checkpoint dict or state_dict) into ``module``
    (SupervisedTransformer / UnsupervisedTransformer3, or the MaCow wrappers UnsupervisedMaCowTransformer3 /
    SupervisedMacowTransformer).  The packed parameters are rebuilt at the next call."""
    sd = flow_entries(_state_dict_of(ckpt), prefix)
    if not sd:
        raise KeyError(f"no entries under prefix '{prefix}' in the checkpoint")
    out = module.load_state_dict(sd, strict=strict)
    flow = module.flow
    (flow.engine if hasattr(flow, "engine") else flow).invalidate()       # flat INN: the engine; MaCow: the flow's program
    return out


def export_flow_to_lightning(module, ckpt=None, prefix="flow.", path=None):
    """Write ``module``'s entries under ``prefix`` into a Lightning-style checkpoint dict (a new one when ``ckpt`` is
    None; other keys of an existing one are kept) and optionally ``torch.save`` it to ``path``."""
    ckpt = dict(ckpt) if ckpt is not None else {"state_dict": {}}
    sd = dict(ckpt.get("state_dict", {}))
    for k in [k for k in sd if k.startswith(prefix)]:
        del sd[k]
    for k, v in module.state_dict().items():
        sd[prefix + k] = v.detach().cpu().clone()
    ckpt["state_dict"] = sd
    if path is not None:
        torch.save(ckpt, path)
    return ckpt
