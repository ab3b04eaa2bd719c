"""torch.autograd glue: one Function per flow call so that ``loss.backward()`` of the caller (Lightning's
training loop in the reference, models/second_stage_video_fc.py:448-490) reaches the fused backward kernels."""
import torch


def _needs_grad(engine, x, cond):
    if not torch.is_grad_enabled():
        return False
    if x.requires_grad or (cond is not None and cond.requires_grad):
        return True
    return any(p.requires_grad for p in engine.flow.parameters())


def flow_forward(engine, x, cond):
    """-> z [B,C], logdet [B], block_outs [n,B,C] | None, block_logdets [n,B] | None."""
    if _needs_grad(engine, x, cond):
        from .backward import FlowFunction
        if getattr(engine, "flat_grads_only", False):
            # the optimizer reads the flat gradient buffer itself: ONE differentiable anchor stands in for the 336
            # parameters as the Function's input, so that backward() runs without per-parameter autograd edges
            return FlowFunction.apply(engine, x, cond, engine.grad_anchor(x.device))
        params = engine.grad_layout()[3]
        return FlowFunction.apply(engine, x, cond, *params)
    with torch.no_grad():
        return engine.forward(x, cond, want_blocks=False)      # per-block lists: lazily, see flow_blocks._FusedFlowMixin


def flow_reverse(engine, z, cond):
    if torch.is_grad_enabled() and (z.requires_grad or (cond is not None and cond.requires_grad)):
        raise NotImplementedError("ipoke_inn_b200: gradients through the reverse (sampling) direction are not "
                                  "implemented; the reference only samples under torch.no_grad() "
                                  "(models/second_stage_video_fc.py:359)")
    with torch.no_grad():
        return engine.reverse(z, cond)
