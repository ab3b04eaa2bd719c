"""FlowLoss with the reference's interface (models/modules/INN/loss.py:8-42, nll :86-97), computed by
``ipoke_inn_flow_loss`` (one reduction kernel that also emits dz / dlogdet for the backward pass)."""
import torch
import torch.nn as nn

from . import _capi


_SCRATCH = {}       # (device, stream) -> zeroed partial-sum scratch of the multi-CTA loss kernel (it leaves it zeroed)


def _scratch(device):
    key = (device, torch.cuda.current_stream(device).cuda_stream)
    buf = _SCRATCH.get(key)
    if buf is None:
        if len(_SCRATCH) >= 16:
            _SCRATCH.pop(next(iter(_SCRATCH)))
        buf = torch.zeros(_capi.lib().ipoke_inn_flow_loss_scratch_bytes() // 8, dtype=torch.float64, device=device)
        _SCRATCH[key] = buf
    return buf


def _loss_kernel(z2, logdet, ref2, nll_w, ld_w, want_grads):
    """One launch: loss, mean nll, -mean logdet, mean nll of the reference sample (+ dz, dlogdet) -> out4."""
    B, C = z2.shape
    out4 = torch.empty(4, dtype=torch.float32, device=z2.device)
    dz = torch.empty_like(z2) if want_grads else None
    dld = torch.empty_like(logdet) if want_grads else None
    stream = torch.cuda.current_stream(z2.device).cuda_stream
    scratch = _scratch(z2.device) if B * C > 32768 else None          # one CTA is enough below that
    _capi.check(_capi.lib().ipoke_inn_flow_loss_ws(z2.data_ptr(), logdet.data_ptr(), ref2.data_ptr(), B, C, float(nll_w),
                                                   float(ld_w), out4.data_ptr(), dz.data_ptr() if want_grads else None,
                                                   dld.data_ptr() if want_grads else None,
                                                   scratch.data_ptr() if scratch is not None else None, stream), "flow_loss")
    return out4, dz, dld


class _FlowLossFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z2, logdet, ref2, nll_w, ld_w):
        out4, dz, dld = _loss_kernel(z2, logdet, ref2, nll_w, ld_w, True)
        ctx.save_for_backward(dz, dld)
        ctx.mark_non_differentiable(ref2)
        return out4

    @staticmethod
    def backward(ctx, g):
        dz, dld = ctx.saved_tensors
        return dz * g[0], dld * g[0], None, None, None      # only out4[0] (the loss) is differentiated


class FlowLoss(nn.Module):
    def __init__(self, spatial_mean=False, logdet_weight=1., nll_weight=1., radial=False):
        super().__init__()
        if spatial_mean or radial:
            raise NotImplementedError("spatial_mean / radial are off in every flat-INN config "
                                      "(reference models/second_stage_video_fc.py:150)")
        self.spatial_mean = spatial_mean
        self.logdet_weight = logdet_weight
        self.nll_weight = nll_weight
        self.radial = radial

    def forward(self, sample, logdet):
        assert len(logdet.shape) == 1
        if sample.device.type != "cuda":
            raise RuntimeError("ipoke_inn_b200.FlowLoss runs on CUDA only")
        B = sample.shape[0]
        z2 = sample.reshape(B, -1).contiguous()
        ld = logdet.contiguous()
        # the reference draws a fresh N(0,1) sample for a logged reference NLL (loss.py:27,32) -- same RNG use; its
        # reduction rides in the loss kernel's launch
        ref2 = torch.randn_like(sample).reshape(B, -1)
        if torch.is_grad_enabled() and (z2.requires_grad or ld.requires_grad):
            out4 = _FlowLossFn.apply(z2, ld, ref2, self.nll_weight, self.logdet_weight)
        else:
            out4, _, _ = _loss_kernel(z2, ld, ref2, self.nll_weight, self.logdet_weight, False)
        loss = out4[0]
        logged = out4.detach()
        log = {"flow_loss": logged[0], "reference_nll_loss": logged[3], "nlogdet_loss": logged[2],
               "nll_loss": logged[1], "logdet_w": self.logdet_weight, "nll_w": self.nll_weight}
        return loss, log
