"""Drop-in replacements for the reference wrappers in models/modules/INN/INN.py:
SupervisedTransformer (:19-88) and UnsupervisedTransformer3 (:250-296).  Same constructor arguments,
forward(input, [conditioning,] reverse=False, train=False), reverse, sample, get_last_layer, attribute and
state_dict names; an extra optional ``precision`` knob ('fp32' parity mode | 'bf16')."""
import torch
import torch.nn as nn

from .flow_blocks import ConditionalFlow, UnconditionalFlow2


class SupervisedTransformer(nn.Module):
    """Vanilla conditional version (reference INN.py:19-88)."""

    def __init__(self, config, precision=None):
        super().__init__()
        self.config = config
        in_channels = config["flow_in_channels"]
        mid_channels = config["flow_mid_channels_factor"] * config["flow_in_channels"]
        hidden_depth = config["flow_hidden_depth"]
        n_flows = config["n_flows"]
        conditioning_option = config["flow_conditioning_option"]
        embedding_channels = (config["flow_embedding_channels"] if "flow_embedding_channels" in config
                              else config["flow_in_channels"])
        self.emb_channels = embedding_channels
        self.in_channels = in_channels
        self.flow = ConditionalFlow(in_channels=in_channels, embedding_dim=self.emb_channels,
                                    hidden_dim=mid_channels, hidden_depth=hidden_depth, n_flows=n_flows,
                                    conditioning_option=conditioning_option,
                                    precision=precision or config.get("precision"))

    def sample(self, shape, label):
        device = torch.device("cuda" if torch.cuda.is_available() else "cpu")
        z_tilde = torch.randn(shape).to(device)
        return self.reverse(z_tilde, label)

    def sample_many(self, n_draws, label, generator=None):
        """n_draws samples per conditioning row in one fused call: the batched form of calling
        ``sample((B, C, 1, 1), label)`` n_draws times (reference INN.py:59-63 under the loop of
        second_stage_video_fc.py:360-364).  -> [n_draws, B, C, 1, 1]"""
        B = label.shape[0]
        z_tilde = torch.randn(n_draws, B, self.in_channels, device=label.device, generator=generator)
        return self.flow.reverse_draws(z_tilde, label, n_draws)

    def forward(self, input, conditioning, reverse=False, train=False):
        if reverse:
            return self.reverse(input, conditioning)
        out, logdet = self.flow(input, conditioning)
        if train:
            return self.flow.last_outs, self.flow.last_logdets
        return out, logdet

    def reverse(self, out, label):
        return self.flow(out, label, reverse=True)

    def get_last_layer(self):
        return getattr(self.flow.sub_layers[-1].coupling.t[-1].main[-1], "weight")


class UnsupervisedTransformer3(nn.Module):
    """Unconditional version, data-independent init (reference INN.py:250-296)."""

    def __init__(self, **kwargs):
        super().__init__()
        in_channels = kwargs["flow_in_channels"]
        mid_channels = kwargs["flow_mid_channels"]
        hidden_depth = kwargs["flow_hidden_depth"]
        n_flows = kwargs["n_flows"]
        self.in_channels = in_channels
        self.flow = UnconditionalFlow2(in_channels=in_channels, hidden_dim=mid_channels, hidden_depth=hidden_depth,
                                       n_flows=n_flows, data_init=False, precision=kwargs.get("precision"))

    def sample(self, shape, device="cpu"):
        z_tilde = torch.randn(shape).to(device)
        sample = self.reverse(z_tilde)
        return sample.squeeze(dim=-1).squeeze(dim=-1)

    def forward(self, input, reverse=False, train=False):
        if len(input.shape) == 2:
            input = input[:, :, None, None]
        if reverse:
            return self.reverse(input)
        out, logdet = self.flow(input)
        if train:
            return self.flow.last_outs, self.flow.last_logdets
        return out, logdet

    def reverse(self, out):
        if len(out.shape) == 2:
            out = out[:, :, None, None]
        return self.flow(out, reverse=True)

    def get_last_layer(self):
        return getattr(self.flow.sub_layers[-1].coupling.t[-1].main[-1], "weight")
