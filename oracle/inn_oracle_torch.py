"""TEST INFRASTRUCTURE, NOT PRODUCT -- the timed CPU baseline of bench.py.

The same restatement as ``inn_oracle.py`` (forward and reverse only), written with the torch CPU operators the
reference itself executes -- ``F.linear`` / ``F.leaky_relu`` / ``tanh`` / ``exp`` / ``chunk`` / ``cat`` / advanced
indexing (models/modules/INN/modules.py:171-188, flow_blocks.py:132-153, 322-326) -- so that the CPU arm runs at the
speed of the reference's own eager path (MKL GEMMs, threaded elementwise kernels) and not at numpy's.  The reference
cannot travel to the GPU box (it is Python under /root/reference); this file can.  Only ``tests/`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import it.  Pinned by tests/test_oracle_golden.py against the same
live-reference fixtures as the numpy oracle.

``sd``: the wrapper's state_dict as CPU torch tensors (reference key names).
"""
import torch
import torch.nn.functional as F


def _linears(sd, prefix):
    idx = sorted({int(k[len(prefix) + 6:].split(".")[0]) for k in sd
                  if k.startswith(prefix + ".main.") and k.endswith(".weight")})
    return [(sd[f"{prefix}.main.{i}.weight"], sd[f"{prefix}.main.{i}.bias"]) for i in idx]


def basic_fc_net(sd, prefix, x, use_tanh):
    """BasicFullyConnectedNet.forward -- modules.py:165-191."""
    lin = _linears(sd, prefix)
    h = x
    for n, (w, b) in enumerate(lin):
        h = F.linear(h, w, b)
        if n + 1 < len(lin):
            h = F.leaky_relu(h, 0.01)
    return torch.tanh(h) if use_tanh else h


def _u(xa, cond, mode):
    return torch.cat((xa, cond), dim=1) if mode == "normal" else (cond if mode == "cond" else xa)


def _mode(i, conditional):
    return ("cond" if i % 2 == 0 else "normal") if conditional else "uncond"      # flow_blocks.py:27


def _inv_leaky(x, alpha, reverse):
    scaling = (x >= 0).to(x) + (x < 0).to(x) * alpha                                # flow_blocks.py:377-387
    return x / scaling if reverse else x * scaling


def block_forward(sd, i, x, cond, conditional):
    """flow_blocks.py:255-267 / :513-523."""
    p = f"flow.sub_layers.{i}"
    loc, scale = sd[f"{p}.norm_layer.loc"].reshape(1, -1), sd[f"{p}.norm_layer.scale"].reshape(1, -1)
    h = scale * (x + loc)                                                           # modules.py:292-301
    logdet = torch.sum(torch.log(torch.abs(scale))) * torch.ones(x.shape[0])
    if conditional:
        h = _inv_leaky(h, 0.9, False)
    mode = _mode(i, conditional)
    for j in range(2):                                                              # flow_blocks.py:132-142
        if j % 2 != 0:
            h = torch.cat(torch.chunk(h, 2, dim=1)[::-1], dim=1)
        xa, xk = torch.chunk(h, 2, dim=1)
        u = _u(xa, cond, mode)
        s = basic_fc_net(sd, f"{p}.coupling.s.{j}", u, True)
        t = basic_fc_net(sd, f"{p}.coupling.t.{j}", u, False)
        h = torch.cat((xa, xk * s.exp() + t), dim=1)
        logdet = logdet + s.sum(dim=1)
    return h[:, sd[f"{p}.shuffle.forward_shuffle_idx"]], logdet                     # flow_blocks.py:322-324


def block_reverse(sd, i, y, cond, conditional):
    """flow_blocks.py:268-274 / :524-529."""
    p = f"flow.sub_layers.{i}"
    loc, scale = sd[f"{p}.norm_layer.loc"].reshape(1, -1), sd[f"{p}.norm_layer.scale"].reshape(1, -1)
    h = y[:, sd[f"{p}.shuffle.backward_shuffle_idx"]]
    mode = _mode(i, conditional)
    for j in reversed(range(2)):                                                    # flow_blocks.py:144-153
        if j % 2 == 0:
            h = torch.cat(torch.chunk(h, 2, dim=1)[::-1], dim=1)
        xa, xk = torch.chunk(h, 2, dim=1)
        u = _u(xa, cond, mode)
        s = basic_fc_net(sd, f"{p}.coupling.s.{j}", u, True)
        t = basic_fc_net(sd, f"{p}.coupling.t.{j}", u, False)
        h = torch.cat((xa, (xk - t) * s.neg().exp()), dim=1)
    if conditional:
        h = _inv_leaky(h, 0.9, True)
    return h / scale - loc                                                          # modules.py:312


def n_flows_of(sd):
    n = 0
    while f"flow.sub_layers.{n}.norm_layer.loc" in sd:
        n += 1
    return n


@torch.no_grad()
def flow_forward(sd, x, cond=None):
    """ConditionalFlow.forward (flow_blocks.py:37-57) / UnconditionalFlow2.forward (:448-459): z [B,C], logdet [B]."""
    logdet = torch.zeros(x.shape[0])
    for i in range(n_flows_of(sd)):
        x, ld = block_forward(sd, i, x, cond, cond is not None)
        logdet = logdet + ld
    return x, logdet


@torch.no_grad()
def flow_reverse(sd, z, cond=None):
    """flow_blocks.py:58-63 / :460-462."""
    for i in reversed(range(n_flows_of(sd))):
        z = block_reverse(sd, i, z, cond, cond is not None)
    return z
