#!/usr/bin/env python
"""Stage-by-stage diagnosis on the GPU box: run ONE flow block and compare every scratch buffer of the
workspace (MLP input, hidden activations, state) with numpy.  Test infrastructure (uses the oracle)."""
import ctypes
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import ipoke_inn_b200 as P
from ipoke_inn_b200 import _capi
from oracle import inn_oracle as O
from conftest import load_golden
from test_host_logic import build


def bf16_view(buf, off, rows, ld):
    t = buf[off:off + rows * ld * 2].view(torch.bfloat16).reshape(rows, ld)
    return t.float().cpu().numpy()


def run(name, precision, block=0):
    dev = torch.device("cuda:0")
    sd, d = load_golden(name)
    net = build(name, precision=precision)
    net.load_state_dict({k: torch.from_numpy(v) for k, v in sd.items()})
    net = net.to(dev)
    eng = net.flow.engine
    x = torch.from_numpy(d["x"]).to(dev)
    cond = torch.from_numpy(d["cond"]).to(dev) if "cond" in d else None
    B, C = d["x"].shape
    _, x2, c2 = eng._prep(x, cond)
    eng._ensure(dev)
    eng._init_epoch = sum(b.norm_layer._state_epoch for b in net.flow.sub_layers)
    eng.pack()
    # input of `block` from the oracle
    cnp = None if cond is None else d["cond"].reshape(B, -1)
    cur = d["x"]
    for k in range(block):
        cur, _ = O.block_forward(sd, k, cur, cnp, cnp is not None)
    xin = torch.from_numpy(np.ascontiguousarray(cur)).to(dev)
    z = torch.empty(B, C, device=dev)
    ld = torch.empty(B, device=dev)
    eng._forward_raw(xin, c2, z, ld, None, None, block, block + 1)
    torch.cuda.synchronize()
    lay = (ctypes.c_longlong * 20)()
    _capi.lib().ipoke_inn_debug_layout(eng._plan, B, lay, 20)
    S0, S1, Uh, Ul, A0h, A0l, A1h, A1l, LD, TOT, ldU, ldACT, Hp, C2p, Ep, Np, ks, Gmax = [int(v) for v in lay[:18]]
    ws = eng._workspace(B)
    print(f"== {name} precision={precision} block={block} B={B} C={C} ldU={ldU} ldACT={ldACT} Hp={Hp} ksplit={ks} "
          f"launches={eng.launches}")
    z_ref, ld_ref = O.block_forward(sd, block, cur, cnp, cnp is not None)
    print("   block out  max|err| = %.3e (max|ref| %.3e)" % (np.abs(z.cpu().numpy() - z_ref).max(), np.abs(z_ref).max()))
    print("   logdet     max|err| = %.3e (max|ref| %.3e)" % (np.abs(ld.cpu().numpy() - ld_ref).max(), np.abs(ld_ref).max()))
    # recompute the coupling by hand to check U / ACT of the LAST chain that ran
    p = f"flow.sub_layers.{block}"
    loc = sd[p + ".norm_layer.loc"].reshape(-1)
    scale = sd[p + ".norm_layer.scale"].reshape(-1)
    h = scale * (cur + loc)
    conditional = cnp is not None
    if conditional:
        h = O.inv_leaky_forward(h)
    mode = O._block_mode(block, conditional)
    half = C // 2
    xa, xb = h[:, :half], h[:, half:]
    u0 = O._coupling_input(xa, cnp, mode)
    s0 = O.basic_fc_net(sd, p + ".coupling.s.0", u0, True)
    t0 = O.basic_fc_net(sd, p + ".coupling.t.0", u0, False)
    xb2 = xb * np.exp(s0) + t0
    u1 = O._coupling_input(xb2, cnp, mode)
    s1 = O.basic_fc_net(sd, p + ".coupling.s.1", u1, True)
    t1 = O.basic_fc_net(sd, p + ".coupling.t.1", u1, False)
    xa2 = xa * np.exp(s1) + t1
    st = ws[S0:S0 + B * C * 4].view(torch.float32).reshape(B, C).cpu().numpy()
    print("   state[:, :half] (xa') max|err| = %.3e ; state[:, half:] (xb') max|err| = %.3e" %
          (np.abs(st[:, :half] - xa2).max(), np.abs(st[:, half:] - xb2).max()))
    U = bf16_view(ws, Uh, B, ldU)
    if ks == 2:
        U = U + bf16_view(ws, Ul, B, ldU)
    if mode != "cond":
        print("   U x-part (xb')  max|err| = %.3e ; pad max = %.3e" %
              (np.abs(U[:, :half] - xb2).max(), np.abs(U[:, half:C2p]).max() if C2p > half else 0.0))
    if conditional:
        print("   U cond part     max|err| = %.3e" % np.abs(U[:, C2p:C2p + cnp.shape[1]] - cnp).max())
    # hidden activations of the last chain: layer index depth -> ACT[depth & 1]
    idx = O._linear_indices(sd, p + ".coupling.s.1")
    depth = len(idx) - 2
    H = sd[p + ".coupling.s.1.main.0.weight"].shape[0]

    def hidden(prefix, u, upto):
        a = u
        for n in range(upto + 1):
            w = sd[f"{prefix}.main.{idx[n]}.weight"]
            b = sd[f"{prefix}.main.{idx[n]}.bias"]
            a = a @ w.T + b
            a = np.where(a >= 0, a, 0.01 * a)
        return a
    for lyr in (depth, depth - 1):
        if lyr < 0:
            continue
        offh, offl = (A0h, A0l) if (lyr & 1) == 0 else (A1h, A1l)
        A = bf16_view(ws, offh, B, ldACT)
        if ks == 2:
            A = A + bf16_view(ws, offl, B, ldACT)
        if mode == "cond":
            groups = [("s.0", u0, 0), ("t.0", u0, 1), ("s.1", u1, 2), ("t.1", u1, 3)]
        else:
            groups = [("s.1", u1, 0), ("t.1", u1, 1)]
        for nm, u, g in groups:
            ref = hidden(p + ".coupling." + nm, u, lyr)
            got = A[:, g * Hp:g * Hp + H]
            print("   ACT layer %d net %s  max|err| = %.3e (max|ref| %.3e)  pad max %.3e" %
                  (lyr, nm, np.abs(got - ref).max(), np.abs(ref).max(),
                   np.abs(A[:, g * Hp + H:(g + 1) * Hp]).max() if Hp > H else 0.0))


if __name__ == "__main__":
    for name in ("uncond_small", "cfgS_cond"):
        for prec in ("bf16", "fp32"):
            for blk in (0, 1):
                try:
                    run(name, prec, blk)
                except Exception as e:       # keep going: this is a diagnosis tool
                    print(f"!! {name} {prec} block {blk}: {type(e).__name__}: {e}")
                    if "CUDA" in str(e) or "cuda" in str(e):
                        sys.exit(1)
