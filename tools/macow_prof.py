#!/usr/bin/env python
"""Two forward and two inverse calls of one MaCow unit (the bench's z32 8x8 unit) for an ncu capture:
ncu --set full --import-source on -k regex:macow_unit_kernel -c 4 python tools/macow_prof.py --batch 20"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ipoke_inn_b200 import macow  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=20)
    ap.add_argument("--channels", type=int, default=32)
    ap.add_argument("--size", type=int, default=8)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.manual_seed(1)
    unit = macow.MaCowUnit(a.channels, [2, 3], activation="elu")
    with torch.no_grad():
        for k, v in unit.state_dict().items():
            if k.endswith("initialized"):
                v.fill_(1)
            elif k.endswith("weight_g"):
                v.copy_(0.15 + 0.3 * torch.rand_like(v))
            elif k.endswith("bias") or k.endswith("log_scale"):
                v.copy_(0.1 * torch.randn_like(v))
    unit = unit.to(dev).eval()
    x = torch.randn(a.batch, a.channels, a.size, a.size, device=dev)
    with torch.no_grad():
        for _ in range(2):
            z, _ = unit(x)
            unit(z, reverse=True)
    torch.cuda.synchronize()
    print("ok")


if __name__ == "__main__":
    main()
