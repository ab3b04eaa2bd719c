#!/usr/bin/env python
"""Per-kernel totals of an ncu launch list (--metrics gpu__time_duration.sum --csv) of the MaCow flow tools."""
import collections
import csv
import re
import sys


def main(path, skip_pack=True):
    hdr, agg = None, collections.defaultdict(list)
    for r in csv.reader(open(path)):
        if r and r[0] == "ID":
            hdr = r
            continue
        if hdr and len(r) == len(hdr):
            d = dict(zip(hdr, r))
            val = float(d["Metric Value"].replace(",", ""))
            val = {"ns": val / 1000, "us": val, "ms": val * 1000}.get(d["Metric Unit"], val)
            name = d["Kernel Name"]
            m = re.search(r"nice_gemm_kernel<(\d+), *(\d+), *(\d+), *(\d+)", name)
            key = f"nice_gemm<src={m.group(1)},epi={m.group(2)},{m.group(3)}x{m.group(4)}>" if m else name.split("(")[0][:64]
            agg[key].append(val)
    tot = sum(sum(v) for v in agg.values())
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print(f"{k:66s} n={len(v):5d} total {sum(v):11.1f} us  avg {sum(v) / len(v):9.1f} us  share {sum(v) / tot:.3f}")


if __name__ == "__main__":
    main(sys.argv[1])
