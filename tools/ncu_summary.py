#!/usr/bin/env python
"""Turn ncu outputs brought back in gpurun_out/ into the committed summaries under profiles/ (runs without a GPU).

  python tools/ncu_summary.py launches <launch-list.csv> <out.csv>      last full step (130 launches) of the list:
                                                                        per-kernel-kind count, total us, share
  python tools/ncu_summary.py full <report.ncu-rep> <out.csv> [traffic.json key]
                                                                        headline metrics per captured launch (+ the
                                                                        average dram traffic of the hidden GEMM)
"""
import collections
import csv
import io
import json
import subprocess
import sys

KIND = [("inn_gemm_kernel<1, 0", "hidden"), ("inn_gemm_kernel<2, 0", "hidden"), ("inn_hidden_pair", "hidden"),
        ("inn_tail_cluster", "tail"),
        ("inn_gemm_kernel<1, 1", "tail"), ("inn_gemm_kernel<1, 2", "tail"), ("inn_gemm_kernel<2, 1", "tail"),
        ("inn_gemm_kernel<2, 2", "tail"), ("inn_boundary", "boundary")]


def kind_of(name):
    for pat, k in KIND:
        if pat in name:
            return k
    return None


def launches(src, dst, per_step=130):
    rows = list(csv.reader(open(src)))
    h0 = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    h = rows[h0]
    ours = [(r[h.index("Kernel Name")], r[h.index("Grid Size")], float(r[-1]) / 1e3) for r in rows[h0 + 1:]
            if len(r) == len(h) and kind_of(r[h.index("Kernel Name")])]
    step = ours[-per_step:]
    agg = collections.OrderedDict()
    for name, grid, us in step:
        k = (kind_of(name), name.split("(")[0][-40:], grid)
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += us
    tot = sum(a[1] for a in agg.values())
    with open(dst, "w") as f:
        w = csv.writer(f)
        w.writerow(["kind", "kernel", "grid", "launches_per_step", "total_us (ncu: cold cache, serialised)", "share_of_step"])
        for (kind, name, grid), (n, us) in agg.items():
            w.writerow([kind, name, grid, n, f"{us:.1f}", f"{us / tot:.3f}"])
        w.writerow(["all", "", "", len(step), f"{tot:.1f}", "1.000"])
    print(open(dst).read())


WANT = ["ID", "Kernel Name", "Grid Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "l1tex__m_xbar2l1tex_read_bytes.sum", "lts__t_sectors.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__cycles_elapsed.avg.per_second", "launch__registers_per_thread"]


def full(rep, dst, traffic_key=None):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h, units = rows[0], rows[1]
    idx = [(w, h.index(w)) for w in WANT if w in h]
    hidden = []
    with open(dst, "w") as f:
        w = csv.writer(f)
        w.writerow([f"{n} [{units[i]}]" if units[i] else n for n, i in idx])
        for r in rows[2:]:
            w.writerow([r[i][:70] for _, i in idx])
            if kind_of(r[h.index("Kernel Name")]) == "hidden":
                def val(name):
                    v = float(r[h.index(name)])
                    u = units[h.index(name)].lower()
                    return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
                hidden.append(val("dram__bytes_read.sum") + val("dram__bytes_write.sum"))
    print(open(dst).read()[:3000])
    if traffic_key and hidden:
        path = "profiles/r02_traffic.json"
        try:
            d = json.load(open(path))
        except Exception:
            d = {}
        d[traffic_key] = {"hidden_gemm_dram_bytes_per_launch": sum(hidden) / len(hidden), "launches_captured": len(hidden),
                          "source": dst}
        json.dump(d, open(path, "w"), indent=1)
        print("traffic", d[traffic_key])


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2], sys.argv[3])
    else:
        full(sys.argv[2], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else None)
