#!/usr/bin/env python
"""Read an .ncu-rep here (no GPU): headline metrics + top stalled SASS lines of one kernel id."""
import csv, subprocess, sys, io
rep = sys.argv[1]; kid = sys.argv[2] if len(sys.argv) > 2 else None; topn = int(sys.argv[3]) if len(sys.argv) > 3 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw))); hdr = rows[0]
want = ['ID', 'Kernel Name', 'gpu__time_duration.sum', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_sectors.sum', 'l1tex__m_xbar2l1tex_read_bytes.sum', 'launch__grid_size', 'launch__registers_per_thread',
        'sm__cycles_elapsed.avg.per_second', 'lts__t_sectors_op_write.sum', 'lts__t_sectors_op_read.sum', 'lts__throughput.avg.pct_of_peak_sustained_elapsed']
idx = [(w, hdr.index(w)) for w in want if w in hdr]
for r in rows[2:]:
    print({w.split('.')[0][-28:]: r[i][:34] for w, i in idx})
if kid is None: sys.exit()
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-id", f":::{kid}"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src))); hdr = rows[1]
iS = hdr.index('# Samples'); isrc = hdr.index('Source'); iex = hdr.index('Instructions Executed')
stall = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
data = []
for r in rows[2:]:
    if len(r) != len(hdr) or not r[iS].isdigit(): break
    data.append(r)
print('total samples', sum(int(r[iS]) for r in data), 'warp instr', sum(int(r[iex]) for r in data))
for n, r in sorted(enumerate(data), key=lambda kv: -int(kv[1][iS]))[:topn]:
    st = sorted(((hdr[i][6:], int(r[i])) for i in stall if int(r[i]) > 0), key=lambda kv: -kv[1])[:3]
    print(n, r[iS], r[iex], r[isrc].strip()[:64], st)
