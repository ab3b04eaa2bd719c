#!/usr/bin/env python
"""Per-CUDA-source-line instruction counts and stall samples of one kernel in an .ncu-rep (needs -lineinfo)."""
import csv, io, subprocess, sys
rep, kid = sys.argv[1], sys.argv[2]; topn = int(sys.argv[3]) if len(sys.argv) > 3 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv", "--kernel-id", f":::{kid}"],
                     capture_output=True, text=True).stdout
cur = None; out = []
def num(x):
    try: return int(x)
    except ValueError: return 0
for r in csv.reader(io.StringIO(txt)):
    if len(r) == 2 and r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if len(r) > 8 and r[0].isdigit() and r[2] == '-':
        out.append((cur, int(r[0]), r[1].strip()[:90], num(r[6]), num(r[7])))
tot = sum(o[4] for o in out); ts = sum(o[3] for o in out)
print('total warp instr', tot, 'samples', ts)
for o in sorted(out, key=lambda o: -o[4])[:topn]:
    print(f"{o[0]}:{o[1]:4d} instr {o[4]:9d} ({100*o[4]/tot:4.1f}%) samples {o[3]:5d} ({100*o[3]/max(ts,1):4.1f}%) | {o[2]}")
