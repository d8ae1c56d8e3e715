"""Aggregate `ncu --page source --csv --print-source cuda,sass` by CUDA source line:
warp-instructions executed and stall samples per line (top N)."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
n_top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = None
out = []
for r in rows:
    if r and r[0] == "Line No":
        hdr = r
        ii = hdr.index("Instructions Executed")
        si = hdr.index("# Samples")
        continue
    if hdr and len(r) > ii and r[0].strip().isdigit():
        try:
            out.append((int(r[0]), r[1].strip()[:110], int(r[ii] or 0), int(r[si] or 0)))
        except ValueError:
            pass
tot = sum(o[2] for o in out)
tots = sum(o[3] for o in out)
print(f"total warp-instructions {tot}, samples {tots}")
for ln, src, n, s in sorted(out, key=lambda o: -o[2])[:n_top]:
    print(f"{ln:5d} {100 * n / tot:5.1f}% inst {100 * s / max(tots, 1):5.1f}% smp  {src}")
