"""Summarise `ncu -i X.ncu-rep --page source --csv --print-source sass` output (first kernel):
total warp instructions, stall breakdown, and the instructions with most stall samples."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
n_top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
blk = rows[starts[0]:(starts[1] if len(starts) > 1 else len(rows))]
print(blk[0][1])
hdr, data = blk[1], [r for r in blk[2:] if len(r) > 10]
ix = {h: i for i, h in enumerate(hdr)}
I = lambda r, k: int(r[ix[k]] or 0)  # noqa: E731
tot_inst = sum(I(r, "Instructions Executed") for r in data)
tot_samp = sum(I(r, "# Samples") for r in data)
n_warps = I(data[0], "Instructions Executed")
print(f"sass lines {len(data)}  warp-instructions {tot_inst}  per warp {tot_inst / n_warps:.0f}  samples {tot_samp}")
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = {s: sum(I(r, s) for r in data) for s in stalls}
print("stall breakdown:", [(k, v, f"{100 * v / tot_samp:.1f}%") for k, v in sorted(agg.items(), key=lambda x: -x[1])[:8]])
print("top instructions by samples:  samples  executed  avg_threads  sass   [top stalls]")
for r in sorted(data, key=lambda r: -I(r, "# Samples"))[:n_top]:
    st = sorted([(I(r, s), s) for s in stalls], reverse=True)[:2]
    print(r[ix["# Samples"]].rjust(6), r[ix["Instructions Executed"]].rjust(8), r[ix["Avg. Threads Executed"]].rjust(5),
          r[ix["Source"]].strip()[:72].ljust(72), st)
