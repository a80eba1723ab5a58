"""Summarise the source page of an .ncu-rep: top stalled SASS instructions and stall-reason totals per kernel.
    python tools/ncu_stalls.py gpurun_out/x.ncu-rep [kernel-index] [top-n]
"""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]
which = int(sys.argv[2]) if len(sys.argv) > 2 else 0
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
blocks, cur = [], None
for row in csv.reader(io.StringIO(out)):
    if row and row[0] == "Kernel Name":
        cur = {"name": row[1], "hdr": None, "rows": []}
        blocks.append(cur)
    elif cur is not None and cur["hdr"] is None:
        cur["hdr"] = row
    elif cur is not None and row:
        cur["rows"].append(row)
b = blocks[which]
h = {n: i for i, n in enumerate(b["hdr"])}
print("kernel:", b["name"], " (", len(blocks), "kernels in report )")
stall_cols = [n for n in b["hdr"] if n.startswith("stall_") and "Not Issued" not in n]
tot = collections.Counter()
samples = 0
for r in b["rows"]:
    samples += int(r[h["# Samples"]] or 0)
    for c in stall_cols:
        tot[c] += int(r[h[c]] or 0)
print("total samples", samples)
for c, v in tot.most_common(10):
    print(f"  {c:28s} {v:8d}  {100.0 * v / max(1, samples):5.1f}%")
rows = sorted(b["rows"], key=lambda r: -int(r[h["# Samples"]] or 0))[:topn]
print("top instructions by samples:")
for r in rows:
    n = int(r[h["# Samples"]] or 0)
    reasons = sorted(((int(r[h[c]] or 0), c) for c in stall_cols), reverse=True)[:3]
    rs = ", ".join(f"{c[6:]}={v}" for v, c in reasons if v)
    print(f"  {n:7d} {100.0 * n / max(1, samples):5.1f}%  exec={r[h['Instructions Executed']]:>9s}  {r[h['Source']].strip()[:70]:70s} {rs}")
