"""Per-source-line stall samples of an .ncu-rep (needs -lineinfo and --import-source on).
    python tools/ncu_lines.py gpurun_out/x.ncu-rep [top-n]"""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"],
                     capture_output=True, text=True).stdout
hdr, agg, src = None, collections.Counter(), {}
reasons = collections.defaultdict(collections.Counter)
for row in csv.reader(io.StringIO(out)):
    if row and row[0] == "Line No":
        hdr = row
        iS = hdr.index("# Samples")
        stall = [(i, n) for i, n in enumerate(hdr) if n.startswith("stall_")]
        continue
    if hdr is None or len(row) != len(hdr) or not row[0].isdigit():
        continue
    n = int(row[iS] or 0)
    ln = int(row[0])
    agg[ln] += n
    src[ln] = row[1].strip()
    for i, nm in stall:
        v = int(row[i] or 0)
        if v:
            reasons[ln][nm[6:]] += v
tot = sum(agg.values())
print("total samples", tot)
for ln, n in agg.most_common(topn):
    rs = ", ".join(f"{k}={v}" for k, v in reasons[ln].most_common(3))
    print(f"{n:7d} {100.0 * n / max(1, tot):5.1f}%  L{ln:<4d} {src[ln][:80]:80s} {rs}")
