"""Per-CUDA-source-line stall samples of an .ncu-rep (needs -lineinfo + --import-source on)."""
import csv, io, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "sass,cuda", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = None; lines = []
for r in rows:
    if r and r[0] == "Line No": hdr = r; continue
    if hdr and r and r[0].isdigit():
        f=lambda v: int(v) if v.isdigit() else 0; lines.append((f(r[hdr.index("# Samples")]), f(r[hdr.index("Instructions Executed")]), int(r[0]), r[1]))
tot = sum(l[0] for l in lines)
for s, n, ln, src in sorted(lines, reverse=True)[:top]:
    print(f"{100*s/max(tot,1):5.1f}% {s:6d} {n:10d}  L{ln:<4d} {src.strip()[:120]}")
