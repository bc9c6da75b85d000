"""Per-CUDA-source-line instruction counts of one kernel from an .ncu-rep captured with --import-source on (-lineinfo build).
usage: srclines.py rep.ncu-rep kernel-regex [min-percent]"""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name", "regex:" + sys.argv[2]],
                     capture_output=True, text=True).stdout
minpct = float(sys.argv[3]) if len(sys.argv) > 3 else 0.5
fname, rows, hdr = "", [], None
for r in csv.reader(out.splitlines()):
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]
    elif r[0] == "Line No":
        hdr = r
    elif hdr and r[0].isdigit() and len(r) > 8:
        i = hdr.index("Instructions Executed"); t = hdr.index("Thread Instructions Executed"); s = hdr.index("# Samples")
        if r[i].isdigit():
            rows.append((fname, int(r[0]), r[1].strip(), int(r[i]), int(r[t]), int(r[s]) if r[s].isdigit() else 0))
seen = {}
for f, ln, src, ex, th, sm in rows:      # a line appears once per file section; keep the first
    seen.setdefault((f, ln), (src, ex, th, sm))
tot = sum(v[1] for v in seen.values()); tsm = sum(v[3] for v in seen.values())
print("total warp instructions", tot, "samples", tsm)
for (f, ln), (src, ex, th, sm) in sorted(seen.items(), key=lambda kv: -kv[1][1]):
    if 100.0 * ex / tot < minpct:
        break
    print(f"{100.0 * ex / tot:5.1f}%  smp {100.0 * sm / max(tsm, 1):5.1f}%  lanes {th / max(ex, 1):4.1f}  {f}:{ln}  {src[:110]}")
