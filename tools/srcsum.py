"""Region summary of an ncu source-page dump: usage: srcsum.py rep.ncu-rep [kernel-regex] [chunk]"""
import csv, subprocess, sys
args = ["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "sass"]
if len(sys.argv) > 2 and sys.argv[2]:
    args += ["--kernel-name", "regex:" + sys.argv[2]]
chunk = int(sys.argv[3]) if len(sys.argv) > 3 else 16
rows = list(csv.reader(subprocess.run(args, capture_output=True, text=True).stdout.splitlines()))
hdr = next(r for r in rows if r and r[0] == "Address")
data = [r for r in rows if len(r) == len(hdr) and r[0].startswith("0x")]
ia, isrc, iex, ith = (hdr.index(k) for k in ("Address", "Source", "Instructions Executed", "Thread Instructions Executed"))
seen = {}
for r in data:
    seen.setdefault(r[ia], r)
data = list(seen.values())
base = int(data[0][ia], 16)
tot = sum(int(r[iex]) for r in data); tth = sum(int(r[ith]) for r in data)
print(len(data), "instructions;", tot, "warp-inst executed; avg lanes %.2f" % (tth / tot))
acc = n = th = 0; start = None; first = ""
for r in data:
    if start is None:
        start = int(r[ia], 16) - base; first = r[isrc].strip()[:40]
    acc += int(r[iex]); th += int(r[ith]); n += 1
    if n == chunk:
        if acc > tot * 0.004:
            print(hex(start).ljust(8), "%6.2f%%" % (100 * acc / tot), "lanes %4.1f" % (th / max(acc, 1)), " ", first)
        acc = n = th = 0; start = None
