"""Summarise an `ncu --page source --csv --print-source sass` dump: per-instruction share of executed warp
instructions and average active lanes. usage: srcpage.py file.csv [min_share_pct]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.2
hdr = next(r for r in rows if r and r[0] == "Address")
data = [r for r in rows if len(r) == len(hdr) and r[0].startswith("0x")]
ia, isrc, iex, ith, ismp = (hdr.index(k) for k in ("Address", "Source", "Instructions Executed", "Thread Instructions Executed", "# Samples"))
base = int(data[0][ia], 16)
tot = sum(int(r[iex]) for r in data); tth = sum(int(r[ith]) for r in data)
print("total warp inst", tot, "avg lanes %.2f" % (tth / tot))
for r in data:
    ex, th = int(r[iex]), int(r[ith])
    if ex > tot * thr / 100:
        print(hex(int(r[ia], 16) - base).ljust(7), r[isrc].strip()[:56].ljust(56), "%6.2f%%" % (100 * ex / tot), "lanes %4.1f" % (th / max(ex, 1)), "smp", r[ismp])
