"""Top SASS instructions by executed count and by stall samples from `ncu --page source --csv` output.
python tools/ncu_hot.py file.csv [n]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
hdr = rows[1]
ia, isrc, isamp, iex = hdr.index("Address"), hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
data = []
for r in rows[2:]:
    try:
        data.append((int(r[isamp] or 0), int(r[iex] or 0), r[ia], r[isrc]))
    except (ValueError, IndexError):
        pass
tot_s, tot_e = sum(d[0] for d in data), sum(d[1] for d in data)
print(f"total samples {tot_s}, total warp-instructions {tot_e}")
print("--- by samples")
for s, e, a, src in sorted(data, reverse=True)[:n]:
    print(f"{s:8d} {100*s/max(tot_s,1):5.1f}%  ex {e:10d}  {a[-5:]}  {src[:110]}")
print("--- by executed")
for s, e, a, src in sorted(data, key=lambda d: -d[1])[:n]:
    print(f"{s:8d}  ex {e:10d} {100*e/max(tot_e,1):5.1f}%  {a[-5:]}  {src[:110]}")
