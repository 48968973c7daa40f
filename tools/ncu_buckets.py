"""Stall samples of `ncu --page source --csv` (SASS view) summed over buckets of consecutive instructions, with each
bucket's most common opcodes: shows which region (= warp role) of a warp-specialised kernel the time goes to.
python tools/ncu_buckets.py file.csv [bucket=48]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
B = int(sys.argv[2]) if len(sys.argv) > 2 else 48
hdr = rows[1]
ia, isrc, isamp, iex = hdr.index("Address"), hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
seen, data = set(), []
for r in rows[2:]:
    try:
        if r[ia] in seen:
            continue
        seen.add(r[ia])
        data.append((int(r[ia], 16) if r[ia].startswith("0x") else int(r[ia]), int(r[isamp] or 0), int(r[iex] or 0), r[isrc]))
    except (ValueError, IndexError):
        pass
data.sort()
tot = sum(d[1] for d in data)
print(f"{len(data)} instructions, {tot} samples, {sum(d[2] for d in data)} warp-instructions executed")
for i in range(0, len(data), B):
    chunk = data[i:i + B]
    s = sum(d[1] for d in chunk)
    if s < 0.01 * tot:
        continue
    ops = collections.Counter()
    for d in chunk:
        t = d[3].split()
        op = t[1] if t and t[0].startswith("@") and len(t) > 1 else (t[0] if t else "")
        ops[op.split(".")[0]] += d[1]
    print(f"[{i:5d}..{i+len(chunk):5d}] {s:6d} {100*s/tot:5.1f}%  ex/instr {sum(d[2] for d in chunk)//len(chunk):9d}  " +
          " ".join(f"{k}:{v}" for k, v in ops.most_common(6)))
