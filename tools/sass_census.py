"""Per-kernel census of the Blackwell-specific SASS opcodes in libmmvae_b200.so (cuobjdump -sass): tcgen05 MMAs (UTCHMMA),
TMA loads (UTMALDG), TMEM loads / stores (LDTM / STTM), tcgen05 commits (UTCBAR), bulk copies (UBLKCP).
    python tools/sass_census.py > profiles/r2_sass_census.txt"""
import collections, os, re, subprocess
lib = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "mmvae_b200", "csrc", "libmmvae_b200.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
ops = ["UTCHMMA", "UTMALDG", "LDTM", "STTM", "UTCBAR", "UBLKCP", "FFMA", "HFMA2"]
cnt, cur = collections.OrderedDict(), None
for ln in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", ln)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = re.sub(r"\(.*$", "", cur).replace("void ", "")
        cnt[cur] = collections.Counter()
        continue
    if cur:
        for o in ops:
            if re.search(r"\b" + o + r"\b", ln) or (o + ".") in ln:
                cnt[cur][o] += 1
print(f"{'kernel':64s} " + " ".join(f"{o:>8s}" for o in ops))
for k, c in cnt.items():
    if any(c[o] for o in ops[:6]):
        print(f"{k[:64]:64s} " + " ".join(f"{c[o]:8d}" for o in ops))
tot = collections.Counter()
for c in cnt.values():
    tot.update(c)
print(f"{'TOTAL (all ' + str(len(cnt)) + ' kernels)':64s} " + " ".join(f"{tot[o]:8d}" for o in ops))
