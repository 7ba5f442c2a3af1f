"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: time and launches per kernel, shares of the total."""
import collections, csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, x in enumerate(rows) if x and x[0] == "ID"][0]
H = rows[hdr]; ki = H.index("Kernel Name"); vi = H.index("Metric Value"); ui = H.index("Metric Unit")
tot = collections.Counter(); cnt = collections.Counter()
for x in rows[hdr + 1:]:
    if len(x) <= vi:
        continue
    full = x[ki]
    m = re.search(r"\b(k_[a-z0-9_]+)", full)
    if m:
        name = m.group(1)
    else:
        m = re.search(r"cub::[A-Za-z0-9_:]*?(Device[A-Za-z]+Kernel|Device[A-Za-z0-9]+)", full)
        name = "cub::" + m.group(1) if m else re.sub(r"<.*", "", full)[:60]
    v = float(x[vi].replace(",", ""))
    v = v / 1e3 if x[ui] in ("ns", "nsecond") else (v * 1e3 if x[ui] in ("ms", "msecond") else v)
    tot[name] += v; cnt[name] += 1
T = sum(tot.values())
print("launches %d, total %.1f ms (cold-cache, serialised: compare shares, not absolutes)" % (sum(cnt.values()), T / 1e3))
for k, v in tot.most_common(40):
    print("%-36s %10.1f us %6.2f%%  n=%d" % (k[:36], v, 100 * v / T, cnt[k]))
