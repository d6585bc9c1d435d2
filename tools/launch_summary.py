#!/usr/bin/env python
"""Per-kernel totals of an ncu launch list (`ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file x.csv ...`):
python tools/launch_summary.py x.csv "command that was profiled" > summary.md"""
import collections
import csv
import sys

path, cmd = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "?")
rows = [l for l in open(path) if l.startswith('"')]
tot, cnt = collections.defaultdict(float), collections.defaultdict(int)
r = csv.reader(rows)
head = next(r)
ki, vi = head.index("Kernel Name"), head.index("Metric Value")
for row in r:
    name = row[ki].replace("<unnamed>::", "")
    tot[name] += float(row[vi].replace(",", ""))
    cnt[name] += 1
allt = sum(tot.values())
print("# ncu launch list of `%s` (1x B200)\n" % cmd)
print("Command: `ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv` (raw CSV: `%s`), run after the same command had exited 0 without ncu." % path.split("/")[-1])
print("Per-launch times are cold-cache and serialised: compare SHARES with bench.py, not absolutes. Unit: ns.\n")
print("| kernel | launches | total | mean | share |\n|---|---:|---:|---:|---:|")
for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
    print("| `%s` | %d | %d | %d | %.4f |" % (k[:110], cnt[k], v, v / cnt[k], v / allt))
# one CG step = one full-size streaming launch + one finishing launch: compare per-launch means (the list also holds the
# short streaming launches of the cold / binaryblock legs, which are left out: only launches above half the longest count)
r = csv.reader(rows)
next(r)
sl, fl = [], []
for row in r:
    name, v = row[ki], float(row[vi].replace(",", ""))
    if "rowstream_kernel<0" in name:
        sl.append(v)
    elif "finish_exchange_kernel" in name and ", 1>" in name:
        fl.append(v)
if sl and fl:
    big = [v for v in sl if v > 0.5 * max(sl)]
    ms, mf = sum(big) / len(big), sum(fl) / len(fl)
    print("\nA CG step is two launches: `rowstream_kernel<0,...>` (mmchain XtwXv) and `finish_exchange_kernel<..., CG>` (fold of the per-CTA partials + the m-vector tail).")
    print("Per-launch means under ncu: %d full-size streaming launches %.0f ns, %d finishing launches %.0f ns: the streaming kernel's share of a step is %.4f." % (len(big), ms, len(fl), mf, ms / (ms + mf)))
