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
stream = sum(v for k, v in tot.items() if k.startswith("void rowstream_kernel<0"))
fin = sum(v for k, v in tot.items() if "finish_exchange_kernel" in k and ", 1>" in k)
if stream and fin:
    print("\nA CG step is two launches: `rowstream_kernel<0,...>` (mmchain XtwXv) and `finish_exchange_kernel<..., CG>` (fold of the per-CTA partials + the m-vector tail).")
    print("The streaming kernel's share of those two under ncu: %.4f." % (stream / (stream + fin)))
