"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per kernel count, total, mean, share.
    python tools/launch_shares.py gpurun_out/launches.csv "<command that was profiled>" > profiles/rNN_launch_shares.txt"""
import collections, csv, sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = rows[0]
iK, iV, iU = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot, cnt = collections.Counter(), collections.Counter()
for r in rows[1:]:
    try:
        v = float(r[iV].replace(",", ""))
    except ValueError:
        continue
    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[iU], 1e-3)
    tot[r[iK]] += v
    cnt[r[iK]] += 1
T = sum(tot.values())
print(f"ncu --metrics gpu__time_duration.sum --clock-control none : {sys.argv[2] if len(sys.argv) > 2 else ''}  (cold-cache, serialised: compare shares)")
for k, v in tot.most_common():
    print(f"{k[:70]:70s} n={cnt[k]:4d} total={v:10.1f}us avg={v / cnt[k]:9.2f}us share={v / T:6.3f}")
