"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` log of scripts/profile_resnet_step.py: per-kernel count and GPU time
of the LAST step (from the last stem im2col launch to the end)."""
import collections, csv, sys
path = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/resnet_launches.csv"
marker = sys.argv[2] if len(sys.argv) > 2 else "im2col_stem"
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
rows = [(r["Kernel Name"], float(r["Metric Value"])) for r in csv.DictReader(lines)]
starts = [i for i, (n, _) in enumerate(rows) if marker in n]
a = starts[-1]
agg = collections.defaultdict(lambda: [0, 0.0])
for n, v in rows[a:]:
    k = n.split("(")[0][:72]
    agg[k][0] += 1
    agg[k][1] += v / 1000
for k, (c, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:74s} {c:4d} {v:9.1f} us")
print("total us %.1f launches %d" % (sum(v for _, v in agg.values()), len(rows) - a))
