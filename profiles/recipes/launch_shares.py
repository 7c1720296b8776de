"""Shares of a step from an ncu launch list (--metrics gpu__time_duration.sum --csv).
usage: launch_shares.py launches.csv "label" id_lo id_hi n_steps ["label" id_lo id_hi n_steps ...]"""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
hdr = rows[hi]
body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
kn, mv = hdr.index("Kernel Name"), hdr.index("Metric Value")
seq = {int(r[0]): (r[kn], float(r[mv].replace(",", "")) / 1e6) for r in body}
args = sys.argv[2:]
for q in range(0, len(args), 4):
    label, lo, hi_, steps = args[q], int(args[q + 1]), int(args[q + 2]), float(args[q + 3])
    agg = collections.OrderedDict()
    for i in range(lo, hi_ + 1):
        k, v = seq[i]
        a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v
    tot = sum(t for _, t in agg.values())
    print(f"# {label} (ids {lo}-{hi_}, {steps} steps, {tot / steps:.4f} ms of kernels per step)")
    for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"  {100 * t / tot:6.2f} %  {t / steps:9.4f} ms/step  x{c / steps:g}  {k[:120]}")
    print()
