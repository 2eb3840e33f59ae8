"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: last training step only."""
import collections, csv, re, sys
path = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/launches.csv"
lines = [l for l in open(path) if not l.startswith("==")]
rows = list(csv.DictReader(lines))
names = [r["Kernel Name"] for r in rows]
firsts = [i for i, n in enumerate(names) if "im2col" in n or "conv3x3_c3_fwd" in n]
start = max(firsts) if firsts else 0
agg = collections.OrderedDict()
total = 0.0
verbose = "-v" in sys.argv
for r in rows[start:]:
    v = float(r["Metric Value"].replace(",", ""))
    us = v / 1000 if r["Metric Unit"] in ("ns", "nsecond") else v
    short = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "").replace("b200dp::", "")[:48]
    if verbose:
        print(f"{us:8.1f} us  {short:48s} grid={r['Grid Size']}")
    a = agg.setdefault(short, [0, 0.0])
    a[0] += 1
    a[1] += us
    total += us
print("--- per kernel family (one step, serialised under ncu) ---")
for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{us:8.1f} us  {100 * us / total:5.1f}%  x{n:<3d} {k}")
print(f"{total:8.1f} us  total")
