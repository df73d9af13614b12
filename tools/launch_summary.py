"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel. Usage: python tools/launch_summary.py file.csv [skip] [count]"""
import collections, csv, re, sys
path = sys.argv[1]
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
count = int(sys.argv[3]) if len(sys.argv) > 3 else 10**9
lines = [l for l in open(path) if not l.startswith("==")]
rows = list(csv.DictReader(lines))[skip:skip + count]
tot, cnt = collections.defaultdict(float), collections.Counter()
for row in rows:
    full = row["Kernel Name"]
    name = re.sub(r"\(.*", "", full)
    if "at::" in full or "cutlass" in full or "cublas" in full.lower() or "nvjet" in full: name = "torch:" + name[:70]
    else: name = re.sub(r".*::", "", name)
    v = float(row["Metric Value"].replace(",", ""))
    unit = row["Metric Unit"]
    v = v / 1e3 if unit in ("ns", "nsecond") else v * 1e3 if unit in ("ms", "msecond") else v
    tot[name] += v; cnt[name] += 1
T = sum(tot.values())
print(f"launches {sum(cnt.values())}  total {T/1e3:.3f} ms")
for k, v in sorted(tot.items(), key=lambda x: -x[1])[:30]:
    print(f"{v:10.1f} us {100*v/T:5.1f}%  n={cnt[k]:5d}  avg={v/cnt[k]:8.1f}  {k}")
