"""Summarise `ncu -i X.ncu-rep --page source --csv` (SASS view): per kernel instance, executed warp-instructions by
opcode, and the instructions with most stall samples.  Usage: python tools/ncu_sass_summary.py src.csv [kernel_index]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
kernels, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "rows": []}
        kernels.append(cur)
    elif cur is not None and cur["hdr"] is None:
        cur["hdr"] = r
    elif cur is not None and r:
        cur["rows"].append(r)
sel = [int(sys.argv[2])] if len(sys.argv) > 2 else range(len(kernels))
for ki in sel:
    k = kernels[ki]
    h = {n: i for i, n in enumerate(k["hdr"])}
    ie, smp = h["Instructions Executed"], h["# Samples"]
    tot = sum(int(r[ie]) for r in k["rows"])
    by = collections.Counter()
    for r in k["rows"]:
        op = r[h["Source"]].split()
        op = [t for t in op if not t.startswith("@")]
        by[op[0].split(".")[0] if op else "?"] += int(r[ie])
    print(f"== kernel {ki}: {k['name'][:60]}  warp-instructions executed {tot}  samples {sum(int(r[smp]) for r in k['rows'])}")
    print("   by opcode:", ", ".join(f"{o} {n} ({100*n/tot:.1f}%)" for o, n in by.most_common(16)))
    top = sorted(k["rows"], key=lambda r: -int(r[smp]))[:14]
    stall_cols = [n for n in k["hdr"] if n.startswith("stall_") and "Not Issued" not in n]
    for r in top:
        st = sorted(((int(r[h[c]]), c) for c in stall_cols), reverse=True)[:2]
        print(f"   {r[smp]:>6} samples  exec {r[ie]:>8}  {r[h['Source']].strip()[:70]:70s} {st}")
