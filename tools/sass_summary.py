"""Per-kernel SASS census of libm3ae_b200.so: which kernels use tcgen05 (UTCHMMA / UTCBAR / LDTM / STTM), TMA (UTMALDG /
UTMASTG / UTMAREDG), mbarriers (SYNCS), legacy tensor-core MMA (HMMA) — the evidence that the hot kernels are sm_100a code.

    python tools/sass_summary.py > profiles/sass_summary.txt
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "mm-vqa-healthcare_b200", "libm3ae_b200.so")
MNEMONICS = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTCATOMSWS", "UTMALDG", "UTMASTG", "UTMAREDG", "UTMAPF", "SYNCS", "HMMA", "FFMA", "MUFU",
             "LDG", "STG", "LDS", "STS", "BAR", "REDG", "ATOMG", "LDL", "STL"]


def demangle(names):
    out = subprocess.run(["c++filt"], input="\n".join(names), capture_output=True, text=True).stdout.splitlines()
    return [re.sub(r"\(anonymous namespace\)::|m3ae::|<unnamed>::", "", o) for o in out]


def main():
    lib = sys.argv[1] if len(sys.argv) > 1 else LIB
    sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    kernels, cur = collections.OrderedDict(), None
    arch = set(re.findall(r"arch = (sm_\w+)", sass))
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)", line)
        if m:
            op = m.group(1)
            kernels[cur]["_total"] += 1
            for mn in MNEMONICS:
                if op == mn or op.startswith(mn + "."):
                    kernels[cur][mn] += 1
    names = demangle(list(kernels))
    print(f"# cuobjdump -sass {os.path.relpath(lib, ROOT)}   arch: {', '.join(sorted(arch))}   kernels: {len(kernels)}")
    print("# static instruction counts per kernel (mnemonic families); tcgen05 = UTCHMMA (MMA) / UTCBAR (commit) / LDTM,STTM (TMEM ld/st);")
    print("# TMA = UTMALDG / UTMASTG / UTMAREDG; mbarrier = SYNCS; legacy warp MMA = HMMA")
    for (k, c), n in zip(kernels.items(), names):
        short = n if len(n) < 150 else n[:147] + "..."
        tags = "  ".join(f"{mn}={c[mn]}" for mn in MNEMONICS if c[mn])
        print(f"{short}\n    instr={c['_total']}  {tags}")
    tc = [n for (k, c), n in zip(kernels.items(), names) if c["UTCHMMA"]]
    print(f"# kernels with tcgen05.mma: {len(tc)}")
    for n in tc:
        print("#   " + n[:140])


if __name__ == "__main__":
    main()
