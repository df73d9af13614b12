"""ncu --set full capture of six consecutive in-step GEMM launches (forward of one fusion layer pair) -> per-launch DRAM
traffic next to the algorithmic bytes; bench.py reports the average as roofline.traffic.
Usage (here, after the GPU run):  ncu -i gpurun_out/X.ncu-rep --page raw --csv > /tmp/raw.csv
                                  python tools/gemm_traffic.py /tmp/raw.csv profiles/r1_gemm_traffic.json "<capture command>" """
import csv, json, sys

raw, out, cmd = sys.argv[1], sys.argv[2], (sys.argv[3] if len(sys.argv) > 3 else "")
rows = list(csv.reader(open(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}


def mb(d, key):
    v, u = float(d[ix[key]].replace(",", "")), units[ix[key]]
    return v * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}[u]


M, H, FF = 22848, 768, 3072  # text + image token rows at batch 64
names = ["fwd QKV", "fwd out-proj (+dropout+residual)", "fwd cross Q + KV", "fwd cross out-proj (+dropout+residual)",
         "fwd FFN-up (+GELU, gelu' side output)", "fwd FFN-down (+dropout+residual)"]
alg = [(M * H + 3 * H * H + M * 3 * H) * 2, (M * H + H * H + 2 * M * H) * 2, (2 * M * H + 3 * H * H + M * 3 * H) * 2,
       (M * H + H * H + 2 * M * H) * 2, (M * H + FF * H + 2 * M * FF) * 2, (M * FF + H * FF + 2 * M * H) * 2]
assert len(data) == 6, f"expected 6 launches, got {len(data)}"
launches = []
for n, a, d in zip(names, alg, data):
    t = mb(d, "dram__bytes_read.sum") + mb(d, "dram__bytes_write.sum")
    launches.append({"name": n, "us": float(d[ix["gpu__time_duration.sum"]]), "dram_mbytes": round(t, 1), "algorithmic_mbytes": round(a / 1e6, 1),
                     "tensor_pipe_active_pct": round(float(d[ix["sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]]), 1)})
res = {"source": cmd, "launches": launches,
       "avg_dram_bytes_per_launch": round(sum(l["dram_mbytes"] for l in launches) / 6 * 1e6),
       "avg_algorithmic_bytes_per_launch": round(sum(alg) / 6)}
json.dump(res, open(out, "w"), indent=1)
print(json.dumps(res, indent=1))
