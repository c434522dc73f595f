"""gpurun_out/gemm_traffic_{base,conformer}.csv (scripts/gpu_traffic.sh) -> profiles/gemm_traffic_r02.json"""
import csv, json, os, sys
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = {}
for w in ("base", "conformer"):
    path = os.path.join(REPO, "gpurun_out", f"gemm_traffic_{w}.csv")
    if not os.path.exists(path):
        continue
    with open(path) as f:
        rows = list(csv.DictReader(l for l in f if l.startswith('"')))
    per = {}
    for r in rows:
        per.setdefault(r["ID"], {})[r["Metric Name"]] = (float(r["Metric Value"].replace(",", "")), r["Metric Unit"])
    def b(v):
        x, u = v
        return x * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]
    def us(v):
        x, u = v
        return x * {"ns": 1e-3, "us": 1, "ms": 1e3, "usecond": 1, "nsecond": 1e-3, "msecond": 1e3}[u]
    n = len(per)
    rd = sum(b(m["dram__bytes_read.sum"]) for m in per.values()) / n
    wr = sum(b(m["dram__bytes_write.sum"]) for m in per.values()) / n
    t = sum(us(m["gpu__time_duration.sum"]) for m in per.values()) / n
    out[w] = {"launches_captured": n, "dram_bytes_per_launch": rd + wr, "dram_read_bytes_per_launch": rd,
              "dram_write_bytes_per_launch": wr, "avg_launch_us_under_ncu": t,
              "source": "ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum -k regex:gemm_tc_kernel "
                        f"-s 3000 -c 2400 on `python bench.py --workload {w} --steps 1 --warmup 3` (scripts/gpu_traffic.sh): average "
                        "over the captured gemm_tc_kernel launches (settling + timed steps of the round-2 pair kernel)"}
json.dump(out, open(os.path.join(REPO, "profiles", "gemm_traffic_r02.json"), "w"), indent=1)
print(json.dumps({k: {a: v[a] for a in ("launches_captured", "dram_bytes_per_launch", "avg_launch_us_under_ncu")} for k, v in out.items()}))
