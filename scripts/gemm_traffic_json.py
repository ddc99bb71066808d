"""ncu csv (dram__bytes_read.sum, dram__bytes_write.sum per gemm_tcgen05_kernel launch of one step) -> the JSON bench.py reads for
`roofline.traffic`:  python scripts/gemm_traffic_json.py gpurun_out/TAG_gemm_traffic.csv profiles/r02_gemm_traffic.json"""
import csv, json, sys
src, dst = sys.argv[1], sys.argv[2]
lines = [l for l in open(src) if not l.startswith("==")]
rd = wr = 0.0
ids = set()
for row in csv.DictReader(lines):
    v = float(row["Metric Value"].replace(",", ""))
    u = row["Metric Unit"]
    v *= {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1.0)
    if row["Metric Name"] == "dram__bytes_read.sum": rd += v; ids.add(row["ID"])
    elif row["Metric Name"] == "dram__bytes_write.sum": wr += v
n = len(ids)
json.dump({"source": f"{src} (ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum over the {n} gemm_tcgen05_kernel launches of one MAE ViT-B/16 step, B=256)",
           "launches": n, "dram_bytes_read": rd, "dram_bytes_write": wr, "traffic_per_launch": (rd + wr) / max(n, 1)}, open(dst, "w"), indent=1)
print(n, rd / 1e9, wr / 1e9, (rd + wr) / max(n, 1) / 1e6)
