"""profiles/ncu_traffic.json from `ncu --set full` captures of bench.py's timed region: DRAM bytes (read + write) per
launch of the dominant kernels at the bench shape.  bench.py puts them into `roofline.traffic`.
usage: python profiles/make_traffic.py T V rep1.ncu-rep [rep2.ncu-rep ...]"""
import csv, io, json, os, subprocess, sys
T, V = int(sys.argv[1]), int(sys.argv[2])
out_path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ncu_traffic.json")
tab = json.load(open(out_path)) if os.path.exists(out_path) else {}
unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
for rep in sys.argv[3:]:
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h, u = rows[0], rows[1]
    col = {k: i for i, k in enumerate(h)}
    per = {}
    for r in rows[2:]:
        name = r[col["Kernel Name"]]
        short = next((k for k in ("fused_rowres_kernel", "logits_fwd_tma_kernel", "logits_bwd_tma_kernel") if k in name), None)
        if short is None:
            continue
        rd = float(r[col["dram__bytes_read.sum"]]) * unit[u[col["dram__bytes_read.sum"]]]
        wr = float(r[col["dram__bytes_write.sum"]]) * unit[u[col["dram__bytes_write.sum"]]]
        per.setdefault(short, []).append((rd, wr, float(r[col["gpu__time_duration.sum"]])))
    for k, v in per.items():
        rd = sum(x[0] for x in v) / len(v)
        wr = sum(x[1] for x in v) / len(v)
        tab[f"{k}:{T}x{V}"] = {"dram_bytes": rd + wr, "dram_read": rd, "dram_write": wr, "launches": len(v),
                               "source": "profiles/" + os.path.basename(rep).replace(".ncu-rep", ".txt") + " (ncu --set full, bench.py timed region)"}
json.dump(tab, open(out_path, "w"), indent=1)
print(json.dumps(tab, indent=1))
